"""GPU parity tests for the MAS kernel: bit-exact against the golden vectors (made from the
reference's compiled core.pyx) and against the C oracle on seeded inputs.  Everything goes
through the C ABI (ctypes) behind matcha_b200.monotonic_align."""
import numpy as np
import pytest
import torch

from oracle import mas as omas
from tests.mas_cases import check_path_properties, gen, golden_case, golden_full

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ma():
    assert torch.cuda.is_available()
    from matcha_b200 import monotonic_align
    return monotonic_align


def _run(ma, v, m, **kw):
    dev = torch.device("cuda:0")
    out = ma.maximum_path(torch.from_numpy(v).to(dev), torch.from_numpy(m).to(dev), **kw)
    torch.cuda.synchronize()
    assert out.dtype == torch.float32 and out.shape == v.shape
    o = out.cpu().numpy()
    assert np.isin(o, (0.0, 1.0)).all()
    return o.astype(np.int32)


@pytest.mark.parametrize("name", sorted(gen.cases().keys()))
def test_golden_full_mask_mode(ma, mas_golden, name):
    v, m, p = golden_case(mas_golden, name)
    assert np.array_equal(_run(ma, v, m, mask_mode="full"), p)


@pytest.mark.parametrize("name", sorted(k for k in gen.cases().keys() if k != "holey_mask"))
def test_golden_prefix_mask_mode(ma, mas_golden, name):
    v, m, p = golden_case(mas_golden, name)
    assert np.array_equal(_run(ma, v, m, mask_mode="prefix"), p)


def test_golden_config1_full_size(ma, mas_golden):
    v, m, p = golden_full(mas_golden)
    assert np.array_equal(_run(ma, v, m, mask_mode="prefix"), p)
    assert np.array_equal(_run(ma, v, m, mask_mode="full"), p)


@pytest.mark.parametrize("name", ["randn_ragged", "quantised_ties", "tile_edges", "tall_450x520"])
def test_host_buffer_c_abi_matches_golden(ma, mas_golden, name):
    # compute_batch_alignments(paths, values, t_xs, t_ys) with numpy buffers, as core.pyx:101-107
    v, m, p = golden_case(mas_golden, name)
    vals = np.ascontiguousarray((v * m).astype(np.float32))
    keep = vals.copy()
    paths = np.zeros(v.shape, dtype=np.int32)            # pre-zeroed by the caller, as in __init__.py:49
    t_x, t_y = m.sum(1)[:, 0].astype(np.int32), m.sum(2)[:, 0].astype(np.int32)
    bad = ma.compute_batch_alignments(paths, vals, t_x, t_y)
    assert bad == 0 and np.array_equal(paths, p)
    assert np.array_equal(vals.view(np.uint32), keep.view(np.uint32)), "values must not be clobbered"


SHAPES = [  # (B, T_x, T_y): every rows-per-lane variant, TMA and non-TMA (T_y % 4 != 0), multi row block
    (3, 20, 64), (5, 70, 333), (4, 130, 404), (3, 200, 1000), (2, 260, 801), (2, 330, 700),
    (2, 400, 2000), (2, 450, 1203), (1, 900, 1500), (9, 31, 32), (2, 1, 5), (3, 97, 97),
]


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("ties", [False, True])
def test_random_vs_oracle(ma, shape, ties):
    B, T_x, T_y = shape
    rng = np.random.default_rng(hash((B, T_x, T_y, ties)) % (2 ** 31))
    v = rng.standard_normal((B, T_x, T_y), dtype=np.float32)
    if ties:
        v = np.round(v * 2).astype(np.float32)
    m = gen.ragged_mask(rng, B, T_x, T_y, lo=0.5)
    want = omas.maximum_path_oracle(v, m)
    for mode in ("prefix", "full"):
        got = _run(ma, v, m, mask_mode=mode)
        assert np.array_equal(got, want), f"{shape} ties={ties} mode={mode}"


def test_int32_paths_and_unaligned_base(ma):
    rng = np.random.default_rng(5)
    B, T_x, T_y = 3, 40, 120
    dev = torch.device("cuda:0")
    big = torch.from_numpy(rng.standard_normal(B * T_x * T_y + 1, dtype=np.float32)).to(dev)
    v = big[1:].view(B, T_x, T_y)                       # 4-byte aligned only -> cp.async path
    assert v.data_ptr() % 16 != 0
    m = gen.ragged_mask(rng, B, T_x, T_y)
    t_x = torch.from_numpy(m.sum(1)[:, 0].astype(np.int32)).to(dev)
    t_y = torch.from_numpy(m.sum(2)[:, 0].astype(np.int32)).to(dev)
    got = ma.maximum_path_from_lengths(v, t_x, t_y, out_dtype=torch.int32)
    want = omas.maximum_path_oracle(v.cpu().numpy(), m)
    assert got.dtype == torch.int32 and np.array_equal(got.cpu().numpy(), want)


def test_out_of_domain_lengths_give_zero_path_and_status(ma):
    dev = torch.device("cuda:0")
    v = torch.randn(4, 8, 8, device=dev)
    t_x = torch.tensor([0, 8, 3, 9], dtype=torch.int32, device=dev)
    t_y = torch.tensor([8, 7, 8, 8], dtype=torch.int32, device=dev)
    out = ma.maximum_path_from_lengths(v, t_x, t_y)
    assert out[[0, 1, 3]].abs().sum().item() == 0 and out[2].sum().item() == 8
    with pytest.raises(ValueError):
        ma.maximum_path_from_lengths(v, t_x, t_y, check_lengths=True)


def test_half_precision_and_cpu_tensor_inputs(ma):
    rng = np.random.default_rng(11)
    v = rng.standard_normal((2, 12, 40), dtype=np.float32)
    m = gen.ragged_mask(rng, 2, 12, 40)
    vh = torch.from_numpy(v).half()
    want = omas.maximum_path_oracle((vh * torch.from_numpy(m).half()).float().numpy(), m)
    out = ma.maximum_path(vh.cuda(), torch.from_numpy(m).half().cuda())
    assert out.dtype == torch.float16 and np.array_equal(out.float().cpu().numpy().astype(np.int32), want)
    out_cpu = ma.maximum_path(torch.from_numpy(v), torch.from_numpy(m))   # CPU in -> CPU out, computed on the GPU
    assert out_cpu.device.type == "cpu"
    assert np.array_equal(out_cpu.numpy().astype(np.int32), omas.maximum_path_oracle(v, m))


@pytest.mark.parametrize("shape", [(64, 400, 2000), (512, 128, 800)])
def test_full_size_properties_and_sampled_parity(ma, shape):
    """BASELINE.json configs[1] sizes: invariants on every item, oracle parity on a sample."""
    B, T_x, T_y = shape
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev); g.manual_seed(3)
    v = torch.randn(B, T_x, T_y, device=dev, generator=g)
    rng = np.random.default_rng(9)
    t_x = np.maximum(1, (T_x * (0.7 + 0.3 * rng.random(B))).astype(np.int32))
    t_y = np.maximum(t_x, (T_y * (0.7 + 0.3 * rng.random(B))).astype(np.int32)).astype(np.int32)
    out = ma.maximum_path_from_lengths(v, torch.from_numpy(t_x).to(dev), torch.from_numpy(t_y).to(dev),
                                       out_dtype=torch.int32)
    again = ma.maximum_path_from_lengths(v, torch.from_numpy(t_x).to(dev), torch.from_numpy(t_y).to(dev),
                                         out_dtype=torch.int32)
    assert torch.equal(out, again), "deterministic"
    # column sums / durations on the device, then the full invariants on a sample
    col = out.sum(1)
    yy = torch.arange(T_y, device=dev)[None, :]
    assert torch.equal(col, (yy < torch.from_numpy(t_y).to(dev)[:, None]).to(col.dtype))
    idx = list(range(0, B, max(1, B // 16)))
    sub = out[idx].cpu().numpy()
    check_path_properties(sub, t_x[idx], t_y[idx])
    vs = np.ascontiguousarray(v[idx].cpu().numpy())
    want = np.zeros(vs.shape, np.int32)
    omas.oracle_compute_batch_alignments(want, vs, t_x[idx], t_y[idx])
    assert np.array_equal(sub, want)


# Paths that run along the band edges: a strongly preferred token keeps the optimum on one row until the
# band (core.pyx:58) forces it off, and log-prior-sized scores (-150 per cell, matcha_tts.py:277-282) make
# the sentinel arithmetic non-absorbing.  (t_y - t_x) % 32 == 0 puts the hand-over between two DP warps
# exactly on a chunk boundary (regression: the value the lower row block reads there was never sent).
BAND_SHAPES = [(1, 128, 800), (4, 128, 800), (3, 200, 1000), (2, 64, 200), (2, 96, 96 + 64), (1, 400, 2000),
               (2, 160, 160 + 32), (3, 130, 130 + 320), (2, 33, 65),
               # larger rows-per-lane variants of the kernel (T_x 450 / 900 / 1300), chunk-aligned band edges
               (2, 450, 450 + 96), (1, 900, 900 + 640), (1, 1300, 1300 + 32)]


@pytest.mark.parametrize("shape", BAND_SHAPES)
@pytest.mark.parametrize("kind", ["bonus_row", "log_prior_scale", "both", "late_bonus"])
def test_band_edge_paths_vs_oracle(ma, shape, kind):
    B, T_x, T_y = shape
    rng = np.random.default_rng(hash((B, T_x, T_y, kind)) % (2 ** 31))
    v = rng.standard_normal((B, T_x, T_y), dtype=np.float32)
    if kind in ("log_prior_scale", "both"):
        v = (v * 10 - 150).astype(np.float32)
    if kind in ("bonus_row", "both"):
        v[:, T_x // 3, :] += 30.0 if kind == "both" else 5.0
    if kind == "late_bonus":                      # the optimum hugs the upper edge (x == y), then one row
        v[:, T_x - 2, :] += 5.0
    m = np.ones_like(v)
    m[1:] = gen.ragged_mask(rng, B, T_x, T_y, lo=0.6)[1:]          # item 0 keeps the full lengths
    want = omas.maximum_path_oracle(v, m)
    for mode in ("prefix", "full"):
        got = _run(ma, v, m, mask_mode=mode)
        assert np.array_equal(got, want), f"{shape} {kind} mode={mode}"


def test_band_edge_all_offsets(ma):
    """t_y - t_x swept over a whole chunk period with a preferred row: every alignment of the band edge to the
    32-frame chunks and to the row blocks."""
    rng = np.random.default_rng(123)
    T_x, B = 100, 33
    T_y = T_x + 64 + B
    v = rng.standard_normal((B, T_x, T_y), dtype=np.float32)
    v[:, 20, :] += 4.0
    m = np.zeros_like(v)
    for b in range(B):
        m[b, :T_x, : T_x + 64 + b] = 1
    want = omas.maximum_path_oracle(v, m)
    assert np.array_equal(_run(ma, v, m, mask_mode="prefix"), want)


def test_band_edge_many_utterances_per_cta(ma):
    """More utterances than resident CTAs (the steady-state writer path, bit words recycled between utterances)
    with every item on the chunk-aligned band edge."""
    rng = np.random.default_rng(77)
    B, T_x, T_y = 700, 64, 192
    v = (rng.standard_normal((B, T_x, T_y), dtype=np.float32) * 10 - 150).astype(np.float32)
    v[np.arange(B), rng.integers(0, T_x, B), :] += 25.0
    m = np.ones_like(v)
    m[::3] = gen.ragged_mask(rng, B, T_x, T_y, lo=0.6)[::3]
    want = omas.maximum_path_oracle(v, m)
    assert np.array_equal(_run(ma, v, m, mask_mode="prefix"), want)


# --------------------------------------------------------------------------------------------------------------------
# Inputs taller than the wavefront kernel takes (T_x > 1344): the one-CTA-per-utterance fallback (mas_tall_kernel).  The
# reference accepts any T_x (core.pyx:16-96), so these are parity cases like any other.
@pytest.mark.parametrize("mask_mode", ["prefix", "full"])
def test_tall_inputs_take_the_fallback_kernel(ma, mask_mode):
    rng = np.random.default_rng(77)
    B, T_x, T_y = 3, 1400, 1500
    v = rng.standard_normal((B, T_x, T_y), dtype=np.float32)
    m = np.zeros((B, T_x, T_y), np.float32)
    lens = [(1400, 1500), (1377, 1377), (901, 1432)]                               # full length, t_x == t_y, ragged
    for b, (tx, ty) in enumerate(lens):
        m[b, :tx, :ty] = 1
    want = omas.maximum_path_oracle(v, m)
    got = _run(ma, v, m, mask_mode=mask_mode)
    assert np.array_equal(got, want)
    check_path_properties(got, [l[0] for l in lens], [l[1] for l in lens])


def test_tall_inputs_ties_and_durations(ma):
    """Dense ties (quantised scores) through the fallback, and the per-token durations it emits."""
    from matcha_b200 import alignment as al
    rng = np.random.default_rng(78)
    B, T_x, T_y = 2, 1350, 1400
    v = np.round(rng.standard_normal((B, T_x, T_y)) * 2).astype(np.float32)
    tx, ty = np.array([1350, 1203], np.int32), np.array([1400, 1399], np.int32)
    m = np.zeros((B, T_x, T_y), np.float32)
    for b in range(B):
        m[b, :tx[b], :ty[b]] = 1
    want = omas.maximum_path_oracle(v, m)
    dev = torch.device("cuda:0")
    path, dur = al.maximum_path_with_durations(torch.from_numpy(v).to(dev), torch.from_numpy(tx).to(dev), torch.from_numpy(ty).to(dev))
    torch.cuda.synchronize()
    assert np.array_equal(path.cpu().numpy().astype(np.int32), want)
    assert np.array_equal(dur.cpu().numpy(), want.sum(-1).astype(np.int32))
