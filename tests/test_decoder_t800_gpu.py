"""GPU parity tests for the CFM decoder at the BENCHMARKED configuration (BASELINE configs[2]: T_mel = 800, Euler n = 10)
and through the sub-batch stream split the solve takes for B >= 8 (csrc/decoder.cu::solve_split).

Checkers: (i) tests/golden/decoder_golden_T800.npz, the output of the reference's own modules at B = 2, T = 800, n = 10
(make_decoder_golden.py --t800); (ii) oracle/decoder.py in fp64 on the same seeded inputs (moved to the GPU: the same
torch restatement, only the device differs).  Tolerances from BASELINE.json north_star: relative L2 <= 1e-5 in fp32
mode, <= 2e-3 in the 16-bit tensor-core mode."""
import os
import subprocess
import sys
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from oracle import decoder as odec

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL = {"fp32": 1e-5, "fp16": 2e-3}
TOL_BF16_FUNCTIONAL = 1.5e-2       # plain bf16 operands: a functional bound, NOT the north_star bar (see the xfail below)
PARAMS = dict(channels=(256, 256), dropout=0.05, attention_head_dim=64, n_blocks=1, num_mid_blocks=2, num_heads=4)


def rel_l2(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float(torch.linalg.norm(a - b) / torch.linalg.norm(b))


def make_cfm(weights, precision):
    from matcha_b200.flow_matching import CFM
    m = CFM(160, 80, SimpleNamespace(solver="euler", sigma_min=1e-4), PARAMS, precision=precision)
    m.load_state_dict({"estimator." + k: v for k, v in weights.items()})
    return m


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(ROOT, "tests", "golden", "decoder_golden_T800.npz"))


@pytest.fixture(scope="module")
def weights(gold):
    return odec.make_weights(int(gold["weight_seed"][0]), "rich")


@pytest.fixture(scope="module")
def w64(weights):
    return odec.cast_weights({k: v.cuda() for k, v in weights.items()}, torch.float64)


@pytest.fixture(scope="module")
def cfms(weights):
    return {p: make_cfm(weights, p) for p in ("fp32", "fp16", "bf16")}


def inputs(seed, B, T, lengths):
    g = torch.Generator().manual_seed(seed)
    mu, z = torch.randn(B, 80, T, generator=g), torch.randn(B, 80, T, generator=g)
    mask = torch.zeros(B, 1, T)
    for i, n in enumerate(lengths):
        mask[i, :, : int(n)] = 1
    return z.cuda(), mu.cuda(), mask.cuda()


def oracle_solve(w64, z, t_span, mu, mask):
    return odec.solve_euler(w64, z.double(), t_span.cpu(), mu.double(), mask.double())


# ------------------------------------------------------------------ (i) the reference's own output at T = 800, n = 10
@pytest.mark.parametrize("precision", ["fp32", "fp16"])
def test_t800_n10_matches_reference_golden(gold, cfms, precision):
    z, mu, mask = inputs(int(gold["seed"][0]), 2, 800, gold["lengths"])
    t_span = torch.linspace(0, 1, int(gold["n_timesteps"][0]) + 1).cuda()
    out = cfms[precision].solve_ode_euler(z, t_span, mu, mask, None, None)
    err = rel_l2(out, torch.from_numpy(gold["out"]))
    print("T800 golden", precision, err)
    assert err < TOL[precision]
    pad = mask[:, 0, :] == 0
    assert torch.equal(out.transpose(1, 2)[pad], z.transpose(1, 2)[pad]), "padded frames keep the initial noise"


def test_t800_n10_bf16_functional(gold, cfms):
    z, mu, mask = inputs(int(gold["seed"][0]), 2, 800, gold["lengths"])
    out = cfms["bf16"].solve_ode_euler(z, torch.linspace(0, 1, 11).cuda(), mu, mask, None, None)
    assert rel_l2(out, torch.from_numpy(gold["out"])) < TOL_BF16_FUNCTIONAL


@pytest.mark.xfail(strict=True, reason="north_star asks <= 2e-3 for the bf16 mode; plain bf16 operands (8 mantissa bits, ~50 GEMMs deep) "
                                       "measure ~5e-3.  The fp16-operand mode is the one that meets the bar (DESIGN.md, Precision).")
def test_t800_n10_bf16_meets_north_star_tolerance(gold, cfms):
    z, mu, mask = inputs(int(gold["seed"][0]), 2, 800, gold["lengths"])
    out = cfms["bf16"].solve_ode_euler(z, torch.linspace(0, 1, 11).cuda(), mu, mask, None, None)
    assert rel_l2(out, torch.from_numpy(gold["out"])) < 2e-3


# ------------------------------------------------------------------ (ii) fp64 oracle, ragged and full masks, T = 800, n = 10
@pytest.mark.parametrize("lengths", [[800, 800, 800, 800], [800, 705, 571, 402]], ids=["full", "ragged"])
@pytest.mark.parametrize("precision", ["fp32", "fp16"])
def test_b4_t800_n10_vs_fp64_oracle(w64, cfms, precision, lengths):
    z, mu, mask = inputs(4800, 4, 800, lengths)
    t_span = torch.linspace(0, 1, 11).cuda()
    want = oracle_solve(w64, z, t_span, mu, mask)
    got = cfms[precision].solve_ode_euler(z, t_span, mu, mask, None, None)
    err = rel_l2(got, want)
    print("B4 T800", precision, lengths[1], err)
    assert err < TOL[precision]


# ------------------------------------------------------------------ the sub-batch split (B >= 8): forward and solve
@pytest.mark.parametrize("B,T", [(8, 136), (9, 201), (13, 64)])
@pytest.mark.parametrize("precision", ["fp32", "fp16"])
def test_split_forward_per_sample_t_vs_fp64_oracle(w64, cfms, precision, B, T):
    """B = 8 / 9 / 13: even and uneven sub-batches; per-sample t (the compute_loss call), ragged masks, odd T."""
    g = torch.Generator().manual_seed(B * 131 + T)
    lengths = [T] + [int(T * (0.4 + 0.6 * float(torch.rand((), generator=g)))) for _ in range(B - 1)]
    x, mu, mask = inputs(B * 7 + T, B, T, lengths)
    t = torch.rand(B, generator=g).cuda()
    want = odec.decoder_forward(w64, x.double(), mask.double(), mu.double(), t.double())
    for graph in (True, False):
        cfms[precision].estimator.graph_forward = graph
        for _ in range(3 if graph else 1):          # a shape is captured the second time it is seen, replayed after
            got = cfms[precision].estimator(x, mask, mu, t)
        cfms[precision].estimator.graph_forward = True
        err = rel_l2(got, want)
        print("split forward", precision, B, T, graph, err)
        assert err < TOL[precision]


@pytest.mark.parametrize("B,T", [(8, 96), (9, 130)])
@pytest.mark.parametrize("precision", ["fp32", "fp16"])
def test_split_solve_vs_fp64_oracle(w64, cfms, precision, B, T):
    g = torch.Generator().manual_seed(B + T)
    lengths = [T] + [int(T * (0.5 + 0.5 * float(torch.rand((), generator=g)))) for _ in range(B - 1)]
    z, mu, mask = inputs(B * 11 + T, B, T, lengths)
    t_span = torch.linspace(0, 1, 4).cuda()
    want = oracle_solve(w64, z, t_span, mu, mask)
    for use_graph in (True, False):
        got = cfms[precision].estimator.solve_euler(z, t_span, mu, mask, use_graph=use_graph)
        err = rel_l2(got, want)
        print("split solve", precision, B, T, use_graph, err)
        assert err < TOL[precision]


@pytest.mark.parametrize("precision", ["fp32", "fp16"])
def test_b32_t800_n10_slices_vs_fp64_oracle(w64, cfms, precision):
    """The bench configuration itself (B = 32: four sub-batch streams, CUDA-graph replay): samples of the first,
    a middle and the last sub-batch against the fp64 oracle run on those samples alone (batch items never interact)."""
    g = torch.Generator().manual_seed(32800)
    lengths = [800] + [int(800 * (0.7 + 0.3 * float(torch.rand((), generator=g)))) for _ in range(31)]
    z, mu, mask = inputs(32800, 32, 800, lengths)
    t_span = torch.linspace(0, 1, 11).cuda()
    got = cfms[precision].solve_ode_euler(z, t_span, mu, mask, None, None)
    got2 = cfms[precision].solve_ode_euler(z, t_span, mu, mask, None, None)       # graph replay
    idx = [0, 7, 8, 19, 31]
    want = oracle_solve(w64, z[idx], t_span, mu[idx], mask[idx])
    err, err2 = rel_l2(got[idx], want), rel_l2(got2[idx], want)
    print("B32 T800", precision, err, err2)
    assert err < TOL[precision] and err2 < TOL[precision]


@pytest.mark.parametrize("precision", ["fp16", "fp32"])
def test_replayed_b32_solves_are_bit_identical(cfms, precision):
    """Stress for the programmatic-dependent-launch chain (round-2 review): the bench configuration (B = 32, T = 800, n = 10,
    four sub-batch streams, graph replay) solved over and over must return the very same bits every time.  A kernel that
    reads a chain-written tensor before its prerequisite has flushed it, or out of a stale L1 line, shows up as a replay
    that differs on some box (the text encoder's first PDL build did, profiles/r2_te_pdl_race.log)."""
    g = torch.Generator().manual_seed(4242)
    lengths = [800] + [int(800 * (0.7 + 0.3 * float(torch.rand((), generator=g)))) for _ in range(31)]
    z, mu, mask = inputs(4242, 32, 800, lengths)
    t_span = torch.linspace(0, 1, 11)
    n_rep = 200 if precision == "fp16" else 12                 # ~9 ms and ~150 ms per solve
    first = cfms[precision].solve_ode_euler(z, t_span, mu, mask, None, None).clone()
    assert bool(torch.isfinite(first).all())
    differing = 0
    for _ in range(n_rep):
        differing += int(not torch.equal(cfms[precision].solve_ode_euler(z, t_span, mu, mask, None, None), first))
    assert differing == 0, f"{differing} of {n_rep} replays differ from the first solve"


def test_graph_replay_survives_other_shapes(cfms):
    """A cached graph owns its workspace: plain launches and captures of other shapes in between must not disturb a
    later replay (round-1 bug: the shared workspace was freed while a graph still pointed at it)."""
    est = cfms["fp16"].estimator
    zA, muA, maskA = inputs(1, 8, 96, [96] * 8)
    tsA = torch.linspace(0, 1, 3).cuda()
    first = est.solve_euler(zA, tsA, muA, maskA).clone()
    zB, muB, maskB = inputs(2, 3, 200, [200, 150, 99])
    est.solve_euler(zB, torch.linspace(0, 1, 5).cuda(), muB, maskB, use_graph=False)
    est.solve_euler(zB, torch.linspace(0, 1, 2).cuda(), muB, maskB)
    for _ in range(3):
        est(zB, maskB, muB, torch.rand(3).cuda())
    junk = [torch.randn(64 << 20, device="cuda") for _ in range(4)]       # reuse whatever the allocator freed
    again = est.solve_euler(zA, tsA, muA, maskA)
    del junk
    assert rel_l2(again, first) < 1e-6


def test_split_off_equals_default():
    """MB200_SPLIT=1 (one stream, whole batch) against the default sub-batch split: same kernels on the same rows, only
    the GroupNorm atomics' order differs."""
    code = (
        "import sys, torch\n"
        f"sys.path[:0] = [{ROOT!r}, {os.path.join(ROOT, 'matcha-tts-etu-upmc-ensam_b200')!r}]\n"
        "from types import SimpleNamespace\n"
        "from oracle import decoder as odec\n"
        "from matcha_b200.flow_matching import CFM\n"
        f"m = CFM(160, 80, SimpleNamespace(solver='euler', sigma_min=1e-4), {PARAMS!r}, precision=sys.argv[1])\n"
        "m.load_state_dict(odec.make_weights(11, 'rich'))\n"
        "g = torch.Generator().manual_seed(5)\n"
        "mu, z = torch.randn(12, 80, 160, generator=g).cuda(), torch.randn(12, 80, 160, generator=g).cuda()\n"
        "mask = torch.ones(12, 1, 160).cuda(); mask[3, :, 100:] = 0; mask[11, :, 41:] = 0\n"
        "out = m.solve_ode_euler(z, torch.linspace(0, 1, 4).cuda(), mu, mask, None, None)\n"
        "torch.save(out.cpu(), sys.argv[2])\n")
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        for prec in ("fp32", "fp16"):
            outs = []
            for split in ("1", "4"):
                path = os.path.join(d, f"{prec}_{split}.pt")
                env = dict(os.environ, MB200_SPLIT=split)
                subprocess.run([sys.executable, "-c", code, prec, path], check=True, env=env, timeout=600)
                outs.append(torch.load(path))
            err = rel_l2(outs[0], outs[1])
            print("split 1 vs 4", prec, err)
            assert err < 1e-6
