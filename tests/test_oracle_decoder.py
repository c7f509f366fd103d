"""CPU tests: pin the torch restatement of the CFM decoder path (oracle/decoder.py) against golden
outputs produced by the reference's own modules (tests/golden/make_decoder_golden.py)."""
import os

import numpy as np
import pytest
import torch

from oracle import decoder as odec

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "decoder_golden.npz"))


@pytest.fixture(scope="module")
def weights(golden):
    return odec.make_weights(int(golden["weight_seed"][0]), "rich")


def rel_l2(a, b):
    return float(torch.linalg.norm((a - b).double()) / torch.linalg.norm(b.double()))


def T(golden, key):
    return torch.from_numpy(golden[key])


def test_param_count_matches_reference():
    n = sum(int(np.prod(s)) for s in odec.param_shapes().values())
    assert n == 11_782_992          # SURVEY §3.3 [probe]


@pytest.mark.parametrize("name", ["eval_T37_tvec", "eval_T44_full"])
def test_decoder_eval_matches_reference(golden, weights, name):
    torch.set_num_threads(4)
    out = odec.decoder_forward(weights, T(golden, name + "/x"), T(golden, name + "/mask"), T(golden, name + "/mu"),
                               T(golden, name + "/t"))
    assert rel_l2(out, T(golden, name + "/out")) < 2e-6
    w64 = odec.cast_weights(weights, torch.float64)
    out64 = odec.decoder_forward(w64, T(golden, name + "/x").double(), T(golden, name + "/mask").double(),
                                 T(golden, name + "/mu").double(), T(golden, name + "/t").double())
    assert rel_l2(out64.float(), T(golden, name + "/out")) < 5e-6   # the fp32 reference's own rounding noise


@pytest.mark.parametrize("name", ["solve_T32_n2", "solve_T40_n4"])
def test_euler_solve_matches_reference(golden, weights, name):
    torch.set_num_threads(4)
    out = odec.solve_euler(weights, T(golden, name + "/z"), T(golden, name + "/t_span"), T(golden, name + "/mu"),
                           T(golden, name + "/mask"))
    ref = T(golden, name + "/out")
    assert rel_l2(out, ref) < 2e-6
    # padded frames keep the initial noise: v is masked to 0 there
    pad = T(golden, name + "/mask")[:, 0, :] == 0
    assert torch.equal(ref.transpose(1, 2)[pad], T(golden, name + "/z").transpose(1, 2)[pad])


def test_compute_loss_matches_reference(golden, weights):
    loss, phi = odec.compute_loss(weights, T(golden, "loss_T32/x1"), T(golden, "loss_T32/mask"), T(golden, "loss_T32/mu"),
                                  T(golden, "loss_T32/t"), T(golden, "loss_T32/z"))
    assert torch.equal(phi, T(golden, "loss_T32/phi_t"))
    assert abs(float(loss) - float(golden["loss_T32/loss"])) < 1e-5 * abs(float(golden["loss_T32/loss"]))


def test_euler_time_grid_is_the_reference_loop():
    t_span = torch.linspace(0, 1, 11)
    ts, dts = odec.euler_time_grid(t_span)
    assert ts.dtype == torch.float32 and len(ts) == 10
    t = t_span[0]; dt = t_span[1] - t_span[0]
    for k in range(10):
        assert ts[k] == t and dts[k] == dt
        t = t + dt
        if k < 9:
            dt = t_span[k + 2] - t


def test_sdpa_form_equals_explicit_softmax(weights):
    """oracle.decoder.USE_SDPA (the form diffusers' AttnProcessor2_0 calls, used by bench.py's same-box GPU baseline)
    is the same arithmetic as the explicit softmax the golden fixtures pin."""
    g = torch.Generator().manual_seed(3)
    x, mu = torch.randn(2, 80, 40, generator=g), torch.randn(2, 80, 40, generator=g)
    mask = torch.ones(2, 1, 40); mask[1, :, 29:] = 0
    t = torch.tensor(0.3)
    a = odec.decoder_forward(weights, x, mask, mu, t)
    odec.USE_SDPA = True
    try:
        b = odec.decoder_forward(weights, x, mask, mu, t)
    finally:
        odec.USE_SDPA = False
    assert rel_l2(b, a) < 2e-6


def test_t800_golden_matches_reference():
    """The benchmarked configuration's size (T_mel = 800, 10 Euler steps): the restatement against the output of the
    reference's own modules (tests/golden/decoder_golden_T800.npz, make_decoder_golden.py --t800)."""
    gold = np.load(os.path.join(ROOT, "tests", "golden", "decoder_golden_T800.npz"))
    w = odec.make_weights(int(gold["weight_seed"][0]), "rich")
    z, mu, mask, t_span = t800_inputs(gold)
    torch.set_num_threads(os.cpu_count() or 4)
    out = odec.solve_euler(w, z, t_span, mu, mask)
    assert rel_l2(out, torch.from_numpy(gold["out"])) < 2e-6


def t800_inputs(gold):
    g = torch.Generator().manual_seed(int(gold["seed"][0]))
    B, T = len(gold["lengths"]), 800
    mu, z = torch.randn(B, 80, T, generator=g), torch.randn(B, 80, T, generator=g)
    mask = torch.zeros(B, 1, T)
    for i, n in enumerate(gold["lengths"]):
        mask[i, :, : int(n)] = 1
    return z, mu, mask, torch.linspace(0, 1, int(gold["n_timesteps"][0]) + 1)
