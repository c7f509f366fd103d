"""GPU parity tests for the CFM decoder path (fp32 mode): against golden outputs of the reference's own
modules (tests/golden/decoder_golden.npz) and against the torch oracle (oracle/decoder.py) in fp64 on
seeded inputs.  Tolerance from BASELINE.json north_star: relative L2 <= 1e-5 in fp32 mode."""
import os
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from oracle import decoder as odec

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL_F32 = 1e-5


def rel_l2(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float(torch.linalg.norm(a - b) / torch.linalg.norm(b))


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "decoder_golden.npz"))


@pytest.fixture(scope="module")
def weights(golden):
    return odec.make_weights(int(golden["weight_seed"][0]), "rich")


@pytest.fixture(scope="module")
def cfm(weights):
    from matcha_b200.flow_matching import CFM
    m = CFM(in_channels=160, out_channel=80, cfm_params=SimpleNamespace(solver="euler", sigma_min=1e-4),
            decoder_params=dict(channels=(256, 256), dropout=0.05, attention_head_dim=64, n_blocks=1,
                                num_mid_blocks=2, num_heads=4))
    m.load_state_dict({"estimator." + k: v for k, v in weights.items()})
    return m


def G(golden, key):
    return torch.from_numpy(golden[key]).cuda()


@pytest.mark.parametrize("name", ["eval_T37_tvec", "eval_T44_full"])
def test_decoder_forward_matches_reference_golden(golden, cfm, name):
    out = cfm.estimator(G(golden, name + "/x"), G(golden, name + "/mask"), G(golden, name + "/mu"), G(golden, name + "/t"))
    assert out.shape == golden[name + "/out"].shape
    assert rel_l2(out, G(golden, name + "/out")) < TOL_F32


@pytest.mark.parametrize("name", ["solve_T32_n2", "solve_T40_n4"])
def test_solve_euler_matches_reference_golden(golden, cfm, name):
    z, mask = G(golden, name + "/z"), G(golden, name + "/mask")
    out = cfm.solve_ode_euler(z, G(golden, name + "/t_span"), G(golden, name + "/mu"), mask, None, None)
    assert rel_l2(out, G(golden, name + "/out")) < TOL_F32
    pad = (mask[:, 0, :] == 0)
    assert torch.equal(out.transpose(1, 2)[pad], z.transpose(1, 2)[pad]), "padded frames keep the initial noise"


def test_compute_loss_matches_reference_golden(golden, cfm):
    torch.manual_seed(0)
    x1, mask, mu = G(golden, "loss_T32/x1"), G(golden, "loss_T32/mask"), G(golden, "loss_T32/mu")
    t, z = G(golden, "loss_T32/t"), G(golden, "loss_T32/z")
    phi_t = (1 - (1 - 1e-4) * t) * z + t * x1
    u_pred = cfm.estimator(phi_t, mask, mu, t.squeeze())
    u_target = x1 - (1 - 1e-4) * z
    loss = torch.nn.functional.mse_loss(u_pred, u_target, reduction="sum") / (mask.sum() * 80)
    assert abs(float(loss) - float(golden["loss_T32/loss"])) < 1e-5 * abs(float(golden["loss_T32/loss"]))
    l2, phi2 = cfm.compute_loss(x1, mask, mu)          # API shape check (own RNG draws)
    assert l2.ndim == 0 and phi2.shape == x1.shape and torch.isfinite(l2)


@pytest.mark.parametrize("B,T,lengths,tvec", [(2, 64, [64, 50], False), (3, 101, [101, 77, 33], True), (1, 130, [130], False),
                                             (2, 2, [2, 1], False), (4, 200, [200, 180, 120, 64], True)])
def test_decoder_forward_vs_fp64_oracle(weights, cfm, B, T, lengths, tvec):
    g = torch.Generator().manual_seed(B * 1000 + T)
    x, mu = torch.randn(B, 80, T, generator=g), torch.randn(B, 80, T, generator=g)
    mask = torch.zeros(B, 1, T)
    for i, n in enumerate(lengths):
        mask[i, :, :n] = 1
    t = torch.rand(B, generator=g) if tvec else torch.rand((), generator=g)
    w64 = odec.cast_weights(weights, torch.float64)
    torch.set_num_threads(8)
    want = odec.decoder_forward(w64, x.double(), mask.double(), mu.double(), t.double())
    got = cfm.estimator(x.cuda(), mask.cuda(), mu.cuda(), t.cuda())
    assert rel_l2(got, want) < TOL_F32


def test_solve_ten_steps_vs_fp64_oracle(weights, cfm):
    g = torch.Generator().manual_seed(77)
    B, T = 2, 96
    z, mu = torch.randn(B, 80, T, generator=g), torch.randn(B, 80, T, generator=g)
    mask = torch.zeros(B, 1, T); mask[0, :, :96] = 1; mask[1, :, :61] = 1
    t_span = torch.linspace(0, 1, 11)
    w64 = odec.cast_weights(weights, torch.float64)
    torch.set_num_threads(8)
    want = odec.solve_euler(w64, z.double(), t_span, mu.double(), mask.double())
    got = cfm.solve_ode_euler(z.cuda(), t_span.cuda(), mu.cuda(), mask.cuda(), None, None)
    assert rel_l2(got, want) < TOL_F32


def test_cfm_forward_api_shapes_and_determinism(cfm):
    mu = torch.randn(2, 80, 48, device="cuda")
    mask = torch.ones(2, 1, 48, device="cuda")
    torch.manual_seed(3); a = cfm(mu, mask, n_timesteps=2, temperature=0.667)
    torch.manual_seed(3); b = cfm(mu, mask, n_timesteps=2, temperature=0.667)
    assert a.shape == mu.shape and torch.equal(a, b) and torch.isfinite(a).all()


def test_state_dict_ingest_rejects_wrong_shapes(weights):
    from matcha_b200.decoder import Decoder
    bad = dict(weights); bad["final_proj.weight"] = torch.zeros(80, 128, 1)
    with pytest.raises(KeyError):
        Decoder().load_state_dict(bad)


# ------------------------------------------------------------------ 16-bit tensor-core modes
# BASELINE.json north_star: relative L2 <= 2e-3 against the fp32 reference for the tensor-core mode.
# "fp16" (IEEE half operands, fp32 accumulate/residual/statistics) meets it; plain "bf16" operands
# measure 5e-3..9e-3 (8 mantissa bits per operand, ~1.6e-3 per GEMM, ~50 GEMMs deep) and are held
# to a FUNCTIONAL bound only (1.5e-2: catches a broken kernel, not the north_star bar) -- see DESIGN.md §Decoder
# precision; the 2e-3 assertion for bf16 lives in tests/test_decoder_t800_gpu.py as xfail(strict=True), so the suite
# says so the day a bf16 path meets it.
TOL_16 = {"fp16": 2e-3, "bf16": 1.5e-2}


@pytest.fixture(scope="module", params=["fp16", "bf16"])
def cfm16(request, weights):
    from matcha_b200.flow_matching import CFM
    m = CFM(in_channels=160, out_channel=80, cfm_params=SimpleNamespace(solver="euler", sigma_min=1e-4),
            decoder_params=dict(channels=(256, 256), dropout=0.05, attention_head_dim=64, n_blocks=1,
                                num_mid_blocks=2, num_heads=4), precision=request.param)
    m.load_state_dict({"estimator." + k: v for k, v in weights.items()})
    return m


@pytest.mark.parametrize("name", ["eval_T37_tvec", "eval_T44_full"])
def test_half_decoder_forward_matches_reference_golden(golden, cfm16, name):
    out = cfm16.estimator(G(golden, name + "/x"), G(golden, name + "/mask"), G(golden, name + "/mu"), G(golden, name + "/t"))
    err = rel_l2(out, G(golden, name + "/out"))
    print(name, cfm16.estimator.precision, "rel-L2", err)
    assert err < TOL_16[cfm16.estimator.precision]


@pytest.mark.parametrize("name", ["solve_T32_n2", "solve_T40_n4"])
def test_half_solve_euler_matches_reference_golden(golden, cfm16, name):
    z, mask = G(golden, name + "/z"), G(golden, name + "/mask")
    out = cfm16.solve_ode_euler(z, G(golden, name + "/t_span"), G(golden, name + "/mu"), mask, None, None)
    err = rel_l2(out, G(golden, name + "/out"))
    print(name, cfm16.estimator.precision, "rel-L2", err)
    assert err < TOL_16[cfm16.estimator.precision]
    pad = (mask[:, 0, :] == 0)
    assert torch.equal(out.transpose(1, 2)[pad], z.transpose(1, 2)[pad])


def test_half_solve_ten_steps_vs_fp64_oracle(weights, cfm16):
    g = torch.Generator().manual_seed(77)
    B, T = 2, 96
    z, mu = torch.randn(B, 80, T, generator=g), torch.randn(B, 80, T, generator=g)
    mask = torch.zeros(B, 1, T); mask[0, :, :96] = 1; mask[1, :, :61] = 1
    t_span = torch.linspace(0, 1, 11)
    w64 = odec.cast_weights(weights, torch.float64)
    torch.set_num_threads(8)
    want = odec.solve_euler(w64, z.double(), t_span, mu.double(), mask.double())
    got = cfm16.solve_ode_euler(z.cuda(), t_span.cuda(), mu.cuda(), mask.cuda(), None, None)
    err = rel_l2(got, want)
    print("10-step", cfm16.estimator.precision, "rel-L2", err)
    assert err < TOL_16[cfm16.estimator.precision]
