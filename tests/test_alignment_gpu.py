"""GPU parity tests for csrc/align.cu and the durations side output of the MAS kernel, through
matcha_b200.alignment / flow_matching (C ABI underneath): against tensors recorded from the reference's
own MatchaTTS.forward / synthesise (tests/golden/alignment_golden.npz) and against oracle/alignment.py
on seeded inputs.  Bars: bit-exact for paths, durations, mu_y, masks, lengths and phi_t; the K=80
contraction of the log-prior and the two loss sums differ from torch only by summation order and are
held to the tolerances written below."""
import os
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from oracle import alignment as oal
from oracle import decoder as odec

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LOG_PRIOR_ATOL = 1e-3      # values are O(-100 .. -1000): a few fp32 ulps of the K=80 sums
LOSS_RTOL = 2e-6           # fp64 accumulation here vs torch's fp32 pairwise sums


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(ROOT, "tests", "golden", "alignment_golden.npz"))


def G(g, key):
    return torch.from_numpy(g[key]).cuda()


def ragged(B, T_x, T_y, seed):
    rng = np.random.RandomState(seed)
    tx = np.maximum(1, (T_x * rng.uniform(0.7, 1.0, B)).astype(np.int64))
    ty = np.maximum(tx, (T_y * rng.uniform(0.7, 1.0, B)).astype(np.int64))
    tx[0], ty[0] = T_x, T_y
    return torch.from_numpy(tx), torch.from_numpy(ty)


@pytest.mark.parametrize("case", ["train_a", "train_b"])
def test_log_prior_matches_reference_golden(gold, case):
    from matcha_b200 import alignment as al
    got = al.log_prior(G(gold, f"{case}/mu_x"), G(gold, f"{case}/y"))
    want = G(gold, f"{case}/log_prior")
    assert got.shape == want.shape
    assert float((got - want).abs().max()) < LOG_PRIOR_ATOL
    assert float(torch.linalg.norm((got - want).double()) / torch.linalg.norm(want.double())) < 1e-6


@pytest.mark.parametrize("B,C,T_x,T_y", [(2, 80, 64, 128), (3, 80, 65, 131), (1, 80, 1, 1), (2, 7, 130, 515), (4, 80, 128, 800)])
def test_log_prior_vs_oracle_shapes(B, C, T_x, T_y):
    from matcha_b200 import alignment as al
    g = torch.Generator().manual_seed(B * 100 + T_x)
    mu_x, y = torch.randn(B, C, T_x, generator=g), torch.randn(B, C, T_y, generator=g) * 1.5
    want = oal.log_prior(mu_x, y)
    got = al.log_prior(mu_x.cuda(), y.cuda()).cpu()
    assert float((got - want).abs().max()) < LOG_PRIOR_ATOL
    # with lengths: the valid block is written, and is the same
    tx, ty = ragged(B, T_x, T_y, 5)
    got2 = al.log_prior(mu_x.cuda(), y.cuda(), tx, ty).cpu()
    for b in range(B):
        assert torch.equal(got2[b, : tx[b], : ty[b]], got[b, : tx[b], : ty[b]])


@pytest.mark.parametrize("case", ["train_a", "train_b"])
def test_align_matches_reference_golden(gold, case):
    from matcha_b200 import alignment as al
    mu_x, y = G(gold, f"{case}/mu_x"), G(gold, f"{case}/y")
    xl, yl = G(gold, f"{case}/x_lengths"), G(gold, f"{case}/y_lengths")
    r = al.align(mu_x, y, xl, yl)
    # MAS on the reference's own log-prior: bit-exact path and durations
    path, dur = al.maximum_path_with_durations(G(gold, f"{case}/log_prior"), xl, yl)
    assert torch.equal(path, G(gold, f"{case}/attn"))
    assert torch.equal(dur, G(gold, f"{case}/attn").sum(-1).to(torch.int32))
    # the whole seam from mu_x, y (log-prior differs from torch's matmul by ulps; no near-ties in this data)
    assert torch.equal(r["attn"], G(gold, f"{case}/attn"))
    assert torch.equal(r["mu_y"], G(gold, f"{case}/mu_y"))
    torch.testing.assert_close(r["logw_"], G(gold, f"{case}/logw_"), rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(r["prior_loss"], G(gold, f"{case}/prior_loss"), rtol=LOSS_RTOL, atol=0)


@pytest.mark.parametrize("B,T_x,T_y,ties", [(5, 40, 90, False), (4, 128, 800, False), (6, 33, 70, True), (3, 200, 1000, False)])
def test_align_vs_oracle_on_its_own_log_prior(B, T_x, T_y, ties):
    from matcha_b200 import alignment as al
    g = torch.Generator().manual_seed(T_x * 7 + T_y)
    mu_x, y = torch.randn(B, 80, T_x, generator=g), torch.randn(B, 80, T_y, generator=g)
    if ties:                                   # quantised inputs: many exactly equal scores
        mu_x, y = torch.round(mu_x), torch.round(y)
    tx, ty = ragged(B, T_x, T_y, T_x)
    r = al.align(mu_x.cuda(), y.cuda(), tx, ty)
    lp = al.log_prior(mu_x.cuda(), y.cuda()).cpu()
    x_mask = oal.sequence_mask(tx, T_x).float()
    y_mask = oal.sequence_mask(ty, T_y).float()
    attn = oal.maximum_path(lp, x_mask.unsqueeze(-1) * y_mask.unsqueeze(1))
    assert torch.equal(r["attn"].cpu(), attn)
    assert torch.equal(r["durations"].cpu(), attn.sum(-1).to(torch.int32))
    mu_y = torch.matmul(attn.transpose(1, 2), mu_x.transpose(1, 2)).transpose(1, 2)
    assert torch.equal(r["mu_y"].cpu(), mu_y)
    logw_ = torch.log(1e-8 + attn.sum(-1)).unsqueeze(1) * x_mask.unsqueeze(1)
    torch.testing.assert_close(r["logw_"].cpu(), logw_, rtol=1e-6, atol=1e-6)
    prior = torch.sum(0.5 * ((y - mu_y) ** 2 + np.log(2 * np.pi)) * y_mask.unsqueeze(1)) / (y_mask.sum() * 80)
    torch.testing.assert_close(r["prior_loss"].cpu(), prior, rtol=LOSS_RTOL, atol=0)


def test_durations_of_out_of_domain_items_are_zero():
    from matcha_b200 import alignment as al
    v = torch.randn(3, 16, 40).cuda()
    tx, ty = torch.tensor([16, 20, 0]), torch.tensor([40, 10, 40])      # item 1: t_x > t_y, item 2: t_x = 0
    path, dur = al.maximum_path_with_durations(v, tx, ty)
    assert int(dur[0].sum()) == 40 and int(path[0].sum()) == 40
    assert int(dur[1:].abs().sum()) == 0 and float(path[1:].abs().sum()) == 0


@pytest.mark.parametrize("case", ["synth_a", "synth_b"])
def test_synthesis_prestep_matches_reference_golden(gold, case):
    from matcha_b200 import alignment as al
    mu_x, logw, x_mask = (G(gold, f"{case}/{k}") for k in ("mu_x", "logw", "x_mask"))
    r = al.synthesis_prestep(mu_x, logw, x_mask, float(gold[f"{case}/length_scale"]))
    assert torch.equal(r["y_lengths"], G(gold, f"{case}/mel_lengths"))
    assert torch.equal(r["y_mask"], G(gold, f"{case}/y_mask"))
    assert torch.equal(r["attn"].squeeze(1), G(gold, f"{case}/gp_path"))
    assert torch.equal(r["mu_y"], G(gold, f"{case}/mu_y"))
    got = al.generate_path(G(gold, f"{case}/gp_duration"), G(gold, f"{case}/gp_mask"))
    assert torch.equal(got, G(gold, f"{case}/gp_path"))


@pytest.mark.parametrize("B,T_x,T_y", [(64, 400, 2000), (3, 1, 4), (2, 300, 257)])
def test_expand_properties_at_size(B, T_x, T_y):
    """Size-independent properties at BASELINE's largest shape: every valid frame has exactly one
    token, row sums give the durations back, mu_y is the gather."""
    from matcha_b200 import alignment as al
    g = torch.Generator().manual_seed(B + T_x)
    tx, _ = ragged(B, T_x, T_y, 9)
    dur = torch.zeros(B, T_x, dtype=torch.int32)
    ty = torch.zeros(B, dtype=torch.int64)
    for b in range(B):
        d = torch.randint(0, max(2, 2 * T_y // T_x), (int(tx[b]),), generator=g)
        while int(d.sum()) > T_y:
            d = d // 2
        dur[b, : tx[b]] = d
        ty[b] = int(d.sum())
    mu_x = torch.randn(B, 80, T_x, generator=g)
    r = al.expand(dur.cuda(), mu_x.cuda(), tx, ty, T_y=T_y, want_attn=True, want_logw=True)
    attn = r["attn"]
    assert torch.equal(attn.sum(-1).to(torch.int32).cpu(), dur)
    y_mask = (torch.arange(T_y).unsqueeze(0) < ty.unsqueeze(1)).float()
    assert torch.equal(attn.sum(1).cpu(), y_mask)
    idx = attn.argmax(1)                                                   # [B,T_y]
    want = torch.gather(mu_x.cuda(), 2, idx.unsqueeze(1).expand(B, 80, T_y)) * y_mask.cuda().unsqueeze(1)
    assert torch.equal(r["mu_y"], want)
    assert torch.equal(al.generate_path(dur.float().cuda(), (oal.sequence_mask(tx, T_x).float().unsqueeze(-1)
                                                              * y_mask.unsqueeze(1)).cuda()), attn)


def test_cfm_loss_kernels_vs_oracle():
    from matcha_b200 import flow_matching as fm
    g = torch.Generator().manual_seed(4)
    B, T = 3, 50
    x1, z = torch.randn(B, 80, T, generator=g), torch.randn(B, 80, T, generator=g)
    mu = torch.randn(B, 80, T, generator=g)
    t = torch.rand(B, 1, 1, generator=g)
    mask = torch.zeros(B, 1, T); mask[0] = 1; mask[1, :, :31] = 1; mask[2, :, :8] = 1
    w = odec.make_weights(11, "rich")
    cfm = fm.CFM(in_channels=160, out_channel=80, cfm_params=SimpleNamespace(solver="euler", sigma_min=1e-4),
                 decoder_params=dict(channels=(256, 256), dropout=0.05, attention_head_dim=64, n_blocks=1,
                                     num_mid_blocks=2, num_heads=4))
    cfm.load_state_dict({"estimator." + k: v for k, v in w.items()})
    loss, phi_t = cfm.compute_loss(x1.cuda(), mask.cuda(), mu.cuda(), t=t.cuda(), z=z.cuda())
    phi_want, u_want = oal.cfm_mix(x1, z, t)
    assert torch.equal(phi_t.cpu(), phi_want)                               # flow_matching.py:138, bit-exact
    u_pred = cfm.estimator(phi_t, mask.cuda(), mu.cuda(), t.reshape(B).cuda()).cpu()
    want = oal.cfm_mse(u_pred, u_want, mask)
    torch.testing.assert_close(loss.cpu(), want, rtol=LOSS_RTOL, atol=0)
    torch.set_num_threads(8)
    full = odec.compute_loss(odec.cast_weights(w, torch.float64), x1.double(), mask.double(), mu.double(),
                             t.double(), z.double())
    full = full[0] if isinstance(full, (tuple, list)) else full
    assert abs(float(loss) - float(full)) < 1e-5 * abs(float(full))


@pytest.mark.parametrize("precision", ["fp16", "bf16"])
@pytest.mark.parametrize("B,T,lengths", [(3, 50, [50, 31, 8]), (9, 131, [131, 120, 99, 80, 64, 50, 33, 9, 2]), (1, 64, [40])])
def test_cfm_loss_fused_into_decoder_eval(precision, B, T, lengths):
    """in the tensor-core modes compute_loss is ONE call -- phi_t formed by the decoder's input-packing kernel
    (bit-exact against flow_matching.py:138), the squared error summed in the final projection's epilogue.  Checked against
    the same sums formed from the decoder's own stored prediction (tight: only the summation order differs) and against the
    fp64 oracle's loss (the mode's tolerance).  B = 9 runs the sub-batch streams (uneven split)."""
    from matcha_b200 import flow_matching as fm
    g = torch.Generator().manual_seed(B * 100 + T)
    x1, z = torch.randn(B, 80, T, generator=g), torch.randn(B, 80, T, generator=g)
    mu = torch.randn(B, 80, T, generator=g)
    t = torch.rand(B, 1, 1, generator=g)
    mask = torch.zeros(B, 1, T)
    for i, n in enumerate(lengths):
        mask[i, :, :n] = 1
    w = odec.make_weights(11, "rich")
    cfm = fm.CFM(in_channels=160, out_channel=80, cfm_params=SimpleNamespace(solver="euler", sigma_min=1e-4),
                 decoder_params=dict(channels=(256, 256), dropout=0.05, attention_head_dim=64, n_blocks=1,
                                     num_mid_blocks=2, num_heads=4), precision=precision)
    cfm.load_state_dict({"estimator." + k: v for k, v in w.items()})
    loss, phi_t = cfm.compute_loss(x1.cuda(), mask.cuda(), mu.cuda(), t=t.cuda(), z=z.cuda())
    phi_want, u_want = oal.cfm_mix(x1, z, t)
    assert torch.equal(phi_t.cpu(), phi_want)
    # the same call with the prediction written out: identical sums up to atomics order; the stored prediction reproduces them
    sums, phi2, pred = cfm.estimator.loss_forward(x1.cuda(), z.cuda(), t.reshape(B).cuda(), mu.cuda(), mask.cuda(), 1e-4, want_pred=True)
    assert torch.equal(phi2.cpu(), phi_want)
    assert float(sums[1]) == float(mask.sum())
    want = oal.cfm_mse(pred.cpu(), u_want, mask)
    torch.testing.assert_close(loss.cpu(), want, rtol=LOSS_RTOL, atol=0)
    torch.testing.assert_close((sums[0] / (sums[1] * 80)).float().cpu(), want, rtol=LOSS_RTOL, atol=0)
    # and the separate forward on phi_t gives the same prediction (same kernels, same inputs)
    u_pred = cfm.estimator(phi_t, mask.cuda(), mu.cuda(), t.reshape(B).cuda())
    assert float((u_pred - pred).abs().max()) <= 1e-6 * float(pred.abs().max()) + 1e-6
    torch.set_num_threads(8)
    full = odec.compute_loss(odec.cast_weights(w, torch.float64), x1.double(), mask.double(), mu.double(), t.double(), z.double())
    full = full[0] if isinstance(full, (tuple, list)) else full
    tol = {"fp16": 2e-3, "bf16": 1.5e-2}[precision] * 2       # loss ~ |u_pred - u|^2: twice the relative error of the prediction
    assert abs(float(loss) - float(full)) < tol * abs(float(full))
