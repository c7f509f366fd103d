"""GPU tests of the assembled flows (matcha_b200.matcha_tts.MatchaTTS): synthesise and the training-step forward values
against the torch restatements composed the way matcha_tts.py composes the reference modules."""
import math

import pytest
import torch

from oracle import alignment as oal
from oracle import decoder as odec
from oracle import text_encoder as oenc

pytestmark = pytest.mark.gpu
N_VOCAB = 178


def rel_l2(a, b):
    return float(torch.linalg.norm((a.double().cpu() - b.double().cpu())) / torch.linalg.norm(b.double().cpu()))


@pytest.fixture(scope="module")
def weights():
    return oenc.make_weights(41, "rich", N_VOCAB), odec.make_weights(42, "rich")


def build(weights, precision):
    from matcha_b200.matcha_tts import MatchaTTS
    we, wd = weights
    sd = {"encoder." + k: v for k, v in we.items()}
    sd.update({"decoder.estimator." + k: v for k, v in wd.items()})
    return MatchaTTS(N_VOCAB, precision=precision).load_state_dict(sd)


@pytest.mark.parametrize("precision,tol", [("fp32", 2e-5), ("fp16", 2e-3)])
def test_synthesise_matches_oracle_composition(weights, precision, tol):
    """matcha_tts.py:212-249.  The durations are ceil(exp(logw)): a 1e-6 (fp32) or 1e-3 (fp16 mode) difference in logw may
    move one across an integer, so the encoder is compared through mu_x / logw tolerances in tests/test_text_encoder_gpu.py
    and, here, the steps AFTER it are checked from the model's own encoder outputs; in fp32 mode the lengths and mu_y must
    also equal the all-oracle composition.  The CUDA path draws z with torch.randn_like on the device, so the oracle re-draws
    it from the same seed."""
    we, wd = weights
    model = build(weights, precision)
    g = torch.Generator().manual_seed(1)
    B, T = 3, 24
    x = torch.randint(0, N_VOCAB, (B, T), generator=g)
    x_lengths = torch.tensor([24, 17, 9])
    torch.manual_seed(77)
    out = model.synthesise(x.cuda(), x_lengths.cuda(), n_timesteps=3, temperature=0.667, length_scale=1.1)
    # the model's own encoder outputs, then the oracle's pre-step (matcha_tts.py:214-229) on them
    mu_x, logw, x_mask = (t.cpu() for t in model.encoder(x.cuda(), x_lengths.cuda()))
    w_ceil = torch.ceil(torch.exp(logw) * x_mask) * 1.1
    y_lengths = torch.clamp_min(torch.sum(w_ceil, [1, 2]), 1).long()
    y_max = int(y_lengths.max())
    T_y = oal.fix_len_compatibility(y_max)
    y_mask = oal.sequence_mask(y_lengths, T_y).unsqueeze(1).float()
    attn = oal.generate_path(w_ceil.squeeze(1), (x_mask.unsqueeze(-1) * y_mask.unsqueeze(2)).squeeze(1))
    mu_y = torch.matmul(attn.transpose(1, 2), mu_x.transpose(1, 2)).transpose(1, 2)
    assert torch.equal(out["mel_lengths"].cpu(), y_lengths)
    assert out["encoder_outputs"].shape == (B, 80, y_max)
    assert torch.equal(out["encoder_outputs"].cpu(), mu_y[:, :, :y_max])          # a gather: exact
    omu_x, ologw, _ = oenc.text_encoder_forward(we, x, x_lengths)
    assert rel_l2(mu_x, omu_x) < (1e-5 if precision == "fp32" else 2e-3)
    assert rel_l2(logw, ologw) < (1e-5 if precision == "fp32" else 2e-3)
    if precision == "fp32":                                                       # all-oracle composition gives the same lengths
        ow = torch.ceil(torch.exp(ologw) * x_mask) * 1.1
        assert torch.equal(torch.clamp_min(torch.sum(ow, [1, 2]), 1).long(), y_lengths)
    torch.manual_seed(77)
    z = (torch.randn(B, 80, T_y, device="cuda") * 0.667).cpu()
    want = odec.solve_euler(wd, z, torch.linspace(0, 1, 4), mu_y, y_mask)[:, :, :y_max]
    assert rel_l2(out["decoder_outputs"], want) < tol
    assert torch.equal(out["mel"], out["decoder_outputs"])               # mel_mean 0, mel_std 1
    assert out["attn"].shape[0] == B and out["attn"].shape[1] == 1


def test_training_forward_values_match_oracle_composition(weights):
    """matcha_tts.py:271-322 (no out_size): encoder -> log-prior -> MAS -> logw_ / mu_y -> the three losses."""
    we, wd = weights
    model = build(weights, "fp32")
    g = torch.Generator().manual_seed(2)
    B, T_x, T_y = 3, 20, 96
    x = torch.randint(0, N_VOCAB, (B, T_x), generator=g)
    x_lengths, y_lengths = torch.tensor([20, 15, 11]), torch.tensor([96, 80, 61])
    y = torch.randn(B, 80, T_y, generator=g)
    t, z = torch.rand(B, 1, 1, generator=g), torch.randn(B, 80, T_y, generator=g)
    dur, prior, diff, attn = model(x.cuda(), x_lengths.cuda(), y.cuda(), y_lengths.cuda(), t=t.cuda(), z=z.cuda())
    mu_x, logw, x_mask = oenc.text_encoder_forward(we, x, x_lengths)
    y_mask = oal.sequence_mask(y_lengths, T_y).unsqueeze(1).float()
    lp = oal.log_prior(mu_x, y)
    oattn = oal.maximum_path(lp, (x_mask.unsqueeze(-1) * y_mask.unsqueeze(2)).squeeze(1))
    assert torch.equal(attn.cpu(), oattn)                                # same path although mu_x differs by ~1e-6
    logw_ = torch.log(1e-8 + torch.sum(oattn.unsqueeze(1), -1)) * x_mask
    odur = torch.sum((logw - logw_) ** 2) / torch.sum(x_lengths)
    mu_y = torch.matmul(oattn.transpose(1, 2), mu_x.transpose(1, 2)).transpose(1, 2)
    oprior = torch.sum(0.5 * ((y - mu_y) ** 2 + math.log(2 * math.pi)) * y_mask) / (torch.sum(y_mask) * 80)
    odiff, _ = odec.compute_loss(wd, y, y_mask, mu_y, t, z)
    assert abs(float(dur) - float(odur)) < 1e-5 * abs(float(odur))
    assert abs(float(prior) - float(oprior)) < 1e-5 * abs(float(oprior))
    assert abs(float(diff) - float(odiff)) < 2e-5 * abs(float(odiff))


@pytest.mark.parametrize("data", ["aligned", "random"])
def test_fp16_training_forward_against_oracle_composition(weights, data):
    """The benchmarked training step runs the fp16 tensor-core encoder and decoder (bench.py training_step); this is its
    end-to-end check against the fp32 oracle composition of matcha_tts.py:271-322.  mu_x moves by ~3e-4 in fp16 mode, so MAS
    may legitimately pick a different token for some frames; the test reports that fraction and bounds it, and bounds the three
    losses.  "aligned": y is mu_x stretched along a random monotone alignment plus noise (what training data looks like once
    the encoder has learnt something); "random": y is pure noise, the flattest score landscape MAS can be given."""
    we, wd = weights
    model = build(weights, "fp16")
    g = torch.Generator().manual_seed(7)
    B, T_x, T_y = 6, 48, 320
    x = torch.randint(0, N_VOCAB, (B, T_x), generator=g)
    x_lengths = torch.tensor([48, 44, 40, 33, 25, 12])
    y_lengths = torch.tensor([320, 300, 251, 200, 163, 90])
    mu_x, logw, x_mask = oenc.text_encoder_forward(we, x, x_lengths)
    y = torch.randn(B, 80, T_y, generator=g)
    if data == "aligned":
        for b in range(B):
            tx, ty = int(x_lengths[b]), int(y_lengths[b])
            cuts = torch.sort(torch.randperm(ty - 1, generator=g)[:tx - 1] + 1).values
            tok = torch.bucketize(torch.arange(ty), cuts, right=True)           # frame -> token, monotone, every token used
            y[b, :, :ty] = mu_x[b][:, tok] + 0.3 * y[b, :, :ty]
    t, z = torch.rand(B, 1, 1, generator=g), torch.randn(B, 80, T_y, generator=g)
    dur, prior, diff, attn = model(x.cuda(), x_lengths.cuda(), y.cuda(), y_lengths.cuda(), t=t.cuda(), z=z.cuda())
    y_mask = oal.sequence_mask(y_lengths, T_y).unsqueeze(1).float()
    oattn = oal.maximum_path(oal.log_prior(mu_x, y), (x_mask.unsqueeze(-1) * y_mask.unsqueeze(2)).squeeze(1))
    logw_ = torch.log(1e-8 + torch.sum(oattn.unsqueeze(1), -1)) * x_mask
    odur = torch.sum((logw - logw_) ** 2) / torch.sum(x_lengths)
    mu_y = torch.matmul(oattn.transpose(1, 2), mu_x.transpose(1, 2)).transpose(1, 2)
    oprior = torch.sum(0.5 * ((y - mu_y) ** 2 + math.log(2 * math.pi)) * y_mask) / (torch.sum(y_mask) * 80)
    odiff, _ = odec.compute_loss(wd, y, y_mask, mu_y, t, z)
    got = attn.cpu()
    assert torch.equal(got.sum(1), y_mask.squeeze(1))                           # still a valid alignment: one token per valid frame
    frames_moved = float(((got.argmax(1) != oattn.argmax(1)) & (y_mask.squeeze(1) > 0)).sum() / y_mask.sum())
    rel = [abs(float(a) - float(b)) / abs(float(b)) for a, b in ((dur, odur), (prior, oprior), (diff, odiff))]
    print(f"fp16 training step [{data}]: frames with another token {frames_moved:.4f}; rel. error dur {rel[0]:.2e} prior {rel[1]:.2e} diff {rel[2]:.2e}")
    assert frames_moved <= FP16_STEP_TOL[data]["frames"]
    assert rel[0] <= FP16_STEP_TOL[data]["dur"] and rel[1] <= FP16_STEP_TOL[data]["prior"] and rel[2] <= FP16_STEP_TOL[data]["diff"]


# tolerances of the fp16 training step against the fp32 oracle composition.  Measured on B200 (profiles/r3_fp16_training_step.log):
# aligned: 0 frames moved, dur 2.6e-5, prior 6e-8, diff 4.7e-5;  random: 0.23 % of frames moved, dur 2.2e-3, prior 2.1e-5, diff 6.0e-5
FP16_STEP_TOL = {"aligned": {"frames": 0.005, "dur": 1e-3, "prior": 1e-4, "diff": 1e-3},
                 "random": {"frames": 0.02, "dur": 2e-2, "prior": 5e-4, "diff": 1e-3}}


def test_training_forward_with_out_size_cut(weights):
    """matcha_tts.py:290-312: the random segment cut.  The offsets come from Python's `random` in the reference's order, so the
    oracle composition re-seeds it and draws the same ones."""
    import random
    we, wd = weights
    model = build(weights, "fp32")
    g = torch.Generator().manual_seed(3)
    B, T_x, T_y, out_size = 3, 20, 96, 32                               # out_size divisible by 4 (two U-Net downsamplings)
    x = torch.randint(0, N_VOCAB, (B, T_x), generator=g)
    x_lengths, y_lengths = torch.tensor([20, 15, 11]), torch.tensor([96, 70, 28])      # the last one is shorter than out_size
    y = torch.randn(B, 80, T_y, generator=g)
    t, z = torch.rand(B, 1, 1, generator=g), torch.randn(B, 80, out_size, generator=g)
    random.seed(5)
    dur, prior, diff, attn = model(x.cuda(), x_lengths.cuda(), y.cuda(), y_lengths.cuda(), out_size=out_size, t=t.cuda(), z=z.cuda())
    # oracle composition
    mu_x, logw, x_mask = oenc.text_encoder_forward(we, x, x_lengths)
    y_mask = oal.sequence_mask(y_lengths, T_y).unsqueeze(1).float()
    oattn = oal.maximum_path(oal.log_prior(mu_x, y), (x_mask.unsqueeze(-1) * y_mask.unsqueeze(2)).squeeze(1))
    random.seed(5)
    max_offset = (y_lengths - out_size).clamp(0)
    offs = [random.choice(range(0, int(e))) if int(e) > 0 else 0 for e in max_offset]
    attn_cut, y_cut, lens = torch.zeros(B, T_x, out_size), torch.zeros(B, 80, out_size), []
    for i, off in enumerate(offs):
        n = out_size + min(int(y_lengths[i]) - out_size, 0)
        lens.append(n)
        y_cut[i, :, :n] = y[i, :, off:off + n]
        attn_cut[i, :, :n] = oattn[i, :, off:off + n]
    cut_mask = oal.sequence_mask(torch.tensor(lens), out_size).unsqueeze(1).float()
    mu_y = torch.matmul(attn_cut.transpose(1, 2), mu_x.transpose(1, 2)).transpose(1, 2)
    oprior = torch.sum(0.5 * ((y_cut - mu_y) ** 2 + math.log(2 * math.pi)) * cut_mask) / (torch.sum(cut_mask) * 80)
    odiff, _ = odec.compute_loss(wd, y_cut, cut_mask, mu_y, t, z)
    assert attn.shape == (B, T_x, out_size) and torch.equal(attn.cpu(), attn_cut)
    assert abs(float(prior) - float(oprior)) < 1e-5 * abs(float(oprior))
    assert abs(float(diff) - float(odiff)) < 2e-5 * abs(float(odiff))
