"""fp16-operand mode: range of the residual stream (csrc/decoder.cu, "fp16 range guard").

IEEE half tops out at 65504.  Everything the decoder feeds the tensor cores sits behind a GroupNorm / LayerNorm / softmax
except the residual-stream tensors (stage outputs and conv inputs), whose size the weights set.  The fp16 mode stores those
scaled by 2^-4 and un-scales in the consuming GEMM's epilogue, so it keeps the north_star tolerance (relative L2 <= 2e-3
against the fp32 reference) for stream magnitudes up to 65504 * 16 ~ 1.04e6, and SATURATES beyond that (finite output)
rather than emitting inf / NaN.  Checker: oracle/decoder.py in fp64 on the same weights and inputs."""
from types import SimpleNamespace

import pytest
import torch

from oracle import decoder as odec

pytestmark = pytest.mark.gpu
PARAMS = dict(channels=(256, 256), dropout=0.05, attention_head_dim=64, n_blocks=1, num_mid_blocks=2, num_heads=4)
FFN2 = "Downsampling_Blocks.0.1.0.ff.net.2."


def rel_l2(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float(torch.linalg.norm(a - b) / torch.linalg.norm(b))


def make_cfm(weights, precision):
    from matcha_b200.flow_matching import CFM
    m = CFM(160, 80, SimpleNamespace(solver="euler", sigma_min=1e-4), PARAMS, precision=precision)
    m.load_state_dict({"estimator." + k: v for k, v in weights.items()})
    return m


def inputs(B, T, lengths, seed):
    g = torch.Generator().manual_seed(seed)
    x, mu = torch.randn(B, 80, T, generator=g).cuda(), torch.randn(B, 80, T, generator=g).cuda()
    mask = torch.zeros(B, 1, T)
    for i, n in enumerate(lengths):
        mask[i, :, :n] = 1
    return x, mu, mask.cuda(), torch.rand(B, generator=g).cuda()


class _ConvProbe:
    """Stands in for torch.nn.functional inside the oracle and records max |input| over every convolution: the conv inputs
    are exactly the residual-stream tensors (stage inputs / outputs, skip connections) plus a few normalised ones."""

    def __init__(self):
        self.peak = 0.0

    def __getattr__(self, name):
        return getattr(torch.nn.functional, name)

    def conv1d(self, x, *a, **k):
        self.peak = max(self.peak, float(x.abs().max()))
        return torch.nn.functional.conv1d(x, *a, **k)

    def conv_transpose1d(self, x, *a, **k):
        self.peak = max(self.peak, float(x.abs().max()))
        return torch.nn.functional.conv_transpose1d(x, *a, **k)


def stream_peak(w64, x, mu, mask, t):
    probe, real = _ConvProbe(), odec.F
    odec.F = probe
    try:
        odec.decoder_forward(w64, x.double(), mask.double(), mu.double(), t.double())
    finally:
        odec.F = real
    return probe.peak


def blown_up_weights(target_peak, B, T, lengths, seed):
    """The 'rich' random weights with stage 0's second feed-forward matrix scaled until the largest residual-stream value of
    the whole U-Net (fp64 oracle, these inputs) sits near target_peak."""
    base = odec.make_weights(23, "rich")
    x, mu, mask, t = inputs(B, T, lengths, seed)
    f, peak = 1.0, None
    for _ in range(4):
        w = dict(base)
        for leaf in ("weight", "bias"):
            w[FFN2 + leaf] = base[FFN2 + leaf] * f
        w64 = odec.cast_weights({k: v.cuda() for k, v in w.items()}, torch.float64)
        peak = stream_peak(w64, x, mu, mask, t)
        if 0.8 * target_peak < peak < 1.2 * target_peak:
            break
        f *= target_peak / peak
    return w, w64, peak


@pytest.mark.parametrize("T,lengths", [(160, [160, 97, 160]), (131, [131, 131, 60])], ids=["evenT", "oddT"])
def test_fp16_mode_holds_tolerance_with_stream_beyond_half_range(T, lengths):
    """Stream peak ~5e5 (8x the largest finite half): the un-guarded round-1 kernels returned NaN here.  Even T takes the
    tensor-core stride-2 conv, odd T the fp32 fallback conv and the nearest resample (both write stream tensors)."""
    w, w64, peak = blown_up_weights(5.0e5, 3, T, lengths, seed=T)
    assert 65504 * 4 < peak < 65504 * 16 * 0.8, peak
    x, mu, mask, t = inputs(3, T, lengths, seed=T)
    want = odec.decoder_forward(w64, x.double(), mask.double(), mu.double(), t.double())
    got32 = make_cfm(w, "fp32").estimator(x, mask, mu, t)
    got16 = make_cfm(w, "fp16").estimator(x, mask, mu, t)
    e32, e16 = rel_l2(got32, want), rel_l2(got16, want)
    print("stream peak", peak, "fp32", e32, "fp16", e16)
    assert e32 < 1e-5
    assert torch.isfinite(got16).all()
    assert e16 < 2e-3


@pytest.mark.parametrize("target", [1.0, 3.0e4])
def test_fp16_control_in_range(target):
    """Control: the same construction with the stream inside half range (the unscaled weights, and a 3e4 peak)."""
    if target == 1.0:
        w = odec.make_weights(23, "rich")
        w64 = odec.cast_weights({k: v.cuda() for k, v in w.items()}, torch.float64)
        peak = None
    else:
        w, w64, peak = blown_up_weights(target, 3, 160, [160, 97, 160], seed=160)
    x, mu, mask, t = inputs(3, 160, [160, 97, 160], seed=160)
    want = odec.decoder_forward(w64, x.double(), mask.double(), mu.double(), t.double())
    e16 = rel_l2(make_cfm(w, "fp16").estimator(x, mask, mu, t), want)
    print("control peak", peak, "fp16", e16)
    assert e16 < 2e-3


def test_fp16_mode_solve_with_large_stream():
    w, w64, peak = blown_up_weights(4.0e5, 2, 96, [96, 61], seed=5)
    x, mu, mask, _ = inputs(2, 96, [96, 61], seed=5)
    t_span = torch.linspace(0, 1, 5).cuda()
    want = odec.solve_euler(w64, x.double(), t_span.cpu(), mu.double(), mask.double())
    got = make_cfm(w, "fp16").solve_ode_euler(x, t_span, mu, mask, None, None)
    err = rel_l2(got, want)
    print("solve, stream peak", peak, "fp16", err)
    assert err < 2e-3


def test_fp16_mode_saturates_instead_of_nan_beyond_guard():
    """Stream peak ~1e8, far past 65504 * 16: accuracy is gone, but the packs saturate -- the output stays finite."""
    w, _, peak = blown_up_weights(1.0e8, 2, 64, [64, 40], seed=9)
    assert peak > 65504 * 16 * 8
    x, mu, mask, t = inputs(2, 64, [64, 40], seed=9)
    got = make_cfm(w, "fp16").estimator(x, mask, mu, t)
    assert torch.isfinite(got).all()


def test_bf16_mode_unaffected_by_large_stream():
    """bf16 has fp32's exponent range: no guard is applied (scale 1) and the functional bound holds."""
    w, w64, _ = blown_up_weights(2.5e5, 2, 96, [96, 70], seed=3)
    x, mu, mask, t = inputs(2, 96, [96, 70], seed=3)
    want = odec.decoder_forward(w64, x.double(), mask.double(), mu.double(), t.double())
    got = make_cfm(w, "bf16").estimator(x, mask, mu, t)
    assert rel_l2(got, want) < 1.5e-2
