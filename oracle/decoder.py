"""oracle/decoder.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatement (plain torch ops, any float dtype) of the reference's CFM decoder path:
  sinusoidal_embedding  <- SinusoidalPosEmb.forward            decoder.py:22-31
  time_mlp              <- TimeStepEmbeddingNet.forward        decoder.py:44-49
  block1d / resnet1d    <- Block1D / Resnet1D                  decoder.py:64-66, 80-86
  transformer_block     <- BasicTransformerBlock (diffusers branch, self-attention + FFN only)
                           transformer.py:297-370, with diffusers' Attention/GELU restated from
                           their 0.25-era semantics (third-party, un-vendored: see SURVEY.md F2,
                           Appendix A): to_q/k/v bias-free, 4 heads x 64, logits q.k/8 + mask01[key]
                           (float mask ADDED), erf-GELU.
  decoder_forward       <- Decoder.forward                     decoder.py:271-371
  solve_euler           <- BaseConditionalFlowMatching.solve_ode_euler   flow_matching.py:67-104
  compute_loss          <- BaseConditionalFlowMatching.compute_loss      flow_matching.py:106-151
(paths under /root/reference/matcha/models/components/).

Parity: the reference has no test or golden vector for this path ("parity unpinned" by the
reference itself).  This restatement is pinned against the reference's own modules, imported
unmodified in the build container with the diffusers stand-in of oracle/_shims, through the
fixtures of tests/golden/decoder_golden.npz (tests/golden/make_decoder_golden.py,
tests/test_oracle_decoder.py).

Weights are a flat dict with the reference's state_dict names (without the "estimator." prefix).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.
"""
import math
from collections import OrderedDict

import torch
import torch.nn.functional as F

C = 256          # channels=(256, 256)                       matcha_tts.py:146-153
HEADS, DHEAD = 4, 64
TIME_DIM = 1024  # channels[0] * 4                           decoder.py:166
IN_CH, OUT_CH = 160, 80


# ----------------------------------------------------------------------------- weights
def param_shapes():
    """state_dict keys/shapes of Decoder(in_channels=160, out_channels=80, channels=(256,256),
    n_blocks=1, num_mid_blocks=2, num_heads=4, attention_head_dim=64) (SURVEY §8 b-names)."""
    s = OrderedDict()

    def lin(name, o, i, bias=True):
        s[name + ".weight"] = (o, i)
        if bias:
            s[name + ".bias"] = (o,)

    def conv(name, o, i, k):
        s[name + ".weight"] = (o, i, k)
        s[name + ".bias"] = (o,)

    def norm(name):
        s[name + ".weight"] = (C,)
        s[name + ".bias"] = (C,)

    def resnet(p, cin):
        lin(p + ".mlp.1", C, TIME_DIM)
        conv(p + ".block1.block.0", C, cin, 3); norm(p + ".block1.block.1")
        conv(p + ".block2.block.0", C, C, 3); norm(p + ".block2.block.1")
        conv(p + ".res_conv", C, cin, 1)

    def tblock(p):
        norm(p + ".norm1")
        for n in ("to_q", "to_k", "to_v"):
            lin(p + ".attn1." + n, C, C, bias=False)
        lin(p + ".attn1.to_out.0", C, C)
        norm(p + ".norm3")
        lin(p + ".ff.net.0.proj", 4 * C, C)
        lin(p + ".ff.net.2", C, 4 * C)

    lin("time_mlp.linear_1", TIME_DIM, IN_CH)
    lin("time_mlp.linear_2", TIME_DIM, TIME_DIM)
    resnet("Downsampling_Blocks.0.0", IN_CH); tblock("Downsampling_Blocks.0.1.0"); conv("Downsampling_Blocks.0.2.conv", C, C, 3)
    resnet("Downsampling_Blocks.1.0", C); tblock("Downsampling_Blocks.1.1.0"); conv("Downsampling_Blocks.1.2", C, C, 3)
    for i in range(2):
        resnet(f"Mid_Blocks.{i}.0", C); tblock(f"Mid_Blocks.{i}.1.0")
    resnet("Upsampling_Blocks.0.0", 2 * C); tblock("Upsampling_Blocks.0.1.0"); conv("Upsampling_Blocks.0.2.conv", C, C, 4)
    resnet("Upsampling_Blocks.1.0", 2 * C); tblock("Upsampling_Blocks.1.1.0"); conv("Upsampling_Blocks.1.2", C, C, 3)
    conv("final_conv", C, C, 3); norm("final_norm"); conv("final_proj", OUT_CH, C, 1)
    return s


def make_weights(seed=0, style="rich"):
    """Deterministic fp32 weights (CPU generator, key order of param_shapes()).
    style="init": the distribution of Decoder.initialize_weights (decoder.py:255-268): Kaiming-normal
                  (relu) matrices, zero biases, unit norms -- the "random-init" of the benchmarks.
    style="rich": same matrices but non-zero biases and non-trivial norm affines, so parity tests
                  exercise every bias/affine path."""
    g = torch.Generator().manual_seed(seed)
    w = OrderedDict()
    for name, shape in param_shapes().items():
        is_norm = ".norm" in name or name.startswith("final_norm") or ".block.1." in name
        if name.endswith(".weight") and not is_norm:
            fan_in = 1
            for d in shape[1:]:
                fan_in *= d
            w[name] = torch.randn(shape, generator=g) * math.sqrt(2.0 / fan_in)
        elif name.endswith(".weight"):
            w[name] = torch.ones(shape) + (0.1 * torch.randn(shape, generator=g) if style == "rich" else 0.0)
        else:
            w[name] = 0.1 * torch.randn(shape, generator=g) if style == "rich" else torch.zeros(shape)
    return w


def cast_weights(w, dtype):
    return OrderedDict((k, v.to(dtype)) for k, v in w.items())


# ----------------------------------------------------------------------------- layers
def sinusoidal_embedding(t, dim=IN_CH, scale=1000):
    """decoder.py:22-31.  The frequency table is built in fp32 exactly as the reference does."""
    if t.ndim < 1:
        t = t.unsqueeze(0)
    half = dim // 2
    emb = torch.log(torch.tensor(10000.0)) / (half - 1)
    emb = torch.exp(torch.arange(half, dtype=torch.float32) * -emb)
    emb = scale * t.unsqueeze(1) * emb.unsqueeze(0).to(device=t.device, dtype=t.dtype)
    return torch.cat((emb.sin(), emb.cos()), dim=-1)


def time_mlp(w, temb):
    h = F.linear(temb, w["time_mlp.linear_1.weight"], w["time_mlp.linear_1.bias"])
    return F.linear(F.silu(h), w["time_mlp.linear_2.weight"], w["time_mlp.linear_2.bias"])


def block1d(w, p, x, mask):
    h = F.conv1d(x * mask, w[p + ".block.0.weight"], w[p + ".block.0.bias"], padding=1)     # decoder.py:59,65
    h = F.group_norm(h, 8, w[p + ".block.1.weight"], w[p + ".block.1.bias"], eps=1e-5)      # :60
    return F.mish(h) * mask                                                                 # :61,66


def resnet1d(w, p, x, mask, temb):
    h = block1d(w, p + ".block1", x, mask)                                                  # :81
    h = h + F.linear(F.mish(temb), w[p + ".mlp.1.weight"], w[p + ".mlp.1.bias"])[:, :, None]  # :82-83
    h = block1d(w, p + ".block2", h, mask)                                                  # :84
    return h + F.conv1d(x * mask, w[p + ".res_conv.weight"], w[p + ".res_conv.bias"])       # :85 (unmasked)


# True: the attention core goes through F.scaled_dot_product_attention with the float mask as attn_mask, which is
# how diffusers' AttnProcessor2_0 calls it (the same-box GPU baseline of bench.py uses this form so that torch may
# pick its fused kernels); False: the explicit softmax below (the form the golden fixtures pin).  Same arithmetic.
USE_SDPA = False


def transformer_block(w, p, x, mask_bt):
    """x [B,T,C]; mask_bt [B,T] float 0/1 (transformer.py:316-368 with attn2=None)."""
    B, T, _ = x.shape
    h = F.layer_norm(x, (C,), w[p + ".norm1.weight"], w[p + ".norm1.bias"], eps=1e-5)
    q = F.linear(h, w[p + ".attn1.to_q.weight"]).view(B, T, HEADS, DHEAD).transpose(1, 2)
    k = F.linear(h, w[p + ".attn1.to_k.weight"]).view(B, T, HEADS, DHEAD).transpose(1, 2)
    v = F.linear(h, w[p + ".attn1.to_v.weight"]).view(B, T, HEADS, DHEAD).transpose(1, 2)
    if USE_SDPA:
        a = F.scaled_dot_product_attention(q, k, v, attn_mask=mask_bt[:, None, None, :].to(q.dtype), dropout_p=0.0, is_causal=False)
    else:
        logits = q @ k.transpose(-1, -2) * (DHEAD ** -0.5) + mask_bt[:, None, None, :]      # float mask is ADDED
        a = torch.softmax(logits, dim=-1) @ v
    a = a.transpose(1, 2).reshape(B, T, C)
    x = x + F.linear(a, w[p + ".attn1.to_out.0.weight"], w[p + ".attn1.to_out.0.bias"])
    h = F.layer_norm(x, (C,), w[p + ".norm3.weight"], w[p + ".norm3.bias"], eps=1e-5)
    h = F.gelu(F.linear(h, w[p + ".ff.net.0.proj.weight"], w[p + ".ff.net.0.proj.bias"]))   # erf form
    return x + F.linear(h, w[p + ".ff.net.2.weight"], w[p + ".ff.net.2.bias"])


def _stage(w, p, x, mask, temb):
    x = resnet1d(w, p + ".0", x, mask, temb)
    x = transformer_block(w, p + ".1.0", x.transpose(1, 2), mask[:, 0, :]).transpose(1, 2)
    return x


def decoder_forward(w, x, mask, mu, t):
    """Decoder.forward (decoder.py:271-371).  x, mu [B,80,T]; mask [B,1,T] float; t 0-d or [B]."""
    temb = time_mlp(w, sinusoidal_embedding(t).to(x.dtype))                                 # :285-286
    x = torch.cat([x, mu], dim=1)                                                           # :288
    hiddens, masks = [], [mask]
    # down                                                                                  :293-316
    m = masks[-1]
    x = _stage(w, "Downsampling_Blocks.0", x, m, temb); hiddens.append(x)
    x = F.conv1d(x * m, w["Downsampling_Blocks.0.2.conv.weight"], w["Downsampling_Blocks.0.2.conv.bias"], stride=2, padding=1)
    masks.append(m[:, :, : (m.shape[-1] + 1) // 2])                                          # prefix truncation, F4
    m = masks[-1]
    x = _stage(w, "Downsampling_Blocks.1", x, m, temb); hiddens.append(x)
    x = F.conv1d(x * m, w["Downsampling_Blocks.1.2.weight"], w["Downsampling_Blocks.1.2.bias"], padding=1)
    masks.append(m)
    masks = masks[:-1]                                                                      # :318
    m_mid = masks[-1]
    for i in range(2):                                                                      # :321-332
        x = _stage(w, f"Mid_Blocks.{i}", x, m_mid, temb)
    # up                                                                                    :334-364
    for i in range(2):
        m_up = masks.pop()
        hidden = hiddens.pop()
        if x.shape[-1] != hidden.shape[-1]:
            x = F.interpolate(x, size=hidden.shape[-1], mode="nearest")
        x = torch.cat([x, hidden], dim=1)
        x = _stage(w, f"Upsampling_Blocks.{i}", x, m_up, temb)
        if i == 0:
            x = F.conv_transpose1d(x * m_up, w["Upsampling_Blocks.0.2.conv.weight"], w["Upsampling_Blocks.0.2.conv.bias"], stride=2, padding=1)
            m_up = F.interpolate(m_up, size=m_up.shape[-1] * 2, mode="nearest")
        else:
            x = F.conv1d(x * m_up, w["Upsampling_Blocks.1.2.weight"], w["Upsampling_Blocks.1.2.bias"], padding=1)
    x = F.conv1d(x * m_up, w["final_conv.weight"], w["final_conv.bias"], padding=1)         # :366
    x = F.mish(F.group_norm(x, 8, w["final_norm.weight"], w["final_norm.bias"], eps=1e-5))  # :367-368
    out = F.conv1d(x * m_up, w["final_proj.weight"], w["final_proj.bias"])                  # :369
    return out * mask                                                                       # :371


def euler_time_grid(t_span):
    """The (t_k, dt_k) the reference's loop produces, in t_span's own dtype (flow_matching.py:85-102)."""
    ts, dts = [], []
    t = t_span[0]
    dt = t_span[1] - t_span[0]
    n = len(t_span) - 1
    for step in range(1, n + 1):
        ts.append(t.clone()); dts.append(dt.clone())
        t = t + dt
        if step < n:
            dt = t_span[step + 1] - t
    return torch.stack(ts), torch.stack(dts)


def solve_euler(w, x, t_span, mu, mask):
    """solve_ode_euler (flow_matching.py:67-104); time arithmetic stays in t_span's dtype."""
    ts, dts = euler_time_grid(t_span)
    for t, dt in zip(ts, dts):
        v = decoder_forward(w, x, mask, mu, t.to(device=x.device, dtype=x.dtype))
        x = x + dt.to(device=x.device, dtype=x.dtype) * v
    return x


def compute_loss(w, x1, mask, mu, t, z, sigma_min=1e-4):
    """compute_loss (flow_matching.py:106-151) with the random draws t [B,1,1] and z supplied."""
    phi_t = (1 - (1 - sigma_min) * t) * z + t * x1
    u_target = x1 - (1 - sigma_min) * z
    u_pred = decoder_forward(w, phi_t, mask, mu, t.squeeze())
    loss = F.mse_loss(u_pred, u_target, reduction="sum") / (torch.sum(mask) * u_target.shape[1])
    return loss, phi_t
