"""oracle/text_encoder.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatement (plain torch ops, any float dtype) of the reference's text encoder
(/root/reference/matcha/models/components/text_encoder.py; SURVEY.md §8f row 4, BASELINE configs[3]
"text-encoder mu_x -> MAS -> CFM loss"):
  conv_relu_norm      <- ConvReluNorm.forward                  text_encoder.py:48-57
  duration_predictor  <- DurationPredictor.forward             text_encoder.py:80-97
  rope                <- RotaryPositionalEmbeddings            text_encoder.py:110-148
  attention           <- MultiHeadAttention.forward / _compute_attention   text_encoder.py:191-229
  ffn                 <- FFN.forward                           text_encoder.py:257-260
  encoder             <- Encoder.forward                       text_encoder.py:307-331
  text_encoder_forward<- TextEncoder.forward                   text_encoder.py:377-402
Configuration: the one MatchaTTS builds (matcha_tts.py:122-139): 192 channels, 768 filter channels, 2 heads,
6 layers, kernel 3, prenet on; duration predictor 256 filter channels, kernel 3; 80 output features.
Dropout is the identity (eval mode).

Parity: the reference has no test or golden vector for this module.  This restatement is pinned against
the reference's own TextEncoder, imported unmodified in the build container, through
tests/golden/text_encoder_golden.npz (tests/golden/make_text_encoder_golden.py, tests/test_oracle_text_encoder.py):
four cases (ragged, full-length, a one-token item, T=160), reproduced bit for bit in fp32.

Weights are a flat dict with the reference's state_dict names.  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / gpu_baseline legs may import this module.
"""
import math
from collections import OrderedDict

import torch
import torch.nn.functional as F

C = 192            # n_channels                                  matcha_tts.py:124
FILTER = 768       # filter_channels                             :125
HEADS = 2          # n_heads                                     :126
DHEAD = C // HEADS
LAYERS = 6         # n_layers                                    :127
KERNEL = 3         # kernel_size                                 :128
DP_FILTER = 256    # filter_channels_dp                          :134
DP_KERNEL = 3
N_FEATS = 80
PRENET_KERNEL, PRENET_LAYERS = 5, 3                             # text_encoder.py:352-359
ROPE_DIM = int(DHEAD * 0.5)                                     # text_encoder.py:178


def param_shapes(n_vocab=178):
    """state_dict keys / shapes of TextEncoder(encoder_type, encoder_params, duration_predictor_params, n_vocab)."""
    s = OrderedDict()

    def conv(name, o, i, k):
        s[name + ".weight"] = (o, i, k)
        s[name + ".bias"] = (o,)

    def norm(name, c):
        s[name + ".weight"] = (c,)
        s[name + ".bias"] = (c,)

    s["embedding.weight"] = (n_vocab, C)
    for i in range(PRENET_LAYERS):
        conv(f"prenet.convolutions.{i}", C, C, PRENET_KERNEL)
    for i in range(PRENET_LAYERS):
        norm(f"prenet.normalizations.{i}", C)
    conv("prenet.projection", C, C, 1)
    for i in range(LAYERS):
        for n in ("query_conv", "key_conv", "value_conv", "output_conv"):
            conv(f"encoder.attention_layers.{i}.{n}", C, C, 1)
    for i in range(LAYERS):
        norm(f"encoder.norm_layers_1.{i}", C)
    for i in range(LAYERS):
        conv(f"encoder.ffn_layers.{i}.conv_net.0", FILTER, C, KERNEL)
        conv(f"encoder.ffn_layers.{i}.conv_net.3", C, FILTER, KERNEL)
    for i in range(LAYERS):
        norm(f"encoder.norm_layers_2.{i}", C)
    conv("mean_projection", N_FEATS, C, 1)
    conv("duration_predictor.conv_layer_1", DP_FILTER, C, DP_KERNEL); norm("duration_predictor.norm_layer_1", DP_FILTER)
    conv("duration_predictor.conv_layer_2", DP_FILTER, DP_FILTER, DP_KERNEL); norm("duration_predictor.norm_layer_2", DP_FILTER)
    conv("duration_predictor.output_projection", 1, DP_FILTER, 1)
    return s


def make_weights(seed=0, style="rich", n_vocab=178):
    """Deterministic fp32 weights (CPU generator, key order of param_shapes()).
    style="init": roughly what the reference constructor leaves (embedding N(0, C^-0.5), uniform fan-in conv
                  weights and biases, unit norms, ZERO prenet projection, text_encoder.py:44-46, 347).
    style="rich": every matrix, bias and affine non-trivial (incl. the prenet projection) so the parity tests
                  exercise every term."""
    g = torch.Generator().manual_seed(seed)
    w = OrderedDict()
    for name, shape in param_shapes(n_vocab).items():
        is_norm = "norm" in name
        if name == "embedding.weight":
            w[name] = torch.randn(shape, generator=g) * C ** -0.5
        elif name.startswith("prenet.projection") and style == "init":
            w[name] = torch.zeros(shape)
        elif is_norm and name.endswith(".weight"):
            w[name] = torch.ones(shape) + (0.1 * torch.randn(shape, generator=g) if style == "rich" else 0.0)
        elif is_norm:
            w[name] = 0.1 * torch.randn(shape, generator=g) if style == "rich" else torch.zeros(shape)
        else:
            wshape = param_shapes(n_vocab)[name.rsplit(".", 1)[0] + ".weight"]
            fan_in = wshape[1] * wshape[2]
            bound = 1.0 / math.sqrt(fan_in)
            if name.endswith(".weight") and style == "rich":
                w[name] = torch.randn(shape, generator=g) * math.sqrt(1.0 / fan_in)
            else:
                w[name] = (torch.rand(shape, generator=g) * 2 - 1) * bound
    return w


def cast_weights(w, dtype):
    return OrderedDict((k, v.to(dtype)) for k, v in w.items())


def sequence_mask(length, max_length):
    """utils/model.py:13-34"""
    x = torch.arange(max_length, dtype=length.dtype, device=length.device)
    return x.unsqueeze(0) < length.unsqueeze(1)


def _ln(x_bct, w, p):
    """nn.LayerNorm over channels of a [B,C,T] tensor (the transposes of text_encoder.py:52-54)."""
    c = x_bct.shape[1]
    return F.layer_norm(x_bct.transpose(1, 2), (c,), w[p + ".weight"], w[p + ".bias"], eps=1e-5).transpose(1, 2)


def conv_relu_norm(w, x, x_mask):
    """ConvReluNorm.forward, text_encoder.py:48-57."""
    residual = x
    for i in range(PRENET_LAYERS):
        x = F.conv1d(x * x_mask, w[f"prenet.convolutions.{i}.weight"], w[f"prenet.convolutions.{i}.bias"],
                     padding=PRENET_KERNEL // 2)                                             # :51
        x = _ln(x, w, f"prenet.normalizations.{i}")                                          # :52-54
        x = torch.relu(x)                                                                    # :55 (dropout = identity)
    x = residual + F.conv1d(x, w["prenet.projection.weight"], w["prenet.projection.bias"])   # :56
    return x * x_mask                                                                        # :57


def duration_predictor(w, x, x_mask):
    """DurationPredictor.forward, text_encoder.py:80-97."""
    p = "duration_predictor."
    x = F.conv1d(x * x_mask, w[p + "conv_layer_1.weight"], w[p + "conv_layer_1.bias"], padding=DP_KERNEL // 2)
    x = _ln(torch.relu(x), w, p + "norm_layer_1")
    x = F.conv1d(x * x_mask, w[p + "conv_layer_2.weight"], w[p + "conv_layer_2.bias"], padding=DP_KERNEL // 2)
    x = _ln(torch.relu(x), w, p + "norm_layer_2")
    x = F.conv1d(x * x_mask, w[p + "output_projection.weight"], w[p + "output_projection.bias"])
    return x * x_mask


def rope_tables(seq_len, device=None):
    """The cos / sin caches of RotaryPositionalEmbeddings._build_cache (text_encoder.py:110-128), built in fp32
    with the reference's own torch ops: [seq_len, ROPE_DIM] each (theta duplicated over the two halves)."""
    theta = 1.0 / (10_000 ** (torch.arange(0, ROPE_DIM, 2).float() / ROPE_DIM))
    pos = torch.arange(seq_len).float()
    pt = torch.einsum("n,d->nd", pos, theta)
    pt2 = torch.cat([pt, pt], dim=1)
    cos, sin = pt2.cos(), pt2.sin()
    if device is not None:
        cos, sin = cos.to(device), sin.to(device)
    return cos, sin


def rope(x):
    """RotaryPositionalEmbeddings.forward on x [B,H,T,D] (text_encoder.py:135-148): rotates the first ROPE_DIM
    features, passes the rest through."""
    T = x.shape[2]
    cos, sin = rope_tables(T, x.device)
    cos, sin = cos.to(x.dtype)[None, None], sin.to(x.dtype)[None, None]
    xr, xp = x[..., :ROPE_DIM], x[..., ROPE_DIM:]
    half = ROPE_DIM // 2
    neg_half = torch.cat([-xr[..., half:], xr[..., :half]], dim=-1)                          # :130-133
    return torch.cat([xr * cos + neg_half * sin, xp], dim=-1)


def attention(w, p, x, attn_mask):
    """MultiHeadAttention.forward with context = x (text_encoder.py:191-229)."""
    B, _, T = x.shape
    q = F.conv1d(x, w[p + ".query_conv.weight"], w[p + ".query_conv.bias"])
    k = F.conv1d(x, w[p + ".key_conv.weight"], w[p + ".key_conv.bias"])
    v = F.conv1d(x, w[p + ".value_conv.weight"], w[p + ".value_conv.bias"])
    q = rope(q.view(B, HEADS, DHEAD, T).transpose(2, 3))                                     # "b (h c) t -> b h t c"
    k = rope(k.view(B, HEADS, DHEAD, T).transpose(2, 3))
    v = v.view(B, HEADS, DHEAD, T).transpose(2, 3)
    scores = torch.matmul(q, k.transpose(-2, -1)) / math.sqrt(DHEAD)                          # :213
    scores = scores.masked_fill(attn_mask == 0, -1e4)                                        # :220
    out = torch.matmul(torch.softmax(scores, dim=-1), v)                                     # :222-224
    out = out.transpose(2, 3).contiguous().view(B, C, T)                                     # :225
    return F.conv1d(out, w[p + ".output_conv.weight"], w[p + ".output_conv.bias"])


def ffn(w, p, x, x_mask):
    """FFN.forward, text_encoder.py:257-260 (only the first conv's input and the result are masked)."""
    h = F.conv1d(x * x_mask, w[p + ".conv_net.0.weight"], w[p + ".conv_net.0.bias"], padding=KERNEL // 2)
    h = F.conv1d(torch.relu(h), w[p + ".conv_net.3.weight"], w[p + ".conv_net.3.bias"], padding=KERNEL // 2)
    return h * x_mask


def encoder(w, x, x_mask):
    """Encoder.forward, text_encoder.py:307-331."""
    attn_mask = x_mask.unsqueeze(2) * x_mask.unsqueeze(-1)                                   # :308
    for i in range(LAYERS):
        x = x * x_mask                                                                       # :311
        a = attention(w, f"encoder.attention_layers.{i}", x, attn_mask)
        x = _ln(x + a, w, f"encoder.norm_layers_1.{i}")                                      # :315-318
        f = ffn(w, f"encoder.ffn_layers.{i}", x, x_mask)
        x = _ln(x + f, w, f"encoder.norm_layers_2.{i}")                                      # :322-326
    return x * x_mask                                                                        # :328


def text_encoder_forward(w, text_input, text_lengths):
    """TextEncoder.forward (text_encoder.py:377-402): ids [B,T] long, lengths [B] ->
    (mu [B,80,T], log_duration [B,1,T], mask [B,1,T])."""
    dtype = w["embedding.weight"].dtype
    x = F.embedding(text_input, w["embedding.weight"]) * math.sqrt(C)                        # :390
    x = torch.transpose(x, 1, -1)
    x_mask = torch.unsqueeze(sequence_mask(text_lengths, x.size(2)), 1).to(dtype)            # :392
    x = conv_relu_norm(w, x, x_mask)                                                         # :394
    x = encoder(w, x, x_mask)                                                                # :396
    mu = F.conv1d(x, w["mean_projection.weight"], w["mean_projection.bias"]) * x_mask        # :398
    logw = duration_predictor(w, x, x_mask)                                                  # :400-401
    return mu, logw, x_mask
