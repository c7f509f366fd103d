"""Drop-in for ``matcha.models.components.text_encoder.TextEncoder`` backed by the sm_100a kernels of
csrc/text_encoder.cu (BASELINE configs[3] "text-encoder mu_x -> MAS -> CFM loss").

Mirrors /root/reference/matcha/models/components/text_encoder.py:338-402:
  ``TextEncoder(encoder_type, encoder_params, duration_predictor_params, n_vocab).forward(text_input, text_lengths)``
  -> ``(mean_output [B,80,T], log_duration [B,1,T], text_mask [B,1,T])``
with the reference's ``state_dict`` names, so a Lightning checkpoint (``encoder.*``) or a freshly built reference
module can be copied over 1:1 with ``load_state_dict``.  Inference / eval only (dropout is the identity, no autograd);
all arithmetic happens in ``lib/libmatcha_b200.so`` -- no CPU fallback.
"""
import math
import os
from collections import OrderedDict
from types import SimpleNamespace

import torch

from ._lib import check, load_library
from .decoder import PREC_F16, PREC_F32, _stream_ptr

_PRECISIONS = {"fp32": PREC_F32, "f32": PREC_F32, "fp16": PREC_F16, "f16": PREC_F16}

C, FILTER, HEADS, LAYERS, KERNEL = 192, 768, 2, 6, 3
DP_FILTER, DP_KERNEL, N_FEATS = 256, 3, 80
PRENET_KERNEL, PRENET_LAYERS = 5, 3
ROPE_DIM = int((C // HEADS) * 0.5)                       # text_encoder.py:178


def default_params():
    """The (encoder_params, duration_predictor_params) MatchaTTS builds (matcha_tts.py:122-139)."""
    return (SimpleNamespace(n_feats=N_FEATS, n_channels=C, filter_channels=FILTER, n_heads=HEADS, n_layers=LAYERS,
                            kernel_size=KERNEL, p_dropout=0.1, prenet=True),
            SimpleNamespace(filter_channels_dp=DP_FILTER, kernel_size=DP_KERNEL, p_dropout=0.1))


def expected_keys(n_vocab):
    """The reference's state_dict keys and shapes for the configured text encoder."""
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


def _taps_major(w):
    """Conv1d weight [out, cin, k] -> [out, k, cin]: GEMM K index = tap * cin + channel."""
    return w.permute(0, 2, 1).contiguous()


def pack_weights(sd, n_vocab):
    """Packs a reference state_dict into the flat fp32 blob of csrc/text_encoder.cu::build_weights (same order)."""
    w = {k: v.detach().to("cpu", torch.float32) for k, v in sd.items()}
    parts = [w["embedding.weight"]]
    for i in range(PRENET_LAYERS):
        parts += [_taps_major(w[f"prenet.convolutions.{i}.weight"]), w[f"prenet.convolutions.{i}.bias"],
                  w[f"prenet.normalizations.{i}.weight"], w[f"prenet.normalizations.{i}.bias"]]
    parts += [w["prenet.projection.weight"][:, :, 0], w["prenet.projection.bias"]]
    for i in range(LAYERS):
        a = f"encoder.attention_layers.{i}."
        parts += [torch.cat([w[a + "query_conv.weight"][:, :, 0], w[a + "key_conv.weight"][:, :, 0], w[a + "value_conv.weight"][:, :, 0]], 0),
                  torch.cat([w[a + "query_conv.bias"], w[a + "key_conv.bias"], w[a + "value_conv.bias"]], 0),
                  w[a + "output_conv.weight"][:, :, 0], w[a + "output_conv.bias"],
                  w[f"encoder.norm_layers_1.{i}.weight"], w[f"encoder.norm_layers_1.{i}.bias"]]
        f = f"encoder.ffn_layers.{i}.conv_net."
        parts += [_taps_major(w[f + "0.weight"]), w[f + "0.bias"], _taps_major(w[f + "3.weight"]), w[f + "3.bias"],
                  w[f"encoder.norm_layers_2.{i}.weight"], w[f"encoder.norm_layers_2.{i}.bias"]]
    parts += [w["mean_projection.weight"][:, :, 0], w["mean_projection.bias"]]
    d = "duration_predictor."
    parts += [_taps_major(w[d + "conv_layer_1.weight"]), w[d + "conv_layer_1.bias"], w[d + "norm_layer_1.weight"], w[d + "norm_layer_1.bias"],
              _taps_major(w[d + "conv_layer_2.weight"]), w[d + "conv_layer_2.bias"], w[d + "norm_layer_2.weight"], w[d + "norm_layer_2.bias"],
              w[d + "output_projection.weight"].reshape(-1),
              torch.cat([w[d + "output_projection.bias"].reshape(-1), torch.zeros(3)])]      # bias padded to 4 floats
    blob = torch.cat([p.contiguous().reshape(-1) for p in parts])
    n = int(load_library().matcha_b200_text_encoder_blob_floats(n_vocab))
    assert blob.numel() == n, f"packed {blob.numel()} floats, library expects {n}"
    return blob


def rope_tables(seq_len):
    """cos / sin caches of RotaryPositionalEmbeddings._build_cache (text_encoder.py:110-128) with the reference's own
    torch ops in fp32 on the CPU: two [seq_len, 48] tensors (theta duplicated over the two halves)."""
    theta = 1.0 / (10_000 ** (torch.arange(0, ROPE_DIM, 2).float() / ROPE_DIM))
    pos = torch.arange(seq_len).float()
    pt = torch.einsum("n,d->nd", pos, theta)
    pt2 = torch.cat([pt, pt], dim=1)
    return pt2.cos().contiguous(), pt2.sin().contiguous()


class TextEncoder(torch.nn.Module):
    """Same constructor signature and ``forward`` as the reference TextEncoder (text_encoder.py:341-402)."""

    def __init__(self, encoder_type="RoPE Encoder", encoder_params=None, duration_predictor_params=None, n_vocab=178,
                 precision="fp32"):
        """``precision``: "fp32" (SIMT kernels, rel-L2 <= 1e-5 against the reference) or "fp16" (every Conv1d on the
        tcgen05 tensor cores with fp16 operands and fp32 accumulation, rel-L2 <= 2e-3)."""
        super().__init__()
        if precision not in _PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(_PRECISIONS)}")
        self.precision = precision
        self.check_inputs = os.environ.get("MB200_CHECK_INPUTS", "0") == "1"    # validate ids / lengths like nn.Embedding (costs a sync)
        ep, dp = default_params()
        encoder_params = encoder_params or ep
        duration_predictor_params = duration_predictor_params or dp
        got = (encoder_params.n_feats, encoder_params.n_channels, encoder_params.filter_channels, encoder_params.n_heads,
               encoder_params.n_layers, encoder_params.kernel_size, bool(encoder_params.prenet),
               duration_predictor_params.filter_channels_dp, duration_predictor_params.kernel_size)
        if got != (N_FEATS, C, FILTER, HEADS, LAYERS, KERNEL, True, DP_FILTER, DP_KERNEL):
            raise NotImplementedError("matcha_b200.TextEncoder implements the reference's configured architecture only "
                                      "(matcha_tts.py:122-139)")
        self.encoder_type = encoder_type
        self.vocab_size = int(n_vocab)
        self.feature_dim, self.channel_dim = N_FEATS, C
        self._weights = None
        self._blob = {}               # device -> packed fp32 blob
        self._blob16 = {}             # device -> the same blob in fp16 (tensor-core mode)
        self._rope = {}               # device -> (seq_len, cos, sin)
        self._ws = {}                 # one plain-launch workspace kept
        self._graphs = OrderedDict()  # (device, B, T) -> (graph, buffers); every entry owns its workspace and I/O tensors
        self._seen = {}
        self.graph_cache_size = 4
        self.use_graph = True

    # -- weights ----------------------------------------------------------------------------
    def initialize_weights(self, seed=0):
        """Random-init weights with the distributions the reference constructor leaves behind: embedding N(0, 192^-0.5)
        (text_encoder.py:347), xavier-uniform q / k / v (:187-192), torch's default U(+-1/sqrt(fan_in)) for the other
        convolutions and all biases, LayerNorm 1 / 0, ZERO prenet projection (:44-46).  CPU generator."""
        g = torch.Generator().manual_seed(seed)
        sd = OrderedDict()
        exp = expected_keys(self.vocab_size)
        for name, shape in exp.items():
            if name == "embedding.weight":
                sd[name] = torch.randn(shape, generator=g) * C ** -0.5
            elif "norm" in name:
                sd[name] = torch.ones(shape) if name.endswith(".weight") else torch.zeros(shape)
            elif name.startswith("prenet.projection"):
                sd[name] = torch.zeros(shape)
            else:
                wshape = exp[name.rsplit(".", 1)[0] + ".weight"]
                fan_in, fan_out = wshape[1] * wshape[2], wshape[0] * wshape[2]
                bound = 1.0 / math.sqrt(fan_in)
                if name.endswith(".weight") and any(n in name for n in ("query_conv", "key_conv", "value_conv")):
                    bound = math.sqrt(6.0 / (fan_in + fan_out))
                sd[name] = (torch.rand(shape, generator=g) * 2 - 1) * bound
        self.load_state_dict(sd)
        return self

    def load_state_dict(self, state_dict, strict=True):
        sd = OrderedDict()
        for k, v in state_dict.items():
            if k.startswith("encoder.") and not k.startswith(("encoder.attention_layers", "encoder.norm_layers", "encoder.ffn_layers")):
                k = k[len("encoder."):]                     # MatchaTTS checkpoint: the module lives under "encoder."
            sd[k] = v
        exp = expected_keys(self.vocab_size)
        missing = [k for k in exp if k not in sd]
        bad = [k for k in exp if k in sd and tuple(sd[k].shape) != exp[k]]
        if missing or bad:
            raise KeyError(f"state_dict mismatch: missing {missing[:4]}..., wrong shape {bad[:4]}...")
        self._weights = OrderedDict((k, sd[k].detach().to("cpu", torch.float32).clone()) for k in exp)
        self._blob.clear()
        self._blob16.clear()
        self._graphs.clear()          # captured graphs hold pointers into the old weight blob
        self._seen.clear()
        return self

    def state_dict(self, *a, **k):
        return OrderedDict(self._weights or {})

    def _blob_on(self, device):
        if self._weights is None:
            raise RuntimeError("matcha_b200.TextEncoder has no weights: call load_state_dict first")
        key = str(device)
        if key not in self._blob:
            self._blob[key] = pack_weights(self._weights, self.vocab_size).to(device)
        return self._blob[key]

    def _blob16_on(self, device):
        """fp16 twin of the blob, element for element (the C side addresses a tensor at the same element offset)."""
        if _PRECISIONS[self.precision] == PREC_F32:
            return None
        key = str(device)
        if key not in self._blob16:
            self._blob16[key] = self._blob_on(device).to(torch.float16)
        return self._blob16[key]

    def _rope_on(self, device, T):
        key = str(device)
        cur = self._rope.get(key)
        if cur is None or cur[0] < T:
            n = max(T, 512)
            cos, sin = rope_tables(n)
            cur = (n, cos.to(device), sin.to(device))
            self._rope[key] = cur
            self._graphs.clear()      # graphs hold pointers into the old tables
        return cur[1], cur[2]

    def _new_workspace(self, device, B, T):
        n = int(load_library().matcha_b200_text_encoder_workspace_bytes(B, T, _PRECISIONS[self.precision]))
        return torch.empty(n, dtype=torch.uint8, device=device)

    # -- forward ----------------------------------------------------------------------------
    @torch.no_grad()
    def forward(self, text_input, text_lengths):
        """text_input [B,T] integer ids, text_lengths [B] -> (mu [B,80,T], log_duration [B,1,T], mask [B,1,T]), fp32."""
        if not text_input.is_cuda:
            raise RuntimeError("matcha_b200.TextEncoder needs CUDA tensors (sm_100a); there is no CPU fallback")
        lib = load_library()
        dev = text_input.device
        B, T = text_input.shape
        ids = text_input.to(torch.int64).contiguous()
        lens = text_lengths.to(dev, torch.int32).contiguous()
        if self.check_inputs:
            # nn.Embedding raises on ids outside [0, n_vocab) (text_encoder.py:377-379); the kernel clamps them instead, so a symbol
            # table that does not match n_vocab would give plausible but wrong mu / logw.  This check reads two scalars back
            # (one device synchronisation), hence a switch: check_inputs=True or MB200_CHECK_INPUTS=1.
            lo, hi, lmax = int(ids.min()), int(ids.max()), int(lens.max())
            if lo < 0 or hi >= self.vocab_size:
                raise IndexError(f"token ids span [{lo}, {hi}] but n_vocab = {self.vocab_size} (index out of range in embedding)")
            if lmax > T or int(lens.min()) < 0:
                raise ValueError(f"text_lengths up to {lmax} exceed the padded length T = {T}")
        blob, blob16 = self._blob_on(dev), self._blob16_on(dev)
        cos, sin = self._rope_on(dev, T)
        prec = _PRECISIONS[self.precision]

        def launch(ids, lens, mu, logw, mask, ws):
            with torch.cuda.device(dev):
                self.last_launches = check(lib.matcha_b200_text_encoder_forward(
                    ids.data_ptr(), lens.data_ptr(), blob.data_ptr(), blob16.data_ptr() if blob16 is not None else None,
                    self.vocab_size, cos.data_ptr(), sin.data_ptr(), mu.data_ptr(), logw.data_ptr(), mask.data_ptr(), B, T, prec,
                    ws.data_ptr(), ws.numel(), _stream_ptr(dev)), "text_encoder_forward")

        def outputs():
            return (torch.empty(B, N_FEATS, T, device=dev), torch.empty(B, 1, T, device=dev), torch.empty(B, 1, T, device=dev))

        key = (str(dev), B, T, self.precision)
        graphable = self.use_graph and not torch.cuda.is_current_stream_capturing()
        entry = None
        if graphable:
            entry = self._graphs.get(key)
            if entry is not None:
                self._graphs.move_to_end(key)
            elif self._seen.get(key, 0) < 1:           # capture a shape the second time it is seen
                if len(self._seen) > 64:
                    self._seen.clear()
                self._seen[key] = 1
                graphable = False
        if not graphable:
            wkey = (str(dev), B, T, self.precision)
            if wkey not in self._ws:
                self._ws = {wkey: self._new_workspace(dev, B, T)}
            mu, logw, mask = outputs()
            launch(ids, lens, mu, logw, mask, self._ws[wkey])
            return mu, logw, mask
        if entry is None:
            mu, logw, mask = outputs()
            st = {"ids": ids.clone(), "lens": lens.clone(), "mu": mu, "logw": logw, "mask": mask, "ws": self._new_workspace(dev, B, T)}
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                launch(st["ids"], st["lens"], st["mu"], st["logw"], st["mask"], st["ws"])     # warm-up outside the capture
            torch.cuda.current_stream(dev).wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                launch(st["ids"], st["lens"], st["mu"], st["logw"], st["mask"], st["ws"])
            entry = (graph, st)
            self._graphs[key] = entry
            while len(self._graphs) > max(1, int(self.graph_cache_size)):
                self._graphs.popitem(last=False)
        graph, st = entry
        st["ids"].copy_(ids); st["lens"].copy_(lens)
        graph.replay()
        return st["mu"].clone(), st["logw"].clone(), st["mask"].clone()
