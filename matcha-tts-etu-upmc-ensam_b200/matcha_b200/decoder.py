"""Drop-in for ``matcha.models.components.decoder.Decoder`` backed by the sm_100a decoder kernels.

Mirrors /root/reference/matcha/models/components/decoder.py:118-371:
  ``Decoder(in_channels=160, out_channels=80, channels=(256,256), ...).forward(x, mask, mu, t, cond=None)``
with the reference's ``state_dict`` names (SURVEY.md §8 b-names), so a Lightning checkpoint
(``decoder.estimator.*``) or a freshly built reference module can be copied over 1:1 with
``load_state_dict``.  Inference only (the reference runs it under ``torch.inference_mode``); all
arithmetic happens in ``lib/libmatcha_b200.so`` -- no CPU fallback.
"""
import ctypes
import math
from collections import OrderedDict

import torch

from ._lib import check, load_library

PREC_F32, PREC_BF16, PREC_F16 = 0, 1, 2
_PRECISIONS = {"fp32": PREC_F32, "f32": PREC_F32, "bf16": PREC_BF16, "fp16": PREC_F16, "f16": PREC_F16}

C, TIME_DIM, IN_CH, OUT_CH = 256, 1024, 160, 80
_STAGES = ("Downsampling_Blocks.0", "Downsampling_Blocks.1", "Mid_Blocks.0", "Mid_Blocks.1",
           "Upsampling_Blocks.0", "Upsampling_Blocks.1")
_STAGE_CIN = (IN_CH, C, C, C, 2 * C, 2 * C)


def _stream_ptr(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def expected_keys():
    """The reference's state_dict keys and shapes for the configured decoder."""
    s = OrderedDict()
    s["time_mlp.linear_1.weight"] = (TIME_DIM, IN_CH); s["time_mlp.linear_1.bias"] = (TIME_DIM,)
    s["time_mlp.linear_2.weight"] = (TIME_DIM, TIME_DIM); s["time_mlp.linear_2.bias"] = (TIME_DIM,)
    for p, cin in zip(_STAGES, _STAGE_CIN):
        r, t = p + ".0", p + ".1.0"
        s[r + ".mlp.1.weight"] = (C, TIME_DIM); s[r + ".mlp.1.bias"] = (C,)
        s[r + ".block1.block.0.weight"] = (C, cin, 3); s[r + ".block1.block.0.bias"] = (C,)
        s[r + ".block1.block.1.weight"] = (C,); s[r + ".block1.block.1.bias"] = (C,)
        s[r + ".block2.block.0.weight"] = (C, C, 3); s[r + ".block2.block.0.bias"] = (C,)
        s[r + ".block2.block.1.weight"] = (C,); s[r + ".block2.block.1.bias"] = (C,)
        s[r + ".res_conv.weight"] = (C, cin, 1); s[r + ".res_conv.bias"] = (C,)
        s[t + ".norm1.weight"] = (C,); s[t + ".norm1.bias"] = (C,)
        for n in ("to_q", "to_k", "to_v"):
            s[t + f".attn1.{n}.weight"] = (C, C)
        s[t + ".attn1.to_out.0.weight"] = (C, C); s[t + ".attn1.to_out.0.bias"] = (C,)
        s[t + ".norm3.weight"] = (C,); s[t + ".norm3.bias"] = (C,)
        s[t + ".ff.net.0.proj.weight"] = (4 * C, C); s[t + ".ff.net.0.proj.bias"] = (4 * C,)
        s[t + ".ff.net.2.weight"] = (C, 4 * C); s[t + ".ff.net.2.bias"] = (C,)
    s["Downsampling_Blocks.0.2.conv.weight"] = (C, C, 3); s["Downsampling_Blocks.0.2.conv.bias"] = (C,)
    s["Downsampling_Blocks.1.2.weight"] = (C, C, 3); s["Downsampling_Blocks.1.2.bias"] = (C,)
    s["Upsampling_Blocks.0.2.conv.weight"] = (C, C, 4); s["Upsampling_Blocks.0.2.conv.bias"] = (C,)
    s["Upsampling_Blocks.1.2.weight"] = (C, C, 3); s["Upsampling_Blocks.1.2.bias"] = (C,)
    s["final_conv.weight"] = (C, C, 3); s["final_conv.bias"] = (C,)
    s["final_norm.weight"] = (C,); s["final_norm.bias"] = (C,)
    s["final_proj.weight"] = (OUT_CH, C, 1); s["final_proj.bias"] = (OUT_CH,)
    return s


def _conv_taps_major(w):
    """Conv1d weight [out, cin, k] -> [out, k, cin]: GEMM K index = tap * cin + channel."""
    return w.permute(0, 2, 1).contiguous()


def pack_weights(sd):
    """Packs a reference state_dict (fp32, any device) into the flat fp32 blob the C ABI expects.
    Order must match csrc/decoder.cu::build_weights."""
    w = {k: v.detach().to("cpu", torch.float32) for k, v in sd.items()}
    parts = []
    half = IN_CH // 2
    freq = torch.log(torch.tensor(10000.0)) / (half - 1)                       # decoder.py:26
    parts.append(torch.exp(torch.arange(half, dtype=torch.float32) * -freq))   # decoder.py:27
    for n in ("time_mlp.linear_1", "time_mlp.linear_2"):
        parts += [w[n + ".weight"], w[n + ".bias"]]
    for i, p in enumerate(_STAGES):
        r, t = p + ".0", p + ".1.0"
        parts += [w[r + ".mlp.1.weight"], w[r + ".mlp.1.bias"]]
        parts += [_conv_taps_major(w[r + ".block1.block.0.weight"]), w[r + ".block1.block.0.bias"],
                  w[r + ".block1.block.1.weight"], w[r + ".block1.block.1.bias"]]
        parts += [_conv_taps_major(w[r + ".block2.block.0.weight"]), w[r + ".block2.block.0.bias"],
                  w[r + ".block2.block.1.weight"], w[r + ".block2.block.1.bias"]]
        parts += [w[r + ".res_conv.weight"][:, :, 0], w[r + ".res_conv.bias"]]
        parts += [w[t + ".norm1.weight"], w[t + ".norm1.bias"],
                  torch.cat([w[t + ".attn1.to_q.weight"], w[t + ".attn1.to_k.weight"], w[t + ".attn1.to_v.weight"]], 0)]
        parts += [w[t + ".attn1.to_out.0.weight"], w[t + ".attn1.to_out.0.bias"]]
        parts += [w[t + ".norm3.weight"], w[t + ".norm3.bias"], w[t + ".ff.net.0.proj.weight"], w[t + ".ff.net.0.proj.bias"],
                  w[t + ".ff.net.2.weight"], w[t + ".ff.net.2.bias"]]
        if i == 0:
            parts += [_conv_taps_major(w[p + ".2.conv.weight"]), w[p + ".2.conv.bias"]]
        elif i in (1, 5):
            parts += [_conv_taps_major(w[p + ".2.weight"]), w[p + ".2.bias"]]
        elif i == 4:
            wt = w[p + ".2.conv.weight"]                  # ConvTranspose1d: [cin, cout, 4]
            even = torch.stack([wt[:, :, 1].t(), wt[:, :, 3].t()], 1)      # [cout][2][cin]: taps in[s], in[s-1]
            odd = torch.stack([wt[:, :, 0].t(), wt[:, :, 2].t()], 1)       #                  taps in[s+1], in[s]
            parts += [even, odd, w[p + ".2.conv.bias"]]
    parts += [_conv_taps_major(w["final_conv.weight"]), w["final_conv.bias"], w["final_norm.weight"], w["final_norm.bias"],
              w["final_proj.weight"][:, :, 0], w["final_proj.bias"]]
    blob = torch.cat([p.contiguous().reshape(-1) for p in parts])
    n = int(load_library().matcha_b200_decoder_blob_floats())
    assert blob.numel() == n, f"packed {blob.numel()} floats, library expects {n}"
    return blob


class Decoder(torch.nn.Module):
    """Same constructor signature and ``forward`` as the reference Decoder (decoder.py:149-159, 271)."""

    def __init__(self, in_channels=IN_CH, out_channels=OUT_CH, channels=(256, 256), dropout=0.05,
                 attention_head_dim=64, n_blocks=1, num_mid_blocks=2, num_heads=4, precision="fp32"):
        super().__init__()
        if (in_channels, out_channels, tuple(channels), attention_head_dim, n_blocks, num_mid_blocks, num_heads) != \
                (IN_CH, OUT_CH, (256, 256), 64, 1, 2, 4):
            raise NotImplementedError("matcha_b200.Decoder implements the reference's configured architecture only "
                                      "(matcha_tts.py:146-153, 167-172)")
        self.in_channels, self.out_channels = in_channels, out_channels
        self.precision = precision
        self._weights = None          # reference-named fp32 tensors (CPU)
        self._blob = {}               # device -> packed fp32 blob
        self._blob16 = {}             # device -> bf16 GEMM weights (bf16 mode)
        self._ws = {}                 # (device, B, T, rows) -> workspace of the plain-launch path (one shape kept)
        self._graphs = OrderedDict()      # captured Euler solves, LRU; every entry OWNS its workspace and I/O tensors
        self._grids = {}                  # (t_k, dt_k) per host t_span, so a repeated solve does no tensor arithmetic on the host
        self._fwd_graphs = OrderedDict()  # captured single evals (forward), LRU, same ownership
        self._fwd_seen = {}           # forward() shapes seen once: captured only when they come back
        self.graph_cache_size = 4     # shapes kept per cache
        self.use_graph = True
        self.graph_forward = True     # replay forward() from a graph too (off: plain launches, e.g. for one-off shapes)

    # -- weights ----------------------------------------------------------------------------
    def initialize_weights(self, seed=0):
        """Random-init weights with the distribution the reference constructor leaves behind
        (decoder.py:255-268): Kaiming-normal (fan_in, relu gain) for every Conv1d / Linear weight, zero
        biases, GroupNorm 1/0; LayerNorm keeps its 1/0 default and the ConvTranspose1d of Upsample1D
        (decoder.py:112), which initialize_weights does not match, keeps torch's default
        U(-1/sqrt(fan_in), 1/sqrt(fan_in)) with fan_in = weight.size(1) * k.  CPU generator, so the values
        depend on `seed` only.  Used by bench.py ("random-init Matcha LJSpeech-size decoder")."""
        g = torch.Generator().manual_seed(seed)
        sd = OrderedDict()
        for name, shape in expected_keys().items():
            is_norm = ".norm" in name or name.startswith("final_norm") or ".block.1." in name
            if name.startswith("Upsampling_Blocks.0.2.conv."):
                bound = 1.0 / math.sqrt(shape[0] * 4 if len(shape) == 1 else shape[1] * shape[2])
                sd[name] = (torch.rand(shape, generator=g) * 2 - 1) * bound
            elif name.endswith(".weight") and not is_norm:
                fan_in = 1
                for d in shape[1:]:
                    fan_in *= d
                sd[name] = torch.randn(shape, generator=g) * math.sqrt(2.0 / fan_in)
            elif name.endswith(".weight"):
                sd[name] = torch.ones(shape)
            else:
                sd[name] = torch.zeros(shape)
        self.load_state_dict(sd)
        return self

    def load_state_dict(self, state_dict, strict=True):
        sd = OrderedDict()
        for k, v in state_dict.items():
            for prefix in ("decoder.estimator.", "estimator."):
                if k.startswith(prefix):
                    k = k[len(prefix):]
            sd[k] = v
        exp = expected_keys()
        missing = [k for k in exp if k not in sd]
        bad = [k for k in exp if k in sd and tuple(sd[k].shape) != exp[k]]
        if missing or bad:
            raise KeyError(f"state_dict mismatch: missing {missing[:4]}..., wrong shape {bad[:4]}...")
        self._weights = OrderedDict((k, sd[k].detach().to("cpu", torch.float32).clone()) for k in exp)
        self._blob.clear()
        self._blob16.clear()
        self._graphs.clear()          # captured graphs hold pointers into the old weight blobs
        self._fwd_graphs.clear()
        self._fwd_seen.clear()
        return self

    def state_dict(self, *a, **k):
        return OrderedDict(self._weights or {})

    def _blob_on(self, device):
        if self._weights is None:
            raise RuntimeError("matcha_b200.Decoder has no weights: call load_state_dict first")
        key = str(device)
        if key not in self._blob:
            self._blob[key] = pack_weights(self._weights).to(device)
        return self._blob[key]

    def _blob16_on(self, device):
        """bf16 GEMM weights for the tcgen05 path (None in fp32 mode)."""
        if _PRECISIONS[self.precision] == PREC_F32:
            return None
        key = str(device)
        if key not in self._blob16:
            lib = load_library()
            buf = torch.empty(int(lib.matcha_b200_decoder_blob_bf16_bytes()), dtype=torch.uint8, device=device)
            with torch.cuda.device(device):
                check(lib.matcha_b200_decoder_pack_bf16(self._blob_on(device).data_ptr(), buf.data_ptr(),
                                                        _PRECISIONS[self.precision], _stream_ptr(device)),
                      "decoder_pack_bf16")
            self._blob16[key] = buf
        return self._blob16[key]

    def _new_workspace(self, device, B, T, rows):
        n = int(load_library().matcha_b200_decoder_workspace_bytes(B, T, rows, _PRECISIONS[self.precision]))
        return torch.empty(n, dtype=torch.uint8, device=device)

    def _workspace(self, device, B, T, rows):
        """Scratch of the plain-launch path only.  Captured graphs never point here: each graph entry owns its own
        workspace (a pointer baked into a graph must outlive every replay), so replacing this one is always safe."""
        key = (str(device), B, T, rows, self.precision)
        if key not in self._ws:
            self._ws = {key: self._new_workspace(device, B, T, rows)}               # keep one shape alive
        return self._ws[key]

    def _cache_get(self, cache, key):
        entry = cache.get(key)
        if entry is not None:
            cache.move_to_end(key)
        return entry

    def _cache_put(self, cache, key, entry):
        cache[key] = entry
        while len(cache) > max(1, int(self.graph_cache_size)):
            cache.popitem(last=False)          # least recently used; its graph, workspace and I/O tensors go together

    # -- forward ----------------------------------------------------------------------------
    @torch.no_grad()
    def forward(self, x, mask, mu, t, cond=None):
        """x, mu [B,80,T]; mask [B,1,T] float; t 0-d or [B].  Returns the velocity [B,80,T] (fp32)."""
        if not x.is_cuda:
            raise RuntimeError("matcha_b200.Decoder needs CUDA tensors (sm_100a); there is no CPU fallback")
        lib = load_library()
        B, _, T = x.shape
        dev = x.device
        xf, muf = x.float().contiguous(), mu.float().contiguous()
        mf = mask.float().reshape(B, T).contiguous()
        tf = t.detach().to(dev, torch.float32).reshape(-1).contiguous()
        assert tf.numel() in (1, B)
        n_t = tf.numel()
        blob, b16 = self._blob_on(dev), self._blob16_on(dev)

        def launch(xf, muf, mf, tf, out, ws):
            with torch.cuda.device(dev):
                self.last_launches = check(lib.matcha_b200_decoder_forward(
                    xf.data_ptr(), muf.data_ptr(), mf.data_ptr(), tf.data_ptr(), n_t,
                    out.data_ptr(), blob.data_ptr(), b16.data_ptr() if b16 is not None else None,
                    ws.data_ptr(), ws.numel(), B, T, _PRECISIONS[self.precision], _stream_ptr(dev)), "decoder_forward")

        # One eval is ~46 launches on each of 4 sub-batch streams: replayed from a CUDA graph per (B, T, n_t) like
        # the solve (t is read from device memory, so one graph serves every time value).  A shape is captured the
        # second time it is seen: a one-off shape (variable-length training batches through compute_loss) costs a
        # warm-up launch + capture + fresh buffers, more than the launches a single replay saves.
        key = (str(dev), B, T, n_t, self.precision)
        graphable = self.use_graph and self.graph_forward and not torch.cuda.is_current_stream_capturing()
        entry = self._cache_get(self._fwd_graphs, key) if graphable else None
        if graphable and entry is None and self._fwd_seen.get(key, 0) < 1:
            if len(self._fwd_seen) > 64:
                self._fwd_seen.clear()
            self._fwd_seen[key] = 1
            graphable = False
        if not graphable:
            out = torch.empty_like(xf)
            launch(xf, muf, mf, tf, out, self._workspace(dev, B, T, n_t))
            return out.to(x.dtype)
        if entry is None:
            st = {"x": xf.clone(), "mu": muf.clone(), "mask": mf.clone(), "t": tf.clone(), "out": torch.empty_like(xf),
                  "ws": self._new_workspace(dev, B, T, n_t)}
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                launch(st["x"], st["mu"], st["mask"], st["t"], st["out"], st["ws"])   # warm-up outside the capture
            torch.cuda.current_stream(dev).wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                launch(st["x"], st["mu"], st["mask"], st["t"], st["out"], st["ws"])
            entry = (graph, st)
            self._cache_put(self._fwd_graphs, key, entry)
        graph, st = entry
        st["x"].copy_(xf); st["mu"].copy_(muf); st["mask"].copy_(mf); st["t"].copy_(tf)
        graph.replay()
        return st["out"].to(x.dtype, copy=True)

    @torch.no_grad()
    def loss_forward(self, x1, z, t, mu, mask, sigma_min, want_pred=False):
        """compute_loss's device work in ONE C call (flow_matching.py:130-149): phi_t, the decoder eval with per-sample t,
        and the two sums of the masked MSE.  Returns (sums[2] fp64 on the device = [sum sq. error, sum(mask)], phi_t,
        u_pred or None).  In the 16-bit modes phi_t is formed by the input-packing kernel and the squared error by the final
        projection's epilogue: u_pred is only written when ``want_pred`` asks for it."""
        if not x1.is_cuda:
            raise RuntimeError("matcha_b200.Decoder needs CUDA tensors (sm_100a); there is no CPU fallback")
        lib = load_library()
        B, _, T = x1.shape
        dev = x1.device
        x1f, zf, muf = x1.float().contiguous(), z.float().contiguous(), mu.float().contiguous()
        mf = mask.float().reshape(B, T).contiguous()
        tf = t.detach().to(dev, torch.float32).reshape(-1).contiguous()
        assert tf.numel() == B
        blob, b16 = self._blob_on(dev), self._blob16_on(dev)
        phi = torch.empty_like(x1f)
        pred = torch.empty_like(x1f) if (want_pred or self.precision == "fp32") else None
        sums = torch.empty(2, dtype=torch.float64, device=dev)
        ws = self._workspace(dev, B, T, B)
        with torch.cuda.device(dev):
            self.last_launches = check(lib.matcha_b200_cfm_loss_forward(
                x1f.data_ptr(), zf.data_ptr(), tf.data_ptr(), muf.data_ptr(), mf.data_ptr(), float(sigma_min), phi.data_ptr(),
                pred.data_ptr() if pred is not None else None, sums.data_ptr(), blob.data_ptr(),
                b16.data_ptr() if b16 is not None else None, ws.data_ptr(), ws.numel(), B, T, _PRECISIONS[self.precision],
                _stream_ptr(dev)), "cfm_loss_forward")
        return sums, phi, pred

    @torch.no_grad()
    def solve_euler(self, x, t_span, mu, mask, use_graph=None):
        """The whole Euler loop of flow_matching.py:67-104 in one C call.  With ``use_graph`` (default: on for
        CUDA inputs) the ~95 kernel launches per step are captured once per (B, T, n_steps) into a CUDA graph
        and replayed, which removes the launch gaps between the small kernels."""
        lib = load_library()
        B, _, T = x.shape
        dev = x.device
        # a host t_span (what CFM.forward passes) costs nothing here; a device one is read back, which synchronises the device
        tkey = tuple(t_span.tolist()) if not t_span.is_cuda else None
        grid = self._grids.get(tkey) if tkey is not None else None
        if grid is None:
            grid = euler_time_grid(t_span.detach().to("cpu", torch.float32))
            if tkey is not None and len(self._grids) < 64:
                self._grids[tkey] = grid
        ts, dts = grid
        n = len(ts)
        if use_graph is None:
            use_graph = self.use_graph
        blob, b16 = self._blob_on(dev), self._blob16_on(dev)
        t_arr = (ctypes.c_float * n)(*[float(v) for v in ts])
        dt_arr = (ctypes.c_float * n)(*[float(v) for v in dts])

        def launch(xf, muf, mf, out, ws):
            with torch.cuda.device(dev):
                self.last_launches = check(lib.matcha_b200_cfm_solve_euler(
                    xf.data_ptr(), muf.data_ptr(), mf.data_ptr(), out.data_ptr(), t_arr, dt_arr, n, blob.data_ptr(),
                    b16.data_ptr() if b16 is not None else None, ws.data_ptr(), ws.numel(), B, T,
                    _PRECISIONS[self.precision], _stream_ptr(dev)), "cfm_solve_euler")

        if not use_graph:
            xf, muf = x.float().contiguous(), mu.float().contiguous()
            mf = mask.float().reshape(B, T).contiguous()
            out = torch.empty_like(xf)
            launch(xf, muf, mf, out, self._workspace(dev, B, T, n))
            return out.to(x.dtype)

        key = (str(dev), B, T, tuple(float(v) for v in ts), tuple(float(v) for v in dts), self.precision)
        entry = self._cache_get(self._graphs, key)
        if entry is None:
            st = {"x": torch.empty(B, 80, T, device=dev), "mu": torch.empty(B, 80, T, device=dev),
                  "mask": torch.empty(B, T, device=dev), "out": torch.empty(B, 80, T, device=dev),
                  "ws": self._new_workspace(dev, B, T, n)}            # owned by the graph entry (see _workspace)
            st["x"].copy_(x); st["mu"].copy_(mu); st["mask"].copy_(mask.reshape(B, T))
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                launch(st["x"], st["mu"], st["mask"], st["out"], st["ws"])      # warm-up outside the capture (sets attributes)
            torch.cuda.current_stream(dev).wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                launch(st["x"], st["mu"], st["mask"], st["out"], st["ws"])
            entry = (graph, st)
            self._cache_put(self._graphs, key, entry)
        graph, st = entry
        st["x"].copy_(x); st["mu"].copy_(mu); st["mask"].copy_(mask.reshape(B, T))
        graph.replay()
        return st["out"].to(x.dtype, copy=True)


def euler_time_grid(t_span):
    """(t_k, dt_k) exactly as the reference loop produces them in fp32 (flow_matching.py:85-102)."""
    ts, dts = [], []
    t = t_span[0]
    dt = t_span[1] - t_span[0]
    n = len(t_span) - 1
    for step in range(1, n + 1):
        ts.append(t.clone()); dts.append(dt.clone())
        t = t + dt
        if step < n:
            dt = t_span[step + 1] - t
    return ts, dts
