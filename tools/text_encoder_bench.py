"""Times the text encoder (matcha_b200.text_encoder.TextEncoder, CUDA-graph replay) at a few shapes and compares it with
the torch restatement in eager mode on the same GPU (profiling aid).  usage: python tools/text_encoder_bench.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200"))
import torch
from matcha_b200.text_encoder import TextEncoder
from oracle import text_encoder as oenc

dev = torch.device("cuda:0")
encs = {p: TextEncoder(n_vocab=178, precision=p).initialize_weights(0) for p in ("fp32", "fp16")}
w = {k: v.to(dev) for k, v in encs["fp32"].state_dict().items()}
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
FLOP_PER_TOKEN = 2 * (3 * 5 * 192 * 192 + 192 * 192 + 6 * (4 * 192 * 192 + 2 * 3 * 192 * 768) + 80 * 192 + 3 * 192 * 256 + 3 * 256 * 256 + 256)
for B, T in ((32, 128), (64, 128), (32, 400), (8, 128), (1, 128)):
    g = torch.Generator().manual_seed(B + T)
    ids = torch.randint(0, 178, (B, T), generator=g).to(dev)
    lens = torch.tensor([T] + [int(T * (0.7 + 0.3 * float(torch.rand((), generator=g)))) for _ in range(B - 1)]).to(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    res = {}
    for prec, enc in encs.items():
        for _ in range(4):
            out = enc(ids, lens)
        torch.cuda.synchronize()
        n = 20
        e0.record()
        for _ in range(n):
            out = enc(ids, lens)
        e1.record(); torch.cuda.synchronize()
        res[prec] = (e0.elapsed_time(e1) / n, out)
    ms, out = res["fp32"]
    attn_flop = 2 * 2 * 6 * B * 2 * T * T * 96
    flop = FLOP_PER_TOKEN * B * T + attn_flop
    with torch.no_grad():
        for _ in range(2):
            ref = oenc.text_encoder_forward(w, ids, lens)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(5):
            ref = oenc.text_encoder_forward(w, ids, lens)
        e1.record(); torch.cuda.synchronize()
    ems = e0.elapsed_time(e1) / 5
    err = float(torch.linalg.norm((out[0] - ref[0]).double()) / torch.linalg.norm(ref[0].double()))
    ms16, out16 = res["fp16"]
    err16 = float(torch.linalg.norm((out16[0] - ref[0]).double()) / torch.linalg.norm(ref[0].double()))
    print(f"B={B:3d} T={T:3d}: fp32 {ms*1e3:8.1f} us {flop/ms/1e9:6.1f} TFLOP/s (rel-L2 vs eager {err:.1e}) | fp16 tensor-core {ms16*1e3:8.1f} us {flop/ms16/1e9:6.1f} TFLOP/s "
          f"(rel-L2 {err16:.1e}) | torch eager fp32, TF32 off, same GPU {ems*1e3:8.1f} us = x{ems/ms16:5.2f} of fp16 | launches {enc.last_launches}", flush=True)
