"""One text-encoder forward per precision without CUDA graphs (for ncu launch lists): python tools/te_one.py [B T prec]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200"))
import torch
from matcha_b200.text_encoder import TextEncoder
B, T = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (32, 128)
prec = sys.argv[3] if len(sys.argv) > 3 else "fp16"
enc = TextEncoder(n_vocab=178, precision=prec).initialize_weights(0)
enc.use_graph = False
g = torch.Generator().manual_seed(1)
ids = torch.randint(0, 178, (B, T), generator=g).cuda()
lens = torch.tensor([T] + [int(T * (0.7 + 0.3 * float(torch.rand((), generator=g)))) for _ in range(B - 1)]).cuda()
for _ in range(3):
    out = enc(ids, lens)
torch.cuda.synchronize()
print("ok", enc.last_launches, float(out[0].abs().mean()))
