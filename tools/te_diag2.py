"""Stage-wise diagnosis of the text-encoder forward against the torch restatement (reads the workspace buffers)."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200"))
import math
import torch, torch.nn.functional as F
from matcha_b200._lib import load_library
from matcha_b200 import text_encoder as te
from oracle import text_encoder as oenc
lib = load_library()
dev = torch.device("cuda:0")
def rel(a, b): return float(torch.linalg.norm(a.double() - b.double()) / torch.linalg.norm(b.double()))
NV = 178
w = oenc.make_weights(23, "rich", NV)
wd = oenc.cast_weights({k: v.to(dev) for k, v in w.items()}, torch.float64)
B, T = int(sys.argv[1]), int(sys.argv[2])
PREC = int(sys.argv[3]) if len(sys.argv) > 3 else 0          # 0 fp32, 2 fp16
g = torch.Generator().manual_seed(5)
ids = torch.randint(0, NV, (B, T), generator=g).to(dev)
lens = torch.tensor([T] + [max(1, int(T * 0.7))] * (B - 1), dtype=torch.int32).to(dev)
blob = te.pack_weights(w, NV).to(dev)
blob16 = blob.half()
cos, sin = te.rope_tables(max(T, 512)); cos, sin = cos.to(dev), sin.to(dev)
M = B * T
def r64(n): return (n + 63) // 64 * 64
names = [("x0", 192), ("x", 192), ("y", 192), ("qkv", 576)]
for nl in (0, 1, 2, 6):
    os.environ["MB200_TE_LAYERS"] = str(nl)
    ws = torch.zeros(int(lib.matcha_b200_text_encoder_workspace_bytes(B, T, PREC)) // 4, device=dev)
    mu, logw, mask = torch.empty(B, 80, T, device=dev), torch.empty(B, 1, T, device=dev), torch.empty(B, 1, T, device=dev)
    rc = lib.matcha_b200_text_encoder_forward(ids.data_ptr(), lens.data_ptr(), blob.data_ptr(), blob16.data_ptr(), NV, cos.data_ptr(), sin.data_ptr(),
                                              mu.data_ptr(), logw.data_ptr(), mask.data_ptr(), B, T, PREC, ws.data_ptr(), ws.numel() * 4, None)
    torch.cuda.synchronize()
    buf, off = {}, 0
    for n, c in names:
        buf[n] = ws[off:off + M * c].view(B, T, c); off += r64(M * c)
    # oracle, truncated at nl layers
    x = F.embedding(ids, wd["embedding.weight"]) * math.sqrt(192)
    x = x.transpose(1, 2)
    xm = oenc.sequence_mask(lens.long(), T).unsqueeze(1).double()
    e_x0 = rel(buf["x0"], x.transpose(1, 2))
    x = oenc.conv_relu_norm(wd, x, xm)
    am = xm.unsqueeze(2) * xm.unsqueeze(-1)
    for i in range(nl):
        x = x * xm
        a = oenc.attention(wd, f"encoder.attention_layers.{i}", x, am)
        x = oenc._ln(x + a, wd, f"encoder.norm_layers_1.{i}")
        f = oenc.ffn(wd, f"encoder.ffn_layers.{i}", x, xm)
        x = oenc._ln(x + f, wd, f"encoder.norm_layers_2.{i}")
    x = x * xm
    print(f"PDL={os.environ.get('MB200_PDL','1')} B={B} T={T} layers={nl}: rc={rc} x0 {e_x0:.1e}  x {rel(buf['x'], x.transpose(1,2)):.2e}", flush=True)
