"""Repeats the text-encoder forward (CUDA-graph replay, the path MatchaTTS uses) and checks that every repeat is
bit-identical to the first and that the first agrees with the fp64 torch restatement.  A programmatic-dependent-launch
ordering or visibility bug shows up as a repeat that differs (profiles/r2_te_pdl_race.log).
usage: python tools/te_pdl_stress.py [repeats]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200"))
import torch
from matcha_b200.text_encoder import TextEncoder
from oracle import text_encoder as oenc

N = int(sys.argv[1]) if len(sys.argv) > 1 else 300
dev = torch.device("cuda:0")
bad_total = 0
for prec, tol in (("fp32", 1e-5), ("fp16", 2e-3)):
    enc = TextEncoder(n_vocab=178, precision=prec).initialize_weights(0)
    w64 = oenc.cast_weights({k: v.to(dev) for k, v in enc.state_dict().items()}, torch.float64)
    for B, T in ((3, 50), (8, 128), (32, 128), (32, 400)):
        g = torch.Generator().manual_seed(B * 1000 + T)
        ids = torch.randint(0, 178, (B, T), generator=g).to(dev)
        lens = torch.tensor([T] + [int(T * (0.6 + 0.4 * float(torch.rand((), generator=g)))) for _ in range(B - 1)]).to(dev)
        first = [o.clone() for o in enc(ids, lens)]
        with torch.no_grad():
            ref = oenc.text_encoder_forward(w64, ids, lens)
        err = float(torch.linalg.norm(first[0].double() - ref[0]) / torch.linalg.norm(ref[0]))
        bad = 0
        for _ in range(N):
            out = enc(ids, lens)
            bad += int(not all(torch.equal(a, b) for a, b in zip(out, first)))
        bad_total += bad + int(err > tol)
        print(f"{prec} B={B:2d} T={T:3d}: rel-L2 vs fp64 {err:.2e} (tol {tol:g}); {bad} of {N} repeats differ from the first", flush=True)
print("TE_PDL_STRESS", "FAIL" if bad_total else "OK")
sys.exit(1 if bad_total else 0)
