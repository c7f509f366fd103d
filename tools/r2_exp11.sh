#!/bin/bash
# MAS writer warps per CTA: 4 (default) vs 2 vs 1 -- fewer threads per CTA let more CTAs share an SM at the small shapes
cd $GRAFT_REPO_ROOT
L=$GRAFT_REPO_ROOT/matcha-tts-etu-upmc-ensam_b200/lib
{
for lib in libmatcha_b200.so libmatcha_b200_w2.so libmatcha_b200_w1.so; do
  echo "== $lib"
  MB200_LIB=$L/$lib timeout 300 python tools/mas_sweep.py 1024x128x800 1024x200x1000 1024x400x2000 16x128x800 | grep '"ragged": true'
  MB200_LIB=$L/$lib python - <<PY
import os, sys, ctypes
sys.path[:0] = [os.getcwd(), os.path.join(os.getcwd(), "matcha-tts-etu-upmc-ensam_b200")]
from matcha_b200._lib import load_library
lib = load_library()
for B, tx, ty in ((1024,128,800),(1024,200,1000),(1024,400,2000)):
    out = (ctypes.c_int * 8)()
    rc = lib.matcha_b200_mas_query_launch(B, tx, ty, 0, 1, out)
    print(tx, ty, "grid", out[0], "threads", out[1], "smem", out[2], "R", out[3], "nstage", out[4], "ctas/SM", out[6], "ndp", out[7])
PY
done
MB200_LIB=$L/libmatcha_b200_w2.so timeout 600 python -m pytest tests/test_mas_gpu.py -q -x 2>&1 | tail -2
} > gpurun_out/r2_exp11.log 2>&1
cat gpurun_out/r2_exp11.log
