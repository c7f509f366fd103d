#!/bin/bash
# MAS: the 5-CTAs-per-SM variant for <= 2 DP warps (T_x <= 192), A/B on one box
cd $GRAFT_REPO_ROOT
{
for sm in 1 0 1 0; do
  echo "== MB200_MAS_SMALL=$sm"
  MB200_MAS_SMALL=$sm timeout 300 python tools/mas_sweep.py 1024x128x800 512x128x800 1024x160x1000 16x128x800 | grep '"ragged": true'
done
MB200_MAS_SMALL=1 python - <<PY
import os, sys, ctypes
sys.path[:0] = [os.getcwd(), os.path.join(os.getcwd(), "matcha-tts-etu-upmc-ensam_b200")]
from matcha_b200._lib import load_library
lib = load_library()
for B, tx, ty in ((1024,128,800),(1024,160,1000),(1024,200,1000)):
    out = (ctypes.c_int * 8)()
    rc = lib.matcha_b200_mas_query_launch(B, tx, ty, 0, 1, out)
    print(tx, ty, "grid", out[0], "threads", out[1], "smem", out[2], "R", out[3], "nstage", out[4], "ctas/SM", out[6], "ndp", out[7])
PY
timeout 600 python -m pytest tests/test_mas_gpu.py tests/test_alignment_gpu.py -q -x 2>&1 | tail -2
} > gpurun_out/r2_exp12.log 2>&1
cat gpurun_out/r2_exp12.log
