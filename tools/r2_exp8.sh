#!/bin/bash
# MAS at the LJSpeech-like shapes, B=1024: rows-per-lane and ring-depth sweep (CTAs per SM follow from the shared memory per CTA)
cd $GRAFT_REPO_ROOT
{
for shape in 1024x128x800 1024x200x1000; do
for R in 0 1 3; do for NS in 0 2 3 4 6; do
  echo "== $shape R=$R NSTAGE=$NS"
  MB200_MAS_R=$R MB200_MAS_NSTAGE=$NS timeout 120 python - <<PY
import os, sys, ctypes
sys.path[:0] = [os.getcwd(), os.path.join(os.getcwd(), "matcha-tts-etu-upmc-ensam_b200")]
from matcha_b200._lib import load_library
lib = load_library()
B, tx, ty = (int(v) for v in "$shape".split("x"))
out = (ctypes.c_int * 8)()
rc = lib.matcha_b200_mas_query_launch(B, tx, ty, 0, 1, out)
print("rc", rc, "grid", out[0], "threads", out[1], "smem", out[2], "R", out[3], "nstage", out[4], "ctas/SM", out[6], "ndp", out[7])
PY
  MB200_MAS_R=$R MB200_MAS_NSTAGE=$NS timeout 120 python tools/mas_sweep.py $shape 2>&1 | grep '"ragged": true'
done; done; done
} > gpurun_out/r2_exp8.log 2>&1
grep -c ragged gpurun_out/r2_exp8.log
