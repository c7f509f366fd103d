#!/bin/bash
# MAS latency experiments on the small configurations (rows-per-lane R, ring depth) with clock64 stamps
cd $GRAFT_REPO_ROOT
{
for R in 0 1 3; do
  echo "== stamps 16x128x800 MB200_MAS_R=$R"; MB200_MAS_R=$R python tools/mas_stamps.py 16 128 800
done
for R in 0 1 3 5; do
  echo "== sweep MB200_MAS_R=$R"; MB200_MAS_R=$R python tools/mas_sweep.py 16x128x800 64x128x800 128x128x800 256x128x800 1024x128x800 64x200x1000 64x400x2000 148x400x2000 2>&1 | grep -v Warn
done
for NS in 2 3 4; do
  echo "== sweep R=1 NSTAGE=$NS"; MB200_MAS_R=1 MB200_MAS_NSTAGE=$NS python tools/mas_sweep.py 16x128x800 1024x128x800 2>&1 | grep -v Warn
done
} > gpurun_out/r2_exp3.log 2>&1
tail -5 gpurun_out/r2_exp3.log
