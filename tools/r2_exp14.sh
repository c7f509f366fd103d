#!/bin/bash
# same-box A/B: previous commit's library (one 128-row TMA box per tile, tiles per sample) vs slab tiling
cd $GRAFT_REPO_ROOT
OLD=$GRAFT_REPO_ROOT/matcha-tts-etu-upmc-ensam_b200/lib/libmatcha_b200_old.so
{
for rep in 1 2; do for b in 32 64 8; do
  echo "== B=$b old"; MB200_LIB=$OLD timeout 120 python tools/decoder_one.py $b 800 10 fp16 4 | tail -1
  echo "== B=$b FLAT=1"; timeout 120 python tools/decoder_one.py $b 800 10 fp16 4 | tail -1
  echo "== B=$b FLAT=0"; MB200_FLAT=0 timeout 120 python tools/decoder_one.py $b 800 10 fp16 4 | tail -1
done; done
} > gpurun_out/r2_exp14.log 2>&1
cat gpurun_out/r2_exp14.log
