#!/bin/bash
# per-kernel same-box A/B at the solve's sub-batch shapes (B=8): previous library vs slab tiling in both modes
cd $GRAFT_REPO_ROOT
OLD=$GRAFT_REPO_ROOT/matcha-tts-etu-upmc-ensam_b200/lib/libmatcha_b200_old.so
{
for T in 400 800; do
for v in old flat0 flat1; do
  case $v in old) export MB200_LIB=$OLD; export MB200_FLAT=1;; flat0) unset MB200_LIB; export MB200_FLAT=0;; flat1) unset MB200_LIB; export MB200_FLAT=1;; esac
  echo "== T=$T $v gemm"; GB_B=8 GB_T=$T timeout 120 python tools/gemm_bench.py | grep "K="
  echo "== T=$T $v abuild/ffn"; AB_B=8 AB_T=$T timeout 120 python tools/abuild_bench.py; FB_T=$T timeout 120 python tools/ffn_outproj_bench.py | cut -c1-40
done; done
} > gpurun_out/r2_exp15.log 2>&1
cat gpurun_out/r2_exp15.log
