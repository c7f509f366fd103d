#!/bin/bash
cd $GRAFT_REPO_ROOT
OLD=$GRAFT_REPO_ROOT/matcha-tts-etu-upmc-ensam_b200/lib/libmatcha_b200_old.so
run() { timeout "$@" || { echo "FAILED/TIMEOUT: $*"; cat gpurun_out/r2_exp17.log; exit 1; }; }
{
for T in 400 800; do
echo "== old abuild T=$T"; MB200_LIB=$OLD AB_B=8 AB_T=$T run 60 python tools/abuild_bench.py
echo "== flat1 abuild T=$T"; AB_B=8 AB_T=$T run 60 python tools/abuild_bench.py
done
run 300 python -m pytest tests/test_gemm_abuild_gpu.py tests/test_decoder_gpu.py -q -x 2>&1 | tail -2
for rep in 1 2; do for b in 32 8; do
  echo "== B=$b old"; MB200_LIB=$OLD run 60 python tools/decoder_one.py $b 800 10 fp16 4 | tail -1
  echo "== B=$b FLAT=1"; run 60 python tools/decoder_one.py $b 800 10 fp16 4 | tail -1
done; done
} > gpurun_out/r2_exp17.log 2>&1
cat gpurun_out/r2_exp17.log
