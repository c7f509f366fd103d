#!/bin/bash
cd $GRAFT_REPO_ROOT
OLD=$GRAFT_REPO_ROOT/matcha-tts-etu-upmc-ensam_b200/lib/libmatcha_b200_old.so
run() { timeout "$@" || { echo "FAILED/TIMEOUT: $*"; cat gpurun_out/r2_exp20.log; exit 1; }; }
{
run 600 python -m pytest tests/test_decoder_gpu.py tests/test_decoder_t800_gpu.py tests/test_alignment_gpu.py tests/test_decoder_range_gpu.py -q -x 2>&1 | tail -3
MB200_SPLIT=1 run 300 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file gpurun_out/r2_exp20_warm.csv python tools/decoder_one.py 8 800 1 fp16 1 > /dev/null 2>&1
for rep in 1 2; do for b in 32 64 8; do
  echo "== B=$b old"; MB200_LIB=$OLD run 60 python tools/decoder_one.py $b 800 10 fp16 4 | tail -1
  echo "== B=$b new"; run 60 python tools/decoder_one.py $b 800 10 fp16 4 | tail -1
done; done
} > gpurun_out/r2_exp20.log 2>&1
cat gpurun_out/r2_exp20.log
