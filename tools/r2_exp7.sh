#!/bin/bash
cd $GRAFT_REPO_ROOT
{
timeout 300 python -m pytest tests/test_ffn_fused_gpu.py -x -q 2>&1 | tail -8
timeout 600 python -m pytest tests/test_decoder_gpu.py tests/test_decoder_t800_gpu.py -x -q 2>&1 | tail -8
for o in 0 1; do echo "== solve OUTPROJ=$o"; MB200_OUTPROJ=$o timeout 120 python tools/decoder_one.py 32 800 10 fp16 4 | tail -2; done
echo "== B=64"; timeout 120 python tools/decoder_one.py 64 800 10 fp16 3 | tail -1
echo "== B=8"; timeout 120 python tools/decoder_one.py 8 800 10 fp16 3 | tail -1
} > gpurun_out/r2_exp7.log 2>&1
cat gpurun_out/r2_exp7.log
