#!/bin/bash
cd $GRAFT_REPO_ROOT
{
for st in 0 1; do for s in 2 4 8; do echo "== solve SPLIT=$s STAGGER=$st"; MB200_STAGGER=$st MB200_SPLIT=$s python tools/decoder_one.py 32 800 10 fp16 4 | tail -2; done; done
MB200_STAGGER=1 python -m pytest tests/test_decoder_t800_gpu.py -x -q -k "b32 or golden" 2>&1 | tail -3
} > gpurun_out/r2_exp5.log 2>&1
cat gpurun_out/r2_exp5.log
