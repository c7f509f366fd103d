#!/bin/bash
cd $GRAFT_REPO_ROOT
{
python -m pytest tests/test_gemm_bf16_gpu.py tests/test_decoder_gpu.py tests/test_decoder_t800_gpu.py -x -q 2>&1 | tail -15
for s in 1 2 4 8; do echo "== solve SPLIT=$s ABUILD=1"; MB200_SPLIT=$s python tools/decoder_one.py 32 800 10 fp16 4 | tail -2; done
for s in 1 4; do echo "== solve SPLIT=$s ABUILD=0"; MB200_SPLIT=$s MB200_ABUILD=0 python tools/decoder_one.py 32 800 10 fp16 4 | tail -2; done
echo "== gemm_bench B=32 T=800"; GB_B=32 GB_T=800 python tools/gemm_bench.py 2>&1 | grep -v stamps
} > gpurun_out/r2_exp2.log 2>&1
tail -40 gpurun_out/r2_exp2.log
