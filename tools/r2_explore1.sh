#!/bin/bash
# exploration: warm per-kernel times at the full batch, and the solve's sensitivity to the sub-batch split / PDL
cd $GRAFT_REPO_ROOT
{
echo "== gemm_bench B=32 T=800"; GB_B=32 GB_T=800 python tools/gemm_bench.py
echo "== gemm_bench B=32 T=400"; GB_B=32 GB_T=400 python tools/gemm_bench.py
echo "== gemm_bench B=8 T=800"; GB_B=8 GB_T=800 python tools/gemm_bench.py
echo "== ffn_bench"; python tools/ffn_bench.py
echo "== abuild_bench"; python tools/abuild_bench.py
echo "== attn_bench"; python tools/attn_bench.py
for s in 1 2 4 8; do for p in 1 0; do echo "== solve SPLIT=$s PDL=$p"; MB200_SPLIT=$s MB200_PDL=$p python tools/decoder_one.py 32 800 10 fp16 4; done; done
echo "== solve fp32"; python tools/decoder_one.py 32 800 10 fp32 2
echo "== solve bf16"; python tools/decoder_one.py 32 800 10 bf16 3
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
nproc; lscpu | grep 'Model name'
} > gpurun_out/r2_explore1.log 2>&1
tail -5 gpurun_out/r2_explore1.log
