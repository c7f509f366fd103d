#!/bin/bash
# slab-based M tiling (tiles may straddle samples): parity suites under both tilings, then A/B timing on one box
cd $GRAFT_REPO_ROOT
{
for f in 1 0; do
echo "== tests MB200_FLAT=$f"
MB200_FLAT=$f timeout 1200 python -m pytest tests/test_gemm_bf16_gpu.py tests/test_gemm_abuild_gpu.py tests/test_ffn_fused_gpu.py tests/test_attention_gpu.py tests/test_decoder_gpu.py tests/test_decoder_t800_gpu.py tests/test_decoder_range_gpu.py tests/test_alignment_gpu.py -q -x 2>&1 | tail -12
done
for rep in 1 2; do for b in 32 64 8; do for f in 1 0; do
  echo "== B=$b FLAT=$f"; MB200_FLAT=$f timeout 120 python tools/decoder_one.py $b 800 10 fp16 4 | tail -2
done; done; done
} > gpurun_out/r2_exp13.log 2>&1
cat gpurun_out/r2_exp13.log
