#!/bin/bash
# cheaper GELU in the 16-bit epilogues: parity suites, then solve time and the fused-FFN micro-benchmark
cd $GRAFT_REPO_ROOT
{
timeout 900 python -m pytest tests/test_ffn_fused_gpu.py tests/test_gemm_bf16_gpu.py tests/test_decoder_gpu.py tests/test_decoder_t800_gpu.py tests/test_decoder_range_gpu.py -x -q -s 2>&1 | grep -i "rel-L2\|T800\|B32\|passed\|failed\|error\|10-step" | tail -40
for b in 32 64 8; do echo "== B=$b"; timeout 120 python tools/decoder_one.py $b 800 10 fp16 4 | tail -2; done
timeout 120 python tools/ffn_outproj_bench.py | tail -4
} > gpurun_out/r2_exp9.log 2>&1
cat gpurun_out/r2_exp9.log
