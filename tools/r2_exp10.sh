#!/bin/bash
# A/B on one box: fitted sigmoid-form GELU (default library) against the round-1 A&S form (libmatcha_b200_gelu_as.so)
cd $GRAFT_REPO_ROOT
ALT=$GRAFT_REPO_ROOT/matcha-tts-etu-upmc-ensam_b200/lib/libmatcha_b200_gelu_as.so
{
timeout 900 python -m pytest tests/test_ffn_fused_gpu.py tests/test_gemm_bf16_gpu.py tests/test_decoder_gpu.py tests/test_decoder_t800_gpu.py tests/test_decoder_range_gpu.py -q -s 2>&1 | grep -i "rel-L2\|T800\|B32\|passed\|failed\|error\|10-step" | tail -40
for rep in 1 2; do
for b in 32 64; do
  echo "== B=$b new"; timeout 120 python tools/decoder_one.py $b 800 10 fp16 4 | tail -2
  echo "== B=$b A&S"; MB200_LIB=$ALT timeout 120 python tools/decoder_one.py $b 800 10 fp16 4 | tail -2
done; done
echo "== ffn new"; timeout 120 python tools/ffn_outproj_bench.py | tail -3
echo "== ffn A&S"; MB200_LIB=$ALT timeout 120 python tools/ffn_outproj_bench.py | tail -3
echo "== MAS ring policy"; timeout 300 python tools/mas_sweep.py 1024x128x800 1024x200x1000 1024x400x2000 512x128x800 256x128x800 | grep '"ragged": true'
} > gpurun_out/r2_exp10.log 2>&1
cat gpurun_out/r2_exp10.log
