#!/bin/bash
# Text-encoder PDL A/B after the self-recursion fix: shipped library (trigger, then wait) vs wait-first variant (-DTE_PDL_EARLY=0).
V=$PWD/matcha-tts-etu-upmc-ensam_b200/lib_var/libmb_te_waitfirst.so
O=gpurun_out/r3_te_pdl.log
{
nvidia-smi --query-gpu=name,driver_version --format=csv,noheader
echo "== shipped (TE_PDL_EARLY=1: launch_dependents, then wait; L2-only activation loads)"
timeout 300 python tools/te_pdl_stress.py 300
timeout 200 python tools/text_encoder_bench.py
echo "== variant (TE_PDL_EARLY=0: wait, then launch_dependents)"
MB200_LIB=$V timeout 300 python tools/te_pdl_stress.py 300
MB200_LIB=$V timeout 200 python tools/text_encoder_bench.py
echo "== shipped again (order effects)"
timeout 200 python tools/text_encoder_bench.py
} > $O 2>&1
python -m pytest tests/test_text_encoder_gpu.py tests/test_matcha_tts_gpu.py -x -q -m gpu >> $O 2>&1
tail -40 $O
