#!/bin/bash
cd $GRAFT_REPO_ROOT
run() { timeout "$@" || { echo "FAILED/TIMEOUT: $*"; exit 1; }; }
MB200_SPLIT=1 run 60 python tools/decoder_one.py 8 800 1 fp16 1 > /dev/null
MB200_FLAT=0 MB200_SPLIT=1 run 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_exp18_flat0.csv python tools/decoder_one.py 8 800 1 fp16 1 > /dev/null 2>&1
MB200_FLAT=0 MB200_SPLIT=1 run 300 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file gpurun_out/r2_exp18_flat0_warm.csv python tools/decoder_one.py 8 800 1 fp16 1 > /dev/null 2>&1
MB200_FLAT=1 MB200_SPLIT=1 run 300 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file gpurun_out/r2_exp18_flat1_warm.csv python tools/decoder_one.py 8 800 1 fp16 1 > /dev/null 2>&1
OLD=$GRAFT_REPO_ROOT/matcha-tts-etu-upmc-ensam_b200/lib/libmatcha_b200_old.so
MB200_LIB=$OLD MB200_SPLIT=1 run 300 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file gpurun_out/r2_exp18_old_warm.csv python tools/decoder_one.py 8 800 1 fp16 1 > /dev/null 2>&1
ls gpurun_out/r2_exp18*
