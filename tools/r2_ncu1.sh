#!/bin/bash
cd $GRAFT_REPO_ROOT
python tools/decoder_one.py 32 800 1 fp16 2 > gpurun_out/r2_ncu1_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_decoder_fp16_launches.csv python tools/decoder_one.py 32 800 1 fp16 2 > gpurun_out/r2_ncu1.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/r2_ncu1.log; wc -l gpurun_out/r2_decoder_fp16_launches.csv
