#!/bin/bash
cd $GRAFT_REPO_ROOT
{
for cfg in "32 4" "64 8" "64 4" "128 8" "16 4" "8 4"; do set -- $cfg; echo "== solve B=$1 SPLIT=$2"; MB200_SPLIT=$2 python tools/decoder_one.py $1 800 10 fp16 3 | tail -1; done
} > gpurun_out/r2_exp6.log 2>&1
cat gpurun_out/r2_exp6.log
