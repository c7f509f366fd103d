#!/bin/bash
cd $GRAFT_REPO_ROOT
{
for rep in 1 2; do for b in 32 64 16; do for s in 2 3 4 6 8; do
  echo -n "B=$b SPLIT=$s rep=$rep: "; MB200_SPLIT=$s timeout 60 python tools/decoder_one.py $b 800 10 fp16 4 | tail -1
done; done; done
} > gpurun_out/r2_split_sweep2.log 2>&1
cat gpurun_out/r2_split_sweep2.log
