#!/bin/bash
cd $GRAFT_REPO_ROOT
python -m pytest tests/test_alignment_gpu.py tests/test_decoder_gpu.py -m gpu -q -x 2>&1 | tail -25 > gpurun_out/r2_f2.log
tail -25 gpurun_out/r2_f2.log
