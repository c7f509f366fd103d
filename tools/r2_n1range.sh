#!/bin/bash
# fp16 range guard: range tests, decoder suites, and the solve time beside the previous build's
cd $GRAFT_REPO_ROOT
python -m pytest tests/test_decoder_range_gpu.py -m gpu -q -s 2>&1 | grep -v "^$" | tail -60 > gpurun_out/r2_n1range.log
tail -40 gpurun_out/r2_n1range.log
