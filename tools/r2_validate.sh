#!/bin/bash
# round-2 validation: GPU tests, the bench line, the reference arm, the sub-batch split sweep
cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_gpu.log
tail -5 gpurun_out/r2_pytest_gpu.log
( time timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err ) 2>> gpurun_out/r2_bench.err; echo "bench rc=$?"
tail -4 gpurun_out/r2_bench.err
( time timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err ) 2>> gpurun_out/r2_bench_ref.err; echo "ref rc=$?"
tail -4 gpurun_out/r2_bench_ref.err
for s in 1 2 3 4 6 8; do echo "== MB200_SPLIT=$s"; MB200_SPLIT=$s timeout 60 python tools/decoder_one.py 32 800 10 fp16 4 | tail -1; done > gpurun_out/r2_split_sweep.log 2>&1
cat gpurun_out/r2_split_sweep.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
