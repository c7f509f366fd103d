#!/bin/bash
# round-2 validation: GPU tests, the bench line, the reference arm
cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_gpu.log
tail -15 gpurun_out/r2_pytest_gpu.log
( time python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err ) 2>> gpurun_out/r2_bench.err; echo "bench rc=$?"
tail -5 gpurun_out/r2_bench.err
( time python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err ) 2>> gpurun_out/r2_bench_ref.err; echo "ref rc=$?"
tail -5 gpurun_out/r2_bench_ref.err
head -c 600 gpurun_out/r2_bench_ref.json
