#!/bin/bash
cd $GRAFT_REPO_ROOT
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; echo "bench n2 rc=$?"
grep -v Warn gpurun_out/r2_bench_n2.err | tail -5
head -c 400 gpurun_out/r2_bench_n2.json
