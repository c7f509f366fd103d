#!/bin/bash
cd $GRAFT_REPO_ROOT
{
echo "== stamps 16x128x800"; python tools/mas_stamps.py 16 128 800
echo "== stamps R=1"; MB200_MAS_R=1 python tools/mas_stamps.py 16 128 800
python tools/mas_sweep.py 16x128x800 64x128x800 1024x128x800 1024x400x2000 2>&1 | grep -v Warn
python -m pytest tests/test_mas_gpu.py tests/test_sharding_gpu.py -x -q 2>&1 | tail -3
} > gpurun_out/r2_exp4.log 2>&1
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; echo "bench n2 rc=$?" >> gpurun_out/r2_exp4.log
tail -3 gpurun_out/r2_bench_n2.err >> gpurun_out/r2_exp4.log
cat gpurun_out/r2_exp4.log
