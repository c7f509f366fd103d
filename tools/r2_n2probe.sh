#!/bin/bash
cd $GRAFT_REPO_ROOT
NCCL_DEBUG=WARN timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/n2_gather_probe.py 2>&1 | grep -v "Warn\|\*\*\*\|OMP_NUM" | tee gpurun_out/r2_n2_gather_probe.log
