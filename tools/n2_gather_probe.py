"""(A CUDA-graph capture of kernel + NCCL gather measured the same as plain launches -- B=1024: 1.055 vs 1.050 ms -- and made
destroy_process_group hang; removed.)
N > 1 probe: MAS kernel + compact all-gather of durations at several batch sizes, and the all-gather alone (NCCL latency curve).
Launch: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 tools/n2_gather_probe.py"""
import ctypes, datetime, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "matcha-tts-etu-upmc-ensam_b200")):
    sys.path.insert(0, p)
import numpy as np
import torch
import torch.distributed as dist
from matcha_b200._lib import load_library
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=120))
lib = load_library()
T_x, T_y = 400, 2000
st = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

def timed(fn, n=20, warm=5):
    for _ in range(warm): fn()
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / n], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())

for B in (128, 512, 1024):
    g = torch.Generator(device=dev); g.manual_seed(1000 + rank)
    value = torch.randn(B, T_x, T_y, device=dev, generator=g)
    rng = np.random.default_rng(1000 + rank)
    ty = rng.integers(int(0.7 * T_y), T_y + 1, B).astype(np.int32); ty[0] = T_y
    tx = np.minimum((ty * (T_x / T_y)).astype(np.int32), T_x).clip(1); tx[0] = T_x
    t_x, t_y = torch.from_numpy(tx).to(dev), torch.from_numpy(ty).to(dev)
    path = torch.empty(B, T_x, T_y, device=dev)
    dur = torch.empty(B, T_x, dtype=torch.int32, device=dev)
    dur_all = torch.empty(world * B, T_x, dtype=torch.int32, device=dev)
    def kernel():
        rc = lib.matcha_b200_mas_align(path.data_ptr(), 0, dur.data_ptr(), value.data_ptr(), None, t_x.data_ptr(), t_y.data_ptr(),
                                       B, T_x, T_y, ctypes.c_float(-1e9), None, st())
        assert rc == 0
    def gather():
        dist.all_gather_into_tensor(dur_all, dur)
    def both():
        kernel(); gather()
    tk, tg, tb = timed(kernel), timed(gather), timed(both)
    if rank == 0:
        print(f"B={B}: kernel {tk:.4f} ms  gather alone {tg:.4f} ms ({B * T_x * 4 / 1e6:.2f} MB/rank)  kernel+gather {tb:.4f} ms", flush=True)
    del value, path
dist.destroy_process_group()
