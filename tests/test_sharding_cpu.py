"""CPU tests of the N>1 path: world_size-2 gloo processes shard a batch, run the (oracle) MAS on their slice
and gather compact durations / dense tensors; the result must equal the single-process answer."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from matcha_b200 import sharding
from oracle import mas as omas
from tests.mas_cases import gen


def test_batch_slice_partitions():
    for total in (1, 5, 16, 17, 1024):
        for world in (1, 2, 3, 8):
            cuts = [sharding.batch_slice(total, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))
            sizes = [hi - lo for lo, hi in cuts]
            assert max(sizes) - min(sizes) <= 1


def test_durations_roundtrip():
    rng = np.random.default_rng(3)
    v = rng.standard_normal((5, 12, 40), dtype=np.float32)
    m = gen.ragged_mask(rng, 5, 12, 40)
    p = torch.from_numpy(omas.maximum_path_oracle(v, m)).float()
    assert torch.equal(sharding.path_from_durations(sharding.durations_from_path(p), 40), p)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    return port


def _worker(rank, world, port, total, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(11)                      # same inputs on every rank
    v = rng.standard_normal((total, 9, 31), dtype=np.float32)
    m = gen.ragged_mask(rng, total, 9, 31)
    lo, hi = sharding.batch_slice(total, rank, world)
    local = torch.from_numpy(omas.maximum_path_oracle(v[lo:hi], m[lo:hi])).float()   # stand-in for the GPU kernel
    full = sharding.all_gather_alignments(local, total)
    dense = sharding.all_gather_batch(local, total)
    want = torch.from_numpy(omas.maximum_path_oracle(v, m)).float()
    ok = torch.equal(full, want) and torch.equal(dense, want)
    with open(os.path.join(out_dir, f"rank{rank}.txt"), "w") as f:
        f.write("ok" if ok else "mismatch")
    dist.destroy_process_group()


@pytest.mark.parametrize("total", [5, 8])
def test_two_rank_gloo_gather_matches_single_process(tmp_path, total):
    mp.spawn(_worker, args=(2, _free_port(), total, str(tmp_path)), nprocs=2, join=True)
    assert all(open(tmp_path / f"rank{r}.txt").read() == "ok" for r in range(2))
