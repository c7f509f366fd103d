"""Batch sharding of the two hot paths across the GPUs of one box.

Utterances are independent in both paths (core.pyx:121 prange; GroupNorm / LayerNorm / attention are
per sample), so rank r owns a contiguous batch slice and runs the kernels on it without any
data-path collective.  Callers that need every rank to see the whole batch gather afterwards:
alignments as compact per-token durations (1.6 KB per utterance instead of a dense 3.2 MB
[T_x, T_y] slab), mels densely.  One process per GPU, ``torch.distributed`` (NCCL on GPUs, gloo in
the CPU tests) is the only plumbing.
"""
import torch
import torch.distributed as dist


def batch_slice(total, rank, world):
    """Contiguous slice [lo, hi) of a batch of ``total`` items owned by ``rank`` (sizes differ by <= 1)."""
    base, rem = divmod(total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def _all_gather_ragged(local, total, dim0_sizes, group=None):
    world = dist.get_world_size(group)
    mx = max(dim0_sizes)
    pad = torch.zeros((mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad, group=group)
    return torch.cat([b[:n] for b, n in zip(bufs, dim0_sizes)], 0)


def all_gather_batch(local, total, group=None):
    """Gathers per-rank batch slices (made with batch_slice) back into the full batch on every rank."""
    world = dist.get_world_size(group)
    sizes = [batch_slice(total, r, world)[1] - batch_slice(total, r, world)[0] for r in range(world)]
    assert local.shape[0] == sizes[dist.get_rank(group)]
    return _all_gather_ragged(local.contiguous(), total, sizes, group)


def durations_from_path(path):
    """[b, t_x, t_y] 0/1 alignment -> int32 frames per token [b, t_x] (what matcha_tts.py:287 sums)."""
    return path.sum(-1).to(torch.int32)


def path_from_durations(durations, t_y):
    """Inverse of durations_from_path for a monotonic alignment (generate_path, model.py:77-114)."""
    b, t_x = durations.shape
    ends = torch.cumsum(durations.to(torch.int64), 1)
    starts = ends - durations
    y = torch.arange(t_y, device=durations.device).view(1, 1, t_y)
    return ((y >= starts.unsqueeze(-1)) & (y < ends.unsqueeze(-1))).to(torch.float32)


def all_gather_alignments(local_path, total, group=None):
    """Every rank receives the full [total, t_x, t_y] alignment; only durations cross the wire."""
    t_y = local_path.shape[-1]
    dur = all_gather_batch(durations_from_path(local_path), total, group)
    return path_from_durations(dur, t_y).to(local_path.dtype)
