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
    """[b, t_x, t_y] 0/1 alignment -> int32 frames per token [b, t_x] (what matcha_tts.py:287 sums).  Host-side
    helper for CPU tensors (the gloo tests); on the GPU the MAS kernel emits the durations itself
    (``alignment.maximum_path_with_durations``), so the dense path is never re-read."""
    return path.sum(-1).to(torch.int32)


def path_from_durations(durations, t_y):
    """Inverse of durations_from_path for a monotonic alignment (generate_path, model.py:77-114).  CUDA int32
    durations go through ``expand_kernel`` (csrc/align.cu: one pass, streaming stores); CPU tensors through torch ops."""
    if durations.is_cuda:
        from . import alignment
        d = durations if durations.dtype == torch.int32 else durations.to(torch.int32)
        return alignment.expand(d, None, None, None, T_y=t_y, want_attn=True, want_mu_y=False)["attn"]
    b, t_x = durations.shape
    ends = torch.cumsum(durations.to(torch.int64), 1)
    starts = ends - durations
    y = torch.arange(t_y, device=durations.device).view(1, 1, t_y)
    return ((y >= starts.unsqueeze(-1)) & (y < ends.unsqueeze(-1))).to(torch.float32)


def all_gather_durations(local_durations, total, group=None, out=None):
    """The alignment in its compact form on every rank: int32 frames per token [total, t_x] (1.6 KB per utterance at
    T_text = 400 instead of a dense 3.2 MB slab).  Equal slices (total % world == 0) take one
    ``all_gather_into_tensor`` (NCCL: ncclAllGather over NVLink) straight into ``out``; ragged slices are padded."""
    world = dist.get_world_size(group)
    d = local_durations.contiguous()
    if total % world == 0:
        if out is None:
            out = torch.empty((total,) + tuple(d.shape[1:]), dtype=d.dtype, device=d.device)
        dist.all_gather_into_tensor(out, d, group=group)
        return out
    return all_gather_batch(d, total, group)


def all_gather_alignments(local, total, group=None, t_y=None, dense=True):
    """Every rank receives the whole batch's alignment; only durations cross the wire.
    ``local`` is either the kernel's int32 durations [b, t_x] (then ``t_y`` must be given) or a dense 0/1 path
    [b, t_x, t_y].  With ``dense=False`` the compact durations [total, t_x] are returned as they are (what
    matcha_tts.py:287 and :314 reduce the alignment to); otherwise they are expanded to [total, t_x, t_y]."""
    if local.dim() == 3:
        t_y = local.shape[-1]
        out_dtype = local.dtype
        local = durations_from_path(local)
    else:
        assert t_y is not None, "durations input needs t_y"
        out_dtype = torch.float32
    dur = all_gather_durations(local.to(torch.int32), total, group)
    if not dense:
        return dur
    return path_from_durations(dur, t_y).to(out_dtype)
