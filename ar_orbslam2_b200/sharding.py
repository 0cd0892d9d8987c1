"""Multi-GPU partitioning of the two workloads that shard.  One process per GPU, torch.distributed
for the plumbing:
  * extraction shards by frame -- independent units, no collective;
  * brute-force descriptor-database matching shards by keyframe (row) range with ONE exchange step: an
    all-gather of the per-shard top-2 lists (16-24 B per query and rank), merged in shard order so the result is
    identical to a single scan.
The merge itself is injected (`merge_fn`): orb_top2_merge on the device in production, the oracle in CPU tests.
"""
from __future__ import annotations

import numpy as np


def shard_range(n: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous range [begin, end) of `n` units owned by `rank` (sizes differ by at most one)."""
    base, rem = divmod(n, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def frames_of_rank(n_frames: int, rank: int, world: int) -> range:
    b, e = shard_range(n_frames, rank, world)
    return range(b, e)


def allgather_top2(local, group=None):
    """local = (best_dist i32[nq], best_idx i64[nq], second_dist i32[nq], second_idx i64[nq]) as torch tensors on the
    backend's device (CUDA for nccl, CPU for gloo).  Returns the four [world, nq] stacks, ranks in order."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    packed = torch.stack([local[0].to(torch.int64), local[1], local[2].to(torch.int64), local[3]])   # [4, nq]
    gathered = [torch.empty_like(packed) for _ in range(world)]
    dist.all_gather(gathered, packed, group=group)
    g = torch.stack(gathered)                                                                         # [world, 4, nq]
    return (g[:, 0].to(torch.int32).contiguous(), g[:, 1].contiguous(), g[:, 2].to(torch.int32).contiguous(), g[:, 3].contiguous())


def sharded_top2(local, merge_fn, group=None):
    """all-gather + merge: every rank ends with the global top-2 (== one scan over the whole database)."""
    parts = allgather_top2(local, group)
    return merge_fn(*parts)


def shard_keyframes(n_keyframes: int, rows_per_keyframe: int, rank: int, world: int) -> tuple[int, int]:
    """Row range [begin, end) of the keyframe-range shard of `rank`: whole keyframes only (BASELINE config 5)."""
    b, e = shard_range(n_keyframes, rank, world)
    return b * rows_per_keyframe, e * rows_per_keyframe


def allgather_keyframe_counts(local_counts, n_keyframes: int, group=None):
    """Per-keyframe accepted-match counts are local to a shard: concatenation in rank order is the whole answer.
    local_counts: int32 torch tensor of this rank's keyframes; shards may differ by one keyframe, so the gather is padded."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    sizes = [shard_range(n_keyframes, r, world) for r in range(world)]
    width = max(e - b for b, e in sizes)
    padded = torch.zeros(width, dtype=torch.int32, device=local_counts.device)
    padded[: local_counts.numel()] = local_counts
    gathered = torch.empty(world * width, dtype=torch.int32, device=local_counts.device)
    dist.all_gather_into_tensor(gathered, padded, group=group)
    return torch.cat([gathered[r * width: r * width + (e - b)] for r, (b, e) in enumerate(sizes)])
