"""Multi-GPU plumbing: row groups shard across ranks, only the aggregates cross NVLink.

Segments never span row groups and carry their own frame of reference / width, so the path has no data
exchange: rank r owns row groups {g : g mod world == r} (round-robin keeps Zipf-hot key ranges spread; the
reference parallelises the same way, one row group per task, row_group_collection.cpp:119-153).  The single
collective is an all-gather of per-rank {sum_lo, sum_hi, count} triples, combined exactly on the host
(an all-reduce on two u64 lanes would lose the carry between the limbs).
"""
from __future__ import annotations

import numpy as np

ROW_GROUP_SIZE = 122880


def shard_row_groups(n_row_groups: int, world: int, rank: int) -> np.ndarray:
    """Indices of the row groups rank `rank` owns (round-robin)."""
    return np.arange(rank, n_row_groups, world, dtype=np.int64)


def owner_of_row(row: np.ndarray | int, world: int):
    """Rank that owns a global row id (host-side routing of fetch_row / point look-ups)."""
    return (np.asarray(row) // ROW_GROUP_SIZE) % world


def local_row(row: np.ndarray | int, world: int):
    """Row id inside the owner's shard (its row groups are stored back to back)."""
    row = np.asarray(row)
    g = row // ROW_GROUP_SIZE
    return (g // world) * ROW_GROUP_SIZE + row % ROW_GROUP_SIZE


def agg_to_int(sum_lo: int, sum_hi: int, signed: bool) -> int:
    v = ((int(sum_hi) & 0xFFFFFFFFFFFFFFFF) << 64) | (int(sum_lo) & 0xFFFFFFFFFFFFFFFF)
    if signed and v >= 1 << 127:
        v -= 1 << 128
    return v


def combine_aggregates(gathered, n_aggs: int, signed: bool):
    """gathered: int64 array/tensor of shape [world * n_aggs * 4] ({sum_lo, sum_hi, count, reserved} per
    aggregate per rank).  Returns [(sum, count)] * n_aggs with exact 128-bit arithmetic."""
    arr = np.asarray(gathered.cpu() if hasattr(gathered, "cpu") else gathered).astype(np.int64).reshape(-1, n_aggs, 4)
    out = []
    for a in range(n_aggs):
        total, count = 0, 0
        for r in range(arr.shape[0]):
            total += agg_to_int(int(arr[r, a, 0]), int(arr[r, a, 1]), signed)
            count += int(arr[r, a, 2]) & 0xFFFFFFFFFFFFFFFF
        lo, hi = -(1 << 127), (1 << 127) - 1
        if signed:
            total = (total - lo) % (1 << 128) + lo  # HUGEINT wrap-around
        else:
            total %= 1 << 128
        assert lo <= total <= max(hi, (1 << 128) - 1)
        out.append((total, count))
    return out


def all_gather_aggregates(local, group=None):
    """torch.distributed all-gather of the per-rank aggregate tensor (NCCL on GPUs, gloo in CPU tests)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    gathered = torch.empty(world * local.numel(), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(gathered, local.contiguous(), group=group)
    return gathered
