"""CPU, world_size 2 over gloo: the multi-GPU plumbing -- row groups shard round-robin, the only collective
is the all-gather of {sum_lo, sum_hi, count} triples, combined exactly."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from duckdb_adaptive_compression_b200 import sharding
from tests.cases import width_values

N_ROWS = 5 * sharding.ROW_GROUP_SIZE + 777
N_AGGS = 3


def column(a: int) -> np.ndarray:
    return width_values(np.int64, 20 + 20 * a, N_ROWS, -5 if a == 1 else 9, seed=a)


def worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n_rg = (N_ROWS + sharding.ROW_GROUP_SIZE - 1) // sharding.ROW_GROUP_SIZE
    mine = sharding.shard_row_groups(n_rg, world, rank)
    local = torch.zeros(N_AGGS * 4, dtype=torch.int64)
    for a in range(N_AGGS):
        v = column(a)
        total, cnt = 0, 0
        for g in mine:
            part = v[g * sharding.ROW_GROUP_SIZE:(g + 1) * sharding.ROW_GROUP_SIZE]
            total += int(part.astype(object).sum())
            cnt += len(part)
        total %= 1 << 128
        lo, hi = total & (2**64 - 1), total >> 64
        local[a * 4 + 0] = lo - (1 << 64) if lo >= 1 << 63 else lo
        local[a * 4 + 1] = hi - (1 << 64) if hi >= 1 << 63 else hi
        local[a * 4 + 2] = cnt
    gathered = sharding.all_gather_aggregates(local)
    res = sharding.combine_aggregates(gathered, N_AGGS, signed=True)
    if rank == 0:
        q.put(res)
    dist.barrier()
    dist.destroy_process_group()


def test_round_robin_partition_covers_every_row_group_once():
    for world in (1, 2, 4, 8):
        seen = np.concatenate([sharding.shard_row_groups(37, world, r) for r in range(world)])
        assert sorted(seen) == list(range(37))
    rows = np.array([0, 122879, 122880, 5 * 122880 + 3])
    assert list(sharding.owner_of_row(rows, 2)) == [0, 0, 1, 1]
    assert list(sharding.local_row(rows, 2)) == [0, 122879, 0, 2 * 122880 + 3]


def test_all_gather_of_aggregates_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for a in range(N_AGGS):
        v = column(a)
        assert res[a] == (int(v.astype(object).sum()), N_ROWS)


def test_combine_handles_carries_and_negative_sums():
    # two ranks whose low limbs overflow 64 bits together, and a negative total
    g = np.array([[-1, 0, 5, 0], [-1, 0, 7, 0]], dtype=np.int64).reshape(-1)  # lo = 2^64-1 twice
    (total, count), = sharding.combine_aggregates(g, 1, signed=False)
    assert total == 2 * (2**64 - 1) and count == 12
    neg = (-12345) % (1 << 128)
    g = np.array([[neg & (2**64 - 1), neg >> 64, 1, 0]], dtype=np.uint64).astype(np.int64).reshape(-1)
    (total, count), = sharding.combine_aggregates(g, 1, signed=True)
    assert total == -12345


def test_numa_placement_helper_is_best_effort():
    """duckdb_adaptive_compression_b200/numa.py: cpulist parsing, and binding never raises (no GPU here: the dict says why)."""
    import os
    from duckdb_adaptive_compression_b200 import numa
    assert numa._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert numa._parse_cpulist("") == set()
    before = os.sched_getaffinity(0)
    info = numa.bind_to_gpu_node(0, 0, 1)
    assert isinstance(info, dict) and "applied" in info
    assert os.sched_getaffinity(0) == before or info.get("applied")
    os.sched_setaffinity(0, before)
