"""Multi-GPU host logic on CPU: partitioning and the world_size-2 gloo gather of substream sizes
(the only cross-rank step of the path)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from entropy_coding_b200 import sharding
from entropy_coding_b200 import workload as W


def test_partition_units_balanced_and_contiguous():
    rng = np.random.default_rng(3)
    cost = rng.integers(1, 1000, 257)
    for world in (1, 2, 4, 8):
        b = sharding.partition_units(cost, world)
        assert b[0] == 0 and b[-1] == len(cost) and (np.diff(b) >= 0).all()
        shares = [cost[b[r]:b[r + 1]].sum() for r in range(world)]
        assert max(shares) - min(shares) <= 2 * cost.max()
    b = sharding.partition_units([5, 5], 4)  # fewer units than ranks: empty blocks, nothing lost
    assert b[-1] == 2 and (np.diff(b) >= 0).all()


def test_shard_workload_covers_everything_once():
    wl = W.random_substreams(48, 9, tus_per_sub=(1, 5))
    seen = []
    for r in range(4):
        part, (lo, hi) = sharding.shard_workload(wl, r, 4, substreams_per_unit=4)
        assert (hi - lo) % 4 == 0 and part.n_sub == hi - lo
        for i in range(part.n_sub):
            a, b = int(part.tu_offsets[i]), int(part.tu_offsets[i + 1])
            ga, gb = int(wl.tu_offsets[lo + i]), int(wl.tu_offsets[lo + i + 1])
            assert b - a == gb - ga
            o, go = int(part.tus[a]["coeff_offset"]), int(wl.tus[ga]["coeff_offset"])
            n = 1 << (int(part.tus[a]["log2_w"]) + int(part.tus[a]["log2_h"]))
            assert (part.coeffs[o:o + n] == wl.coeffs[go:go + n]).all()
        seen += list(range(lo, hi))
    assert seen == list(range(wl.n_sub))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, sizes_all, bounds, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = bounds[rank], bounds[rank + 1]
    ep, first = sharding.gather_entry_points(sizes_all[lo:hi])
    q.put((rank, ep.tolist(), first))
    dist.barrier()
    dist.destroy_process_group()


def test_gather_entry_points_gloo_world2():
    rng = np.random.default_rng(11)
    sizes_all = rng.integers(1, 5000, 37).astype(np.int64)
    bounds = [0, 15, 37]  # ragged shares
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, sizes_all, bounds, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = np.concatenate([[0], np.cumsum(sizes_all)]).tolist()
    for rank, ep, first in res:
        assert ep == want and first == bounds[rank]


def test_gather_entry_points_single_process():
    ep, first = sharding.gather_entry_points([3, 4, 5])
    assert ep.tolist() == [0, 3, 7, 12] and first == 0
