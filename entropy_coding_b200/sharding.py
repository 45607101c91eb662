"""Multi-GPU partitioning of a batch of substreams.

Substreams are independent (VTM EncSlice.cpp:2046-2054, 2091-2131), so a batch is split
into contiguous blocks of "units" (frames: the atomic shard for WPP, where the rows of a
frame are chained by context hand-off) with about equal estimated cost, one block per rank.
There is NO data-path collective.  The only cross-rank step is the host-side gather of
per-substream byte counts, from which every rank derives the global entry-point table
(the slice header's entry points, VTM EncGOP.cpp:3811-3824).

Host logic only (numpy + torch.distributed on CPU tensors): runs under `gloo` in the CPU
tests and next to NCCL in bench.py.
"""
from __future__ import annotations

import numpy as np


def partition_units(unit_costs, world_size: int) -> np.ndarray:
    """Contiguous partition of units into world_size blocks with balanced cost.

    Returns bounds[world_size + 1]; rank r owns units bounds[r] .. bounds[r + 1].  Boundaries are
    placed where the running cost crosses r / world_size of the total (each block within one
    unit's cost of the ideal share); blocks may be empty when there are fewer units than ranks.
    """
    c = np.asarray(unit_costs, np.float64)
    n = len(c)
    csum = np.concatenate([[0.0], np.cumsum(c)])
    total = csum[-1]
    bounds = np.zeros(world_size + 1, np.int64)
    bounds[-1] = n
    for r in range(1, world_size):
        target = total * r / world_size
        k = int(np.searchsorted(csum, target, side="left"))
        if k > 0 and abs(csum[k - 1] - target) <= abs(csum[min(k, n)] - target):
            k -= 1
        bounds[r] = min(max(k, bounds[r - 1]), n)
    return bounds


def substream_costs(tus: np.ndarray, tu_offsets: np.ndarray) -> np.ndarray:
    """Cheap cost proxy per substream: coefficient positions of its TUs."""
    area = (1 << (tus["log2_w"].astype(np.int64) + tus["log2_h"].astype(np.int64)))
    csum = np.concatenate([[0], np.cumsum(area)])
    off = np.asarray(tu_offsets, np.int64)
    return csum[off[1:]] - csum[off[:-1]]


def shard_workload(wl, rank: int, world_size: int, substreams_per_unit: int = 1):
    """This rank's share of a workload (re-based onto its own slab) and its global substream range."""
    n_units = wl.n_sub // substreams_per_unit
    assert n_units * substreams_per_unit == wl.n_sub
    cost = substream_costs(wl.tus, wl.tu_offsets).reshape(n_units, substreams_per_unit).sum(axis=1)
    b = partition_units(cost, world_size)
    lo, hi = int(b[rank]) * substreams_per_unit, int(b[rank + 1]) * substreams_per_unit
    return wl.subset(range(lo, hi)), (lo, hi)


def gather_entry_points(local_sizes, group=None):
    """Host-side gather of per-substream byte counts -> global entry points.

    Every rank passes the sizes of its own substreams (in global order within the rank) and
    receives (entry_points[n_total + 1], first): the byte offset of every substream in the
    concatenated payload and the index of this rank's first substream.  Works on any backend
    (CPU tensors for gloo; with NCCL the caller passes a device and this stays off the hot path).
    """
    import torch
    import torch.distributed as dist
    sizes = torch.as_tensor(np.asarray(local_sizes, np.int64))
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        ep = np.concatenate([[0], np.cumsum(sizes.numpy())]).astype(np.uint64)
        return ep, 0
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    counts = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(counts, torch.tensor([sizes.numel()], dtype=torch.int64, device=dev), group=group)
    counts = [int(c.item()) for c in counts]
    pad = max(counts) if counts else 0
    mine = torch.zeros(pad, dtype=torch.int64, device=dev)
    mine[:sizes.numel()] = sizes.to(dev)
    bufs = [torch.zeros(pad, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(bufs, mine, group=group)
    allsz = np.concatenate([b[:n].cpu().numpy() for b, n in zip(bufs, counts)]) if pad else np.zeros(0, np.int64)
    ep = np.concatenate([[0], np.cumsum(allsz)]).astype(np.uint64)
    return ep, int(sum(counts[:rank]))
