"""Dev tool: what the box's host side can move.  Every rank copies the bench step's end-to-end traffic
(1.05 GB host->device, 0.88 GB device->host, pinned buffers, plain cudaMemcpyAsync, no kernels) and the
ranks run concurrently; prints per-rank and aggregate GB/s.  Launch: python -m torch.distributed.run
--nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 tests/tools/pcie_ceiling.py"""
import json
import os
import time

import torch
import torch.distributed as dist

rank, lrank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(lrank)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lrank))
H2D, D2H = 1_052_000_000, 883_000_000
h_in = torch.empty(H2D, dtype=torch.uint8).pin_memory()
h_out = torch.empty(D2H, dtype=torch.uint8).pin_memory()
d_in = torch.empty(H2D, dtype=torch.uint8, device="cuda")
d_out = torch.empty(D2H, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
res = {}
for mode in ("h2d", "d2h", "both"):
    for it in range(4):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        if mode in ("h2d", "both"):
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
        if mode in ("d2h", "both"):
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    nbytes = (H2D if mode != "d2h" else 0) + (D2H if mode != "h2d" else 0)
    t = torch.tensor([dt], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res[mode] = {"ms": 1e3 * float(t), "gb_s_per_rank": nbytes / float(t) / 1e9, "gb_s_total": world * nbytes / float(t) / 1e9}
if rank == 0:
    print(json.dumps({"ranks": world, **res}), flush=True)
if world > 1:
    dist.destroy_process_group()
