"""Dev tool: the literal configs[2] batch (N 1080p frames as N slices) through the sliced encoder, a few passes;
run under `ncu --metrics gpu__time_duration.sum` for the per-kernel split."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import entropy_coding_b200 as E
from entropy_coding_b200 import workload as W

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 16
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
wl0 = W.frames_as_ctu_substreams(frames, 1920, 1080, seed0=1000)
wl = W.merge_substreams(wl0, wl0.n_sub // frames)
dev = torch.device("cuda", 0)
eng = E.CabacEngine(0)
st = torch.cuda.Stream()
torch.cuda.set_stream(st)
eng.set_stream(st.cuda_stream)
B = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
cap = wl.out_capacity()
d_tus, d_off, d_par, d_cap = B(wl.tus), B(wl.tu_offsets), B(wl.params), B(cap)
d_co = torch.from_numpy(wl.coeffs).to(dev)
d_out = torch.empty(int(cap[-1]), dtype=torch.uint8, device=dev)
d_sizes = torch.zeros(wl.n_sub, dtype=torch.int32, device=dev)
d_bins = torch.zeros_like(d_sizes)
for it in range(reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    eng.encode_residual_dev(d_tus, d_off, d_par, wl.n_sub, d_co, d_out, d_cap, d_sizes, d_bins)
    e1.record(st)
    torch.cuda.synchronize()
    print(f"pass {it}: {e0.elapsed_time(e1):.2f} ms, {int(d_bins.sum())} bins, {int(d_bins.sum()) / e0.elapsed_time(e1) / 1e6:.2f} Gbins/s", flush=True)
