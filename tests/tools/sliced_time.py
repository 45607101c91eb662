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
# decode of the same slices: lockstep (1) vs thread (2) kernel
d_pk = torch.empty_like(d_out)
d_ent = torch.zeros(wl.n_sub + 1, dtype=torch.int64, device=dev)
eng.pack_dev(d_out, d_cap, d_sizes, wl.n_sub, d_pk, d_ent)
d_dec = torch.empty_like(d_co)
d_db = torch.zeros_like(d_bins)
d_st = torch.zeros_like(d_bins)
for mode in (1, 2):
    eng.set_decoder_mode(mode)
    for it in range(2):
        d_dec.fill_(0x5a5a5a5a)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        eng.decode_residual_dev(d_pk, d_ent, d_tus, d_off, d_par, wl.n_sub, d_dec, d_db, status=d_st)
        e1.record(st)
        torch.cuda.synchronize()
        ok = bool(torch.equal(d_dec, d_co)) and int(d_st.abs().max()) == 0 and bool(torch.equal(d_db, d_bins))
        ms = e0.elapsed_time(e1)
        print(f"decode mode {mode} pass {it}: {ms:.2f} ms, {int(d_db.sum()) / ms / 1e3:.1f} Mbins/s, ok={ok}", flush=True)
