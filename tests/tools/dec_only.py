"""Dev tool: the residual decoder alone on the bench batch (A/B timing)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import entropy_coding_b200 as E
from entropy_coding_b200 import workload as W

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 64
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
wl = W.frames_as_ctu_substreams(frames, 1920, 1080, seed0=1000)
cap = wl.out_capacity()
dev = torch.device("cuda", 0)
eng = E.CabacEngine(0)
st = torch.cuda.Stream(); torch.cuda.set_stream(st); eng.set_stream(st.cuda_stream)
B = lambda a: torch.from_numpy(np.frombuffer(a.tobytes(), np.uint8).copy()).to(dev)
d_tus, d_tuoff, d_par, d_cap = B(wl.tus), B(wl.tu_offsets), B(wl.params), B(cap)
n_sub, n_co = wl.n_sub, wl.coeffs.size
d_co = torch.from_numpy(wl.coeffs).to(dev)
d_out = torch.empty(int(cap[-1]), dtype=torch.uint8, device=dev)
d_sizes = torch.zeros(n_sub, dtype=torch.int32, device=dev); d_bins = torch.zeros_like(d_sizes)
d_packed = torch.empty_like(d_out); d_entry = torch.zeros(n_sub + 1, dtype=torch.int64, device=dev)
d_dec = torch.empty(n_co, dtype=torch.int32, device=dev)
eng.encode_residual_dev(d_tus, d_tuoff, d_par, n_sub, d_co, d_out, d_cap, d_sizes, d_bins)
eng.pack_dev(d_out, d_cap, d_sizes, n_sub, d_packed, d_entry)
ts = []
for it in range(reps):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st); eng.decode_residual_dev(d_packed, d_entry, d_tus, d_tuoff, d_par, n_sub, d_dec, d_bins); b.record(st)
    torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
ok = bool((d_dec == d_co).all().item())
print(f"{n_sub} substreams: decode best {min(ts):.2f} ms, all {[round(x, 2) for x in ts]}, round trip ok={ok}")
