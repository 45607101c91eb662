"""Dev tool: the residual encoder alone on the bench batch (for ncu captures and A/B timing)."""
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
d_co = torch.from_numpy(wl.coeffs).to(dev)
d_out = torch.empty(int(cap[-1]), dtype=torch.uint8, device=dev)
d_sizes = torch.zeros(wl.n_sub, dtype=torch.int32, device=dev); d_bins = torch.zeros_like(d_sizes)
ts = []
for it in range(reps):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st); eng.encode_residual_dev(d_tus, d_tuoff, d_par, wl.n_sub, d_co, d_out, d_cap, d_sizes, d_bins); b.record(st)
    torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
bins = int(d_bins.sum().item())
t = min(ts)
print(f"{wl.n_sub} substreams, {bins / 1e6:.1f} Mbins: encode best {t:.2f} ms ({bins / t / 1e6:.2f} Gbins/s), all {[round(x, 2) for x in ts]}, "
      f"bytes {int(d_sizes.sum().item())}")
