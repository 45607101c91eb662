"""Dev tool: device time of the two hot kernels with the coefficient slab in HBM vs in pinned host memory."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import entropy_coding_b200 as E
from entropy_coding_b200 import workload as W

wl = W.frames_as_ctu_substreams(64, 1920, 1080, seed0=1000)
cap = wl.out_capacity()
dev = torch.device("cuda", 0)
eng = E.CabacEngine(0)
st = torch.cuda.Stream(); torch.cuda.set_stream(st); eng.set_stream(st.cuda_stream)
B = lambda a: torch.from_numpy(np.frombuffer(a.tobytes(), np.uint8).copy()).to(dev)
d_tus, d_tuoff, d_par, d_cap = B(wl.tus), B(wl.tu_offsets), B(wl.params), B(cap)
n_sub, n_co = wl.n_sub, wl.coeffs.size
d_co = torch.from_numpy(wl.coeffs).to(dev)
h_co = eng.pinned_copy(wl.coeffs); h_dec = eng.host_array(n_co, np.int32)
d_out = torch.empty(int(cap[-1]), dtype=torch.uint8, device=dev)
d_sizes = torch.zeros(n_sub, dtype=torch.int32, device=dev); d_bins = torch.zeros_like(d_sizes)
d_packed = torch.empty_like(d_out); d_entry = torch.zeros(n_sub + 1, dtype=torch.int64, device=dev)
d_dec = torch.empty(n_co, dtype=torch.int32, device=dev)
ev = lambda: torch.cuda.Event(enable_timing=True)
def run(co_enc, co_dec, tag):
    te = td = 0
    for it in range(5):
        a, b, c, d = ev(), ev(), ev(), ev()
        a.record(st); eng.encode_residual_dev(d_tus, d_tuoff, d_par, n_sub, co_enc, d_out, d_cap, d_sizes, d_bins); b.record(st)
        eng.pack_dev(d_out, d_cap, d_sizes, n_sub, d_packed, d_entry)
        c.record(st); eng.decode_residual_dev(d_packed, d_entry, d_tus, d_tuoff, d_par, n_sub, co_dec, d_bins); d.record(st)
        torch.cuda.synchronize()
        if it >= 2: te += a.elapsed_time(b); td += c.elapsed_time(d)
    print(f"{tag}: encode {te / 3:.2f} ms  decode {td / 3:.2f} ms")
run(d_co, d_dec, "slab in HBM        ")
run(int(h_co.ctypes.data), int(h_dec.ctypes.data), "slab in pinned host ")
assert (h_dec == wl.coeffs).all()
