"""Dev tool: thread-per-substream decoder vs the lockstep decoder (parity + timing at N substreams)."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import entropy_coding_b200 as E
from entropy_coding_b200 import workload as W

eng = E.CabacEngine(0)
# parity on a mixed batch (TS, BDPCM, SBT, sign hiding, ISP shapes come from the test-suite; here the random generator)
wl = W.random_substreams(300, seed=5, tus_per_sub=(1, 12), ts_fraction=0.25, bdpcm_fraction=0.05, sbt_fraction=0.05)
cap = wl.out_capacity(bytes_per_coeff=5.0)
r = eng.encode_residual_packed(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, cap)
for mode in (1, 2):
    eng.set_decoder_mode(mode)
    d = eng.decode_residual(r["packed"], r["entry_points"], wl.tus, wl.tu_offsets, wl.params, wl.coeffs.size, want_ctx=True,
                            raise_on_status=False)
    print("mode", mode, "coeffs ok", bool((d["coeffs"] == wl.coeffs).all()), "bins ok", bool((d["num_bins"] == r["num_bins"]).all()),
          "status", int(d["status"].max()), "bits", int(d["bits_read"].sum()), "flags", int(d["tu_flags"].sum()),
          "ctx", int(d["ctx_out"]["state0"].astype(np.int64).sum()))
    if not (d["coeffs"] == wl.coeffs).all():
        bad = np.nonzero(d["coeffs"] != wl.coeffs)[0]
        print("  first bad coeff idx", bad[:5], d["coeffs"][bad[:5]], wl.coeffs[bad[:5]])
# timing at scale (device-resident)
tiles = int(sys.argv[1]) if len(sys.argv) > 1 else 10
base = W.sweep_substreams(10_000, 42 + 10_000)
cap1 = base.out_capacity()
dev = torch.device("cuda", 0)
st = torch.cuda.Stream(); torch.cuda.set_stream(st); eng.set_stream(st.cuda_stream)
B = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
tus, tu_off, params, capt = W.tile_descriptors(base, tiles, cap1)
d_tus, d_off, d_par, d_cap = B(tus), B(tu_off), B(params), B(capt)
d_base = torch.from_numpy(base.coeffs).to(dev)
n1 = base.n_sub
d_out = torch.empty(int(cap1[-1]), dtype=torch.uint8, device=dev)
d_sz = torch.zeros(n1, dtype=torch.int32, device=dev); d_b = torch.zeros_like(d_sz)
d_pk = torch.empty_like(d_out); d_e1 = torch.zeros(n1 + 1, dtype=torch.int64, device=dev)
eng.set_decoder_mode(0)
eng.encode_residual_dev(d_tus, d_off, d_par, n1, d_base, d_out, d_cap, d_sz, d_b)
eng.pack_dev(d_out, d_cap, d_sz, n1, d_pk, d_e1)
torch.cuda.synchronize()
b1 = int(d_e1[-1])
d_in = d_pk[:b1].repeat(tiles)
d_ent = torch.cat([(d_e1[:-1][None, :] + (torch.arange(tiles, device=dev) * b1)[:, None]).reshape(-1), torch.tensor([b1 * tiles], device=dev)])
n = n1 * tiles
d_dec = torch.empty(base.coeffs.size * tiles, dtype=torch.int32, device=dev)
d_db = torch.zeros(n, dtype=torch.int32, device=dev); d_st = torch.zeros(n, dtype=torch.int32, device=dev)
import os
modes = (int(os.environ["TDEC_MODE"]),) if "TDEC_MODE" in os.environ else (1, 2)
for mode in modes:
    eng.set_decoder_mode(mode)
    ts = []
    for it in range(int(os.environ.get("TDEC_ITERS", "3"))):
        d_dec.fill_(0x5a5a5a5a)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        eng.decode_residual_dev(d_in, d_ent, d_tus, d_off, d_par, n, d_dec, d_db, status=d_st)
        e1.record(st)
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ok = bool((d_dec.view(tiles, -1) == d_base[None, :]).all()) and int(d_st.abs().max()) == 0
    bins = int(d_db.to(torch.int64).sum())
    print(f"decoder mode {mode}: {n} substreams, {min(ts):.2f} ms, {bins / min(ts) / 1e6:.2f} Gbins/s, ok={ok}", flush=True)
