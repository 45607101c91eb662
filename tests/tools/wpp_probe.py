"""Dev tool: WPP encode / decode on a tiny batch with timing and status (run with CABAC_B200_WPP_SPIN=4096)."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import entropy_coding_b200 as E
from entropy_coding_b200 import workload as W
from entropy_coding_b200.wpp import sync_records

eng = E.CabacEngine(0)
nf = int(sys.argv[1]) if len(sys.argv) > 1 else 1
wl, rows, first = W.frames_as_ctu_rows(nf, 512, 384, seed0=77)
cap = wl.out_capacity(bytes_per_coeff=5.0)
w = sync_records(wl.n_sub, rows, first)
print("rows", rows, "first", first, "ntu", np.diff(wl.tu_offsets.astype(np.int64)), flush=True)
t = time.time()
r = eng.encode_residual_wpp(wl.tus, wl.tu_offsets, wl.params, w, wl.coeffs, cap, raise_on_status=False)
print("encode", time.time() - t, r["status"], r["sizes"], flush=True)
# plain decode with no WPP (wrong contexts for rows > 0, but must terminate)
t = time.time()
d0 = eng.decode_residual(r["packed"], r["entry_points"], wl.tus, wl.tu_offsets, wl.params, wl.coeffs.size, raise_on_status=False)
print("plain decode", time.time() - t, d0["status"], flush=True)
t = time.time()
d = eng.decode_residual_wpp(r["packed"], r["entry_points"], wl.tus, wl.tu_offsets, wl.params, w, wl.coeffs.size,
                            raise_on_status=False)
print("wpp decode", time.time() - t, d["status"], (d["coeffs"] == wl.coeffs).all(), flush=True)
