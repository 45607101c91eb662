"""Dev tool: the sliced encoder (CABAC_ENCODER_SLICED) against the fused kernel on a small batch."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import entropy_coding_b200 as E
from entropy_coding_b200 import workload as W

eng = E.CabacEngine(0)
n_sub = int(sys.argv[1]) if len(sys.argv) > 1 else 3
tus = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (5, 40)
wl = W.random_substreams(n_sub, seed=11, tus_per_sub=tus, ts_fraction=0.2)
cap = wl.out_capacity(bytes_per_coeff=5.0)
eng.set_encoder_mode(1)
a = eng.encode_residual(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, cap, want_ctx=True)
eng.set_encoder_mode(4)
t = time.time()
b = eng.encode_residual(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, cap, want_ctx=True, raise_on_status=False)
print("sliced call", time.time() - t, "status", b["status"][:8])
print("sizes", a["sizes"][:8], b["sizes"][:8])
print("bins ", a["num_bins"][:8], b["num_bins"][:8])
ok = True
for i in range(n_sub):
    x = a["out"][int(cap[i]):int(cap[i]) + int(a["sizes"][i])]
    y = b["out"][int(cap[i]):int(cap[i]) + int(b["sizes"][i])]
    if x.tobytes() != y.tobytes():
        ok = False
        k = int(np.argmax(x[:min(len(x), len(y))] != y[:min(len(x), len(y))])) if len(x) and len(y) else -1
        print(f"substream {i}: differ, sizes {len(x)} {len(y)}, first diff at byte {k}: {x[max(0,k-2):k+6].tolist()} vs {y[max(0,k-2):k+6].tolist()}")
        if i > 4:
            break
print("bytes equal:", ok, "| ctx equal:", bool((a["ctx_out"] == b["ctx_out"]).all()), "| flags equal:", bool((a["tu_flags"] == b["tu_flags"]).all()))
