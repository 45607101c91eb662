"""Dev tool: randomised A/B of the warp-specialised encoder against the fused resolve + walk kernel
(same library, byte-for-byte), many batch shapes; looks for rare hand-off races."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import entropy_coding_b200 as E
from entropy_coding_b200 import workload as W

eng = E.CabacEngine(0)
rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 1)
budget = float(sys.argv[2]) if len(sys.argv) > 2 else 60.0
t0, runs, bins = time.time(), 0, 0
while time.time() - t0 < budget:
    n_sub = int(rng.choice([1, 7, 19, 20, 148, 500, 3000]))
    kw = dict(tus_per_sub=(1, int(rng.choice([2, 6, 30]))), ts_fraction=float(rng.choice([0.0, 0.2, 0.9])),
              bdpcm_fraction=float(rng.choice([0.0, 0.2])))
    preset = rng.choice([None, "qp0", "qp32", "qp51"])
    flags = rng.choice([None, 1, 0, 1 | 8, 16, 2, 3])
    wl = W.random_substreams(n_sub, int(rng.integers(1 << 30)), preset=None if preset is None else str(preset),
                             flags=None if flags is None else int(flags), **kw)
    cap = wl.out_capacity(bytes_per_coeff=5.0)
    eng.set_encoder_mode(1)
    a = eng.encode_residual(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, cap, want_ctx=True)
    for rep in range(3):
        eng.set_encoder_mode(3)
        b = eng.encode_residual(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, cap, want_ctx=True)
        ok = (a["sizes"] == b["sizes"]).all() and a["out"].tobytes() == b["out"].tobytes() and \
            (a["num_bins"] == b["num_bins"]).all() and (a["ctx_out"] == b["ctx_out"]).all() and \
            (a["status"] == b["status"]).all() and (a["tu_flags"] == b["tu_flags"]).all()
        if not ok:
            bad = np.nonzero(a["sizes"] != b["sizes"])[0][:5]
            print("MISMATCH", n_sub, kw, preset, flags, "first bad substreams", bad)
            sys.exit(1)
    runs += 1
    bins += int(a["num_bins"].sum())
print(f"ws_stress ok: {runs} batches x 3 repeats, {bins / 1e6:.1f} Mbins each pass, {time.time() - t0:.0f} s")
