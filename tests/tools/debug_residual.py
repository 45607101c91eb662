"""Dev tool (test infrastructure): find the first TU whose GPU encoding differs from the oracle."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import entropy_coding_b200 as E
from entropy_coding_b200 import workload as W
from entropy_coding_b200.types import TU_DT, PARAMS_DT
from oracle.checker import Checker

def main():
    seed = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    nsub = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    eng = E.CabacEngine(0)
    chk = Checker("oracle")
    wl = W.random_substreams(nsub, seed=seed, tus_per_sub=(2, 8))
    # every TU as its own substream
    ntu = len(wl.tus)
    sub_of = np.searchsorted(wl.tu_offsets, np.arange(ntu), side="right") - 1
    params = wl.params[sub_of].copy()
    offs = np.arange(ntu + 1, dtype=np.uint64)
    w1 = W.Workload(wl.tus, offs, params, wl.coeffs)
    cap = w1.out_capacity(bytes_per_coeff=5.0)
    r = eng.encode_residual(w1.tus, w1.tu_offsets, w1.params, w1.coeffs, cap, raise_on_status=False)
    bad = 0
    for t in range(ntu):
        p = params[t]
        eb, nb, _ = chk.encode_residual(wl.tus[t:t + 1], wl.coeffs, int(p["qp"]), int(p["init_id"]), int(p["flags"]))
        got = r["out"][int(cap[t]):int(cap[t]) + int(r["sizes"][t])]
        if got.tobytes() != eb.tobytes() or int(r["num_bins"][t]) != nb:
            d = wl.tus[t]
            w, h = 1 << int(d["log2_w"]), 1 << int(d["log2_h"])
            print(f"TU {t}: {w}x{h} comp {d['comp_id']} flags {d['flags']} qp {p['qp']} subflags {p['flags']} "
                  f"bins gpu {r['num_bins'][t]} oracle {nb} bytes gpu {len(got)} oracle {len(eb)} status {r['status'][t]}")
            if bad < 3 and w * h <= 256:
                blk = wl.coeffs[int(d["coeff_offset"]):int(d["coeff_offset"]) + w * h].reshape(h, w)
                print(blk)
                print("gpu   ", got.tobytes().hex())
                print("oracle", eb.tobytes().hex())
            bad += 1
    print(f"{bad} of {ntu} TUs differ")

main()
