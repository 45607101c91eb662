#!/usr/bin/env python
"""Generates tests/golden/gen_slices.npz: legal all-intra slices from the stream generator (oracle/syntax_host.cpp:
syn_generate_slice - the CTU-level parser on an engine that picks and codes every bin), committed so that bench.py can
time cabac_decode_slices on picture-sized, dense input without touching oracle/ at run time.

TEST / BENCH INPUT, not conformance data (that is tests/golden/vtm_*.npz).  Per slice: payload bytes, the 32
cabac_slice_params words, QP, and what the generating walk recorded as ground truth: bins, CUs, TUs, coefficient
positions and a checksum sum((i + 1) * coeff[i]) mod 2^64 over the coefficients in coding order.

    python tests/golden/make_gen_slices.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
import test_syntax as T  # noqa: E402

SETS = {  # name: (width, height, slices, split_p, density)
    "hd": (1920, 1080, 2, 0.45, 0.5),      # BASELINE configs[2]'s shape: one slice per 1080p picture
    "wqvga": (416, 240, 8, 0.45, 0.5),
}


# the same picture shape with entropy_coding_sync_enabled: nine CTU-row substreams per slice (gen_slices_wpp.npz)
SETS_WPP = {"hdwpp": (1920, 1080, 2, 0.45, 0.5)}


def main():
    lib = T._host_lib()
    write(lib, SETS, T.ALL_TOOLS, "gen_slices.npz")
    write(lib, SETS_WPP, T.ALL_TOOLS | T.SL_WPP, "gen_slices_wpp.npz")


def write(lib, sets, tools, fname):
    d = {}
    for name, (w, h, n, split_p, density) in sets.items():
        sp = T.make_slice_params(w, h, 7, tools=tools)
        for k in range(n):
            g = T.generate_slice(lib, 7000 + 17 * k + len(name), sp, split_p=split_p, density=density)
            co = g["coeffs"].astype(np.int64).astype(np.uint64)
            cs = (np.arange(1, co.size + 1, dtype=np.uint64) * co).sum(dtype=np.uint64)
            key = f"{name}{k}"
            d[key + "_payload"] = g["payload"]
            d[key + "_slice"] = np.array([sp], T.SLICE_DT).view(np.uint32).copy()
            d[key + "_meta"] = np.array([g["qp"], g["init_id"], g["flags"], g["num_bins"], len(g["cus"]), len(g["tus"]),
                                         g["coeffs"].size], np.int64)
            d[key + "_checksum"] = np.uint64(cs)
            print(key, g["payload"].size, "bytes", g["num_bins"], "bins", len(g["cus"]), "CUs", len(g["tus"]), "TUs")
    d["sets"] = np.array([f"{k}:{v[2]}" for k, v in sets.items()])
    path = os.path.join(HERE, fname)
    np.savez_compressed(path, **d)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
