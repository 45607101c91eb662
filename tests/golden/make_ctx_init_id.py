#!/usr/bin/env python3
"""tests/golden/make_ctx_init_id.py - writes tests/golden/ctx_init_id.npz from the COMPILED REFERENCE.

For the residual cases already in residual.npz (inputs are read from there), each substream coded as
a B and as a P slice: BinEncoder_Std::getNumBins(ctxId) for all 379 contexts after
CABACWriter::residual_coding over the TU list, and CABACWriter::getCtxInitId(slice)
(cabac_writer.cpp:41-81) evaluated by the reference on that coder.  Build container only
(needs oracle/_ref/libcabac_ref.so).
"""
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")
sys.path.insert(0, ROOT)
from oracle.checker import Checker  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = ("default", "signhide", "depquant_pb", "tsrc_disabled", "plain", "small_shapes")  # the non-MTS cases


def main():
    ref = Checker("ref")
    g = np.load(os.path.join(HERE, "residual.npz"))
    out = {"names": np.asarray(CASES)}
    for name in CASES:
        tus, offs, params, coeffs = (g[f"{name}_{k}"] for k in ("tus", "tu_offsets", "params", "coeffs"))
        bins, choice = [], []
        for i in range(len(params)):
            a, b = int(offs[i]), int(offs[i + 1])
            p = params[i]
            for init_id in (0, 1):
                bn, _, ch = ref.residual_ctx_bins(tus[a:b], coeffs, int(p["qp"]), init_id, int(p["flags"]))
                bins.append(bn)
                choice.append(ch)
        out[f"{name}_ctx_bins"] = np.stack(bins).reshape(len(params), 2, -1)
        out[f"{name}_choice"] = np.asarray(choice, np.uint8).reshape(len(params), 2)
        print(name, len(params), "substreams x {B, P}: choices", np.bincount(choice, minlength=2))
    np.savez_compressed(os.path.join(HERE, "ctx_init_id.npz"), **out)
    print("ctx_init_id.npz", os.path.getsize(os.path.join(HERE, "ctx_init_id.npz")), "bytes")


if __name__ == "__main__":
    main()
