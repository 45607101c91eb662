#!/usr/bin/env python3
"""tests/golden/make_golden.py - writes tests/golden/*.npz from the COMPILED REFERENCE.

Run in the build container only (needs oracle/_ref/libcabac_ref.so, i.e.
/root/reference).  Every expected output in the fixtures comes from the
unmodified reference sources (BinEncoder_Std / BinDecoder_Std / CABACWriter /
CABACReader) through oracle/ref_driver.cpp; inputs are stored alongside so the
fixtures do not depend on numpy's random streams.
"""
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")
sys.path.insert(0, ROOT)
from entropy_coding_b200 import workload as W  # noqa: E402
from entropy_coding_b200.types import (SUB_ALIGN, SUB_DEP_QUANT, SUB_SIGN_HIDING, SUB_TERMINATE,  # noqa: E402
                                       SUB_TSRC_DISABLED, mk_align, mk_ctx, mk_ep, mk_rem, mk_trm)
from oracle.checker import Checker  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def random_cmds(rng, n, with_align):
    cmds = []
    p1 = rng.random()
    for _ in range(n):
        k = rng.random()
        if k < 0.6:
            cmds.append(mk_ctx(int(rng.integers(0, 379)), int(rng.random() < p1)))
        elif k < 0.8:
            nb = int(rng.integers(0, 25))
            cmds.append(mk_ep(int(rng.integers(0, 1 << nb)) if nb else 0, nb))
        elif k < 0.93:
            rice = int(rng.integers(0, 5))
            vmax = ((4100 + (1 << (15 - rice))) << rice) - 1  # largest in-contract remainder
            cmds.append(mk_rem(min(int(rng.integers(0, 1 << int(rng.integers(1, 18)))), vmax), rice))
        elif k < 0.97 or not with_align:
            cmds.append(mk_trm(0))
        else:
            cmds.append(mk_align())
    return np.asarray(cmds, np.uint32)


def main():
    ref = Checker("ref")
    T = SUB_TERMINATE | SUB_ALIGN

    # ---- known-answer vectors -----------------------
    kats = {}
    kats["kat0"] = (np.zeros(0, np.uint32), 32, 2)
    kats["kat1"] = (np.asarray([mk_ctx(90 + i % 12, i & 1) for i in range(64)] + [mk_ep(0x5A5, 11)], np.uint32), 32, 2)
    c = []
    for i in range(200):
        c.append(mk_ctx(150 + i % 21, int((i * 7 % 5) == 0)))
        if i % 10 == 9:
            c.append(mk_rem(i, i % 4))
    kats["kat2"] = (np.asarray(c, np.uint32), 0, 2)
    kats["kat3"] = (np.asarray([mk_ctx(246, 1)] * 300 + [mk_ep(1, 1)] * 40 + [mk_trm(0)], np.uint32), 51, 0)
    out = {}
    for name, (cmds, qp, iid) in kats.items():
        b, nb = ref.encode_bins(cmds, qp, iid, T)
        out[name + "_cmds"], out[name + "_bytes"] = cmds, b
        out[name + "_meta"] = np.asarray([qp, iid, nb], np.int64)
        print(name, b.tobytes().hex(), nb)
    np.savez_compressed(os.path.join(HERE, "kat.npz"), **out)

    # ---- context initialisation --------------------------------------------
    out = {}
    for qp in (-5, 0, 9, 17, 26, 32, 40, 51, 63, 70):
        for iid in range(3):
            st, rates = ref.ctx_init(qp, iid)
            out[f"q{qp}_i{iid}"] = np.stack([st["state0"], st["state1"]], axis=1)
    out["rates"] = rates
    for r in range(4):
        out[f"table{r}"] = ref.init_table(r)
    out["ctx_offsets"] = ref.ctx_offsets()
    np.savez_compressed(os.path.join(HERE, "ctx_init.npz"), **out)

    # ---- scan orders ---------------------------------------------------------
    out = {}
    for lw in range(1, 7):
        for lh in range(1, 7):
            tab, cw, ch = ref.scan_table(lw, lh)
            out[f"s{lw}_{lh}"] = tab.astype(np.uint16)
            out[f"cg{lw}_{lh}"] = np.asarray([cw, ch], np.uint8)
    np.savez_compressed(os.path.join(HERE, "scan.npz"), **out)

    # ---- random bin-level streams ----------------------------------------------
    rng = np.random.default_rng(20260101)
    out = {}
    n_streams = 48
    for i in range(n_streams):
        with_align = i % 4 == 0
        cmds = random_cmds(rng, int(rng.integers(0, 600)), with_align)
        qp, iid = int(rng.integers(0, 64)), int(rng.integers(0, 3))
        b, nb, ctx = ref.encode_bins(cmds, qp, iid, T, want_ctx=True)
        vals, bits, st = ref.decode_bins(b, cmds, qp, iid)
        out[f"cmds{i}"], out[f"bytes{i}"], out[f"vals{i}"] = cmds, b, vals
        out[f"ctx{i}"] = np.stack([ctx["state0"], ctx["state1"]], axis=1)
        out[f"meta{i}"] = np.asarray([qp, iid, nb, bits, st], np.int64)
    out["n"] = np.asarray([n_streams])
    np.savez_compressed(os.path.join(HERE, "bins.npz"), **out)

    # ---- residual-level streams ----------------------------------------------
    cases = [
        ("default", dict(seed=11), None, False),
        ("signhide", dict(seed=12, flags=T | SUB_SIGN_HIDING), None, False),
        ("depquant_pb", dict(seed=13, flags=T | SUB_DEP_QUANT), None, False),
        ("tsrc_disabled", dict(seed=14, flags=T | SUB_DEP_QUANT | SUB_TSRC_DISABLED, ts_fraction=0.4), None, False),
        ("sbt_bdpcm", dict(seed=15, flags=T | SUB_DEP_QUANT, sbt_fraction=0.25, bdpcm_fraction=0.15), None, True),
        ("plain", dict(seed=16, flags=T, ts_fraction=0.3), None, False),
        ("small_shapes", dict(seed=17, shapes=[(2, 2), (2, 4), (4, 2), (2, 8), (8, 2), (2, 16), (16, 2), (32, 2),
                                                   (2, 32), (4, 4), (4, 32), (64, 4), (4, 64), (64, 64), (64, 32)]),
         None, False),
    ]
    out = {}
    names = []
    for name, kw, _, use_mts in cases:
        wl = W.random_substreams(10, tus_per_sub=(1, 12), **kw)
        names.append(name)
        out[f"{name}_tus"] = wl.tus
        out[f"{name}_tu_offsets"] = wl.tu_offsets
        out[f"{name}_params"] = wl.params
        assert np.abs(wl.coeffs).max() <= 32768
        out[f"{name}_coeffs"] = wl.coeffs.astype(np.int32)
        blobs, metas, flags_all, ctxs, dec_all = [], [], [], [], np.zeros_like(wl.coeffs)
        for i in range(wl.n_sub):
            a, b = int(wl.tu_offsets[i]), int(wl.tu_offsets[i + 1])
            p = wl.params[i]
            by, nb, tuf, ctx = ref.encode_residual(wl.tus[a:b], wl.coeffs, int(p["qp"]), int(p["init_id"]),
                                                   int(p["flags"]), want_ctx=True, use_mts=use_mts)
            dec, bits, tuf2, st = ref.decode_residual(by, wl.tus[a:b], wl.coeffs.size, int(p["qp"]),
                                                      int(p["init_id"]), int(p["flags"]), use_mts=use_mts)
            assert st == 0 and (tuf == tuf2).all()
            lo = int(wl.tus[a]["coeff_offset"])
            last = wl.tus[b - 1]
            hi = int(last["coeff_offset"]) + (1 << (int(last["log2_w"]) + int(last["log2_h"])))
            dec_all[lo:hi] = dec[lo:hi]
            blobs.append(by)
            metas.append([nb, bits, len(by)])
            flags_all.append(tuf)
            ctxs.append(np.stack([ctx["state0"], ctx["state1"]], axis=1))
        out[f"{name}_bytes"] = np.concatenate(blobs)
        out[f"{name}_meta"] = np.asarray(metas, np.int64)
        out[f"{name}_tu_flags"] = np.concatenate(flags_all)
        out[f"{name}_ctx"] = np.stack(ctxs)
        out[f"{name}_decoded"] = dec_all
        print(name, wl.n_sub, "substreams", len(wl.tus), "TUs", sum(m[0] for m in metas), "bins")
    out["names"] = np.asarray(names)
    np.savez_compressed(os.path.join(HERE, "residual.npz"), **out)
    # ---- RDO bit estimates of the same residual cases (BitEstimator_Std behind CABACWriter) ----
    est = {}
    for (name, _, _, use_mts) in cases:
        tus, offs, params, coeffs = (out[f"{name}_{k}"] for k in ("tus", "tu_offsets", "params", "coeffs"))
        totals, per_tu = [], []
        for i in range(len(params)):
            a, b = int(offs[i]), int(offs[i + 1])
            p = params[i]
            tot, pt = ref.estimate_residual(tus[a:b], coeffs, int(p["qp"]), int(p["init_id"]), int(p["flags"]),
                                            use_mts=use_mts)
            totals.append(tot)
            per_tu.append(pt)
        est[f"{name}_est_bits"] = np.asarray(totals, np.uint64)
        est[f"{name}_est_per_tu"] = np.concatenate(per_tu)
    est["frac_bits"] = ref.frac_bits()
    np.savez_compressed(os.path.join(HERE, "estimate.npz"), **est)
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    main()
