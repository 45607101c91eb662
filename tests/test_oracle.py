"""CPU suite: pins the C oracle against (a) golden vectors produced by the compiled
reference, (b) the compiled reference itself when oracle/_ref is present."""
import numpy as np
import pytest

from conftest import load_golden
from entropy_coding_b200 import workload as W
from entropy_coding_b200.types import (CTX_DT, SUB_ALIGN, SUB_DEP_QUANT, SUB_SIGN_HIDING, SUB_TERMINATE,
                                       SUB_TSRC_DISABLED, TU_DT, PARAMS_DT, mk_ctx, mk_ep, mk_trm)
from helpers import T, ctx_pairs, expected_values, random_cmds, sub_range

KAT_HEX = {
    "kat0": "fe80",
    "kat1": "dbdf9f8c19caa1e870",
    "kat2": "e17ad7809c79df2838b55a456920d7fc885c553a085c7a9c295f14e61947b14cf706f1e138a791311e63525281a3bea8"
            "22151f1bc1b9692dbb372f80",
    "kat3": "fefe00000000000000005c7fffffffff40",
}
KAT_BINS = {"kat0": 1, "kat1": 76, "kat2": 490, "kat3": 342}


def test_kats(oracle):
    g = load_golden("kat")
    for name, hx in KAT_HEX.items():
        qp, iid, nb = (int(v) for v in g[name + "_meta"])
        assert g[name + "_bytes"].tobytes().hex() == hx
        out, bins = oracle.encode_bins(g[name + "_cmds"], qp, iid, T)
        assert out.tobytes().hex() == hx and bins == nb == KAT_BINS[name]


def test_ctx_init_golden(oracle):
    g = load_golden("ctx_init")
    for r in range(4):
        assert (oracle.init_table(r) == g[f"table{r}"]).all()
    assert (oracle.ctx_offsets() == g["ctx_offsets"]).all()
    for qp in (-5, 0, 9, 17, 26, 32, 40, 51, 63, 70):
        for iid in range(3):
            st, rates = oracle.ctx_init(qp, iid)
            assert (ctx_pairs(st) == g[f"q{qp}_i{iid}"]).all()
            assert (rates == g["rates"]).all()
    # spot values quoted in SURVEY.md section 8c (ctx 90, I slices)
    assert tuple(g["q0_i2"][90]) == (6912, 6912) and tuple(g["q32_i2"][90]) == (2816, 2816)
    assert tuple(g["q51_i2"][90]) == (256, 256) and g["rates"][90] == 0x58


def test_scan_golden(oracle):
    g = load_golden("scan")
    for lw in range(1, 7):
        for lh in range(1, 7):
            tab, cw, ch = oracle.scan_table(lw, lh)
            assert (tab == g[f"s{lw}_{lh}"]).all() and (cw, ch) == tuple(int(v) for v in g[f"cg{lw}_{lh}"])


def test_bins_golden(oracle):
    g = load_golden("bins")
    for i in range(int(g["n"][0])):
        qp, iid, nb, bits, st = (int(v) for v in g[f"meta{i}"])
        cmds = g[f"cmds{i}"]
        out, bins, ctx = oracle.encode_bins(cmds, qp, iid, T, want_ctx=True)
        assert out.tobytes() == g[f"bytes{i}"].tobytes() and bins == nb
        assert (ctx_pairs(ctx) == g[f"ctx{i}"]).all()
        vals, b, s = oracle.decode_bins(g[f"bytes{i}"], cmds, qp, iid)
        assert (vals == g[f"vals{i}"]).all() and b == bits and s == st
        if not ((cmds >> 29) == 4).any():
            assert (vals == expected_values(cmds)).all() and s == 0


def test_residual_golden(oracle):
    g = load_golden("residual")
    for name in g["names"]:
        tus, offs, params, coeffs = g[f"{name}_tus"], g[f"{name}_tu_offsets"], g[f"{name}_params"], g[f"{name}_coeffs"]
        meta, blob = g[f"{name}_meta"], g[f"{name}_bytes"]
        pos, dec_all = 0, np.zeros_like(coeffs)
        for i in range(len(params)):
            a, b = int(offs[i]), int(offs[i + 1])
            p = params[i]
            nb, bits, nbytes = (int(v) for v in meta[i])
            out, bins, tuf, ctx = oracle.encode_residual(tus[a:b], coeffs, int(p["qp"]), int(p["init_id"]),
                                                         int(p["flags"]), want_ctx=True)
            exp = blob[pos:pos + nbytes]
            pos += nbytes
            assert out.tobytes() == exp.tobytes(), (name, i)
            assert bins == nb and (tuf == g[f"{name}_tu_flags"][a:b]).all()
            assert (ctx_pairs(ctx) == g[f"{name}_ctx"][i]).all()
            dec, b2, tuf2, st = oracle.decode_residual(exp, tus[a:b], coeffs.size, int(p["qp"]), int(p["init_id"]),
                                                       int(p["flags"]))
            assert st == 0 and b2 == bits and (tuf2 == tuf).all()
            lo = int(tus[a]["coeff_offset"])
            hi = int(tus[b - 1]["coeff_offset"]) + (1 << (int(tus[b - 1]["log2_w"]) + int(tus[b - 1]["log2_h"])))
            dec_all[lo:hi] = dec[lo:hi]
        assert (dec_all == g[f"{name}_decoded"]).all()
        if not (params["flags"] & SUB_SIGN_HIDING).any():
            assert (dec_all == coeffs).all()  # decode(encode(x)) == x


# ---- differential tests against the compiled reference (build container only) ----
def test_oracle_vs_ref_tables(oracle, ref):
    assert (oracle.ctx_offsets() == ref.ctx_offsets()).all()
    for r in range(4):
        assert (oracle.init_table(r) == ref.init_table(r)).all()
    for qp in range(-2, 66, 3):
        for iid in range(3):
            a, b = oracle.ctx_init(qp, iid), ref.ctx_init(qp, iid)
            assert (a[0] == b[0]).all() and (a[1] == b[1]).all()
    for lw in range(0, 7):
        for lh in range(0, 7):
            a, b = oracle.scan_table(lw, lh), ref.scan_table(lw, lh)
            assert (a[0] == b[0]).all() and a[1:] == b[1:]


def test_oracle_vs_ref_bins(oracle, ref):
    rng = np.random.default_rng(7)
    for it in range(120):
        cmds = random_cmds(rng, int(rng.integers(0, 500)), with_align=(it % 3 == 0))
        qp, iid = int(rng.integers(0, 64)), int(rng.integers(0, 3))
        a = ref.encode_bins(cmds, qp, iid, T, want_ctx=True)
        b = oracle.encode_bins(cmds, qp, iid, T, want_ctx=True)
        assert a[0].tobytes() == b[0].tobytes() and a[1] == b[1] and (a[2] == b[2]).all()
        da = ref.decode_bins(a[0], cmds, qp, iid, want_ctx=True)
        db = oracle.decode_bins(a[0], cmds, qp, iid, want_ctx=True)
        assert (da[0] == db[0]).all() and da[1:3] == db[1:3] and (da[3] == db[3]).all()


@pytest.mark.parametrize("seed,flags,kw", [
    (101, None, {}),
    (102, T | SUB_SIGN_HIDING, {}),
    (103, T | SUB_DEP_QUANT | SUB_SIGN_HIDING, {"ts_fraction": 0.3}),
    (104, T | SUB_DEP_QUANT | SUB_TSRC_DISABLED, {"ts_fraction": 0.4}),
    (105, T | SUB_DEP_QUANT, {"sbt_fraction": 0.2, "bdpcm_fraction": 0.1}),
])
def test_oracle_vs_ref_residual(oracle, ref, seed, flags, kw):
    wl = W.random_substreams(12, seed, flags=flags, **kw)
    mts = "sbt_fraction" in kw
    for i in range(wl.n_sub):
        a, b, lo, hi = sub_range(wl, i)
        p = wl.params[i]
        args = (int(p["qp"]), int(p["init_id"]), int(p["flags"]))
        ra = ref.encode_residual(wl.tus[a:b], wl.coeffs, *args, want_ctx=True, use_mts=mts)
        oa = oracle.encode_residual(wl.tus[a:b], wl.coeffs, *args, want_ctx=True)
        assert ra[0].tobytes() == oa[0].tobytes() and ra[1] == oa[1]
        assert (ra[2] == oa[2]).all() and (ra[3] == oa[3]).all()
        rd = ref.decode_residual(ra[0], wl.tus[a:b], wl.coeffs.size, *args, want_ctx=True, use_mts=mts)
        od = oracle.decode_residual(ra[0], wl.tus[a:b], wl.coeffs.size, *args, want_ctx=True)
        assert (rd[0] == od[0]).all() and rd[1] == od[1] and (rd[2] == od[2]).all() and rd[3] == od[3] == 0
        assert (rd[4] == od[4]).all()
        assert (np.abs(od[0][lo:hi]) == np.abs(wl.coeffs[lo:hi])).all()
        if not (int(p["flags"]) & SUB_SIGN_HIDING):
            assert (od[0][lo:hi] == wl.coeffs[lo:hi]).all()


def test_oracle_edge_cases(oracle):
    # empty command stream, terminate only (KAT0); zero-length EP; single-coefficient TUs of every shape
    out, nb = oracle.encode_bins([mk_ep(0, 0)], 32, 2, T)
    assert out.tobytes().hex() == "fe80" and nb == 1
    for lw in range(0, 7):
        for lh in range(0, 7):
            tus = np.zeros(1, TU_DT)
            tus["log2_w"], tus["log2_h"] = lw, lh
            for pos in (0, (1 << (min(lw, 5))) - 1):
                coeffs = np.zeros(1 << (lw + lh), np.int32)
                coeffs[pos] = -3
                out, nb, tuf = oracle.encode_residual(tus, coeffs, 30, 2, T | SUB_DEP_QUANT)
                dec, bits, tuf2, st = oracle.decode_residual(out, tus, coeffs.size, 30, 2, T | SUB_DEP_QUANT)
                assert st == 0 and (dec == coeffs).all() and (tuf == tuf2).all()
    # extreme levels round-trip (escape codes at the dynamic-range limit)
    tus = np.zeros(1, TU_DT)
    tus["log2_w"], tus["log2_h"] = 2, 2
    coeffs = np.asarray([32767, -32768, 1, 0, 0, 0, -32768, 0, 5, 0, 0, 0, 0, 0, 0, 32767], np.int32)
    out, nb, tuf = oracle.encode_residual(tus, coeffs, 0, 2, T | SUB_DEP_QUANT)
    dec, bits, tuf2, st = oracle.decode_residual(out, tus, 16, 0, 2, T | SUB_DEP_QUANT)
    assert st == 0 and (dec == coeffs).all()


def test_oracle_vs_ref_isp_shapes(oracle, ref):
    """1xN / 2xN / Nx1 / Nx2 blocks (intra sub-partitions; coefficient groups 1x16, 2x8, 16x1, 8x2, 2x2 of
    g_log2SbbSize, rom.cpp:41-50): restatement == compiled reference on dense random levels, encode + decode."""
    rng = np.random.default_rng(4321)
    shapes = [(0, 4), (0, 5), (0, 6), (4, 0), (5, 0), (6, 0), (1, 3), (1, 4), (1, 5), (3, 1), (4, 1), (5, 1), (6, 1),
              (1, 6), (1, 1), (1, 2), (2, 1), (0, 2), (2, 0), (0, 3), (3, 0), (0, 1), (1, 0), (0, 0)]
    for rep in range(8):
        tus, coeffs, cur = np.zeros(len(shapes), TU_DT), [], 0
        for k, (lw, lh) in enumerate(shapes):
            w, h = 1 << lw, 1 << lh
            c = np.rint(rng.laplace(0, [0.6, 2.0, 9.0, 300.0][rep % 4], (h, w))).astype(np.int32)
            c[min(h, 32):, :] = 0
            c[:, min(w, 32):] = 0
            if not c.any():
                c[0, 0] = -1
            tus[k]["coeff_offset"], tus[k]["log2_w"], tus[k]["log2_h"], tus[k]["comp_id"] = cur, lw, lh, rep % 3
            coeffs.append(c.reshape(-1))
            cur += c.size
        coeffs = np.concatenate(coeffs)
        flags = T | (SUB_DEP_QUANT if rep % 2 else SUB_SIGN_HIDING)
        qp = [0, 22, 37, 51][rep % 4]
        ro = oracle.encode_residual(tus, coeffs, qp, 2, flags, want_ctx=True)
        rr = ref.encode_residual(tus, coeffs, qp, 2, flags, want_ctx=True)
        assert ro[0].tobytes() == rr[0].tobytes() and ro[1] == rr[1] and (ro[2] == rr[2]).all()
        assert (ro[3] == rr[3]).all()
        od = oracle.decode_residual(ro[0], tus, coeffs.size, qp, 2, flags)
        rd = ref.decode_residual(ro[0], tus, coeffs.size, qp, 2, flags)
        assert (od[0] == rd[0]).all() and od[1] == rd[1] and od[3] == rd[3] == 0


def test_stream_oracle_vs_reference(oracle, ref):
    """Mixed command streams (BinEncIf calls + residual_coding markers): restatement == compiled reference."""
    from entropy_coding_b200 import workload as W
    from helpers import mixed_stream
    rng = np.random.default_rng(77)
    wl = W.random_substreams(24, 77, tus_per_sub=(0, 7), ts_fraction=0.2)
    for i in range(wl.n_sub):
        a, b = int(wl.tu_offsets[i]), int(wl.tu_offsets[i + 1])
        cmds = mixed_stream(rng, b - a)
        p = wl.params[i]
        args = (cmds, wl.tus[a:b], wl.coeffs, int(p["qp"]), int(p["init_id"]), int(p["flags"]))
        ro, rr = oracle.encode_stream(*args, want_ctx=True), ref.encode_stream(*args, want_ctx=True)
        assert ro[0].tobytes() == rr[0].tobytes() and ro[1] == rr[1]
        assert (ro[2] == rr[2]).all() and (ro[3] == rr[3]).all()


# ---- the RDO bit estimator (BitEstimator_Std behind the same writer) ----
def test_frac_bits_table_vs_ref(oracle, ref):
    assert (oracle.frac_bits() == ref.frac_bits()).all()


@pytest.mark.parametrize("seed,flags,kw", [
    (201, None, {}),
    (202, SUB_SIGN_HIDING, {}),
    (203, SUB_DEP_QUANT | SUB_SIGN_HIDING, {"ts_fraction": 0.3}),
    (204, SUB_DEP_QUANT | SUB_TSRC_DISABLED, {"ts_fraction": 0.4, "bdpcm_fraction": 0.1}),
])
def test_oracle_vs_ref_estimate(oracle, ref, seed, flags, kw):
    wl = W.random_substreams(10, seed, flags=flags, **kw)
    for i in range(wl.n_sub):
        a, b, _, _ = sub_range(wl, i)
        p = wl.params[i]
        args = (int(p["qp"]), int(p["init_id"]), int(p["flags"]))
        r = ref.estimate_residual(wl.tus[a:b], wl.coeffs, *args, want_ctx=True)
        o = oracle.estimate_residual(wl.tus[a:b], wl.coeffs, *args, want_ctx=True)
        assert r[0] == o[0] and (r[1] == o[1]).all() and (r[2] == o[2]).all()
        # the estimator leaves the contexts exactly where the real encoder does
        e = oracle.encode_residual(wl.tus[a:b], wl.coeffs, *args, want_ctx=True)
        assert (e[3] == o[2]).all()
        # and its estimate tracks the real payload (15-bit fixed point) to within the coder's slack
        bits = o[0] / 32768.0
        assert abs(bits - 8 * len(e[0])) <= 0.02 * 8 * len(e[0]) + 64


def test_estimate_golden(oracle):
    """The oracle's estimator against the reference-generated estimate.npz (BitEstimator_Std totals)."""
    g, e = load_golden("residual"), load_golden("estimate")
    assert (oracle.frac_bits() == e["frac_bits"]).all()
    for name in g["names"]:
        tus, offs, params, coeffs = g[f"{name}_tus"], g[f"{name}_tu_offsets"], g[f"{name}_params"], g[f"{name}_coeffs"]
        for i in range(len(params)):
            a, b = int(offs[i]), int(offs[i + 1])
            p = params[i]
            tot, per_tu = oracle.estimate_residual(tus[a:b], coeffs, int(p["qp"]), int(p["init_id"]), int(p["flags"]))
            assert tot == int(e[f"{name}_est_bits"][i]), (name, i)
            assert (per_tu == e[f"{name}_est_per_tu"][a:b]).all()


def test_ctx_init_id_golden(oracle):
    """BinCounter per context and xGetCtxInitId (cabac_writer.cpp:41-81) of the restatement against the
    reference-generated ctx_init_id.npz (getNumBins(k) and CABACWriter::getCtxInitId of the compiled reference)."""
    g, c = load_golden("residual"), load_golden("ctx_init_id")
    seen = set()
    for name in c["names"]:
        tus, offs, params, coeffs = g[f"{name}_tus"], g[f"{name}_tu_offsets"], g[f"{name}_params"], g[f"{name}_coeffs"]
        for i in range(len(params)):
            a, b = int(offs[i]), int(offs[i + 1])
            p = params[i]
            for init_id in (0, 1):
                bins, ctx, choice = oracle.residual_ctx_bins(tus[a:b], coeffs, int(p["qp"]), init_id, int(p["flags"]))
                assert (bins == c[f"{name}_ctx_bins"][i, init_id]).all(), (name, i, init_id)
                assert choice == int(c[f"{name}_choice"][i, init_id]), (name, i, init_id)
                seen.add(choice)
    assert seen == {0, 1}  # both tables get chosen somewhere in the fixture


@pytest.mark.parametrize("seed", [231, 232])
def test_oracle_vs_ref_ctx_init_id(oracle, ref, seed):
    wl = W.random_substreams(8, seed, tus_per_sub=(1, 10), ts_fraction=0.2)
    for i in range(wl.n_sub):
        a, b, _, _ = sub_range(wl, i)
        p = wl.params[i]
        for init_id in (0, 1, 2):
            for qp in (int(p["qp"]), 17, 40):
                o = oracle.residual_ctx_bins(wl.tus[a:b], wl.coeffs, qp, init_id, int(p["flags"]))
                r = ref.residual_ctx_bins(wl.tus[a:b], wl.coeffs, qp, init_id, int(p["flags"]))
                assert (o[0] == r[0]).all() and (o[1] == r[1]).all() and o[2] == r[2], (seed, i, init_id, qp)
