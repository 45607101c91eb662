"""GPU suite (-m gpu): the CUDA path, called through the C ABI, against the CPU
oracle and the reference-generated golden fixtures.  Bit-exact everywhere: this
path is integer/byte work."""
import os

import numpy as np
import pytest

from conftest import load_golden
from entropy_coding_b200 import workload as W
from entropy_coding_b200.types import (CTX_DT, PARAMS_DT, SS_EMPTY_TU, SS_OUT_OVERFLOW, SUB_ALIGN, SUB_CTX_IN,
                                       SUB_DEP_QUANT, SUB_SIGN_HIDING, SUB_TERMINATE, SUB_TSRC_DISABLED, TU_DT,
                                       mk_ctx, mk_ep, mk_trm)
from helpers import T, ctx_pairs, expected_values, mixed_stream, random_cmds, sub_range

pytestmark = pytest.mark.gpu


@pytest.fixture(params=[0, 2], ids=["auto_fused", "two_stage"])
def stream_mode(request, engine):
    """Recorded streams run on the fused lockstep coder (AUTO below 32 768 substreams) and on stage P + stage D."""
    engine.set_encoder_mode(request.param)
    yield request.param
    engine.set_encoder_mode(0)


def offsets_of(lengths, align=1):
    lengths = (np.asarray(lengths, np.int64) + align - 1) // align * align
    return np.concatenate([[0], np.cumsum(lengths)]).astype(np.uint64)


def split(out, offsets, sizes):
    return [out[int(offsets[i]):int(offsets[i]) + int(sizes[i])] for i in range(len(sizes))]


# ---------------------------------------------------------------- (2) contexts
def test_ctx_init_matches_golden_and_oracle(engine, oracle):
    g = load_golden("ctx_init")
    qps = (-5, 0, 9, 17, 26, 32, 40, 51, 63, 70)
    params = np.zeros(len(qps) * 3, PARAMS_DT)
    k = 0
    for qp in qps:
        for iid in range(3):
            params[k] = (qp, iid, 0, 0)
            k += 1
    out = engine.ctx_init(params)
    k = 0
    for qp in qps:
        for iid in range(3):
            assert (ctx_pairs(out[k]) == g[f"q{qp}_i{iid}"]).all()
            assert (ctx_pairs(out[k]) == ctx_pairs(oracle.ctx_init(qp, iid)[0])).all()
            k += 1
    assert (engine.ctx_rates() == g["rates"]).all()


@pytest.mark.usefixtures("stream_mode")
# ---------------------------------------------------------------- (1) bin level
def test_kats_on_gpu(engine):
    g = load_golden("kat")
    names = ["kat0", "kat1", "kat2", "kat3"]
    cmds = [g[n + "_cmds"] for n in names]
    params = np.zeros(4, PARAMS_DT)
    for i, n in enumerate(names):
        qp, iid, _ = (int(v) for v in g[n + "_meta"])
        params[i] = (qp, iid, T, 0)
    coff = offsets_of([len(c) for c in cmds])
    ooff = offsets_of([128] * 4)
    r = engine.encode_bins(np.concatenate(cmds), coff, params, ooff)
    for i, n in enumerate(names):
        got = split(r["out"], ooff, r["sizes"])[i]
        assert got.tobytes() == g[n + "_bytes"].tobytes(), n
        assert int(r["num_bins"][i]) == int(g[n + "_meta"][2])
    assert not r["status"].any()


@pytest.mark.usefixtures("stream_mode")
def test_bins_golden_on_gpu(engine):
    g = load_golden("bins")
    n = int(g["n"][0])
    cmds = [g[f"cmds{i}"] for i in range(n)]
    params = np.zeros(n, PARAMS_DT)
    for i in range(n):
        params[i] = (int(g[f"meta{i}"][0]), int(g[f"meta{i}"][1]), T, 0)
    coff = offsets_of([len(c) for c in cmds])
    ooff = offsets_of([len(g[f"bytes{i}"]) + 32 for i in range(n)], 16)
    r = engine.encode_bins(np.concatenate(cmds), coff, params, ooff, want_ctx=True)
    outs = split(r["out"], ooff, r["sizes"])
    for i in range(n):
        assert outs[i].tobytes() == g[f"bytes{i}"].tobytes(), i
        assert int(r["num_bins"][i]) == int(g[f"meta{i}"][2])
        assert (ctx_pairs(r["ctx_out"][i]) == g[f"ctx{i}"]).all()
    # decode the reference's bytes
    blobs = [g[f"bytes{i}"] for i in range(n)]
    ioff = offsets_of([len(b) for b in blobs])
    params["flags"] = SUB_TERMINATE
    d = engine.decode_bins(np.concatenate(blobs), ioff, np.concatenate(cmds), coff, params, want_ctx=True)
    for i in range(n):
        a, b = int(coff[i]), int(coff[i + 1])
        assert (d["values"][a:b] == g[f"vals{i}"]).all(), i
        assert int(d["bits_read"][i]) == int(g[f"meta{i}"][3]) and int(d["status"][i]) == int(g[f"meta{i}"][4])
        assert (ctx_pairs(d["ctx_out"][i]) == g[f"ctx{i}"]).all()


@pytest.mark.usefixtures("stream_mode")
def test_bins_random_vs_oracle(engine, oracle):
    rng = np.random.default_rng(2024)
    n = 300
    cmds = [random_cmds(rng, int(rng.integers(0, 900))) for _ in range(n)]
    cmds[0] = np.zeros(0, np.uint32)  # empty substream
    params = np.zeros(n, PARAMS_DT)
    params["qp"] = rng.integers(0, 64, n)
    params["init_id"] = rng.integers(0, 3, n)
    params["flags"] = T
    coff = offsets_of([len(c) for c in cmds])
    ooff = offsets_of([len(c) * 5 + 64 for c in cmds], 16)
    r = engine.encode_bins(np.concatenate(cmds), coff, params, ooff, want_ctx=True)
    outs = split(r["out"], ooff, r["sizes"])
    for i in range(n):
        eb, nb, ectx = oracle.encode_bins(cmds[i], int(params["qp"][i]), int(params["init_id"][i]), T, want_ctx=True)
        assert outs[i].tobytes() == eb.tobytes(), i
        assert int(r["num_bins"][i]) == nb and (r["ctx_out"][i] == ectx).all()
    ioff = offsets_of(r["sizes"])
    params["flags"] = SUB_TERMINATE
    d = engine.decode_bins(np.concatenate(outs), ioff, np.concatenate(cmds), coff, params)
    assert (d["values"] == expected_values(np.concatenate(cmds))).all()
    assert not d["status"].any()
    assert (d["bits_read"] <= 8 * r["sizes"]).all()


@pytest.mark.usefixtures("stream_mode")
def test_bins_ctx_handoff(engine, oracle):
    """Encoding in two launches with exported/imported contexts (WPP-style hand-off) keeps the models exact."""
    rng = np.random.default_rng(5)
    c1, c2 = random_cmds(rng, 400), random_cmds(rng, 400)
    p = np.zeros(1, PARAMS_DT)
    p[0] = (27, 1, 0, 0)
    r1 = engine.encode_bins(c1, offsets_of([len(c1)]), p, offsets_of([4096]), want_ctx=True)
    p2 = p.copy()
    p2["flags"] = T | SUB_CTX_IN
    r2 = engine.encode_bins(c2, offsets_of([len(c2)]), p2, offsets_of([4096]), ctx_in=r1["ctx_out"], want_ctx=True)
    _, _, ectx1 = oracle.encode_bins(c1, 27, 1, 0, want_ctx=True)
    assert (r1["ctx_out"][0] == ectx1).all()
    eb, nb, ectx2 = oracle.encode_bins(c2, 27, 1, T, ctx_in=ectx1, want_ctx=True)
    assert r2["out"][:r2["sizes"][0]].tobytes() == eb.tobytes() and (r2["ctx_out"][0] == ectx2).all()


@pytest.mark.usefixtures("stream_mode")
def test_long_ff_run_and_overflow(engine, oracle):
    """Carry propagation through a long outstanding-byte run (numBufferedBytes), and the overflow status."""
    cmds = np.asarray([mk_ctx(246, 1)] * 300 + [mk_ep(1, 1)] * 3000 + [mk_trm(0)], np.uint32)
    p = np.zeros(1, PARAMS_DT)
    p[0] = (51, 0, T, 0)
    r = engine.encode_bins(cmds, offsets_of([len(cmds)]), p, offsets_of([1024]))
    eb, nb = oracle.encode_bins(cmds, 51, 0, T)
    assert r["out"][:r["sizes"][0]].tobytes() == eb.tobytes() and len(eb) > 300
    r = engine.encode_bins(cmds, offsets_of([len(cmds)]), p, offsets_of([64]), raise_on_status=False)
    assert int(r["status"][0]) & SS_OUT_OVERFLOW and int(r["sizes"][0]) == len(eb)
    assert r["out"][:64].tobytes() == eb[:64].tobytes()


# ---------------------------------------------------------------- (3) residual level
def run_residual(engine, wl):
    cap = wl.out_capacity(bytes_per_coeff=5.0)
    r = engine.encode_residual(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, cap, want_ctx=True)
    return r, split(r["out"], cap, r["sizes"])


def test_residual_golden_on_gpu(engine):
    g = load_golden("residual")
    for name in g["names"]:
        wl = W.Workload(g[f"{name}_tus"], g[f"{name}_tu_offsets"], g[f"{name}_params"], g[f"{name}_coeffs"], name)
        r, outs = run_residual(engine, wl)
        meta, blob = g[f"{name}_meta"], g[f"{name}_bytes"]
        pos = 0
        for i in range(wl.n_sub):
            nb, bits, nbytes = (int(v) for v in meta[i])
            assert outs[i].tobytes() == blob[pos:pos + nbytes].tobytes(), (name, i)
            pos += nbytes
            assert int(r["num_bins"][i]) == nb
            assert (ctx_pairs(r["ctx_out"][i]) == g[f"{name}_ctx"][i]).all()
        assert (r["tu_flags"] == g[f"{name}_tu_flags"]).all()
        # decode the reference's bytes
        ioff = offsets_of(meta[:, 2])
        d = engine.decode_residual(blob, ioff, wl.tus, wl.tu_offsets, wl.params, wl.coeffs.size, want_ctx=True)
        assert (d["coeffs"] == g[f"{name}_decoded"]).all(), name
        assert (d["bits_read"] == meta[:, 1]).all() and (d["num_bins"] == meta[:, 0]).all()
        assert (d["tu_flags"] == g[f"{name}_tu_flags"]).all() and not d["status"].any()
        assert (ctx_pairs(d["ctx_out"]) == g[f"{name}_ctx"]).all()


@pytest.mark.parametrize("seed,flags,kw", [
    (201, None, {}),
    (202, T | SUB_SIGN_HIDING, {}),
    (203, T | SUB_DEP_QUANT | SUB_SIGN_HIDING, {"ts_fraction": 0.3}),
    (204, T | SUB_DEP_QUANT | SUB_TSRC_DISABLED, {"ts_fraction": 0.4}),
    (205, T | SUB_DEP_QUANT, {"sbt_fraction": 0.2, "bdpcm_fraction": 0.1}),
    (206, T, {"preset": "qp0", "tus_per_sub": (1, 3)}),
    (207, None, {"preset": "qp51", "tus_per_sub": (30, 60)}),
])
def test_residual_random_vs_oracle(engine, oracle, seed, flags, kw):
    wl = W.random_substreams(40, seed, flags=flags, **kw)
    r, outs = run_residual(engine, wl)
    assert not r["status"].any()
    for i in range(wl.n_sub):
        a, b, lo, hi = sub_range(wl, i)
        p = wl.params[i]
        eb, nb, tuf, ectx = oracle.encode_residual(wl.tus[a:b], wl.coeffs, int(p["qp"]), int(p["init_id"]),
                                                   int(p["flags"]), want_ctx=True)
        assert outs[i].tobytes() == eb.tobytes(), (seed, i)
        assert int(r["num_bins"][i]) == nb and (r["tu_flags"][a:b] == tuf).all()
        assert (r["ctx_out"][i] == ectx).all()
    ioff = offsets_of(r["sizes"])
    d = engine.decode_residual(np.concatenate(outs), ioff, wl.tus, wl.tu_offsets, wl.params, wl.coeffs.size,
                               want_ctx=True)
    assert not d["status"].any()
    assert (d["num_bins"] == r["num_bins"]).all() and (d["tu_flags"] == r["tu_flags"]).all()
    assert (d["ctx_out"] == r["ctx_out"]).all()
    assert (np.abs(d["coeffs"]) == np.abs(wl.coeffs)).all()
    if not (wl.params["flags"] & SUB_SIGN_HIDING).any():
        assert (d["coeffs"] == wl.coeffs).all()
    else:  # hidden signs follow the parity rule: compare with the oracle's decode
        for i in range(0, wl.n_sub, 7):
            a, b, lo, hi = sub_range(wl, i)
            p = wl.params[i]
            od = oracle.decode_residual(outs[i], wl.tus[a:b], wl.coeffs.size, int(p["qp"]), int(p["init_id"]),
                                        int(p["flags"]))
            assert (od[0][lo:hi] == d["coeffs"][lo:hi]).all() and od[1] == int(d["bits_read"][i])


@pytest.mark.parametrize("mode", [1, 2, 3], ids=["fused", "two_stage", "warp_specialised"])
@pytest.mark.parametrize("seed,flags,kw", [
    (301, None, {}),
    (302, T | SUB_DEP_QUANT | SUB_SIGN_HIDING, {"ts_fraction": 0.3, "bdpcm_fraction": 0.1}),
    (303, SUB_DEP_QUANT, {"preset": "qp0", "tus_per_sub": (1, 4)}),        # not terminated: bytes held back
    (304, SUB_ALIGN, {"preset": "qp32"}),
    (305, T, {"preset": "qp0", "tus_per_sub": (20, 40), "ts_fraction": 0.0}),
])
def test_residual_encoder_variants(engine, oracle, mode, seed, flags, kw):
    """The residual encoders (fused lockstep kernel, two-stage P + D, warp-specialised) against the oracle, byte for byte."""
    engine.set_encoder_mode(mode)
    try:
        wl = W.random_substreams(70, seed, flags=flags, **kw)
        r, outs = run_residual(engine, wl)
        assert not r["status"].any()
        for i in range(wl.n_sub):
            a, b, lo, hi = sub_range(wl, i)
            p = wl.params[i]
            eb, nb, tuf, ectx = oracle.encode_residual(wl.tus[a:b], wl.coeffs, int(p["qp"]), int(p["init_id"]),
                                                       int(p["flags"]), want_ctx=True)
            assert outs[i].tobytes() == eb.tobytes(), (mode, seed, i)
            assert int(r["num_bins"][i]) == nb and (r["tu_flags"][a:b] == tuf).all()
            assert (r["ctx_out"][i] == ectx).all()
        # overflow is reported, sizes stay exact, the in-capacity prefix is intact
        small = offsets_of([48] * wl.n_sub, align=16)
        r2 = engine.encode_residual(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, small, raise_on_status=False)
        assert (r2["sizes"] == r["sizes"]).all()
        over = r["sizes"] > 48
        assert ((r2["status"] & SS_OUT_OVERFLOW) != 0)[over].all() and not r2["status"][~over].any()
        for i in range(wl.n_sub):
            k = min(48, int(r["sizes"][i]))
            assert r2["out"][int(small[i]):int(small[i]) + k].tobytes() == outs[i][:k].tobytes()
    finally:
        engine.set_encoder_mode(0)


@pytest.mark.usefixtures("stream_mode")
@pytest.mark.parametrize("seed,flags", [(401, T | SUB_DEP_QUANT), (402, T | SUB_SIGN_HIDING), (403, SUB_DEP_QUANT)])
def test_stream_mixed_vs_oracle(engine, oracle, seed, flags):
    """cabac_encode_stream: recorded BinEncIf calls interleaved with residual_coding() markers."""
    rng = np.random.default_rng(seed)
    wl = W.random_substreams(48, seed, flags=flags, tus_per_sub=(0, 9), ts_fraction=0.2)
    streams = [mixed_stream(rng, int(wl.tu_offsets[i + 1] - wl.tu_offsets[i])) for i in range(wl.n_sub)]
    coff = offsets_of([len(c) for c in streams])
    cap = offsets_of([5 * len(c) + 64 for c in streams], align=16) + wl.out_capacity(bytes_per_coeff=5.0)
    r = engine.encode_stream(np.concatenate(streams), coff, wl.tus, wl.tu_offsets, wl.params, wl.coeffs, cap,
                             want_ctx=True)
    outs = split(r["out"], cap, r["sizes"])
    assert not r["status"].any()
    for i in range(wl.n_sub):
        a, b = int(wl.tu_offsets[i]), int(wl.tu_offsets[i + 1])
        p = wl.params[i]
        eb, nb, tuf, ectx = oracle.encode_stream(streams[i], wl.tus[a:b], wl.coeffs, int(p["qp"]), int(p["init_id"]),
                                                 int(p["flags"]), want_ctx=True)
        assert outs[i].tobytes() == eb.tobytes(), (seed, i)
        assert int(r["num_bins"][i]) == nb and (r["tu_flags"][a:b] == tuf).all()
        assert (r["ctx_out"][i] == ectx).all()
    # a marker without a TU left is a malformed stream
    bad = engine.encode_stream(np.asarray([W.types.mk_tu()] if hasattr(W, "types") else [5 << 29], np.uint32),
                               offsets_of([1]), np.zeros(0, TU_DT), np.zeros(2, np.uint64), wl.params[:1],
                               np.zeros(1, np.int32), offsets_of([64]), raise_on_status=False)
    assert int(bad["status"][0]) & 4


def test_residual_all_shapes_single_coeff(engine, oracle):
    """Every block shape 1..64 x 1..64 (1- and 2-wide: ISP sub-partitions), one level at DC / at the last
    coded column, extreme magnitudes."""
    tus, coeffs, cur = [], [], 0
    for lw in range(0, 7):
        for lh in range(0, 7):
            for pos, val in ((0, -3), ((1 << min(lw, 5)) - 1, 32767), (0, -32768)):
                t = np.zeros(1, TU_DT)
                t["coeff_offset"], t["log2_w"], t["log2_h"], t["comp_id"] = cur, lw, lh, (lw + lh) % 3
                c = np.zeros(1 << (lw + lh), np.int32)
                c[pos] = val
                tus.append(t)
                coeffs.append(c)
                cur += c.size
    tus, coeffs = np.concatenate(tus), np.concatenate(coeffs)
    n = len(tus)
    params = np.zeros(n, PARAMS_DT)
    params["qp"], params["init_id"], params["flags"] = 30, 2, T | SUB_DEP_QUANT
    wl = W.Workload(tus, np.arange(n + 1, dtype=np.uint64), params, coeffs)
    r, outs = run_residual(engine, wl)
    for i in range(n):
        eb, nb, tuf = oracle.encode_residual(tus[i:i + 1], coeffs, 30, 2, T | SUB_DEP_QUANT)
        assert outs[i].tobytes() == eb.tobytes(), (i, tus[i])
    d = engine.decode_residual(np.concatenate(outs), offsets_of(r["sizes"]), tus, wl.tu_offsets, params, coeffs.size)
    assert (d["coeffs"] == coeffs).all()
    # the other encoder kernels and the bit estimator on the same shapes
    for mode in (2, 3, 4):
        engine.set_encoder_mode(mode)
        try:
            r2, outs2 = run_residual(engine, wl)
        finally:
            engine.set_encoder_mode(0)
        assert (r2["sizes"] == r["sizes"]).all() and all(a.tobytes() == b.tobytes() for a, b in zip(outs, outs2)), mode
    est = engine.estimate_residual_bits(tus, wl.tu_offsets, params, coeffs)
    for i in range(n):
        tot, _ = oracle.estimate_residual(tus[i:i + 1], coeffs, 30, 2, T | SUB_DEP_QUANT)
        assert int(est["est_bits"][i]) == tot, (i, tus[i])


def test_residual_narrow_isp_shapes_random(engine, oracle):
    """Intra sub-partition blocks (1xN, 2xN, Nx1, Nx2: coefficient groups 1x16 / 2x8 / 16x1 / 8x2 / 2x2,
    rom.cpp:41-50) with dense random levels: every encoder kernel and the decoder against the oracle."""
    rng = np.random.default_rng(1234)
    shapes = [(0, 4), (0, 5), (0, 6), (4, 0), (5, 0), (6, 0), (1, 3), (1, 4), (1, 5), (3, 1), (4, 1), (5, 1), (6, 1),
              (1, 6), (1, 1), (1, 2), (2, 1), (0, 2), (2, 0), (0, 3), (3, 0), (0, 1), (1, 0)]
    tus, coeffs, cur = [], [], 0
    for rep in range(6):
        for lw, lh in shapes:
            t = np.zeros(1, TU_DT)
            t["coeff_offset"], t["log2_w"], t["log2_h"], t["comp_id"] = cur, lw, lh, 0 if rep < 4 else 1 + rep % 2
            w, h = 1 << lw, 1 << lh
            c = np.rint(rng.laplace(0, [0.6, 2.0, 9.0, 200.0][rep % 4], (h, w))).astype(np.int32)
            c[min(h, 32):, :] = 0
            c[:, min(w, 32):] = 0
            if not c.any():
                c[0, 0] = 2
            tus.append(t)
            coeffs.append(c.reshape(-1))
            cur += c.size
    tus, coeffs = np.concatenate(tus), np.concatenate(coeffs)
    per = len(shapes)
    n = len(tus) // per
    params = np.zeros(n, PARAMS_DT)
    params["qp"], params["init_id"] = [0, 22, 37, 51, 30, 30], 2
    params["flags"] = [T | SUB_DEP_QUANT, T | SUB_SIGN_HIDING, T, T | SUB_DEP_QUANT, T | SUB_DEP_QUANT, T]
    wl = W.Workload(tus, (np.arange(n + 1) * per).astype(np.uint64), params, coeffs)
    want = []
    for i in range(n):
        p = params[i]
        want.append(oracle.encode_residual(tus[i * per:(i + 1) * per], coeffs, int(p["qp"]), 2, int(p["flags"]), want_ctx=True))
    for mode in (1, 2, 3, 4):
        engine.set_encoder_mode(mode)
        try:
            r, outs = run_residual(engine, wl)
        finally:
            engine.set_encoder_mode(0)
        for i in range(n):
            assert outs[i].tobytes() == want[i][0].tobytes(), (mode, i)
            assert int(r["num_bins"][i]) == want[i][1] and (r["tu_flags"][i * per:(i + 1) * per] == want[i][2]).all()
    d = engine.decode_residual(np.concatenate(outs), offsets_of(r["sizes"]), tus, wl.tu_offsets, params, coeffs.size)
    for i in range(n):
        dec = oracle.decode_residual(want[i][0], tus[i * per:(i + 1) * per], coeffs.size, int(params[i]["qp"]), 2,
                                     int(params[i]["flags"]))[0]
        a = int(tus[i * per]["coeff_offset"])
        b = int(tus[(i + 1) * per - 1]["coeff_offset"]) + (1 << int(tus[(i + 1) * per - 1]["log2_w"] + tus[(i + 1) * per - 1]["log2_h"]))
        assert (d["coeffs"][a:b] == dec[a:b]).all(), i


def test_residual_errors(engine):
    """Empty TU and output overflow are reported per substream, as the reference's CHECKs would throw."""
    tus = np.zeros(2, TU_DT)
    tus["log2_w"], tus["log2_h"] = 3, 3
    tus["coeff_offset"] = [0, 64]
    coeffs = np.zeros(128, np.int32)
    coeffs[64 + 5] = 9
    params = np.zeros(2, PARAMS_DT)
    params["qp"], params["init_id"], params["flags"] = 32, 2, T
    r = engine.encode_residual(tus, np.asarray([0, 1, 2], np.uint64), params, coeffs, offsets_of([64, 64]),
                               raise_on_status=False)
    assert int(r["status"][0]) & SS_EMPTY_TU and int(r["status"][1]) == 0
    import entropy_coding_b200 as E
    with pytest.raises(E.CabacError):
        engine.encode_residual(tus, np.asarray([0, 1, 2], np.uint64), params, coeffs, offsets_of([64, 64]))
    wl = W.random_substreams(2, 3, preset="qp0")
    r = engine.encode_residual(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, offsets_of([32, 32]),
                               raise_on_status=False)
    assert (r["status"] & SS_OUT_OVERFLOW).all() and (r["sizes"] > 32).all()


def test_empty_batches(engine):
    z64 = np.zeros(1, np.uint64)
    r = engine.encode_residual(np.zeros(0, TU_DT), z64, np.zeros(0, PARAMS_DT), np.zeros(0, np.int32), z64)
    assert len(r["sizes"]) == 0
    # a substream with no TUs still terminates: trm(1) + finish + alignment == KAT0
    p = np.zeros(1, PARAMS_DT)
    p[0] = (32, 2, T, 0)
    r = engine.encode_residual(np.zeros(0, TU_DT), np.zeros(2, np.uint64), p, np.zeros(0, np.int32),
                               offsets_of([16]))
    assert r["out"][:r["sizes"][0]].tobytes().hex() == "fe80"


def test_pack_and_entry_points(engine):
    wl = W.random_substreams(257, 77, tus_per_sub=(1, 6))
    r, outs = run_residual(engine, wl)
    packed, entry = engine.pack(r["out"], wl.out_capacity(bytes_per_coeff=5.0), r["sizes"])
    assert packed.tobytes() == b"".join(o.tobytes() for o in outs)
    assert (entry == offsets_of(r["sizes"])).all()


# ---------------------------------------------------------------- full-size properties
def test_full_size_round_trip_and_checksum(engine, oracle):
    """At bench scale the oracle is too slow for every substream: check size-independent properties
    (decode(encode(x)) == x, bins conserved) on everything and the oracle on a sample."""
    wl = W.frames_as_ctu_substreams(4, 1920, 1080)  # 540 substreams, ~12.4 M coefficients
    cap = wl.out_capacity()
    r = engine.encode_residual(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, cap)
    outs = split(r["out"], cap, r["sizes"])
    ioff = offsets_of(r["sizes"])
    d = engine.decode_residual(np.concatenate(outs), ioff, wl.tus, wl.tu_offsets, wl.params, wl.coeffs.size)
    assert (d["coeffs"] == wl.coeffs).all()
    assert (d["num_bins"] == r["num_bins"]).all()
    assert (d["bits_read"] <= 8 * r["sizes"]).all() and (d["bits_read"] > 8 * (r["sizes"].astype(np.int64) - 3)).all()
    for i in range(0, wl.n_sub, 45):
        a, b, lo, hi = sub_range(wl, i)
        p = wl.params[i]
        eb, nb, _ = oracle.encode_residual(wl.tus[a:b], wl.coeffs, int(p["qp"]), int(p["init_id"]), int(p["flags"]))
        assert outs[i].tobytes() == eb.tobytes() and nb == int(r["num_bins"][i])


# ---------------------------------------------------------------- chunked copy/compute pipeline
@pytest.mark.parametrize("layout", ["ordered", "shuffled"])
def test_pipelined_host_path_matches_single_launch(monkeypatch, oracle, layout):
    """cabac_encode_residual_packed / cabac_decode_residual stream the coefficient slab over PCIe in
    chunks while earlier chunks are coded.  Results must equal the one-copy-one-launch path for any
    slab layout (TU blocks in substream order or scattered, with gaps no descriptor covers)."""
    import entropy_coding_b200 as E
    wl = W.random_substreams(700, 321, tus_per_sub=(2, 9))
    tus, coeffs = wl.tus.copy(), wl.coeffs
    if layout == "shuffled":  # scatter the TU blocks over a larger slab, leaving holes
        rng = np.random.default_rng(5)
        sizes = 1 << (tus["log2_w"].astype(np.int64) + tus["log2_h"].astype(np.int64))
        perm = rng.permutation(len(tus))
        gaps = rng.integers(0, 64, len(tus)) * 4
        new_off = np.zeros(len(tus), np.int64)
        new_off[perm] = np.concatenate([[0], np.cumsum(sizes[perm] + gaps[perm])[:-1]]) + 16
        total = int((new_off + sizes).max()) + 32
        slab = np.zeros(total, np.int32)
        for t in range(len(tus)):
            o, n = int(tus["coeff_offset"][t]), int(sizes[t])
            slab[int(new_off[t]):int(new_off[t]) + n] = coeffs[o:o + n]
        tus["coeff_offset"] = new_off.astype(np.uint64)
        coeffs = slab
    cap = wl.out_capacity(bytes_per_coeff=5.0)
    results = {}
    for name, chunks in (("single", "1"), ("pipelined", "6")):
        monkeypatch.setenv("CABAC_B200_PIPE_CHUNKS", chunks)
        monkeypatch.setenv("CABAC_B200_PIPE_MIN_COEFFS", "1")
        monkeypatch.setenv("CABAC_B200_PIPE_MIN_SUBSTREAMS", "1")
        eng = E.CabacEngine(0)
        n0 = eng.launch_count
        r = eng.encode_residual_packed(tus, wl.tu_offsets, wl.params, coeffs, cap, want_ctx=True)
        n1 = eng.launch_count
        d = eng.decode_residual(r["packed"], r["entry_points"], tus, wl.tu_offsets, wl.params, coeffs.size,
                                want_ctx=True)
        results[name] = (r, d, n1 - n0, eng.launch_count - n1)
        eng.close()
    (r1, d1, le1, ld1), (r2, d2, le2, ld2) = results["single"], results["pipelined"]
    assert le2 > le1 and ld2 > ld1, "the pipelined path did not run"
    assert r1["packed"].tobytes() == r2["packed"].tobytes()
    for k in ("entry_points", "sizes", "num_bins", "tu_flags"):
        assert (r1[k] == r2[k]).all(), k
    assert (ctx_pairs(r1["ctx_out"]) == ctx_pairs(r2["ctx_out"])).all() and (ctx_pairs(d1["ctx_out"]) == ctx_pairs(d2["ctx_out"])).all()
    assert (d1["coeffs"] == d2["coeffs"]).all() and (d2["coeffs"] == coeffs).all()
    for k in ("num_bins", "bits_read", "tu_flags"):
        assert (d1[k] == d2[k]).all(), k
    # and against the oracle on a sample
    ep = r2["entry_points"]
    for i in range(0, wl.n_sub, 70):
        a, b = int(wl.tu_offsets[i]), int(wl.tu_offsets[i + 1])
        p = wl.params[i]
        eb, nb, _ = oracle.encode_residual(tus[a:b], coeffs, int(p["qp"]), int(p["init_id"]), int(p["flags"]))
        assert r2["packed"][int(ep[i]):int(ep[i + 1])].tobytes() == eb.tobytes() and nb == int(r2["num_bins"][i])


# ---------------------------------------------------------------- configs[4]: decode sweep
def _tiny_substreams(n_sub, seed):
    """Many short substreams (1-3 small TUs each), mixed QP presets, built without a Python loop per TU."""
    rng = np.random.default_rng(seed)
    ntu = rng.integers(1, 4, n_sub)
    tu_offsets = np.concatenate([[0], np.cumsum(ntu)]).astype(np.uint64)
    n = int(tu_offsets[-1])
    tus = np.zeros(n, TU_DT)
    tus["log2_w"], tus["log2_h"] = rng.integers(2, 4, n), rng.integers(2, 4, n)
    tus["comp_id"] = rng.integers(0, 3, n)
    tus["flags"] = np.where(rng.random(n) < 0.15, 3, 2)  # TS allowed, 15 % transform skip
    area = 1 << (tus["log2_w"].astype(np.int64) + tus["log2_h"].astype(np.int64))
    tus["coeff_offset"] = np.concatenate([[0], np.cumsum(area)[:-1]]).astype(np.uint64)
    total = int(area.sum())
    scale = np.repeat(rng.choice([0.6, 2.5, 30.0], n), area)
    coeffs = np.rint(rng.laplace(0, 1, total) * scale).astype(np.int32)
    first = tus["coeff_offset"].astype(np.int64)
    empty = np.add.reduceat(np.abs(coeffs), first) == 0
    coeffs[first[empty]] = 1  # residual_coding is only called for coded blocks
    params = np.zeros(n_sub, PARAMS_DT)
    params["qp"] = rng.choice([0, 32, 51], n_sub)
    params["init_id"] = rng.integers(0, 3, n_sub)
    params["flags"] = SUB_DEP_QUANT | T
    return W.Workload(tus, tu_offsets, params, coeffs, f"tiny-{n_sub}")


@pytest.mark.parametrize("n_sub", [10_000, 40_000])
def test_decode_sweep_many_substreams(engine, oracle, n_sub):
    """10^4 .. 4x10^4 independent substreams of mixed QP per launch (40 000 also crosses the threshold
    where the AUTO encoder switches to the two-stage coder): GPU decode of GPU bitstreams round-trips,
    a sample of substreams equals the oracle byte for byte, and the oracle's bitstreams decode."""
    wl = _tiny_substreams(n_sub, 42 + n_sub)
    cap = wl.out_capacity(bytes_per_coeff=6.0, slack=64)
    n0 = engine.launch_count
    r = engine.encode_residual_packed(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, cap)
    assert engine.launch_count - n0 == (7 if n_sub >= 32768 else 6)  # two-stage = one more launch
    ep = r["entry_points"]
    d = engine.decode_residual(r["packed"], ep, wl.tus, wl.tu_offsets, wl.params, wl.coeffs.size)
    assert (d["coeffs"] == wl.coeffs).all() and (d["num_bins"] == r["num_bins"]).all()
    sample = range(0, n_sub, n_sub // 50)
    blobs = []
    for i in sample:
        a, b = int(wl.tu_offsets[i]), int(wl.tu_offsets[i + 1])
        p = wl.params[i]
        eb, nb, _ = oracle.encode_residual(wl.tus[a:b], wl.coeffs, int(p["qp"]), int(p["init_id"]), int(p["flags"]),
                                           cap=4096)
        assert r["packed"][int(ep[i]):int(ep[i + 1])].tobytes() == eb.tobytes() and nb == int(r["num_bins"][i])
        blobs.append(eb)
    # decode-only leg on bitstreams produced by the CPU reference restatement
    sub = wl.subset(list(sample))
    d2 = engine.decode_residual(np.concatenate(blobs), offsets_of([len(b) for b in blobs]), sub.tus, sub.tu_offsets,
                                sub.params, sub.coeffs.size)
    assert (d2["coeffs"] == sub.coeffs).all()


def test_zero_copy_host_path_matches_staged(monkeypatch, oracle):
    """Pinned (device-mapped) host slabs are read / written by the kernels directly over PCIe; results
    must equal the staged path (H2D copy, device slab, D2H copy) bit for bit."""
    import entropy_coding_b200 as E
    wl = W.frames_as_ctu_substreams(2, 1920, 1080, seed0=2024)  # 270 substreams, 6.2 M coefficients
    cap = wl.out_capacity()
    res = {}
    for name, zc in (("zero_copy", "1"), ("staged", "0")):
        monkeypatch.setenv("CABAC_B200_ZERO_COPY", zc)
        monkeypatch.setenv("CABAC_B200_PIPE_CHUNKS", "1")
        eng = E.CabacEngine(0)
        h_co = eng.pinned_copy(wl.coeffs)
        r = eng.encode_residual_packed(wl.tus, wl.tu_offsets, wl.params, h_co, cap, want_ctx=True)
        h_dec = eng.host_array(wl.coeffs.size, np.int32)
        h_dec[:] = 0x5A5A5A5A
        n0 = eng.launch_count
        d = eng.decode_residual(r["packed"], r["entry_points"], wl.tus, wl.tu_offsets, wl.params, wl.coeffs.size,
                                coeffs_out=h_dec)
        res[name] = (r, d["coeffs"].copy(), d["num_bins"], eng.launch_count - n0)
        eng.close()
    (r1, c1, b1, l1), (r2, c2, b2, l2) = res["zero_copy"], res["staged"]
    assert l1 == l2 + 1, "the zero-copy decode path (coverage check + direct stores) did not run"
    assert r1["packed"].tobytes() == r2["packed"].tobytes()
    for k in ("entry_points", "sizes", "num_bins", "tu_flags"):
        assert (r1[k] == r2[k]).all(), k
    assert (ctx_pairs(r1["ctx_out"]) == ctx_pairs(r2["ctx_out"])).all()
    assert (c1 == wl.coeffs).all() and (c2 == wl.coeffs).all() and (b1 == b2).all()
    ep = r1["entry_points"]
    for i in range(0, wl.n_sub, 45):
        a, b, lo, hi = sub_range(wl, i)
        p = wl.params[i]
        eb, nb, _ = oracle.encode_residual(wl.tus[a:b], wl.coeffs, int(p["qp"]), int(p["init_id"]), int(p["flags"]))
        assert r1["packed"][int(ep[i]):int(ep[i + 1])].tobytes() == eb.tobytes() and nb == int(r1["num_bins"][i])


# ---------------------------------------------------------------- RDO bit estimator
def test_frac_bits_table(engine, oracle):
    assert (engine.frac_bits_table() == oracle.frac_bits()).all()


def test_estimate_golden_on_gpu(engine):
    g, e = load_golden("residual"), load_golden("estimate")
    assert (engine.frac_bits_table() == e["frac_bits"]).all()
    for name in g["names"]:
        r = engine.estimate_residual_bits(g[f"{name}_tus"], g[f"{name}_tu_offsets"], g[f"{name}_params"],
                                          g[f"{name}_coeffs"])
        assert (r["est_bits"] == e[f"{name}_est_bits"]).all(), name
        assert (r["est_per_tu"] == e[f"{name}_est_per_tu"]).all(), name
        assert (r["tu_flags"] == g[f"{name}_tu_flags"]).all()


@pytest.mark.parametrize("seed,flags,kw", [
    (601, None, {}),
    (602, SUB_SIGN_HIDING, {}),
    (603, T | SUB_DEP_QUANT | SUB_SIGN_HIDING, {"ts_fraction": 0.3}),
    (604, SUB_DEP_QUANT | SUB_TSRC_DISABLED, {"ts_fraction": 0.4, "bdpcm_fraction": 0.1}),
    (605, SUB_DEP_QUANT, {"sbt_fraction": 0.2}),
    (606, None, {"preset": "qp0", "tus_per_sub": (1, 3)}),
    (607, None, {"preset": "qp51", "tus_per_sub": (30, 60)}),
])
def test_estimate_vs_oracle(engine, oracle, seed, flags, kw):
    """cabac_estimate_residual_bits == BitEstimator_Std behind the writer: totals, per-TU running totals,
    bins, TU side effects and the contexts it leaves."""
    wl = W.random_substreams(40, seed, flags=flags, **kw)
    r = engine.estimate_residual_bits(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, want_ctx=True)
    assert not r["status"].any()
    enc, _ = run_residual(engine, wl)
    for i in range(wl.n_sub):
        a, b, _, _ = sub_range(wl, i)
        p = wl.params[i]
        tot, per_tu, ectx = oracle.estimate_residual(wl.tus[a:b], wl.coeffs, int(p["qp"]), int(p["init_id"]),
                                                     int(p["flags"]), want_ctx=True)
        assert int(r["est_bits"][i]) == tot, (seed, i)
        assert (r["est_per_tu"][a:b] == per_tu).all()
        assert (r["ctx_out"][i] == ectx).all()
    # the estimator and the real coder agree on everything but the bytes
    trm = (wl.params["flags"] & SUB_TERMINATE) != 0  # the coder's count includes its terminating bin
    assert (r["num_bins"] + trm == enc["num_bins"]).all() and (r["tu_flags"] == enc["tu_flags"]).all()
    assert (r["ctx_out"] == enc["ctx_out"]).all()


@pytest.mark.parametrize("seed,kw", [
    (621, {}),
    (622, {"ts_fraction": 0.3, "flags": SUB_DEP_QUANT | SUB_SIGN_HIDING}),
    (623, {"preset": "qp0", "tus_per_sub": (1, 4)}),
])
def test_ctx_bins_and_init_id_vs_oracle(engine, oracle, seed, kw):
    """cabac_residual_ctx_bins == BinEncIf::getNumBins(k) after residual_coding, and cabac_ctx_init_id ==
    CABACWriter::getCtxInitId (cabac_writer.cpp:41-81): choice and both cost sums."""
    wl = W.random_substreams(48, seed, **kw)
    params = wl.params.copy()
    params["init_id"] = np.arange(wl.n_sub) % 3  # B, P and I slices
    r = engine.residual_ctx_bins(wl.tus, wl.tu_offsets, params, wl.coeffs)
    assert not r["status"].any()
    est = engine.estimate_residual_bits(wl.tus, wl.tu_offsets, params, wl.coeffs, want_ctx=True)
    assert (est["ctx_out"] == r["ctx_out"]).all() and (est["num_bins"] == r["num_bins"]).all()
    choice, costs = engine.ctx_init_id(params, r["ctx_out"], r["ctx_bins"])
    for i in range(wl.n_sub):
        a, b, _, _ = sub_range(wl, i)
        p = params[i]
        bins, ctx, ch = oracle.residual_ctx_bins(wl.tus[a:b], wl.coeffs, int(p["qp"]), int(p["init_id"]),
                                                 int(p["flags"]))
        assert (r["ctx_bins"][i] == bins).all(), (seed, i)
        assert (r["ctx_out"][i] == ctx).all()
        assert int(choice[i]) == ch, (seed, i)
        _, oc = oracle.ctx_init_id(int(p["qp"]), ctx, bins)
        assert (costs[i] == oc).all()
    assert (r["ctx_bins"].sum(axis=1) <= r["num_bins"]).all()


def test_ctx_init_id_golden(engine):
    """The device path against ctx_init_id.npz (getNumBins(k) + getCtxInitId of the compiled reference)."""
    g, c = load_golden("residual"), load_golden("ctx_init_id")
    for name in c["names"]:
        tus, offs, params, coeffs = g[f"{name}_tus"], g[f"{name}_tu_offsets"], g[f"{name}_params"], g[f"{name}_coeffs"]
        for init_id in (0, 1):
            p = params.copy()
            p["init_id"] = init_id
            r = engine.residual_ctx_bins(tus, offs, p, coeffs)
            assert (r["ctx_bins"] == c[f"{name}_ctx_bins"][:, init_id]).all(), (name, init_id)
            choice, _ = engine.ctx_init_id(p, r["ctx_out"], r["ctx_bins"])
            assert (choice == c[f"{name}_choice"][:, init_id]).all(), (name, init_id)


def test_estimate_ctx_handoff_and_candidates(engine, oracle):
    """RDO use: many candidate TU lists priced from ONE saved context set (ctx_in), and an estimate
    continued across two calls equals the estimate of the concatenation."""
    wl = W.random_substreams(24, 611, flags=SUB_DEP_QUANT, tus_per_sub=(4, 9))
    full = engine.estimate_residual_bits(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, want_ctx=True)
    # split every list in two: first halves, then second halves starting from the contexts of the first
    mid = (wl.tu_offsets[:-1] + (wl.tu_offsets[1:] - wl.tu_offsets[:-1]) // 2).astype(np.uint64)
    order1 = np.concatenate([np.arange(int(wl.tu_offsets[i]), int(mid[i])) for i in range(wl.n_sub)])
    order2 = np.concatenate([np.arange(int(mid[i]), int(wl.tu_offsets[i + 1])) for i in range(wl.n_sub)])
    off1 = offsets_of(mid - wl.tu_offsets[:-1])
    off2 = offsets_of(wl.tu_offsets[1:] - mid)
    h1 = engine.estimate_residual_bits(wl.tus[order1], off1, wl.params, wl.coeffs, want_ctx=True)
    p2 = wl.params.copy()
    p2["flags"] |= SUB_CTX_IN
    h2 = engine.estimate_residual_bits(wl.tus[order2], off2, p2, wl.coeffs, ctx_in=h1["ctx_out"], want_ctx=True)
    assert (h1["est_bits"] + h2["est_bits"] == full["est_bits"]).all()
    assert (h2["ctx_out"] == full["ctx_out"]).all()
    # the same candidate priced n times from one context set gives one answer, equal to the oracle's
    k = 7
    a, b, _, _ = sub_range(wl, k)
    n = 300
    tus = np.tile(wl.tus[a:b], n)
    offs = (np.arange(n + 1) * (b - a)).astype(np.uint64)
    params = np.repeat(p2[k:k + 1], n)
    ctx = np.repeat(h1["ctx_out"][k:k + 1], n, axis=0)
    r = engine.estimate_residual_bits(tus, offs, params, wl.coeffs, ctx_in=ctx)
    tot, _ = oracle.estimate_residual(wl.tus[a:b], wl.coeffs, int(p2[k]["qp"]), int(p2[k]["init_id"]),
                                      int(p2[k]["flags"]), ctx_in=h1["ctx_out"][k])
    assert (r["est_bits"] == tot).all()


def test_estimate_empty_tu_and_empty_batch(engine):
    tus = np.zeros(2, TU_DT)
    tus["log2_w"], tus["log2_h"] = 3, 3
    tus["coeff_offset"] = [0, 64]
    coeffs = np.zeros(128, np.int32)
    coeffs[64 + 5] = 3
    params = np.zeros(1, PARAMS_DT)
    params["qp"] = 32
    r = engine.estimate_residual_bits(tus, np.array([0, 2], np.uint64), params, coeffs, raise_on_status=False)
    assert int(r["status"][0]) & SS_EMPTY_TU
    assert int(r["est_per_tu"][0]) == 0 and int(r["est_per_tu"][1]) == int(r["est_bits"][0]) > 0
    r = engine.estimate_residual_bits(np.zeros(0, TU_DT), np.zeros(1, np.uint64), np.zeros(0, PARAMS_DT),
                                      np.zeros(0, np.int32))
    assert r["est_bits"].size == 0


def test_warp_specialised_matches_fused_randomised(engine):
    """Hand-off stress: batch shapes around the CTA geometry (19 producers + 1 walker), every output of
    the warp-specialised kernel against the fused kernel's, repeated (tests/tools/ws_stress.py, bounded)."""
    rng = np.random.default_rng(77)
    try:
        for it in range(24):
            n_sub = int(rng.choice([1, 7, 19, 20, 148, 500, 3000]))
            flags = rng.choice([None, 1, 0, 1 | 8, 16, 2, 3])
            wl = W.random_substreams(n_sub, int(rng.integers(1 << 30)), flags=None if flags is None else int(flags),
                                     tus_per_sub=(1, int(rng.choice([2, 6, 30]))),
                                     ts_fraction=float(rng.choice([0.0, 0.2, 0.9])))
            cap = wl.out_capacity(bytes_per_coeff=5.0)
            engine.set_encoder_mode(1)
            a = engine.encode_residual(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, cap, want_ctx=True)
            engine.set_encoder_mode(3)
            for rep in range(2):
                b = engine.encode_residual(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, cap, want_ctx=True)
                assert (a["sizes"] == b["sizes"]).all() and a["out"].tobytes() == b["out"].tobytes(), (it, n_sub)
                assert (a["num_bins"] == b["num_bins"]).all() and (a["ctx_out"] == b["ctx_out"]).all()
                assert (a["status"] == b["status"]).all() and (a["tu_flags"] == b["tu_flags"]).all()
    finally:
        engine.set_encoder_mode(0)


def test_bench_scale_properties_all_encoders(engine):
    """BASELINE.json's configs[1] batch in full (64 frames 1080p, 8 640 CTU substreams, 199 M coefficient
    positions), host-pointer API: the size-independent properties the domain offers -
    decode(encode(x)) == x, bins conserved, every encoder kernel byte-identical (sizes + a checksum of
    the packed bitstream), and the bit estimate within 1 % of the bits actually written."""
    import zlib
    wl = W.frames_as_ctu_substreams(64, 1920, 1080, seed0=3000)
    assert wl.n_sub >= 5624
    cap = wl.out_capacity()
    ref = None
    try:
        for mode in (0, 1, 2, 3):
            engine.set_encoder_mode(mode)
            r = engine.encode_residual_packed(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, cap)
            assert not r["status"].any()
            sig = (zlib.crc32(r["packed"].tobytes()), zlib.crc32(r["sizes"].tobytes()), zlib.crc32(r["num_bins"].tobytes()),
                   zlib.crc32(r["tu_flags"].tobytes()))
            if ref is None:
                ref, first = sig, r
            assert sig == ref, f"encoder mode {mode} differs from AUTO"
    finally:
        engine.set_encoder_mode(0)
    r = first
    d = engine.decode_residual(r["packed"], r["entry_points"], wl.tus, wl.tu_offsets, wl.params, wl.coeffs.size)
    assert not d["status"].any()
    assert (d["coeffs"] == wl.coeffs).all() and (d["num_bins"] == r["num_bins"]).all()
    est = engine.estimate_residual_bits(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, want_per_tu=False)
    trm = (wl.params["flags"] & SUB_TERMINATE) != 0
    assert (est["num_bins"] + trm == r["num_bins"]).all()
    written = 8.0 * float(r["sizes"].astype(np.int64).sum())
    assert abs(float(est["est_bits"].astype(np.float64).sum()) / 32768.0 / written - 1.0) < 0.01


def test_million_substreams_round_trip():
    """BASELINE configs[4] upper end: 10^4, 10^5 and 10^6 substreams in one launch (a seeded base batch tiled on the
    device, tests/tools/substream_sweep.py): decode(encode(x)) == x on every tile and all tiles code identically."""
    import json
    import subprocess
    import sys
    tool = os.path.join(os.path.dirname(os.path.abspath(__file__)), "tools", "substream_sweep.py")
    out = subprocess.run([sys.executable, tool, "100"], capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stderr[-2000:]
    rows = [json.loads(l) for l in out.stdout.splitlines() if l.startswith("{")]
    assert [r["substreams"] for r in rows] == [10_000, 100_000, 1_000_000]
    for r in rows:
        assert r["round_trip_ok"] and r["tiles_identical"], r
        assert r["bins"] == rows[0]["bins"] * (r["substreams"] // 10_000)


def test_int16_coefficient_slabs(engine, oracle):
    """cabac_encode_residual_packed_i16 / cabac_decode_residual_i16: the slab crosses the ABI as int16; bytes,
    bin counts and decoded levels equal the int32 calls' (every encoder kernel; staged and zero-copy host paths,
    unaligned block starts, 64-wide blocks, ISP shapes)."""
    wl = W.random_substreams(200, seed=321, tus_per_sub=(1, 12), ts_fraction=0.2)
    # odd coefficient offsets: shift every block by a different small gap so that 2- / 4- / 16-byte paths all run
    tus = wl.tus.copy()
    gaps = (np.arange(len(tus)) % 5).astype(np.uint64)
    new_off = np.concatenate([[0], np.cumsum(gaps[:-1] + (1 << (tus["log2_w"].astype(np.uint64) + tus["log2_h"].astype(np.uint64)))[:-1])]).astype(np.uint64) + gaps[0]
    slab = np.zeros(int(new_off[-1]) + (1 << int(tus["log2_w"][-1] + tus["log2_h"][-1])) + 8, np.int32)
    for t, o in zip(wl.tus, new_off):
        a = 1 << int(t["log2_w"] + t["log2_h"])
        slab[int(o):int(o) + a] = wl.coeffs[int(t["coeff_offset"]):int(t["coeff_offset"]) + a]
    tus["coeff_offset"] = new_off
    cap = wl.out_capacity(bytes_per_coeff=5.0)
    slab16 = slab.astype(np.int16)
    ref = engine.encode_residual_packed(tus, wl.tu_offsets, wl.params, slab, cap)
    for mode in (0, 1, 2, 3, 4):
        engine.set_encoder_mode(mode)
        try:
            r = engine.encode_residual_packed(tus, wl.tu_offsets, wl.params, slab16, cap, i16=True)
        finally:
            engine.set_encoder_mode(0)
        assert r["packed"].tobytes() == ref["packed"].tobytes() and (r["num_bins"] == ref["num_bins"]).all(), mode
        assert (r["tu_flags"] == ref["tu_flags"]).all()
    d32 = engine.decode_residual(ref["packed"], ref["entry_points"], tus, wl.tu_offsets, wl.params, slab.size)
    d16 = engine.decode_residual(ref["packed"], ref["entry_points"], tus, wl.tu_offsets, wl.params, slab.size, i16=True)
    assert d16["coeffs"].dtype == np.int16 and (d16["coeffs"].astype(np.int32) == d32["coeffs"]).all()
    assert (d16["num_bins"] == d32["num_bins"]).all() and (d32["coeffs"] == slab).all()
    # the oracle on a sample, from the int16 call's bytes
    for i in range(0, wl.n_sub, 25):
        a, b = int(wl.tu_offsets[i]), int(wl.tu_offsets[i + 1])
        p = wl.params[i]
        eb = oracle.encode_residual(tus[a:b], slab, int(p["qp"]), int(p["init_id"]), int(p["flags"]))[0]
        e = ref["entry_points"]
        assert ref["packed"][int(e[i]):int(e[i + 1])].tobytes() == eb.tobytes()
    # zero-copy host path (pinned slabs >= 4 Mi coefficients): int16 in, int16 out
    big = W.frames_as_ctu_substreams(2, 1920, 1080, seed0=5150)
    capb = big.out_capacity()
    h16 = engine.pinned_copy(big.coeffs.astype(np.int16))
    rb32 = engine.encode_residual_packed(big.tus, big.tu_offsets, big.params, engine.pinned_copy(big.coeffs), capb)
    rb16 = engine.encode_residual_packed(big.tus, big.tu_offsets, big.params, h16, capb, i16=True)
    assert rb16["packed"].tobytes() == rb32["packed"].tobytes()
    out16 = engine.host_array(big.coeffs.size, np.int16)
    out16[:] = 0x5a5a
    engine.decode_residual(rb16["packed"], rb16["entry_points"], big.tus, big.tu_offsets, big.params, big.coeffs.size,
                           coeffs_out=out16, i16=True)
    assert (out16.astype(np.int32) == big.coeffs).all()


def test_sliced_encoder_long_substreams(engine, oracle):
    """CABAC_ENCODER_SLICED (few, long substreams; every stage parallel inside a substream): bytes, bin counts,
    final contexts and TU flags equal the fused kernel's and the oracle's - substreams of hundreds of
    segments, handed-in contexts, sign hiding, transform skip, one substream of a single TU, an output slab
    that is too small, and AUTO picking the path for a frame-sized slice."""
    wl = W.random_substreams(7, seed=77, tus_per_sub=(200, 900), ts_fraction=0.15)
    tiny = W.random_substreams(2, seed=78, tus_per_sub=(1, 1))
    # append the two one-TU substreams
    tus2 = tiny.tus.copy()
    tus2["coeff_offset"] += np.uint64(wl.coeffs.size)
    tus = np.concatenate([wl.tus, tus2])
    tu_off = np.concatenate([wl.tu_offsets, wl.tu_offsets[-1] + tiny.tu_offsets[1:]]).astype(np.uint64)
    params = np.concatenate([wl.params, tiny.params])
    params["flags"][1] = T | SUB_SIGN_HIDING
    params["flags"][2] = T
    params["init_id"][3] = 0
    coeffs = np.concatenate([wl.coeffs, tiny.coeffs])
    batch = W.Workload(tus, tu_off, params, coeffs)
    n = batch.n_sub
    ctx_in = np.zeros((n, 379), CTX_DT)
    ctx_in[4] = oracle.ctx_init(17, 1)[0]
    params["flags"][4] |= SUB_CTX_IN
    cap = batch.out_capacity(bytes_per_coeff=5.0)
    engine.set_encoder_mode(1)
    a = engine.encode_residual(tus, tu_off, params, coeffs, cap, ctx_in=ctx_in, want_ctx=True)
    engine.set_encoder_mode(4)
    try:
        b = engine.encode_residual(tus, tu_off, params, coeffs, cap, ctx_in=ctx_in, want_ctx=True)
        assert (a["sizes"] == b["sizes"]).all() and (a["num_bins"] == b["num_bins"]).all()
        for i in range(n):
            x = a["out"][int(cap[i]):int(cap[i]) + int(a["sizes"][i])]
            y = b["out"][int(cap[i]):int(cap[i]) + int(b["sizes"][i])]
            assert x.tobytes() == y.tobytes(), i
        assert (a["ctx_out"] == b["ctx_out"]).all() and (a["tu_flags"] == b["tu_flags"]).all()
        for i in (0, 4, n - 1):
            p = params[i]
            t0, t1 = int(tu_off[i]), int(tu_off[i + 1])
            eb, nb, _ = oracle.encode_residual(tus[t0:t1], coeffs, int(p["qp"]), int(p["init_id"]), int(p["flags"]),
                                               ctx_in=ctx_in[i] if i == 4 else None)
            assert b["out"][int(cap[i]):int(cap[i]) + int(b["sizes"][i])].tobytes() == eb.tobytes() and int(b["num_bins"][i]) == nb
        # an output slab that is too small is reported, nothing is written past it
        small = cap.copy()
        small[1:] = np.cumsum(np.maximum(64, (np.diff(cap.astype(np.int64)) // 40) // 16 * 16))
        c = engine.encode_residual(tus, tu_off, params, coeffs, small, ctx_in=ctx_in, raise_on_status=False)
        assert (c["status"][:7] & SS_OUT_OVERFLOW).all()
    finally:
        engine.set_encoder_mode(0)
    # AUTO: a frame-sized slice takes the sliced path (same bytes as the CTU-substream kernels on a sample check)
    big = W.frames_as_ctu_substreams(2, 1920, 1080, seed0=99)  # 2 slices of ~86 000 TUs
    slices = W.merge_substreams(big, big.n_sub // 2)
    capb = slices.out_capacity()
    l0 = engine.launch_count
    r0 = engine.encode_residual(slices.tus, slices.tu_offsets, slices.params, slices.coeffs, capb)
    assert engine.launch_count - l0 > 15  # the multi-kernel sliced pipeline, not one fused launch
    engine.set_encoder_mode(1)
    try:
        r1 = engine.encode_residual(slices.tus, slices.tu_offsets, slices.params, slices.coeffs, capb)
    finally:
        engine.set_encoder_mode(0)
    assert (r0["sizes"] == r1["sizes"]).all() and r0["out"].tobytes() == r1["out"].tobytes()
    d = engine.decode_residual(np.concatenate([r0["out"][int(capb[i]):int(capb[i]) + int(r0["sizes"][i])] for i in range(2)]),
                               offsets_of(r0["sizes"]), slices.tus, slices.tu_offsets, slices.params, slices.coeffs.size)
    assert (d["coeffs"] == slices.coeffs).all()


@pytest.mark.parametrize("seed,flags,kw", [
    (901, None, {"ts_fraction": 0.25, "bdpcm_fraction": 0.05, "sbt_fraction": 0.05}),
    (902, T | SUB_DEP_QUANT | SUB_SIGN_HIDING, {"ts_fraction": 0.3}),
    (903, T | SUB_DEP_QUANT | SUB_TSRC_DISABLED, {"ts_fraction": 0.4}),
    (904, T, {"preset": "qp0", "tus_per_sub": (1, 3)}),
])
def test_thread_decoder_matches_lockstep_and_oracle(engine, oracle, seed, flags, kw):
    """CABAC_DECODER_THREAD (one thread per substream, contexts in a global row): levels, bin counts, bits read,
    TU flags, final contexts and status words equal the lockstep decoder's; a sample against the oracle; int16
    slabs; a truncated substream reports the same status."""
    wl = W.random_substreams(300, seed, flags=flags, **kw)
    r = engine.encode_residual_packed(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, wl.out_capacity(bytes_per_coeff=5.0))
    ent = r["entry_points"]
    res = {}
    for mode in (1, 2):
        engine.set_decoder_mode(mode)
        try:
            res[mode] = engine.decode_residual(r["packed"], ent, wl.tus, wl.tu_offsets, wl.params, wl.coeffs.size,
                                               want_ctx=True)
            res[mode, 16] = engine.decode_residual(r["packed"], ent, wl.tus, wl.tu_offsets, wl.params, wl.coeffs.size,
                                                   i16=True)
        finally:
            engine.set_decoder_mode(0)
    a, b = res[1], res[2]
    for k in ("coeffs", "num_bins", "bits_read", "tu_flags", "status"):
        assert (a[k] == b[k]).all(), k
    assert (a["ctx_out"] == b["ctx_out"]).all() and (b["num_bins"] == r["num_bins"]).all()
    assert (res[2, 16]["coeffs"].astype(np.int32) == b["coeffs"]).all()
    for i in range(0, wl.n_sub, 29):
        t0, t1, lo, hi = sub_range(wl, i)
        p = wl.params[i]
        od = oracle.decode_residual(r["packed"][int(ent[i]):int(ent[i + 1])], wl.tus[t0:t1], wl.coeffs.size, int(p["qp"]),
                                    int(p["init_id"]), int(p["flags"]))
        assert (od[0][lo:hi] == b["coeffs"][lo:hi]).all() and od[1] == int(b["bits_read"][i])
    # every substream cut to half its bytes: same status words from both kernels
    cut = np.concatenate([[0], np.cumsum((np.diff(ent) + 1) // 2)]).astype(np.uint64)
    half = np.concatenate([r["packed"][int(ent[i]):int(ent[i]) + int(cut[i + 1] - cut[i])] for i in range(wl.n_sub)])
    st = {}
    for mode in (1, 2):
        engine.set_decoder_mode(mode)
        try:
            st[mode] = engine.decode_residual(half, cut, wl.tus, wl.tu_offsets, wl.params, wl.coeffs.size,
                                              raise_on_status=False)["status"]
        finally:
            engine.set_decoder_mode(0)
    assert (st[1] == st[2]).all() and st[2].any()


def test_malformed_descriptors_are_flagged_not_followed(engine):
    """A cabac_tu_desc outside the legal geometry (side > 64 samples, component > Cr: the reference CHECKs both) sets
    CABAC_SS_BAD_DESC on its substream in every kernel; the kernels clamp the fields instead of indexing with them,
    and every other substream of the batch codes exactly as in the clean batch."""
    from entropy_coding_b200.types import SS_BAD_DESC
    wl = W.random_substreams(60, seed=888, flags=T | SUB_DEP_QUANT, tus_per_sub=(3, 9), ts_fraction=0.2)
    cap = wl.out_capacity(bytes_per_coeff=5.0)
    clean = engine.encode_residual(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, cap)
    clean_outs = split(clean["out"], cap, clean["sizes"])
    tus = wl.tus.copy()
    bad = {7: ("log2_w", 9), 20: ("comp_id", 5), 33: ("log2_h", 255)}
    for sub, (field, val) in bad.items():
        t = int(wl.tu_offsets[sub]) + 1
        tus[field][t] = val
        tus["coeff_offset"][t] = wl.coeffs.size  # the clamped (up to 64 x 64) block lives in the padding below
    slab = np.concatenate([wl.coeffs, np.zeros(1 << 13, np.int32)])
    good = np.array([i not in bad for i in range(wl.n_sub)])
    for mode in (0, 1, 2, 3, 4):
        engine.set_encoder_mode(mode)
        try:
            r = engine.encode_residual(tus, wl.tu_offsets, wl.params, slab, cap, raise_on_status=False)
        finally:
            engine.set_encoder_mode(0)
        assert all(int(r["status"][i]) & SS_BAD_DESC for i in bad), mode
        assert not r["status"][good].any(), mode
        outs = split(r["out"], cap, r["sizes"])
        assert all(outs[i].tobytes() == clean_outs[i].tobytes() for i in range(wl.n_sub) if good[i]), mode
    est = engine.estimate_residual_bits(tus, wl.tu_offsets, wl.params, slab, raise_on_status=False)
    assert all(int(est["status"][i]) & SS_BAD_DESC for i in bad) and not est["status"][good].any()
    ioff = offsets_of(clean["sizes"])
    data = np.concatenate(clean_outs)
    for mode in (1, 2):
        engine.set_decoder_mode(mode)
        try:
            d = engine.decode_residual(data, ioff, tus, wl.tu_offsets, wl.params, slab.size, raise_on_status=False)
        finally:
            engine.set_decoder_mode(0)
        assert all(int(d["status"][i]) & SS_BAD_DESC for i in bad), mode
        assert not d["status"][good].any(), mode
        for i in range(wl.n_sub):
            if good[i]:
                a, b, lo, hi = sub_range(wl, i)
                assert (d["coeffs"][lo:hi] == wl.coeffs[lo:hi]).all(), (mode, i)
