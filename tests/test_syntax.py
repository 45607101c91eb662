"""the CTU-level syntax of all-intra substreams, decoded from the bytes alone.

The pin is the reference itself: tests/golden/vtm_*.npz hold, per substream of a VTM-produced stream,
every BinEncIf call the reference's CABACWriter made (`cmds`: the non-residual calls with a CABAC_CMD_TU
marker wherever it called residual_coding), the TU descriptors and TCoeff blocks behind the markers,
one record per coding_unit() call (`cus`), the slice parameters (`slice`), the final contexts, the bin
count and the payload.  A decoder that walks the payload must issue the mirror image of that call
sequence - the same context for every bin, the same EP bin counts, residual_coding at the same places -
and must come out with the same CUs, TUs, coefficients and contexts.

  CPU (-m "not gpu"): the product's parser header (csrc/cabac_syntax.cuh) compiled for the host on the
                      oracle's scalar arithmetic decoder (oracle/syntax_host.cpp) produces exactly that
                      trace; context-set offsets equal the compiled reference's.
  GPU (-m gpu)      : cabac_decode_slices (decode_slice_kernel, one warp per substream) returns the same
                      records, coefficients, contexts, bin counts and bits read, for all substreams of
                      all fixtures in one launch.
"""
import ctypes as C
import glob
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.join(HERE, "..")
sys.path.insert(0, os.path.join(HERE, "golden"))
from vtm_streams import load_fixture  # noqa: E402

from entropy_coding_b200.types import (CTU_DT, CTX_DT, CU_CCLM, CU_DT, CU_FROM_DIR, CU_MIP, CU_MIP_TRANSPOSED, CU_MPM, PARAMS_DT, SL_WPP, SLICE_DT,
                                       SS_BAD_DESC, SS_CAPACITY, SS_UNSUPPORTED, SUB_ALIGN, SUB_CTX_IN, SUB_TERMINATE, TU_DT)

FIXTURES = sorted(os.path.basename(p) for p in glob.glob(os.path.join(HERE, "golden", "vtm_*.npz")))


def _merge_rows(rows):
    """The CTU-row substreams of one WPP slice back to back, as they lie in the slice data."""
    m = dict(rows[0])
    for k in ("payload", "cmds", "coeffs", "cus"):
        m[k] = np.concatenate([r[k] for r in rows])
    tus, base = [], 0
    for r in rows:
        t = r["tus"].copy()
        t["coeff_offset"] += base
        base += r["coeffs"].size
        tus.append(t)
    m["tus"] = np.concatenate(tus)
    m["num_bins"] = sum(r["num_bins"] for r in rows)
    m["ctx_out"] = rows[-1]["ctx_out"]
    m["row_bytes"] = [int(r["payload"].size) for r in rows]
    sp = rows[0]["slice"].astype(np.uint32).copy()
    sp[5] = rows[-1]["slice"][5]  # y1
    m["slice"] = sp
    return m


def _substreams():
    """Every fixture substream the built scope covers: slices (whole frames); of the WPP fixture each frame as ONE
    slice of three CTU-row substreams (contexts handed over after the first CTU of the row above), and the first
    row of each frame on its own as a plain 64x64-CTU slice."""
    out = []
    for name in FIXTURES:
        subs, _ = load_fixture(name)
        subs = [s for s in subs if "slice" in s and len(s["slice"])]
        i = 0
        while i < len(subs):
            s = subs[i]
            sp = s["slice"].astype(np.uint32).copy()
            if int(sp[30]) & SL_WPP:
                j = i + 1
                while j < len(subs) and int(subs[j]["slice"][3]) != 0:
                    j += 1
                m = _merge_rows(subs[i:j])
                out.append((f"{name}:{i}-{j - 1}(wpp)", m, m["slice"].view(SLICE_DT)[0]))
                sp[30] &= ~np.uint32(SL_WPP)
                out.append((f"{name}:{i}(row0)", s, sp.view(SLICE_DT)[0]))
                i = j
            else:
                out.append((f"{name}:{i}", s, sp.view(SLICE_DT)[0]))
                i += 1
    return out


def _flat(cmds):
    """Command words -> one token per bin (EP words are split differently by writer and reader)."""
    out = []
    for w in cmds:
        w = int(w)
        k = w >> 29
        if k == 0:
            out.append(("ctx", w & 511, (w >> 9) & 1))
        elif k == 1:
            n, v = (w >> 24) & 31, w & 0xFFFFFF
            out.extend(("ep", (v >> i) & 1) for i in range(n - 1, -1, -1))
        elif k == 3:
            out.append(("trm", w & 1))
        elif k == 5:
            out.append(("tu",))
        else:
            out.append(("other", w))
    return out


def _host_lib():
    path = os.path.join(ROOT, "oracle", "libsyntax_host.so")
    if not os.path.exists(path):
        import subprocess
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "oracle", "syntax"])
    lib = C.CDLL(path)
    lib.syn_decode_slice.restype = C.c_long
    return lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _host_decode(lib, s, sp):
    params = np.array([(s["qp"], s["init_id"], s["flags"] | SUB_TERMINATE | SUB_CTX_IN, 0)], PARAMS_DT)
    ctx_in = np.zeros(379, CTX_DT)
    ctx_in["state0"], ctx_in["state1"] = s["ctx_in"][:, 0], s["ctx_in"][:, 1]
    area = int(sp["x1"] - sp["x0"]) * int(sp["y1"] - sp["y0"])
    cap_cu = cap_tu = area // 16 * 2 + 8
    cap_co = area * 3 // 2 + 64
    cus, tus, tuf = np.zeros(cap_cu, CU_DT), np.zeros(cap_tu, TU_DT), np.zeros(cap_tu, np.uint8)
    ctus, coeffs = np.zeros(64, CTU_DT), np.zeros(cap_co, np.int32)
    counts, nb, st = np.zeros(3, np.uint32), np.zeros(1, np.uint32), np.zeros(1, np.uint32)
    ctx_out, trace, nt = np.zeros(379, CTX_DT), np.zeros(1 << 20, np.uint32), C.c_long(0)
    payload = np.ascontiguousarray(s["payload"])
    spa = np.array([sp], SLICE_DT)
    bits = lib.syn_decode_slice(_p(payload), C.c_long(payload.size), _p(params), _p(spa), _p(ctx_in), _p(cus),
                                C.c_long(cap_cu), _p(tus), _p(tuf), C.c_long(cap_tu), _p(ctus), C.c_long(64), _p(coeffs),
                                C.c_long(cap_co), _p(counts), _p(nb), _p(ctx_out), _p(st), _p(trace),
                                C.c_long(trace.size), C.byref(nt))
    return dict(cus=cus[:counts[0]], tus=tus[:counts[1]], tu_flags=tuf[:counts[1]], coeffs=coeffs[:counts[2]],
                ctus=ctus, num_bins=int(nb[0]), status=int(st[0]), ctx_out=ctx_out, trace=trace[:nt.value], bits=int(bits))


def _check_against_fixture(tag, r, s):
    """CUs, TUs, coefficients, contexts and bin count of one decoded substream vs what the reference's writer recorded."""
    assert r["status"] == 0, f"{tag}: status {r['status']}"
    assert r["num_bins"] == s["num_bins"], tag
    t, g = r["tus"], s["tus"]
    assert len(t) == len(g), f"{tag}: {len(t)} TUs, the writer coded {len(g)}"
    for f in ("log2_w", "log2_h", "comp_id", "flags"):
        assert (t[f] == g[f]).all(), f"{tag}: TU {f}"
    assert r["coeffs"].size == s["coeffs"].size and (r["coeffs"] == s["coeffs"]).all(), f"{tag}: coefficients"
    assert (r["ctx_out"]["state0"] == s["ctx_out"][:, 0]).all() and (r["ctx_out"]["state1"] == s["ctx_out"][:, 1]).all(), tag
    c, w = r["cus"], s["cus"]
    assert len(c) == len(w), f"{tag}: {len(c)} CUs, the writer coded {len(w)}"
    for f, g_ in (("x", "x"), ("y", "y"), ("log2_w", "log2w"), ("log2_h", "log2h"), ("ch_type", "ch_type"),
                  ("qt_depth", "qt_depth"), ("mt_depth", "mt_depth"), ("bdpcm", "bdpcm"), ("mrl_idx", "mrl"),
                  ("isp", "isp"), ("lfnst_idx", "lfnst")):
        assert (c[f] == w[g_]).all(), f"{tag}: CU field {f} at {np.nonzero(c[f] != w[g_])[0][:4]}"
    # mtsIdx[Y] of the first TU: the reader only sets it inside residual_coding (ts_flag, cabac_reader.cpp:2737-2752),
    # so a BDPCM CU without coded luma keeps DCT2 there while the encoder's object says "skip": a don't-care
    care = ~((c["bdpcm"] > 0) & (c["n_tu"] == 0))
    assert (c["mts_idx"][care] == w["mts"][care]).all(), f"{tag}: CU field mts_idx"
    both = CU_MIP | CU_MIP_TRANSPOSED
    assert ((c["flags"] & both) == (w["flags"] & both)).all(), f"{tag}: MIP flags"
    # pu.intraDir[chType] as the reference's objects held it: MPM-derived luma directions, MIP modes, chroma modes
    assert (c["intra_dir"] == w["dir"]).all(), \
        f"{tag}: intra direction at CUs {np.nonzero(c['intra_dir'] != w['dir'])[0][:6]}: " \
        f"{c['intra_dir'][c['intra_dir'] != w['dir']][:6]} vs {w['dir'][c['intra_dir'] != w['dir']][:6]}"
    # the stop bit follows the terminating bin: everything but the alignment was consumed
    assert 0 <= s["payload"].size * 8 - r["bits"] <= 8, f"{tag}: bits read {r['bits']} of {s['payload'].size * 8}"


def test_syntax_fixtures_present():
    tags = [t for t, _, _ in _substreams()]
    assert any(t.startswith("vtm_qp51") for t in tags) and any(t.startswith("vtm_qp32") for t in tags), tags
    assert any(t.startswith("vtm_wpp") for t in tags), "CTU-64 rows"


def test_syntax_ctx_offsets_equal_reference(ref):
    lib = _host_lib()
    mine, theirs = np.zeros(29, np.uint32), np.zeros(40, np.uint32)
    lib.syn_ctx_offsets(_p(mine))
    n = ref.lib.ref_syntax_ctx_offsets(_p(theirs))
    assert n == 29 and (mine == theirs[:29]).all()


def test_host_parser_issues_the_writers_call_sequence():
    lib = _host_lib()
    seen_ctx = set()
    n_cus = n_tus = 0
    for tag, s, sp in _substreams():
        r = _host_decode(lib, s, sp)
        got, want = _flat(r["trace"]), _flat(s["cmds"])
        bad = next((k for k in range(min(len(got), len(want))) if got[k] != want[k]), None)
        assert bad is None and len(got) == len(want), \
            f"{tag}: call {bad}: parser {got[bad - 2:bad + 3] if bad is not None else len(got)} " \
            f"writer {want[bad - 2:bad + 3] if bad is not None else len(want)}"
        _check_against_fixture(tag, r, s)
        seen_ctx |= {t[1] for t in got if t[0] == "ctx"}
        n_cus += len(r["cus"])
        n_tus += len(r["tus"])
    # coverage note: which non-residual context sets the fixtures exercise
    assert n_cus > 200 and n_tus > 200, (n_cus, n_tus)
    for base, what in ((0, "split_cu_flag"), (9, "split_qt_flag"), (15, "mtt_split_cu_vertical_flag"),
                       (20, "mtt_split_cu_binary_flag"), (46, "intra_mip_flag"), (40, "intra_luma_mpm_flag"),
                       (43, "cclm_mode_flag"), (45, "intra_chroma_pred_mode"), (38, "intra_luma_ref_idx"),
                       (316, "intra_subpartitions_mode_flag"), (77, "tu_y_coded_flag"), (81, "tu_cb_coded_flag"),
                       (83, "tu_cr_coded_flag"), (354, "tu_joint_cbcr_residual_flag"), (312, "mts_idx"),
                       (296, "lfnst_idx"), (294, "sao_merge"), (295, "sao_type_idx"), (332, "alf_ctb_flag")):
        assert any(base <= c < base + 3 for c in seen_ctx), f"no fixture substream codes {what}"


def test_host_parser_refuses_what_is_out_of_scope():
    lib = _host_lib()
    tag, s, sp = _substreams()[0]
    for bit in (20, 21, 22):  # IBC, palette, ACT
        sp2 = sp.copy()
        sp2["tools"] |= np.uint32(1 << bit)
        assert _host_decode(lib, s, sp2)["status"] & SS_UNSUPPORTED
    sp2 = sp.copy()
    sp2["tools"] &= ~np.uint32(1 << 23)  # P / B slice
    assert _host_decode(lib, s, sp2)["status"] & SS_UNSUPPORTED
    sp2 = sp.copy()
    sp2["tools"] &= ~np.uint32(1)  # single tree
    assert _host_decode(lib, s, sp2)["status"] & SS_UNSUPPORTED


@pytest.mark.gpu
def test_gpu_decode_slices_equals_reference_records(engine):
    """All fixture substreams in one launch through the C ABI, from the payload bytes alone."""
    items = _substreams()
    lib = _host_lib()
    n = len(items)
    params = np.zeros(n, PARAMS_DT)
    sps = np.zeros(n, SLICE_DT)
    ctx_in = np.zeros((n, 379), CTX_DT)
    data, off = [], [0]
    for k, (tag, s, sp) in enumerate(items):
        params[k] = (s["qp"], s["init_id"], s["flags"] | SUB_TERMINATE | SUB_CTX_IN, 0)
        sps[k] = sp
        ctx_in[k]["state0"], ctx_in[k]["state1"] = s["ctx_in"][:, 0], s["ctx_in"][:, 1]
        data.append(s["payload"])
        off.append(off[-1] + s["payload"].size)
    before = engine.launch_count
    r = engine.decode_slices(np.concatenate(data), np.asarray(off, np.uint64), params, sps, ctx_in=ctx_in, want_ctx=True)
    assert engine.launch_count > before
    cap = r["cap"]
    for k, (tag, s, sp) in enumerate(items):
        ncu, ntu, nco = (int(v) for v in r["counts"][k])
        c0, t0, k0 = int(cap["cu"][k]), int(cap["tu"][k]), int(cap["coeff"][k])
        tus = r["tus"][t0:t0 + ntu].copy()
        assert (tus["coeff_offset"] >= k0).all() and (tus["coeff_offset"] < int(cap["coeff"][k + 1])).all(), tag
        one = dict(cus=r["cus"][c0:c0 + ncu], tus=tus, tu_flags=r["tu_flags"][t0:t0 + ntu], coeffs=r["coeffs"][k0:k0 + nco],
                   num_bins=int(r["num_bins"][k]), status=int(r["status"][k]), ctx_out=r["ctx_out"][k],
                   bits=int(r["bits_read"][k]))
        _check_against_fixture(tag, one, s)
        # and field by field against the host run of the same parser (records the fixtures do not hold)
        h = _host_decode(lib, s, sp)
        assert one["cus"].tobytes() == h["cus"].tobytes(), f"{tag}: CU records differ from the host run"
        assert (one["tu_flags"] == h["tu_flags"]).all(), tag
        nctu = int(cap["ctu"][k + 1] - cap["ctu"][k])
        assert r["ctus"][int(cap["ctu"][k]):int(cap["ctu"][k + 1])].tobytes() == h["ctus"][:nctu].tobytes(), f"{tag}: CTU records"
        assert one["bits"] == h["bits"], tag


@pytest.mark.gpu
def test_gpu_decode_slices_many_copies_and_errors(engine):
    """4096 substreams in one launch (every fixture slice many times over): identical results per copy;
    a truncated payload and an out-of-scope tool are flagged per substream, the rest of the batch is unharmed."""
    items = _substreams()
    reps = 4096 // len(items) + 1
    order = [k % len(items) for k in range(reps * len(items))][:4096]
    n = len(order)
    params, sps, ctx_in = np.zeros(n, PARAMS_DT), np.zeros(n, SLICE_DT), np.zeros((n, 379), CTX_DT)
    data, off = [], [0]
    for j, k in enumerate(order):
        tag, s, sp = items[k]
        params[j] = (s["qp"], s["init_id"], s["flags"] | SUB_TERMINATE | SUB_CTX_IN, 0)
        sps[j] = sp
        ctx_in[j]["state0"], ctx_in[j]["state1"] = s["ctx_in"][:, 0], s["ctx_in"][:, 1]
        pl = s["payload"]
        if j == 5:
            pl = pl[:pl.size // 2]  # truncated
        data.append(pl)
        off.append(off[-1] + pl.size)
    sps[7]["tools"] |= np.uint32(1 << 20)  # IBC
    r = engine.decode_slices(np.concatenate(data), np.asarray(off, np.uint64), params, sps, ctx_in=ctx_in,
                             raise_on_status=False)
    assert r["status"][7] & SS_UNSUPPORTED and r["status"][5] != 0
    cap = r["cap"]
    for j, k in enumerate(order):
        if j in (5, 7):
            continue
        tag, s, sp = items[k]
        assert r["status"][j] == 0 and int(r["num_bins"][j]) == s["num_bins"], f"copy {j} of {tag}"
        ncu, ntu, nco = (int(v) for v in r["counts"][j])
        assert ntu == len(s["tus"]) and ncu == len(s["cus"])
        k0 = int(cap["coeff"][j])
        assert (r["coeffs"][k0:k0 + nco] == s["coeffs"]).all(), f"copy {j} of {tag}"


def _garbage_cases(n, seed):
    """Random bytes and corrupted real payloads under real slice parameters."""
    rng = np.random.default_rng(seed)
    items = _substreams()
    cases = []
    for k in range(n):
        tag, s, sp = items[k % len(items)]
        pl = s["payload"].copy()
        kind = k % 3
        if kind == 0:
            pl = rng.integers(0, 256, int(rng.integers(2, 400)), dtype=np.uint8)
        elif kind == 1:
            pos = rng.integers(0, pl.size, 4)
            pl[pos] ^= rng.integers(1, 256, 4, dtype=np.uint8)
        else:
            pl = pl[:int(rng.integers(1, pl.size))]
        cases.append((s, sp, pl))
    return cases


def test_host_parser_survives_garbage():
    """Malformed slice data must end in status bits or a (wrong but bounded) parse - never in a crash or a loop."""
    lib = _host_lib()
    for s, sp, pl in _garbage_cases(120, 7):
        s2 = dict(s)
        s2["payload"] = pl
        r = _host_decode(lib, s2, sp)
        assert r["num_bins"] < 10_000_000


@pytest.mark.gpu
@pytest.mark.timeout(120)
def test_gpu_decode_slices_survives_garbage(engine):
    """The same on the device: 600 malformed substreams in one launch return (no hang, no fault), every region stays
    inside its capacity, and a well-formed batch decoded right after is still exact."""
    cases = _garbage_cases(600, 11)
    n = len(cases)
    params, sps, ctx_in = np.zeros(n, PARAMS_DT), np.zeros(n, SLICE_DT), np.zeros((n, 379), CTX_DT)
    data, off = [], [0]
    for j, (s, sp, pl) in enumerate(cases):
        params[j] = (s["qp"], s["init_id"], s["flags"] | SUB_TERMINATE | SUB_CTX_IN, 0)
        sps[j] = sp
        ctx_in[j]["state0"], ctx_in[j]["state1"] = s["ctx_in"][:, 0], s["ctx_in"][:, 1]
        data.append(pl)
        off.append(off[-1] + pl.size)
    r = engine.decode_slices(np.concatenate(data), np.asarray(off, np.uint64), params, sps, ctx_in=ctx_in,
                             raise_on_status=False)
    cap = r["cap"]
    for j in range(n):
        ncu, ntu, nco = (int(v) for v in r["counts"][j])
        assert ncu <= int(cap["cu"][j + 1] - cap["cu"][j]) or r["status"][j] != 0
        assert ntu <= int(cap["tu"][j + 1] - cap["tu"][j]) and nco <= int(cap["coeff"][j + 1] - cap["coeff"][j])
    assert (r["status"] != 0).sum() > n // 2  # most corrupted substreams are caught (overrun / no terminating bin)
    tag, s, sp = _substreams()[0]
    par = np.array([(s["qp"], s["init_id"], s["flags"] | SUB_TERMINATE | SUB_CTX_IN, 0)], PARAMS_DT)
    ci = np.zeros((1, 379), CTX_DT)
    ci[0]["state0"], ci[0]["state1"] = s["ctx_in"][:, 0], s["ctx_in"][:, 1]
    ok = engine.decode_slices(s["payload"], np.array([0, s["payload"].size], np.uint64), par, np.array([sp], SLICE_DT), ctx_in=ci)
    assert int(ok["num_bins"][0]) == s["num_bins"] and (ok["coeffs"][:s["coeffs"].size] == s["coeffs"]).all()


# ---- generated streams: the parser on an engine that picks every bin (oracle/syntax_host.cpp: syn_generate_slice) ----
# Not conformance (generator and parser share the walk - conformance is the VTM fixtures above): symmetry, capacity,
# every CTU size and picture shape, and the tools no fixture exercises (BDPCM, cu_qp_delta, chroma QP offsets, ALF
# chroma / CC-ALF CTU syntax, 32-sample maximum transform size, WPP rows at every CTU size).
from entropy_coding_b200.types import (SL_ALF_CB, SL_ALF_CR, SL_ALF_Y, SL_BDPCM, SL_CCALF_CB, SL_CCALF_CR, SL_CCLM,  # noqa: E402
                                       SL_CHROMA_QP_ADJ, SL_DQP, SL_DUAL_TREE, SL_INTRA, SL_ISP, SL_JCCR, SL_LFNST, SL_MIP,
                                       SL_MRL, SL_MTS, SL_MTS_INTRA, SL_SAO_CHROMA, SL_SAO_LUMA, SL_TS, SUB_DEP_QUANT)

ALL_TOOLS = (SL_DUAL_TREE | SL_SAO_LUMA | SL_SAO_CHROMA | SL_ALF_Y | SL_ALF_CB | SL_ALF_CR | SL_CCALF_CB | SL_CCALF_CR |
             SL_MTS_INTRA | SL_LFNST | SL_ISP | SL_MIP | SL_MRL | SL_TS | SL_BDPCM | SL_JCCR | SL_CCLM | SL_DQP |
             SL_CHROMA_QP_ADJ | SL_INTRA | SL_MTS)


def make_slice_params(w, h, log2_ctu=7, tools=ALL_TOOLS, **kw):
    sp = np.zeros(1, SLICE_DT)[0]
    sp["pic_w"], sp["pic_h"], sp["x1"], sp["y1"], sp["log2_ctu"] = w, h, w, h, log2_ctu
    mx = min(log2_ctu, 6)
    sp["log2_min_qt_luma"] = sp["log2_min_qt_chroma"] = 3
    sp["log2_max_bt_luma"], sp["log2_max_bt_chroma"] = min(5, mx), mx
    sp["log2_max_tt_luma"] = sp["log2_max_tt_chroma"] = min(5, mx)
    sp["max_mtt_depth_luma"] = sp["max_mtt_depth_chroma"] = 3
    for f in ("log2_min_bt_luma", "log2_min_bt_chroma", "log2_min_tt_luma", "log2_min_tt_chroma"):
        sp[f] = 2
    sp["log2_max_tb"], sp["log2_max_ts"], sp["bit_depth"], sp["chroma_format"] = 6, 5, 10, 1
    sp["alf_num_aps_luma"], sp["alf_chroma_num_alts"] = 2, 3
    sp["ccalf_filter_count_cb"], sp["ccalf_filter_count_cr"], sp["chroma_qp_offset_list_len"] = 3, 2, 3
    sp["tools"] = tools
    for k, v in kw.items():
        sp[k] = v
    return sp


def generate_slice(lib, seed, sp, qp=32, split_p=0.35, density=0.08):
    """One legal slice for `sp` and what the generating walk recorded (CUs, TUs, CTUs, coefficients, contexts)."""
    lib.syn_generate_slice.restype = C.c_long
    area = int(sp["x1"] - sp["x0"]) * int(sp["y1"] - sp["y0"])
    cap, capco = area // 16 * 2 + 8, area * 3 // 2 + 64
    ctu = 1 << int(sp["log2_ctu"])
    nctu = -(-int(sp["x1"] - sp["x0"]) // ctu) * -(-int(sp["y1"] - sp["y0"]) // ctu)
    par = np.array([(qp, 2, SUB_DEP_QUANT | SUB_TERMINATE, 0)], PARAMS_DT)
    out, cus, tus, tuf = np.zeros(area * 4 + 4096, np.uint8), np.zeros(cap, CU_DT), np.zeros(cap, TU_DT), np.zeros(cap, np.uint8)
    ctus, co, cnt, nb = np.zeros(nctu, CTU_DT), np.zeros(capco, np.int32), np.zeros(3, np.uint32), np.zeros(1, np.uint32)
    cout, st = np.zeros(379, CTX_DT), np.zeros(1, np.uint32)
    n = lib.syn_generate_slice(C.c_uint64(seed), C.c_double(split_p), C.c_double(density), _p(par), _p(np.array([sp], SLICE_DT)),
                               _p(out), C.c_long(out.size), _p(cus), C.c_long(cap), _p(tus), _p(tuf), C.c_long(cap), _p(ctus),
                               C.c_long(nctu), _p(co), C.c_long(capco), _p(cnt), _p(nb), _p(cout), _p(st))
    assert int(st[0]) == 0 and 0 < n <= out.size, f"generator status {int(st[0])}"
    init = np.zeros((379, 2), np.uint16)  # placeholder: generated slices start from Ctx::init (no SUB_CTX_IN)
    return dict(payload=out[:n].copy(), cus=cus[:cnt[0]].copy(), tus=tus[:cnt[1]].copy(), tu_flags=tuf[:cnt[1]].copy(),
                coeffs=co[:cnt[2]].copy(), ctus=ctus, num_bins=int(nb[0]), ctx_out=cout, qp=qp, init_id=2,
                flags=SUB_DEP_QUANT, ctx_in=init, fresh=True)


GEN_CASES = [  # (w, h, log2_ctu, extra slice parameters, WPP)
    (176, 144, 7, {}, False), (416, 240, 7, dict(log2_max_tb=5), True), (200, 120, 6, dict(cu_qp_delta_subdiv=2), False),
    (96, 80, 5, dict(cu_chroma_qp_offset_subdiv=2), True), (640, 360, 7, dict(cu_qp_delta_subdiv=4, log2_max_ts=2), False),
    (132, 68, 6, dict(tools=ALL_TOOLS & ~(SL_BDPCM | SL_DQP | SL_ALF_CB | SL_CCALF_CR)), False),
    (64, 64, 6, dict(tools=SL_DUAL_TREE | SL_INTRA), True), (3840, 136, 7, {}, False),
]


def _gen_items(lib, reps=1):
    items = []
    for rep in range(reps):
        for k, (w, h, l2, kw, wpp) in enumerate(GEN_CASES):
            sp = make_slice_params(w, h, l2, **kw)
            if wpp:
                sp["tools"] |= np.uint32(SL_WPP)
            items.append((f"gen{rep}:{w}x{h}/ctu{1 << l2}", generate_slice(lib, 1000 + 31 * rep + k, sp), sp))
    return items


def _host_decode_fresh(lib, s, sp):
    """_host_decode without handed-over contexts (Ctx::init from qp / init table)."""
    s2 = dict(s)
    r = None
    from oracle.checker import Checker
    init, _ = Checker("oracle").ctx_init(s["qp"], s["init_id"])
    s2["ctx_in"] = np.stack([init["state0"], init["state1"]], axis=1)
    r = _host_decode(lib, s2, sp)
    return r


def _check_against_generator(tag, r, g):
    assert r["status"] == 0 and r["num_bins"] == g["num_bins"], f"{tag}: status {r['status']}, bins {r['num_bins']} / {g['num_bins']}"
    assert r["cus"].tobytes() == g["cus"].tobytes(), f"{tag}: CU records"
    t, w = r["tus"], g["tus"]
    assert len(t) == len(w) and all((t[f] == w[f]).all() for f in ("log2_w", "log2_h", "comp_id", "flags")), f"{tag}: TUs"
    assert (r["tu_flags"] == g["tu_flags"]).all(), f"{tag}: TU result flags"
    assert r["coeffs"].size == g["coeffs"].size and (r["coeffs"] == g["coeffs"]).all(), f"{tag}: coefficients"
    assert r["ctx_out"].tobytes() == g["ctx_out"].tobytes(), f"{tag}: final contexts"


def test_generated_streams_decode_on_host():
    lib = _host_lib()
    for tag, g, sp in _gen_items(lib):
        r = _host_decode_fresh(lib, g, sp)
        _check_against_generator(tag, r, g)
        n = len(g["ctus"])
        assert r["ctus"][:n].tobytes() == g["ctus"].tobytes(), f"{tag}: CTU records"
        assert 0 <= g["payload"].size * 8 - r["bits"] <= 8, tag
    # capacity: a region that is too small ends in CABAC_SS_CAPACITY, not in a write beyond it
    tag, g, sp = _gen_items(lib)[0]
    sp_small = sp.copy()
    sp_small["x1"] = sp_small["pic_w"] = 64  # claims a smaller rectangle than the stream codes: regions shrink with it
    assert _host_decode_fresh(lib, g, sp_small)["status"] != 0


@pytest.mark.gpu
@pytest.mark.timeout(300)
def test_gpu_decode_slices_on_generated_streams(engine):
    """Every CTU size, odd picture shapes, a 3840-wide strip, WPP rows, and the tools no VTM fixture exercises, three
    seeds each, in one launch: records, coefficients, contexts and bin counts equal the generating walk's."""
    lib = _host_lib()
    items = _gen_items(lib, reps=3)
    n = len(items)
    params, sps = np.zeros(n, PARAMS_DT), np.zeros(n, SLICE_DT)
    data, off = [], [0]
    for k, (tag, g, sp) in enumerate(items):
        params[k] = (g["qp"], g["init_id"], g["flags"] | SUB_TERMINATE, 0)
        sps[k] = sp
        data.append(g["payload"])
        off.append(off[-1] + g["payload"].size)
    r = engine.decode_slices(np.concatenate(data), np.asarray(off, np.uint64), params, sps, want_ctx=True)
    cap = r["cap"]
    for k, (tag, g, sp) in enumerate(items):
        ncu, ntu, nco = (int(v) for v in r["counts"][k])
        c0, t0, k0 = int(cap["cu"][k]), int(cap["tu"][k]), int(cap["coeff"][k])
        one = dict(cus=r["cus"][c0:c0 + ncu], tus=r["tus"][t0:t0 + ntu], tu_flags=r["tu_flags"][t0:t0 + ntu],
                   coeffs=r["coeffs"][k0:k0 + nco], num_bins=int(r["num_bins"][k]), status=int(r["status"][k]),
                   ctx_out=r["ctx_out"][k])
        _check_against_generator(tag, one, g)
        a, b = int(cap["ctu"][k]), int(cap["ctu"][k + 1])
        assert r["ctus"][a:b].tobytes() == g["ctus"].tobytes(), f"{tag}: CTU records"


# ---- writer direction: records -> the writer's call sequence (syn_emit_slice on the host, cabac_encode_slices on the GPU)
def _from_dir(cus):
    """The records an encoder has: decisions (intra_dir), no syntax element values for the prediction mode."""
    cus = cus.copy()
    cus["flags"] = (cus["flags"] & ~np.uint8(CU_MPM | CU_CCLM)) | np.uint8(CU_FROM_DIR)
    cus["pred_syntax"] = 0xAA
    return cus


def _host_emit(lib, sp, r, from_dir=False):
    lib.syn_emit_slice.restype = C.c_long
    cmds, st = np.zeros(1 << 20, np.uint32), np.zeros(1, np.uint32)
    cus, tus, tuf, ctus = (np.ascontiguousarray(r[k]) for k in ("cus", "tus", "tu_flags", "ctus"))
    if from_dir:
        cus = _from_dir(cus)
    n = lib.syn_emit_slice(_p(np.array([sp], SLICE_DT)), _p(cus), C.c_long(len(cus)), _p(tus), _p(tuf), C.c_long(len(tus)),
                           _p(ctus), C.c_long(len(ctus)), _p(cmds), C.c_long(cmds.size), _p(st))
    return cmds[:n], int(st[0])


def test_host_writer_direction_reproduces_the_writers_calls():
    """Decode a VTM slice to records, then walk the records in the writer direction: the emitted BinEncIf call
    sequence is the one the reference's CABACWriter made (so cabac_encode_stream turns it into VTM's bytes)."""
    lib = _host_lib()
    n = 0
    for tag, s, sp in _substreams():
        r = _host_decode(lib, s, sp)
        nctu = -(-int(sp["x1"] - sp["x0"]) >> int(sp["log2_ctu"])) * -(-int(sp["y1"] - sp["y0"]) >> int(sp["log2_ctu"]))
        r["ctus"] = r["ctus"][:nctu]
        cmds, st = _host_emit(lib, sp, r)
        assert st == 0, f"{tag}: emit status {st}"
        got, want = _flat(cmds), _flat(s["cmds"])
        bad = next((k for k in range(min(len(got), len(want))) if got[k] != want[k]), None)
        assert bad is None and len(got) == len(want), f"{tag}: call {bad}: emitted {got[bad - 2:bad + 3] if bad is not None else len(got)} writer {want[bad - 2:bad + 3] if bad is not None else len(want)}"
        # and from the decisions alone (CABAC_CU_FROM_DIR: intra_dir in, MPM index / remainder / chroma candidate derived)
        cmds2, st2 = _host_emit(lib, sp, r, from_dir=True)
        assert st2 == 0 and np.array_equal(cmds2, cmds), f"{tag}: from intra_dir"
        n += 1
    assert n >= 6
    # and on generated slices (all tools): emit(decode(x)) is the generating walk's sequence again
    for tag, g, sp in _gen_items(lib):
        r = _host_decode_fresh(lib, g, sp)
        r["ctus"] = r["ctus"][:len(g["ctus"])]
        cmds, st = _host_emit(lib, sp, r)
        assert st == 0 and _flat(cmds) == _flat(r["trace"]), tag
        cmds2, st2 = _host_emit(lib, sp, r, from_dir=True)
        assert st2 == 0 and np.array_equal(cmds2, cmds), f"{tag}: from intra_dir"


def _compact(r, n):
    """decode_slices output (capacity-spaced regions) -> exact-offset records for the writer direction."""
    cap, cnt = r["cap"], r["counts"]
    cus = np.concatenate([r["cus"][int(cap["cu"][k]):int(cap["cu"][k]) + int(cnt[k][0])] for k in range(n)])
    tus = np.concatenate([r["tus"][int(cap["tu"][k]):int(cap["tu"][k]) + int(cnt[k][1])] for k in range(n)])
    tuf = np.concatenate([r["tu_flags"][int(cap["tu"][k]):int(cap["tu"][k]) + int(cnt[k][1])] for k in range(n)])
    off = lambda col: np.concatenate([[0], np.cumsum(cnt[:, col])]).astype(np.uint64)  # noqa: E731
    return dict(cus=cus, cu_offsets=off(0), tus=tus, tu_flags=tuf, tu_offsets=off(1), ctus=r["ctus"], ctu_offsets=cap["ctu"],
                coeffs=r["coeffs"])


@pytest.mark.gpu
def test_gpu_encode_slices_reproduces_the_reference_payloads(engine):
    """The writer direction through the C ABI: records (from cabac_decode_slices of the VTM payloads) in, bytes out -
    identical with what the reference writer produced, contexts and bin counts included; likewise on generated
    slices of every tool; bad records and a small command scratch are flagged per slice."""
    lib = _host_lib()
    vtm = [(t, s, sp, s["flags"] | SUB_TERMINATE | SUB_CTX_IN, np.stack([s["ctx_in"][:, 0], s["ctx_in"][:, 1]], 1))
           for t, s, sp in _substreams() if not int(sp["tools"]) & SL_WPP]
    gen = [(t, g, sp, g["flags"] | SUB_TERMINATE, None) for t, g, sp in _gen_items(lib, reps=2) if not int(sp["tools"]) & SL_WPP]
    assert len(vtm) >= 6 and len(gen) >= 10
    for items in (vtm, gen):
        n = len(items)
        params, sps = np.zeros(n, PARAMS_DT), np.zeros(n, SLICE_DT)
        ctx_in = np.zeros((n, 379), CTX_DT) if items[0][4] is not None else None
        data, off = [], [0]
        for k, (tag, s, sp, fl, cin) in enumerate(items):
            params[k] = (s["qp"], s["init_id"], fl | SUB_ALIGN, 0)  # the writer side also writes the rbsp trailing bits
            sps[k] = sp
            if cin is not None:
                ctx_in[k]["state0"], ctx_in[k]["state1"] = cin[:, 0], cin[:, 1]
            data.append(s["payload"])
            off.append(off[-1] + s["payload"].size)
        d = engine.decode_slices(np.concatenate(data), np.asarray(off, np.uint64), params, sps, ctx_in=ctx_in, want_ctx=True)
        rec = _compact(d, n)
        out_off = np.concatenate([[0], np.cumsum([p.size + 64 for p in data])]).astype(np.uint64)
        before = engine.launch_count
        w = engine.encode_slices(params, sps, rec["cus"], rec["cu_offsets"], rec["tus"], rec["tu_flags"], rec["tu_offsets"],
                                 rec["ctus"], rec["ctu_offsets"], rec["coeffs"], out_off, ctx_in=ctx_in, want_ctx=True)
        assert engine.launch_count >= before + 7
        for k, (tag, s, sp, fl, cin) in enumerate(items):
            assert int(w["status"][k]) == 0, f"{tag}: status {int(w['status'][k])}"
            got = w["out"][int(out_off[k]):int(out_off[k]) + int(w["sizes"][k])]
            assert got.tobytes() == s["payload"].tobytes(), f"{tag}: {got.size} bytes vs the writer's {s['payload'].size}"
            assert int(w["num_bins"][k]) == int(d["num_bins"][k]), tag
            assert w["ctx_out"][k].tobytes() == d["ctx_out"][k].tobytes(), tag
    # the generated batch once more from the decisions alone (CABAC_CU_FROM_DIR)
    w2 = engine.encode_slices(params, sps, _from_dir(rec["cus"]), rec["cu_offsets"], rec["tus"], rec["tu_flags"], rec["tu_offsets"],
                              rec["ctus"], rec["ctu_offsets"], rec["coeffs"], out_off)
    assert (w2["sizes"] == w["sizes"]).all()
    for k, (tag, s, sp, fl, cin) in enumerate(items):
        a = int(out_off[k])
        assert w2["out"][a:a + int(w2["sizes"][k])].tobytes() == s["payload"].tobytes(), f"{tag}: from intra_dir"
    # errors, on the generated batch: a missing CU record, a TU record of another size, records of another slice
    n = len(gen)
    bad = {k: v.copy() if isinstance(v, np.ndarray) else v for k, v in rec.items()}
    cu_off = bad["cu_offsets"].copy()
    cu_off[1:] -= 1  # slice 0 loses its last CU; every later slice starts one record early
    bad["tus"]["log2_w"][int(rec["tu_offsets"][2])] ^= 1
    w = engine.encode_slices(params, sps, bad["cus"][:-1], cu_off, bad["tus"], bad["tu_flags"], bad["tu_offsets"], bad["ctus"],
                             bad["ctu_offsets"], bad["coeffs"], out_off, raise_on_status=False)
    assert int(w["status"][0]) & SS_BAD_DESC and int(w["status"][2]) & SS_BAD_DESC
    assert all(int(v) != 0 for v in w["status"][:n - 1]), "shifted records must be noticed"
    # command scratch too small for the batch: CABAC_SS_CAPACITY on the slices past the end, nothing encoded
    w = engine.encode_slices(params, sps, rec["cus"], rec["cu_offsets"], rec["tus"], rec["tu_flags"], rec["tu_offsets"],
                             rec["ctus"], rec["ctu_offsets"], rec["coeffs"], out_off, cmd_words=1000, raise_on_status=False)
    assert int(w["status"][-1]) & SS_CAPACITY and not w["sizes"].any()
    # WPP slices are refused in this direction
    t, s, sp = next(x for x in _substreams() if int(x[2]["tools"]) & SL_WPP)
    h = _host_decode(lib, s, sp)
    one = np.array([(s["qp"], s["init_id"], s["flags"] | SUB_TERMINATE | SUB_ALIGN, 0)], PARAMS_DT)
    nctu = len(h["ctus"])
    w = engine.encode_slices(one, np.array([sp], SLICE_DT), h["cus"], [0, len(h["cus"])], h["tus"], h["tu_flags"], [0, len(h["tus"])],
                             h["ctus"], [0, nctu], h["coeffs"], [0, s["payload"].size + 64], raise_on_status=False)
    assert int(w["status"][0]) & SS_UNSUPPORTED


@pytest.mark.gpu
def test_gpu_encode_slices_wpp_rows(engine):
    """cabac_encode_slices_wpp: WPP slices written as one substream per CTU row, contexts handed over on the device -
    the VTM WPP frames come back row by row with the reference's row sizes; mixed in one call with every other
    fixture slice and with generated slices (WPP at every CTU size)."""
    lib = _host_lib()
    vtm = [(t, s, sp, s["flags"] | SUB_TERMINATE | SUB_CTX_IN, np.stack([s["ctx_in"][:, 0], s["ctx_in"][:, 1]], 1))
           for t, s, sp in _substreams()]
    gen = [(t, g, sp, g["flags"] | SUB_TERMINATE, None) for t, g, sp in _gen_items(lib, reps=2)]
    assert sum(1 for x in vtm if int(x[2]["tools"]) & SL_WPP) >= 2 and sum(1 for x in gen if int(x[2]["tools"]) & SL_WPP) >= 3
    for items in (vtm, gen):
        n = len(items)
        params, sps = np.zeros(n, PARAMS_DT), np.zeros(n, SLICE_DT)
        ctx_in = np.zeros((n, 379), CTX_DT) if items[0][4] is not None else None
        data, off = [], [0]
        for k, (tag, s, sp, fl, cin) in enumerate(items):
            params[k] = (s["qp"], s["init_id"], fl | SUB_ALIGN, 0)
            sps[k] = sp
            if cin is not None:
                ctx_in[k]["state0"], ctx_in[k]["state1"] = cin[:, 0], cin[:, 1]
            data.append(s["payload"])
            off.append(off[-1] + s["payload"].size)
        d = engine.decode_slices(np.concatenate(data), np.asarray(off, np.uint64), params, sps, ctx_in=ctx_in, want_ctx=True)
        rec = _compact(d, n)
        rows = engine.slice_rows(sps)
        cap = np.concatenate([[data[k].size + 64] * int(rows[k + 1] - rows[k]) for k in range(n)])
        out_off = np.concatenate([[0], np.cumsum(cap)]).astype(np.uint64)
        w = engine.encode_slices_wpp(params, sps, rec["cus"], rec["cu_offsets"], rec["tus"], rec["tu_flags"], rec["tu_offsets"],
                                     rec["ctus"], rec["ctu_offsets"], rec["coeffs"], out_off, ctx_in=ctx_in, want_ctx=True)
        assert (w["row_offsets"] == rows).all()
        for k, (tag, s, sp, fl, cin) in enumerate(items):
            a, b = int(rows[k]), int(rows[k + 1])
            assert not w["status"][a:b].any(), f"{tag}: status {w['status'][a:b]}"
            got = np.concatenate([w["out"][int(out_off[r]):int(out_off[r]) + int(w["sizes"][r])] for r in range(a, b)])
            assert got.tobytes() == s["payload"].tobytes(), f"{tag}: {got.size} bytes vs the writer's {s['payload'].size}"
            if "row_bytes" in s:
                assert [int(v) for v in w["sizes"][a:b]] == s["row_bytes"], tag
            assert int(w["num_bins"][a:b].sum()) == int(d["num_bins"][k]), tag
            assert w["ctx_out"][b - 1].tobytes() == d["ctx_out"][k].tobytes(), tag
    # a row count that does not match the rectangle is refused for that slice only
    k = next(i for i, x in enumerate(gen) if int(x[2]["tools"]) & SL_WPP and int(rows[i + 1] - rows[i]) > 1)
    sps2 = sps.copy()
    sps2[k]["tools"] &= ~np.uint32(SL_WPP)  # one row announced (slice_rows), records of a slice with several
    rows2 = engine.slice_rows(sps2)
    sps2[k]["tools"] |= np.uint32(SL_WPP)
    cap2 = np.concatenate([[data[i].size + 64] * int(rows2[i + 1] - rows2[i]) for i in range(n)])
    out2 = np.concatenate([[0], np.cumsum(cap2)]).astype(np.uint64)
    lib_rc = engine.lib.cabac_encode_slices_wpp
    sizes, bins, status = np.zeros(int(rows2[-1]), np.uint32), np.zeros(int(rows2[-1]), np.uint32), np.zeros(int(rows2[-1]), np.uint32)
    out = np.zeros(int(out2[-1]), np.uint8)
    P = lambda a: np.ascontiguousarray(a).ctypes.data_as(C.c_void_p)  # noqa: E731
    keep = [np.ascontiguousarray(rec[x]) for x in ("cus", "cu_offsets", "tus", "tu_flags", "tu_offsets", "ctus", "ctu_offsets", "coeffs")]
    rc = lib_rc(engine.h, P(params), P(sps2), n, P(keep[0]), P(keep[1]), P(keep[2]), P(keep[3]), P(keep[4]), P(keep[5]), P(keep[6]),
                P(keep[7]), keep[7].size, None, 1 << 22, P(rows2), int(rows2[-1]), P(out), P(out2), P(sizes), P(bins), None, P(status))
    assert rc == -3 and int(status[int(rows2[k])]) & SS_BAD_DESC
    assert not np.delete(status, int(rows2[k])).any()


@pytest.mark.gpu
def test_gpu_encode_slices_survives_garbage_records(engine):
    """Records with random bytes flipped (CU fields, TU shapes and flags, CTU parameters; coefficient offsets stay valid)
    come back with status bits or with some legal stream - no hang, no write outside a slice's output region."""
    lib = _host_lib()
    items = [(t, g, sp) for t, g, sp in _gen_items(lib, reps=1)]
    n = len(items)
    params, sps = np.zeros(n, PARAMS_DT), np.zeros(n, SLICE_DT)
    data, off = [], [0]
    for k, (tag, g, sp) in enumerate(items):
        params[k] = (g["qp"], g["init_id"], g["flags"] | SUB_TERMINATE | SUB_ALIGN, 0)
        sps[k] = sp
        data.append(g["payload"])
        off.append(off[-1] + g["payload"].size)
    d = engine.decode_slices(np.concatenate(data), np.asarray(off, np.uint64), params, sps)
    rec = _compact(d, n)
    rows = engine.slice_rows(sps)
    cap = np.concatenate([[data[k].size * 2 + 256] * int(rows[k + 1] - rows[k]) for k in range(n)])
    out_off = np.concatenate([[0], np.cumsum(cap)]).astype(np.uint64)
    rng = np.random.default_rng(77)
    flagged = 0
    for trial in range(6):
        cus, tus, ctus = rec["cus"].copy(), rec["tus"].copy(), rec["ctus"].copy()
        cb = cus.view(np.uint8).reshape(len(cus), -1)
        for _ in range(40 << trial):
            cb[rng.integers(len(cus)), rng.integers(cb.shape[1])] ^= np.uint8(1 << rng.integers(8))
        tb = tus.view(np.uint8).reshape(len(tus), -1)
        for _ in range(10 << trial):
            tb[rng.integers(len(tus)), 8 + rng.integers(4)] ^= np.uint8(1 << rng.integers(8))  # log2_w, log2_h, comp_id, flags
        ub = ctus.view(np.uint8).reshape(len(ctus), -1)
        for _ in range(10 << trial):
            ub[rng.integers(len(ctus)), rng.integers(ub.shape[1])] ^= np.uint8(1 << rng.integers(8))
        w = engine.encode_slices_wpp(params, sps, cus, rec["cu_offsets"], tus, rec["tu_flags"], rec["tu_offsets"], ctus,
                                     rec["ctu_offsets"], rec["coeffs"], out_off, raise_on_status=False)
        assert (w["sizes"].astype(np.int64) <= cap).all()
        flagged += int((w["status"] != 0).sum())
    assert flagged > 0
    # a TU record pointing outside the coefficient slab is refused by the host-pointer call before anything runs
    tus = rec["tus"].copy()
    tus["coeff_offset"][len(tus) // 2] = rec["coeffs"].size - 3
    with pytest.raises(Exception, match="outside the coefficient slab"):
        engine.encode_slices_wpp(params, sps, rec["cus"], rec["cu_offsets"], tus, rec["tu_flags"], rec["tu_offsets"], rec["ctus"],
                                 rec["ctu_offsets"], rec["coeffs"], out_off)
    # and the clean records still code to the payloads afterwards
    w = engine.encode_slices_wpp(params, sps, rec["cus"], rec["cu_offsets"], rec["tus"], rec["tu_flags"], rec["tu_offsets"],
                                 rec["ctus"], rec["ctu_offsets"], rec["coeffs"], out_off)
    for k, (tag, g, sp) in enumerate(items):
        a, b = int(rows[k]), int(rows[k + 1])
        got = np.concatenate([w["out"][int(out_off[r]):int(out_off[r]) + int(w["sizes"][r])] for r in range(a, b)])
        assert got.tobytes() == g["payload"].tobytes(), tag


@pytest.mark.gpu
def test_gpu_slice_round_trip_at_1080p(engine):
    """BASELINE configs[2]'s substream shape at f2 level: whole 1920x1080 pictures as one slice each (3.0 Mbins per slice,
    tests/golden/gen_slices.npz) - parsed from the bytes, written back from the records: the same bytes, bins, contexts."""
    z = np.load(os.path.join(HERE, "golden", "gen_slices.npz"))
    n = 2
    params, sps = np.zeros(n, PARAMS_DT), np.zeros(n, SLICE_DT)
    data, off = [], [0]
    for k in range(n):
        meta = z[f"hd{k}_meta"]
        params[k] = (int(meta[0]), int(meta[1]), int(meta[2]) | SUB_TERMINATE | SUB_ALIGN, 0)
        sps[k] = z[f"hd{k}_slice"].view(SLICE_DT)[0]
        data.append(z[f"hd{k}_payload"])
        off.append(off[-1] + data[-1].size)
    assert int(sps[0]["pic_w"]) == 1920 and int(sps[0]["pic_h"]) == 1080
    d = engine.decode_slices(np.concatenate(data), np.asarray(off, np.uint64), params, sps, want_ctx=True)
    for k in range(n):
        meta = z[f"hd{k}_meta"]
        assert int(d["status"][k]) == 0 and int(d["num_bins"][k]) == int(meta[3]) and int(d["counts"][k][0]) == int(meta[4])
    rec = _compact(d, n)
    out_off = np.concatenate([[0], np.cumsum([p.size + 256 for p in data])]).astype(np.uint64)
    w = engine.encode_slices(params, sps, _from_dir(rec["cus"]), rec["cu_offsets"], rec["tus"], rec["tu_flags"], rec["tu_offsets"],
                             rec["ctus"], rec["ctu_offsets"], rec["coeffs"], out_off, want_ctx=True)
    for k in range(n):
        got = w["out"][int(out_off[k]):int(out_off[k]) + int(w["sizes"][k])]
        assert got.tobytes() == data[k].tobytes(), f"1080p slice {k}"
        assert int(w["num_bins"][k]) == int(d["num_bins"][k]) and w["ctx_out"][k].tobytes() == d["ctx_out"][k].tobytes()
    # the same pictures with entropy_coding_sync_enabled (gen_slices_wpp.npz): nine CTU-row substreams per slice, parsed
    # as one unit, written as rows with the context hand-over on the device
    zw = np.load(os.path.join(HERE, "golden", "gen_slices_wpp.npz"))
    data, off = [], [0]
    for k in range(n):
        meta = zw[f"hdwpp{k}_meta"]
        params[k] = (int(meta[0]), int(meta[1]), int(meta[2]) | SUB_TERMINATE | SUB_ALIGN, 0)
        sps[k] = zw[f"hdwpp{k}_slice"].view(SLICE_DT)[0]
        data.append(zw[f"hdwpp{k}_payload"])
        off.append(off[-1] + data[-1].size)
    assert int(sps[0]["tools"]) & SL_WPP
    d = engine.decode_slices(np.concatenate(data), np.asarray(off, np.uint64), params, sps, want_ctx=True)
    rec = _compact(d, n)
    rows = engine.slice_rows(sps)
    assert int(rows[-1]) == 18
    out_off = np.concatenate([[0], np.cumsum([data[k].size + 256 for k in range(n) for _ in range(9)])]).astype(np.uint64)
    w = engine.encode_slices_wpp(params, sps, _from_dir(rec["cus"]), rec["cu_offsets"], rec["tus"], rec["tu_flags"],
                                 rec["tu_offsets"], rec["ctus"], rec["ctu_offsets"], rec["coeffs"], out_off, want_ctx=True)
    for k in range(n):
        a, b = int(rows[k]), int(rows[k + 1])
        got = np.concatenate([w["out"][int(out_off[q]):int(out_off[q]) + int(w["sizes"][q])] for q in range(a, b)])
        assert got.tobytes() == data[k].tobytes(), f"1080p WPP slice {k}"
        assert int(w["num_bins"][a:b].sum()) == int(d["num_bins"][k]) and w["ctx_out"][b - 1].tobytes() == d["ctx_out"][k].tobytes()


# ---- what the VTM fixtures pin: the non-residual contexts the reference's writer actually used in them ----
WALK_CTX_SETS = [("split_cu_flag", 0, 9), ("split_qt_flag", 9, 6), ("mtt_split_cu_vertical_flag", 15, 5),
                 ("mtt_split_cu_binary_flag", 20, 4), ("multi_ref_line_idx", 38, 2), ("intra_luma_mpm_flag", 40, 1),
                 ("intra_luma_planar_flag", 41, 2), ("cclm_mode_flag", 43, 1), ("cclm_mode_idx", 44, 1),
                 ("intra_chroma_pred_mode", 45, 1), ("mip_flag", 46, 4), ("cu_qp_delta_abs", 50, 2), ("bdpcm_mode", 71, 4),
                 ("cbf_y", 77, 4), ("cbf_cb", 81, 2), ("cbf_cr", 83, 3), ("sao_merge_flag", 294, 1), ("sao_type_idx", 295, 1),
                 ("lfnst_idx (separate trees)", 297, 2), ("mts_idx", 312, 4), ("isp_mode", 316, 2),
                 ("cu_chroma_qp_offset_flag", 325, 1), ("cu_chroma_qp_offset_idx", 326, 1), ("alf_ctb_flag", 332, 9),
                 ("alf_ctb_filter_alt_idx", 341, 2), ("alf_use_aps_flag", 343, 1), ("cc_alf_ctb_idc", 344, 6),
                 ("tu_joint_cbcr_residual_flag", 354, 3)]
# DESIGN.md section 2 names these as reached by generated streams only
NOT_REACHED_BY_VTM_FIXTURES = {314, 315}  # mts_idx beyond the second bin


def test_fixture_context_coverage_is_what_design_md_says():
    """Every context the CTU-level walk can select, except the documented handful, is used by at least one call the
    reference's writer made while producing the fixtures - so the pin of the walk is not narrower than claimed."""
    hit = np.zeros(512, np.int64)
    for name in FIXTURES:
        for s in load_fixture(name)[0]:
            c = np.asarray(s["cmds"], np.uint32)
            np.add.at(hit, c[(c >> 29) == 0] & 511, 1)
    walk = [a + i for _, a, n in WALK_CTX_SETS for i in range(n)]
    assert len(walk) == 84
    missing = {k for k in walk if hit[k] == 0}
    assert missing <= NOT_REACHED_BY_VTM_FIXTURES, f"contexts no fixture reaches: {sorted(missing)}"
