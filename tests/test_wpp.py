"""WPP (one substream per CTU row, contexts handed down after the first CTU of the row above):
the wavefront scheduler of entropy_coding_b200/wpp.py against the oracle run sequentially."""
import numpy as np
import pytest

from entropy_coding_b200 import workload as W
from entropy_coding_b200.types import SUB_ALIGN, SUB_CTX_IN, SUB_TERMINATE


def test_row_workload_layout():
    wl, rows, first = W.frames_as_ctu_rows(2, 640, 360)
    assert rows == 3 and wl.n_sub == 6 and len(first) == 6
    assert (first > 0).all() and (first <= np.diff(wl.tu_offsets.astype(np.int64))).all()
    assert (wl.params["qp"][:3] == wl.params["qp"][0]).all()


def oracle_wpp(oracle, wl, rows, first):
    """Reference order: row r of a frame starts from the contexts of row r-1 after its first CTU."""
    outs, bins = [], []
    for f in range(wl.n_sub // rows):
        sync = None
        for r in range(rows):
            i = f * rows + r
            a, b = int(wl.tu_offsets[i]), int(wl.tu_offsets[i + 1])
            p = wl.params[i]
            flags = int(p["flags"]) | (SUB_CTX_IN if sync is not None else 0)
            eb, nb, _ = oracle.encode_residual(wl.tus[a:b], wl.coeffs, int(p["qp"]), int(p["init_id"]), flags, ctx_in=sync)
            outs.append(eb)
            bins.append(nb)
            # contexts after the first CTU (not terminated) seed the next row
            *_, ctx = oracle.encode_residual(wl.tus[a:a + int(first[i])], wl.coeffs, int(p["qp"]), int(p["init_id"]),
                                               flags & ~(SUB_TERMINATE | SUB_ALIGN), ctx_in=sync, want_ctx=True)
            sync = ctx
    return outs, bins


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [0, 1, 3], ids=["auto", "fused", "warp_specialised"])
def test_wpp_encode_decode_vs_oracle(engine, oracle, mode):
    """One launch through cabac_encode_residual_wpp / cabac_decode_residual_wpp, every CTU coded once."""
    from entropy_coding_b200.wpp import WppCoder
    wl, rows, first = W.frames_as_ctu_rows(3, 512, 384, seed0=77)  # 3 frames x 3 rows x 4 CTUs
    cap = wl.out_capacity(bytes_per_coeff=5.0)
    coder = WppCoder(engine, rows)
    engine.set_encoder_mode(mode)
    try:
        r = coder.encode(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, cap, first)
    finally:
        engine.set_encoder_mode(0)
    want, want_bins = oracle_wpp(oracle, wl, rows, first)
    ep = r["entry_points"]
    outs = [r["packed"][int(ep[i]):int(ep[i + 1])] for i in range(wl.n_sub)]
    for i in range(wl.n_sub):
        assert outs[i].tobytes() == want[i].tobytes(), f"row substream {i}"
        assert int(r["num_bins"][i]) == want_bins[i]
    assert r["launches"] <= 6  # cost, order, slab scan, ONE coding kernel, entry points + pack
    # hand-off matters: without it rows > 0 would differ
    eb0, _, _ = oracle.encode_residual(wl.tus[int(wl.tu_offsets[1]):int(wl.tu_offsets[2])], wl.coeffs,
                                       int(wl.params[1]["qp"]), 2, int(wl.params[1]["flags"]))
    assert eb0.tobytes() != want[1].tobytes()
    d = coder.decode(r["packed"], ep, wl.tus, wl.tu_offsets, wl.params, wl.coeffs.size, first)
    assert (d["coeffs"] == wl.coeffs).all()
    assert (d["num_bins"] == r["num_bins"]).all()
    assert d["launches"] <= 4


@pytest.mark.gpu
def test_wpp_final_contexts_and_ragged_rows(engine, oracle):
    """ctx_out of every row, rows whose first CTU is the whole row (publish at the end), a row without
    dependants, and a frame whose first row starts from handed-in contexts (CABAC_SUB_CTX_IN + WPP)."""
    from entropy_coding_b200.wpp import sync_records
    from entropy_coding_b200.types import CTX_DT
    wl, rows, first = W.frames_as_ctu_rows(2, 384, 384, seed0=5)
    first = first.copy()
    ntu = np.diff(wl.tu_offsets.astype(np.int64))
    first[1] = ntu[1] + 7          # sync point beyond the row: publish at its end
    first[4] = ntu[4]              # exactly the row
    w = sync_records(wl.n_sub, rows, first)
    cap = wl.out_capacity(bytes_per_coeff=5.0)
    params = wl.params.copy()
    ctx_in = np.zeros((wl.n_sub, 379), CTX_DT)
    seed_ctx, _ = oracle.ctx_init(12, 0)
    ctx_in[rows] = seed_ctx        # first row of frame 1 starts from other contexts
    params["flags"][rows] |= SUB_CTX_IN
    r = engine.encode_residual_wpp(wl.tus, wl.tu_offsets, params, w, wl.coeffs, cap, ctx_in=ctx_in, want_ctx=True)
    ep = r["entry_points"]
    for f in range(wl.n_sub // rows):
        sync = seed_ctx if f == 1 else None
        for rr in range(rows):
            i = f * rows + rr
            a, b = int(wl.tu_offsets[i]), int(wl.tu_offsets[i + 1])
            p = wl.params[i]
            flags = int(p["flags"])
            eb, nb, _, cfin = oracle.encode_residual(wl.tus[a:b], wl.coeffs, int(p["qp"]), int(p["init_id"]), flags,
                                                     ctx_in=sync, want_ctx=True)
            assert r["packed"][int(ep[i]):int(ep[i + 1])].tobytes() == eb.tobytes(), f"row {i}"
            assert (r["ctx_out"][i]["state0"] == cfin["state0"]).all() and (r["ctx_out"][i]["state1"] == cfin["state1"]).all()
            k = min(int(first[i]), b - a)
            *_, sync = oracle.encode_residual(wl.tus[a:a + k], wl.coeffs, int(p["qp"]), int(p["init_id"]),
                                              flags & ~(SUB_TERMINATE | SUB_ALIGN), ctx_in=sync, want_ctx=True)
    d = engine.decode_residual_wpp(r["packed"], ep, wl.tus, wl.tu_offsets, params, w, wl.coeffs.size, ctx_in=ctx_in,
                                   want_ctx=True)
    assert (d["coeffs"] == wl.coeffs).all()
    assert (d["ctx_out"]["state0"] == r["ctx_out"]["state0"]).all() and (d["ctx_out"]["state1"] == r["ctx_out"]["state1"]).all()


@pytest.mark.gpu
def test_wpp_rejects_forward_dependency(engine):
    from entropy_coding_b200.engine import CabacError
    from entropy_coding_b200.wpp import sync_records
    wl, rows, first = W.frames_as_ctu_rows(1, 256, 256, seed0=9)
    w = sync_records(wl.n_sub, rows, first)
    w["dep"][0] = 1  # a row may only depend on an earlier substream (claim order = index order)
    with pytest.raises(CabacError):
        engine.encode_residual_wpp(wl.tus, wl.tu_offsets, wl.params, w, wl.coeffs, wl.out_capacity())


@pytest.mark.gpu
def test_wpp_4k_round_trip(engine):
    """configs[3] geometry (3840x2160: 17 rows of 30 CTUs) on a few frames: round trip + bin conservation."""
    from entropy_coding_b200.wpp import WppCoder
    wl, rows, first = W.frames_as_ctu_rows(4, 3840, 2160, seed0=4000)
    assert rows == 17
    cap = wl.out_capacity()
    coder = WppCoder(engine, rows)
    r = coder.encode(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, cap, first)
    d = coder.decode(r["packed"], r["entry_points"], wl.tus, wl.tu_offsets, wl.params, wl.coeffs.size, first)
    assert (d["coeffs"] == wl.coeffs).all() and (d["num_bins"] == r["num_bins"]).all()
    assert r["launches"] <= 6 and d["launches"] <= 4
