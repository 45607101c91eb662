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
def test_wpp_encode_decode_vs_oracle(engine, oracle):
    from entropy_coding_b200.wpp import WppCoder
    wl, rows, first = W.frames_as_ctu_rows(3, 512, 384, seed0=77)  # 3 frames x 3 rows x 4 CTUs
    cap = wl.out_capacity(bytes_per_coeff=5.0)
    coder = WppCoder(engine, rows)
    r = coder.encode(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, cap, first)
    want, want_bins = oracle_wpp(oracle, wl, rows, first)
    outs = [r["out"][int(cap[i]):int(cap[i]) + int(r["sizes"][i])] for i in range(wl.n_sub)]
    for i in range(wl.n_sub):
        assert outs[i].tobytes() == want[i].tobytes(), f"row substream {i}"
        assert int(r["num_bins"][i]) == want_bins[i]
    # hand-off matters: without it rows > 0 would differ
    eb0, _, _ = oracle.encode_residual(wl.tus[int(wl.tu_offsets[1]):int(wl.tu_offsets[2])], wl.coeffs,
                                       int(wl.params[1]["qp"]), 2, int(wl.params[1]["flags"]))
    assert eb0.tobytes() != want[1].tobytes()
    ioff = np.concatenate([[0], np.cumsum(r["sizes"])]).astype(np.uint64)
    d = coder.decode(np.concatenate(outs), ioff, wl.tus, wl.tu_offsets, wl.params, wl.coeffs.size, first)
    assert (d["coeffs"] == wl.coeffs).all()
    assert (d["num_bins"] == r["num_bins"]).all()


@pytest.mark.gpu
def test_wpp_4k_round_trip(engine):
    """configs[3] geometry (3840x2160: 17 rows of 30 CTUs) on a few frames: round trip + bin conservation."""
    from entropy_coding_b200.wpp import WppCoder
    wl, rows, first = W.frames_as_ctu_rows(4, 3840, 2160, seed0=4000)
    assert rows == 17
    cap = wl.out_capacity()
    coder = WppCoder(engine, rows)
    r = coder.encode(wl.tus, wl.tu_offsets, wl.params, wl.coeffs, cap, first)
    outs = [r["out"][int(cap[i]):int(cap[i]) + int(r["sizes"][i])] for i in range(wl.n_sub)]
    ioff = np.concatenate([[0], np.cumsum(r["sizes"])]).astype(np.uint64)
    d = coder.decode(np.concatenate(outs), ioff, wl.tus, wl.tu_offsets, wl.params, wl.coeffs.size, first)
    assert (d["coeffs"] == wl.coeffs).all() and (d["num_bins"] == r["num_bins"]).all()
    assert r["launches"] >= 4 * (rows - 1) + 4  # R-1 dependent phase-A launches + the full-width pass
