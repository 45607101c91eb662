"""CPU suite: the synthetic workload generator (host logic)."""
import numpy as np

from entropy_coding_b200 import workload as W


def test_random_substreams_deterministic_and_valid():
    a = W.random_substreams(6, 5)
    b = W.random_substreams(6, 5)
    assert (a.coeffs == b.coeffs).all() and (a.tus == b.tus).all()
    assert a.tu_offsets[0] == 0 and a.tu_offsets[-1] == len(a.tus)
    area = 1 << (a.tus["log2_w"].astype(int) + a.tus["log2_h"].astype(int))
    assert (np.diff(a.tus["coeff_offset"].astype(np.int64)) == area[:-1]).all()
    assert a.coeffs.size == int(a.tus["coeff_offset"][-1]) + int(area[-1])
    for t in a.tus:  # every TU is coded (cbf = 1)
        n = 1 << (int(t["log2_w"]) + int(t["log2_h"]))
        assert a.coeffs[int(t["coeff_offset"]):int(t["coeff_offset"]) + n].any()
    assert np.abs(a.coeffs).max() <= 32768


def test_frame_tiling_covers_frame():
    wl = W.frames_as_ctu_substreams(1, 256, 192, ctu=128)
    assert wl.n_sub == 4  # 2 x 2 CTUs (second row 64 high)
    luma = wl.tus[wl.tus["comp_id"] == 0]
    assert int((1 << (luma["log2_w"].astype(int) + luma["log2_h"].astype(int))).sum()) == 256 * 192
    chroma = wl.tus[wl.tus["comp_id"] > 0]
    assert int((1 << (chroma["log2_w"].astype(int) + chroma["log2_h"].astype(int))).sum()) == 2 * 128 * 96
    per = W.frames_as_ctu_substreams(2, 256, 192, per_frame=True)
    assert per.n_sub == 2 and len(per.tus) > 0


def test_subset_rebases():
    wl = W.random_substreams(8, 9)
    s = wl.subset([5, 2])
    assert s.n_sub == 2 and s.tus["coeff_offset"][0] == 0
    a, b = int(wl.tu_offsets[5]), int(wl.tu_offsets[6])
    n = int(s.tu_offsets[1])
    assert n == b - a
    lo = int(wl.tus[a]["coeff_offset"])
    assert (s.coeffs[:16] == wl.coeffs[lo:lo + 16]).all()
    cap = wl.out_capacity()
    assert (cap % 16 == 0).all() and len(cap) == wl.n_sub + 1
