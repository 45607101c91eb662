"""Dev tool: wall time of the two host-pointer calls of bench.py's e2e leg, separately."""
import ctypes as C, sys, time, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import entropy_coding_b200 as E
from entropy_coding_b200 import workload as W

wl = W.frames_as_ctu_substreams(64, 1920, 1080, seed0=1000)
cap = wl.out_capacity()
eng = E.CabacEngine(0)
lib, h = eng.lib, eng.h
P = lambda a: a.ctypes.data_as(C.c_void_p)
n_sub, n_tu, n_co = wl.n_sub, len(wl.tus), wl.coeffs.size
h_tus, h_tuoff, h_par, h_cap, h_co = (eng.pinned_copy(x) for x in (wl.tus, wl.tu_offsets, wl.params, cap, wl.coeffs))
pc = 100 << 20
h_packed, h_entry = eng.host_array(pc, np.uint8), eng.host_array(n_sub + 1, np.uint64)
h_sizes, h_ebins, h_est, h_dbins, h_dbits, h_dst = (eng.host_array(n_sub, np.uint32) for _ in range(6))
h_eflags, h_dflags, h_dec = eng.host_array(n_tu, np.uint8), eng.host_array(n_tu, np.uint8), eng.host_array(n_co, np.int32)
te = td = 0.0
for it in range(6):
    t0 = time.perf_counter()
    rc = lib.cabac_encode_residual_packed(h, P(h_tus), P(h_tuoff), P(h_par), n_sub, P(h_co), n_co, None, P(h_cap),
                                          P(h_packed), pc, P(h_entry), P(h_sizes), P(h_ebins), P(h_eflags), None, P(h_est))
    t1 = time.perf_counter()
    rc2 = lib.cabac_decode_residual(h, P(h_packed), P(h_entry), P(h_tus), P(h_tuoff), P(h_par), n_sub, None, P(h_dec),
                                    n_co, P(h_dbins), P(h_dbits), P(h_dflags), None, P(h_dst))
    t2 = time.perf_counter()
    assert rc == 0 and rc2 == 0
    if it >= 2:
        te += t1 - t0; td += t2 - t1
print(f"encode_packed {te / 4 * 1e3:.2f} ms   decode {td / 4 * 1e3:.2f} ms   ok={bool((h_dec == wl.coeffs).all())}")
