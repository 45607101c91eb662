"""Debug aid: decode one fixture substream with a -DCABAC_SLICE_TRACE build of the library and show where the
device's call sequence leaves the host run's (oracle/syntax_host.cpp).  Usage:
  nvcc ... -DCABAC_SLICE_TRACE -o entropy_coding_b200/libcabac_b200_trace.so ...
  CABAC_B200_LIB=entropy_coding_b200/libcabac_b200_trace.so python tests/tools/slice_trace.py [index]"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, os.path.join(HERE, "..", "golden"))
import entropy_coding_b200 as E  # noqa: E402
import test_syntax as T  # noqa: E402
from entropy_coding_b200.types import CTX_DT, PARAMS_DT, SLICE_DT, SUB_CTX_IN, SUB_TERMINATE  # noqa: E402

idx = int(sys.argv[1]) if len(sys.argv) > 1 else 0
tag, s, sp = T._substreams()[idx]
eng = E.CabacEngine(0)
params = np.array([(s["qp"], s["init_id"], s["flags"] | SUB_TERMINATE | SUB_CTX_IN, 0)], PARAMS_DT)
ctx_in = np.zeros((1, 379), CTX_DT)
ctx_in[0]["state0"], ctx_in[0]["state1"] = s["ctx_in"][:, 0], s["ctx_in"][:, 1]
r = eng.decode_slices(s["payload"], np.array([0, s["payload"].size], np.uint64), params, np.array([sp], SLICE_DT),
                      ctx_in=ctx_in, raise_on_status=False)
print(tag, "status", r["status"], "bins", r["num_bins"], "expected", s["num_bins"], "counts", r["counts"])
def fetch(which):
    tr, n = np.zeros(1 << 17, np.uint32), C.c_uint32(0)
    assert eng.lib.cabac_debug_slice_trace(which, tr.ctypes.data_as(C.c_void_p), C.c_size_t(tr.size), C.byref(n)) == 0
    return tr[:n.value]


dev, dev5 = fetch(0), fetch(1)
k5 = next((k for k in range(min(len(dev), len(dev5))) if dev[k] != dev5[k]), None)
print("lane 0 recorded", len(dev), "calls, lane 5", len(dev5), "- first difference between the lanes at", k5)
if k5 is not None:
    print(" lane 0:", T._flat(dev[max(0, k5 - 4):k5 + 3]), [hex(int(w)) for w in dev[max(0, k5 - 2):k5 + 2]])
    print(" lane 5:", T._flat(dev5[max(0, k5 - 4):k5 + 3]), [hex(int(w)) for w in dev5[max(0, k5 - 2):k5 + 2]])
host = T._host_decode(T._host_lib(), s, sp)["trace"]
# TU markers of the device trace carry the running bin count in the low bits
fd = T._flat(dev & np.where((dev >> 29) == 5, np.uint32(0xE0000000), np.uint32(0xFFFFFFFF)))
fh = T._flat(host)
bad = next((k for k in range(min(len(fd), len(fh))) if fd[k] != fh[k]), None)
print("device calls", len(fd), "host calls", len(fh), "first difference at", bad)
if bad is not None:
    print(" device:", fd[max(0, bad - 6):bad + 4])
    print(" host  :", fh[max(0, bad - 6):bad + 4])
    ntu = sum(1 for t in fd[:bad] if t[0] == "tu")
    print(" TUs before:", ntu, "device bins after each TU:", [int(w & 0xFFFFFF) for w in dev if (w >> 29) == 5][:ntu + 1])
