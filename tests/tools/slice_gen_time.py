"""Timing of cabac_decode_slices_dev on GENERATED slices (oracle/syntax_host.cpp: syn_generate_slice): 1080p pictures as
one slice each, dense residuals - the shape of BASELINE configs[2] for the whole-slice decoder.  Dev tool, B200 only.
  python tests/tools/slice_gen_time.py [n_slices=64] [density=0.5] [split_p=0.45]"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
import entropy_coding_b200 as E  # noqa: E402
import test_syntax as T  # noqa: E402
from entropy_coding_b200.types import PARAMS_DT, SLICE_DT, SUB_TERMINATE  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
density = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
split_p = float(sys.argv[3]) if len(sys.argv) > 3 else 0.45
w, h = (int(v) for v in (sys.argv[4].split("x") if len(sys.argv) > 4 else (1920, 1080)))
lib = T._host_lib()
eng = E.CabacEngine(0)
stream = torch.cuda.Stream(device=0)
torch.cuda.set_stream(stream)
eng.set_stream(stream.cuda_stream)
base = [T.generate_slice(lib, 500 + k, T.make_slice_params(w, h, 7), split_p=split_p, density=density) for k in range(min(n, 4))]
sp = T.make_slice_params(w, h, 7)
params, sps = np.zeros(n, PARAMS_DT), np.zeros(n, SLICE_DT)
data, off, bins, ncoef = [], [0], 0, 0
for j in range(n):
    g = base[j % len(base)]
    params[j] = (g["qp"], g["init_id"], g["flags"] | SUB_TERMINATE, 0)
    sps[j] = sp
    data.append(g["payload"])
    off.append(off[-1] + g["payload"].size)
    bins += g["num_bins"]
    ncoef += g["coeffs"].size
cap = E.CabacEngine.slice_capacities(sps)
dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()  # noqa: E731
d = {k: dev(v) for k, v in (("in", np.concatenate(data)), ("off", np.asarray(off, np.uint64)), ("p", params), ("sp", sps),
                            ("cu", cap["cu"]), ("tu", cap["tu"]), ("ctu", cap["ctu"]), ("co", cap["coeff"]))}
cus = torch.empty(int(cap["cu"][-1]) * 32, dtype=torch.uint8, device="cuda")
tus = torch.empty(int(cap["tu"][-1]) * 16, dtype=torch.uint8, device="cuda")
tuf = torch.empty(int(cap["tu"][-1]), dtype=torch.uint8, device="cuda")
ctus = torch.empty(int(cap["ctu"][-1]) * 32, dtype=torch.uint8, device="cuda")
co = torch.zeros(int(cap["coeff"][-1]), dtype=torch.int32, device="cuda")
cnt = torch.zeros(n * 3, dtype=torch.int32, device="cuda")
nb = torch.zeros(n, dtype=torch.int32, device="cuda")
st = torch.zeros(n, dtype=torch.int32, device="cuda")


def run():
    eng.decode_slices_dev(d["in"], d["off"], d["p"], d["sp"], n, d["cu"], d["tu"], d["ctu"], d["co"], cus, tus, tuf, ctus, co,
                          counts=cnt, num_bins=nb, status=st)


run()
torch.cuda.synchronize()
assert int(st.abs().sum()) == 0 and int(nb.to(torch.int64).sum()) == bins
k0 = int(cap["coeff"][0])
assert (co[k0:k0 + base[0]["coeffs"].size].cpu().numpy() == base[0]["coeffs"]).all()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3):
    run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 3
print(f"decode_slices, generated {w}x{h} slices: n={n}, {bins / n / 1e6:.2f} Mbins and {off[-1] / n / 1e3:.0f} kB per slice, "
      f"{bins / ncoef:.2f} bins per coefficient position: {ms:.2f} ms, {bins / ms / 1e6:.3f} Gbins/s", flush=True)
