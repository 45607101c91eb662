"""Reader for the VTM-produced stream fixtures (tests/golden/vtm_*.npz) and for the raw
records tests/golden/vtm_record/recording_bin_encoder.hpp appends to $VTM_RECORD_OUT.

TEST INFRASTRUCTURE.  A fixture is what the reference's slice writer really emitted for a
seeded synthetic QCIF clip through EncoderApp (STANDALONE_ENTROPY_CODEC=ON): per substream the
recorded BinEncIf calls with CABAC_CMD_TU markers, the TU descriptors + TCoeff blocks behind the
markers, slice QP / init table / tool flags, contexts before and after, per-context bin counts
and the payload bytes (FIFO after finish() + writeByteAlignment()).
"""
from __future__ import annotations

import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
MAGIC = 0x43425243   # 'CBRC': 12 header words
MAGIC2 = 0x32524243  # 'CBR2': + n_slice, n_cus (slice parameters and CU records for the CTU parser)
CU_DT = np.dtype([("x", "<u2"), ("y", "<u2"), ("log2w", "u1"), ("log2h", "u1"), ("ch_type", "u1"), ("qt_depth", "u1"),
                  ("mt_depth", "u1"), ("flags", "u1"), ("bdpcm", "u1"), ("mrl", "u1"), ("isp", "u1"), ("lfnst", "u1"),
                  ("mts", "u1"), ("dir", "u1")])
NUM_CTX = 379
TU_DT = np.dtype([("coeff_offset", "<u8"), ("log2_w", "u1"), ("log2_h", "u1"), ("comp_id", "u1"),
                  ("flags", "u1"), ("reserved", "<u4")])


def parse_records(path: str) -> list[dict]:
    raw = np.fromfile(path, np.uint8)
    pos, out = 0, []

    def take(dt, n):
        nonlocal pos
        a = raw[pos:pos + n * np.dtype(dt).itemsize].view(dt).copy()
        pos += n * np.dtype(dt).itemsize
        return a

    while pos < raw.size:
        h = take("<u4", 12)
        assert int(h[0]) in (MAGIC, MAGIC2), "bad record magic"
        n_slice, n_cus = (int(v) for v in take("<u4", 2)) if int(h[0]) == MAGIC2 else (0, 0)
        qp = int(np.int32(h[1]))
        n_cmds, n_tus, n_coeffs, n_payload, n_ctu, n_full = (int(h[4]), int(h[5]), int(h[6]), int(h[7]), int(h[9]),
                                                              int(h[11]))
        out.append(dict(
            qp=qp, init_id=int(h[2]), flags=int(h[3]), num_bins=int(h[8]), tools=int(h[10]),
            cmds=take("<u4", n_cmds), tus=take(TU_DT, n_tus), coeffs=take("<i4", n_coeffs),
            cuctx_before=take("u1", n_tus), cuctx_after=take("u1", n_tus),
            ctu_marks=take("<u4", 2 * n_ctu).reshape(-1, 2),
            ctx_in=take("<u2", 2 * NUM_CTX).reshape(-1, 2), ctx_out=take("<u2", 2 * NUM_CTX).reshape(-1, 2),
            ctx_bins=take("<u4", NUM_CTX), payload=take("u1", n_payload), full_cmds=take("<u4", n_full),
            slice=take("<u4", n_slice), cus=take(CU_DT, n_cus)))
    return out


_KEYS = ("cmds", "tus", "coeffs", "cuctx_before", "cuctx_after", "ctu_marks", "ctx_in", "ctx_out", "ctx_bins",
         "payload", "full_cmds", "slice", "cus")
_SCALARS = ("qp", "init_id", "flags", "num_bins", "tools")


def save_fixture(path: str, subs: list[dict], extra: dict) -> None:
    d = {"n_sub": np.int64(len(subs))}
    for i, s in enumerate(subs):
        for k in _KEYS:
            a = s[k]
            # coefficient blocks are mostly zero: store int16 when they fit (they do for 10-bit video)
            if k == "coeffs" and a.size and np.abs(a).max() < 32768:
                a = a.astype(np.int16)
            d[f"s{i}_{k}"] = a
        d[f"s{i}_meta"] = np.array([s[k] for k in _SCALARS], np.int64)
    for k, v in extra.items():
        d[k] = v
    np.savez_compressed(path, **d)


def load_fixture(name: str) -> tuple[list[dict], dict]:
    z = np.load(os.path.join(HERE, name))
    n = int(z["n_sub"])
    subs = []
    for i in range(n):
        s = {k: z[f"s{i}_{k}"] for k in _KEYS if f"s{i}_{k}" in z.files}
        if "cus" in s and s["cus"].dtype != CU_DT:
            s["cus"] = s["cus"].view(CU_DT)
        s["coeffs"] = s["coeffs"].astype(np.int32)
        s["tus"] = s["tus"].view(TU_DT) if s["tus"].dtype != TU_DT else s["tus"]
        s.update({k: int(v) for k, v in zip(_SCALARS, z[f"s{i}_meta"])})
        subs.append(s)
    extra = {k: z[k] for k in z.files if not k.startswith("s") or not k[1:2].isdigit()}
    return subs, extra
