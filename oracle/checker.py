"""oracle/checker.py - TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes front-ends for the two CPU checkers:

* ``Checker("ref")``    -> oracle/_ref/libcabac_ref.so: the UNMODIFIED reference
  sources (/root/reference/src) behind oracle/ref_driver.cpp.
* ``Checker("oracle")`` -> oracle/libcabac_oracle.so: the plain-C restatement
  (oracle/cabac_oracle.c).

Both export the same entry points (prefix ``ref_`` / ``orc_``).  Only tests/,
``__graft_entry__.smoke()`` and bench.py's CPU legs may import this module; the
product package (entropy_coding_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
NUM_CTX = 379

# numpy mirrors of include/cabac_b200.h
PARAMS_DT = np.dtype([("qp", "i1"), ("init_id", "u1"), ("flags", "u1"), ("reserved", "u1")])
TU_DT = np.dtype(
    [("coeff_offset", "<u8"), ("log2_w", "u1"), ("log2_h", "u1"), ("comp_id", "u1"),
     ("flags", "u1"), ("reserved", "<u4")]
)
CTX_DT = np.dtype([("state0", "<u2"), ("state1", "<u2")])
assert PARAMS_DT.itemsize == 4 and TU_DT.itemsize == 16 and CTX_DT.itemsize == 4

SUB_DEP_QUANT, SUB_SIGN_HIDING, SUB_TSRC_DISABLED = 1, 2, 4
SUB_TERMINATE, SUB_ALIGN, SUB_CTX_IN = 8, 16, 32
TU_TS, TU_TS_ALLOWED, TU_BDPCM, TU_SBT_ZEROOUT = 1, 2, 4, 8
CMD_CTX, CMD_EP, CMD_REM, CMD_TRM, CMD_ALIGN, CMD_TU = 0, 1, 2, 3, 4, 5


def mk_ctx(ctx, b):
    return (CMD_CTX << 29) | ((b & 1) << 9) | (ctx & 511)


def mk_ep(bins, n):
    return (CMD_EP << 29) | ((n & 31) << 24) | (bins & 0xFFFFFF)


def mk_rem(value, rice):
    return (CMD_REM << 29) | ((rice & 31) << 24) | (value & 0xFFFFFF)


def mk_trm(b):
    return (CMD_TRM << 29) | (b & 1)


def mk_align():
    return CMD_ALIGN << 29


def mk_tu():
    return CMD_TU << 29


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def lib_path(kind: str) -> str:
    if kind == "ref":
        return os.path.join(HERE, "_ref", "libcabac_ref.so")
    return os.path.join(HERE, "libcabac_oracle.so")


def available(kind: str) -> bool:
    return os.path.exists(lib_path(kind))


class NalRef:
    """VTM's own NALwrite.cpp / NALread.cpp (oracle/_ref/libnal_ref.so, built by `make -C oracle nalref`)."""
    PATH = os.path.join(HERE, "_ref", "libnal_ref.so")

    @classmethod
    def available(cls) -> bool:
        return os.path.exists(cls.PATH)

    def __init__(self):
        self.lib = C.CDLL(self.PATH)
        self.lib.nalref_insert.restype = C.c_long
        self.lib.nalref_remove.restype = C.c_long

    def insert(self, rbsp) -> np.ndarray:
        rbsp = np.ascontiguousarray(rbsp, np.uint8)
        out = np.zeros(2 * rbsp.size + 2, np.uint8)
        n = self.lib.nalref_insert(_ptr(rbsp), C.c_long(rbsp.size), _ptr(out))
        return out[:n].copy()

    def remove(self, nal):
        """NAL unit incl. its 2 header bytes -> (rbsp incl. header, epb_locations) or None when VTM throws."""
        nal = np.ascontiguousarray(nal, np.uint8)
        rbsp = np.zeros(nal.size + 1, np.uint8)
        epb = np.zeros(nal.size // 3 + 2, np.uint32)
        k = C.c_long(0)
        n = self.lib.nalref_remove(_ptr(nal), C.c_long(nal.size), _ptr(rbsp), _ptr(epb), C.byref(k))
        return None if n < 0 else (rbsp[:n].copy(), epb[:k.value].copy())


class Checker:
    def __init__(self, kind: str = "oracle"):
        assert kind in ("ref", "oracle")
        self.kind = kind
        self.pfx = "ref_" if kind == "ref" else "orc_"
        path = lib_path(kind)
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} missing: run `make -C oracle` (or __graft_entry__.build())")
        # The reference's loggers open bin_log.txt / bit_log.txt in the CWD at
        # static-init time (src/log/log.cpp:3-4): load from a scratch directory.
        cwd = os.getcwd()
        with tempfile.TemporaryDirectory() as tmp:
            os.chdir(tmp)
            try:
                self.lib = C.CDLL(path)
            finally:
                os.chdir(cwd)
        for name in ("encode_bins", "decode_bins", "encode_residual", "decode_residual", "encode_stream"):
            getattr(self.lib, self.pfx + name).restype = C.c_long
        for name in ("time_encode_residual", "time_decode_residual"):
            getattr(self.lib, self.pfx + name).restype = C.c_double

    def _f(self, name):
        return getattr(self.lib, self.pfx + name)

    # ---- tables ---------------------------------------------------------
    def init_table(self, row: int) -> np.ndarray:
        out = np.zeros(NUM_CTX, np.uint8)
        self._f("init_table")(C.c_int(row), _ptr(out))
        return out

    def ctx_init(self, qp: int, init_id: int):
        out = np.zeros(NUM_CTX, CTX_DT)
        rates = np.zeros(NUM_CTX, np.uint8)
        self._f("ctx_init")(C.c_int(qp), C.c_int(init_id), _ptr(out), _ptr(rates))
        return out, rates

    def scan_table(self, log2w: int, log2h: int):
        out = np.zeros(1 << (log2w + log2h), np.uint32)
        r = self._f("scan_table")(C.c_int(log2w), C.c_int(log2h), _ptr(out))
        return out, r & 255, r >> 8

    def ctx_offsets(self) -> np.ndarray:
        out = np.zeros(25, np.uint32)
        self._f("ctx_offsets")(_ptr(out))
        return out

    # ---- start-code emulation ----------------------------------------------
    def count_emulations(self, data) -> int:
        data = np.ascontiguousarray(data, np.uint8)
        f = self._f("count_emulations")
        f.restype = C.c_long
        return int(f(_ptr(data), C.c_long(data.size)))

    def insert_emulation_prevention(self, data) -> np.ndarray:
        """Oracle only (the function lives in the VTM submodule, not in the reference's src/)."""
        data = np.ascontiguousarray(data, np.uint8)
        out = np.zeros(data.size + data.size // 2 + 2, np.uint8)
        f = self.lib.orc_insert_emulation_prevention
        f.restype = C.c_long
        return out[:int(f(_ptr(data), C.c_long(data.size), _ptr(out)))].copy()

    def remove_emulation_prevention(self, nal, vcl=False):
        """Oracle only: (rbsp, epb_locations) or None where VTM's CHECKs would throw (NALread.cpp:59-106)."""
        nal = np.ascontiguousarray(nal, np.uint8)
        rbsp = np.zeros(nal.size + 1, np.uint8)
        epb = np.zeros(nal.size // 3 + 2, np.uint32)
        k = C.c_long(0)
        f = self.lib.orc_remove_emulation_prevention
        f.restype = C.c_long
        n = int(f(_ptr(nal), C.c_long(nal.size), C.c_int(int(vcl)), _ptr(rbsp), _ptr(epb), C.byref(k)))
        return None if n < 0 else (rbsp[:n].copy(), epb[:k.value].copy())

    def entry_point_syntax(self, sizes, emulations):
        """Oracle only: (sh_entry_offset_len_minus1, sh_entry_point_offset_minus1[n_sub - 1])."""
        sizes = np.ascontiguousarray(sizes, np.uint32)
        emulations = np.ascontiguousarray(emulations, np.uint32)
        out = np.zeros(max(len(sizes) - 1, 0), np.uint32)
        ln = self.lib.orc_entry_point_syntax(_ptr(sizes), _ptr(emulations), C.c_long(len(sizes)), _ptr(out))
        return int(ln), out

    def substream_ranges(self, offset_minus1, epb, hdr_end, rbsp_size):
        """Oracle only: RBSP byte ranges of the substreams of a slice (in_offsets[n_entry + 2])."""
        off = np.ascontiguousarray(offset_minus1, np.uint32)
        epb = np.ascontiguousarray(epb, np.uint32)
        out = np.zeros(len(off) + 2, np.uint64)
        self.lib.orc_substream_ranges(_ptr(off), C.c_long(len(off)), _ptr(epb), C.c_long(len(epb)),
                                      C.c_uint64(int(hdr_end)), C.c_uint64(int(rbsp_size)), _ptr(out))
        return out

    # ---- bin level --------------------------------------------------------
    def encode_bins(self, cmds, qp, init_id, flags=SUB_TERMINATE | SUB_ALIGN, ctx_in=None,
                    want_ctx=False, cap=None):
        cmds = np.ascontiguousarray(cmds, np.uint32)
        p = np.zeros(1, PARAMS_DT)
        p["qp"], p["init_id"], p["flags"] = qp, init_id, flags | (SUB_CTX_IN if ctx_in is not None else 0)
        cap = cap or (len(cmds) * 5 + 64)
        out = np.zeros(cap, np.uint8)
        nb = np.zeros(1, np.uint32)
        ctx_out = np.zeros(NUM_CTX, CTX_DT) if want_ctx else None
        n = self._f("encode_bins")(_ptr(cmds), C.c_long(len(cmds)), _ptr(p), _ptr(ctx_in), _ptr(out),
                                   C.c_long(cap), _ptr(nb), _ptr(ctx_out))
        assert 0 <= n <= cap, n
        res = (out[:n].copy(), int(nb[0]))
        return res + (ctx_out,) if want_ctx else res

    def decode_bins(self, data, cmds, qp, init_id, flags=SUB_TERMINATE, ctx_in=None, want_ctx=False):
        data = np.ascontiguousarray(data, np.uint8)
        cmds = np.ascontiguousarray(cmds, np.uint32)
        p = np.zeros(1, PARAMS_DT)
        p["qp"], p["init_id"], p["flags"] = qp, init_id, flags | (SUB_CTX_IN if ctx_in is not None else 0)
        vals = np.zeros(len(cmds), np.uint32)
        st = np.zeros(1, np.uint32)
        ctx_out = np.zeros(NUM_CTX, CTX_DT) if want_ctx else None
        bits = self._f("decode_bins")(_ptr(data), C.c_long(len(data)), _ptr(cmds), C.c_long(len(cmds)), _ptr(p),
                                      _ptr(ctx_in), _ptr(vals), _ptr(ctx_out), _ptr(st))
        res = (vals, int(bits), int(st[0]))
        return res + (ctx_out,) if want_ctx else res

    # ---- residual level ---------------------------------------------------
    def encode_residual(self, tus, coeffs, qp, init_id, flags, ctx_in=None, want_ctx=False, cap=None,
                        use_mts=False):
        tus = np.ascontiguousarray(tus, TU_DT)
        coeffs = np.ascontiguousarray(coeffs, np.int32)
        p = np.zeros(1, PARAMS_DT)
        p["qp"], p["init_id"], p["flags"] = qp, init_id, flags | (SUB_CTX_IN if ctx_in is not None else 0)
        cap = cap or (coeffs.size * 5 + 64 * len(tus) + 64)
        out = np.zeros(cap, np.uint8)
        nb = np.zeros(1, np.uint32)
        tuf = np.zeros(len(tus), np.uint8)
        ctx_out = np.zeros(NUM_CTX, CTX_DT) if want_ctx else None
        n = self._f("encode_residual")(_ptr(tus), C.c_long(len(tus)), _ptr(p), _ptr(coeffs), _ptr(ctx_in),
                                       _ptr(out), C.c_long(cap), _ptr(nb), _ptr(tuf), _ptr(ctx_out),
                                       C.c_int(int(use_mts)))
        assert 0 <= n <= cap, n
        res = (out[:n].copy(), int(nb[0]), tuf)
        return res + (ctx_out,) if want_ctx else res

    def frac_bits(self) -> np.ndarray:
        """BinProbModelBase::m_binFracBits flattened [state][bin] (contexts.cpp:791-878)."""
        out = np.zeros(512, np.uint32)
        self._f("frac_bits")(_ptr(out))
        return out

    def estimate_residual(self, tus, coeffs, qp, init_id, flags=0, ctx_in=None, want_ctx=False, use_mts=False):
        """BitEstimator_Std behind CABACWriter::residual_coding: (total, per-TU running totals[, ctx])."""
        tus = np.ascontiguousarray(tus, TU_DT)
        coeffs = np.ascontiguousarray(coeffs, np.int32)
        p = np.zeros(1, PARAMS_DT)
        p["qp"], p["init_id"], p["flags"] = qp, init_id, flags | (SUB_CTX_IN if ctx_in is not None else 0)
        per_tu = np.zeros(len(tus), np.uint64)
        ctx_out = np.zeros(NUM_CTX, CTX_DT) if want_ctx else None
        f = self._f("estimate_residual")
        f.restype = C.c_uint64
        tot = int(f(_ptr(tus), C.c_long(len(tus)), _ptr(p), _ptr(coeffs), _ptr(ctx_in), _ptr(ctx_out),
                    _ptr(per_tu), C.c_int(int(use_mts))))
        assert tot != 0xFFFFFFFFFFFFFFFF
        return (tot, per_tu, ctx_out) if want_ctx else (tot, per_tu)

    def residual_ctx_bins(self, tus, coeffs, qp, init_id, flags=0, ctx_in=None):
        """BinCounter per context after residual_coding over the list: (ctx_bins[379], ctx_out, choice).
        choice = CABACWriter::getCtxInitId on that coder (cabac_writer.cpp:41-81): the compiled
        reference evaluates it itself, the restatement through orc_ctx_init_id."""
        tus = np.ascontiguousarray(tus, TU_DT)
        coeffs = np.ascontiguousarray(coeffs, np.int32)
        p = np.zeros(1, PARAMS_DT)
        p["qp"], p["init_id"], p["flags"] = qp, init_id, flags | (SUB_CTX_IN if ctx_in is not None else 0)
        bins = np.zeros(NUM_CTX, np.uint32)
        ctx_out = np.zeros(NUM_CTX, CTX_DT)
        f = self._f("residual_ctx_bins")
        f.restype = C.c_long
        if self.pfx == "ref_":
            ch = C.c_int(-1)
            n = f(_ptr(tus), C.c_long(len(tus)), _ptr(p), _ptr(coeffs), _ptr(ctx_in), _ptr(ctx_out), _ptr(bins),
                  C.byref(ch))
            choice = ch.value
        else:
            n = f(_ptr(tus), C.c_long(len(tus)), _ptr(p), _ptr(coeffs), _ptr(ctx_in), _ptr(ctx_out), _ptr(bins))
            choice = 2 if init_id == 2 else self.ctx_init_id(qp, ctx_out, bins)[0]
        assert n >= 0
        return bins, ctx_out, choice

    def ctx_init_id(self, qp, ctx_ref, ctx_bins):
        """xGetCtxInitId for a non-intra slice (restatement only): (choice, costs[2])."""
        assert self.pfx == "orc_"
        costs = np.zeros(2, np.uint64)
        ctx_ref = np.ascontiguousarray(ctx_ref, CTX_DT)
        ctx_bins = np.ascontiguousarray(ctx_bins, np.uint32)
        ch = self.lib.orc_ctx_init_id(C.c_int(int(qp)), _ptr(ctx_ref), _ptr(ctx_bins), _ptr(costs))
        return int(ch), costs

    def encode_stream(self, cmds, tus, coeffs, qp, init_id, flags, ctx_in=None, want_ctx=False, cap=None):
        """Recorded commands with CMD_TU markers (marker k codes tus[k])."""
        cmds = np.ascontiguousarray(cmds, np.uint32)
        tus = np.ascontiguousarray(tus, TU_DT)
        coeffs = np.ascontiguousarray(coeffs, np.int32)
        p = np.zeros(1, PARAMS_DT)
        p["qp"], p["init_id"], p["flags"] = qp, init_id, flags | (SUB_CTX_IN if ctx_in is not None else 0)
        cap = cap or (coeffs.size * 5 + 64 * len(tus) + 5 * len(cmds) + 64)
        out = np.zeros(cap, np.uint8)
        nb = np.zeros(1, np.uint32)
        tuf = np.zeros(len(tus), np.uint8)
        ctx_out = np.zeros(NUM_CTX, CTX_DT) if want_ctx else None
        n = self._f("encode_stream")(_ptr(cmds), C.c_long(len(cmds)), _ptr(tus), C.c_long(len(tus)), _ptr(p),
                                     _ptr(coeffs), _ptr(ctx_in), _ptr(out), C.c_long(cap), _ptr(nb), _ptr(tuf),
                                     _ptr(ctx_out))
        assert 0 <= n <= cap, n
        res = (out[:n].copy(), int(nb[0]), tuf)
        return res + (ctx_out,) if want_ctx else res

    def decode_residual(self, data, tus, n_coeffs, qp, init_id, flags, ctx_in=None, want_ctx=False,
                        use_mts=False):
        data = np.ascontiguousarray(data, np.uint8)
        tus = np.ascontiguousarray(tus, TU_DT)
        p = np.zeros(1, PARAMS_DT)
        p["qp"], p["init_id"], p["flags"] = qp, init_id, flags | (SUB_CTX_IN if ctx_in is not None else 0)
        coeffs = np.zeros(n_coeffs, np.int32)
        tuf = np.zeros(len(tus), np.uint8)
        st = np.zeros(1, np.uint32)
        ctx_out = np.zeros(NUM_CTX, CTX_DT) if want_ctx else None
        bits = self._f("decode_residual")(_ptr(data), C.c_long(len(data)), _ptr(tus), C.c_long(len(tus)), _ptr(p),
                                          _ptr(ctx_in), _ptr(coeffs), _ptr(tuf), _ptr(ctx_out), _ptr(st),
                                          C.c_int(int(use_mts)))
        res = (coeffs, int(bits), tuf, int(st[0]))
        return res + (ctx_out,) if want_ctx else res

    # ---- timed batches (CPU baseline) ------------------------------------
    def time_encode_residual(self, tus, tu_offsets, params, coeffs, nthreads, out=None, out_offsets=None,
                             out_sizes=None):
        tus = np.ascontiguousarray(tus, TU_DT)
        tu_offsets = np.ascontiguousarray(tu_offsets, np.uint64)
        params = np.ascontiguousarray(params, PARAMS_DT)
        coeffs = np.ascontiguousarray(coeffs, np.int32)
        bins = np.zeros(1, np.uint64)
        nbytes = np.zeros(1, np.uint64)
        t = self._f("time_encode_residual")(_ptr(tus), _ptr(tu_offsets), _ptr(params), C.c_long(len(params)),
                                            _ptr(coeffs), C.c_int(nthreads), _ptr(bins), _ptr(nbytes), _ptr(out),
                                            _ptr(out_offsets), _ptr(out_sizes))
        return float(t), int(bins[0]), int(nbytes[0])

    def time_decode_residual(self, data, in_offsets, in_sizes, tus, tu_offsets, params, coeffs_out, nthreads):
        t = self._f("time_decode_residual")(_ptr(data), _ptr(in_offsets), _ptr(in_sizes), _ptr(tus),
                                            _ptr(tu_offsets), _ptr(params), C.c_long(len(params)),
                                            _ptr(coeffs_out), C.c_int(nthreads))
        return float(t)
