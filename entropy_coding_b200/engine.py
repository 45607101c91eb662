"""Python face of the C ABI (include/cabac_b200.h).

`CabacEngine` owns one handle (one CUDA stream on one device).  Methods named
after the C entry points take numpy arrays (host pointers; copies happen inside
the library) or, for the `*_dev` variants, torch CUDA tensors / raw device
pointers.  Nothing here computes: every call lands in a CUDA kernel, and fails
loudly when the library or a GPU is missing.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .types import (CTU_DT, CTX_DT, CU_DT, ERR_SUBSTREAM, NUM_CTX, OK, PARAMS_DT, SL_WPP, SLICE_DT, TU_DT, WPP_DT)


class CabacError(RuntimeError):
    """Mirrors the reference's Common::Exception (CHECK/THROW, type_def.hpp:295-329)."""

    def __init__(self, code, msg, status=None):
        super().__init__(f"cabac_b200 error {code}: {msg}")
        self.code = code
        self.status = status


def _np(a, dt):
    if a is None:
        return None
    a = np.ascontiguousarray(a, dt)
    return a


def _hp(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _dp(t):
    """Device pointer of a torch tensor (or a raw int / None)."""
    if t is None:
        return None
    if isinstance(t, int):
        return C.c_void_p(t)
    assert t.is_cuda and t.is_contiguous(), "device tensors must be contiguous CUDA tensors"
    return C.c_void_p(t.data_ptr())


class CabacEngine:
    def __init__(self, device: int = -1):
        self.lib = _lib.load()
        self.h = self.lib.cabac_create(device)
        if not self.h:
            raise CabacError(-1, "cabac_create failed: no usable CUDA device (there is no CPU fallback)")

    def close(self):
        if getattr(self, "h", None):
            self.lib.cabac_destroy(self.h)
            self.h = None
            for p in getattr(self, "_pinned", []):
                self.lib.cabac_host_free(p)
            self._pinned = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- plumbing ---------------------------------------------------------
    def _check(self, rc, status=None):
        if rc == OK:
            return
        if rc == ERR_SUBSTREAM:
            raise CabacError(rc, "one or more substreams failed", status)
        raise CabacError(rc, self.lib.cabac_last_error(self.h).decode())

    def set_stream(self, cuda_stream_ptr: int | None):
        self._check(self.lib.cabac_set_stream(self.h, C.c_void_p(cuda_stream_ptr or 0)))

    def set_encoder_mode(self, mode: int):
        """0 = auto, 1 = fused lockstep kernel, 2 = two-stage (stage P + stage D)."""
        self._check(self.lib.cabac_set_encoder_mode(self.h, int(mode)))

    def set_decoder_mode(self, mode: int):
        """0 = auto, 1 = lockstep (one warp per substream), 2 = one thread per substream."""
        self._check(self.lib.cabac_set_decoder_mode(self.h, int(mode)))

    def sync(self):
        self._check(self.lib.cabac_sync(self.h))

    @property
    def launch_count(self) -> int:
        return int(self.lib.cabac_launch_count(self.h))

    # ---- pinned host buffers -------------------------------------------------
    def host_array(self, n: int, dtype) -> np.ndarray:
        """numpy array over page-locked memory (cabac_host_alloc); freed with the engine."""
        dt = np.dtype(dtype)
        p = self.lib.cabac_host_alloc(max(1, n * dt.itemsize))
        if not p:
            raise CabacError(-1, "cabac_host_alloc failed")
        self._pinned = getattr(self, "_pinned", [])
        self._pinned.append(p)
        buf = (C.c_uint8 * max(1, n * dt.itemsize)).from_address(p)
        return np.frombuffer(buf, dtype=dt, count=n)

    def pinned_copy(self, a: np.ndarray) -> np.ndarray:
        out = self.host_array(a.size, a.dtype)
        out[...] = a.reshape(-1)
        return out

    # ---- tables -------------------------------------------------------------
    def ctx_rates(self) -> np.ndarray:
        out = np.zeros(NUM_CTX, np.uint8)
        self._check(self.lib.cabac_ctx_rates(_hp(out)))
        return out

    def ctx_init_table(self, row: int) -> np.ndarray:
        out = np.zeros(NUM_CTX, np.uint8)
        self._check(self.lib.cabac_ctx_init_table(row, _hp(out)))
        return out

    def frac_bits_table(self) -> np.ndarray:
        """m_binFracBits as [2 * state + bin] (contexts.cpp:791-878)."""
        out = np.zeros(512, np.uint32)
        self._check(self.lib.cabac_frac_bits_table(_hp(out)))
        return out

    def ctx_init(self, params) -> np.ndarray:
        params = _np(params, PARAMS_DT)
        out = np.zeros((len(params), NUM_CTX), CTX_DT)
        self._check(self.lib.cabac_ctx_init(self.h, _hp(params), len(params), _hp(out)))
        return out

    # ---- host-pointer calls ---------------------------------------------------
    def encode_bins(self, cmds, cmd_offsets, params, out_offsets, ctx_in=None, want_ctx=False, raise_on_status=True):
        cmds, cmd_offsets = _np(cmds, np.uint32), _np(cmd_offsets, np.uint64)
        params, out_offsets = _np(params, PARAMS_DT), _np(out_offsets, np.uint64)
        n = len(params)
        out = np.zeros(int(out_offsets[-1]) if n else 0, np.uint8)
        sizes, bins, status = np.zeros(n, np.uint32), np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        ctx_in = _np(ctx_in, CTX_DT)
        ctx_out = np.zeros((n, NUM_CTX), CTX_DT) if want_ctx else None
        rc = self.lib.cabac_encode_bins(self.h, _hp(cmds), _hp(cmd_offsets), _hp(params), n, _hp(ctx_in), _hp(out),
                                        _hp(out_offsets), _hp(sizes), _hp(bins), _hp(ctx_out), _hp(status))
        if rc != ERR_SUBSTREAM or raise_on_status:
            self._check(rc, status)
        return dict(out=out, sizes=sizes, num_bins=bins, status=status, ctx_out=ctx_out)

    def encode_stream(self, cmds, cmd_offsets, tus, tu_offsets, params, coeffs, out_offsets, ctx_in=None,
                      want_ctx=False, raise_on_status=True):
        """Recorded command streams with CABAC_CMD_TU markers (the CABACWriter seam)."""
        cmds, cmd_offsets = _np(cmds, np.uint32), _np(cmd_offsets, np.uint64)
        tus, tu_offsets = _np(tus, TU_DT), _np(tu_offsets, np.uint64)
        params, coeffs, out_offsets = _np(params, PARAMS_DT), _np(coeffs, np.int32), _np(out_offsets, np.uint64)
        n = len(params)
        out = np.zeros(int(out_offsets[-1]) if n else 0, np.uint8)
        sizes, bins, status = np.zeros(n, np.uint32), np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        tu_flags = np.zeros(len(tus), np.uint8)
        ctx_in = _np(ctx_in, CTX_DT)
        ctx_out = np.zeros((n, NUM_CTX), CTX_DT) if want_ctx else None
        rc = self.lib.cabac_encode_stream(self.h, _hp(cmds), _hp(cmd_offsets), _hp(tus), _hp(tu_offsets), _hp(params),
                                          n, _hp(coeffs), coeffs.size, _hp(ctx_in), _hp(out), _hp(out_offsets),
                                          _hp(sizes), _hp(bins), _hp(tu_flags), _hp(ctx_out), _hp(status))
        if rc != ERR_SUBSTREAM or raise_on_status:
            self._check(rc, status)
        return dict(out=out, sizes=sizes, num_bins=bins, tu_flags=tu_flags, status=status, ctx_out=ctx_out)

    def decode_bins(self, data, in_offsets, cmds, cmd_offsets, params, ctx_in=None, want_ctx=False,
                    raise_on_status=True):
        data, in_offsets = _np(data, np.uint8), _np(in_offsets, np.uint64)
        cmds, cmd_offsets, params = _np(cmds, np.uint32), _np(cmd_offsets, np.uint64), _np(params, PARAMS_DT)
        n = len(params)
        values = np.zeros(len(cmds), np.uint32)
        bits, status = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        ctx_in = _np(ctx_in, CTX_DT)
        ctx_out = np.zeros((n, NUM_CTX), CTX_DT) if want_ctx else None
        rc = self.lib.cabac_decode_bins(self.h, _hp(data), _hp(in_offsets), _hp(cmds), _hp(cmd_offsets), _hp(params),
                                        n, _hp(ctx_in), _hp(values), _hp(bits), _hp(ctx_out), _hp(status))
        if rc != ERR_SUBSTREAM or raise_on_status:
            self._check(rc, status)
        return dict(values=values, bits_read=bits, status=status, ctx_out=ctx_out)

    def encode_residual(self, tus, tu_offsets, params, coeffs, out_offsets, ctx_in=None, want_ctx=False,
                        raise_on_status=True):
        tus, tu_offsets = _np(tus, TU_DT), _np(tu_offsets, np.uint64)
        params, coeffs, out_offsets = _np(params, PARAMS_DT), _np(coeffs, np.int32), _np(out_offsets, np.uint64)
        n = len(params)
        out = np.zeros(int(out_offsets[-1]) if n else 0, np.uint8)
        sizes, bins, status = np.zeros(n, np.uint32), np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        tu_flags = np.zeros(len(tus), np.uint8)
        ctx_in = _np(ctx_in, CTX_DT)
        ctx_out = np.zeros((n, NUM_CTX), CTX_DT) if want_ctx else None
        rc = self.lib.cabac_encode_residual(self.h, _hp(tus), _hp(tu_offsets), _hp(params), n, _hp(coeffs),
                                            coeffs.size, _hp(ctx_in), _hp(out), _hp(out_offsets), _hp(sizes),
                                            _hp(bins), _hp(tu_flags), _hp(ctx_out), _hp(status))
        if rc != ERR_SUBSTREAM or raise_on_status:
            self._check(rc, status)
        return dict(out=out, sizes=sizes, num_bins=bins, tu_flags=tu_flags, status=status, ctx_out=ctx_out)

    def estimate_residual_bits(self, tus, tu_offsets, params, coeffs, ctx_in=None, want_ctx=False,
                               want_per_tu=True, raise_on_status=True):
        """CABACWriter::residual_coding on BitEstimator_Std for every TU list (candidate) of the batch:
        est_bits in 1/32768 bit per list and, per TU, the running total after it."""
        tus, tu_offsets = _np(tus, TU_DT), _np(tu_offsets, np.uint64)
        params, coeffs = _np(params, PARAMS_DT), _np(coeffs, np.int32)
        n = len(params)
        est, bins, status = np.zeros(n, np.uint64), np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        per_tu = np.zeros(len(tus), np.uint64) if want_per_tu else None
        tu_flags = np.zeros(len(tus), np.uint8)
        ctx_in = _np(ctx_in, CTX_DT)
        ctx_out = np.zeros((n, NUM_CTX), CTX_DT) if want_ctx else None
        rc = self.lib.cabac_estimate_residual_bits(self.h, _hp(tus), _hp(tu_offsets), _hp(params), n, _hp(coeffs),
                                                   coeffs.size, _hp(ctx_in), _hp(est), _hp(per_tu), _hp(bins),
                                                   _hp(tu_flags), _hp(ctx_out), _hp(status))
        if rc != ERR_SUBSTREAM or raise_on_status:
            self._check(rc, status)
        return dict(est_bits=est, est_per_tu=per_tu, num_bins=bins, tu_flags=tu_flags, status=status,
                    ctx_out=ctx_out)

    def residual_ctx_bins(self, tus, tu_offsets, params, coeffs, ctx_in=None, raise_on_status=True):
        """BinEncIf::getNumBins(ctxId) per TU list after residual_coding (BinCounter, arith_codec.hpp:72-93):
        ctx_bins[n, 379], with the final contexts - the two inputs of CABACWriter::getCtxInitId."""
        tus, tu_offsets = _np(tus, TU_DT), _np(tu_offsets, np.uint64)
        params, coeffs = _np(params, PARAMS_DT), _np(coeffs, np.int32)
        n = len(params)
        ctx_bins = np.zeros((n, NUM_CTX), np.uint32)
        bins, status = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        ctx_in = _np(ctx_in, CTX_DT)
        ctx_out = np.zeros((n, NUM_CTX), CTX_DT)
        rc = self.lib.cabac_residual_ctx_bins(self.h, _hp(tus), _hp(tu_offsets), _hp(params), n, _hp(coeffs),
                                              coeffs.size, _hp(ctx_in), _hp(ctx_bins), _hp(bins), _hp(ctx_out),
                                              _hp(status))
        if rc != ERR_SUBSTREAM or raise_on_status:
            self._check(rc, status)
        return dict(ctx_bins=ctx_bins, num_bins=bins, ctx_out=ctx_out, status=status)

    def ctx_init_id(self, params, ctx_ref, ctx_bins):
        """CABACWriter::getCtxInitId (cabac_writer.cpp:41-81) for n slices: (choice[n], costs[n, 2])."""
        params, ctx_ref = _np(params, PARAMS_DT), _np(ctx_ref, CTX_DT)
        ctx_bins = _np(ctx_bins, np.uint32)
        n = len(params)
        choice, costs = np.zeros(n, np.uint8), np.zeros((n, 2), np.uint64)
        self._check(self.lib.cabac_ctx_init_id(self.h, _hp(params), _hp(ctx_ref), _hp(ctx_bins), n, _hp(choice),
                                               _hp(costs)))
        return choice, costs

    def encode_residual_packed(self, tus, tu_offsets, params, coeffs, out_offsets, packed=None, ctx_in=None,
                               want_ctx=False, raise_on_status=True, i16=False):
        """Encode + pack: returns the substreams concatenated in order plus the entry-point table.
        i16: the coefficient slab crosses the ABI as int16 (cabac_encode_residual_packed_i16)."""
        tus, tu_offsets = _np(tus, TU_DT), _np(tu_offsets, np.uint64)
        params, out_offsets = _np(params, PARAMS_DT), _np(out_offsets, np.uint64)
        coeffs = _np(coeffs, np.int16 if i16 else np.int32)
        n = len(params)
        if packed is None:
            packed = np.zeros(int(out_offsets[-1]) if n else 0, np.uint8)
        entry = np.zeros(n + 1, np.uint64)
        sizes, bins, status = np.zeros(n, np.uint32), np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        tu_flags = np.zeros(len(tus), np.uint8)
        ctx_in = _np(ctx_in, CTX_DT)
        ctx_out = np.zeros((n, NUM_CTX), CTX_DT) if want_ctx else None
        fn = self.lib.cabac_encode_residual_packed_i16 if i16 else self.lib.cabac_encode_residual_packed
        rc = fn(self.h, _hp(tus), _hp(tu_offsets), _hp(params), n, _hp(coeffs),
                                                   coeffs.size, _hp(ctx_in), _hp(out_offsets), _hp(packed),
                                                   packed.size, _hp(entry), _hp(sizes), _hp(bins), _hp(tu_flags),
                                                   _hp(ctx_out), _hp(status))
        if rc != ERR_SUBSTREAM or raise_on_status:
            self._check(rc, status)
        return dict(packed=packed[:int(entry[-1])], entry_points=entry, sizes=sizes, num_bins=bins,
                    tu_flags=tu_flags, status=status, ctx_out=ctx_out)

    def decode_residual(self, data, in_offsets, tus, tu_offsets, params, n_coeffs, ctx_in=None, want_ctx=False,
                        raise_on_status=True, coeffs_out=None, i16=False):
        data, in_offsets = _np(data, np.uint8), _np(in_offsets, np.uint64)
        tus, tu_offsets, params = _np(tus, TU_DT), _np(tu_offsets, np.uint64), _np(params, PARAMS_DT)
        n = len(params)
        cdt = np.int16 if i16 else np.int32
        coeffs = np.zeros(n_coeffs, cdt) if coeffs_out is None else coeffs_out
        assert coeffs.dtype == cdt and coeffs.size >= n_coeffs and coeffs.flags.c_contiguous
        bins, bits, status = np.zeros(n, np.uint32), np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        tu_flags = np.zeros(len(tus), np.uint8)
        ctx_in = _np(ctx_in, CTX_DT)
        ctx_out = np.zeros((n, NUM_CTX), CTX_DT) if want_ctx else None
        fn = self.lib.cabac_decode_residual_i16 if i16 else self.lib.cabac_decode_residual
        rc = fn(self.h, _hp(data), _hp(in_offsets), _hp(tus), _hp(tu_offsets),
                                            _hp(params), n, _hp(ctx_in), _hp(coeffs), n_coeffs, _hp(bins), _hp(bits),
                                            _hp(tu_flags), _hp(ctx_out), _hp(status))
        if rc != ERR_SUBSTREAM or raise_on_status:
            self._check(rc, status)
        return dict(coeffs=coeffs, num_bins=bins, bits_read=bits, tu_flags=tu_flags, status=status, ctx_out=ctx_out)

    # ---- whole substreams: CTU-level syntax on the device (cabac_decode_slices) ----
    @staticmethod
    def slice_capacities(slice_params):
        """Record / coefficient capacities that always suffice for each substream's CTU rectangle."""
        sp = _np(slice_params, SLICE_DT)
        area = (sp["x1"].astype(np.int64) - sp["x0"]) * (sp["y1"].astype(np.int64) - sp["y0"])
        ctu = np.int64(1) << sp["log2_ctu"].astype(np.int64)
        n_ctu = (-(-(sp["x1"].astype(np.int64) - sp["x0"]) // ctu)) * (-(-(sp["y1"].astype(np.int64) - sp["y0"]) // ctu))

        def off(counts):
            return np.concatenate([[0], np.cumsum(counts)]).astype(np.uint64)
        return dict(cu=off(area // 16 * 2 + 8), tu=off(area // 16 * 2 + 8), ctu=off(n_ctu), coeff=off(area * 3 // 2 + 64))

    def decode_slices(self, data, in_offsets, params, slice_params, ctx_in=None, want_ctx=False, raise_on_status=True):
        data, in_offsets = _np(data, np.uint8), _np(in_offsets, np.uint64)
        params, sp = _np(params, PARAMS_DT), _np(slice_params, SLICE_DT)
        n = len(params)
        cap = self.slice_capacities(sp)
        cus, tus = np.zeros(int(cap["cu"][-1]), CU_DT), np.zeros(int(cap["tu"][-1]), TU_DT)
        tu_flags, ctus = np.zeros(int(cap["tu"][-1]), np.uint8), np.zeros(int(cap["ctu"][-1]), CTU_DT)
        coeffs = np.zeros(int(cap["coeff"][-1]), np.int32)
        counts = np.zeros((n, 3), np.uint32)
        bins, bits, status = np.zeros(n, np.uint32), np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        ctx_in = _np(ctx_in, CTX_DT)
        ctx_out = np.zeros((n, NUM_CTX), CTX_DT) if want_ctx else None
        rc = self.lib.cabac_decode_slices(self.h, _hp(data), _hp(in_offsets), _hp(params), _hp(sp), n, _hp(ctx_in),
                                          _hp(cap["cu"]), _hp(cap["tu"]), _hp(cap["ctu"]), _hp(cap["coeff"]), _hp(cus),
                                          _hp(tus), _hp(tu_flags), _hp(ctus), _hp(coeffs), _hp(counts), _hp(bins),
                                          _hp(bits), _hp(ctx_out), _hp(status))
        if rc != ERR_SUBSTREAM or raise_on_status:
            self._check(rc, status)
        return dict(cus=cus, tus=tus, tu_flags=tu_flags, ctus=ctus, coeffs=coeffs, counts=counts, cap=cap, num_bins=bins,
                    bits_read=bits, status=status, ctx_out=ctx_out)

    def decode_slices_dev(self, data, in_offsets, params, slice_params, n_sub, cu_offsets, tu_offsets, ctu_offsets,
                          coeff_offsets, cus, tus, tu_flags, ctus, coeffs, counts=None, num_bins=None, bits_read=None,
                          ctx_in=None, ctx_out=None, status=None):
        self._check(self.lib.cabac_decode_slices_dev(self.h, _dp(data), _dp(in_offsets), _dp(params), _dp(slice_params),
                                                     n_sub, _dp(ctx_in), _dp(cu_offsets), _dp(tu_offsets),
                                                     _dp(ctu_offsets), _dp(coeff_offsets), _dp(cus), _dp(tus),
                                                     _dp(tu_flags), _dp(ctus), _dp(coeffs), _dp(counts), _dp(num_bins),
                                                     _dp(bits_read), _dp(ctx_out), _dp(status)))

    def encode_slices(self, params, slice_params, cus, cu_offsets, tus, tu_flags, tu_offsets, ctus, ctu_offsets, coeffs,
                      out_offsets, cmd_words=None, ctx_in=None, want_ctx=False, raise_on_status=True):
        """cabac_encode_slices: the writer direction - CU / TU / CTU records (as decode_slices returns them, compacted
        to exact offsets) to slice payloads."""
        params, sp = _np(params, PARAMS_DT), _np(slice_params, SLICE_DT)
        cus, tus, ctus = _np(cus, CU_DT), _np(tus, TU_DT), _np(ctus, CTU_DT)
        tu_flags, coeffs = _np(tu_flags, np.uint8), _np(coeffs, np.int32)
        cu_offsets, tu_offsets = _np(cu_offsets, np.uint64), _np(tu_offsets, np.uint64)
        ctu_offsets, out_offsets = _np(ctu_offsets, np.uint64), _np(out_offsets, np.uint64)
        n = len(params)
        if cmd_words is None:  # the bound the library itself stages with
            cmd_words = 80 * len(cus) + len(tus) + 450 * len(ctus) + 64 * n + 64
        out = np.zeros(int(out_offsets[-1]) if n else 0, np.uint8)
        sizes, bins, status = np.zeros(n, np.uint32), np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        ctx_in = _np(ctx_in, CTX_DT)
        ctx_out = np.zeros((n, NUM_CTX), CTX_DT) if want_ctx else None
        rc = self.lib.cabac_encode_slices(self.h, _hp(params), _hp(sp), n, _hp(cus), _hp(cu_offsets), _hp(tus),
                                          _hp(tu_flags), _hp(tu_offsets), _hp(ctus), _hp(ctu_offsets), _hp(coeffs),
                                          coeffs.size, _hp(ctx_in), int(cmd_words), _hp(out), _hp(out_offsets), _hp(sizes),
                                          _hp(bins), _hp(ctx_out), _hp(status))
        if rc != ERR_SUBSTREAM or raise_on_status:
            self._check(rc, status)
        return dict(out=out, sizes=sizes, num_bins=bins, status=status, ctx_out=ctx_out)

    @staticmethod
    def slice_rows(slice_params):
        """row_offsets for cabac_encode_slices_wpp: one row substream per CTU row of a WPP slice, one per other slice."""
        sp = _np(slice_params, SLICE_DT)
        ctu = np.int64(1) << sp["log2_ctu"].astype(np.int64)
        rows = np.where(sp["tools"] & SL_WPP, -(-(sp["y1"].astype(np.int64) - sp["y0"]) // ctu), 1)
        return np.concatenate([[0], np.cumsum(rows)]).astype(np.uint64)

    def encode_slices_wpp(self, params, slice_params, cus, cu_offsets, tus, tu_flags, tu_offsets, ctus, ctu_offsets, coeffs,
                          out_offsets, cmd_words=None, ctx_in=None, want_ctx=False, raise_on_status=True):
        """cabac_encode_slices_wpp: like encode_slices for any mix of WPP and other slices; out_offsets, sizes, bins,
        status and contexts are per ROW substream (row_offsets = slice_rows(slice_params), also returned)."""
        params, sp = _np(params, PARAMS_DT), _np(slice_params, SLICE_DT)
        cus, tus, ctus = _np(cus, CU_DT), _np(tus, TU_DT), _np(ctus, CTU_DT)
        tu_flags, coeffs = _np(tu_flags, np.uint8), _np(coeffs, np.int32)
        cu_offsets, tu_offsets = _np(cu_offsets, np.uint64), _np(tu_offsets, np.uint64)
        ctu_offsets, out_offsets = _np(ctu_offsets, np.uint64), _np(out_offsets, np.uint64)
        n = len(params)
        row_offsets = self.slice_rows(sp)
        n_rows = int(row_offsets[-1])
        assert len(out_offsets) == n_rows + 1, "out_offsets is per row substream"
        if cmd_words is None:  # the bound the library itself stages with
            cmd_words = 80 * len(cus) + len(tus) + 450 * len(ctus) + 64 * n + 64
        out = np.zeros(int(out_offsets[-1]) if n else 0, np.uint8)
        sizes, bins, status = np.zeros(n_rows, np.uint32), np.zeros(n_rows, np.uint32), np.zeros(n_rows, np.uint32)
        ctx_in = _np(ctx_in, CTX_DT)
        ctx_out = np.zeros((n_rows, NUM_CTX), CTX_DT) if want_ctx else None
        rc = self.lib.cabac_encode_slices_wpp(self.h, _hp(params), _hp(sp), n, _hp(cus), _hp(cu_offsets), _hp(tus),
                                              _hp(tu_flags), _hp(tu_offsets), _hp(ctus), _hp(ctu_offsets), _hp(coeffs),
                                              coeffs.size, _hp(ctx_in), int(cmd_words), _hp(row_offsets), n_rows, _hp(out),
                                              _hp(out_offsets), _hp(sizes), _hp(bins), _hp(ctx_out), _hp(status))
        if rc != ERR_SUBSTREAM or raise_on_status:
            self._check(rc, status)
        return dict(out=out, sizes=sizes, num_bins=bins, status=status, ctx_out=ctx_out, row_offsets=row_offsets)

    def encode_slices_dev(self, params, slice_params, n_sub, cus, cu_offsets, tus, tu_flags, tu_offsets, ctus,
                          ctu_offsets, coeffs, cmd_scratch, cmd_scratch_words, out, out_offsets, out_sizes, status,
                          num_bins=None, ctx_in=None, ctx_out=None):
        self._check(self.lib.cabac_encode_slices_dev(self.h, _dp(params), _dp(slice_params), n_sub, _dp(cus),
                                                     _dp(cu_offsets), _dp(tus), _dp(tu_flags), _dp(tu_offsets), _dp(ctus),
                                                     _dp(ctu_offsets), _dp(coeffs), _dp(ctx_in), _dp(cmd_scratch),
                                                     int(cmd_scratch_words), _dp(out), _dp(out_offsets), _dp(out_sizes),
                                                     _dp(num_bins), _dp(ctx_out), _dp(status)))

    def encode_slices_wpp_dev(self, params, slice_params, n_sub, cus, cu_offsets, tus, tu_flags, tu_offsets, ctus,
                              ctu_offsets, coeffs, cmd_scratch, cmd_scratch_words, row_offsets, n_rows, out, out_offsets,
                              out_sizes, status, num_bins=None, ctx_in=None, ctx_out=None):
        """Device pointers; out_offsets / out_sizes / num_bins / ctx_out / status are per row substream."""
        self._check(self.lib.cabac_encode_slices_wpp_dev(self.h, _dp(params), _dp(slice_params), n_sub, _dp(cus),
                                                         _dp(cu_offsets), _dp(tus), _dp(tu_flags), _dp(tu_offsets),
                                                         _dp(ctus), _dp(ctu_offsets), _dp(coeffs), _dp(ctx_in),
                                                         _dp(cmd_scratch), int(cmd_scratch_words), _dp(row_offsets),
                                                         int(n_rows), _dp(out), _dp(out_offsets), _dp(out_sizes),
                                                         _dp(num_bins), _dp(ctx_out), _dp(status)))

    # ---- WPP: CTU rows with context hand-over in one launch (cabac_*_wpp) ----
    def encode_residual_wpp(self, tus, tu_offsets, params, wpp, coeffs, out_offsets, packed=None, ctx_in=None,
                            want_ctx=False, raise_on_status=True):
        """cabac_encode_residual_wpp: like encode_residual_packed, substream i starting from the contexts
        substream wpp[i].dep published after wpp[dep].sync_tus TUs."""
        tus, tu_offsets = _np(tus, TU_DT), _np(tu_offsets, np.uint64)
        params, coeffs, out_offsets = _np(params, PARAMS_DT), _np(coeffs, np.int32), _np(out_offsets, np.uint64)
        wpp = _np(wpp, WPP_DT)
        n = len(params)
        assert len(wpp) == n
        if packed is None:
            packed = np.zeros(int(out_offsets[-1]) if n else 0, np.uint8)
        entry = np.zeros(n + 1, np.uint64)
        sizes, bins, status = np.zeros(n, np.uint32), np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        tu_flags = np.zeros(len(tus), np.uint8)
        ctx_in = _np(ctx_in, CTX_DT)
        ctx_out = np.zeros((n, NUM_CTX), CTX_DT) if want_ctx else None
        rc = self.lib.cabac_encode_residual_wpp(self.h, _hp(tus), _hp(tu_offsets), _hp(params), _hp(wpp), n,
                                                _hp(coeffs), coeffs.size, _hp(ctx_in), _hp(out_offsets), _hp(packed),
                                                packed.size, _hp(entry), _hp(sizes), _hp(bins), _hp(tu_flags),
                                                _hp(ctx_out), _hp(status))
        if rc != ERR_SUBSTREAM or raise_on_status:
            self._check(rc, status)
        return dict(packed=packed[:int(entry[-1])], entry_points=entry, sizes=sizes, num_bins=bins,
                    tu_flags=tu_flags, status=status, ctx_out=ctx_out)

    def decode_residual_wpp(self, data, in_offsets, tus, tu_offsets, params, wpp, n_coeffs, ctx_in=None,
                            want_ctx=False, raise_on_status=True, coeffs_out=None):
        data, in_offsets = _np(data, np.uint8), _np(in_offsets, np.uint64)
        tus, tu_offsets, params = _np(tus, TU_DT), _np(tu_offsets, np.uint64), _np(params, PARAMS_DT)
        wpp = _np(wpp, WPP_DT)
        n = len(params)
        assert len(wpp) == n
        coeffs = np.zeros(n_coeffs, np.int32) if coeffs_out is None else coeffs_out
        assert coeffs.dtype == np.int32 and coeffs.size >= n_coeffs and coeffs.flags.c_contiguous
        bins, bits, status = np.zeros(n, np.uint32), np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        tu_flags = np.zeros(len(tus), np.uint8)
        ctx_in = _np(ctx_in, CTX_DT)
        ctx_out = np.zeros((n, NUM_CTX), CTX_DT) if want_ctx else None
        rc = self.lib.cabac_decode_residual_wpp(self.h, _hp(data), _hp(in_offsets), _hp(tus), _hp(tu_offsets),
                                                _hp(params), _hp(wpp), n, _hp(ctx_in), _hp(coeffs), n_coeffs,
                                                _hp(bins), _hp(bits), _hp(tu_flags), _hp(ctx_out), _hp(status))
        if rc != ERR_SUBSTREAM or raise_on_status:
            self._check(rc, status)
        return dict(coeffs=coeffs, num_bins=bins, bits_read=bits, tu_flags=tu_flags, status=status, ctx_out=ctx_out)

    def encode_stream_wpp(self, cmds, cmd_offsets, tus, tu_offsets, params, wpp, coeffs, out_offsets, ctx_in=None,
                          want_ctx=False, raise_on_status=True):
        """Recorded command streams with CABAC_CMD_TU markers and WPP hand-over after wpp[i].sync_cmds words."""
        cmds, cmd_offsets = _np(cmds, np.uint32), _np(cmd_offsets, np.uint64)
        tus, tu_offsets = _np(tus, TU_DT), _np(tu_offsets, np.uint64)
        params, coeffs, out_offsets = _np(params, PARAMS_DT), _np(coeffs, np.int32), _np(out_offsets, np.uint64)
        wpp = _np(wpp, WPP_DT)
        n = len(params)
        out = np.zeros(int(out_offsets[-1]) if n else 0, np.uint8)
        sizes, bins, status = np.zeros(n, np.uint32), np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        tu_flags = np.zeros(len(tus), np.uint8)
        ctx_in = _np(ctx_in, CTX_DT)
        ctx_out = np.zeros((n, NUM_CTX), CTX_DT) if want_ctx else None
        rc = self.lib.cabac_encode_stream_wpp(self.h, _hp(cmds), _hp(cmd_offsets), _hp(tus), _hp(tu_offsets),
                                              _hp(params), _hp(wpp), n, _hp(coeffs), coeffs.size, _hp(ctx_in),
                                              _hp(out), _hp(out_offsets), _hp(sizes), _hp(bins), _hp(tu_flags),
                                              _hp(ctx_out), _hp(status))
        if rc != ERR_SUBSTREAM or raise_on_status:
            self._check(rc, status)
        return dict(out=out, sizes=sizes, num_bins=bins, tu_flags=tu_flags, status=status, ctx_out=ctx_out)

    def count_emulations(self, data, offsets):
        """countStartCodeEmulations per segment (reference bit_stream.cpp:157-181)."""
        data, offsets = _np(data, np.uint8), _np(offsets, np.uint64)
        n = len(offsets) - 1
        counts = np.zeros(n, np.uint32)
        self._check(self.lib.cabac_count_emulations(self.h, _hp(data), _hp(offsets), n, _hp(counts)))
        return counts

    def insert_emulation_prevention(self, rbsp):
        """RBSP -> NAL payload bytes (VTM NALwrite.cpp:67-126)."""
        rbsp = _np(rbsp, np.uint8)
        out = np.zeros(rbsp.size + rbsp.size // 2 + 1, np.uint8)
        size = np.zeros(1, np.uint64)
        self._check(self.lib.cabac_insert_emulation_prevention(self.h, _hp(rbsp), rbsp.size, _hp(out), out.size,
                                                               _hp(size)))
        return out[:int(size[0])]

    def remove_emulation_prevention(self, nal, vcl=False, raise_on_violation=True):
        """NAL unit bytes -> dict(rbsp, epb_locations, violations) (VTM NALread.cpp:59-106)."""
        nal = _np(nal, np.uint8)
        rbsp = np.zeros(max(nal.size, 1), np.uint8)
        sizes = np.zeros(2, np.uint64)
        epb = np.zeros(nal.size // 3 + 1, np.uint32)
        viol = np.zeros(1, np.uint32)
        rc = self.lib.cabac_remove_emulation_prevention(self.h, _hp(nal), nal.size, int(vcl), _hp(rbsp), _hp(sizes),
                                                        _hp(epb), _hp(viol))
        if rc != OK and (raise_on_violation or not int(viol[0])):
            self._check(rc)
        return dict(rbsp=rbsp[:int(sizes[0])], epb_locations=epb[:int(sizes[1])], violations=int(viol[0]))

    def entry_point_syntax(self, sizes, emulations=None):
        """(sh_entry_offset_len_minus1, sh_entry_point_offset_minus1[n_sub - 1]) (VLCWriter.cpp:2833-2866)."""
        sizes = _np(sizes, np.uint32)
        emulations = _np(emulations, np.uint32)
        off = np.zeros(max(len(sizes) - 1, 1), np.uint32)
        ln = np.zeros(1, np.uint32)
        self._check(self.lib.cabac_entry_point_syntax(self.h, _hp(sizes), _hp(emulations), len(sizes), _hp(off), _hp(ln)))
        return int(ln[0]), off[:max(len(sizes) - 1, 0)]

    def substream_ranges(self, offset_minus1, epb_locations, slice_data_pos, rbsp_size):
        """RBSP byte ranges of a slice's substreams from the signalled entry points (VLCReader.cpp:4346-4381)."""
        off = _np(offset_minus1, np.uint32)
        epb = _np(epb_locations, np.uint32)
        out = np.zeros(len(off) + 2, np.uint64)
        self._check(self.lib.cabac_substream_ranges(self.h, _hp(off) if len(off) else None, len(off),
                                                    _hp(epb) if len(epb) else None, len(epb), int(slice_data_pos),
                                                    int(rbsp_size), _hp(out)))
        return out

    def pack(self, out, out_offsets, sizes):
        out, out_offsets, sizes = _np(out, np.uint8), _np(out_offsets, np.uint64), _np(sizes, np.uint32)
        n = len(sizes)
        total = int(sizes.astype(np.uint64).sum())
        packed = np.zeros(total, np.uint8)
        entry = np.zeros(n + 1, np.uint64)
        self._check(self.lib.cabac_pack(self.h, _hp(out), _hp(out_offsets), _hp(sizes), n, _hp(packed), total,
                                        _hp(entry)))
        return packed, entry

    # ---- device-pointer calls (torch tensors), asynchronous ------------------
    def encode_residual_dev(self, tus, tu_offsets, params, n_sub, coeffs, out, out_offsets, out_sizes, num_bins=None,
                            tu_flags=None, ctx_in=None, ctx_out=None, status=None):
        self._check(self.lib.cabac_encode_residual_dev(self.h, _dp(tus), _dp(tu_offsets), _dp(params), n_sub,
                                                       _dp(coeffs), _dp(ctx_in), _dp(out), _dp(out_offsets),
                                                       _dp(out_sizes), _dp(num_bins), _dp(tu_flags), _dp(ctx_out),
                                                       _dp(status)))

    def estimate_residual_bits_dev(self, tus, tu_offsets, params, n_sub, coeffs, est_bits, est_per_tu=None,
                                   num_bins=None, tu_flags=None, ctx_in=None, ctx_out=None, status=None):
        self._check(self.lib.cabac_estimate_residual_bits_dev(self.h, _dp(tus), _dp(tu_offsets), _dp(params), n_sub,
                                                              _dp(coeffs), _dp(ctx_in), _dp(est_bits),
                                                              _dp(est_per_tu), _dp(num_bins), _dp(tu_flags),
                                                              _dp(ctx_out), _dp(status)))

    def residual_ctx_bins_dev(self, tus, tu_offsets, params, n_sub, coeffs, ctx_bins, num_bins=None, ctx_in=None,
                              ctx_out=None, status=None):
        self._check(self.lib.cabac_residual_ctx_bins_dev(self.h, _dp(tus), _dp(tu_offsets), _dp(params), n_sub,
                                                         _dp(coeffs), _dp(ctx_in), _dp(ctx_bins), _dp(num_bins),
                                                         _dp(ctx_out), _dp(status)))

    def ctx_init_id_dev(self, params, ctx_ref, ctx_bins, n, choice, costs=None):
        self._check(self.lib.cabac_ctx_init_id_dev(self.h, _dp(params), _dp(ctx_ref), _dp(ctx_bins), n, _dp(choice),
                                                   _dp(costs)))

    def encode_residual_wpp_dev(self, tus, tu_offsets, params, wpp, n_sub, coeffs, out, out_offsets, out_sizes,
                                num_bins=None, tu_flags=None, ctx_in=None, ctx_out=None, status=None):
        self._check(self.lib.cabac_encode_residual_wpp_dev(self.h, _dp(tus), _dp(tu_offsets), _dp(params), _dp(wpp),
                                                           n_sub, _dp(coeffs), _dp(ctx_in), _dp(out), _dp(out_offsets),
                                                           _dp(out_sizes), _dp(num_bins), _dp(tu_flags), _dp(ctx_out),
                                                           _dp(status)))

    def decode_residual_wpp_dev(self, data, in_offsets, tus, tu_offsets, params, wpp, n_sub, coeffs, num_bins=None,
                                bits_read=None, tu_flags=None, ctx_in=None, ctx_out=None, status=None):
        self._check(self.lib.cabac_decode_residual_wpp_dev(self.h, _dp(data), _dp(in_offsets), _dp(tus),
                                                           _dp(tu_offsets), _dp(params), _dp(wpp), n_sub, _dp(ctx_in),
                                                           _dp(coeffs), _dp(num_bins), _dp(bits_read), _dp(tu_flags),
                                                           _dp(ctx_out), _dp(status)))

    def decode_residual_dev(self, data, in_offsets, tus, tu_offsets, params, n_sub, coeffs, num_bins=None,
                            bits_read=None, tu_flags=None, ctx_in=None, ctx_out=None, status=None):
        self._check(self.lib.cabac_decode_residual_dev(self.h, _dp(data), _dp(in_offsets), _dp(tus), _dp(tu_offsets),
                                                       _dp(params), n_sub, _dp(ctx_in), _dp(coeffs), _dp(num_bins),
                                                       _dp(bits_read), _dp(tu_flags), _dp(ctx_out), _dp(status)))

    def decode_residual_static_dev(self, data, in_offsets, tus, tu_offsets, params, n_sub, coeffs, num_bins=None,
                                   bits_read=None, tu_flags=None, ctx_in=None, ctx_out=None, status=None):
        """cabac_decode_residual_static_dev: the same call without any allocation (context rows in a static device
        array, at most types.STATIC_MAX_SUBSTREAMS substreams; calls that overlap on the GPU race)."""
        self._check(self.lib.cabac_decode_residual_static_dev(self.h, _dp(data), _dp(in_offsets), _dp(tus),
                                                              _dp(tu_offsets), _dp(params), n_sub, _dp(ctx_in),
                                                              _dp(coeffs), _dp(num_bins), _dp(bits_read), _dp(tu_flags),
                                                              _dp(ctx_out), _dp(status)))

    def encode_bins_dev(self, cmds, cmd_offsets, params, n_sub, out, out_offsets, out_sizes, num_bins=None,
                        ctx_in=None, ctx_out=None, status=None):
        self._check(self.lib.cabac_encode_bins_dev(self.h, _dp(cmds), _dp(cmd_offsets), _dp(params), n_sub,
                                                   _dp(ctx_in), _dp(out), _dp(out_offsets), _dp(out_sizes),
                                                   _dp(num_bins), _dp(ctx_out), _dp(status)))

    def decode_bins_dev(self, data, in_offsets, cmds, cmd_offsets, params, n_sub, values, bits_read=None,
                        ctx_in=None, ctx_out=None, status=None):
        self._check(self.lib.cabac_decode_bins_dev(self.h, _dp(data), _dp(in_offsets), _dp(cmds), _dp(cmd_offsets),
                                                   _dp(params), n_sub, _dp(ctx_in), _dp(values), _dp(bits_read),
                                                   _dp(ctx_out), _dp(status)))

    def pack_dev(self, out, out_offsets, out_sizes, n_sub, packed, entry_points):
        self._check(self.lib.cabac_pack_dev(self.h, _dp(out), _dp(out_offsets), _dp(out_sizes), n_sub, _dp(packed),
                                            _dp(entry_points)))

    def ctx_init_dev(self, params, n, out):
        self._check(self.lib.cabac_ctx_init_dev(self.h, _dp(params), n, _dp(out)))
