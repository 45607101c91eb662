"""Wavefront (WPP) scheduling with context hand-off (BASELINE.json configs[3]).

With entropy_coding_sync_enabled every CTU row is its own substream (own arithmetic coder, own
entry point), but row r starts from the contexts row r-1 held after ITS FIRST CTU
(VTM EncSlice.cpp:2057-2072, 2085-2089; DecSlice.cpp:194-210, 241-245).  The dependency is on
contexts only, one CTU deep, so the schedule is:

  phase A  for r = 0 .. R-2: code the first CTU of row r of EVERY frame (one launch, n_frames
           substreams, not terminated) from the contexts step r-1 produced; keep ctx_out
  phase B  code all rows of all frames in one launch, row r > 0 starting from phase A's
           contexts (CABAC_SUB_CTX_IN)

Phase A is R-1 short dependent launches over n_frames substreams, phase B the full-width batch.
The first CTU of a row is coded twice (1 / ctus_per_row of the work); everything is exact.
Device-resident: inputs are uploaded once, all launches use the `_dev` entry points.
"""
from __future__ import annotations

import numpy as np

from .types import CTX_DT, NUM_CTX, PARAMS_DT, SUB_ALIGN, SUB_CTX_IN, SUB_TERMINATE, TU_DT


def _dev_bytes(a: np.ndarray, dev):
    import torch
    return torch.from_numpy(np.frombuffer(np.ascontiguousarray(a).tobytes(), np.uint8).copy()).to(dev)


def _first_ctu_batch(tus, tu_offsets, first_ctu_tus, rows, dev):
    """TU list / offsets holding only the first CTU of the given substreams."""
    off = np.asarray(tu_offsets, np.int64)
    cnt = np.asarray(first_ctu_tus, np.int64)[rows]
    idx = np.concatenate([np.arange(off[s], off[s] + c) for s, c in zip(rows, cnt)]) if len(rows) else np.zeros(0, np.int64)
    sub_off = np.concatenate([[0], np.cumsum(cnt)]).astype(np.uint64)
    return _dev_bytes(tus[idx], dev), _dev_bytes(sub_off, dev)


class WppCoder:
    """Encode / decode batches of WPP frames: substream f * rows + r is CTU row r of frame f."""

    def __init__(self, engine, rows_per_frame: int):
        import torch
        self.eng, self.R, self.torch = engine, int(rows_per_frame), torch
        self.dev = torch.device("cuda", torch.cuda.current_device())
        # a real stream shared by torch's copies and the engine's launches (handle 0, the legacy
        # default stream, would make cabac_set_stream fall back to the engine's own stream)
        self.stream = torch.cuda.Stream(device=self.dev)

    # ------------------------------------------------------------------ shared phase A driver
    def _sync_contexts(self, n_sub, params, launch_first_ctu):
        """Runs phase A; returns (ctx_in[n_sub, NUM_CTX] device tensor, params with CTX_IN set)."""
        torch, R = self.torch, self.R
        n_frames = n_sub // R
        assert n_frames * R == n_sub
        ctx_all = torch.zeros(n_sub * NUM_CTX * CTX_DT.itemsize, dtype=torch.uint8, device=self.dev)
        ctx_rows = ctx_all.view(n_frames, R, NUM_CTX * CTX_DT.itemsize)
        prev = None
        for r in range(R - 1):
            rows = np.arange(n_frames, dtype=np.int64) * R + r
            p = params[rows].copy()
            p["flags"] &= np.uint8(0xFF ^ (SUB_TERMINATE | SUB_ALIGN | SUB_CTX_IN))
            if r > 0:
                p["flags"] |= np.uint8(SUB_CTX_IN)
            ctx_out = torch.empty(n_frames * NUM_CTX * CTX_DT.itemsize, dtype=torch.uint8, device=self.dev)
            launch_first_ctu(rows, _dev_bytes(p, self.dev), prev, ctx_out)
            ctx_rows[:, r + 1, :] = ctx_out.view(n_frames, -1)
            prev = ctx_out
        p_all = params.copy()
        if R > 1:
            sel = (np.arange(n_sub) % R) != 0
            p_all["flags"][sel] |= np.uint8(SUB_CTX_IN)
        return ctx_all, p_all

    # ------------------------------------------------------------------ encode
    def encode(self, tus, tu_offsets, params, coeffs, out_offsets, first_ctu_tus):
        """Returns dict(out, sizes, num_bins, launches): per-row bitstreams in the capacity slabs."""
        torch, eng = self.torch, self.eng
        tus, params = np.asarray(tus, TU_DT), np.asarray(params, PARAMS_DT)
        n_sub = len(params)
        eng.set_stream(self.stream.cuda_stream)
        try:
          with torch.cuda.stream(self.stream):
            launches0 = eng.launch_count
            d_tus, d_tuoff = _dev_bytes(tus, self.dev), _dev_bytes(np.asarray(tu_offsets, np.uint64), self.dev)
            d_co = torch.from_numpy(np.ascontiguousarray(coeffs, np.int32)).to(self.dev)
            d_cap = _dev_bytes(np.asarray(out_offsets, np.uint64), self.dev)
            d_out = torch.empty(int(out_offsets[-1]), dtype=torch.uint8, device=self.dev)
            d_sizes = torch.zeros(n_sub, dtype=torch.int32, device=self.dev)
            d_bins = torch.zeros(n_sub, dtype=torch.int32, device=self.dev)
            d_status = torch.zeros(n_sub, dtype=torch.int32, device=self.dev)
            cap = np.asarray(out_offsets, np.uint64)

            def first_ctu(rows, d_p, ctx_in, ctx_out):
                t, toff = _first_ctu_batch(tus, tu_offsets, first_ctu_tus, rows, self.dev)
                n = len(rows)
                # the bytes of this trial run are discarded: only ctx_out matters
                scratch_off = np.concatenate([[0], np.cumsum((cap[rows + 1] - cap[rows]).astype(np.int64))]).astype(np.uint64)
                d_scr = torch.empty(int(scratch_off[-1]), dtype=torch.uint8, device=self.dev)
                d_ssz = torch.zeros(n, dtype=torch.int32, device=self.dev)
                eng.encode_residual_dev(t, toff, d_p, n, d_co, d_scr, _dev_bytes(scratch_off, self.dev), d_ssz,
                                        ctx_in=ctx_in, ctx_out=ctx_out)

            ctx_in, p_all = self._sync_contexts(n_sub, params, first_ctu)
            eng.encode_residual_dev(d_tus, d_tuoff, _dev_bytes(p_all, self.dev), n_sub, d_co, d_out, d_cap, d_sizes,
                                    num_bins=d_bins, ctx_in=ctx_in, status=d_status)
            self.stream.synchronize()
            assert int(d_status.abs().max()) == 0, "substream status set"
            return dict(out=d_out.cpu().numpy(), sizes=d_sizes.cpu().numpy().astype(np.uint32),
                        num_bins=d_bins.cpu().numpy().astype(np.uint32), launches=eng.launch_count - launches0)
        finally:
            eng.set_stream(None)

    # ------------------------------------------------------------------ decode
    def decode(self, data, in_offsets, tus, tu_offsets, params, n_coeffs, first_ctu_tus):
        """Returns dict(coeffs, num_bins, launches)."""
        torch, eng = self.torch, self.eng
        tus, params = np.asarray(tus, TU_DT), np.asarray(params, PARAMS_DT)
        n_sub = len(params)
        eng.set_stream(self.stream.cuda_stream)
        try:
          with torch.cuda.stream(self.stream):
            launches0 = eng.launch_count
            d_tus, d_tuoff = _dev_bytes(tus, self.dev), _dev_bytes(np.asarray(tu_offsets, np.uint64), self.dev)
            d_in = torch.from_numpy(np.ascontiguousarray(data, np.uint8)).to(self.dev)
            ioff = np.asarray(in_offsets, np.uint64)
            d_ioff = _dev_bytes(ioff, self.dev)
            d_co = torch.zeros(n_coeffs, dtype=torch.int32, device=self.dev)
            d_bins = torch.zeros(n_sub, dtype=torch.int32, device=self.dev)
            d_status = torch.zeros(n_sub, dtype=torch.int32, device=self.dev)

            def first_ctu(rows, d_p, ctx_in, ctx_out):
                t, toff = _first_ctu_batch(tus, tu_offsets, first_ctu_tus, rows, self.dev)
                # the rows' byte ranges are not contiguous: gather them into one small payload
                lens = (ioff[rows + 1] - ioff[rows]).astype(np.int64)
                sub_off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
                idx = torch.from_numpy(np.concatenate([np.arange(int(ioff[s]), int(ioff[s + 1])) for s in rows])).to(self.dev)
                eng.decode_residual_dev(d_in[idx].contiguous(), _dev_bytes(sub_off, self.dev), t, toff, d_p, len(rows),
                                        d_co, ctx_in=ctx_in, ctx_out=ctx_out)

            ctx_in, p_all = self._sync_contexts(n_sub, params, first_ctu)
            eng.decode_residual_dev(d_in, d_ioff, d_tus, d_tuoff, _dev_bytes(p_all, self.dev), n_sub, d_co,
                                    num_bins=d_bins, ctx_in=ctx_in, status=d_status)
            self.stream.synchronize()
            assert int(d_status.abs().max()) == 0, "substream status set"
            return dict(coeffs=d_co.cpu().numpy(), num_bins=d_bins.cpu().numpy().astype(np.uint32),
                        launches=eng.launch_count - launches0)
        finally:
            eng.set_stream(None)
