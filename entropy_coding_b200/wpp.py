"""Wavefront (WPP) coding with context hand-off (BASELINE.json configs[3]).

With entropy_coding_sync_enabled every CTU row is its own substream (own arithmetic coder, own
entry point), but row r starts from the contexts row r-1 held after ITS FIRST CTU
(VTM EncSlice.cpp:2057-2072, 2085-2089; DecSlice.cpp:194-210, 241-245).

The hand-over runs on the device inside ONE launch (cabac_encode_residual_wpp /
cabac_decode_residual_wpp, include/cabac_b200.h): the warp coding row r publishes its contexts
after the row's first CTU, the warp coding row r+1 waits for them.  This module only builds the
`cabac_wpp_sync` records for the common layout "substream f * rows + r is CTU row r of frame f"
and forwards to the C ABI; nothing here computes.
"""
from __future__ import annotations

import numpy as np

from .types import WPP_DT, WPP_NONE


def sync_records(n_sub: int, rows_per_frame: int, first_ctu_tus, first_ctu_cmds=None) -> np.ndarray:
    """cabac_wpp_sync per substream: row r > 0 of a frame depends on row r - 1; every row publishes
    after its first CTU (first_ctu_tus TUs / first_ctu_cmds recorded command words)."""
    assert n_sub % rows_per_frame == 0
    w = np.zeros(n_sub, WPP_DT)
    i = np.arange(n_sub, dtype=np.int64)
    w["dep"] = np.where(i % rows_per_frame == 0, WPP_NONE, i - 1).astype(np.uint32)
    w["sync_tus"] = np.asarray(first_ctu_tus, np.uint32)
    if first_ctu_cmds is not None:
        w["sync_cmds"] = np.asarray(first_ctu_cmds, np.uint32)
    return w


class WppCoder:
    """Encode / decode batches of WPP frames through the host-pointer WPP entry points."""

    def __init__(self, engine, rows_per_frame: int):
        self.eng, self.R = engine, int(rows_per_frame)

    def encode(self, tus, tu_offsets, params, coeffs, out_offsets, first_ctu_tus):
        """dict(packed, entry_points, sizes, num_bins, launches): the rows' bitstreams concatenated in order."""
        l0 = self.eng.launch_count
        w = sync_records(len(params), self.R, first_ctu_tus)
        r = self.eng.encode_residual_wpp(tus, tu_offsets, params, w, coeffs, out_offsets)
        r["launches"] = self.eng.launch_count - l0
        return r

    def decode(self, data, in_offsets, tus, tu_offsets, params, n_coeffs, first_ctu_tus):
        """dict(coeffs, num_bins, launches)."""
        l0 = self.eng.launch_count
        w = sync_records(len(params), self.R, first_ctu_tus)
        r = self.eng.decode_residual_wpp(data, in_offsets, tus, tu_offsets, params, w, n_coeffs)
        r["launches"] = self.eng.launch_count - l0
        return r
