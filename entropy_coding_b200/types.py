"""numpy mirrors of include/cabac_b200.h (structs, flags, command words)."""
from __future__ import annotations

import numpy as np

NUM_CTX = 379  # Ctx::NumberOfContexts, reference src/common/contexts.cpp:773-774

PARAMS_DT = np.dtype([("qp", "i1"), ("init_id", "u1"), ("flags", "u1"), ("reserved", "u1")])
TU_DT = np.dtype([("coeff_offset", "<u8"), ("log2_w", "u1"), ("log2_h", "u1"), ("comp_id", "u1"),
                  ("flags", "u1"), ("reserved", "<u4")])
CTX_DT = np.dtype([("state0", "<u2"), ("state1", "<u2")])
WPP_DT = np.dtype([("dep", "<u4"), ("sync_tus", "<u4"), ("sync_cmds", "<u4"), ("reserved", "<u4")])  # cabac_wpp_sync
WPP_NONE = 0xFFFFFFFF
assert PARAMS_DT.itemsize == 4 and TU_DT.itemsize == 16 and CTX_DT.itemsize == 4 and WPP_DT.itemsize == 16

# status codes
OK, ERR_CUDA, ERR_ARG, ERR_SUBSTREAM = 0, -1, -2, -3
# per-substream status bits
SS_OUT_OVERFLOW, SS_EMPTY_TU, SS_BAD_CMD, SS_IN_OVERRUN, SS_COEFF_RANGE, SS_NO_TERMINATE = 1, 2, 4, 8, 16, 32
SS_WPP_TIMEOUT = 64
SS_BAD_DESC = 128
STATIC_MAX_SUBSTREAMS = 16384  # CABAC_STATIC_MAX_SUBSTREAMS (cabac_decode_residual_static_dev)
# substream flags
SUB_DEP_QUANT, SUB_SIGN_HIDING, SUB_TSRC_DISABLED = 1, 2, 4
SUB_TERMINATE, SUB_ALIGN, SUB_CTX_IN = 8, 16, 32
# TU descriptor flags / result flags
TU_TS, TU_TS_ALLOWED, TU_BDPCM, TU_SBT_ZEROOUT = 1, 2, 4, 8
TUR_TS, TUR_VIOLATES_LFNST, TUR_LFNST_LAST_POS, TUR_MTS_LAST_POS, TUR_VIOLATES_MTS = 1, 2, 4, 8, 16
# command kinds
CMD_CTX, CMD_EP, CMD_REM, CMD_TRM, CMD_ALIGN, CMD_TU = 0, 1, 2, 3, 4, 5


def mk_ctx(ctx: int, b: int) -> int:
    return (CMD_CTX << 29) | ((b & 1) << 9) | (ctx & 511)


def mk_ep(bins: int, n: int) -> int:
    return (CMD_EP << 29) | ((n & 31) << 24) | (bins & 0xFFFFFF)


def mk_rem(value: int, rice: int) -> int:
    return (CMD_REM << 29) | ((rice & 31) << 24) | (value & 0xFFFFFF)


def mk_trm(b: int) -> int:
    return (CMD_TRM << 29) | (b & 1)


def mk_align() -> int:
    return CMD_ALIGN << 29


def mk_tu() -> int:
    return CMD_TU << 29
