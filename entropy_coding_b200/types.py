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

# cabac_decode_slices (CTU-level syntax on the device): cabac_slice_params, cabac_cu_desc, cabac_ctu_desc
SLICE_FIELDS = ("pic_w", "pic_h", "x0", "y0", "x1", "y1", "log2_ctu", "log2_min_qt_luma", "log2_min_qt_chroma",
                "log2_max_bt_luma", "log2_max_bt_chroma", "log2_max_tt_luma", "log2_max_tt_chroma",
                "max_mtt_depth_luma", "max_mtt_depth_chroma", "log2_min_bt_luma", "log2_min_bt_chroma",
                "log2_min_tt_luma", "log2_min_tt_chroma", "log2_max_tb", "log2_max_ts", "bit_depth",
                "alf_num_aps_luma", "alf_chroma_num_alts", "ccalf_filter_count_cb", "ccalf_filter_count_cr",
                "cu_qp_delta_subdiv", "cu_chroma_qp_offset_subdiv", "chroma_qp_offset_list_len", "chroma_format",
                "tools", "reserved")
SLICE_DT = np.dtype([(f, "<u4") for f in SLICE_FIELDS])
CU_DT = np.dtype([("x", "<u2"), ("y", "<u2"), ("log2_w", "u1"), ("log2_h", "u1"), ("ch_type", "u1"), ("qt_depth", "u1"),
                  ("mt_depth", "u1"), ("flags", "u1"), ("bdpcm", "u1"), ("mrl_idx", "u1"), ("isp", "u1"),
                  ("lfnst_idx", "u1"), ("mts_idx", "u1"), ("pred_syntax", "u1"), ("first_tu", "<u4"), ("n_tu", "<u2"),
                  ("qp_delta", "i1"), ("chroma_qp_adj", "u1"), ("intra_dir", "u1"), ("reserved", "u1"), ("cbf_mask", "<u2"), ("split_series", "<u4")])
CTU_DT = np.dtype([("sao_mode", "u1", 3), ("sao_type", "u1", 3), ("sao_band", "u1", 3), ("sao_offset", "i1", (3, 4)),
                   ("alf_flag", "u1", 3), ("alf_filter_idx", "u1"), ("alf_alt", "u1", 2), ("ccalf_idc", "u1", 2),
                   ("reserved", "u1", 3)])
assert SLICE_DT.itemsize == 128 and CU_DT.itemsize == 32 and CTU_DT.itemsize == 32
SL_DUAL_TREE, SL_SAO_LUMA, SL_SAO_CHROMA, SL_ALF_Y, SL_ALF_CB, SL_ALF_CR, SL_CCALF_CB, SL_CCALF_CR = (1 << i for i in range(8))
SL_MTS_INTRA, SL_LFNST, SL_ISP, SL_MIP, SL_MRL, SL_TS, SL_BDPCM, SL_JCCR = (1 << i for i in range(8, 16))
SL_CCLM, SL_DQP, SL_CHROMA_QP_ADJ, SL_WPP, SL_IBC, SL_PLT, SL_ACT, SL_INTRA, SL_MTS = (1 << i for i in range(16, 25))
CU_MIP, CU_MIP_TRANSPOSED, CU_MPM, CU_CCLM, CU_JCCR, CU_DQP, CU_FROM_DIR = 1, 2, 4, 8, 16, 32, 64
SS_UNSUPPORTED, SS_CAPACITY = 256, 512

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
