// cabac_syntax.cuh - the CTU-level syntax of an all-intra, dual-tree VVC substream (SURVEY.md
// section 8, row f2): coding_tree_unit -> sao / alf -> coding_tree (split_cu_mode with the
// QTBT partitioner's legality rules) -> coding_unit -> intra modes -> transform_tree ->
// transform_unit (cbf, cu_qp_delta, joint_cb_cr, residual_coding) -> lfnst / mts.
//
// The parser is control flow only: every bin goes through the engine E, which owns the arithmetic
// decoder, the 379 contexts and the residual decoder.  In the product E is the warp-lockstep device
// engine of decode_slice_kernel (cabac_kernels.cu): all 32 lanes of a warp run this code on the
// same values (warp-uniform control flow) and take part in every bin and every coefficient group.
// The test tree compiles the same header for the host with a scalar engine (syntax_host.cpp), so the
// walk can be checked on the CPU against the call sequence the reference's writer recorded.
//
// Reference (src/entropy_codec/cabac_reader.cpp unless noted): coding_tree_unit :103-209, sao
// :284-419, coding_tree :429-655, split_cu_mode :677-751, coding_unit :764-825, bdpcm_mode :1063,
// xReadTruncBinCode :1162, extend_ref_line :1188, intra_luma_pred_modes :1231, intra_chroma_pred_mode
// :1353, cu_residual :1396, transform_tree :2297, cbf_comp :2353, transform_unit :2456, cu_qp_delta
// :2578, cu_chroma_qp_offset :2602, joint_cb_cr :2632, mts_idx :2754, isp_mode :2785,
// residual_lfnst_mode :2808, mip_flag / mip_pred_mode :3381-3408; QTBTPartitioner::canSplit /
// getImplicitSplit / splitCurrArea (src/common/unit_partitioner.cpp:303-470, 94-190);
// DeriveCtx::CtxSplit / CtxMipFlag / CtxQtCbf (src/common/context_modelling.cpp:108-230, 317-331);
// CodingUnit::checkCCLMAllowed (src/common/unit.cpp:289-372); CU::canUseISP / getISPSplitDim /
// canUseLfnstWithISP / isMTSAllowed / bdpcmAllowed, TU::isTSAllowed (src/common/unit_tools.cpp).
#pragma once
#include <stdint.h>

#include "cabac_b200.h"

#if defined(__CUDACC__)
#define SX_FN __device__ __forceinline__
#define SX_FN_NOINLINE __device__ __noinline__
#else
#define SX_FN inline
#define SX_FN_NOINLINE inline
#endif
// SX_OPAQUE(v): the value goes through an instruction the optimiser cannot fold (a warp shuffle from lane 0; the
// warp is converged and v is uniform).  Needed in one place: with CUDA 12.9 ptxas (-O3 and -O1 alike) turns the test
// "tool bit && (cbf_cb ? 2 : 0 | cbf_cr ? 1 : 0) != 0" in front of tu_joint_cbcr_residual_flag into
// PLOP3.LUT 0x8f over (tool, cbf_cb, cbf_cr) - skip when BOTH flags are set, read when neither is - although the PTX
// is right (nvdisasm -g of the cubin; the device left the reference's call sequence at the first chroma TU with
// both cbfs set, and a printf next to the test made the problem disappear).
#if defined(__CUDA_ARCH__)
#define SX_OPAQUE(v) ((v) = __shfl_sync(0xffffffffu, (v), 0))
#else
#define SX_OPAQUE(v) ((void)0)
#endif

namespace cabac_syntax {

// Context-set offsets (positional in src/common/contexts.cpp; tests/test_syntax.py compares them
// with the compiled reference through ref_driver.cpp:ref_syntax_ctx_offsets).
enum : int {
  SX_SPLIT_FLAG = 0,
  SX_SPLIT_QT_FLAG = 9,
  SX_SPLIT_HV_FLAG = 15,
  SX_SPLIT12_FLAG = 20,
  SX_MODE_CONS_FLAG = 24,
  SX_MIP_FLAG = 46,
  SX_INTRA_LUMA_MPM_FLAG = 40,
  SX_INTRA_LUMA_PLANAR_FLAG = 41,
  SX_CCLM_MODE_FLAG = 43,
  SX_CCLM_MODE_IDX = 44,
  SX_INTRA_CHROMA_PRED_MODE = 45,
  SX_MULTI_REF_LINE_IDX = 38,
  SX_ISP_MODE = 316,
  SX_BDPCM_MODE = 71,
  SX_QT_CBF_Y = 77,
  SX_QT_CBF_CB = 81,
  SX_QT_CBF_CR = 83,
  SX_DELTA_QP = 50,
  SX_CHROMA_QP_ADJ_FLAG = 325,
  SX_CHROMA_QP_ADJ_IDC = 326,
  SX_JOINT_CBCR_FLAG = 354,
  SX_MTS_IDX = 312,
  SX_LFNST_IDX = 296,
  SX_SAO_MERGE_FLAG = 294,
  SX_SAO_TYPE_IDX = 295,
  SX_CTB_ALF_FLAG = 332,
  SX_CTB_ALF_ALTERNATIVE = 341,
  SX_ALF_USE_TEMPORAL_FILT = 343,
  SX_CCALF_FILTER_CONTROL_FLAG = 344
};
constexpr int SX_NUM_OFFSETS = 29;
#if !defined(__CUDACC__)
inline void syntaxCtxOffsets(uint32_t *o) {
  const int v[SX_NUM_OFFSETS] = {SX_SPLIT_FLAG, SX_SPLIT_QT_FLAG, SX_SPLIT_HV_FLAG, SX_SPLIT12_FLAG, SX_MODE_CONS_FLAG,
                                 SX_MIP_FLAG, SX_INTRA_LUMA_MPM_FLAG, SX_INTRA_LUMA_PLANAR_FLAG, SX_CCLM_MODE_FLAG,
                                 SX_CCLM_MODE_IDX, SX_INTRA_CHROMA_PRED_MODE, SX_MULTI_REF_LINE_IDX, SX_ISP_MODE,
                                 SX_BDPCM_MODE, SX_QT_CBF_Y, SX_QT_CBF_CB, SX_QT_CBF_CR, SX_DELTA_QP,
                                 SX_CHROMA_QP_ADJ_FLAG, SX_CHROMA_QP_ADJ_IDC, SX_JOINT_CBCR_FLAG, SX_MTS_IDX,
                                 SX_LFNST_IDX, SX_SAO_MERGE_FLAG, SX_SAO_TYPE_IDX, SX_CTB_ALF_FLAG,
                                 SX_CTB_ALF_ALTERNATIVE, SX_ALF_USE_TEMPORAL_FILT, SX_CCALF_FILTER_CONTROL_FLAG};
  for (int i = 0; i < SX_NUM_OFFSETS; i++) o[i] = uint32_t(v[i]);
}
#endif

// PartSplit, src/common/type_def.hpp (only the CU-level values are used)
enum : int { SPL_CTU_LEVEL = 0, SPL_NONE = 2000, SPL_QUAD = 1, SPL_HORZ = 2, SPL_VERT = 3, SPL_TRIH = 4, SPL_TRIV = 5 };

constexpr int SX_MAX_DEPTH = 12;      // 128 -> 4 is five QT levels; MTT depth <= 4 plus implicit boundary splits
constexpr int SX_MAX_LINE = 2048;     // above-line entries: pictures up to 8192 luma samples wide

// What a decoded CU leaves behind for its right / lower neighbours (CtxSplit, CtxMipFlag): one 16-bit word per
// 4 luma samples of the picture width (above line) and of the CTU height (left column), per tree.
SX_FN uint16_t nbPack(int log2w, int log2h, int qtDepth, int mip) {
  return uint16_t(0x8000 | log2w | (log2h << 3) | (qtDepth << 6) | (mip << 9));
}
SX_FN int nbLog2W(uint16_t v) { return v & 7; }
SX_FN int nbLog2H(uint16_t v) { return (v >> 3) & 7; }
SX_FN int nbQt(uint16_t v) { return (v >> 6) & 7; }
SX_FN int nbMip(uint16_t v) { return (v >> 9) & 1; }

struct Node {  // one level of the coding tree: an area and the partitioner's depth counters for it
  int16_t x, y;
  uint8_t log2w, log2h;
  uint8_t lastSplit;   // the split that produced this node (PartLevel::split of its parent level)
  uint8_t partIdx;     // index among its siblings
  uint8_t qt, mt, bt, implBt;
  uint8_t subdiv;
  uint8_t qg, qgChroma;  // currQgEnable / currQgChromaEnable
  uint8_t split;       // this node's own split once parsed
  int8_t next;         // next child to visit, -1 = not parsed yet
  uint8_t implicit;    // this node's split equals the implicit one (boundary)
};

template <class E>
struct SliceParser {
  // The parser's whole state lives in this object - no arrays or records on the stack - so that the device can
  // keep ONE copy per warp in shared memory (cabac_slice.cuh) instead of 32 identical per-lane copies in local
  // memory: all lanes hold the same values, and a noinline bin call then has little live state to save around it.
  E &e;
  const cabac_slice_params sp;
  Node st[SX_MAX_DEPTH];  // coding_tree: the explicit recursion stack
  Node roots[4];          // the trees of the current CTU
  cabac_cu_desc cuRec;    // the CU being parsed
  cabac_ctu_desc ctuRec;  // the CTU being parsed
  uint32_t nCu = 0, nTu = 0, nCtu = 0;
  uint64_t nCoeff = 0;
  uint32_t status = 0;
  // capacities of this substream's regions
  uint32_t capCu, capTu, capCtu;
  uint64_t capCoeff, coeffBase;
  // per-CU accumulators of CUCtx (cabac_reader.cpp:1417-1421)
  uint32_t cuRes = 0;
  bool isDqpCoded = true, isChromaQpAdjCoded = true;
  // the luma 64x64 node the chroma tree is collocated with (checkCCLMAllowed)
  int lumaRootSplit = SPL_NONE, lumaRootIsp = 0;
  // left / above CTU in-loop filter flags
  uint8_t leftCtuFlags = 0;

  SX_FN SliceParser(E &eng, const cabac_slice_params &p, uint32_t cCu, uint32_t cTu, uint32_t cCtu, uint64_t cCoeff,
                    uint64_t cBase)
      : e(eng), sp(p), capCu(cCu), capTu(cTu), capCtu(cCtu), capCoeff(cCoeff), coeffBase(cBase) {}

  SX_FN bool tool(uint32_t t) const { return (sp.tools & t) != 0; }

  // ---- small binarisations ----
  // Every decision goes through e.bin(ctx, want) / e.ep(n, want): a decoding engine ignores `want` and returns
  // what the stream holds, an emitting engine (E::kEncode: the writer direction, cabac_encode_slices) codes `want`
  // and returns it.  The walk is the same in both directions; in the writer direction `want` comes from the
  // records the caller supplies (the very records the decoder direction produces).
  SX_FN uint32_t truncBin(uint32_t maxSymbol, uint32_t want) {  // xReadTruncBinCode :1162-1186 / xWriteTruncBinCode
    int thresh = 0;                                             // cabac_writer.cpp:854-882 (maxSymbol <= 256)
    while ((2u << thresh) <= maxSymbol) thresh++;  // g_tbMax[maxSymbol] = floor(log2)
    const uint32_t val = 1u << thresh, b = maxSymbol - val;
    const bool wLong = want >= val - b;
    const uint32_t wCode = wLong ? want + (val - b) : want;  // long codes: thresh + 1 bits of (symbol + val - b)
    uint32_t symbol = thresh ? e.ep(thresh, wLong ? wCode >> 1 : wCode) : 0;
    if (symbol >= val - b) symbol = ((symbol << 1) | e.ep(1, wCode & 1u)) - (val - b);
    return symbol;
  }
  SX_FN uint32_t unaryMaxEp(uint32_t maxSymbol, uint32_t want) {  // unary_max_eqprob, :3359-3366
    for (uint32_t k = 0; k < maxSymbol; k++)
      if (!e.ep(1, k < want ? 1u : 0u)) return k;
    return maxSymbol;
  }
  SX_FN uint32_t unaryMaxSymbol(int ctx0, int ctxN, uint32_t maxSymbol, uint32_t want) {  // :3349-3357
    uint32_t ones = 0;
    while (ones < maxSymbol && e.bin(ones == 0 ? ctx0 : ctxN, ones < want ? 1u : 0u)) ones++;
    return ones;
  }
  SX_FN uint32_t expGolombEp(uint32_t count, uint32_t want) {  // exp_golomb_eqprob, :3368-3379 / cabac_writer.cpp:3096-3118
    uint32_t symbol = 0, bit = 1, rem = want;
    while (bit) {
      bit = e.ep(1, rem >= (1u << count) ? 1u : 0u);
      if (bit) rem -= rem >= (1u << count) ? (1u << count) : 0u;
      symbol += bit << count++;
      if (count > 31) {
        status |= CABAC_SS_COEFF_RANGE;
        break;
      }
    }
    if (--count) symbol += e.ep(count, rem & ((1u << count) - 1u));
    return symbol;
  }

  // ---- neighbours ----
  SX_FN bool leftAvail(int x) const { return x > int(sp.x0); }
  SX_FN bool aboveAvail(int y) const { return y > int(sp.y0); }
  SX_FN uint16_t nbLeft(int ch, int x, int y) {
    return leftAvail(x) ? e.leftCol(ch)[(y & ((1 << sp.log2_ctu) - 1)) >> 2] : uint16_t(0);
  }
  SX_FN uint16_t nbAbove(int ch, int x, int y) {
    return aboveAvail(y) ? e.aboveLine(ch)[(x - int(sp.x0)) >> 2] : uint16_t(0);
  }
  // dirNb: the luma direction the CU shows its neighbours and the chroma tree (PU::getIntraDirLuma: PLANAR for MIP)
  SX_FN void nbStore(int ch, const Node &n, int mip, int dirNb) {
    const uint16_t v = nbPack(n.log2w, n.log2h, n.qt, mip);
    const int col = (n.x - int(sp.x0)) >> 2, row = (n.y & ((1 << sp.log2_ctu) - 1)) >> 2;
    const int w4 = (1 << n.log2w) >> 2, h4 = (1 << n.log2h) >> 2;
    e.fill(e.aboveLine(ch) + col, w4, v);
    e.fill(e.leftCol(ch) + row, h4, v);
    if (ch == 0) {
      e.fill8(e.modeAbove() + col, w4, uint8_t(dirNb));
      e.fill8(e.modeLeft() + row, h4, uint8_t(dirNb));
      e.fill(e.edgeLeft() + row, h4, uint16_t(n.x + (1 << n.log2w)));
      e.fillMap((n.x & 63) >> 2, (n.y & 63) >> 2, w4, h4, uint8_t(dirNb));
    }
  }

  // PU::getIntraMPMs, unit_tools.cpp:485-577: left = the PU at the bottom-left corner (x - 1, y + h - 1) if already
  // decoded, above = the PU at the top-right corner (x + w - 1, y - 1) if inside the same CTU
  SX_FN void intraMpms(const Node &n, int *mpm) {
    const int w = 1 << n.log2w, h = 1 << n.log2h, ctuMask = (1 << sp.log2_ctu) - 1;
    int left = 0, above = 0;  // PLANAR_IDX
    const int rowL = ((n.y + h - 1) & ctuMask) >> 2;
    if (leftAvail(n.x) && e.leftCol(0)[rowL] && int(e.edgeLeft()[rowL]) == int(n.x)) left = e.modeLeft()[rowL];
    if (aboveAvail(n.y) && (n.y & ctuMask) != 0) above = e.modeAbove()[(n.x + w - 1 - int(sp.x0)) >> 2];
    const int offset = 61, mod = 64;  // NUM_LUMA_MODE - 6, offset + 3
    mpm[0] = 0;
    mpm[1] = 1;
    mpm[2] = 50;
    mpm[3] = 18;
    mpm[4] = 46;
    mpm[5] = 54;
    if (left == above) {
      if (left > 1) {
        mpm[0] = 0;
        mpm[1] = left;
        mpm[2] = ((left + offset) % mod) + 2;
        mpm[3] = ((left - 1) % mod) + 2;
        mpm[4] = ((left + offset - 1) % mod) + 2;
        mpm[5] = (left % mod) + 2;
      }
    } else if (left > 1 && above > 1) {
      mpm[0] = 0;
      mpm[1] = left;
      mpm[2] = above;
      const int mx = left > above ? left : above, mn = left > above ? above : left;
      if (mx - mn == 1) {
        mpm[3] = ((mn + offset) % mod) + 2;
        mpm[4] = ((mx - 1) % mod) + 2;
        mpm[5] = ((mn + offset - 1) % mod) + 2;
      } else if (mx - mn >= 62) {
        mpm[3] = ((mn - 1) % mod) + 2;
        mpm[4] = ((mx + offset) % mod) + 2;
        mpm[5] = (mn % mod) + 2;
      } else if (mx - mn == 2) {
        mpm[3] = ((mn - 1) % mod) + 2;
        mpm[4] = ((mn + offset) % mod) + 2;
        mpm[5] = ((mx - 1) % mod) + 2;
      } else {
        mpm[3] = ((mn + offset) % mod) + 2;
        mpm[4] = ((mn - 1) % mod) + 2;
        mpm[5] = ((mx + offset) % mod) + 2;
      }
    } else if (left + above >= 2) {
      const int mx = left < above ? above : left;
      mpm[0] = 0;
      mpm[1] = mx;
      mpm[2] = ((mx + offset) % mod) + 2;
      mpm[3] = ((mx - 1) % mod) + 2;
      mpm[4] = ((mx + offset - 1) % mod) + 2;
      mpm[5] = (mx % mod) + 2;
    }
  }

  // ---- QTBTPartitioner::getImplicitSplit, unit_partitioner.cpp:425-470 ----
  SX_FN int implicitSplit(const Node &n, int ch) const {
    const int w = 1 << n.log2w, h = 1 << n.log2h;
    const bool blIn = n.y + h - 1 < int(sp.pic_h), trIn = n.x + w - 1 < int(sp.pic_w);  // x, y themselves are inside
    const int maxBt = 1 << sp.log2_max_bt[ch], minQt = 1 << sp.log2_min_qt[ch];
    const bool btAllowed = w <= maxBt && h <= maxBt && n.mt < sp.max_mtt_depth[ch] + n.implBt;
    const bool qtAllowed = w > minQt && h > minQt && n.bt == 0;
    int split = SPL_NONE;
    if (!blIn && !trIn && qtAllowed)
      split = SPL_QUAD;
    else if (!blIn && btAllowed && w <= 64)
      split = SPL_HORZ;
    else if (!trIn && btAllowed && h <= 64)
      split = SPL_VERT;
    else if (!blIn || !trIn)
      split = SPL_QUAD;
    if (tool(CABAC_SL_DUAL_TREE) && (w > 64 || h > 64)) split = SPL_QUAD;
    return split;
  }

  // ---- QTBTPartitioner::canSplit, unit_partitioner.cpp:303-391; bit i of the result = canSpl[i]
  //      (0 no split, 1 QT, 2 BT hor, 3 BT ver, 4 TT hor, 5 TT ver) ----
  SX_FN uint32_t canSplit(const Node &n, int ch, int implicit) const {
    const int w = 1 << n.log2w, h = 1 << n.log2h;
    const int maxBtd = int(sp.max_mtt_depth[ch]) + n.implBt;
    const int maxBt = 1 << sp.log2_max_bt[ch], minBt = 1 << sp.log2_min_bt[ch];
    const int maxTt = 1 << sp.log2_max_tt[ch], minTt = 1 << sp.log2_min_tt[ch];
    const int minQt = 1 << sp.log2_min_qt[ch];
    const int cw = w >> 1, chh = h >> 1;  // chroma block of a chroma-tree node (4:2:0)
    bool canNo = true, canQt = true, canBh = true, canBv = true, canTh = true, canTv = true;
    bool canBtt = n.mt < maxBtd;
    if (n.lastSplit != SPL_CTU_LEVEL && n.lastSplit != SPL_QUAD) canQt = false;
    if (w <= minQt) canQt = false;
    if (ch && cw <= 4) canQt = false;  // MIN_DUALTREE_CHROMA_WIDTH
    if (implicit != SPL_NONE) {
      canNo = canTh = canTv = false;
      canBh = implicit == SPL_HORZ;
      canBv = implicit == SPL_VERT;
      if (ch && cw == 4) canBv = false;
      if (!canBh && !canBv && !canQt) canQt = true;
      return uint32_t(canNo) | (uint32_t(canQt) << 1) | (uint32_t(canBh) << 2) | (uint32_t(canBv) << 3);
    }
    if ((n.lastSplit == SPL_TRIH || n.lastSplit == SPL_TRIV) && n.partIdx == 1) {
      const int parl = n.lastSplit == SPL_TRIH ? SPL_HORZ : SPL_VERT;
      canBh = parl != SPL_HORZ;
      canBv = parl != SPL_VERT;
    }
    if (canBtt && (w <= minBt && h <= minBt) && (w <= minTt && h <= minTt)) canBtt = false;
    if (canBtt && (w > maxBt || h > maxBt) && (w > maxTt || h > maxTt)) canBtt = false;
    if (!canBtt) return uint32_t(canNo) | (uint32_t(canQt) << 1);
    if (w > maxBt || h > maxBt) canBh = canBv = false;
    if (h <= minBt) canBh = false;
    if (w > 64 && h <= 64) canBh = false;
    if (ch && cw * chh <= 16) canBh = false;  // MIN_DUALTREE_CHROMA_SIZE
    if (w <= minBt) canBv = false;
    if (w <= 64 && h > 64) canBv = false;
    if (ch && (cw * chh <= 16 || cw == 4)) canBv = false;
    if (h <= 2 * minTt || h > maxTt || w > maxTt) canTh = false;
    if (w > 64 || h > 64) canTh = false;
    if (ch && cw * chh <= 32) canTh = false;
    if (w <= 2 * minTt || w > maxTt || h > maxTt) canTv = false;
    if (w > 64 || h > 64) canTv = false;
    if (ch && (cw * chh <= 32 || cw == 8)) canTv = false;
    return uint32_t(canNo) | (uint32_t(canQt) << 1) | (uint32_t(canBh) << 2) | (uint32_t(canBv) << 3) |
           (uint32_t(canTh) << 4) | (uint32_t(canTv) << 5);
  }

  // ---- split_cu_mode (:677-751) with DeriveCtx::CtxSplit (context_modelling.cpp:108-204) ----
  SX_FN int splitCuMode(const Node &n, int ch, int implicit, int want) {
    const uint32_t can = canSplit(n, ch, implicit);
    const bool canNo = can & 1, canQt = can & 2, canBh = can & 4, canBv = can & 8, canTh = can & 16, canTv = can & 32;
    const uint16_t left = nbLeft(ch, n.x, n.y), above = nbAbove(ch, n.x, n.y);
    bool isSplit = canBh || canBv || canTh || canTv || canQt;
    if (canNo && isSplit) {
      int ctx = 0;
      if (left) ctx += nbLog2H(left) < n.log2h;
      if (above) ctx += nbLog2W(above) < n.log2w;
      int numSplit = (canQt ? 2 : 0) + (canBh ? 1 : 0) + (canBv ? 1 : 0) + (canTh ? 1 : 0) + (canTv ? 1 : 0);
      if (numSplit > 0) numSplit--;
      ctx += 3 * (numSplit >> 1);
      isSplit = e.bin(SX_SPLIT_FLAG + ctx, want != SPL_NONE) != 0;
    }
    if (!isSplit) return SPL_NONE;
    const bool canBtt = canBh || canBv || canTh || canTv;
    bool isQt = canQt;
    if (isQt && canBtt) {
      int ctx = (left && nbQt(left) > n.qt) ? 1 : 0;
      ctx += (above && nbQt(above) > n.qt) ? 1 : 0;
      ctx += n.qt < 2 ? 0 : 3;
      isQt = e.bin(SX_SPLIT_QT_FLAG + ctx, want == SPL_QUAD) != 0;
    }
    if (isQt) return SPL_QUAD;
    const bool canHor = canBh || canTh;
    bool isVer = canBv || canTv;
    if (isVer && canHor) {
      int ctx;
      const int numHor = (canBh ? 1 : 0) + (canTh ? 1 : 0), numVer = (canBv ? 1 : 0) + (canTv ? 1 : 0);
      if (numVer == numHor) {
        const int wAbove = above ? (1 << nbLog2W(above)) : 1, hLeft = left ? (1 << nbLog2H(left)) : 1;
        const int depAbove = (1 << n.log2w) / wAbove, depLeft = (1 << n.log2h) / hLeft;
        if (depAbove == depLeft || !left || !above)
          ctx = 0;
        else if (depAbove < depLeft)
          ctx = 1;
        else
          ctx = 2;
      } else {
        ctx = numVer < numHor ? 3 : 4;
      }
      isVer = e.bin(SX_SPLIT_HV_FLAG + ctx, want == SPL_VERT || want == SPL_TRIV) != 0;
    }
    const bool can14 = isVer ? canTv : canTh;
    bool is12 = isVer ? canBv : canBh;
    if (is12 && can14)
      is12 = e.bin(SX_SPLIT12_FLAG + (isVer ? (n.mt <= 1 ? 3 : 2) : (n.mt <= 1 ? 1 : 0)), want == SPL_VERT || want == SPL_HORZ) != 0;
    return isVer ? (is12 ? SPL_VERT : SPL_TRIV) : (is12 ? SPL_HORZ : SPL_TRIH);
  }

  // ---- child areas + depth counters: getCUSubPartitions / splitCurrArea / nextPart,
  //      unit_partitioner.cpp:94-190, 232-276, 604-760 ----
  SX_FN static int numParts(int split) { return split == SPL_QUAD ? 4 : (split == SPL_HORZ || split == SPL_VERT) ? 2 : 3; }
  SX_FN Node child(const Node &p, int i) const {
    Node c = p;
    c.lastSplit = p.split;
    c.partIdx = uint8_t(i);
    c.next = -1;
    c.split = 0;
    c.implicit = 0;
    switch (p.split) {
      case SPL_QUAD:
        c.log2w = p.log2w - 1;
        c.log2h = p.log2h - 1;
        c.x = int16_t(p.x + ((i & 1) << c.log2w));
        c.y = int16_t(p.y + ((i >> 1) << c.log2h));
        c.qt = p.qt + 1;
        c.mt = c.bt = 0;
        c.subdiv = p.subdiv + 2;
        break;
      case SPL_HORZ:
        c.log2h = p.log2h - 1;
        c.y = int16_t(p.y + (i << c.log2h));
        c.mt = p.mt + 1;
        c.bt = p.bt + 1;
        c.implBt = p.implBt + p.implicit;
        c.subdiv = p.subdiv + 1;
        break;
      case SPL_VERT:
        c.log2w = p.log2w - 1;
        c.x = int16_t(p.x + (i << c.log2w));
        c.mt = p.mt + 1;
        c.bt = p.bt + 1;
        c.implBt = p.implBt + p.implicit;
        c.subdiv = p.subdiv + 1;
        break;
      case SPL_TRIH:
        c.log2h = p.log2h - (i == 1 ? 1 : 2);
        c.y = int16_t(p.y + (i == 0 ? 0 : i == 1 ? (1 << (p.log2h - 2)) : (3 << (p.log2h - 2))));
        c.mt = p.mt + 1;
        c.bt = p.bt + (i == 1 ? 1 : 2);
        c.implBt = p.implBt + p.implicit;
        c.subdiv = p.subdiv + (i == 1 ? 1 : 2);
        break;
      default:  // SPL_TRIV
        c.log2w = p.log2w - (i == 1 ? 1 : 2);
        c.x = int16_t(p.x + (i == 0 ? 0 : i == 1 ? (1 << (p.log2w - 2)) : (3 << (p.log2w - 2))));
        c.mt = p.mt + 1;
        c.bt = p.bt + (i == 1 ? 1 : 2);
        c.implBt = p.implBt + p.implicit;
        c.subdiv = p.subdiv + (i == 1 ? 1 : 2);
        break;
    }
    c.qg = p.qg && c.subdiv <= sp.cu_qp_delta_subdiv;
    c.qgChroma = p.qgChroma && c.subdiv <= sp.cu_chroma_qp_offset_subdiv;
    return c;
  }

  // ---- coding_tree (:429-655) for one tree (ch = 0 luma, 1 chroma) rooted at `root` ----
  SX_FN void codingTree(Node root, int ch) {
    int sp_ = 0;
    root.next = -1;
    st[0] = root;
    while (sp_ >= 0) {
      Node &n = st[sp_];
      if (n.next < 0) {
        if (n.x >= int(sp.pic_w) || n.y >= int(sp.pic_h)) {  // cs.area.blocks[chType].contains(pos), :468, :600
          sp_--;
          continue;
        }
        // quantisation-group resets, :436-445
        if (tool(CABAC_SL_DQP) && n.qg && ch == 0) isDqpCoded = false;
        if (tool(CABAC_SL_CHROMA_QP_ADJ) && n.qgChroma) isChromaQpAdjCoded = false;
        const int implicit = implicitSplit(n, ch);
        // writer direction: the split of this node is on the path to the next CU of the list (split_series:
        // three bits per depth below the tree's root, the CU's own depth holding "no split" = 0)
        int want = SPL_NONE;
        if (E::kEncode) {
          const uint32_t ser = (e.getCu(nCu).split_series >> (3 * sp_)) & 7u;
          want = ser ? int(ser) : int(SPL_NONE);
        }
        const int split = splitCuMode(n, ch, implicit, want);
        if (sp_ == 0 && ch == 0) lumaRootSplit = split;
        if (split == SPL_NONE) {
          codingUnit(n, ch, st, sp_);
          sp_--;
          continue;
        }
        n.split = uint8_t(split);
        n.implicit = split == implicit;
        n.next = 0;
      }
      if (n.next >= numParts(n.split) || sp_ + 1 >= SX_MAX_DEPTH) {
        if (sp_ + 1 >= SX_MAX_DEPTH) status |= CABAC_SS_CAPACITY;
        sp_--;
        continue;
      }
      const Node c = child(n, n.next);
      n.next++;
      st[++sp_] = c;
    }
  }

  // ---- CodingUnit::checkCCLMAllowed, unit.cpp:289-372 (dual tree) ----
  SX_FN bool cclmAllowed(const Node &n, const Node *st, int depth) const {
    if (sp.log2_ctu <= 5) return true;
    // the chroma tree's stack starts at the 64x64 node: st[0].split is the split at depthFor64x64Node
    const int s1 = depth >= 1 ? int(st[0].split) : int(SPL_NONE);
    const int s2 = depth >= 2 ? int(st[1].split) : int(SPL_NONE);
    bool allow = false;
    if (s1 == SPL_QUAD || (s1 == SPL_HORZ && s2 == SPL_VERT))
      allow = true;
    else if (s1 == SPL_NONE)
      allow = true;
    else if (s1 == SPL_HORZ && s2 == SPL_NONE)
      allow = true;
    if (allow) {
      // the collocated luma 64x64 node: split by BT / TT, or unsplit with ISP -> no CCLM
      if (lumaRootSplit != SPL_NONE) {
        if (lumaRootSplit != SPL_QUAD) allow = false;
      } else if (lumaRootIsp) {
        allow = false;
      }
    }
    (void)n;
    return allow;
  }

  // ---- coding_unit (:764-825) + cu_pred_data (:1097) + cu_residual (:1396) for an intra CU ----
  SX_FN void codingUnit(const Node &n, int ch, const Node *st, int depth) {
    cabac_cu_desc &cu = cuRec;
    if (E::kEncode) {
      cu = e.getCu(nCu);  // the caller's record: must describe this very node
      if (cu.x != uint16_t(n.x) || cu.y != uint16_t(n.y) || cu.log2_w != n.log2w || cu.log2_h != n.log2h ||
          cu.ch_type != uint8_t(ch) || cu.bdpcm > 2 || cu.mrl_idx > 2 || cu.isp > 2 || cu.lfnst_idx > 2 || cu.mts_idx > 5 ||
          cu.intra_dir > 70 || (!(cu.flags & CABAC_CU_FROM_DIR) && cu.pred_syntax > 60)) {
        status |= CABAC_SS_BAD_DESC;
        e.abort();
        return;
      }
    } else {
      cu.x = uint16_t(n.x);
      cu.y = uint16_t(n.y);
      cu.log2_w = n.log2w;
      cu.log2_h = n.log2h;
      cu.ch_type = uint8_t(ch);
      cu.qt_depth = n.qt;
      cu.mt_depth = n.mt;
      cu.flags = cu.bdpcm = cu.mrl_idx = cu.isp = cu.lfnst_idx = cu.mts_idx = cu.pred_syntax = 0;
      cu.first_tu = nTu;
      cu.n_tu = 0;
      cu.qp_delta = 0;
      cu.chroma_qp_adj = 0;
      cu.intra_dir = 0;
      cu.reserved = 0;
      cu.cbf_mask = 0;
      uint32_t ser = 0;
      for (int d = 0; d < depth && d < 10; d++) ser |= uint32_t(st[d].split & 7u) << (3 * d);
      cu.split_series = ser;
    }
    const int w = 1 << n.log2w, h = 1 << n.log2h;
    const int tsMax = 1 << sp.log2_max_ts, maxTb = 1 << sp.log2_max_tb;
    if (ch == 0) {
      // bdpcm_mode(Y), :1063-1095 with CU::bdpcmAllowed, unit_tools.cpp:226-241
      if (tool(CABAC_SL_BDPCM) && w <= tsMax && h <= tsMax) {
        if (e.bin(SX_BDPCM_MODE + 0, cu.bdpcm != 0)) cu.bdpcm = uint8_t(1 + e.bin(SX_BDPCM_MODE + 1, cu.bdpcm == 2));
      }
      if (E::kEncode && (cu.flags & CABAC_CU_FROM_DIR) && !cu.bdpcm) {
        // writer direction from the decision itself: pu.intraDir[0] -> mpm flag / index / remainder, the way
        // CABACWriter::intra_luma_pred_modes binarises it (cabac_writer.cpp:949-1060)
        if (cu.flags & CABAC_CU_MIP) {
          cu.pred_syntax = cu.intra_dir;
        } else {
          int mpm[6];
          intraMpms(n, mpm);
          int at = -1;
          for (int i = 5; i >= 0; i--) at = mpm[i] == int(cu.intra_dir) ? i : at;
          if (at >= 0) {
            cu.flags |= CABAC_CU_MPM;
            cu.pred_syntax = uint8_t(at);
          } else {
            cu.flags &= uint8_t(~CABAC_CU_MPM);
            int mode = cu.intra_dir;
            for (int i = 0; i < 6; i++) mode -= mpm[i] < int(cu.intra_dir);  // rank among the non-MPM modes
            cu.pred_syntax = uint8_t(mode);
            if (cu.mrl_idx || cu.intra_dir > 66) {  // a reference line other than 0 only goes with an MPM
              status |= CABAC_SS_BAD_DESC;
              e.abort();
              return;
            }
          }
          if (cu.mrl_idx && at == 0) {  // ... and not with PLANAR
            status |= CABAC_SS_BAD_DESC;
            e.abort();
            return;
          }
        }
      }
      if (!cu.bdpcm) {  // intra_luma_pred_modes, :1231-1316
        int mip = 0;
        if (tool(CABAC_SL_MIP)) {  // mip_flag :3381 + DeriveCtx::CtxMipFlag
          const uint16_t left = nbLeft(0, n.x, n.y), above = nbAbove(0, n.x, n.y);
          int ctx = (left ? nbMip(left) : 0) + (above ? nbMip(above) : 0);
          if (w > 2 * h || h > 2 * w) ctx = 3;
          mip = int(e.bin(SX_MIP_FLAG + ctx, (cu.flags & CABAC_CU_MIP) != 0));
        }
        if (mip) {  // mip_pred_mode :3408
          cu.flags |= CABAC_CU_MIP;
          if (e.ep(1, (cu.flags & CABAC_CU_MIP_TRANSPOSED) != 0)) cu.flags |= CABAC_CU_MIP_TRANSPOSED;
          const int numModes = (w == 4 && h == 4) ? 16 : (w == 4 || h == 4 || (w == 8 && h == 8)) ? 8 : 6;
          cu.pred_syntax = uint8_t(truncBin(uint32_t(numModes), cu.pred_syntax));
        } else {
          // extend_ref_line :1188-1229
          if (tool(CABAC_SL_MRL) && (n.y & ((1 << sp.log2_ctu) - 1)) != 0) {
            if (e.bin(SX_MULTI_REF_LINE_IDX + 0, cu.mrl_idx != 0))
              cu.mrl_idx = uint8_t(1 + e.bin(SX_MULTI_REF_LINE_IDX + 1, cu.mrl_idx == 2));
          }
          // isp_mode :2785-2806, CU::canUseISP unit_tools.cpp:290-298
          if (!cu.mrl_idx && tool(CABAC_SL_ISP) && n.log2w + n.log2h > 4 && w <= maxTb && h <= maxTb) {
            if (e.bin(SX_ISP_MODE + 0, cu.isp != 0)) cu.isp = uint8_t(1 + e.bin(SX_ISP_MODE + 1, cu.isp == 2));
          }
          const bool mpm = cu.mrl_idx ? true : e.bin(SX_INTRA_LUMA_MPM_FLAG, (cu.flags & CABAC_CU_MPM) != 0) != 0;
          if (mpm) {
            cu.flags |= CABAC_CU_MPM;
            const uint32_t wi = cu.pred_syntax;
            uint32_t idx = cu.mrl_idx ? 1u : e.bin(SX_INTRA_LUMA_PLANAR_FLAG + (cu.isp ? 0 : 1), wi > 0);
            if (idx) idx += e.ep(1, wi > 1);
            if (idx > 1) idx += e.ep(1, wi > 2);
            if (idx > 2) idx += e.ep(1, wi > 3);
            if (idx > 3) idx += e.ep(1, wi > 4);
            cu.pred_syntax = uint8_t(idx);
          } else {
            cu.pred_syntax = uint8_t(truncBin(61, cu.pred_syntax));  // NUM_LUMA_MODE - NUM_MOST_PROBABLE_MODES
          }
        }
      }
      // pu.intraDir[0], :1240-1316
      if (cu.bdpcm) {
        cu.intra_dir = cu.bdpcm == 2 ? 50 : 18;  // VER_IDX : HOR_IDX
      } else if (cu.flags & CABAC_CU_MIP) {
        cu.intra_dir = cu.pred_syntax;
      } else {
        int mpm[6];
        intraMpms(n, mpm);
        if (cu.flags & CABAC_CU_MPM) {
          cu.intra_dir = uint8_t(mpm[cu.pred_syntax < 6 ? cu.pred_syntax : 5]);
        } else {
          for (int i = 1; i < 6; i++) {  // std::sort(mpm, mpm + 6)
            const int v = mpm[i];
            int j = i - 1;
            while (j >= 0 && mpm[j] > v) {
              mpm[j + 1] = mpm[j];
              j--;
            }
            mpm[j + 1] = v;
          }
          int mode = cu.pred_syntax;
          for (int i = 0; i < 6; i++) mode += mode >= mpm[i];
          cu.intra_dir = uint8_t(mode);
        }
      }
      if (depth == 0) lumaRootIsp = cu.isp;
    } else {
      const int cw = w >> 1, chh = h >> 1;
      if (tool(CABAC_SL_BDPCM) && cw <= tsMax && chh <= tsMax) {
        if (e.bin(SX_BDPCM_MODE + 2, cu.bdpcm != 0)) cu.bdpcm = uint8_t(1 + e.bin(SX_BDPCM_MODE + 3, cu.bdpcm == 2));
      }
      if (E::kEncode && (cu.flags & CABAC_CU_FROM_DIR) && !cu.bdpcm) {
        // pu.intraDir[1] -> cclm flag / index or the candidate's position (cabac_writer.cpp:1167-1210)
        cu.flags &= uint8_t(~CABAC_CU_CCLM);
        if (cu.intra_dir >= 67 && cu.intra_dir <= 69) {
          cu.flags |= CABAC_CU_CCLM;
          cu.pred_syntax = uint8_t(cu.intra_dir - 67);
        } else if (cu.intra_dir == 70) {
          cu.pred_syntax = 0;
        } else {
          const int cand[4] = {0, 50, 18, 1};
          const int lumaMode = e.lumaMap()[((((n.y + (h >> 1)) & 63) >> 2) << 4) + (((n.x + (w >> 1)) & 63) >> 2)];
          int at = -1;
          for (int i = 3; i >= 0; i--) at = (cand[i] == lumaMode ? 66 : cand[i]) == int(cu.intra_dir) ? i : at;
          if (at < 0) {
            status |= CABAC_SS_BAD_DESC;
            e.abort();
            return;
          }
          cu.pred_syntax = uint8_t(1 + at);
        }
      }
      if (!cu.bdpcm) {  // intra_chroma_pred_mode :1353-1394
        bool done = false;
        if (tool(CABAC_SL_CCLM) && cclmAllowed(n, st, depth)) {
          if (e.bin(SX_CCLM_MODE_FLAG, (cu.flags & CABAC_CU_CCLM) != 0)) {
            cu.flags |= CABAC_CU_CCLM;
            const uint32_t ws = cu.pred_syntax;
            uint32_t sym = e.bin(SX_CCLM_MODE_IDX, ws > 0);
            if (sym) sym += e.ep(1, ws > 1);
            cu.pred_syntax = uint8_t(sym);
            done = true;
          }
        }
        if (!done && e.bin(SX_INTRA_CHROMA_PRED_MODE, cu.pred_syntax != 0)) cu.pred_syntax = uint8_t(1 + e.ep(2, (cu.pred_syntax - 1u) & 3u));
      }
      // pu.intraDir[1], :1318-1394 with PU::getIntraChromaCandModes (unit_tools.cpp:460-483)
      if (cu.bdpcm) {
        cu.intra_dir = cu.bdpcm == 2 ? 50 : 18;
      } else if (cu.flags & CABAC_CU_CCLM) {
        cu.intra_dir = uint8_t(67 + cu.pred_syntax);  // LM_CHROMA_IDX, MDLM_L_IDX, MDLM_T_IDX
      } else if (cu.pred_syntax == 0) {
        cu.intra_dir = 70;  // DM_CHROMA_IDX (resolved at prediction time)
      } else {
        const int cand[4] = {0, 50, 18, 1};  // PLANAR, VER, HOR, DC
        // the collocated luma PU: at the centre of the chroma block (PU::getCoLocatedLumaPU, unit_tools.cpp:603-613)
        const int lumaMode = e.lumaMap()[((((n.y + (h >> 1)) & 63) >> 2) << 4) + (((n.x + (w >> 1)) & 63) >> 2)];
        const int m = cand[(cu.pred_syntax - 1) & 3];
        cu.intra_dir = uint8_t(m == lumaMode ? 66 : m);  // VDIA_IDX replaces the candidate equal to the luma mode
      }
    }

    // ---- cu_residual :1396-1437 -> transform_tree :2297-2351 ----
    cuRes = 0;
    bool anyTs = false;
    if (ch == 0 && cu.isp) {
      // TUIntraSubPartitioner: getTUIntraSubPartitions, unit_partitioner.cpp:838-895; CU::getISPSplitDim
      const bool rows = cu.isp == 1;
      const int splitDim = rows ? h : w, other = rows ? w : h;
      const int factor = other < 16 ? (16 >> (rows ? n.log2w : n.log2h)) : 1;
      const int part = (splitDim >> 2) < factor ? factor : (splitDim >> 2);
      const int nParts = splitDim / part;
      int log2part = 0;
      while ((1 << log2part) < part) log2part++;
      bool rootCbfSoFar = false, prevCbf = false;
      for (int i = 0; i < nParts; i++) {
        bool cbf;
        if (i == nParts - 1 && !rootCbfSoFar)
          cbf = true;  // lastCbfIsInferred, :2500-2511
        else
          cbf = e.bin(SX_QT_CBF_Y + 2 + (prevCbf ? 1 : 0), (cu.cbf_mask >> i) & 1u) != 0;
        rootCbfSoFar |= cbf;
        prevCbf = cbf;
        if (cbf) cu.cbf_mask |= uint16_t(1u << i);
        if (cbf) dqp(cu);
        if (cbf) {
          const int tw = rows ? n.log2w : log2part, th = rows ? log2part : n.log2h;
          anyTs |= residual(cu, tw, th, 0, 0);  // TU::isTSAllowed: never with ISP
        }
      }
    } else {
      // TU_MAX_TR_SPLIT tiling (getMaxTuTiling, unit_partitioner.cpp:762-836): raster order
      const int tl2w = n.log2w > int(sp.log2_max_tb) ? int(sp.log2_max_tb) : n.log2w;
      const int tl2h = n.log2h > int(sp.log2_max_tb) ? int(sp.log2_max_tb) : n.log2h;
      const int nx = 1 << (n.log2w - tl2w), ny = 1 << (n.log2h - tl2h);
      for (int ty = 0; ty < ny; ty++)
        for (int tx = 0; tx < nx; tx++) {
          const int tile = (ty * nx + tx) & 3;  // at most 2 x 2 tiles (128 / 32 is not reachable in a dual tree)
          if (ch == 0) {
            const bool cbf = e.bin(SX_QT_CBF_Y + (cu.bdpcm ? 1 : 0), (cu.cbf_mask >> tile) & 1u) != 0;  // cbf_comp :2353-2388
            if (cbf) cu.cbf_mask |= uint16_t(1u << tile);
            if (cbf || w > 64 || h > 64) dqp(cu);
            if (cbf) {
              const bool tsAllowed = tool(CABAC_SL_TS) && !cu.bdpcm && (1 << tl2w) <= tsMax && (1 << tl2h) <= tsMax;
              anyTs |= residual(cu, tl2w, tl2h, 0,
                                uint8_t((tsAllowed ? CABAC_TU_TS_ALLOWED : 0) | (cu.bdpcm ? CABAC_TU_BDPCM : 0)));
            }
          } else {
            const int cl2w = tl2w - 1, cl2h = tl2h - 1;
            const bool cbfCb = e.bin(SX_QT_CBF_CB + (cu.bdpcm ? 1 : 0), (cu.cbf_mask >> (4 + tile)) & 1u) != 0;
            const bool cbfCr = e.bin(SX_QT_CBF_CR + (cu.bdpcm ? 2 : (cbfCb ? 1 : 0)), (cu.cbf_mask >> (8 + tile)) & 1u) != 0;
            cu.cbf_mask |= uint16_t((cbfCb ? 1u << (4 + tile) : 0u) | (cbfCr ? 1u << (8 + tile) : 0u));
            const bool cbfChroma = cbfCb || cbfCr;
            // cu_chroma_qp_offset :2602-2630 (condition at :2551-2563)
            if (tool(CABAC_SL_CHROMA_QP_ADJ) && ((w >> 1) > 64 || (h >> 1) > 64 || cbfChroma) && !isChromaQpAdjCoded) {
              const uint32_t wa = cu.chroma_qp_adj;
              uint32_t adj = e.bin(SX_CHROMA_QP_ADJ_FLAG, wa != 0);
              if (adj && sp.chroma_qp_offset_list_len > 1)
                adj += unaryMaxSymbol(SX_CHROMA_QP_ADJ_IDC, SX_CHROMA_QP_ADJ_IDC, sp.chroma_qp_offset_list_len - 1,
                                      wa ? wa - 1 : 0);
              cu.chroma_qp_adj = uint8_t(adj);
              isChromaQpAdjCoded = true;
            }
            int cbfMask = (cbfCb ? 2 : 0) + (cbfCr ? 1 : 0);
            SX_OPAQUE(cbfMask);
            int joint = 0;
            if (tool(CABAC_SL_JCCR) && cbfMask) {  // joint_cb_cr :2632-2645 (intra CU)
              if (e.bin(SX_JOINT_CBCR_FLAG + cbfMask - 1, (cu.cbf_mask >> (12 + tile)) & 1u)) joint = cbfMask;
            }
            if (joint) {
              cu.flags |= CABAC_CU_JCCR;
              cu.cbf_mask |= uint16_t(1u << (12 + tile));
            }
            const bool tsAllowed = tool(CABAC_SL_TS) && !cu.bdpcm && (1 << cl2w) <= tsMax && (1 << cl2h) <= tsMax;
            const uint8_t fl = uint8_t((tsAllowed ? CABAC_TU_TS_ALLOWED : 0) | (cu.bdpcm ? CABAC_TU_BDPCM : 0));
            if (cbfCb) anyTs |= residual(cu, cl2w, cl2h, 1, fl);
            if (cbfCr && joint != 3) anyTs |= residual(cu, cl2w, cl2h, 2, fl);
          }
        }
    }

    // ---- residual_lfnst_mode :2808-2863 ----
    {
      bool skip = false;
      if (cu.isp) {  // CU::canUseLfnstWithISP, unit_tools.cpp:300-318
        const bool rows = cu.isp == 1;
        const int splitDim = rows ? h : w, other = rows ? w : h;
        const int factor = other < 16 ? (16 >> (rows ? n.log2w : n.log2h)) : 1;
        const int part = (splitDim >> 2) < factor ? factor : (splitDim >> 2);
        const int tw = rows ? w : part, th = rows ? part : h;
        if (!(tw >= 4 && th >= 4)) skip = true;
      }
      if (tool(CABAC_SL_LFNST) && (cu.flags & CABAC_CU_MIP) && !(w >= 16 && h >= 16)) skip = true;
      if (ch == 1 && ((w >> 1) < 4 || (h >> 1) < 4)) skip = true;
      if (w > maxTb || h > maxTb) skip = true;
      if (!skip && tool(CABAC_SL_LFNST)) {
        const bool violates = (cuRes & CABAC_TUR_VIOLATES_LFNST) != 0;
        const bool lastPos = (cuRes & CABAC_TUR_LFNST_LAST_POS) != 0;
        if (!((!lastPos && !cu.isp) || violates || anyTs)) {
          uint32_t idx = e.bin(SX_LFNST_IDX + 1, cu.lfnst_idx != 0);  // isSepTree: always in a dual-tree slice
          if (idx) idx += e.bin(SX_LFNST_IDX + 2, cu.lfnst_idx == 2);
          cu.lfnst_idx = uint8_t(idx);
        }
      }
    }
    // ---- mts_idx :2754-2783, CU::isMTSAllowed unit_tools.cpp:243-260 ----
    if (ch == 0) {
      const bool firstTuTs = anyTs && !cu.isp;  // firstTU->mtsIdx[Y] == MTS_SKIP: MTS only exists for single-TU CUs
      const uint32_t wm = cu.mts_idx;
      cu.mts_idx = firstTuTs ? 1 : 0;
      const bool allowed = tool(CABAC_SL_MTS_INTRA) && w <= 32 && h <= 32 && !cu.isp && !(cu.bdpcm && w <= tsMax && h <= tsMax);
      if (allowed && !(cuRes & CABAC_TUR_VIOLATES_MTS) && (cuRes & CABAC_TUR_MTS_LAST_POS) && cu.lfnst_idx == 0 &&
          !firstTuTs) {
        if (e.bin(SX_MTS_IDX + 0, wm >= 2)) {
          int m = 2;
          for (int i = 0; i < 3; i++) {
            const uint32_t s = e.bin(SX_MTS_IDX + 1 + i, wm > uint32_t(2 + i));
            m += int(s);
            if (!s) break;
          }
          cu.mts_idx = uint8_t(m);
        }
      }
    }
    nbStore(ch, n, (cu.flags & CABAC_CU_MIP) ? 1 : 0, (cu.flags & CABAC_CU_MIP) ? 0 : cu.intra_dir);
    if (nCu < capCu)
      e.putCu(nCu, cu);
    else
      status |= CABAC_SS_CAPACITY;
    nCu++;
  }

  // cu_qp_delta :2578-2600 under the condition of :2537-2545
  SX_FN void dqp(cabac_cu_desc &cu) {
    if (!tool(CABAC_SL_DQP) || isDqpCoded) return;
    const int wq = cu.qp_delta;
    const uint32_t wAbs = uint32_t(wq < 0 ? -wq : wq);
    int dq = int(unaryMaxSymbol(SX_DELTA_QP, SX_DELTA_QP + 1, 5, wAbs));
    if (dq >= 5) dq += int(expGolombEp(0, wAbs >= 5 ? wAbs - 5 : 0));
    if (dq > 0 && e.ep(1, wq < 0)) dq = -dq;
    cu.qp_delta = int8_t(dq);
    cu.flags |= CABAC_CU_DQP;
    isDqpCoded = true;
  }

  // residual_coding of one block; returns whether it was transform-skipped
  SX_FN bool residual(cabac_cu_desc &cu, int log2w, int log2h, int comp, uint8_t flags) {
    cabac_tu_desc d;
    d.coeff_offset = coeffBase + nCoeff;
    d.log2_w = uint8_t(log2w);
    d.log2_h = uint8_t(log2h);
    d.comp_id = uint8_t(comp);
    d.flags = flags;
    d.reserved = 0;
    const uint64_t area = uint64_t(1) << (log2w + log2h);
    if (nTu >= capTu || nCoeff + area > capCoeff) {
      // No room: the bins still have to be consumed to stay in step, but there is nowhere to put the levels.
      status |= CABAC_SS_CAPACITY;
      e.abort();
      return false;
    }
    const uint32_t r = e.residual(nTu, d);
    cuRes |= r;
    nTu++;
    nCoeff += area;
    if (!E::kEncode) cu.n_tu++;
    return (r & CABAC_TUR_TS) != 0;
  }

  // ---- sao :284-419 ----
  SX_FN void sao(cabac_ctu_desc &c, int ctuX, int ctuY) {
    const bool luma = tool(CABAC_SL_SAO_LUMA), chroma = tool(CABAC_SL_SAO_CHROMA);
    if (!luma && !chroma) return;
    int merge = -1;
    // writer direction: the record says "merge" in sao_mode (2) and which neighbour in sao_type (0 left, 1 above)
    const bool wMerge = c.sao_mode[0] == 2;  // the luma entry carries the merge (:330-333), whichever planes use SAO
    const uint32_t wMergeLeft = wMerge && c.sao_type[0] == 0, wMergeUp = wMerge && c.sao_type[0] == 1;
    if (leftAvail(ctuX)) merge += int(e.bin(SX_SAO_MERGE_FLAG, wMergeLeft));
    if (merge < 0 && aboveAvail(ctuY)) merge += int(e.bin(SX_SAO_MERGE_FLAG, wMergeUp)) << 1;
    if (merge >= 0) {
      c.sao_mode[0] = 2;
      c.sao_type[0] = uint8_t(merge);
      if (chroma) {
        c.sao_mode[1] = c.sao_mode[2] = 2;
        c.sao_type[1] = c.sao_type[2] = uint8_t(merge);
      }
      return;
    }
    const int first = luma ? 0 : 1, last = chroma ? 2 : 0;
    const int bd = int(sp.bit_depth) < 10 ? int(sp.bit_depth) : 10;
    const uint32_t maxOffset = (1u << (bd - 5)) - 1;
    for (int comp = first; comp <= last; comp++) {
      const uint8_t wType = c.sao_type[comp];  // writer direction: 0 band, 1..4 = edge class + 1
      int8_t wOff[4];
      for (int k = 0; k < 4; k++) wOff[k] = c.sao_offset[comp][k];
      const uint8_t wBand = c.sao_band[comp];
      if (comp != 2) {
        if (e.bin(SX_SAO_TYPE_IDX, c.sao_mode[comp] == 1)) {
          c.sao_mode[comp] = 1;
          c.sao_type[comp] = e.ep(1, wType != 0) ? 1 : 0;  // 1 = edge (class added below), 0 = band
        } else {
          c.sao_mode[comp] = 0;
        }
      } else {
        c.sao_mode[2] = c.sao_mode[1];
        c.sao_type[2] = c.sao_type[1] ? 1 : 0;
      }
      if (c.sao_mode[comp] == 0) continue;
      int off[4];
      for (int k = 0; k < 4; k++) off[k] = int(unaryMaxEp(maxOffset, uint32_t(wOff[k] < 0 ? -wOff[k] : wOff[k])));
      if (c.sao_type[comp] == 0) {  // band offset
        for (int k = 0; k < 4; k++)
          if (off[k] && e.ep(1, wOff[k] < 0)) off[k] = -off[k];
        c.sao_band[comp] = uint8_t(e.ep(5, wBand));
      } else if (comp != 2) {
        c.sao_type[comp] = uint8_t(1 + e.ep(2, (wType - 1u) & 3u));  // sao_eo_class
      } else {
        c.sao_type[2] = c.sao_type[1];
      }
      for (int k = 0; k < 4; k++) c.sao_offset[comp][k] = int8_t(off[k]);
    }
  }

  // ---- coding_tree_unit :103-209 ----
  SX_FN void codingTreeUnit(int ctuX, int ctuY) {
    cabac_ctu_desc &c = ctuRec;
    if (E::kEncode) {
      c = e.getCtu(nCtu);
    } else {
      for (int i = 0; i < 3; i++) {
        c.sao_mode[i] = c.sao_type[i] = c.sao_band[i] = 0;
        c.alf_flag[i] = 0;
        c.reserved[i] = 0;
        for (int k = 0; k < 4; k++) c.sao_offset[i][k] = 0;
      }
      c.alf_filter_idx = 0;
      c.alf_alt[0] = c.alf_alt[1] = c.ccalf_idc[0] = c.ccalf_idc[1] = 0;
    }
    sao(c, ctuX, ctuY);
    const int col = (ctuX - int(sp.x0)) >> sp.log2_ctu;
    const bool lAvail = leftAvail(ctuX), aAvail = aboveAvail(ctuY);
    const uint8_t aboveFlags = aAvail ? e.ctuLine()[col] : uint8_t(0);
    const uint8_t leftFlags = lAvail ? leftCtuFlags : uint8_t(0);
    uint8_t mine = 0;
    if (tool(CABAC_SL_ALF_Y)) {
      for (int comp = 0; comp < 3; comp++) {
        if (!tool(CABAC_SL_ALF_Y << comp)) continue;
        const int ctx = ((leftFlags >> comp) & 1) + ((aboveFlags >> comp) & 1);
        const uint32_t f = e.bin(SX_CTB_ALF_FLAG + comp * 3 + ctx, c.alf_flag[comp] != 0);
        c.alf_flag[comp] = uint8_t(f);
        mine |= uint8_t(f << comp);
        if (comp == 0 && f) {  // readAlfCtuFilterIndex :211-235
          const uint32_t numAps = sp.alf_num_aps_luma;
          const uint32_t wi = c.alf_filter_idx;
          uint32_t idx = 0;
          if (numAps > 0) {
            if (e.bin(SX_ALF_USE_TEMPORAL_FILT, wi >= 16)) {
              if (numAps > 1) idx = truncBin(numAps, wi >= 16 ? wi - 16 : 0);
              idx += 16;
            } else {
              idx = truncBin(16, wi & 15u);
            }
          } else {
            idx = truncBin(16, wi & 15u);
          }
          c.alf_filter_idx = uint8_t(idx);
        }
        if (comp > 0 && f) {
          const uint32_t wa = c.alf_alt[comp - 1];
          uint32_t decoded = 0;
          while (decoded + 1 < sp.alf_chroma_num_alts && e.bin(SX_CTB_ALF_ALTERNATIVE + comp - 1, decoded < wa)) decoded++;
          c.alf_alt[comp - 1] = uint8_t(decoded);
        }
      }
    }
    for (int comp = 1; comp < 3; comp++) {  // ccAlfFilterControlIdc :237-276
      if (!tool(CABAC_SL_CCALF_CB << (comp - 1))) continue;
      int ctx = ((leftFlags >> (2 + comp)) & 1) + ((aboveFlags >> (2 + comp)) & 1) + (comp == 2 ? 3 : 0);
      const uint32_t wc = c.ccalf_idc[comp - 1];
      uint32_t idc = e.bin(SX_CCALF_FILTER_CONTROL_FLAG + ctx, wc != 0);
      if (idc)
        while (idc != sp.ccalf_filter_count[comp - 1] && e.ep(1, idc < wc)) idc++;
      c.ccalf_idc[comp - 1] = uint8_t(idc);
      if (idc) mine |= uint8_t(4u << comp);
    }
    e.ctuLineStore(col, mine);
    leftCtuFlags = mine;
    if (nCtu < capCtu)
      e.putCtu(nCtu, c);
    else
      status |= CABAC_SS_CAPACITY;
    nCtu++;

    Node root;
    root.x = int16_t(ctuX);
    root.y = int16_t(ctuY);
    root.log2w = root.log2h = uint8_t(sp.log2_ctu);
    root.lastSplit = SPL_CTU_LEVEL;
    root.partIdx = 0;
    root.qt = root.mt = root.bt = root.implBt = 0;
    root.subdiv = 0;
    root.qg = root.qgChroma = 1;
    root.split = 0;
    root.next = -1;
    root.implicit = 0;
    // The trees of this CTU in coding order.  Above 64x64 the implicit quad split is shared and luma and chroma
    // alternate per 64x64 node (:464-517; split_cu_mode of the 128x128 node reads nothing: canNo = false, only QT
    // possible); otherwise the luma tree of the CTU is followed by its chroma tree (:196-208).  One loop, so the
    // tree walk - the whole CU parser - is instantiated once and its state stays in registers.
    int nRoots = 0;
    if (sp.log2_ctu > 6) {
      if (tool(CABAC_SL_DQP)) isDqpCoded = false;
      if (tool(CABAC_SL_CHROMA_QP_ADJ)) isChromaQpAdjCoded = false;
      root.split = SPL_QUAD;
      for (int i = 0; i < 4; i++) {
        const Node q = child(root, i);
        if (q.x >= int(sp.pic_w) || q.y >= int(sp.pic_h)) continue;
        roots[nRoots++] = q;
      }
    } else {
      roots[nRoots++] = root;
    }
    for (int r = 0; r < 2 * nRoots; r++) {
      const int ch = r & 1;
      if (ch == 0) {
        lumaRootSplit = SPL_NONE;
        lumaRootIsp = 0;
      }
      codingTree(roots[r >> 1], ch);
    }
  }

  // ---- the substream: its CTUs in raster order inside the rectangle, then the terminating bin.
  //      With entropy coding sync (WPP) the slice's bytes are its CTU-row substreams back to back
  //      (EncSlice.cpp:2110-2125): every row ends with end_of_subset_one_bit + byte alignment, the next row
  //      restarts the arithmetic decoder at the next byte (the decoder has then read exactly through the row's
  //      last byte: BinDecoderBase::finish, arith_codec.cpp:68-73) and starts from the contexts saved after the
  //      FIRST CTU of the row above (DecSlice.cpp:194-210, 241-245).  Rows are walked one after the other by the
  //      same warp: neighbour lines carry over, so above-neighbours are available as inside any slice. ----
  SX_FN void run(bool terminate) {
    const uint32_t refuse = CABAC_SL_IBC | CABAC_SL_PLT | CABAC_SL_ACT;
    if ((sp.tools & refuse) || !tool(CABAC_SL_INTRA) || !tool(CABAC_SL_DUAL_TREE) || sp.chroma_format != 1 ||
        sp.log2_ctu < 5 || sp.log2_ctu > 7 || ((sp.x1 - sp.x0 + 3) >> 2) > uint32_t(SX_MAX_LINE) || sp.x1 <= sp.x0 ||
        sp.y1 <= sp.y0 || sp.x1 > sp.pic_w || sp.y1 > sp.pic_h || sp.log2_max_tb < 5 || sp.log2_max_tb > 6) {
      status |= CABAC_SS_UNSUPPORTED;
      return;
    }
    if (E::kEncode && tool(CABAC_SL_WPP) && !e.rows()) {  // writer direction: the rows of a WPP slice are separate
      status |= CABAC_SS_UNSUPPORTED;                       // substreams; the engine must be set up to cut them
      return;
    }
    const int ctu = 1 << sp.log2_ctu;
    const bool wpp = tool(CABAC_SL_WPP);
    for (int y = int(sp.y0); y < int(sp.y1) && !e.aborted(); y += ctu) {
      if (wpp && y > int(sp.y0)) {
        if (!e.trm(1)) {  // end_of_subset_one_bit
          status |= CABAC_SS_NO_TERMINATE;
          return;
        }
        e.nextRow();  // byte-aligned restart + contexts of the row above after its first CTU
      }
      leftCtuFlags = 0;
      for (int x = int(sp.x0); x < int(sp.x1) && !e.aborted(); x += ctu) {
        codingTreeUnit(x, y);
        if (wpp && x == int(sp.x0)) e.saveRowContexts();
      }
    }
    if (terminate && !e.aborted() && !e.trm(1)) status |= CABAC_SS_NO_TERMINATE;
  }
};

}  // namespace cabac_syntax
