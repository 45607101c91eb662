// cabac_residual.cuh - residual_coding binarisation and context selection.
//
// Encoder side: everything a bin needs except the arithmetic-coder state is a
// pure function of the coefficient block, so the 16 positions of a coefficient
// group are binarised by 16 lanes at once (template sums, context ids, Rice
// parameters, dependent-quantisation state, context-bin budget, sign pattern
// all come out of ballots and popcounts) and land as command words in a
// shared-memory queue that lane 0's arithmetic coder drains.
//
// Decoder side: every decoded bin selects the next context, so parsing is
// serial; lane 0 parses into a shared-memory tile, the warp zero-fills and
// writes tiles back coalesced.
//
// Reference behaviour restated (paths relative to p-sawicki/entropy_coding):
//   src/common/context_modelling.{hpp,cpp}   CoeffCodingContext
//   src/entropy_codec/cabac_writer.cpp:2424-3062
//   src/entropy_codec/cabac_reader.cpp:2647-3339
//   src/common/rom.cpp                        scan orders, LUTs
#pragma once
#include "cabac_device.cuh"

namespace cabac {

// ---------------------------------------------------------------------------
// Context layout.  Residual coding touches 230 of the 379 contexts
// (contexts.cpp:77-770, positional ids): 86..291, 310..311, 357..378.  The
// residual kernels keep just those in shared memory, re-indexed densely.
// ---------------------------------------------------------------------------
constexpr int RCTX = 230;
__host__ __device__ constexpr int rctx(int id) {
  return id < 310 ? id - 86 : (id < 357 ? id - 310 + 206 : id - 357 + 208);
}
__host__ __device__ constexpr int rctx_inv(int r) { return r < 206 ? r + 86 : (r < 208 ? r - 206 + 310 : r - 208 + 357); }
__host__ __device__ constexpr bool is_residual_ctx(int id) {
  return (id >= 86 && id <= 291) || id == 310 || id == 311 || (id >= 357 && id <= 378);
}

constexpr int R_SIG_GROUP = rctx(86);     // + 2 * chType + (right|below)
constexpr int R_SIG_FLAG = rctx(90);      // sets at 90,102,110,122,130,142
constexpr int R_PAR_FLAG = rctx(150);     // luma 150 (21), chroma 171 (11)
constexpr int R_GTX_FLAG = rctx(182);     // 182,203,214,235
constexpr int R_LAST_X = rctx(246);       // luma 246 (20), chroma 266 (3)
constexpr int R_LAST_Y = rctx(269);       // luma 269 (20), chroma 289 (3)
constexpr int R_TS_FLAG = rctx(310);
constexpr int R_TS_SIG_GROUP = rctx(357);
constexpr int R_TS_SIG_FLAG = rctx(360);
constexpr int R_TS_PAR_FLAG = rctx(363);
constexpr int R_TS_GTX_FLAG = rctx(364);
constexpr int R_TS_LRG1_FLAG = rctx(369);
constexpr int R_TS_SIGN = rctx(373);

__device__ __forceinline__ int sigFlagBase(int ch, int set) {  // SigFlag[ch + 2 * set]
  // offsets 90,102 | 110,122 | 130,142  -> 20 per set, chroma +12
  return R_SIG_FLAG + 20 * set + 12 * ch;
}
__device__ __forceinline__ int parFlagBase(int ch) { return R_PAR_FLAG + 21 * ch; }
__device__ __forceinline__ int gtxFlagBase(int idx) {  // GtxFlag[idx], idx = ch + 2 * x
  // 182, 203, 214, 235
  return R_GTX_FLAG + 32 * (idx >> 1) + 21 * (idx & 1);
}
__device__ __forceinline__ int lastXBase(int ch) { return R_LAST_X + 20 * ch; }
__device__ __forceinline__ int lastYBase(int ch) { return R_LAST_Y + 20 * ch; }

// ---------------------------------------------------------------------------
// ROM (rom.cpp).  Scan tables live in global memory (L1-resident, 1.3 KB).
// ---------------------------------------------------------------------------
struct ScanRom {
  uint8_t cgScan[4][4][64];     // [log2 wig][log2 hig][k] -> cgx | cgy << 4   (SCAN_UNGROUPED / DIAG)
  uint8_t cgScanInv[4][4][64];  // [..][cgy * wig + cgx] -> k
  uint8_t inScan[4][16];        // [shape][p] -> px | py << 4                  (within one CG)
  uint8_t inScanInv[4][16];     // [shape][py * gw + px] -> p
};
// shapes: 0 = 4x4, 1 = 2x2, 2 = 2 wide x 8 high, 3 = 8 wide x 2 high

__device__ ScanRom g_rom;

__device__ __forceinline__ uint32_t minInGroup(uint32_t g) {  // g_minInGroup, rom.cpp:18-19
  // 0,1,2,3,4,6,8,12,16,24,32,48,...
  return g < 4 ? g : ((2u + (g & 1u)) << ((g >> 1) - 1));
}
__device__ __forceinline__ uint32_t groupIdx(uint32_t pos) {  // g_groupIdx, rom.cpp:21-25 (pos < 32)
  return pos < 4 ? pos : (2u * (31 - __clz(pos)) + ((pos >> (30 - __clz(pos))) & 1u));
}
__device__ __forceinline__ uint32_t goRicePar(int sum) {  // g_goRiceParsCoeff[clip(sum,0,31)], rom.cpp:27-29
  return (sum >= 7) + (sum >= 14) + (sum >= 28);
}

// Per-TU constants (CoeffCodingContext ctor, context_modelling.cpp:7-79).
struct TuGeom {
  int log2w, log2h, w, h;     // coded block
  int tw, th, log2tw;         // tile = min(32, w) x min(32, h) (JVET_C0024_ZERO_OUT_TH)
  int log2cgw, log2cgh, log2cg, shape;
  int lwig, lhig, wig, hig, numCG;
  int ch, comp;
  int lastOffsetX, lastOffsetY, lastShiftX, lastShiftY, maxLastX, maxLastY;
  bool sbtZero;
};

__device__ inline TuGeom makeGeom(const cabac_tu_desc &d) {
  TuGeom g;
  g.log2w = d.log2_w;
  g.log2h = d.log2_h;
  g.w = 1 << g.log2w;
  g.h = 1 << g.log2h;
  g.tw = min(32, g.w);
  g.th = min(32, g.h);
  g.log2tw = min(5, g.log2w);
  // g_log2SbbSize, rom.cpp:41-50, for 2 <= w, h
  if (g.log2w >= 2 && g.log2h >= 2) {
    g.log2cgw = 2; g.log2cgh = 2; g.shape = 0;
  } else if (g.log2w == 1) {
    if (g.log2h <= 2) { g.log2cgw = 1; g.log2cgh = 1; g.shape = 1; }
    else              { g.log2cgw = 1; g.log2cgh = 3; g.shape = 2; }
  } else {  // log2h == 1, log2w >= 2
    if (g.log2w == 2) { g.log2cgw = 1; g.log2cgh = 1; g.shape = 1; }
    else              { g.log2cgw = 3; g.log2cgh = 1; g.shape = 3; }
  }
  g.log2cg = g.log2cgw + g.log2cgh;
  g.lwig = min(5, g.log2w) - g.log2cgw;
  g.lhig = min(5, g.log2h) - g.log2cgh;
  g.wig = 1 << g.lwig;
  g.hig = 1 << g.lhig;
  g.numCG = g.wig * g.hig;
  g.comp = d.comp_id;
  g.ch = d.comp_id ? 1 : 0;
  g.sbtZero = (d.flags & CABAC_TU_SBT_ZEROOUT) && g.w <= 32 && g.h <= 32 && g.comp == 0;
  g.maxLastX = groupIdx(g.tw - 1);
  g.maxLastY = groupIdx(g.th - 1);
  if (g.sbtZero) {  // cabac_writer.cpp:2661-2667
    if (g.w == 32) g.maxLastX = 7;
    if (g.h == 32) g.maxLastY = 7;
  }
  if (g.ch) {
    g.lastOffsetX = g.lastOffsetY = 0;
    g.lastShiftX = min(2, g.w >> 3);
    g.lastShiftY = min(2, g.h >> 3);
  } else {
    // prefix_ctx[] = {0,0,0,3,6,10,15,21}, context_modelling.cpp:58
    // i.e. n(n-1)/2 for log2 size n >= 3, else 0
    g.lastOffsetX = g.log2w <= 2 ? 0 : (g.log2w * (g.log2w - 1)) >> 1;
    g.lastOffsetY = g.log2h <= 2 ? 0 : (g.log2h * (g.log2h - 1)) >> 1;
    g.lastShiftX = (g.log2w + 1) >> 2;
    g.lastShiftY = (g.log2h + 1) >> 2;
  }
  return g;
}

// regBinLimit, cabac_writer.cpp:2487-2491 + unit.cpp:465-478
__device__ __forceinline__ int regBinLimit(const TuGeom &g) {
  int zw = g.tw, zh = g.th;
  if (g.sbtZero) {
    if (g.w == 32) zw = 16;
    if (g.h == 32) zh = 16;
  }
  return (zw * zh * 28) >> 4;
}

// scan position -> tile coordinates
__device__ __forceinline__ void scanXY(const TuGeom &g, int scanPos, int &x, int &y) {
  const int cg = scanPos >> g.log2cg, p = scanPos & ((1 << g.log2cg) - 1);
  const uint32_t c = __ldg(&g_rom.cgScan[g.lwig][g.lhig][cg]);
  const uint32_t q = __ldg(&g_rom.inScan[g.shape][p]);
  x = int((c & 15u) << g.log2cgw) + int(q & 15u);
  y = int((c >> 4) << g.log2cgh) + int(q >> 4);
}
__device__ __forceinline__ int xyToScan(const TuGeom &g, int x, int y) {
  const int cgx = x >> g.log2cgw, cgy = y >> g.log2cgh;
  const int px = x & ((1 << g.log2cgw) - 1), py = y & ((1 << g.log2cgh) - 1);
  const int cg = __ldg(&g_rom.cgScanInv[g.lwig][g.lhig][cgy * g.wig + cgx]);
  const int p = __ldg(&g_rom.inScanInv[g.shape][(py << g.log2cgw) + px]);
  return (cg << g.log2cg) + p;
}

// DepQuant state machine.  stateTab 32040 (cabac_writer.cpp:2484, 2808) written
// on the two state bits (hi, lo): one step with parity p maps (hi, lo) to
// (lo ^ p, hi); with stateTab == 0 the state stays 0.
__device__ __forceinline__ int stateStep(int state, int par) {
  return (((state & 1) ^ par) << 1) | (state >> 1);
}

// ---------------------------------------------------------------------------
// Template sums (sigCtxIdAbs / templateAbsSum, context_modelling.hpp:71-157).
// T is the tile element type (int16 encoder tile, int32 decoder tile).
// ---------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void templateSums(const T *tile, const TuGeom &g, int x, int y, int &sumAbs1,
                                             int &numPos, int &sumAll) {
  sumAbs1 = 0; numPos = 0; sumAll = 0;
  const T *p = tile + (y << g.log2tw) + x;
  auto upd = [&](int v) {
    const int a = abs(v);
    sumAbs1 += min(4 + (a & 1), a);
    numPos += a != 0;
    sumAll += a;
  };
  // neighbours outside the 32x32 tile of a 64-point block are zero-out samples
  const bool r1 = x + 1 < g.tw, r2 = x + 2 < g.tw, b1 = y + 1 < g.th, b2 = y + 2 < g.th;
  if (r1) {
    upd(p[1]);
    if (r2) upd(p[2]);
    if (b1) upd(p[g.tw + 1]);
  }
  if (b1) {
    upd(p[g.tw]);
    if (b2) upd(p[2 * g.tw]);
  }
}

__device__ __forceinline__ int sigCtxOffset(const TuGeom &g, int sumAbs1, int diag) {
  int ofs = min((sumAbs1 + 1) >> 1, 3) + (diag < 2 ? 4 : 0);
  if (g.ch == 0) ofs += diag < 5 ? 4 : 0;
  return ofs;
}
// ctxOffsetAbs, context_modelling.hpp:113-124 (for positions where sigCtxIdAbs ran)
__device__ __forceinline__ int gtxCtxOffset(const TuGeom &g, int sumAbs1, int numPos, int diag) {
  int ofs = min(sumAbs1 - numPos, 4) + 1;
  if (diag == 0)
    ofs += g.ch == 0 ? 15 : 5;
  else if (g.ch == 0)
    ofs += diag < 3 ? 10 : (diag < 10 ? 5 : 0);
  return ofs;
}

// ---------------------------------------------------------------------------
// Encoder: command queue drained by lane 0
// ---------------------------------------------------------------------------
constexpr int QUEUE_WORDS = 192;  // worst case: one transform-skip group (178 words)

__device__ __noinline__ uint32_t drainQueue(Encoder &enc, uint32_t *ctx, const uint32_t *queue, int n) {
  uint32_t status = 0;
  for (int i = 0; i < n; i++) {
    const uint32_t w = queue[i];
    switch (CABAC_CMD_KIND(w)) {
      case CABAC_CMD_CTX: enc.encodeBin((w >> 9) & 1u, ctx + (w & 511u)); break;
      case CABAC_CMD_EP: enc.encodeBinsEP(w & 0xFFFFFFu, (w >> 24) & 31u); break;
      case CABAC_CMD_REM: enc.encodeRemAbsEP(w & 0xFFFFFFu, (w >> 24) & 31u); break;
      case CABAC_CMD_TRM: enc.encodeBinTrm(w & 1u); break;
      case CABAC_CMD_ALIGN: enc.range = 256; break;
      default: status |= CABAC_SS_BAD_CMD; break;
    }
  }
  return status;
}

// last_sig_coeff, cabac_writer.cpp:2639-2722 (lane 0)
__device__ inline void encLastSigCoeff(Encoder &enc, uint32_t *ctx, const TuGeom &g, int posX, int posY) {
  const uint32_t gx = groupIdx(posX), gy = groupIdx(posY);
  uint32_t k;
  const int bx = lastXBase(g.ch) + g.lastOffsetX, by = lastYBase(g.ch) + g.lastOffsetY;
  for (k = 0; k < gx; k++) enc.encodeBin(1, ctx + bx + (k >> g.lastShiftX));
  if (gx < uint32_t(g.maxLastX)) enc.encodeBin(0, ctx + bx + (k >> g.lastShiftX));
  for (k = 0; k < gy; k++) enc.encodeBin(1, ctx + by + (k >> g.lastShiftY));
  if (gy < uint32_t(g.maxLastY)) enc.encodeBin(0, ctx + by + (k >> g.lastShiftY));
  if (gx > 3) {
    const uint32_t n = (gx - 2) >> 1;
    enc.encodeBinsEP((posX - minInGroup(gx)) & ((1u << n) - 1u), n);
  }
  if (gy > 3) {
    const uint32_t n = (gy - 2) >> 1;
    enc.encodeBinsEP((posY - minInGroup(gy)) & ((1u << n) - 1u), n);
  }
}

// TS helpers, context_modelling.hpp:246-385
template <typename T>
__device__ __forceinline__ void neighTS(const T *tile, const TuGeom &g, int x, int y, int &left, int &above) {
  const T *p = tile + (y << g.log2tw) + x;
  left = x > 0 ? int(p[-1]) : 0;
  above = y > 0 ? int(p[-g.tw]) : 0;
}
__device__ __forceinline__ int signCtxTS(int left, int above, bool bdpcm) {
  const int sl = (left > 0) - (left < 0), sa = (above > 0) - (above < 0);
  int c;
  if ((sl == 0 && sa == 0) || sl * sa < 0) c = 0;
  else if (sl >= 0 && sa >= 0) c = 1;
  else c = 2;
  return R_TS_SIGN + c + (bdpcm ? 3 : 0);
}
__device__ __forceinline__ int modCoeffTS(int left, int above, int absCoeff, bool bdpcm) {  // deriveModCoeff
  if (absCoeff == 0) return 0;
  int v = absCoeff;
  if (!bdpcm) {
    const int pred = max(abs(above), abs(left));
    if (v == pred) v = 1;
    else if (v < pred) v = v + 1;
  }
  return v;
}
__device__ __forceinline__ int decModCoeffTS(int left, int above, int absCoeff) {  // decDeriveModCoeff
  if (absCoeff == 0) return 0;
  const int pred = max(abs(above), abs(left));
  if (absCoeff == 1 && pred > 0) return pred;
  return absCoeff - (absCoeff <= pred);
}

// residual_codingTS + residual_coding_subblockTS, cabac_writer.cpp:2874-3062 (lane 0, serial;
// transform-skip blocks are a small share of the units).
__device__ __noinline__ void encResidualTS(Encoder &enc, uint32_t *ctx, const int16_t *tile, const TuGeom &g,
                                           bool bdpcm, const uint16_t *cgMask) {
  int remBins = ((g.w * g.h) * 7) >> 2;
  uint64_t sigGroups = 0;  // indexed by raster CG position
  const int cgSize = 1 << g.log2cg;
  const int lastSubSet = g.numCG - 1;
  for (int s = 0; s < g.numCG; s++) {
    const uint32_t c = __ldg(&g_rom.cgScan[g.lwig][g.lhig][s]);
    const int cgx = c & 15, cgy = c >> 4, cgPos = cgy * g.wig + cgx;
    const bool sig = cgMask[s] != 0;
    if (sig) sigGroups |= 1ull << cgPos;
    const int left = cgx > 0 ? int((sigGroups >> (cgPos - 1)) & 1) : 0;
    const int above = cgy > 0 ? int((sigGroups >> (cgPos - g.wig)) & 1) : 0;
    const uint64_t others = sigGroups & ~(1ull << lastSubSet);
    if (s != lastSubSet || others != 0) {
      enc.encodeBin(sig, ctx + R_TS_SIG_GROUP + left + above);
      if (!sig) continue;
    }
    const int minPos = s << g.log2cg, maxPos = minPos + cgSize - 1;
    int numNonZero = 0, lastPass1 = -1, lastPass2 = -1;
    int sp = minPos;
    for (; sp <= maxPos && remBins >= 4; sp++) {
      int x, y, l, a;
      scanXY(g, sp, x, y);
      const int v = tile[(y << g.log2tw) + x];
      const uint32_t sigf = v != 0;
      neighTS(tile, g, x, y, l, a);
      if (numNonZero || sp != maxPos) {
        enc.encodeBin(sigf, ctx + R_TS_SIG_FLAG + (l != 0) + (a != 0));
        remBins--;
      }
      if (sigf) {
        enc.encodeBin(v < 0, ctx + signCtxTS(l, a, bdpcm));
        remBins--;
        numNonZero++;
        int rem = modCoeffTS(l, a, abs(v), bdpcm) - 1;
        const uint32_t gt1 = rem != 0;
        enc.encodeBin(gt1, ctx + R_TS_LRG1_FLAG + (bdpcm ? 3 : (l != 0) + (a != 0)));
        remBins--;
        if (gt1) {
          rem -= 1;
          enc.encodeBin(uint32_t(rem) & 1u, ctx + R_TS_PAR_FLAG);
          remBins--;
        }
      }
      lastPass1 = sp;
    }
    for (sp = minPos; sp <= maxPos && remBins >= 4; sp++) {
      int x, y, l, a;
      scanXY(g, sp, x, y);
      neighTS(tile, g, x, y, l, a);
      const int absLevel = modCoeffTS(l, a, abs(int(tile[(y << g.log2tw) + x])), bdpcm);
      for (int cutoff = 2; cutoff <= 8; cutoff += 2) {
        if (absLevel >= cutoff) {
          enc.encodeBin(absLevel >= cutoff + 2, ctx + R_TS_GTX_FLAG + (cutoff >> 1));
          remBins--;
        }
      }
      lastPass2 = sp;
    }
    for (sp = minPos; sp <= maxPos; sp++) {
      int x, y, l, a;
      scanXY(g, sp, x, y);
      neighTS(tile, g, x, y, l, a);
      const int cutoff = sp <= lastPass2 ? 10 : (sp <= lastPass1 ? 2 : 0);
      const int v = tile[(y << g.log2tw) + x];
      const int absLevel = modCoeffTS(l, a, abs(v), bdpcm || !cutoff);
      if (absLevel >= cutoff) {
        const uint32_t rem = sp <= lastPass1 ? uint32_t(absLevel - cutoff) >> 1 : uint32_t(absLevel);
        enc.encodeRemAbsEP(rem, 1);
        if (absLevel && sp > lastPass1) enc.encodeBinsEP(v < 0, 1);
      }
    }
  }
}

// ---------------------------------------------------------------------------
// Decoder (lane 0): cabac_reader.cpp:2865-3339 on an int32 shared-memory tile.
// ---------------------------------------------------------------------------
__device__ inline int decLastSigCoeff(Decoder &dec, uint32_t *ctx, const TuGeom &g) {
  uint32_t posX = 0, posY = 0;
  const int bx = lastXBase(g.ch) + g.lastOffsetX, by = lastYBase(g.ch) + g.lastOffsetY;
  for (; posX < uint32_t(g.maxLastX); posX++)
    if (!dec.decodeBin(ctx + bx + (posX >> g.lastShiftX))) break;
  for (; posY < uint32_t(g.maxLastY); posY++)
    if (!dec.decodeBin(ctx + by + (posY >> g.lastShiftY))) break;
  if (posX > 3) {
    const uint32_t n = (posX - 2) >> 1;
    uint32_t t = 0;
    for (int i = int(n) - 1; i >= 0; i--) t += dec.decodeBinEP() << i;
    posX = minInGroup(posX) + t;
  }
  if (posY > 3) {
    const uint32_t n = (posY - 2) >> 1;
    uint32_t t = 0;
    for (int i = int(n) - 1; i >= 0; i--) t += dec.decodeBinEP() << i;
    posY = minInGroup(posY) + t;
  }
  return xyToScan(g, int(posX), int(posY));
}

__device__ __forceinline__ void checkConformance(Decoder &dec, int v) {  // cabac_reader.cpp:2942-2946
  if (v < -32768 || v > 32767) dec.status |= CABAC_SS_COEFF_RANGE;
}

// residual_coding regular path, cabac_reader.cpp:2670-2735 + 2948-3128.
// Returns CABAC_TUR_* flags.
__device__ __noinline__ uint32_t decResidualRegular(Decoder &dec, uint32_t *ctx, int32_t *tile, const TuGeom &g,
                                                    uint32_t subFlags, bool ts) {
  uint32_t res = 0;
  const int scanPosLast = decLastSigCoeff(dec, ctx, g);
  if (!ts && g.w >= 4 && g.h >= 4) {
    const int maxLfnstPos = ((g.h == 4 && g.w == 4) || (g.h == 8 && g.w == 8)) ? 7 : 15;
    if (scanPosLast > maxLfnstPos) res |= CABAC_TUR_VIOLATES_LFNST;
    if (scanPosLast >= 1) res |= CABAC_TUR_LFNST_LAST_POS;
  }
  if (g.comp == 0 && !ts && scanPosLast >= 1) res |= CABAC_TUR_MTS_LAST_POS;
  const bool depQuant = (subFlags & CABAC_SUB_DEP_QUANT) != 0;
  const bool signHiding = (subFlags & CABAC_SUB_SIGN_HIDING) != 0;
  int state = 0;
  int remRegBins = regBinLimit(g);
  uint64_t sigGroups = 0;
  const int cgSize = 1 << g.log2cg;
  for (int s = scanPosLast >> g.log2cg; s >= 0; s--) {
    const uint32_t c = __ldg(&g_rom.cgScan[g.lwig][g.lhig][s]);
    const int cgx = c & 15, cgy = c >> 4, cgPos = cgy * g.wig + cgx;
    if (g.sbtZero) {
      if ((g.h == 32 && cgy >= (16 >> g.log2cgh)) || (g.w == 32 && cgx >= (16 >> g.log2cgw))) continue;
    }
    const int minSubPos = s << g.log2cg;
    const bool isLast = (scanPosLast >> g.log2cg) == s;
    const int firstSigPos = isLast ? scanPosLast : minSubPos + cgSize - 1;
    bool sigGroup = isLast || s == 0;
    if (!sigGroup) {
      const int right = (cgx + 1 < g.wig) ? int((sigGroups >> (cgPos + 1)) & 1) : 0;
      const int lower = (cgy + 1 < g.hig) ? int((sigGroups >> (cgPos + g.wig)) & 1) : 0;
      sigGroup = dec.decodeBin(ctx + R_SIG_GROUP + 2 * g.ch + (right | lower));
    }
    if (!sigGroup) continue;
    sigGroups |= 1ull << cgPos;
    if (g.comp == 0 && (cgy > 3 || cgx > 3)) res |= CABAC_TUR_VIOLATES_MTS;

    const int inferSigPos = firstSigPos != scanPosLast ? (s != 0 ? minSubPos : -1) : firstSigPos;
    int firstNZPos = firstSigPos, lastNZPos = -1, numNonZero = 0;
    uint16_t sigIdx[16];  // tile index of every nonzero, in coding order
    int nextSigPos = firstSigPos;
    for (; nextSigPos >= minSubPos && remRegBins >= 4; nextSigPos--) {
      int x, y, s1, np, sa;
      scanXY(g, nextSigPos, x, y);
      const int ti = (y << g.log2tw) + x;
      bool sig = (!numNonZero && nextSigPos == inferSigPos);
      int ctxOff = 0;
      if (!sig || nextSigPos != scanPosLast) {
        templateSums(tile, g, x, y, s1, np, sa);
        ctxOff = gtxCtxOffset(g, s1, np, x + y);
      }
      if (!sig) {
        sig = dec.decodeBin(ctx + sigFlagBase(g.ch, max(0, state - 1)) + sigCtxOffset(g, s1, x + y));
        remRegBins--;
      }
      int val = 0;
      if (sig) {
        sigIdx[numNonZero++] = uint16_t(ti);
        firstNZPos = nextSigPos;
        lastNZPos = max(lastNZPos, nextSigPos);
        const uint32_t gt1 = dec.decodeBin(ctx + gtxFlagBase(g.ch + 2) + ctxOff);
        remRegBins--;
        uint32_t par = 0, gt2 = 0;
        if (gt1) {
          par = dec.decodeBin(ctx + parFlagBase(g.ch) + ctxOff);
          remRegBins--;
          gt2 = dec.decodeBin(ctx + gtxFlagBase(g.ch) + ctxOff);
          remRegBins--;
        }
        val = int(1 + par + gt1 + (gt2 << 1));
        tile[ti] = val;
      }
      if (depQuant) state = stateStep(state, val & 1);
    }
    const int firstPosMode2 = nextSigPos;
    // 2nd pass: remainders of the context-coded positions
    for (int sp = firstSigPos; sp > firstPosMode2; sp--) {
      int x, y, s1, np, sa;
      scanXY(g, sp, x, y);
      const int ti = (y << g.log2tw) + x;
      const int t = tile[ti];
      if (t >= 4) {
        templateSums(tile, g, x, y, s1, np, sa);
        tile[ti] = t + int(dec.decodeRemAbsEP(goRicePar(sa - 20)) << 1);
      }
    }
    // bypass-only positions
    for (int sp = firstPosMode2; sp >= minSubPos; sp--) {
      int x, y, s1, np, sa;
      scanXY(g, sp, x, y);
      templateSums(tile, g, x, y, s1, np, sa);
      const uint32_t rice = goRicePar(sa);
      const uint32_t pos0 = (state < 2 ? 1u : 2u) << rice;
      const uint32_t rem = dec.decodeRemAbsEP(rice);
      const int t = int(rem == pos0 ? 0 : (rem < pos0 ? rem + 1 : rem));
      if (depQuant) state = stateStep(state, t & 1);
      if (t) {
        const int ti = (y << g.log2tw) + x;
        sigIdx[numNonZero++] = uint16_t(ti);
        firstNZPos = sp;
        lastNZPos = max(lastNZPos, sp);
        tile[ti] = t;
      }
    }
    const uint32_t numSigns = (signHiding && lastNZPos - firstNZPos >= 4) ? numNonZero - 1 : numNonZero;
    uint32_t signPattern = dec.decodeBinsEP(numSigns);
    signPattern = numSigns ? signPattern << (32 - numSigns) : 0;
    int sumAbs = 0;
    for (uint32_t k = 0; k < numSigns; k++) {
      const int a = tile[sigIdx[k]];
      sumAbs += a;
      const int v = (signPattern & 0x80000000u) ? -a : a;
      tile[sigIdx[k]] = v;
      signPattern <<= 1;
      checkConformance(dec, v);
    }
    if (uint32_t(numNonZero) > numSigns) {
      const int a = tile[sigIdx[numSigns]];
      sumAbs += a;
      const int v = (sumAbs & 1) ? -a : a;
      tile[sigIdx[numSigns]] = v;
      checkConformance(dec, v);
    }
  }
  return res;
}

// residual_codingTS + subblockTS, cabac_reader.cpp:3130-3339
__device__ __noinline__ void decResidualTS(Decoder &dec, uint32_t *ctx, int32_t *tile, const TuGeom &g, bool bdpcm) {
  int remBins = ((g.w * g.h) * 7) >> 2;
  uint64_t sigGroups = 0;
  const int cgSize = 1 << g.log2cg;
  const int lastSubSet = g.numCG - 1;
  for (int s = 0; s < g.numCG; s++) {
    const uint32_t c = __ldg(&g_rom.cgScan[g.lwig][g.lhig][s]);
    const int cgx = c & 15, cgy = c >> 4, cgPos = cgy * g.wig + cgx;
    bool sigGroup = (s == lastSubSet) && sigGroups == 0;
    if (!sigGroup) {
      const int left = cgx > 0 ? int((sigGroups >> (cgPos - 1)) & 1) : 0;
      const int above = cgy > 0 ? int((sigGroups >> (cgPos - g.wig)) & 1) : 0;
      sigGroup = dec.decodeBin(ctx + R_TS_SIG_GROUP + left + above);
    }
    if (!sigGroup) continue;
    sigGroups |= 1ull << cgPos;
    const int minPos = s << g.log2cg, maxPos = minPos + cgSize - 1;
    int numNonZero = 0, lastPass1 = -1, lastPass2 = -1;
    uint32_t signPattern = 0;
    uint16_t sigIdx[16];
    int sp = minPos;
    for (; sp <= maxPos && remBins >= 4; sp++) {
      int x, y, l, a;
      scanXY(g, sp, x, y);
      const int ti = (y << g.log2tw) + x;
      neighTS(tile, g, x, y, l, a);
      bool sig = (!numNonZero && sp == maxPos);
      if (!sig) {
        sig = dec.decodeBin(ctx + R_TS_SIG_FLAG + (l != 0) + (a != 0));
        remBins--;
      }
      if (sig) {
        const uint32_t sign = dec.decodeBin(ctx + signCtxTS(l, a, bdpcm));
        remBins--;
        signPattern += sign << numNonZero;
        sigIdx[numNonZero++] = uint16_t(ti);
        const uint32_t gt1 = dec.decodeBin(ctx + R_TS_LRG1_FLAG + (bdpcm ? 3 : (l != 0) + (a != 0)));
        remBins--;
        uint32_t par = 0;
        if (gt1) {
          par = dec.decodeBin(ctx + R_TS_PAR_FLAG);
          remBins--;
        }
        tile[ti] = (sign ? -1 : 1) * int(1 + par + gt1);
      }
      lastPass1 = sp;
    }
    for (sp = minPos; sp <= maxPos && remBins >= 4; sp++) {
      int x, y;
      scanXY(g, sp, x, y);
      const int ti = (y << g.log2tw) + x;
      int t = abs(tile[ti]);
      for (int cutoff = 2; cutoff <= 8; cutoff += 2) {
        if (t >= cutoff) {
          t += int(dec.decodeBin(ctx + R_TS_GTX_FLAG + (cutoff >> 1)) << 1);
          remBins--;
        }
      }
      tile[ti] = t;
      lastPass2 = sp;
    }
    for (sp = minPos; sp <= maxPos; sp++) {
      int x, y;
      scanXY(g, sp, x, y);
      const int ti = (y << g.log2tw) + x;
      const int cutoff = sp <= lastPass2 ? 10 : (sp <= lastPass1 ? 2 : 0);
      int t = abs(tile[ti]);
      if (t >= cutoff) {
        const int rem = int(dec.decodeRemAbsEP(1));
        t += (sp <= lastPass1) ? (rem << 1) : rem;
        if (t && sp > lastPass1) {
          const uint32_t sign = dec.decodeBinEP();
          signPattern += sign << numNonZero;
          sigIdx[numNonZero++] = uint16_t(ti);
        }
      }
      if (!bdpcm && cutoff && t > 0) {
        int l, a;
        neighTS(tile, g, x, y, l, a);
        t = decModCoeffTS(l, a, t);
      }
      tile[ti] = t;
    }
    for (int k = 0; k < numNonZero; k++) {
      const int a = tile[sigIdx[k]];
      const int v = (signPattern & 1u) ? -a : a;
      tile[sigIdx[k]] = v;
      signPattern >>= 1;
      checkConformance(dec, v);
    }
  }
}

}  // namespace cabac
