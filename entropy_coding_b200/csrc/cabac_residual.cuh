// cabac_residual.cuh - residual_coding binarisation and context selection.
//
// Geometry, scan ROM, context layout and template arithmetic shared by the residual
// encoder (binariseCG and the producers in cabac_kernels.cu) and the lockstep decoder
// (cabac_lockstep.cuh).
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
  uint8_t inScan[25][16];       // [shape][p] -> px | py << 4                  (within one CG)
  uint8_t inScanInv[25][16];    // [shape][py * gw + px] -> p
};
// shape = 5 * log2 CG width + log2 CG height (g_log2SbbSize, rom.cpp:41-50): 4x4 for blocks with both
// sides >= 4; 2x2, 2x8 / 8x2 for 2-wide / 2-high blocks; 1x16 / 16x1 (and the smaller 1xN / Nx1) for the
// one-sample-wide sub-partitions intra sub-partitioning (ISP) produces

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

// A descriptor the reference would CHECK() away (a block side above 64 samples, a component above Cr) must not
// turn into wild shared-memory indices: every kernel reads descriptors through loadDesc, which clamps the geometry
// into range; the coding kernels report CABAC_SS_BAD_DESC for the substream (its output is then meaningless).
__device__ __forceinline__ bool descBad(const cabac_tu_desc &d) {
  return d.log2_w > 6 || d.log2_h > 6 || d.comp_id > 2;
}
__device__ __forceinline__ cabac_tu_desc loadDesc(const cabac_tu_desc *__restrict__ p) {
  cabac_tu_desc d = *p;
  if (descBad(d)) {
    d.log2_w = d.log2_w > 6 ? 6 : d.log2_w;
    d.log2_h = d.log2_h > 6 ? 6 : d.log2_h;
    d.comp_id = d.comp_id > 2 ? 2 : d.comp_id;
    d.flags |= 0x80;  // CABAC_TU_* uses bits 0-4; bit 7 marks the clamped copy inside the kernels
  }
  return d;
}
constexpr uint8_t DESC_CLAMPED = 0x80;

__device__ inline TuGeom makeGeom(const cabac_tu_desc &d) {
  TuGeom g;
  g.log2w = d.log2_w;
  g.log2h = d.log2_h;
  g.w = 1 << g.log2w;
  g.h = 1 << g.log2h;
  g.tw = min(32, g.w);
  g.th = min(32, g.h);
  g.log2tw = min(5, g.log2w);
  // g_log2SbbSize, rom.cpp:41-50
  if (g.log2w >= 2 && g.log2h >= 2) {
    g.log2cgw = 2; g.log2cgh = 2;
  } else if (g.log2w == 0) {
    g.log2cgw = 0; g.log2cgh = min(g.log2h, 4);
  } else if (g.log2h == 0) {
    g.log2cgw = min(g.log2w, 4); g.log2cgh = 0;
  } else if (g.log2w == 1) {
    g.log2cgw = 1; g.log2cgh = g.log2h <= 2 ? 1 : 3;
  } else {  // log2h == 1, log2w >= 2
    g.log2cgw = g.log2w == 2 ? 1 : 3; g.log2cgh = 1;
  }
  g.shape = 5 * g.log2cgw + g.log2cgh;
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

constexpr int QUEUE_WORDS = 192;  // worst case: one transform-skip group (178 words)

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

}  // namespace cabac
