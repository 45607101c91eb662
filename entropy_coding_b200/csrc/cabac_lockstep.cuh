// cabac_lockstep.cuh - warp-lockstep CABAC decoder for sm_100a.
//
// Decoding is a serial recurrence whose speed is set by the dependent-instruction
// chain per bin and by the instruction-cache footprint, not by lane count.  This
// decoder therefore runs UNIFORMLY on all 32 lanes of the substream's warp (every
// lane holds the same arithmetic-decoder state; shared-memory and bitstream
// accesses are broadcasts) and uses the lanes only where the syntax is parallel:
//   * lane p owns scan position p of the current coefficient group: its template
//     sums (sigCtxIdAbs / templateAbsSum, context_modelling.hpp:71-157) live in its
//     registers and are updated with two predicated adds per decoded level instead
//     of five shared-memory reads per position;
//   * bypass bins are decoded many at a time: n EP bins are the quotient of one
//     division of the offset by the range (the bit-serial loop of
//     arith_codec.cpp:100-151 IS long division), done with an exact 34-bit
//     reciprocal from a 255-entry table; a Golomb-Rice word (prefix, suffix) is one
//     16-bin peek + one commit.
//
// State representation (equivalent to BinDecoderBase, arith_codec.hpp:253-256):
//   value64 = hi:lo, hi = m_Value (16 bits) << 16 | next 16 stream bits, lo = more
//   stream bits; `avail` = valid look-ahead bits below m_Value (>= 32 between ops).
//   scaledRange = range << 23 in this scale (VTM: range << 7).
//   getNumBitsRead (arith_codec.cpp:201-203) = 8 + total renormalisation shift.
#pragma once
#include "cabac_residual.cuh"

namespace cabac {

// ceil(2^34 / r) for r = 256..510: floor(A / r) == (A * M) >> 34 for all A < 2^25.
__device__ __constant__ uint32_t c_recip[256];
// the same table in global memory for the thread-per-substream decoder (a per-lane index is slow in the constant bank)
__device__ uint32_t g_recipTab[256];

// Context entry: everything a bin needs in one 16-byte shared-memory word.
//   x = state1 (bits 0..14) | state0 << 16        BinProbModel_Std::m_state, contexts.hpp:184-190
//   y = (0x7FFF >> r1) & MASK_1 | ((0x7FFF >> r0) & MASK_0) << 16   the "bin == 1" increment
//   z = r0, w = r1                                  setLog2WindowSize, contexts.cpp:915-920
__host__ __device__ inline uint4 ctx_entry(uint32_t s0, uint32_t s1, uint32_t win) {
  const uint32_t r0 = 2u + ((win >> 2) & 3u), r1 = 3u + r0 + (win & 3u);
  uint4 e;
  e.x = (s1 & MASK_1) | ((s0 & MASK_0) << 16);
  e.y = ((0x7FFFu >> r1) & MASK_1) | (((0x7FFFu >> r0) & MASK_0) << 16);
  e.z = r0;
  e.w = r1;
  return e;
}
__host__ __device__ inline uint4 ctx_entry_init(int qp, uint32_t initVal, uint32_t win) {  // contexts.cpp:893-901
  const int slope = int(initVal >> 3) - 4;
  const int offset = int(initVal & 7) * 18 + 1;
  int s = ((slope * (qp - 16)) >> 1) + offset;
  s = s < 1 ? 1 : (s > 127 ? 127 : s);
  return ctx_entry(uint32_t(s) << 8, uint32_t(s) << 8, win);
}

// Two stream bytes, zero beyond the end; by value so that no decoder state goes through memory.
__device__ __noinline__ uint32_t lsFetchHalf(const uint8_t *in, uint32_t nbytes, uint32_t pos) {
  if (pos + 2u <= nbytes) return (uint32_t(__ldg(in + pos)) << 8) | uint32_t(__ldg(in + pos + 1));
  return pos < nbytes ? uint32_t(__ldg(in + pos)) << 8 : 0u;
}

struct LsDecoder {
  const uint8_t *in;
  uint32_t nbytes;
  uint32_t hi, lo;
  int avail;
  uint32_t pos;       // next byte to prefetch
  uint32_t nextHalf;  // bytes pos-2, pos-1, not yet merged
  uint32_t range;
  uint32_t recip, recipRange;
  uint32_t status, nbins;

  __device__ __forceinline__ uint32_t byteAt(uint32_t i) const { return i < nbytes ? uint32_t(__ldg(in + i)) : 0u; }
  __device__ __forceinline__ void prefetch() {
    nextHalf = lsFetchHalf(in, nbytes, pos);
    pos += 2;
  }
  // start(), arith_codec.cpp:60-66
  __device__ __forceinline__ void start(const uint8_t *p, uint32_t n) {
    in = p;
    nbytes = n;
    status = 0;
    nbins = 0;
    range = 510;
    recipRange = 0;
    recip = 0;
    hi = (byteAt(0) << 24) | (byteAt(1) << 16) | (byteAt(2) << 8) | byteAt(3);
    lo = (byteAt(4) << 24) | (byteAt(5) << 16);
    avail = 32;
    pos = 6;
    prefetch();
  }
  __device__ __forceinline__ void refill() {
    if (avail < 32) {
      lo |= nextHalf << (32 - avail);
      avail += 16;
      prefetch();
    }
  }
  __device__ __forceinline__ void shiftIn(int n) {  // value64 <<= n
    hi = __funnelshift_l(lo, hi, n);
    lo <<= n;
    avail -= n;
  }
  // 8 * bytes read + m_bitsNeeded of the reference decoder
  __device__ __forceinline__ uint32_t bitsRead() const { return 8u * (pos - 2u) - 8u - uint32_t(avail); }
  // bytes the reference would have pulled from the InputBitstream (readByte, bit_stream.cpp:268-274)
  __device__ __forceinline__ void checkOverrun() {
    const uint32_t shift = bitsRead() - 8u;
    if (2u + (shift >> 3) > nbytes) status |= CABAC_SS_IN_OVERRUN;
  }

  // TBinDecoder::decodeBin, arith_codec.cpp:242-277, model update contexts.cpp:903-913
  __device__ __forceinline__ uint32_t decodeBin(uint4 *m) {
    const uint4 e = *m;
    const uint32_t q = ((e.x >> 16) + (e.x & 0xFFFFu)) >> 8;
    const uint32_t mps = q >> 7;
    const uint32_t qq = q ^ (0xFFu & (0u - mps));
    const uint32_t lps = (((qq >> 2) * (range >> 5)) >> 1) + 4;
    nbins++;
    const uint32_t rmps = range - lps;
    const uint32_t sr = rmps << 23;
    const bool isLps = hi >= sr;
    const uint32_t bin = mps ^ uint32_t(isLps);
    // update
    const uint32_t d0 = (e.x >> e.z) & (MASK_0 << 16);
    const uint32_t d1 = ((e.x & 0xFFFFu) >> e.w) & MASK_1;
    m->x = e.x - d0 - d1 + (bin ? e.y : 0u);
    // (an early return for "MPS, no renormalisation" - most bins - was measured: 35.6 vs 35.3 ms, the
    // uniform branch costs what the skipped shifts save)
    // one normalisation rule for both paths: rmps >= 128, so clz gives 0 / 1 for the MPS path and
    // the renormTab value (contexts.cpp:787-789) for the LPS path
    const uint32_t nr = isLps ? lps : rmps;
    const int n = __clz(int(nr | 7u)) - 23;
    hi -= isLps ? sr : 0u;
    range = nr << n;
    shiftIn(n);
    refill();
    return bin;
  }

  // n EP bins (1..16) as one quotient, no state change.  decodeBinsEP, arith_codec.cpp:116-151
  __device__ __forceinline__ uint32_t epPeek(int n) {
    if (recipRange != range) {
      recip = c_recip[range - 256u];
      recipRange = range;
    }
    return __umulhi(hi >> (23 - n), recip) >> 2;
  }
  __device__ __forceinline__ void epCommit(int n, uint32_t q) {
    const uint32_t top = __funnelshift_l(lo, hi, n);
    hi = top - ((q * range) << 23);
    lo <<= n;
    avail -= n;
    nbins += uint32_t(n);
    refill();
  }
  __device__ __forceinline__ uint32_t ep(int n) {  // n <= 16
    if (n == 0) return 0;
    const uint32_t q = epPeek(n);
    epCommit(n, q);
    return q;
  }
  __device__ __forceinline__ uint32_t epLong(uint32_t n) {  // any n <= 32, MSB-first like decodeBinsEP
    uint32_t v = 0;
    while (n > 16) {
      v = (v << 16) | ep(16);
      n -= 16;
    }
    return (v << n) | ep(int(n));
  }

  // decodeRemAbsEP(rice, cutoff = 5, maxLog2TrDynamicRange = 15), arith_codec.cpp:153-179.
  // Fast path: the whole word (unary prefix, terminator, suffix) lies inside one 16-bin window.
  __device__ __forceinline__ uint32_t remAbs(uint32_t rice);

  // decodeBinTrm, arith_codec.cpp:181-197
  __device__ __forceinline__ uint32_t decodeBinTrm() {
    nbins++;
    range -= 2;
    if (hi >= (range << 23)) return 1;
    if (range < 256u) {
      range <<= 1;
      shiftIn(1);
      refill();
    }
    return 0;
  }
};

// Rare: words longer than 16 bins (large levels).  Out of line; the state round-trips through
// local memory here only.
__device__ __noinline__ uint32_t lsRemAbsSlow(LsDecoder &d, uint32_t rice, uint32_t q16, uint32_t ones) {
  uint32_t prefix = ones, len;
  if (ones < 16u) {
    d.epCommit(int(ones + 1u), q16 >> (15u - ones));
    len = ones < 5u ? rice : ones - 5u + rice;
  } else {  // 16 ones: the prefix is 16 (a zero follows) or 17 (the limit, no terminator)
    d.epCommit(16, 0xFFFFu);
    prefix = 16u + d.ep(1);
    len = prefix == 17u ? 15u : prefix - 5u + rice;
  }
  const uint32_t base = prefix < 5u ? prefix << rice : ((1u << (prefix - 5u)) + 4u) << rice;
  return base + d.epLong(len);
}
__device__ __forceinline__ uint32_t LsDecoder::remAbs(uint32_t rice) {
  const uint32_t q16 = epPeek(16);
  const uint32_t ones = uint32_t(__clz(int(~(q16 << 16))));  // leading ones of the 16-bin window
  const bool esc = ones >= 5u;
  const uint32_t len = rice + (esc ? ones - 5u : 0u);
  const uint32_t n = ones + 1u + len;
  if (n > 16u) {  // copy in / copy out: the decoder's own address never escapes (it stays in registers)
    LsDecoder tmp = *this;
    const uint32_t r = lsRemAbsSlow(tmp, rice, q16, ones);
    *this = tmp;
    return r;
  }
  const uint32_t bins = q16 >> (16u - n);
  epCommit(int(n), bins);
  const uint32_t base = esc ? (1u << (ones - 5u)) + 4u : ones;
  return (base << rice) + (bins & ((1u << len) - 1u));
}
#define lsDecodeBin(d, m) ((d).decodeBin(m))
#define lsRemAbs(d, rice) ((d).remAbs(rice))

// ---------------------------------------------------------------------------
// last_sig_coeff, cabac_reader.cpp:2865-2940
// ---------------------------------------------------------------------------
__device__ __forceinline__ int lsLastSigCoeff(LsDecoder &dec, uint4 *ctx, const TuGeom &g) {
  uint32_t posXY[2];
#pragma unroll 1
  for (int dim = 0; dim < 2; dim++) {  // one decodeBin site for both prefixes
    const int base = dim ? lastYBase(g.ch) + g.lastOffsetY : lastXBase(g.ch) + g.lastOffsetX;
    const int shift = dim ? g.lastShiftY : g.lastShiftX;
    const uint32_t maxv = uint32_t(dim ? g.maxLastY : g.maxLastX);
    uint32_t v = 0;
    while (v < maxv && dec.decodeBin(ctx + base + (v >> shift))) v++;
    if (dim) posXY[1] = v; else posXY[0] = v;
  }
  uint32_t posX = posXY[0], posY = posXY[1];
  if (posX > 3) posX = minInGroup(posX) + dec.ep(int((posX - 2) >> 1));
  if (posY > 3) posY = minInGroup(posY) + dec.ep(int((posY - 2) >> 1));
  return xyToScan(g, int(posX), int(posY));
}

// Template state of the scan position a lane owns -> the context offsets / Rice
// parameters that position will need (R3-R5 of SURVEY.md section 8a).
//   bits 0..3 sig offset, 4..8 gtx offset.  (The Rice parameters of the 2nd / bypass pass come from
//   the owner lane's absolute sum directly: one shuffle + one goRicePar per remainder.)
__device__ __forceinline__ uint32_t lanePack(int s1, int np, int sigDiag, int gtxDiag) {
  const uint32_t sig = uint32_t(min((s1 + 1) >> 1, 3) + sigDiag);
  const uint32_t gtx = uint32_t(min(s1 - np, 4) + 1 + gtxDiag);
  return sig | (gtx << 4);
}

// ---------------------------------------------------------------------------
// Regular residual, cabac_reader.cpp:2670-2735 + residual_coding_subblock :2948-3128.
// All lanes call; returns CABAC_TUR_* flags.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t lsResidualRegular(LsDecoder &dec, uint4 *ctx, int32_t *tile, const TuGeom &g,
                                                   uint32_t subFlags, bool ts, int lane) {
  uint32_t res = 0;
  const int scanPosLast = lsLastSigCoeff(dec, ctx, g);
  if (!ts && g.w >= 4 && g.h >= 4) {
    const int maxLfnstPos = ((g.h == 4 && g.w == 4) || (g.h == 8 && g.w == 8)) ? 7 : 15;
    if (scanPosLast > maxLfnstPos) res |= CABAC_TUR_VIOLATES_LFNST;
    if (scanPosLast >= 1) res |= CABAC_TUR_LFNST_LAST_POS;
  }
  if (g.comp == 0 && !ts && scanPosLast >= 1) res |= CABAC_TUR_MTS_LAST_POS;
  const bool depQuant = (subFlags & CABAC_SUB_DEP_QUANT) != 0;
  const bool signHiding = (subFlags & CABAC_SUB_SIGN_HIDING) != 0;

  // ---- per-lane constants of this block: the in-group position the lane owns ----
  const int cgSize = 1 << g.log2cg, cgw = 1 << g.log2cgw, cgh = 1 << g.log2cgh;
  const bool laneActive = lane < cgSize;
  const uint32_t qxy = __ldg(&g_rom.inScan[g.shape][lane & 15]);
  const int px = int(qxy & 15u), py = int(qxy >> 4);
  uint32_t tmask = 0;  // group positions inside this lane's template
  {
    const int dx[5] = {1, 2, 1, 0, 0}, dy[5] = {0, 0, 1, 1, 2};
#pragma unroll
    for (int k = 0; k < 5; k++) {
      const int nx = px + dx[k], ny = py + dy[k];
      if (laneActive && nx < cgw && ny < cgh) tmask |= 1u << __ldg(&g_rom.inScanInv[g.shape][(ny << g.log2cgw) + nx]);
    }
  }
  // bit p: position p - 1 lies on another anti-diagonal than p (its template may contain p)
  const int myDiag = px + py;
  const int prevDiag = __shfl_up_sync(FULL, myDiag, 1);
  const uint32_t diagEnd = __ballot_sync(FULL, laneActive && (lane == 0 || prevDiag != myDiag));

  int state = 0;
  int remRegBins = regBinLimit(g);
  uint64_t sigGroups = 0;
  const int lastCG = scanPosLast >> g.log2cg;
  // context-set bases of this block, fixed for all its bins: gt1 | par << 10 | gt2 << 20
  const int sigBase0 = sigFlagBase(g.ch, 0);
  const uint32_t gtPack = uint32_t(gtxFlagBase(g.ch + 2)) | (uint32_t(parFlagBase(g.ch)) << 10) |
                          (uint32_t(gtxFlagBase(g.ch)) << 20);
  for (int s = lastCG; s >= 0; s--) {
    const uint32_t c = __ldg(&g_rom.cgScan[g.lwig][g.lhig][s]);
    const int cgx = c & 15, cgy = c >> 4, cgPos = cgy * g.wig + cgx;
    if (g.sbtZero) {
      if ((g.h == 32 && cgy >= (16 >> g.log2cgh)) || (g.w == 32 && cgx >= (16 >> g.log2cgw))) continue;
    }
    const bool isLast = s == lastCG;
    bool sigGroup = isLast || s == 0;
    if (!sigGroup) {
      const int right = (cgx + 1 < g.wig) ? int((sigGroups >> (cgPos + 1)) & 1) : 0;
      const int lower = (cgy + 1 < g.hig) ? int((sigGroups >> (cgPos + g.wig)) & 1) : 0;
      sigGroup = lsDecodeBin(dec, ctx + R_SIG_GROUP + 2 * g.ch + (right | lower)) != 0;
    }
    if (!sigGroup) continue;
    sigGroups |= 1ull << cgPos;
    if (g.comp == 0 && (cgy > 3 || cgx > 3)) res |= CABAC_TUR_VIOLATES_MTS;

    // ---- parallel set-up: template contributions from already decoded groups ----
    const int x = (cgx << g.log2cgw) + px, y = (cgy << g.log2cgh) + py;
    const int ti = (y << g.log2tw) + x, d = x + y;
    const int sigDiag = (d < 2 ? 4 : 0) + ((g.ch == 0 && d < 5) ? 4 : 0);
    const int gtxDiag = d == 0 ? (g.ch == 0 ? 15 : 5) : (g.ch == 0 ? (d < 3 ? 10 : (d < 10 ? 5 : 0)) : 0);
    int s1 = 0, np = 0, sa = 0;
    if (laneActive) {
      const int dx[5] = {1, 2, 1, 0, 0}, dy[5] = {0, 0, 1, 1, 2};
#pragma unroll
      for (int k = 0; k < 5; k++) {
        const int nx = x + dx[k], ny = y + dy[k];
        const bool inGroup = (px + dx[k] < cgw) && (py + dy[k] < cgh);
        if (!inGroup && nx < g.tw && ny < g.th) {
          const int a = abs(tile[(ny << g.log2tw) + nx]);
          s1 += min(4 + (a & 1), a);
          np += a != 0;
          sa += a;
        }
      }
    }
    int myVal = 0;
    uint32_t myInfo = lanePack(s1, np, sigDiag, gtxDiag);
    const int P0 = isLast ? scanPosLast - (s << g.log2cg) : cgSize - 1;
    const int inferSigPos = isLast ? P0 : (s != 0 ? 0 : -1);
    uint32_t nzMask = 0, remMask = 0;

    // ---- 1st pass: context-coded sig / gt1 / par / gt2 ----
    int p = P0;
    uint32_t info = __shfl_sync(FULL, myInfo, p);
    for (; p >= 0 && remRegBins >= 4; p--) {
      const bool boundary = (diagEnd >> p) & 1u;
      uint32_t infoNext = 0;
      if (!boundary) infoNext = __shfl_sync(FULL, myInfo, p - 1);  // overlaps the bins below
      bool sig = nzMask == 0 && p == inferSigPos;
      const int gtxOfs = (isLast && p == P0) ? 0 : int((info >> 4) & 31u);
      if (!sig) {
        sig = dec.decodeBin(ctx + sigBase0 + 20 * max(0, state - 1) + int(info & 15u)) != 0;
        remRegBins--;
      }
      int val = 0;
      if (sig) {
        // gt1, then par and gt2 when gt1 is set: one decodeBin site, three context sets
        uint32_t gtBits = 0;  // bit 0 gt1, bit 1 par, bit 2 gt2
        uint32_t gp = gtPack;
#pragma unroll 1
        for (int k = 0; k < 3; k++) {
          gtBits |= dec.decodeBin(ctx + (gp & 1023u) + gtxOfs) << k;
          gp >>= 10;
          remRegBins--;
          if (!gtBits) break;
        }
        const uint32_t gt2 = gtBits >> 2;
        val = int(1 + (gtBits & 1u) + ((gtBits >> 1) & 1u) + (gt2 << 1));
        nzMask |= 1u << p;
        remMask |= gt2 << p;
        if (lane == p) myVal = val;
        if ((tmask >> p) & 1u) {  // levels <= 5 here: min(4 + (a & 1), a) == a
          s1 += val;
          np += 1;
          sa += val;
        }
      }
      if (depQuant) state = stateStep(state, val & 1);
      if (boundary && p > 0) {
        myInfo = lanePack(s1, np, sigDiag, gtxDiag);
        infoNext = __shfl_sync(FULL, myInfo, p - 1);
      }
      info = infoNext;
    }
    // ---- 2nd pass (remainders of context-coded levels >= 4) and bypass pass in one descending walk ----
    const int cut = p;  // positions <= cut are bypass coded
    uint32_t visit = remMask | (cut >= 0 ? ((2u << cut) - 1u) : 0u);
    while (visit) {
      const int j = 31 - __clz(int(visit));
      visit &= ~(1u << j);
      const int saj = __shfl_sync(FULL, sa, j);  // templateAbsSum of position j, context_modelling.hpp:135-157
      if (j > cut) {
        const int add = int(dec.remAbs(goRicePar(saj - 20)) << 1);
        if (lane == j) myVal += add;
        if ((tmask >> j) & 1u) sa += add;
      } else {
        const uint32_t rice = goRicePar(saj);
        const uint32_t pos0 = (state < 2 ? 1u : 2u) << rice;
        const uint32_t rem = dec.remAbs(rice);
        const int val = int(rem == pos0 ? 0u : (rem < pos0 ? rem + 1u : rem));
        if (depQuant) state = stateStep(state, val & 1);
        if (val) {
          nzMask |= 1u << j;
          if (lane == j) myVal = val;
          if ((tmask >> j) & 1u) sa += val;
        }
      }
    }
    // ---- signs (coding order = descending position; first coded sign is the pattern's MSB) ----
    const int numNZ = __popc(nzMask);
    const bool hide = signHiding && nzMask && ((31 - __clz(int(nzMask))) - (__ffs(int(nzMask)) - 1) >= 4);
    const int numSigns = numNZ - int(hide);
    const uint32_t pattern = dec.ep(numSigns);
    int v = myVal;
    if (myVal) {
      const int rank = __popc(nzMask >> (lane + 1));
      if (rank < numSigns && ((pattern >> (numSigns - 1 - rank)) & 1u)) v = -myVal;
    }
    if (hide) {  // the last coded (lowest) non-zero takes the parity of the group's sum
      const int sumAbs = __reduce_add_sync(FULL, myVal);
      if (lane == __ffs(int(nzMask)) - 1 && (sumAbs & 1)) v = -myVal;
    }
    if (__any_sync(FULL, v < -32768 || v > 32767)) dec.status |= CABAC_SS_COEFF_RANGE;  // cabac_reader.cpp:2942-2946
    if (laneActive && v) tile[ti] = v;
    __syncwarp();
  }
  return res;
}

// residual_codingTS + subblockTS, cabac_reader.cpp:3130-3339
// Lockstep flavour: every lane executes the same serial parse (transform-skip blocks are a
// small share of the bins); tile stores are same-value broadcasts.
__device__ __forceinline__ void lsResidualTS(LsDecoder &dec, uint4 *ctx, int32_t *tile, const TuGeom &g, bool bdpcm,
                                             int lane) {
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
      sigGroup = lsDecodeBin(dec, ctx + R_TS_SIG_GROUP + left + above);
    }
    if (!sigGroup) continue;
    sigGroups |= 1ull << cgPos;
    const int minPos = s << g.log2cg, maxPos = minPos + cgSize - 1;
    // lane k owns scan position minPos + k: tile index | has-left << 16 | has-above << 17, fetched by
    // shuffle in the passes below (keeps the scan-table loads off the per-bin dependent chain)
    uint32_t myPos;
    {
      int x, y;
      scanXY(g, minPos + (lane & (cgSize - 1)), x, y);
      myPos = uint32_t((y << g.log2tw) + x) | (x > 0 ? 1u << 16 : 0u) | (y > 0 ? 1u << 17 : 0u);
    }
#define TS_POS(sp, ti, l, a)                                              \
  const uint32_t pk_ = __shfl_sync(FULL, myPos, (sp) - minPos);           \
  const int ti = int(pk_ & 0xFFFFu);                                      \
  const int l = (pk_ >> 16) & 1u ? tile[ti - 1] : 0;                      \
  const int a = (pk_ >> 17) & 1u ? tile[ti - g.tw] : 0;
    int numNonZero = 0, lastPass1 = -1, lastPass2 = -1;
    uint32_t signMask = 0;  // bit k: position minPos + k is negative (coded signs arrive in ascending position order)
    int sp = minPos;
    for (; sp <= maxPos && remBins >= 4; sp++) {
      TS_POS(sp, ti, l, a)
      bool sig = (!numNonZero && sp == maxPos);
      if (!sig) {
        sig = lsDecodeBin(dec, ctx + R_TS_SIG_FLAG + (l != 0) + (a != 0));
        remBins--;
      }
      if (sig) {
        const uint32_t sign = lsDecodeBin(dec, ctx + signCtxTS(l, a, bdpcm));
        remBins--;
        signMask |= sign << (sp - minPos);
        numNonZero++;
        const uint32_t gt1 = lsDecodeBin(dec, ctx + R_TS_LRG1_FLAG + (bdpcm ? 3 : (l != 0) + (a != 0)));
        remBins--;
        uint32_t par = 0;
        if (gt1) {
          par = lsDecodeBin(dec, ctx + R_TS_PAR_FLAG);
          remBins--;
        }
        tile[ti] = (sign ? -1 : 1) * int(1 + par + gt1);
      }
      lastPass1 = sp;
    }
    for (sp = minPos; sp <= maxPos && remBins >= 4; sp++) {
      const int ti = int(__shfl_sync(FULL, myPos, sp - minPos) & 0xFFFFu);
      int t = abs(tile[ti]);
      for (int cutoff = 2; cutoff <= 8; cutoff += 2) {
        if (t >= cutoff) {
          t += int(lsDecodeBin(dec, ctx + R_TS_GTX_FLAG + (cutoff >> 1)) << 1);
          remBins--;
        }
      }
      tile[ti] = t;
      lastPass2 = sp;
    }
    for (sp = minPos; sp <= maxPos; sp++) {
      TS_POS(sp, ti, l, a)
      const int cutoff = sp <= lastPass2 ? 10 : (sp <= lastPass1 ? 2 : 0);
      int t = abs(tile[ti]);
      if (t >= cutoff) {
        const int rem = int(lsRemAbs(dec, 1));
        t += (sp <= lastPass1) ? (rem << 1) : rem;
        if (t && sp > lastPass1) {
          signMask |= dec.ep(1) << (sp - minPos);
        }
      }
      if (!bdpcm && cutoff && t > 0) t = decModCoeffTS(l, a, t);
      tile[ti] = t;
    }
#undef TS_POS
    // signs: one position per lane
    __syncwarp();
    int v = 0;
    if (lane < cgSize) {
      const int ti = int(myPos & 0xFFFFu);
      const int a = tile[ti];
      v = (signMask >> lane) & 1u ? -a : a;
      tile[ti] = v;
    }
    if (__any_sync(FULL, v < -32768 || v > 32767)) dec.status |= CABAC_SS_COEFF_RANGE;
    __syncwarp();
  }
}

}  // namespace cabac
