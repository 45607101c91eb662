// cabac_decode_thread.cuh - residual decoder with ONE THREAD per substream (included by cabac_kernels.cu).
//
// The lockstep decoder (cabac_lockstep.cuh) spends a warp on one substream: 32 lanes execute the same
// arithmetic-decoder instruction.  With 10^5 .. 10^6 substreams per launch there is enough independent
// work to give every lane its own substream instead - the decoder counterpart of the encoder's stage D
// (range_code_kernel).  Per thread: the arithmetic decoder in registers (LsDecoder, the same 64-bit value
// window and division-based bypass decoding), its 230 residual contexts as one packed 32-bit word each in a
// private row of global memory (L1 / L2 resident), and the coefficient block it is parsing read back from
// the output slab for the context templates (the blocks are zeroed by a coalesced pre-pass).
//
// Restates, scalar: CABACReader::residual_coding / residual_coding_subblock / residual_codingTS /
// last_sig_coeff (cabac_reader.cpp:2647-3339), CoeffCodingContext (context_modelling.hpp:71-385).
#pragma once

namespace cabac {

template <typename T>
struct TdBlock {  // the block being parsed, in the output slab: row-major, stride = width
  T *p;
  int log2w;
  __device__ __forceinline__ int get(int x, int y) const { return int(p[(y << log2w) + x]); }
  __device__ __forceinline__ void set(int x, int y, int v) const { p[(y << log2w) + x] = T(v); }
};

struct TdCoder : LsDecoder {
  // decodeBin on a packed model word (cabac_device.cuh): state1 | rate bits | state0
  __device__ __forceinline__ uint32_t bin(uint32_t *m) {
    const uint32_t w = *m;
    const uint32_t q = model_q(w);
    const uint32_t mps = q >> 7;
    const uint32_t lps = lps_from_q(q, range);
    nbins++;
    const uint32_t rmps = range - lps;
    const uint32_t sr = rmps << 23;
    const bool isLps = hi >= sr;
    const uint32_t b = mps ^ uint32_t(isLps);
    *m = model_update(w, b);
    const uint32_t nr = isLps ? lps : rmps;
    const int n = __clz(int(nr | 7u)) - 23;
    hi -= isLps ? sr : 0u;
    range = nr << n;
    shiftIn(n);
    refill();
    return b;
  }
  __device__ __forceinline__ uint32_t epPeekT(int n) {
    if (recipRange != range) {
      recip = __ldg(&g_recipTab[range - 256u]);
      recipRange = range;
    }
    return __umulhi(hi >> (23 - n), recip) >> 2;
  }
  __device__ __forceinline__ uint32_t epT(int n) {  // n <= 16
    if (n == 0) return 0;
    const uint32_t q = epPeekT(n);
    epCommit(n, q);
    return q;
  }
  __device__ __forceinline__ uint32_t epLongT(uint32_t n) {
    uint32_t v = 0;
    while (n > 16) {
      v = (v << 16) | epT(16);
      n -= 16;
    }
    return (v << n) | epT(int(n));
  }
  // decodeRemAbsEP(rice, 5, 15), arith_codec.cpp:153-179
  __device__ __forceinline__ uint32_t remAbsT(uint32_t rice) {
    const uint32_t q16 = epPeekT(16);
    const uint32_t ones = uint32_t(__clz(int(~(q16 << 16))));
    if (ones < 16u) {
      const bool esc = ones >= 5u;
      const uint32_t len = rice + (esc ? ones - 5u : 0u);
      const uint32_t n = ones + 1u + len;
      if (n <= 16u) {
        const uint32_t bins = q16 >> (16u - n);
        epCommit(int(n), bins);
        const uint32_t base = esc ? (1u << (ones - 5u)) + 4u : ones;
        return (base << rice) + (bins & ((1u << len) - 1u));
      }
      epCommit(int(ones + 1u), q16 >> (15u - ones));
      const uint32_t base = ones < 5u ? ones << rice : ((1u << (ones - 5u)) + 4u) << rice;
      return base + epLongT(len);
    }
    epCommit(16, 0xFFFFu);
    const uint32_t prefix = 16u + epT(1);
    const uint32_t len = prefix == 17u ? 15u : prefix - 5u + rice;
    return (((1u << (prefix - 5u)) + 4u) << rice) + epLongT(len);
  }
};

template <typename T>
__device__ __forceinline__ void tdTemplate(const TdBlock<T> &b, const TuGeom &g, int x, int y, int &sum1, int &numPos,
                                           int &sumAll) {
  sum1 = numPos = sumAll = 0;
  auto upd = [&](int v) {
    const int a = abs(v);
    sum1 += min(4 + (a & 1), a);
    numPos += a != 0;
    sumAll += a;
  };
  const bool r1 = x + 1 < g.tw, r2 = x + 2 < g.tw, b1 = y + 1 < g.th, b2 = y + 2 < g.th;
  if (r1) {
    upd(b.get(x + 1, y));
    if (r2) upd(b.get(x + 2, y));
    if (b1) upd(b.get(x + 1, y + 1));
  }
  if (b1) {
    upd(b.get(x, y + 1));
    if (b2) upd(b.get(x, y + 2));
  }
}

// residual_coding (regular), cabac_reader.cpp:2670-2735 + residual_coding_subblock :2948-3128
template <typename T>
__device__ uint32_t tdResidualRegular(TdCoder &dec, uint32_t *ctx, const TdBlock<T> &blk, const TuGeom &g,
                                      uint32_t subFlags, bool ts) {
  uint32_t res = 0;
  // last_sig_coeff, cabac_reader.cpp:2865-2940
  uint32_t posX = 0, posY = 0;
  {
    const int bx = lastXBase(g.ch) + g.lastOffsetX, by = lastYBase(g.ch) + g.lastOffsetY;
    while (posX < uint32_t(g.maxLastX) && dec.bin(ctx + bx + (posX >> g.lastShiftX))) posX++;
    while (posY < uint32_t(g.maxLastY) && dec.bin(ctx + by + (posY >> g.lastShiftY))) posY++;
    if (posX > 3) posX = minInGroup(posX) + dec.epT(int((posX - 2) >> 1));
    if (posY > 3) posY = minInGroup(posY) + dec.epT(int((posY - 2) >> 1));
  }
  if (posX >= uint32_t(g.tw) || posY >= uint32_t(g.th)) {  // a position in the zero-out region: malformed stream
    dec.status |= CABAC_SS_COEFF_RANGE;
    return res;
  }
  const int scanPosLast = xyToScan(g, int(posX), int(posY));
  if (!ts && g.w >= 4 && g.h >= 4) {
    const int maxLfnstPos = ((g.h == 4 && g.w == 4) || (g.h == 8 && g.w == 8)) ? 7 : 15;
    if (scanPosLast > maxLfnstPos) res |= CABAC_TUR_VIOLATES_LFNST;
    if (scanPosLast >= 1) res |= CABAC_TUR_LFNST_LAST_POS;
  }
  if (g.comp == 0 && !ts && scanPosLast >= 1) res |= CABAC_TUR_MTS_LAST_POS;
  const bool depQuant = (subFlags & CABAC_SUB_DEP_QUANT) != 0;
  const bool signHiding = (subFlags & CABAC_SUB_SIGN_HIDING) != 0;
  const int cgSize = 1 << g.log2cg;
  int state = 0, remRegBins = regBinLimit(g);
  uint64_t sigGroups = 0;
  const int lastCG = scanPosLast >> g.log2cg;
  for (int s = lastCG; s >= 0; s--) {
    const uint32_t c = __ldg(&g_rom.cgScan[g.lwig][g.lhig][s]);
    const int cgx = c & 15, cgy = c >> 4, cgPos = cgy * g.wig + cgx;
    if (g.sbtZero && ((g.h == 32 && cgy >= (16 >> g.log2cgh)) || (g.w == 32 && cgx >= (16 >> g.log2cgw)))) continue;
    const bool isLast = s == lastCG;
    bool sigGroup = isLast || s == 0;
    if (!sigGroup) {
      const int right = (cgx + 1 < g.wig) ? int((sigGroups >> (cgPos + 1)) & 1) : 0;
      const int lower = (cgy + 1 < g.hig) ? int((sigGroups >> (cgPos + g.wig)) & 1) : 0;
      sigGroup = dec.bin(ctx + R_SIG_GROUP + 2 * g.ch + (right | lower)) != 0;
    }
    if (!sigGroup) continue;
    sigGroups |= 1ull << cgPos;
    if (g.comp == 0 && (cgy > 3 || cgx > 3)) res |= CABAC_TUR_VIOLATES_MTS;
    const int x0 = cgx << g.log2cgw, y0 = cgy << g.log2cgh;
    const int P0 = isLast ? scanPosLast - (s << g.log2cg) : cgSize - 1;
    const int inferSigPos = isLast ? P0 : (s != 0 ? 0 : -1);
    uint32_t nzMask = 0, remMask = 0;
    // ---- 1st pass ----
    int p = P0;
    for (; p >= 0 && remRegBins >= 4; p--) {
      const uint32_t qxy = __ldg(&g_rom.inScan[g.shape][p]);
      const int x = x0 + int(qxy & 15u), y = y0 + int(qxy >> 4);
      int s1, np, sa;
      tdTemplate(blk, g, x, y, s1, np, sa);
      const int d = x + y;
      bool sig = nzMask == 0 && p == inferSigPos;
      if (!sig) {
        sig = dec.bin(ctx + sigFlagBase(g.ch, max(0, state - 1)) + sigCtxOffset(g, s1, d)) != 0;
        remRegBins--;
      }
      int val = 0;
      if (sig) {
        const int ofs = (isLast && p == P0) ? 0 : gtxCtxOffset(g, s1, np, d);
        const uint32_t gt1 = dec.bin(ctx + gtxFlagBase(g.ch + 2) + ofs);
        remRegBins--;
        uint32_t par = 0, gt2 = 0;
        if (gt1) {
          par = dec.bin(ctx + parFlagBase(g.ch) + ofs);
          gt2 = dec.bin(ctx + gtxFlagBase(g.ch) + ofs);
          remRegBins -= 2;
        }
        val = int(1 + par + gt1 + (gt2 << 1));
        nzMask |= 1u << p;
        remMask |= gt2 << p;
        blk.set(x, y, val);
      }
      if (depQuant) state = stateStep(state, val & 1);
    }
    // ---- 2nd pass (remainders of context-coded levels) and bypass pass ----
    const int cut = p;
    uint32_t visit = remMask | (cut >= 0 ? ((2u << cut) - 1u) : 0u);
    while (visit) {
      const int j = 31 - __clz(int(visit));
      visit &= ~(1u << j);
      const uint32_t qxy = __ldg(&g_rom.inScan[g.shape][j]);
      const int x = x0 + int(qxy & 15u), y = y0 + int(qxy >> 4);
      int s1, np, sa;
      tdTemplate(blk, g, x, y, s1, np, sa);
      if (j > cut) {
        blk.set(x, y, blk.get(x, y) + int(dec.remAbsT(goRicePar(sa - 20)) << 1));
      } else {
        const uint32_t rice = goRicePar(sa);
        const uint32_t pos0 = (state < 2 ? 1u : 2u) << rice;
        const uint32_t rem = dec.remAbsT(rice);
        const int val = int(rem == pos0 ? 0u : (rem < pos0 ? rem + 1u : rem));
        if (depQuant) state = stateStep(state, val & 1);
        if (val) {
          nzMask |= 1u << j;
          blk.set(x, y, val);
        }
      }
    }
    // ---- signs: coding order = descending position; the first coded sign is the pattern's MSB ----
    const int numNZ = __popc(nzMask);
    const bool hide = signHiding && nzMask && ((31 - __clz(int(nzMask))) - (__ffs(int(nzMask)) - 1) >= 4);
    const int numSigns = numNZ - int(hide);
    uint32_t pattern = dec.epT(numSigns);
    int sumAbs = 0, k = 0;
    uint32_t m = nzMask;
    while (m) {
      const int j = 31 - __clz(int(m));
      m &= ~(1u << j);
      const uint32_t qxy = __ldg(&g_rom.inScan[g.shape][j]);
      const int x = x0 + int(qxy & 15u), y = y0 + int(qxy >> 4);
      const int a = blk.get(x, y);
      sumAbs += a;
      bool neg;
      if (k < numSigns) neg = (pattern >> (numSigns - 1 - k)) & 1u;
      else neg = (sumAbs & 1) != 0;  // the hidden sign: parity of the group's sum
      const int v = neg ? -a : a;
      if (v < -32768 || v > 32767) dec.status |= CABAC_SS_COEFF_RANGE;
      if (neg) blk.set(x, y, v);
      k++;
    }
  }
  return res;
}

// residual_codingTS + residual_coding_subblockTS, cabac_reader.cpp:3130-3339
template <typename T>
__device__ void tdResidualTS(TdCoder &dec, uint32_t *ctx, const TdBlock<T> &blk, const TuGeom &g, bool bdpcm) {
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
      sigGroup = dec.bin(ctx + R_TS_SIG_GROUP + left + above) != 0;
    }
    if (!sigGroup) continue;
    sigGroups |= 1ull << cgPos;
    const int x0 = cgx << g.log2cgw, y0 = cgy << g.log2cgh;
    int numNonZero = 0, lastPass1 = -1, lastPass2 = -1;
    uint32_t signMask = 0;  // bit k: in-group position k is negative
    auto xy = [&](int k, int &x, int &y) {
      const uint32_t qxy = __ldg(&g_rom.inScan[g.shape][k]);
      x = x0 + int(qxy & 15u);
      y = y0 + int(qxy >> 4);
    };
    int k = 0;
    for (; k < cgSize && remBins >= 4; k++) {
      int x, y;
      xy(k, x, y);
      const int l = x > 0 ? blk.get(x - 1, y) : 0, a = y > 0 ? blk.get(x, y - 1) : 0;
      bool sig = !numNonZero && k == cgSize - 1;
      if (!sig) {
        sig = dec.bin(ctx + R_TS_SIG_FLAG + (l != 0) + (a != 0)) != 0;
        remBins--;
      }
      if (sig) {
        const uint32_t sign = dec.bin(ctx + signCtxTS(l, a, bdpcm));
        signMask |= sign << k;
        numNonZero++;
        const uint32_t gt1 = dec.bin(ctx + R_TS_LRG1_FLAG + (bdpcm ? 3 : (l != 0) + (a != 0)));
        remBins -= 2;
        uint32_t par = 0;
        if (gt1) {
          par = dec.bin(ctx + R_TS_PAR_FLAG);
          remBins--;
        }
        blk.set(x, y, (sign ? -1 : 1) * int(1 + par + gt1));  // signed: the neighbours' sign contexts read it
      }
      lastPass1 = k;
    }
    for (k = 0; k < cgSize && remBins >= 4; k++) {
      int x, y;
      xy(k, x, y);
      int t = abs(blk.get(x, y));
      for (int cutoff = 2; cutoff <= 8; cutoff += 2) {
        if (t >= cutoff) {
          t += int(dec.bin(ctx + R_TS_GTX_FLAG + (cutoff >> 1)) << 1);
          remBins--;
        }
      }
      blk.set(x, y, t);
      lastPass2 = k;
    }
    for (k = 0; k < cgSize; k++) {
      int x, y;
      xy(k, x, y);
      const int cutoff = k <= lastPass2 ? 10 : (k <= lastPass1 ? 2 : 0);
      int t = abs(blk.get(x, y));
      if (t >= cutoff) {
        const int rem = int(dec.remAbsT(1));
        t += (k <= lastPass1) ? (rem << 1) : rem;
        if (t && k > lastPass1) signMask |= dec.epT(1) << k;
      }
      if (!bdpcm && cutoff && t > 0) {
        const int l = x > 0 ? blk.get(x - 1, y) : 0, a = y > 0 ? blk.get(x, y - 1) : 0;
        t = decModCoeffTS(l, a, t);
      }
      blk.set(x, y, t);
    }
    for (k = 0; k < cgSize; k++) {
      if (!((signMask >> k) & 1u)) continue;
      int x, y;
      xy(k, x, y);
      blk.set(x, y, -blk.get(x, y));
    }
    for (k = 0; k < cgSize; k++) {
      int x, y;
      xy(k, x, y);
      const int v = blk.get(x, y);
      if (v < -32768 || v > 32767) dec.status |= CABAC_SS_COEFF_RANGE;
    }
  }
}

// pre-pass: the decoder only stores non-zero levels, so every block is cleared first (one warp per TU, coalesced)
template <typename T>
__global__ void __launch_bounds__(256) td_zero_blocks_kernel(const cabac_tu_desc *__restrict__ tus,
                                                             const uint64_t *__restrict__ tuOffsets, size_t nSub,
                                                             T *__restrict__ coeffs) {
  const size_t nTu = tuOffsets[nSub];
  const int lane = threadIdx.x & 31;
  const size_t warp = (size_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const size_t nWarps = (size_t(gridDim.x) * blockDim.x) >> 5;
  for (size_t t = tuOffsets[0] + warp; t < nTu; t += nWarps) {
    const cabac_tu_desc d = loadDesc(tus + t);
    T *dst = coeffs + d.coeff_offset;
    const int n = 1 << (d.log2_w + d.log2_h);
    for (int i = lane; i < n; i += 32) dst[i] = T(0);
  }
}

template <typename T, int MINB>
__global__ void __launch_bounds__(128, MINB)
decode_residual_thread_kernel(const uint8_t *__restrict__ in, const uint64_t *__restrict__ inOffsets,
                              const cabac_tu_desc *__restrict__ tus, const uint64_t *__restrict__ tuOffsets,
                              const cabac_substream_params *__restrict__ params, size_t nSub,
                              const cabac_ctx_model *__restrict__ ctxIn, T *__restrict__ coeffs,
                              uint32_t *__restrict__ numBins, uint32_t *__restrict__ bitsRead,
                              uint8_t *__restrict__ tuFlags, cabac_ctx_model *__restrict__ ctxOut,
                              uint32_t *__restrict__ status, const uint32_t *__restrict__ order,
                              uint32_t *__restrict__ ctxRows, int spread) {
  // spread: one substream per WARP (lane 0 works) - few, long substreams are latency-bound, every chain gets a
  // scheduler slot of its own instead of sharing a warp's instruction stream with 31 diverged neighbours
  size_t slot = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (spread) {
    if (threadIdx.x & 31) return;
    slot >>= 5;
  }
  if (slot >= nSub) return;
  const size_t sub = order ? size_t(order[slot]) : slot;
  const cabac_substream_params p = params[sub];
  const cabac_ctx_model *src = ((p.flags & CABAC_SUB_CTX_IN) && ctxIn) ? ctxIn + sub * CABAC_NUM_CTX : nullptr;
  uint32_t *ctx = ctxRows + slot * size_t(RCTX);  // this thread's contexts
  {
    const int qp = clipQp(p.qp);
    for (int k = 0; k < RCTX; k++) {
      const int id = rctx_inv(k);
      const uint32_t win = CABAC_CTX_INIT_TABLE[3][id];
      ctx[k] = src ? model_pack(src[id].state0, src[id].state1, win) : model_init(qp, CABAC_CTX_INIT_TABLE[p.init_id][id], win);
    }
  }
  TdCoder dec;
  dec.start(in + inOffsets[sub], uint32_t(inOffsets[sub + 1] - inOffsets[sub]));
  const uint64_t t0 = tuOffsets[sub], t1 = tuOffsets[sub + 1];
  for (uint64_t t = t0; t < t1; t++) {
    const cabac_tu_desc d = loadDesc(tus + t);
    if (d.flags & DESC_CLAMPED) dec.status |= CABAC_SS_BAD_DESC;
    const TuGeom g = makeGeom(d);
    TdBlock<T> blk;
    blk.p = coeffs + d.coeff_offset;
    blk.log2w = g.log2w;
    bool ts = (d.flags & CABAC_TU_BDPCM) ? true : (d.flags & CABAC_TU_TS) != 0;
    if (d.flags & CABAC_TU_TS_ALLOWED) ts = dec.bin(ctx + R_TS_FLAG + (g.comp ? 1 : 0)) != 0;
    uint32_t res = ts ? CABAC_TUR_TS : 0;
    if (ts && !(p.flags & CABAC_SUB_TSRC_DISABLED))
      tdResidualTS(dec, ctx, blk, g, (d.flags & CABAC_TU_BDPCM) != 0);
    else
      res |= tdResidualRegular(dec, ctx, blk, g, p.flags, ts);
    if (tuFlags) tuFlags[t] = uint8_t(res);
  }
  if (p.flags & CABAC_SUB_TERMINATE)
    if (!dec.decodeBinTrm()) dec.status |= CABAC_SS_NO_TERMINATE;
  dec.checkOverrun();
  if (numBins) numBins[sub] = dec.nbins;
  if (bitsRead) bitsRead[sub] = dec.bitsRead();
  if (status) status[sub] = dec.status;
  if (ctxOut) {
    const int qp = clipQp(p.qp);
    for (int id = 0; id < CABAC_NUM_CTX; id++) {
      cabac_ctx_model m;
      if (is_residual_ctx(id)) {
        const uint32_t w = ctx[rctx(id)];
        m.state0 = uint16_t(model_state0(w));
        m.state1 = uint16_t(model_state1(w));
      } else if (src) {
        m = src[id];
      } else {
        const uint32_t w = model_init(qp, CABAC_CTX_INIT_TABLE[p.init_id][id], CABAC_CTX_INIT_TABLE[3][id]);
        m.state0 = uint16_t(model_state0(w));
        m.state1 = uint16_t(model_state1(w));
      }
      ctxOut[sub * CABAC_NUM_CTX + id] = m;
    }
  }
}

}  // namespace cabac
