// cabac_kernels.cu - sm_100a kernels of the B200-native VVC CABAC engine.
//
// One warp per substream; WARPS_PER_BLOCK substreams per CTA.  Each warp owns a
// private slice of shared memory: its probability models, a coefficient tile,
// a command queue and an output staging ring (see cabac_device.cuh).
#include <cuda_runtime.h>

#include <cstdlib>

#include <cstring>
#include <algorithm>

#include "cabac_lockstep_enc.cuh"
#include "cabac_kernels.h"

namespace cabac {

#define CABAC_CTX_TABLE_QUALIFIER __device__ __constant__
#include "cabac_ctx_tables.inc"
#undef CABAC_CTX_TABLE_QUALIFIER
// m_binFracBits of the bit estimator: indexed per lane (divergent), so it lives in global memory
// behind L1 rather than in the constant bank
#define CABAC_CTX_TABLE_QUALIFIER __device__ const
#include "cabac_frac_bits.inc"
#undef CABAC_CTX_TABLE_QUALIFIER
#define CABAC_CTX_TABLE_QUALIFIER static const
namespace hosttab {
#include "cabac_ctx_tables.inc"
#include "cabac_frac_bits.inc"
}

constexpr int WARPS_PER_BLOCK = 4;
// resident CTAs per SM the register allocation of the two hot kernels is held to (occupancy sweep in
// 5 CTAs = 20 warps per SM is the measured optimum of both; tighter register
// caps cost more in extra instructions than the added warps return)
#ifndef ENC_MIN_CTAS
#define ENC_MIN_CTAS 5
#endif
#ifndef DEC_MIN_CTAS
#define DEC_MIN_CTAS 5
#endif

__device__ __forceinline__ int clipQp(int qp) { return qp < 0 ? 0 : (qp > 63 ? 63 : qp); }

// ---------------------------------------------------------------------------
// Work distribution.  Substreams differ in length by orders of magnitude (a
// QP0 CTU carries ~40x the bins of a QP51 one) and each is a serial chain, so
// the makespan is set by when the longest ones START.  Every launch therefore
// (1) estimates a cost per substream, (2) orders substreams longest-first with a
// 1024-bucket counting sort on a log-scale key, and (3) runs persistent warps
// that pull the next substream from an atomic cursor (LPT list scheduling).
// ---------------------------------------------------------------------------
__device__ __forceinline__ bool nextSubstream(const uint32_t *__restrict__ order, uint32_t *cursor, size_t nSub,
                                              int lane, size_t &sub) {
  uint32_t i = 0;
  if (lane == 0) i = atomicAdd(cursor, 1u);
  i = __shfl_sync(FULL, i, 0);
  if (i >= nSub) return false;
  sub = order ? size_t(order[i]) : size_t(i);
  return true;
}

// decode: bitstream bytes are an excellent proxy for bins
__global__ void cost_bytes_kernel(const uint64_t *__restrict__ inOffsets, size_t n, uint32_t *__restrict__ cost) {
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) {
    const uint64_t b = inOffsets[i + 1] - inOffsets[i];
    cost[i] = b > 0xFFFFFFFFull ? 0xFFFFFFFFu : uint32_t(b);
  }
}

// decode: bytes x weight + what every block costs whatever it holds (descriptor handling, clearing and
// storing its coefficient positions).  One CTA per substream over its descriptors.
__global__ void __launch_bounds__(256) cost_decode_kernel(const uint64_t *__restrict__ inOffsets,
                                                          const cabac_tu_desc *__restrict__ tus,
                                                          const uint64_t *__restrict__ tuOffsets, size_t n,
                                                          uint32_t byteWeight, uint32_t tuConst, uint32_t areaShift,
                                                          uint32_t *__restrict__ cost) {
  __shared__ uint32_t sFixed;
  for (size_t s = blockIdx.x; s < n; s += gridDim.x) {
    if (threadIdx.x == 0) sFixed = 0;
    __syncthreads();
    uint32_t fixed = 0;
    for (uint64_t t = tuOffsets[s] + threadIdx.x, te = tuOffsets[s + 1]; t < te; t += blockDim.x) {
      const cabac_tu_desc d = loadDesc(tus + t);
      fixed += tuConst + ((1u << (d.log2_w + d.log2_h)) >> areaShift);
    }
    fixed = __reduce_add_sync(FULL, fixed);
    if ((threadIdx.x & 31) == 0 && fixed) atomicAdd(&sFixed, fixed);
    __syncthreads();
    if (threadIdx.x == 0) {
      const uint64_t b = (inOffsets[s + 1] - inOffsets[s]) * byteWeight + sFixed;
      cost[s] = b > 0xFFFFFFFFull ? 0xFFFFFFFFu : uint32_t(b);
    }
    __syncthreads();
  }
}

// Upper bound of the command words one block can produce: header <= 23,
// <= 2 per coefficient group (group flag, sign word), context-coded words <= 1.75 A + 8 (the
// regular-bin budget), <= 3 per coefficient of Golomb-Rice / sign words in the bypass passes.
__device__ __forceinline__ uint32_t tuCmdCapacity(const cabac_tu_desc &d) {
  const uint32_t lw = d.log2_w < 5 ? d.log2_w : 5, lh = d.log2_h < 5 ? d.log2_h : 5;
  const uint32_t area = 1u << (lw + lh);
  const uint32_t log2cg = (lw >= 2 && lh >= 2) ? 4u : ((lw + lh <= 3) ? 2u : 4u);
  const uint32_t numCG = area >> log2cg ? area >> log2cg : 1u;
  return 5u * area + 2u * numCG + 32u;
}

// encode: one streaming pass over the coefficients, ~2 bins per significant bit
// One CTA (8 warps) per substream: warp w takes the descriptor groups w, w + 8, ... of 32 blocks each
// (one coalesced descriptor load per group, broadcast by shuffle), so a long substream is no longer one
// warp's chain of dependent loads.
__global__ void __launch_bounds__(256) cost_coeff_kernel(const cabac_tu_desc *__restrict__ tus,
                                                         const uint64_t *__restrict__ tuOffsets,
                                                         const int32_t *__restrict__ coeffs, size_t n,
                                                         uint32_t *__restrict__ cost,
                                                         uint32_t *__restrict__ capWords,
                                                         const uint64_t *__restrict__ rawOffsets,
                                                         const cabac_substream_params *__restrict__ params,
                                                         int sampleMask, uint32_t tuConst, uint32_t areaShift,
                                                         int c16) {
  // sampleMask < 0: every block's coefficients are read (slab in HBM).  Otherwise the slab is pinned host
  // memory behind PCIe: the cost is the area / QP / transform-skip estimate of all blocks, scaled by
  // what the estimate misses on the blocks whose index has (index & sampleMask) == 0, which ARE read.
  __shared__ uint32_t sCost, sCap, sReal, sGuess, sFixed;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  for (size_t s = blockIdx.x; s < n; s += gridDim.x) {
    if (threadIdx.x == 0) {
      sCost = 0;
      sCap = 0;
      sReal = 0;
      sGuess = 0;
      sFixed = 0;
    }
    __syncthreads();
    uint32_t c = 0, cap = 0, real = 0, guess = 0, fixed = 0;
    const bool sampling = coeffs != nullptr && sampleMask >= 0;
    if (rawOffsets && threadIdx.x == 0) {  // recorded commands: one word each, REM commands may become two
      const uint32_t nRaw = uint32_t(rawOffsets[s + 1] - rawOffsets[s]);
      cap = 2u * nRaw;
      c = nRaw;
    }
    const uint64_t t0 = tus ? tuOffsets[s] : 0, te = tus ? tuOffsets[s + 1] : 0;
    const int q = 70 - int(params ? params[s].qp : 32);
    for (uint64_t base = t0 + uint64_t(wib) * 32u; base < te; base += uint64_t(blockDim.x)) {
      const bool have = base + uint64_t(lane) < te;
      cabac_tu_desc dl = {};
      if (have) dl = loadDesc(tus + base + uint64_t(lane));
      const uint32_t area = have ? 1u << (dl.log2_w + dl.log2_h) : 0u;
      const bool sampled = have && sampling && ((base + uint64_t(lane) - t0) & uint64_t(sampleMask)) == 0;
      if (have) {
        cap += tuCmdCapacity(dl);
        if (!coeffs || sampling) {
          // zero-copy launches (the slab stays in pinned host memory): no pass over the coefficients;
          // estimate from the block area, the slice QP (bins per coefficient fall ~4x per 20 QP steps)
          // and transform skip (no frequency decay: ~1.5x the bins)
          const uint32_t e = (area * uint32_t(q * q) >> 6) * ((dl.flags & CABAC_TU_TS) ? 3u : 2u);
          c += e;
          if (sampled) guess += e;
          fixed += tuConst + (area >> areaShift);  // the per-block term is known from the descriptor
        } else {
          c += tuConst + (area >> areaShift);
        }
      }
      if (!coeffs) continue;
      const int cnt = int(min(uint64_t(32), te - base));
      uint32_t todo = sampling ? __ballot_sync(FULL, sampled) : (cnt == 32 ? FULL : (1u << cnt) - 1u);
      uint32_t &acc = sampling ? real : c;
      while (todo) {
        const int j = __ffs(int(todo)) - 1;
        todo &= todo - 1u;
        const uint64_t cofs = __shfl_sync(FULL, dl.coeff_offset, j);
        const int32_t *src = coeffs + cofs;
        const int total = int(__shfl_sync(FULL, area, j));
        if (c16) {
          const int16_t *s16 = reinterpret_cast<const int16_t *>(coeffs) + cofs;
          for (int i = lane; i < total; i += 32) {
            const int v = s16[i];
            acc += v ? 2 * (33 - __clz(abs(v))) : 0;
          }
        } else if ((reinterpret_cast<uintptr_t>(src) & 15u) == 0 && total >= 128) {
          for (int i = lane * 4; i < total; i += 128) {
            const int4 v = __ldg(reinterpret_cast<const int4 *>(src + i));
            acc += (v.x ? 2 * (33 - __clz(abs(v.x))) : 0) + (v.y ? 2 * (33 - __clz(abs(v.y))) : 0) +
                 (v.z ? 2 * (33 - __clz(abs(v.z))) : 0) + (v.w ? 2 * (33 - __clz(abs(v.w))) : 0);
          }
        } else {
          for (int i = lane; i < total; i += 32) {
            const int v = __ldg(src + i);
            acc += v ? 2 * (33 - __clz(abs(v))) : 0;
          }
        }
      }
    }
    c = __reduce_add_sync(FULL, c);
    cap = __reduce_add_sync(FULL, cap);
    if (lane == 0 && (c | cap)) {
      atomicAdd(&sCost, c);
      atomicAdd(&sCap, cap);
    }
    if (sampling) {
      real = __reduce_add_sync(FULL, real);
      guess = __reduce_add_sync(FULL, guess);
      fixed = __reduce_add_sync(FULL, fixed);
      if (lane == 0 && (real | guess | fixed)) {
        atomicAdd(&sReal, real);
        atomicAdd(&sGuess, guess);
        atomicAdd(&sFixed, fixed);
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      // sampled blocks: `real` counts what the HBM pass counts (2 per significant bit), `guess` is the
      // estimate of the same blocks; the ratio corrects the estimate of all blocks
      cost[s] = sampling && sGuess
                    ? uint32_t(min(uint64_t(sCost) * (uint64_t(sReal) + 1u) / (uint64_t(sGuess) + 1u) + sFixed,
                                   uint64_t(0xFFFFFFFFu)))
                    : sCost;
      if (capWords) capWords[s] = (sCap + 3u) & ~3u;  // keeps every substream's slab 16-byte aligned
    }
    __syncthreads();
  }
}

constexpr int COST_BUCKETS = 1024;
__device__ __forceinline__ uint32_t costKey(uint32_t c) {  // 5 mantissa bits under the leading one (3 % steps)
  if (c < 32) return c;
  const int e = 31 - __clz(c);
  return uint32_t(e - 4) * 32u + ((c >> (e - 5)) & 31u);  // 32..895
}

__global__ void __launch_bounds__(1024) lpt_order_kernel(const uint32_t *__restrict__ cost, size_t n,
                                                         uint32_t *__restrict__ order, uint32_t *cursor) {
  __shared__ uint32_t hist[COST_BUCKETS], start[COST_BUCKETS];
  if (threadIdx.x < COST_BUCKETS) hist[threadIdx.x] = 0;
  if (threadIdx.x == 0) *cursor = 0;
  __syncthreads();
  for (size_t i = threadIdx.x; i < n; i += blockDim.x) atomicAdd(&hist[costKey(cost[i])], 1u);
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t acc = 0;
    for (int b = COST_BUCKETS - 1; b >= 0; b--) {  // longest first
      start[b] = acc;
      acc += hist[b];
    }
  }
  __syncthreads();
  for (size_t i = threadIdx.x; i < n; i += blockDim.x) order[atomicAdd(&start[costKey(cost[i])], 1u)] = uint32_t(i);
}

// ---------------------------------------------------------------------------
// K1: context initialisation (Ctx::init, contexts.cpp:996-1015)
// ---------------------------------------------------------------------------
__global__ void ctx_init_kernel(const cabac_substream_params *__restrict__ params, size_t n,
                                cabac_ctx_model *__restrict__ out) {
  const size_t total = n * CABAC_NUM_CTX;
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < total; i += size_t(gridDim.x) * blockDim.x) {
    const size_t s = i / CABAC_NUM_CTX;
    const int k = int(i - s * CABAC_NUM_CTX);
    const cabac_substream_params p = params[s];
    const uint32_t w = model_init(clipQp(p.qp), CABAC_CTX_INIT_TABLE[p.init_id][k], CABAC_CTX_INIT_TABLE[3][k]);
    cabac_ctx_model m;
    m.state0 = uint16_t(model_state0(w));
    m.state1 = uint16_t(model_state1(w));
    out[i] = m;
  }
}

// ---------------------------------------------------------------------------
// WPP (wavefront parallel processing) inside one launch.  Substream i may name a predecessor
// (cabac_wpp_sync::dep, the CTU row above): it starts from the contexts the predecessor held after
// its first CTU (EncSlice.cpp:2057-2072, 2085-2089; DecSlice.cpp:194-210, 241-245; the copy is
// Ctx::operator=, contexts.cpp:1096-1108).  The predecessor's warp writes that snapshot to global
// memory at its sync point, fences and raises a flag; the dependant's warp spins on the flag.
// Substreams are claimed in index order and dep < own index, so a predecessor is always already
// claimed by a running warp when somebody waits for it: no deadlock, whatever the grid size.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t ldAcquire(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void stRelease(uint32_t *p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// The row of 379 models a substream starts from: the predecessor's snapshot (WPP), its ctx_in row
// (CABAC_SUB_CTX_IN) or nullptr = Ctx::init(qp, initId).  Blocks until a WPP snapshot is published.
__device__ __forceinline__ const cabac_ctx_model *ctxSource(const cabac_substream_params &p,
                                                            const cabac_ctx_model *ctxIn, size_t sub,
                                                            const WppArgs &wpp, int lane, uint32_t &st) {
  if (wpp.sync) {
    const uint32_t dep = wpp.sync[sub].dep;
    if (dep != CABAC_WPP_NONE && !(wpp.debug & 1u)) {
      // A predecessor that never publishes (a malformed dep chain) ends in a status bit, not in a hung GPU:
      // ~2^24 polls of >= 100 ns are far beyond any legitimate wait (a chain of first CTUs).
      // every lane polls (a warp-uniform loop: no lane-0-only region for the others to wait out)
      const volatile uint32_t *flag = wpp.ready + dep;
      uint32_t spins = 0;
      bool ok = true;
      while (*flag == 0u) {
        __nanosleep(100);
        if (++spins > wpp.spinLimit) {
          ok = false;
          break;
        }
      }
      __syncwarp();
      if (!__all_sync(FULL, ok)) {
        st |= CABAC_SS_WPP_TIMEOUT;
        return nullptr;
      }
      __threadfence();
      (void)ldAcquire(wpp.ready + dep);  // every lane orders its own snapshot loads after the flag
      return wpp.snap + size_t(dep) * CABAC_NUM_CTX;
    }
  }
  return ((p.flags & CABAC_SUB_CTX_IN) && ctxIn) ? ctxIn + sub * CABAC_NUM_CTX : nullptr;
}
__device__ __forceinline__ void wppPublish(const WppArgs &wpp, size_t sub, int lane) {
  if (wpp.debug & 2u) return;
  __threadfence();
  __syncwarp();
  if (lane == 0) stRelease(wpp.ready + sub, 1u);
}

// Residual contexts as 16-byte entries (see cabac_lockstep.cuh).  src: see ctxSource.
__device__ inline void loadCtxEntries(uint4 *ctx, const cabac_substream_params &p, const cabac_ctx_model *src,
                                      int lane) {
  const int qp = clipQp(p.qp);
  for (int k = lane; k < RCTX; k += 32) {
    const int id = rctx_inv(k);
    const uint32_t win = CABAC_CTX_INIT_TABLE[3][id];
    if (src) {
      const cabac_ctx_model m = src[id];
      ctx[k] = ctx_entry(m.state0, m.state1, win);
    } else {
      ctx[k] = ctx_entry_init(qp, CABAC_CTX_INIT_TABLE[p.init_id][id], win);
    }
  }
  __syncwarp();
}

// dst: a row of CABAC_NUM_CTX models (ctx_out of the substream, or its WPP snapshot).  Contexts the
// residual syntax does not touch pass through from the source row.
__device__ inline void storeCtxEntries(const uint4 *ctx, const cabac_substream_params &p, const cabac_ctx_model *src,
                                       cabac_ctx_model *dst, int lane) {
  if (!dst) return;
  __syncwarp();
  const int qp = clipQp(p.qp);
  for (int id = lane; id < CABAC_NUM_CTX; id += 32) {
    cabac_ctx_model m;
    if (is_residual_ctx(id)) {
      const uint32_t x = ctx[rctx(id)].x;
      m.state0 = uint16_t(x >> 16);
      m.state1 = uint16_t(x & 0xFFFFu);
    } else if (src) {
      m = src[id];
    } else {
      const uint32_t w = model_init(qp, CABAC_CTX_INIT_TABLE[p.init_id][id], CABAC_CTX_INIT_TABLE[3][id]);
      m.state0 = uint16_t(model_state0(w));
      m.state1 = uint16_t(model_state1(w));
    }
    dst[id] = m;
  }
}

// All 379 contexts (recorded command streams address them by their reference ids).
__device__ inline void loadCtxEntriesFull(uint4 *ctx, const cabac_substream_params &p, const cabac_ctx_model *src,
                                          int lane) {
  const int qp = clipQp(p.qp);
  for (int id = lane; id < CABAC_NUM_CTX; id += 32) {
    const uint32_t win = CABAC_CTX_INIT_TABLE[3][id];
    if (src) {
      const cabac_ctx_model m = src[id];
      ctx[id] = ctx_entry(m.state0, m.state1, win);
    } else {
      ctx[id] = ctx_entry_init(qp, CABAC_CTX_INIT_TABLE[p.init_id][id], win);
    }
  }
  __syncwarp();
}
__device__ inline void storeCtxEntriesFull(const uint4 *ctx, cabac_ctx_model *dst, int lane) {
  if (!dst) return;
  __syncwarp();
  for (int id = lane; id < CABAC_NUM_CTX; id += 32) {
    cabac_ctx_model m;
    m.state0 = uint16_t(ctx[id].x >> 16);
    m.state1 = uint16_t(ctx[id].x & 0xFFFFu);
    dst[id] = m;
  }
}

// ---------------------------------------------------------------------------
// K2: bin-level encoding is cabac_encode_stream without TU markers (stage P in stream mode +
//     stage D, below).
// K3: bin-level decoder with a known call skeleton (test vehicle: the context of every bin is
//     given by the command words).  Lockstep: every lane runs the same decoder.
// ---------------------------------------------------------------------------
struct __align__(16) BinsDecWarpSmem {
  uint4 ctx[CABAC_NUM_CTX];
};

__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
decode_bins_kernel(const uint8_t *__restrict__ in, const uint64_t *__restrict__ inOffsets,
                   const uint32_t *__restrict__ cmds, const uint64_t *__restrict__ cmdOffsets,
                   const cabac_substream_params *__restrict__ params, size_t nSub,
                   const cabac_ctx_model *__restrict__ ctxIn, uint32_t *__restrict__ values,
                   uint32_t *__restrict__ bitsRead, cabac_ctx_model *__restrict__ ctxOut,
                   uint32_t *__restrict__ status, const uint32_t *__restrict__ order, uint32_t *cursor) {
  __shared__ BinsDecWarpSmem smem[WARPS_PER_BLOCK];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  BinsDecWarpSmem &sm = smem[wib];
  size_t sub;
  while (nextSubstream(order, cursor, nSub, lane, sub)) {
    const cabac_substream_params p = params[sub];
    loadCtxEntriesFull(sm.ctx, p, ((p.flags & CABAC_SUB_CTX_IN) && ctxIn) ? ctxIn + sub * CABAC_NUM_CTX : nullptr, lane);
    LsDecoder dec;
    dec.start(in + inOffsets[sub], uint32_t(inOffsets[sub + 1] - inOffsets[sub]));
    const uint64_t c0 = cmdOffsets[sub], c1 = cmdOffsets[sub + 1];
#pragma unroll 1
    for (uint64_t c = c0; c < c1; c++) {
      const uint32_t w = cmds[c];
      uint32_t v = 0;
      switch (CABAC_CMD_KIND(w)) {
        case CABAC_CMD_CTX:
          if ((w & 511u) < CABAC_NUM_CTX) v = dec.decodeBin(sm.ctx + (w & 511u));
          else dec.status |= CABAC_SS_BAD_CMD;
          break;
        case CABAC_CMD_EP: v = dec.epLong((w >> 24) & 31u); break;
        case CABAC_CMD_REM: v = dec.remAbs((w >> 24) & 31u); break;
        case CABAC_CMD_TRM: v = dec.decodeBinTrm(); break;
        case CABAC_CMD_ALIGN: dec.range = 256; break;
        default: dec.status |= CABAC_SS_BAD_CMD; break;
      }
      if (values && lane == 0) values[c] = v;
    }
    if (p.flags & CABAC_SUB_TERMINATE)
      if (!dec.decodeBinTrm()) dec.status |= CABAC_SS_NO_TERMINATE;
    dec.checkOverrun();
    if (lane == 0) {
      if (bitsRead) bitsRead[sub] = dec.bitsRead();
      if (status) status[sub] = dec.status;
    }
    storeCtxEntriesFull(sm.ctx, ctxOut ? ctxOut + sub * CABAC_NUM_CTX : nullptr, lane);
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------
// K4: residual-level encoder
// ---------------------------------------------------------------------------
template <int kNumCtx>
struct __align__(16) EncWarpSmemT {
  uint4 ctx[kNumCtx];
  uint8_t ring[STAGE_BYTES];  // output staging of the fused variant
  int16_t tile[32 * 32];
  uint32_t queue[QUEUE_WORDS];
  uint16_t cgMask[64];  // nonzero mask of every coefficient group, by scan index
};

// Stages the coded region of a TU into the int16 tile (all lanes).  Returns
// false (warp-uniform) when a level is outside the int16 TCoeff range.
__device__ inline bool loadTile(int16_t *tile, const int32_t *__restrict__ src, const TuGeom &g, int lane,
                                bool c16 = false) {
  bool ok = true;
  const int n = g.tw * g.th;
  if (c16) {  // int16 TCoeff slab (cabac_*_i16): same layout, half the bytes, always in range
    const int16_t *s16 = reinterpret_cast<const int16_t *>(src);
    for (int i = lane; i < n; i += 32) {
      const int y = i >> g.log2tw, x = i & (g.tw - 1);
      tile[i] = s16[(size_t(y) << g.log2w) + x];
    }
    __syncwarp();
    return true;
  }
  if ((g.tw & 3) == 0 && (reinterpret_cast<uintptr_t>(src) & 15u) == 0) {
    for (int i = lane * 4; i < n; i += 128) {
      const int y = i >> g.log2tw, x = i & (g.tw - 1);
      const int4 v = __ldg(reinterpret_cast<const int4 *>(src + (size_t(y) << g.log2w) + x));
      ok &= (v.x == int16_t(v.x)) & (v.y == int16_t(v.y)) & (v.z == int16_t(v.z)) & (v.w == int16_t(v.w));
      *reinterpret_cast<short4 *>(tile + i) = make_short4(short(v.x), short(v.y), short(v.z), short(v.w));
    }
  } else {
    for (int i = lane; i < n; i += 32) {
      const int y = i >> g.log2tw, x = i & (g.tw - 1);
      const int v = __ldg(src + (size_t(y) << g.log2w) + x);
      ok &= v == int16_t(v);
      tile[i] = int16_t(v);
    }
  }
  __syncwarp();
  return __all_sync(FULL, ok);
}

// Blocks of one shape and component share their geometry: consecutive descriptors usually do.
__device__ __forceinline__ uint32_t geomKey(const cabac_tu_desc &d) {
  return uint32_t(d.log2_w) | (uint32_t(d.log2_h) << 8) | (uint32_t(d.comp_id) << 16) |
         (uint32_t(d.flags & CABAC_TU_SBT_ZEROOUT) << 24);
}

// Asynchronous variant for the fused encoder: the coded part (<= 32 x 32) of the NEXT block is
// fetched with cp.async into a raw int32 staging tile while the current block is coded, so the
// HBM - or, for zero-copy launches, PCIe - round trip of a block never sits on the warp's serial path.
__device__ __forceinline__ void prefetchBlock(int32_t *raw, const int32_t *__restrict__ src, const cabac_tu_desc &d,
                                              int lane, bool c16 = false) {
  const int log2w = d.log2_w, log2tw = min(int(d.log2_w), 5), log2th = min(int(d.log2_h), 5);
  const uint32_t dst = uint32_t(__cvta_generic_to_shared(raw));
  if (c16) {
    // int16 slab: the coded part lands in `raw` as tw x th int16 (row-major, stride tw).  Blocks up to 32 wide
    // are contiguous in the slab (stride == width, unit.hpp:551); 64-wide ones contribute 64-byte row heads.
    const char *s8 = reinterpret_cast<const char *>(src);
    const int rowBytes = 2 << log2tw, rows = (log2w <= 5) ? 1 : (1 << log2th);
    const int lineBytes = (log2w <= 5) ? (2 << (log2tw + log2th)) : rowBytes;  // contiguous bytes per source line
    const size_t srcStride = size_t(2) << log2w;
    if ((reinterpret_cast<uintptr_t>(s8) & 15u) == 0 && (lineBytes & 15) == 0) {
      const int perLine = lineBytes >> 4, total = perLine * rows;
      for (int u = lane; u < total; u += 32) {
        const int y = u / perLine, k = u - y * perLine;
        asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst + uint32_t(y * lineBytes + k * 16)),
                     "l"(s8 + y * srcStride + k * 16)
                     : "memory");
      }
    } else if ((reinterpret_cast<uintptr_t>(s8) & 3u) == 0 && (lineBytes & 3) == 0) {
      const int perLine = lineBytes >> 2, total = perLine * rows;
      for (int u = lane; u < total; u += 32) {
        const int y = u / perLine, k = u - y * perLine;
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst + uint32_t(y * lineBytes + k * 4)),
                     "l"(s8 + y * srcStride + k * 4)
                     : "memory");
      }
    } else {  // 2-byte aligned odd-sized pieces: plain loads
      int16_t *r16 = reinterpret_cast<int16_t *>(raw);
      const int16_t *s16 = reinterpret_cast<const int16_t *>(src);
      const int n = 1 << (log2tw + log2th);
      for (int i = lane; i < n; i += 32) r16[i] = s16[(size_t(i >> log2tw) << log2w) + (i & ((1 << log2tw) - 1))];
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    return;
  }
  if ((reinterpret_cast<uintptr_t>(src) & 15u) == 0 && log2tw >= 2) {
    const int total = 1 << (log2tw + log2th - 2);  // 16-byte units
    for (int u = lane; u < total; u += 32) {
      const int y = u >> (log2tw - 2), x = (u & ((1 << (log2tw - 2)) - 1)) << 2;
      asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst + uint32_t(u) * 16u),
                   "l"(src + (size_t(y) << log2w) + x)
                   : "memory");
    }
  } else {
    const int total = 1 << (log2tw + log2th);
    for (int i = lane; i < total; i += 32) {
      const int y = i >> log2tw, x = i & ((1 << log2tw) - 1);
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst + uint32_t(i) * 4u),
                   "l"(src + (size_t(y) << log2w) + x)
                   : "memory");
    }
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
}
// waits for the staged block and narrows it into the int16 tile; false when a level is out of range
__device__ __forceinline__ bool takeBlock(int16_t *tile, const int32_t *raw, const TuGeom &g, int lane,
                                          bool c16 = false) {
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncwarp();
  bool ok = true;
  const int n = g.tw * g.th;
  if (c16) {  // already int16 in the landing tile: move it over
    const int16_t *r16 = reinterpret_cast<const int16_t *>(raw);
    if (n >= 2)
      for (int i = lane * 2; i < n; i += 64) *reinterpret_cast<uint32_t *>(tile + i) = *reinterpret_cast<const uint32_t *>(r16 + i);
    else if (lane == 0)
      tile[0] = r16[0];
    __syncwarp();
    return true;
  }
  if (n < 4) {  // 2 x 1 / 1 x 2 / 1 x 1
    if (lane < n) {
      const int v = raw[lane];
      ok = v == int16_t(v);
      tile[lane] = int16_t(v);
    }
    __syncwarp();
    return __all_sync(FULL, ok);
  }
  for (int i = lane * 4; i < n; i += 128) {
    const int4 v = *reinterpret_cast<const int4 *>(raw + i);
    ok &= (v.x == int16_t(v.x)) & (v.y == int16_t(v.y)) & (v.z == int16_t(v.z)) & (v.w == int16_t(v.w));
    *reinterpret_cast<short4 *>(tile + i) = make_short4(short(v.x), short(v.y), short(v.z), short(v.w));
  }
  __syncwarp();
  return __all_sync(FULL, ok);
}

// One regular-residual coefficient group, binarised by lanes 0..cgSize-1
// (residual_coding_subblock, cabac_writer.cpp:2724-2872).  Returns the number
// of command words written to `queue`.  All lanes must call.
__device__ inline int binariseCG(uint32_t *queue, const int16_t *tile, const TuGeom &g, int lane, int s,
                                 int scanPosLast, bool codeSigGroup, int sigGroupCtx, bool depQuant,
                                 bool signHiding, int &state, int &remRegBins) {
  const int cgSize = 1 << g.log2cg;
  const int minSubPos = s << g.log2cg;
  const bool isLast = (scanPosLast >> g.log2cg) == s;
  const int P0 = isLast ? scanPosLast - minSubPos : cgSize - 1;  // first coded position (in-CG index)
  const bool active = lane <= P0;                                 // implies lane < cgSize
  const uint32_t activeMask = (P0 >= 31) ? FULL : ((2u << P0) - 1u);
  const uint32_t above = activeMask & ~((2u << lane) - 1u);       // positions coded before this lane's
  int nq = 0;
  if (codeSigGroup) {
    if (lane == 0) queue[0] = CABAC_MK_CTX(sigGroupCtx, 1);
    nq = 1;
  }

  int x = 0, y = 0, v = 0, sumAbs1 = 0, numPos = 0, sumAll = 0;
  if (active) {
    const uint32_t c = __ldg(&g_rom.cgScan[g.lwig][g.lhig][s]);
    const uint32_t q = __ldg(&g_rom.inScan[g.shape][lane]);
    x = int((c & 15u) << g.log2cgw) + int(q & 15u);
    y = int((c >> 4) << g.log2cgh) + int(q >> 4);
    v = tile[(y << g.log2tw) + x];
    templateSums(tile, g, x, y, sumAbs1, numPos, sumAll);
  }
  const int a = abs(v);
  const bool nz = a != 0;
  const uint32_t Z = __ballot_sync(FULL, active && nz);
  const uint32_t G = __ballot_sync(FULL, active && a > 1);
  const uint32_t Pm = __ballot_sync(FULL, active && (a & 1));
  const uint32_t Neg = __ballot_sync(FULL, active && v < 0);

  // sig_coeff_flag is inferred at one position when nothing non-zero preceded it
  const int inferLane = isLast ? P0 : (s != 0 ? 0 : -1);
  const bool sigCoded = active && ((Z & above) != 0 || lane != inferLane);
  const uint32_t S = __ballot_sync(FULL, sigCoded);
  // context-coded bins spent before this position if everything so far was context coded
  const int cum = __popc(S & above) + __popc(Z & above) + 2 * __popc(G & above);
  const bool inPass1 = active && (remRegBins - cum >= 4);
  const uint32_t M1 = __ballot_sync(FULL, inPass1);
  const int myBins = inPass1 ? (int(sigCoded) + (nz ? 1 + (a > 1 ? 2 : 0) : 0)) : 0;
  const int pass1Total = __popc(S & M1) + __popc(Z & M1) + 2 * __popc(G & M1);

  // dependent-quantisation state before this position (see stateStep)
  int st = 0;
  int stateEnd = 0;
  if (depQuant) {
    // step t = P0 - pos + 1 for the position `pos`; odd t <=> (P0 - pos) even
    const uint32_t evenDist = (P0 & 1) ? 0xAAAAAAAAu : 0x55555555u;  // positions with (P0 - pos) even
    const int h0 = state >> 1, l0 = state & 1;
    {
      const int k = P0 - lane;  // steps taken before this lane's position
      const int O = __popc(Pm & above & evenDist) & 1, E = __popc(Pm & above & ~evenDist) & 1;
      const int hi = (k & 1) ? (l0 ^ O) : (h0 ^ E);
      const int lo = (k & 1) ? (h0 ^ E) : (l0 ^ O);
      st = (hi << 1) | lo;
    }
    {
      const int k = P0 + 1;
      const int O = __popc(Pm & activeMask & evenDist) & 1, E = __popc(Pm & activeMask & ~evenDist) & 1;
      const int hi = (k & 1) ? (l0 ^ O) : (h0 ^ E);
      const int lo = (k & 1) ? (h0 ^ E) : (l0 ^ O);
      stateEnd = (hi << 1) | lo;
    }
  }

  const int diag = x + y;
  const bool atLast = isLast && lane == P0;
  if (inPass1) {
    int qi = nq + cum;
    if (sigCoded) {
      const int ctxId = sigFlagBase(g.ch, max(0, st - 1)) + sigCtxOffset(g, sumAbs1, diag);
      queue[qi++] = CABAC_MK_CTX(ctxId, nz);
    }
    if (nz) {
      const int ofs = atLast ? 0 : gtxCtxOffset(g, sumAbs1, numPos, diag);
      queue[qi++] = CABAC_MK_CTX(gtxFlagBase(g.ch + 2) + ofs, a > 1);
      if (a > 1) {
        const int rem = a - 2;
        queue[qi++] = CABAC_MK_CTX(parFlagBase(g.ch) + ofs, rem & 1);
        queue[qi++] = CABAC_MK_CTX(gtxFlagBase(g.ch) + ofs, (rem >> 1) != 0);
      }
    }
  }
  nq += pass1Total;
  // 2nd pass (abs_remainder of context-coded positions with |level| >= 4) and bypass-only positions
  // (dec_abs_level): each Golomb-Rice word leaves its lane already packed as one or two EP strings
  const bool hasRem = inPass1 && a >= 4;
  const bool inPass3 = active && !inPass1;
  uint32_t w0 = 0, w1 = 0;
  int cnt = 0;
  if (hasRem) {
    cnt = lsRemWords(uint32_t(a - 4) >> 1, goRicePar(sumAll - 20), w0, w1);
  } else if (inPass3) {
    const uint32_t rice = goRicePar(sumAll);
    const uint32_t pos0 = (st < 2 ? 1u : 2u) << rice;
    const uint32_t rem = a == 0 ? pos0 : (uint32_t(a) <= pos0 ? uint32_t(a) - 1u : uint32_t(a));
    cnt = lsRemWords(rem, rice, w0, w1);
  }
  const uint32_t E1 = __ballot_sync(FULL, cnt >= 1), E2 = __ballot_sync(FULL, cnt == 2);
  if (cnt) {
    const int at = nq + __popc(E1 & above) + __popc(E2 & above);
    queue[at] = w0;
    if (cnt == 2) queue[at + 1] = w1;
  }
  nq += __popc(E1) + __popc(E2);
  // signs: first coded non-zero is the pattern's MSB
  uint32_t numSigns = __popc(Z);
  uint32_t mySign = (active && nz && v < 0) ? (1u << __popc(Z & ((1u << lane) - 1u))) : 0u;
  uint32_t pattern = __reduce_or_sync(FULL, mySign);
  (void)Neg;
  if (signHiding && Z) {
    const int lastNZ = 31 - __clz(Z), firstNZ = __ffs(Z) - 1;
    if (lastNZ - firstNZ >= 4) {  // SBH_THRESHOLD, common_def.hpp:96
      numSigns--;
      pattern >>= 1;
    }
  }
  if (lane == 0) queue[nq] = CABAC_MK_EP(pattern, numSigns);
  nq += 1;
  remRegBins -= pass1Total;
  (void)myBins;
  if (depQuant) state = stateEnd;
  __syncwarp();
  return nq;
}

// last_sig_coeff as command words, cabac_writer.cpp:2639-2722 (same words on every lane)
__device__ __forceinline__ int pushLastSigCoeff(uint32_t *q, const TuGeom &g, int posX, int posY, int lane) {
  // last_sig_coeff_{x,y}_prefix as unary context-coded bins, then the two bypass suffixes
  // (cabac_writer.cpp:2812-2860): at most 10 + 10 + 2 words, one lane each
  const uint32_t gx = groupIdx(posX), gy = groupIdx(posY);
  const int bx = lastXBase(g.ch) + g.lastOffsetX, by = lastYBase(g.ch) + g.lastOffsetY;
  const uint32_t nX = gx + (gx < uint32_t(g.maxLastX) ? 1u : 0u), nY = gy + (gy < uint32_t(g.maxLastY) ? 1u : 0u);
  const uint32_t i = uint32_t(lane);
  if (i < nX) {
    q[i] = CABAC_MK_CTX(bx + (i >> g.lastShiftX), i < gx);
  } else if (i < nX + nY) {
    const uint32_t k = i - nX;
    q[i] = CABAC_MK_CTX(by + (k >> g.lastShiftY), k < gy);
  }
  int n = int(nX + nY);
  if (gx > 3) {
    const uint32_t nb = (gx - 2) >> 1;
    if (lane == 0) q[n] = CABAC_MK_EP((posX - minInGroup(gx)) & ((1u << nb) - 1u), nb);
    n++;
  }
  if (gy > 3) {
    const uint32_t nb = (gy - 2) >> 1;
    if (lane == 0) q[n] = CABAC_MK_EP((posY - minInGroup(gy)) & ((1u << nb) - 1u), nb);
    n++;
  }
  return n;
}

// Exclusive prefix sum over the 16 lanes of a coefficient group (lanes >= 16 contribute 0 and
// receive the total); `total` is the group's sum on every lane.
__device__ __forceinline__ int cgExclusiveScan(int v, int lane, int &total) {
  int inc = v;
#pragma unroll
  for (int o = 1; o < 16; o <<= 1) {
    const int t = __shfl_up_sync(FULL, inc, o);
    if ((lane & 15) >= o) inc += t;
  }
  total = __shfl_sync(FULL, inc, 15);
  return inc - v;
}

// One coefficient group of a transform-skip block as command words
// (residual_coding_subblockTS, cabac_writer.cpp:2924-3062).  The encoder knows every level up
// front, so lane k binarises scan position minPos + k on its own: contexts come from the original
// left / above neighbours, and the only serial quantities - the context-coded bin budget that ends
// pass 1 / pass 2 and each word's slot in the queue - are prefix sums over the lanes.
// Returns the word count (uniform).
__device__ __forceinline__ int produceTsCG(uint32_t *q, const int16_t *tile, const TuGeom &g, bool bdpcm,
                                           const uint16_t *cgMask, int s, int &remBins, uint64_t &sigGroups,
                                           int lane) {
  int n = 0;
  const int cgSize = 1 << g.log2cg;
  const int lastSubSet = g.numCG - 1;
  const uint32_t c = __ldg(&g_rom.cgScan[g.lwig][g.lhig][s]);
  const int cgx = c & 15, cgy = c >> 4, cgPos = cgy * g.wig + cgx;
  const bool sig = cgMask[s] != 0;
  if (sig) sigGroups |= 1ull << cgPos;
  const int left = cgx > 0 ? int((sigGroups >> (cgPos - 1)) & 1) : 0;
  const int above = cgy > 0 ? int((sigGroups >> (cgPos - g.wig)) & 1) : 0;
  const uint64_t others = sigGroups & ~(1ull << lastSubSet);
  if (s != lastSubSet || others != 0) {
    if (lane == 0) q[0] = CABAC_MK_CTX(R_TS_SIG_GROUP + left + above, sig);
    n = 1;
    if (!sig) return n;
  }
  const bool active = lane < cgSize;
  const int k = lane & (cgSize - 1);
  int x, y, l, a;
  scanXY(g, (s << g.log2cg) + k, x, y);
  const int v = active ? int(tile[(y << g.log2tw) + x]) : 0;
  neighTS(tile, g, x, y, l, a);
  const int absV = abs(v);
  const int mod = modCoeffTS(l, a, absV, bdpcm);  // level after prediction from the neighbours
  const uint32_t nzMask = uint32_t(cgMask[s]);
  // ---- pass 1: sig, sign, gt1, par ----
  const bool sigCoded = active && ((nzMask & ((1u << k) - 1u)) != 0 || k != cgSize - 1);
  const int gt1 = mod > 1;
  const int c1 = active ? int(sigCoded) + (v ? 2 + gt1 : 0) : 0;
  int tot1, tot2, tot3;
  const int ex1 = cgExclusiveScan(c1, lane, tot1);
  const bool inP1 = active && remBins - ex1 >= 4;
  const uint32_t p1Mask = __ballot_sync(FULL, inP1);  // a prefix of the group's positions
  const int nP1 = __popc(p1Mask);
  if (inP1) {
    uint32_t *w = q + n + ex1;
    const int nb = (l != 0) + (a != 0);
    if (sigCoded) *w++ = CABAC_MK_CTX(R_TS_SIG_FLAG + nb, v != 0);
    if (v) {
      *w++ = CABAC_MK_CTX(signCtxTS(l, a, bdpcm), v < 0);
      *w++ = CABAC_MK_CTX(R_TS_LRG1_FLAG + (bdpcm ? 3 : nb), gt1);
      if (gt1) *w++ = CABAC_MK_CTX(R_TS_PAR_FLAG, uint32_t(mod - 2) & 1u);
    }
  }
  const int used1 = nP1 == cgSize ? tot1 : __shfl_sync(FULL, ex1, nP1 & 31);  // bins of positions 0..nP1-1
  n += used1;
  remBins -= used1;
  // ---- pass 2: greater-than-2/4/6/8 flags ----
  const int c2 = active ? min(mod >> 1, 4) : 0;
  const int ex2 = cgExclusiveScan(c2, lane, tot2);
  const bool inP2 = active && remBins - ex2 >= 4;
  const int nP2 = __popc(__ballot_sync(FULL, inP2));
  if (inP2) {
    uint32_t *w = q + n + ex2;
    for (int cutoff = 2; cutoff <= 8 && mod >= cutoff; cutoff += 2)
      *w++ = CABAC_MK_CTX(R_TS_GTX_FLAG + (cutoff >> 1), mod >= cutoff + 2);
  }
  const int used2 = nP2 == cgSize ? tot2 : __shfl_sync(FULL, ex2, nP2 & 31);
  n += used2;
  remBins -= used2;
  // ---- pass 3: bypass remainders (+ bypass signs of positions pass 1 did not reach) ----
  const int cutoff = k < nP2 ? 10 : (k < nP1 ? 2 : 0);
  const int absLevel = cutoff ? mod : (bdpcm ? mod : absV);  // modCoeffTS(l, a, |v|, bdpcm || !cutoff)
  uint32_t w0 = 0, w1 = 0;
  int cnt = 0;
  bool signWord = false;
  if (active && absLevel >= cutoff) {
    const uint32_t rem = k < nP1 ? uint32_t(absLevel - cutoff) >> 1 : uint32_t(absLevel);
    cnt = lsRemWords(rem, 1, w0, w1);
    signWord = absLevel && k >= nP1;
  }
  const int c3 = cnt + int(signWord);
  const int ex3 = cgExclusiveScan(c3, lane, tot3);
  if (c3) {
    uint32_t *w = q + n + ex3;
    *w++ = w0;
    if (cnt == 2) *w++ = w1;
    if (signWord) *w++ = CABAC_MK_EP(v < 0, 1);
  }
  n += tot3;
  return n;
}

// ---------------------------------------------------------------------------
// K4a: fused lockstep encoder (produce + drain in one kernel, one warp per substream).
// Chosen for batches with few, long substreams (see launchEncodeResidual).
// ---------------------------------------------------------------------------
struct __align__(16) FusedEncWarpSmem {
  uint4 ctx[RCTX];
  uint8_t ring[STAGE_BYTES];
  int32_t raw[32 * 32];  // cp.async landing tile of the next block
  int16_t tile[32 * 32];
  uint32_t queue[QUEUE_WORDS];
  uint16_t cgMask[64];
};

// kMode FUSED_RESOLVE : probabilities of a batch are resolved lane-parallel (lsResolveQueue) before the
//                       serial walk, which then only carries the interval recurrence.
//       FUSED_LEGACY  : lsDrain codes the raw words (model + interval on the serial path).
//       FUSED_ESTIMATE: the RDO bit estimator (CABACWriter on BitEstimator_Std): the same binarisation
//                       and model walk, no interval coder; 15-bit fixed-point bits per substream
//                       (estBits) and the running total after every TU (estPerTu).  out / outOffsets /
//                       outSizes are unused.
//       FUSED_WS      : warp-specialised: WS_PRODUCERS producer warps per CTA binarise and resolve; ONE
//                       walker warp runs the range recurrence of all of them, one per lane; the
//                       producers then fold their words into `low` in lane-parallel groups
//                       (cabac_lockstep_enc.cuh, "Warp-specialised fused encoder").
enum { FUSED_LEGACY = 0, FUSED_RESOLVE = 1, FUSED_ESTIMATE = 2, FUSED_WS = 3 };
constexpr int WS_THREADS = (WS_PRODUCERS + 1) * 32;
#ifndef CABAC_WS_MIN_CTAS
#define CABAC_WS_MIN_CTAS 1
#endif
constexpr int WS_MIN_CTAS = CABAC_WS_MIN_CTAS;
// Producer warp of the warp-specialised kernel: shared memory per substream in flight is what limits
// the producers per SM, so a context is 4 bytes of state here; what depends on the context index alone
// (the window shifts and the "bin == 1" increment, y/z/w of ctx_entry) is one table per CTA.
struct __align__(16) WsEncWarpSmem {
  uint32_t ctx[(RCTX + 3) & ~3];
  uint8_t ring[STAGE_BYTES];
  int32_t raw[32 * 32];
  int16_t tile[32 * 32];
  uint32_t queue[QUEUE_WORDS];
  uint16_t cgMask[64];
};
constexpr size_t WS_RATES_BYTES = RCTX * sizeof(uint4);
constexpr size_t WS_SMEM_BYTES = WS_PRODUCERS * (sizeof(WsEncWarpSmem) + sizeof(WsChannel)) + WS_RATES_BYTES;

template <bool kWs>
__device__ __forceinline__ auto &fusedEncSmem(int wib) {
  if constexpr (kWs) {
    extern __shared__ __align__(16) unsigned char wsDynSmem[];
    return reinterpret_cast<WsEncWarpSmem *>(wsDynSmem)[wib];
  } else {
    __shared__ FusedEncWarpSmem smem[WARPS_PER_BLOCK];
    return smem[wib];
  }
}

// contexts of a warp-specialised producer: the state word only (x of ctx_entry)
__device__ inline void loadCtxStates(uint32_t *ctx, const cabac_substream_params &p, const cabac_ctx_model *src,
                                     int lane) {
  const int qp = clipQp(p.qp);
  for (int k = lane; k < RCTX; k += 32) {
    const int id = rctx_inv(k);
    if (src) {
      const cabac_ctx_model m = src[id];
      ctx[k] = ctx_entry(m.state0, m.state1, 0).x;
    } else {
      ctx[k] = ctx_entry_init(qp, CABAC_CTX_INIT_TABLE[p.init_id][id], 0).x;
    }
  }
  __syncwarp();
}
__device__ inline void storeCtxStates(const uint32_t *ctx, const cabac_substream_params &p, const cabac_ctx_model *src,
                                      cabac_ctx_model *dst, int lane) {
  if (!dst) return;
  __syncwarp();
  const int qp = clipQp(p.qp);
  for (int id = lane; id < CABAC_NUM_CTX; id += 32) {
    cabac_ctx_model m;
    if (is_residual_ctx(id)) {
      const uint32_t x = ctx[rctx(id)];
      m.state0 = uint16_t(x >> 16);
      m.state1 = uint16_t(x & 0xFFFFu);
    } else if (src) {
      m = src[id];
    } else {
      const uint32_t w = model_init(qp, CABAC_CTX_INIT_TABLE[p.init_id][id], CABAC_CTX_INIT_TABLE[3][id]);
      m.state0 = uint16_t(model_state0(w));
      m.state1 = uint16_t(model_state1(w));
    }
    dst[id] = m;
  }
}

// kC16 (int16 slab, cabac_*_i16) is compile-time like the mode: as a run-time argument it cost the int32 headline
// path 2.4 ms of 22.8 on the bench batch.
template <int kMode, bool kC16 = false>
__global__ void __launch_bounds__(kMode == FUSED_WS ? WS_THREADS : WARPS_PER_BLOCK * 32,
                                  kMode == FUSED_WS ? WS_MIN_CTAS : ENC_MIN_CTAS)
encode_residual_kernel(const cabac_tu_desc *__restrict__ tus, const uint64_t *__restrict__ tuOffsets,
                       const cabac_substream_params *__restrict__ params, size_t nSub,
                       const int32_t *__restrict__ coeffs, const cabac_ctx_model *__restrict__ ctxIn,
                       uint8_t *__restrict__ out, const uint64_t *__restrict__ outOffsets,
                       uint32_t *__restrict__ outSizes, uint32_t *__restrict__ numBins,
                       uint8_t *__restrict__ tuFlags, cabac_ctx_model *__restrict__ ctxOut,
                       uint32_t *__restrict__ status, const uint32_t *__restrict__ order, uint32_t *cursor,
                       uint64_t *__restrict__ estBits, uint64_t *__restrict__ estPerTu,
                       uint32_t *__restrict__ ctxBins, const WppArgs wpp) {
  constexpr bool c16 = kC16;
  constexpr bool kResolve = kMode == FUSED_RESOLVE;
  constexpr bool kEstimate = kMode == FUSED_ESTIMATE;
  constexpr bool kWs = kMode == FUSED_WS;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  WsChannel *chan = nullptr;
  uint4 *wsRates = nullptr;
  uint32_t chHead = 0, drained = 0;
  if constexpr (kWs) {
    WsChannel *chans = reinterpret_cast<WsChannel *>(&fusedEncSmem<true>(WS_PRODUCERS));
    if (threadIdx.x < WS_PRODUCERS) chans[threadIdx.x].head = chans[threadIdx.x].tail = chans[threadIdx.x].want = 0;
    wsRates = reinterpret_cast<uint4 *>(chans + WS_PRODUCERS);
    for (int k = threadIdx.x; k < RCTX; k += WS_THREADS) {
      uint4 e = ctx_entry(0, 0, CABAC_CTX_INIT_TABLE[3][rctx_inv(k)]);
      e.x = 0;
      wsRates[k] = e;
    }
    __syncthreads();
    if (wib == WS_PRODUCERS) {
      wsWalk(chans, lane);
      return;
    }
    chan = chans + wib;
  }
  auto &sm = fusedEncSmem<kWs>(wib);
  size_t sub;
  while (nextSubstream(order, cursor, nSub, lane, sub)) {
    const cabac_substream_params p = params[sub];
    uint32_t st = 0;
    const cabac_ctx_model *ctxSrc = ctxSource(p, ctxIn, sub, wpp, lane, st);
    if constexpr (kWs) loadCtxStates(sm.ctx, p, ctxSrc, lane);
    else loadCtxEntries(sm.ctx, p, ctxSrc, lane);
    if constexpr (kWs) {  // the ring is empty here: every substream ends fully drained
      wsPushWord(*chan, chHead, WS_MK_CTL(WS_OP_START), lane);
      drained = chHead;
    }
    // identical on every lane; only the coder of this instantiation survives dead-code elimination
    LsEncoder enc;
    LsCoder rc;
    uint64_t est = 0;
    uint32_t estBins = 0;
    const bool wsSingle = !(p.flags & CABAC_SUB_TERMINATE);
    // words [drained, chHead) are in the ring; drain 32 at a time once the walker has passed them.
    // minFree: words the next push needs (waits for the walker only when the ring lacks that room)
    auto wsDrain = [&](uint32_t minFree, bool all) {
      for (;;) {
        const uint32_t pending = chHead - drained;
        const bool must = all ? pending != 0 : pending + minFree > WS_RING;
        if (!must && pending < 32u) break;
        const uint32_t m = min(32u, pending);
        uint32_t spins = 0;
        while (int32_t(wsLoad(&chan->tail) - drained) < int32_t(m)) {  // signed: the walker may not have passed START yet
          if (!must) return;
          if (spins == 0 && lane == 0) wsStore(&chan->want, 1u);
          __nanosleep(32);
          if (++spins > WS_SPIN_LIMIT) __trap();
        }
        if (spins && lane == 0) wsStore(&chan->want, 0u);
        __threadfence_block();  // tail observed above; rng[] read by the slice below
        wsDrainSlice(rc, *chan, drained, int(m), lane, wsSingle);
        drained += m;
      }
    };
    if constexpr (kEstimate) {
    } else if constexpr (kWs) {
      rc.start(sm.ring, out + outOffsets[sub], uint32_t(outOffsets[sub + 1] - outOffsets[sub]), lane, false);
      rc.trackPending = wsSingle;  // eager emission either way; the outstanding-byte count only when observable
    } else if constexpr (kResolve) {
      rc.start(sm.ring, out + outOffsets[sub], uint32_t(outOffsets[sub + 1] - outOffsets[sub]), lane,
               (p.flags & CABAC_SUB_TERMINATE) != 0);
    } else {
      enc.start();
      enc.out.ring = sm.ring;
      enc.out.dst = out + outOffsets[sub];
      enc.out.cap = uint32_t(outOffsets[sub + 1] - outOffsets[sub]);
      enc.out.n = 0;
      enc.out.flushed = 0;
      enc.out.lane = lane;
    }
    const bool depQuant = (p.flags & CABAC_SUB_DEP_QUANT) != 0;
    const bool signHiding = (p.flags & CABAC_SUB_SIGN_HIDING) != 0;
    uint64_t t = tuOffsets[sub];
    const uint64_t t1 = tuOffsets[sub + 1];
    // element size of the caller's TCoeff slab: int32 (unit.hpp:551) or int16 (cabac_*_i16)
    auto blockPtr = [&](uint64_t ofs) {
      return c16 ? reinterpret_cast<const int32_t *>(reinterpret_cast<const int16_t *>(coeffs) + ofs) : coeffs + ofs;
    };
    if (t < t1) prefetchBlock(sm.raw, blockPtr(tus[t].coeff_offset), loadDesc(tus + t), lane, c16);
    // WPP: the contexts after this substream's first CTU go to its snapshot row for the row below
    bool wppPending = wpp.sync != nullptr;
    const uint64_t tSync = wppPending ? t + wpp.sync[sub].sync_tus : 0;
    auto wppSnapshot = [&]() {
      cabac_ctx_model *row = wpp.snap + sub * CABAC_NUM_CTX;
      if constexpr (kWs) storeCtxStates(sm.ctx, p, ctxSrc, row, lane);
      else storeCtxEntries(sm.ctx, p, ctxSrc, row, lane);
      wppPublish(wpp, sub, lane);
      wppPending = false;
    };

    // per-block producer state
    bool inTu = false, tsPath = false, bdpcm = false;
    cabac_tu_desc d;
    TuGeom g;
    uint32_t gKey = ~0u;
    uint32_t res = 0;
    int lastCG = 0, scanPosLast = 0, s = 0, state = 0, remRegBins = 0, remBinsTS = 0;
    uint64_t sigGroups = 0;

    // Every iteration produces one batch of command words (block header and/or one coefficient
    // group) and drains it through the single arithmetic-coder site.
    for (;;) {
      int nq = 0;
      if (!inTu) {
        if (wppPending && t >= tSync) wppSnapshot();  // every word of the TUs before t is resolved here
        if (t == t1) break;
        d = loadDesc(tus + t);
        if (d.flags & DESC_CLAMPED) st |= CABAC_SS_BAD_DESC;
        if (geomKey(d) != gKey) {
          g = makeGeom(d);
          gKey = geomKey(d);
        }
        const bool ts = (d.flags & CABAC_TU_TS) != 0;
        tsPath = ts && !(p.flags & CABAC_SUB_TSRC_DISABLED);
        bdpcm = (d.flags & CABAC_TU_BDPCM) != 0;
        res = ts ? CABAC_TUR_TS : 0;
        if (!takeBlock(sm.tile, sm.raw, g, lane, c16)) st |= CABAC_SS_COEFF_RANGE;
        if (t + 1 < t1) {
          const cabac_tu_desc dn = loadDesc(tus + t + 1);
          prefetchBlock(sm.raw, blockPtr(dn.coeff_offset), dn, lane, c16);
        }
        // ---- pre-scan: non-zero mask per coefficient group, last position ----
        lastCG = -1;
        const int cgSize = 1 << g.log2cg;
        const int perIter = 32 >> g.log2cg;  // coefficient groups handled per iteration
        for (int s0 = 0; s0 < g.numCG; s0 += perIter) {
          const int sc = s0 + (lane >> g.log2cg), pl = lane & (cgSize - 1);
          bool nz = false;
          if (sc < g.numCG) {
            int x, y;
            scanXY(g, (sc << g.log2cg) + pl, x, y);
            nz = sm.tile[(y << g.log2tw) + x] != 0;
          }
          const uint32_t b = __ballot_sync(FULL, nz);
          if (pl == 0 && sc < g.numCG)
            sm.cgMask[sc] = uint16_t((b >> (lane & ~(cgSize - 1))) & ((1u << cgSize) - 1u));
          if (b) lastCG = s0 + ((31 - __clz(b)) >> g.log2cg);
        }
        __syncwarp();
        if (lastCG < 0) {  // "Coefficient coding called for empty TU", cabac_writer.cpp:2457
          st |= CABAC_SS_EMPTY_TU;
          if (tuFlags && lane == 0) tuFlags[t] = uint8_t(res);
          if constexpr (kEstimate)
            if (estPerTu && lane == 0) estPerTu[t] = est;
          t++;
          continue;
        }
        // transform_skip_flag (ts_flag, cabac_writer.cpp:2527-2535)
        if (d.flags & CABAC_TU_TS_ALLOWED) sm.queue[nq++] = CABAC_MK_CTX(R_TS_FLAG + (g.comp ? 1 : 0), ts);
        if (tsPath) {
          s = 0;
          remBinsTS = ((g.w * g.h) * 7) >> 2;
        } else {
          scanPosLast = (lastCG << g.log2cg) + (31 - __clz(uint32_t(sm.cgMask[lastCG])));
          if (!ts && g.w >= 4 && g.h >= 4) {  // CUCtx side effects, cabac_writer.cpp:2460-2478
            const int maxLfnstPos = ((g.h == 4 && g.w == 4) || (g.h == 8 && g.w == 8)) ? 7 : 15;
            if (scanPosLast > maxLfnstPos) res |= CABAC_TUR_VIOLATES_LFNST;
            if (scanPosLast >= 1) res |= CABAC_TUR_LFNST_LAST_POS;
          }
          if (g.comp == 0 && !ts && scanPosLast >= 1) res |= CABAC_TUR_MTS_LAST_POS;
          int x, y;
          scanXY(g, scanPosLast, x, y);
          nq += pushLastSigCoeff(sm.queue + nq, g, x, y, lane);
          state = 0;
          remRegBins = regBinLimit(g);
          s = lastCG;
        }
        sigGroups = 0;
        inTu = true;
        __syncwarp();
      }
      bool endTu;
      if (tsPath) {
        nq += produceTsCG(sm.queue + nq, sm.tile, g, bdpcm, sm.cgMask, s, remBinsTS, sigGroups, lane);
        s++;
        endTu = s == g.numCG;
      } else {
        const uint32_t c = __ldg(&g_rom.cgScan[g.lwig][g.lhig][s]);
        const int cgx = c & 15, cgy = c >> 4, cgPos = cgy * g.wig + cgx;
        const bool sig = sm.cgMask[s] != 0;
        // initSubblock, context_modelling.cpp:81-106
        if (sig) sigGroups |= 1ull << cgPos;
        const bool zeroedOut = g.sbtZero && ((g.h == 32 && cgy >= (16 >> g.log2cgh)) ||
                                             (g.w == 32 && cgx >= (16 >> g.log2cgw)));
        if (!zeroedOut) {
          const bool codeSigGroup = s != lastCG && s != 0;
          const int right = (cgx + 1 < g.wig) ? int((sigGroups >> (cgPos + 1)) & 1) : 0;
          const int lower = (cgy + 1 < g.hig) ? int((sigGroups >> (cgPos + g.wig)) & 1) : 0;
          const int sigGroupCtx = R_SIG_GROUP + 2 * g.ch + (right | lower);
          if (codeSigGroup && !sig) {
            sm.queue[nq++] = CABAC_MK_CTX(sigGroupCtx, 0);
          } else {
            nq += binariseCG(sm.queue + nq, sm.tile, g, lane, s, scanPosLast, codeSigGroup, sigGroupCtx, depQuant,
                             signHiding, state, remRegBins);
            if (sig && g.comp == 0 && (cgy > 3 || cgx > 3)) res |= CABAC_TUR_VIOLATES_MTS;
          }
        }
        s--;
        endTu = s < 0;
      }
      if (endTu) {
        if (tuFlags && lane == 0) tuFlags[t] = uint8_t(res);
        t++;
        inTu = false;
      }
      __syncwarp();
      if constexpr (kEstimate) {
        est += lsEstimateQueue(sm.ctx, sm.queue, nq, lane, CABAC_FRAC_BITS, estBins,
                               ctxBins ? ctxBins + size_t(sub) * CABAC_NUM_CTX : nullptr);
        if (endTu && estPerTu && lane == 0) estPerTu[t - 1] = est;
      } else if constexpr (kWs) {
        rc.nbins += lsResolveQueueStates(sm.ctx, wsRates, sm.queue, nq, lane);
        wsDrain(uint32_t(nq), false);
        if (nq) wsPush(*chan, chHead, sm.queue, nq, lane);
      } else if constexpr (kResolve) {
        rc.nbins += lsResolveQueue<false>(sm.ctx, sm.queue, nq, lane);
        lsWalkResolved(rc, sm.queue, nq);
      } else {
        lsDrain(enc, sm.ctx, sm.queue, nq);
        if (enc.out.n - enc.out.flushed >= uint32_t(STAGE_BYTES / 2)) enc.out.flush(false);
      }
      __syncwarp();
    }
    uint32_t total, cap, nb;
    if constexpr (kEstimate) {
      total = cap = 0;
      nb = estBins;
      if (lane == 0 && estBits) estBits[sub] = est;
    } else if constexpr (kWs) {
      // a no-op word (EP, 0 bins) whose recorded range is the one finish() continues from
      wsDrain(1, false);
      wsPushWord(*chan, chHead, uint32_t(CABAC_CMD_EP) << 29, lane);
      wsDrain(0, true);
      rc.range = chan->rng[(chHead - 1u) & (WS_RING - 1u)];
      total = rc.finish((p.flags & CABAC_SUB_TERMINATE) != 0, (p.flags & CABAC_SUB_ALIGN) != 0);
      cap = rc.cap;
      nb = rc.nbins;
    } else if constexpr (kResolve) {
      total = rc.finish((p.flags & CABAC_SUB_TERMINATE) != 0, (p.flags & CABAC_SUB_ALIGN) != 0);
      cap = rc.cap;
      nb = rc.nbins;
    } else {
      enc.finish((p.flags & CABAC_SUB_TERMINATE) != 0, (p.flags & CABAC_SUB_ALIGN) != 0);
      enc.out.flush(true);
      total = enc.out.n;
      cap = enc.out.cap;
      nb = enc.nbins;
    }
    if (lane == 0) {
      if (total > cap) st |= CABAC_SS_OUT_OVERFLOW;
      if constexpr (!kEstimate) outSizes[sub] = total;
      if (numBins) numBins[sub] = nb;
      if (status) status[sub] = st;
    }
    if (wppPending) wppSnapshot();
    if constexpr (kWs) storeCtxStates(sm.ctx, p, ctxSrc, ctxOut ? ctxOut + sub * CABAC_NUM_CTX : nullptr, lane);
    else storeCtxEntries(sm.ctx, p, ctxSrc, ctxOut ? ctxOut + sub * CABAC_NUM_CTX : nullptr, lane);
    __syncwarp();
  }
  if constexpr (kWs) {  // nobody waits for EXIT to be taken: mark the channel urgent for good
    if (lane == 0) wsStore(&chan->want, 1u);
    wsPushWord(*chan, chHead, WS_MK_CTL(WS_OP_EXIT), lane);
  }
}

// ---------------------------------------------------------------------------
// K4b: two-stage encoder, stage P (binarise + resolve probabilities -> command words in HBM)
// ---------------------------------------------------------------------------
// kStream = false: the substream is its TU list (cabac_encode_residual).
// kStream = true : the substream is a recorded command stream (cabac_encode_stream): BinEncIf calls
//                  with raw context ids, and CABAC_CMD_TU markers that each code the next TU of the
//                  substream's list (CABACWriter::residual_coding in the middle of other syntax).
// kFused   = true : the resolved words of every batch are walked at once by the warp's own LsCoder
//                   (few, long substreams: slices, frames) and the bytes leave through the ring;
//                   cmds / cmdOffsets / cmdCounts are unused.
template <bool kStream, bool kFused>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
produce_kernel(const uint32_t *__restrict__ rawCmds, const uint64_t *__restrict__ rawOffsets,
               const cabac_tu_desc *__restrict__ tus, const uint64_t *__restrict__ tuOffsets,
               const cabac_substream_params *__restrict__ params, size_t nSub,
               const int32_t *__restrict__ coeffs, const cabac_ctx_model *__restrict__ ctxIn,
               uint32_t *__restrict__ cmds, const uint64_t *__restrict__ cmdOffsets,
               uint32_t *__restrict__ cmdCounts, uint8_t *__restrict__ tuFlags,
               cabac_ctx_model *__restrict__ ctxOut, uint32_t *__restrict__ status,
               const uint32_t *__restrict__ order, uint32_t *cursor, uint8_t *__restrict__ out,
               const uint64_t *__restrict__ outOffsets, uint32_t *__restrict__ outSizes,
               uint32_t *__restrict__ numBins, const WppArgs wpp, const int c16) {
  typedef EncWarpSmemT<kStream ? CABAC_NUM_CTX : RCTX> Smem;
  __shared__ Smem smem[WARPS_PER_BLOCK];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  Smem &sm = smem[wib];
  size_t sub;
  while (nextSubstream(order, cursor, nSub, lane, sub)) {
    const cabac_substream_params p = params[sub];
    uint32_t st = 0;
    const cabac_ctx_model *ctxSrc = ctxSource(p, ctxIn, sub, wpp, lane, st);
    if (kStream)
      loadCtxEntriesFull(sm.ctx, p, ctxSrc, lane);
    else
      loadCtxEntries(sm.ctx, p, ctxSrc, lane);
    uint64_t rc = kStream ? rawOffsets[sub] : 0;
    const uint64_t rc1 = kStream ? rawOffsets[sub + 1] : 0;
    uint32_t *cmdBase = kFused ? nullptr : cmds + cmdOffsets[sub];  // this substream's resolved command words
    uint32_t cmdN = 0;
    LsCoder coder;
    if constexpr (kFused)
      coder.start(sm.ring, out + outOffsets[sub], uint32_t(outOffsets[sub + 1] - outOffsets[sub]), lane,
               (p.flags & CABAC_SUB_TERMINATE) != 0);
    const bool depQuant = (p.flags & CABAC_SUB_DEP_QUANT) != 0;
    const bool signHiding = (p.flags & CABAC_SUB_SIGN_HIDING) != 0;
    uint64_t t = tus ? tuOffsets[sub] : 0;
    const uint64_t t1 = tus ? tuOffsets[sub + 1] : 0;
    // WPP: snapshot after the first CTU (sync_cmds recorded words in stream mode, sync_tus TUs otherwise)
    bool wppPending = wpp.sync != nullptr;
    const uint64_t syncAt = !wppPending ? 0 : (kStream ? rc + wpp.sync[sub].sync_cmds : t + wpp.sync[sub].sync_tus);
    auto wppSnapshot = [&]() {
      cabac_ctx_model *row = wpp.snap + sub * CABAC_NUM_CTX;
      if (kStream) storeCtxEntriesFull(sm.ctx, row, lane);
      else storeCtxEntries(sm.ctx, p, ctxSrc, row, lane);
      wppPublish(wpp, sub, lane);
      wppPending = false;
    };

    // per-block producer state
    bool inTu = false, tsPath = false, bdpcm = false;
    cabac_tu_desc d;
    TuGeom g;
    uint32_t gKey = ~0u;
    uint32_t res = 0;
    int lastCG = 0, scanPosLast = 0, s = 0, state = 0, remRegBins = 0, remBinsTS = 0;
    uint64_t sigGroups = 0;

    // Every iteration produces one batch of command words (block header and/or one coefficient
    // group, or a run of recorded bin commands), resolves its probabilities and appends it to the
    // substream's slab.
    for (;;) {
      int nq = 0;
      if (!inTu) {
        if (wppPending && (kStream ? rc : t) >= syncAt) wppSnapshot();
        if (kStream) {
          if (rc == rc1) break;
          // up to 32 recorded commands, stopping at the next TU marker (and at the WPP sync point)
          uint64_t rcStop = rc1;
          if (kStream && wppPending && syncAt < rc1) rcStop = syncAt;
          const bool have = rc + lane < rcStop;
          const uint32_t w = have ? rawCmds[rc + lane] : (uint32_t(CABAC_CMD_TU) << 29);
          const uint32_t kind = w >> 29;
          const uint32_t stop = __ballot_sync(FULL, kind == CABAC_CMD_TU);
          const int nRaw = stop ? __ffs(int(stop)) - 1 : 32;
          if (nRaw > 0) {
            // REM -> one or two EP words; everything else is copied (CTX words keep raw context ids)
            uint32_t w0 = w, w1 = 0;
            int cnt = lane < nRaw ? 1 : 0;
            if (cnt && kind == CABAC_CMD_REM) {
              const uint32_t rice = (w >> 24) & 31u;
              if (rice > 12u) {  // suffix words must fit 24 bins
                st |= CABAC_SS_BAD_CMD;
                w0 = CABAC_MK_EP(0, 0);
              } else {
                cnt = lsRemWords(w & 0xFFFFFFu, rice, w0, w1);
              }
            } else if (cnt && (kind > CABAC_CMD_ALIGN || (kind == CABAC_CMD_EP && ((w >> 24) & 31u) > 24u) ||
                               (kind == CABAC_CMD_CTX && (w & 511u) >= CABAC_NUM_CTX))) {
              st |= CABAC_SS_BAD_CMD;
              w0 = CABAC_MK_EP(0, 0);
            }
            const uint32_t below = (1u << lane) - 1u;
            const uint32_t E1 = __ballot_sync(FULL, cnt >= 1), E2 = __ballot_sync(FULL, cnt == 2);
            if (cnt) {
              const int at = __popc(E1 & below) + __popc(E2 & below);
              sm.queue[at] = w0;
              if (cnt == 2) sm.queue[at + 1] = w1;
            }
            nq = __popc(E1) + __popc(E2);
            st = __reduce_or_sync(FULL, st);
            rc += uint64_t(nRaw);
            __syncwarp();
            const uint32_t nb = lsResolveQueue<false>(sm.ctx, sm.queue, nq, lane);
            if constexpr (kFused) {
              coder.nbins += nb;
              lsWalkResolvedAny(coder, sm.queue, nq);
            } else {
              for (int i = lane; i < nq; i += 32) cmdBase[cmdN + i] = sm.queue[i];
              cmdN += uint32_t(nq);
            }
            __syncwarp();
            continue;
          }
          rc++;  // a TU marker: code the next transform unit of the substream's list
          if (t == t1) {
            st |= CABAC_SS_BAD_CMD;
            continue;
          }
        } else if (t == t1) {
          break;
        }
        d = loadDesc(tus + t);
        if (d.flags & DESC_CLAMPED) st |= CABAC_SS_BAD_DESC;
        if (geomKey(d) != gKey) {
          g = makeGeom(d);
          gKey = geomKey(d);
        }
        const bool ts = (d.flags & CABAC_TU_TS) != 0;
        tsPath = ts && !(p.flags & CABAC_SUB_TSRC_DISABLED);
        bdpcm = (d.flags & CABAC_TU_BDPCM) != 0;
        res = ts ? CABAC_TUR_TS : 0;
        if (!loadTile(sm.tile, c16 ? reinterpret_cast<const int32_t *>(reinterpret_cast<const int16_t *>(coeffs) + d.coeff_offset)
                                   : coeffs + d.coeff_offset, g, lane, c16))
          st |= CABAC_SS_COEFF_RANGE;
        // ---- pre-scan: non-zero mask per coefficient group, last position ----
        lastCG = -1;
        const int cgSize = 1 << g.log2cg;
        const int perIter = 32 >> g.log2cg;  // coefficient groups handled per iteration
        for (int s0 = 0; s0 < g.numCG; s0 += perIter) {
          const int sc = s0 + (lane >> g.log2cg), pl = lane & (cgSize - 1);
          bool nz = false;
          if (sc < g.numCG) {
            int x, y;
            scanXY(g, (sc << g.log2cg) + pl, x, y);
            nz = sm.tile[(y << g.log2tw) + x] != 0;
          }
          const uint32_t b = __ballot_sync(FULL, nz);
          if (pl == 0 && sc < g.numCG)
            sm.cgMask[sc] = uint16_t((b >> (lane & ~(cgSize - 1))) & ((1u << cgSize) - 1u));
          if (b) lastCG = s0 + ((31 - __clz(b)) >> g.log2cg);
        }
        __syncwarp();
        if (lastCG < 0) {  // "Coefficient coding called for empty TU", cabac_writer.cpp:2457
          st |= CABAC_SS_EMPTY_TU;
          if (tuFlags && lane == 0) tuFlags[t] = uint8_t(res);
          t++;
          continue;
        }
        // transform_skip_flag (ts_flag, cabac_writer.cpp:2527-2535)
        if (d.flags & CABAC_TU_TS_ALLOWED) sm.queue[nq++] = CABAC_MK_CTX(R_TS_FLAG + (g.comp ? 1 : 0), ts);
        if (tsPath) {
          s = 0;
          remBinsTS = ((g.w * g.h) * 7) >> 2;
        } else {
          scanPosLast = (lastCG << g.log2cg) + (31 - __clz(uint32_t(sm.cgMask[lastCG])));
          if (!ts && g.w >= 4 && g.h >= 4) {  // CUCtx side effects, cabac_writer.cpp:2460-2478
            const int maxLfnstPos = ((g.h == 4 && g.w == 4) || (g.h == 8 && g.w == 8)) ? 7 : 15;
            if (scanPosLast > maxLfnstPos) res |= CABAC_TUR_VIOLATES_LFNST;
            if (scanPosLast >= 1) res |= CABAC_TUR_LFNST_LAST_POS;
          }
          if (g.comp == 0 && !ts && scanPosLast >= 1) res |= CABAC_TUR_MTS_LAST_POS;
          int x, y;
          scanXY(g, scanPosLast, x, y);
          nq += pushLastSigCoeff(sm.queue + nq, g, x, y, lane);
          state = 0;
          remRegBins = regBinLimit(g);
          s = lastCG;
        }
        sigGroups = 0;
        inTu = true;
        __syncwarp();
      }
      bool endTu;
      if (tsPath) {
        nq += produceTsCG(sm.queue + nq, sm.tile, g, bdpcm, sm.cgMask, s, remBinsTS, sigGroups, lane);
        s++;
        endTu = s == g.numCG;
      } else {
        const uint32_t c = __ldg(&g_rom.cgScan[g.lwig][g.lhig][s]);
        const int cgx = c & 15, cgy = c >> 4, cgPos = cgy * g.wig + cgx;
        const bool sig = sm.cgMask[s] != 0;
        // initSubblock, context_modelling.cpp:81-106
        if (sig) sigGroups |= 1ull << cgPos;
        const bool zeroedOut = g.sbtZero && ((g.h == 32 && cgy >= (16 >> g.log2cgh)) ||
                                             (g.w == 32 && cgx >= (16 >> g.log2cgw)));
        if (!zeroedOut) {
          const bool codeSigGroup = s != lastCG && s != 0;
          const int right = (cgx + 1 < g.wig) ? int((sigGroups >> (cgPos + 1)) & 1) : 0;
          const int lower = (cgy + 1 < g.hig) ? int((sigGroups >> (cgPos + g.wig)) & 1) : 0;
          const int sigGroupCtx = R_SIG_GROUP + 2 * g.ch + (right | lower);
          if (codeSigGroup && !sig) {
            sm.queue[nq++] = CABAC_MK_CTX(sigGroupCtx, 0);
          } else {
            nq += binariseCG(sm.queue + nq, sm.tile, g, lane, s, scanPosLast, codeSigGroup, sigGroupCtx, depQuant,
                             signHiding, state, remRegBins);
            if (sig && g.comp == 0 && (cgy > 3 || cgx > 3)) res |= CABAC_TUR_VIOLATES_MTS;
          }
        }
        s--;
        endTu = s < 0;
      }
      if (endTu) {
        if (tuFlags && lane == 0) tuFlags[t] = uint8_t(res);
        t++;
        inTu = false;
      }
      __syncwarp();
      const uint32_t nb = lsResolveQueue<kStream>(sm.ctx, sm.queue, nq, lane);  // the TU producers emit compact context ids
      if constexpr (kFused) {
        coder.nbins += nb;
        lsWalkResolved(coder, sm.queue, nq);
      } else {
        for (int i = lane; i < nq; i += 32) cmdBase[cmdN + i] = sm.queue[i];
        cmdN += uint32_t(nq);
      }
      __syncwarp();
    }
    if constexpr (kFused) {
      const uint32_t total = coder.finish((p.flags & CABAC_SUB_TERMINATE) != 0, (p.flags & CABAC_SUB_ALIGN) != 0);
      if (lane == 0) {
        if (total > coder.cap) st |= CABAC_SS_OUT_OVERFLOW;
        outSizes[sub] = total;
        if (numBins) numBins[sub] = coder.nbins;
        status[sub] = st;
      }
    } else if (lane == 0) {
      cmdCounts[sub] = cmdN;
      status[sub] = st;
    }
    if (wppPending) wppSnapshot();
    if (kStream)
      storeCtxEntriesFull(sm.ctx, ctxOut ? ctxOut + sub * CABAC_NUM_CTX : nullptr, lane);
    else
      storeCtxEntries(sm.ctx, p, ctxSrc, ctxOut ? ctxOut + sub * CABAC_NUM_CTX : nullptr, lane);
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------
// K4b: stage D, the interval coder.  One THREAD per substream (32 per warp, neighbours in
// the longest-first order, so lanes of a warp carry similar lengths).  Resolved command
// words are staged global -> shared with 16-byte cp.async, double buffered: per 32-word
// chunk lane L fetches quarter (L & 7) of substream 4r + (L >> 3), r = 0..7.
// ---------------------------------------------------------------------------
constexpr int RC_WARPS = 4;
constexpr int RC_ROW = 36;  // words per staged row: 32 + padding, rows stay 16-byte aligned

__device__ __forceinline__ void cpAsync16(void *smemDst, const void *gmemSrc) {
  const uint32_t d = uint32_t(__cvta_generic_to_shared(smemDst));
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmemSrc) : "memory");
}

__global__ void __launch_bounds__(RC_WARPS * 32)
range_code_kernel(const uint32_t *__restrict__ cmds, const uint64_t *__restrict__ cmdOffsets,
                  const uint32_t *__restrict__ cmdCounts, const cabac_substream_params *__restrict__ params,
                  size_t nSub, uint8_t *__restrict__ out, const uint64_t *__restrict__ outOffsets,
                  uint32_t *__restrict__ outSizes, uint32_t *__restrict__ numBins, uint32_t *__restrict__ status,
                  const uint32_t *__restrict__ order) {
  __shared__ __align__(16) uint32_t stage[RC_WARPS][2][32 * RC_ROW];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const size_t slot = (size_t(blockIdx.x) * RC_WARPS + wib) * 32 + lane;
  const bool valid = slot < nSub;
  const size_t sub = valid ? (order ? size_t(order[slot]) : slot) : 0;
  const uint32_t count = valid ? cmdCounts[sub] : 0u;
  const uint32_t *src = cmds + (valid ? cmdOffsets[sub] : 0);
  const cabac_substream_params p = valid ? params[sub] : cabac_substream_params{0, 0, 0, 0};
  RangeCoder enc;
  enc.start(out + (valid ? outOffsets[sub] : 0), valid ? uint32_t(outOffsets[sub + 1] - outOffsets[sub]) : 0u);
  const uint32_t maxCount = __reduce_max_sync(FULL, count);
  const uint32_t nChunks = (maxCount + 31u) >> 5;
  uint32_t *st = stage[wib][0];

  auto fetch = [&](uint32_t chunk) {
    uint32_t *dst = st + (chunk & 1u) * (32 * RC_ROW);
#pragma unroll
    for (int r = 0; r < 8; r++) {
      const int j = 4 * r + (lane >> 3);
      const uint32_t cj = __shfl_sync(FULL, count, j);
      const uint64_t pj = __shfl_sync(FULL, reinterpret_cast<uint64_t>(src), j);
      const uint32_t w0 = chunk * 32u + uint32_t(lane & 7) * 4u;
      if (w0 < cj) cpAsync16(dst + j * RC_ROW + (lane & 7) * 4, reinterpret_cast<const uint32_t *>(pj) + w0);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  if (nChunks) fetch(0);
  for (uint32_t c = 0; c < nChunks; c++) {
    if (c + 1 < nChunks) {
      fetch(c + 1);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncwarp();
    const uint32_t *row = st + (c & 1u) * (32 * RC_ROW) + lane * RC_ROW;
    const uint32_t lim = min(32u, maxCount - c * 32u);
#pragma unroll 1
    for (uint32_t i = 0; i < lim; i++) {
      if (c * 32u + i < count) {
        const uint32_t w = row[i];
        const uint32_t kind = w >> 29;
        if (kind <= CABAC_CMD_EP) {
          enc.step(w);  // branch-free for the two hot kinds, so mixed lanes do not serialise
        } else if (kind == CABAC_CMD_TRM) {
          enc.trm(w & 1u);
        } else {  // CABAC_CMD_ALIGN
          enc.range = 256;
        }
        enc.flushBits();
      }
    }
    __syncwarp();
  }
  if (valid) {
    const uint32_t total = enc.finish((p.flags & CABAC_SUB_TERMINATE) != 0, (p.flags & CABAC_SUB_ALIGN) != 0);
    outSizes[sub] = total;
    if (numBins) numBins[sub] = enc.nbins;
    if (total > enc.cap && status) status[sub] |= CABAC_SS_OUT_OVERFLOW;
  }
}

// ---------------------------------------------------------------------------
// K5: residual-level decoder
// ---------------------------------------------------------------------------
struct __align__(16) DecWarpSmem {
  int32_t tile[32 * 32];
  uint4 ctx[RCTX];
};

// WPP (context hand-over between CTU rows) and C16 (int16 slab) are compile-time: the plain int32 launch - the
// headline path - carries neither the per-block hand-over test nor the slab-type branch (with both as run-time
// arguments the bench batch went from 35.7 to 39.6 ms).
template <bool WPP, bool C16>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32, DEC_MIN_CTAS)
decode_residual_kernel(const uint8_t *__restrict__ in, const uint64_t *__restrict__ inOffsets,
                       const cabac_tu_desc *__restrict__ tus, const uint64_t *__restrict__ tuOffsets,
                       const cabac_substream_params *__restrict__ params, size_t nSub,
                       const cabac_ctx_model *__restrict__ ctxIn, int32_t *__restrict__ coeffs,
                       uint32_t *__restrict__ numBins, uint32_t *__restrict__ bitsRead,
                       uint8_t *__restrict__ tuFlags, cabac_ctx_model *__restrict__ ctxOut,
                       uint32_t *__restrict__ status, const uint32_t *__restrict__ order, uint32_t *cursor,
                       const WppArgs wpp) {
  __shared__ DecWarpSmem smem[WARPS_PER_BLOCK];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  DecWarpSmem &sm = smem[wib];
  size_t sub;
  while (nextSubstream(order, cursor, nSub, lane, sub)) {
    const cabac_substream_params p = params[sub];
    uint32_t wppStatus = 0;
    const cabac_ctx_model *ctxSrc;
    if (WPP)
      ctxSrc = ctxSource(p, ctxIn, sub, wpp, lane, wppStatus);
    else
      ctxSrc = ((p.flags & CABAC_SUB_CTX_IN) && ctxIn) ? ctxIn + sub * CABAC_NUM_CTX : nullptr;
    loadCtxEntries(sm.ctx, p, ctxSrc, lane);
    LsDecoder dec;  // identical on every lane
    dec.start(in + inOffsets[sub], uint32_t(inOffsets[sub + 1] - inOffsets[sub]));
    dec.status |= wppStatus;
    const uint64_t t0 = tuOffsets[sub], t1 = tuOffsets[sub + 1];
    bool wppPending = WPP && wpp.sync != nullptr;
    const uint64_t tSync = wppPending ? t0 + wpp.sync[sub].sync_tus : 0;
    for (uint64_t t = t0; t < t1; t++) {
      if (WPP && wppPending && t >= tSync) {  // contexts after the first CTU: the row below starts from them
        storeCtxEntries(sm.ctx, p, ctxSrc, wpp.snap + sub * CABAC_NUM_CTX, lane);
        wppPublish(wpp, sub, lane);
        wppPending = false;
      }
      const cabac_tu_desc d = loadDesc(tus + t);
      if (d.flags & DESC_CLAMPED) dec.status |= CABAC_SS_BAD_DESC;
      const TuGeom g = makeGeom(d);
      const int n = g.tw * g.th;
      for (int i = lane; i < n; i += 32) sm.tile[i] = 0;
      __syncwarp();
      // ts_flag, cabac_reader.cpp:2737-2752
      bool ts = (d.flags & CABAC_TU_BDPCM) ? true : (d.flags & CABAC_TU_TS) != 0;
      if (d.flags & CABAC_TU_TS_ALLOWED) ts = lsDecodeBin(dec, sm.ctx + R_TS_FLAG + (g.comp ? 1 : 0)) != 0;
      uint32_t res = ts ? CABAC_TUR_TS : 0;
      if (ts && !(p.flags & CABAC_SUB_TSRC_DISABLED))
        lsResidualTS(dec, sm.ctx, sm.tile, g, (d.flags & CABAC_TU_BDPCM) != 0, lane);
      else
        res |= lsResidualRegular(dec, sm.ctx, sm.tile, g, p.flags, ts, lane);
      if (tuFlags && lane == 0) tuFlags[t] = uint8_t(res);
      __syncwarp();
      // write the block back (the zero-out region of 64-point blocks as zeros)
      const int total = g.w * g.h;
      if (C16) {  // int16 TCoeff slab (cabac_decode_residual_i16): levels are range-checked above
        int16_t *dst = reinterpret_cast<int16_t *>(coeffs) + d.coeff_offset;
        for (int i = lane; i < total; i += 32) {
          const int y = i >> g.log2w, x = i & (g.w - 1);
          dst[i] = (x < g.tw && y < g.th) ? int16_t(sm.tile[(y << g.log2tw) + x]) : int16_t(0);
        }
      } else {
        int32_t *dst = coeffs + d.coeff_offset;
        for (int i = lane; i < total; i += 32) {
          const int y = i >> g.log2w, x = i & (g.w - 1);
          dst[i] = (x < g.tw && y < g.th) ? sm.tile[(y << g.log2tw) + x] : 0;
        }
      }
      __syncwarp();
    }
    if (p.flags & CABAC_SUB_TERMINATE)
      if (!dec.decodeBinTrm()) dec.status |= CABAC_SS_NO_TERMINATE;
    dec.checkOverrun();
    if (lane == 0) {
      if (numBins) numBins[sub] = dec.nbins;
      if (bitsRead) bitsRead[sub] = dec.bitsRead();
      if (status) status[sub] = dec.status;
    }
    if (WPP && wppPending) {
      storeCtxEntries(sm.ctx, p, ctxSrc, wpp.snap + sub * CABAC_NUM_CTX, lane);
      wppPublish(wpp, sub, lane);
    }
    storeCtxEntries(sm.ctx, p, ctxSrc, ctxOut ? ctxOut + sub * CABAC_NUM_CTX : nullptr, lane);
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------
// K6/K7: substream packing (entry points + compaction)
// ---------------------------------------------------------------------------
// Single-CTA exclusive scan of the substream sizes (n_sub <= ~10^6: a few
// dozen microseconds; the sizes are 4 bytes per substream).
__global__ void __launch_bounds__(1024) entry_points_kernel(const uint32_t *__restrict__ sizes,
                                                            const uint64_t *__restrict__ capOffsets, size_t n,
                                                            uint64_t *__restrict__ entry) {
  __shared__ uint64_t warpSums[32];
  __shared__ uint64_t carry;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (size_t base = 0; base < n; base += blockDim.x) {
    const size_t i = base + threadIdx.x;
    uint64_t v = i < n ? sizes[i] : 0;
    if (capOffsets && i < n && v > capOffsets[i + 1] - capOffsets[i])
      v = capOffsets[i + 1] - capOffsets[i];  // overflowed slab
    uint64_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint64_t t = __shfl_up_sync(FULL, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) warpSums[w] = incl;
    __syncthreads();
    if (w == 0) {
      uint64_t s = warpSums[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint64_t t = __shfl_up_sync(FULL, s, o);
        if (lane >= o) s += t;
      }
      warpSums[lane] = s;  // inclusive over warps
    }
    __syncthreads();
    const uint64_t before = carry + (w ? warpSums[w - 1] : 0) + incl - v;
    if (i < n) entry[i] = before;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry = before + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) entry[n] = carry;
}

// One warp per substream: slab -> packed payload at its entry point.
__global__ void __launch_bounds__(256) pack_kernel(const uint8_t *__restrict__ out,
                                                   const uint64_t *__restrict__ outOffsets,
                                                   const uint32_t *__restrict__ sizes, size_t n,
                                                   uint8_t *__restrict__ packed,
                                                   const uint64_t *__restrict__ entry) {
  const int lane = threadIdx.x & 31;
  const size_t warp = (size_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const size_t nWarps = (size_t(gridDim.x) * blockDim.x) >> 5;
  for (size_t s = warp; s < n; s += nWarps) {
    const uint8_t *src = out + outOffsets[s];
    uint8_t *dst = packed + entry[s];
    uint32_t len = sizes[s];
    if (len > outOffsets[s + 1] - outOffsets[s]) len = uint32_t(outOffsets[s + 1] - outOffsets[s]);
    // head: bytes up to the destination's 16-byte boundary
    uint32_t head = uint32_t((16 - (reinterpret_cast<uintptr_t>(dst) & 15u)) & 15u);
    if (head > len) head = len;
    for (uint32_t i = lane; i < head; i += 32) dst[i] = src[i];
    const uint32_t body = (len - head) & ~15u;
    if ((reinterpret_cast<uintptr_t>(src + head) & 3u) == 0) {
      for (uint32_t i = 16u * lane; i < body; i += 512u) {
        const uint32_t *s4 = reinterpret_cast<const uint32_t *>(src + head + i);
        *reinterpret_cast<uint4 *>(dst + head + i) = make_uint4(s4[0], s4[1], s4[2], s4[3]);
      }
    } else {
      for (uint32_t i = 16u * lane; i < body; i += 512u) {
        uint32_t wv[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
          const uint8_t *b = src + head + i + 4 * k;
          wv[k] = uint32_t(b[0]) | (uint32_t(b[1]) << 8) | (uint32_t(b[2]) << 16) | (uint32_t(b[3]) << 24);
        }
        *reinterpret_cast<uint4 *>(dst + head + i) = make_uint4(wv[0], wv[1], wv[2], wv[3]);
      }
    }
    for (uint32_t i = head + body + lane; i < len; i += 32) dst[i] = src[i];
  }
}

// ---------------------------------------------------------------------------
// host-side launchers (called by cabac_api.cpp)
// ---------------------------------------------------------------------------
static void buildRom(ScanRom &r) {
  auto diag = [](int bw, int bh, uint8_t *xs, uint8_t *ys) {  // ScanGenerator SCAN_DIAG, rom.cpp:75-93
    int k = 0;
    for (int d = 0; d <= bw + bh - 2; d++) {
      int y = d < bh ? d : bh - 1;
      int x = d - y;
      for (; y >= 0 && x < bw; y--, x++) { xs[k] = uint8_t(x); ys[k] = uint8_t(y); k++; }
    }
  };
  uint8_t xs[64], ys[64];
  memset(&r, 0, sizeof(r));
  for (int lw = 0; lw < 4; lw++)
    for (int lh = 0; lh < 4; lh++) {
      const int w = 1 << lw, h = 1 << lh;
      diag(w, h, xs, ys);
      for (int k = 0; k < w * h; k++) {
        r.cgScan[lw][lh][k] = uint8_t(xs[k] | (ys[k] << 4));
        r.cgScanInv[lw][lh][ys[k] * w + xs[k]] = uint8_t(k);
      }
    }
  for (int lcw = 0; lcw <= 4; lcw++)
    for (int lch = 0; lcw + lch <= 4; lch++) {  // every coefficient-group shape of g_log2SbbSize (<= 16 positions)
      const int w = 1 << lcw, h = 1 << lch, sh = 5 * lcw + lch;
      diag(w, h, xs, ys);
      for (int k = 0; k < w * h; k++) {
        r.inScan[sh][k] = uint8_t(xs[k] | (ys[k] << 4));
        r.inScanInv[sh][ys[k] * w + xs[k]] = uint8_t(k);
      }
    }
}

cudaError_t uploadRom() {
  ScanRom r;
  buildRom(r);
  cudaError_t e = cudaMemcpyToSymbol(g_rom, &r, sizeof(r));
  if (e != cudaSuccess) return e;
  uint32_t recip[256] = {0};
  for (uint32_t range = 256; range <= 510; range++)
    recip[range - 256] = uint32_t(((1ull << 34) + range - 1) / range);  // ceil(2^34 / range)
  e = cudaMemcpyToSymbol(c_recip, recip, sizeof(recip));
  if (e != cudaSuccess) return e;
  return cudaMemcpyToSymbol(g_recipTab, recip, sizeof(recip));
}

const unsigned char *hostInitTable(int row) { return hosttab::CABAC_CTX_INIT_TABLE[row]; }
const unsigned int *hostFracBits() { return hosttab::CABAC_FRAC_BITS; }

// Persistent grid: as many CTAs as can be resident (or fewer when there is less work).
template <typename K>
static int persistentGrid(K kernel, size_t nSub, int device) {
  int sms = 148, perSm = 1;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, kernel, WARPS_PER_BLOCK * 32, 0) != cudaSuccess || perSm < 1)
    perSm = 1;
  static const int maxPerSm = [] {  // experiment knob: CABAC_B200_CTAS_PER_SM caps the resident CTAs per SM
    const char *e = getenv("CABAC_B200_CTAS_PER_SM");
    return e ? atoi(e) : 0;
  }();
  if (maxPerSm > 0 && perSm > maxPerSm) perSm = maxPerSm;
  const size_t blocks = (nSub + WARPS_PER_BLOCK - 1) / WARPS_PER_BLOCK;
  const size_t cap = size_t(sms) * perSm;
  return int(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
}

// cost -> longest-first order; resets the cursor.  `bytesCost` selects the decode estimate.
static cudaError_t launchOrder(cudaStream_t s, const Sched &sc, size_t nSub) {
  lpt_order_kernel<<<1, 1024, 0, s>>>(sc.cost, nSub, sc.order, sc.cursor);
  return cudaGetLastError();
}
static int costGrid(size_t n, int perBlock, int maxPerSm = 16) {
  const size_t b = (n + perBlock - 1) / perBlock, cap = size_t(148) * maxPerSm;
  return int(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace cabac
#include "cabac_sliced.cuh"
#include "cabac_decode_thread.cuh"
namespace cabac {

cudaError_t launchCtxInit(cudaStream_t s, int device, const cabac_substream_params *params, size_t n,
                          cabac_ctx_model *out) {
  (void)device;
  if (n == 0) return cudaSuccess;
  const size_t total = n * CABAC_NUM_CTX;
  const size_t wantBlocks = (total + 255) / 256;
  const int blocks = int(wantBlocks < size_t(148 * 16) ? wantBlocks : size_t(148 * 16));
  ctx_init_kernel<<<blocks, 256, 0, s>>>(params, n, out);
  return cudaGetLastError();
}

cudaError_t launchDecodeBins(cudaStream_t s, int device, const Sched &sc, const uint8_t *in,
                             const uint64_t *inOffsets, const uint32_t *cmds, const uint64_t *cmdOffsets,
                             const cabac_substream_params *params, size_t nSub, const cabac_ctx_model *ctxIn,
                             uint32_t *values, uint32_t *bitsRead, cabac_ctx_model *ctxOut, uint32_t *status) {
  if (nSub == 0) return cudaSuccess;
  cost_bytes_kernel<<<costGrid(nSub, 256), 256, 0, s>>>(cmdOffsets, nSub, sc.cost);
  cudaError_t e = launchOrder(s, sc, nSub);
  if (e != cudaSuccess) return e;
  decode_bins_kernel<<<persistentGrid(decode_bins_kernel, nSub, device), WARPS_PER_BLOCK * 32, 0, s>>>(
      in, inOffsets, cmds, cmdOffsets, params, nSub, ctxIn, values, bitsRead, ctxOut, status, sc.order, sc.cursor);
  return cudaGetLastError();
}

// Residual encode, part 1: per-substream cost and command-slab capacity, longest-first order,
// slab offsets (exclusive scan; cmdOffsets[nSub] = words the slab must hold).
cudaError_t launchEncodePrepare(cudaStream_t s, int device, const Sched &sc, const cabac_tu_desc *tus,
                                const uint64_t *tuOffsets, size_t nSub, const int32_t *coeffs,
                                const uint64_t *rawOffsets, const cabac_substream_params *params, int sampleMask) {
  (void)device;
  if (nSub == 0) return cudaSuccess;
  static const int tuConst = getenv("CABAC_B200_ENC_TU_CONST") ? atoi(getenv("CABAC_B200_ENC_TU_CONST")) : 24;
  static const int areaShift = getenv("CABAC_B200_ENC_AREA_SHIFT") ? atoi(getenv("CABAC_B200_ENC_AREA_SHIFT")) : 4;
  // many substreams are short substreams: one warp each instead of eight
  const int costThreads = nSub >= 16384 ? 32 : 256;
  cost_coeff_kernel<<<costGrid(nSub, 1, costThreads == 32 ? 32 : 16), costThreads, 0, s>>>(tus, tuOffsets, coeffs, nSub, sc.cost, sc.capWords,
                                                             rawOffsets, params, sampleMask, uint32_t(tuConst),
                                                             uint32_t(areaShift), sc.coeff16 ? 1 : 0);
  cudaError_t e = launchOrder(s, sc, nSub);
  if (e != cudaSuccess) return e;
  entry_points_kernel<<<1, 1024, 0, s>>>(sc.capWords, nullptr, nSub, sc.cmdOffsets);
  return cudaGetLastError();
}

// Part 2: stage P (binarise + resolve probabilities -> command words) and stage D (interval coder).
cudaError_t launchEncodeResidual(cudaStream_t s, int device, const Sched &sc, const cabac_tu_desc *tus,
                                 const uint64_t *tuOffsets, const cabac_substream_params *params, size_t nSub,
                                 const int32_t *coeffs, const cabac_ctx_model *ctxIn, uint8_t *out,
                                 const uint64_t *outOffsets, uint32_t *outSizes, uint32_t *numBins,
                                 uint8_t *tuFlags, cabac_ctx_model *ctxOut, uint32_t *status) {
  if (nSub == 0) return cudaSuccess;
  if (!sc.twoStage) {
    if (sc.warpSpecialised && nSub < (size_t(1) << 27)) {
      static bool optedIn[64] = {};  // the > 48 KB dynamic shared memory opt-in is per device
      if (device < 0 || device >= 64 || !optedIn[device]) {
        cudaError_t attr = cudaFuncSetAttribute(encode_residual_kernel<FUSED_WS, false>,
                                                cudaFuncAttributeMaxDynamicSharedMemorySize, int(WS_SMEM_BYTES));
        if (attr != cudaSuccess) return attr;
        attr = cudaFuncSetAttribute(encode_residual_kernel<FUSED_WS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    int(WS_SMEM_BYTES));
        if (attr != cudaSuccess) return attr;
        if (device >= 0 && device < 64) optedIn[device] = true;
      }
      int sms = 148;
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
      const size_t blocks = (nSub + WS_PRODUCERS - 1) / WS_PRODUCERS, cap = size_t(sms) * WS_MIN_CTAS;
#define WS_LAUNCH(C)                                                                                                 \
  encode_residual_kernel<FUSED_WS, C><<<unsigned(blocks > cap ? cap : blocks), WS_THREADS, WS_SMEM_BYTES, s>>>(          \
      tus, tuOffsets, params, nSub, coeffs, ctxIn, out, outOffsets, outSizes, numBins, tuFlags, ctxOut,                 \
      status ? status : sc.statusScratch, sc.wpp.sync ? nullptr : sc.order, sc.cursor, nullptr, nullptr, nullptr, sc.wpp)
      if (sc.coeff16) WS_LAUNCH(true); else WS_LAUNCH(false);
#undef WS_LAUNCH
      return cudaGetLastError();
    }
#define FUSED_LAUNCH(M, C)                                                                                            \
  encode_residual_kernel<M, C><<<persistentGrid(encode_residual_kernel<M, C>, nSub, device), WARPS_PER_BLOCK * 32, 0, s>>>( \
      tus, tuOffsets, params, nSub, coeffs, ctxIn, out, outOffsets, outSizes, numBins, tuFlags, ctxOut,                  \
      status ? status : sc.statusScratch, sc.wpp.sync ? nullptr : sc.order, sc.cursor, nullptr, nullptr, nullptr, sc.wpp)
    if (sc.fusedResolve) {
      if (sc.coeff16) FUSED_LAUNCH(FUSED_RESOLVE, true); else FUSED_LAUNCH(FUSED_RESOLVE, false);
    } else {
      if (sc.coeff16) FUSED_LAUNCH(FUSED_LEGACY, true); else FUSED_LAUNCH(FUSED_LEGACY, false);
    }
#undef FUSED_LAUNCH
    return cudaGetLastError();
  }
  produce_kernel<false, false><<<persistentGrid(produce_kernel<false, false>, nSub, device), WARPS_PER_BLOCK * 32, 0, s>>>(
      nullptr, nullptr, tus, tuOffsets, params, nSub, coeffs, ctxIn, sc.cmdBuf, sc.cmdOffsets, sc.cmdCounts, tuFlags, ctxOut,
      status ? status : sc.statusScratch, sc.wpp.sync ? nullptr : sc.order, sc.cursor, nullptr, nullptr, nullptr, nullptr, sc.wpp, sc.coeff16 ? 1 : 0);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const size_t perBlock = RC_WARPS * 32;
  range_code_kernel<<<unsigned((nSub + perBlock - 1) / perBlock), RC_WARPS * 32, 0, s>>>(
      sc.cmdBuf, sc.cmdOffsets, sc.cmdCounts, params, nSub, out, outOffsets, outSizes, numBins,
      status ? status : sc.statusScratch, sc.order);
  return cudaGetLastError();
}

// RDO bit estimate of TU lists (BitEstimator_Std): the fused kernel without its interval coder.
cudaError_t launchEstimateResidual(cudaStream_t s, int device, const Sched &sc, const cabac_tu_desc *tus,
                                   const uint64_t *tuOffsets, const cabac_substream_params *params, size_t nSub,
                                   const int32_t *coeffs, const cabac_ctx_model *ctxIn, uint64_t *estBits,
                                   uint64_t *estPerTu, uint32_t *numBins, uint8_t *tuFlags, cabac_ctx_model *ctxOut,
                                   uint32_t *status, uint32_t *ctxBins) {
  if (nSub == 0) return cudaSuccess;
  if (ctxBins) {  // BinCounter rows: the kernel adds into them
    const cudaError_t e = cudaMemsetAsync(ctxBins, 0, nSub * CABAC_NUM_CTX * sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
  }
#define EST_LAUNCH(C)                                                                                                  \
  encode_residual_kernel<FUSED_ESTIMATE, C><<<persistentGrid(encode_residual_kernel<FUSED_ESTIMATE, C>, nSub, device),     \
                                              WARPS_PER_BLOCK * 32, 0, s>>>(                                             \
      tus, tuOffsets, params, nSub, coeffs, ctxIn, nullptr, nullptr, nullptr, numBins, tuFlags, ctxOut,                  \
      status ? status : sc.statusScratch, sc.order, sc.cursor, estBits, estPerTu, ctxBins, WppArgs())
  if (sc.coeff16) EST_LAUNCH(true); else EST_LAUNCH(false);
#undef EST_LAUNCH
  return cudaGetLastError();
}

// xGetCtxInitId (cabac_writer.cpp:41-73) for n non-intra slices: one warp per slice sums
// numBins(k) * estFracExcessBits(final model k, model k initialised from table B / P)
// (contexts.cpp:965-968) over the contexts and keeps the cheaper table; ties keep B.
__global__ void ctx_init_id_kernel(const cabac_substream_params *__restrict__ params,
                                   const cabac_ctx_model *__restrict__ ctxRef, const uint32_t *__restrict__ ctxBins,
                                   size_t n, uint8_t *__restrict__ choice, uint64_t *__restrict__ costs) {
  const int lane = threadIdx.x & 31;
  const size_t i = (size_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  if (i >= n) return;
  const cabac_substream_params p = params[i];
  const int qp = clipQp(p.qp);
  unsigned long long c0 = 0, c1 = 0;
  for (int k = lane; k < CABAC_NUM_CTX; k += 32) {
    const uint32_t nb = ctxBins[i * CABAC_NUM_CTX + k];
    if (!nb) continue;
    const cabac_ctx_model m = ctxRef[i * CABAC_NUM_CTX + k];
    const unsigned long long ns = 2u * ((uint32_t(m.state0) + m.state1) >> 8) + 1u;
    for (int t = 0; t < 2; t++) {
      const uint32_t w = model_init(qp, CABAC_CTX_INIT_TABLE[t][k], CABAC_CTX_INIT_TABLE[3][k]);
      const uint32_t qt = (model_state0(w) + model_state1(w)) >> 8;
      const unsigned long long x = ((512u - ns) * CABAC_FRAC_BITS[2u * qt] + ns * CABAC_FRAC_BITS[2u * qt + 1u] + 256u) >> 9;
      (t ? c1 : c0) += nb * x;
    }
  }
  for (int o = 16; o; o >>= 1) {
    c0 += __shfl_xor_sync(0xFFFFFFFFu, c0, o);
    c1 += __shfl_xor_sync(0xFFFFFFFFu, c1, o);
  }
  if (lane == 0) {
    choice[i] = p.init_id == 2 ? uint8_t(2) : uint8_t(c1 < c0 ? 1 : 0);  // an intra slice keeps I_SLICE
    if (costs) costs[2 * i] = c0, costs[2 * i + 1] = c1;
  }
}

cudaError_t launchCtxInitId(cudaStream_t s, const cabac_substream_params *params, const cabac_ctx_model *ctxRef,
                            const uint32_t *ctxBins, size_t n, uint8_t *choice, uint64_t *costs) {
  if (n == 0) return cudaSuccess;
  ctx_init_id_kernel<<<unsigned((n + 3) / 4), 128, 0, s>>>(params, ctxRef, ctxBins, n, choice, costs);
  return cudaGetLastError();
}

// Recorded command streams (BinEncIf calls + TU markers): stage P in stream mode, then stage D.
cudaError_t launchEncodeStream(cudaStream_t s, int device, const Sched &sc, const uint32_t *rawCmds,
                               const uint64_t *rawOffsets, const cabac_tu_desc *tus, const uint64_t *tuOffsets,
                               const cabac_substream_params *params, size_t nSub, const int32_t *coeffs,
                               const cabac_ctx_model *ctxIn, uint8_t *out, const uint64_t *outOffsets,
                               uint32_t *outSizes, uint32_t *numBins, uint8_t *tuFlags, cabac_ctx_model *ctxOut,
                               uint32_t *status) {
  if (nSub == 0) return cudaSuccess;
  if (!sc.twoStage) {  // few, long substreams: one warp each, bytes straight out of the producer kernel
    produce_kernel<true, true><<<persistentGrid(produce_kernel<true, true>, nSub, device), WARPS_PER_BLOCK * 32, 0, s>>>(
        rawCmds, rawOffsets, tus, tuOffsets, params, nSub, coeffs, ctxIn, nullptr, nullptr, nullptr, tuFlags, ctxOut,
        status ? status : sc.statusScratch, sc.wpp.sync ? nullptr : sc.order, sc.cursor, out, outOffsets, outSizes, numBins, sc.wpp, sc.coeff16 ? 1 : 0);
    return cudaGetLastError();
  }
  produce_kernel<true, false><<<persistentGrid(produce_kernel<true, false>, nSub, device), WARPS_PER_BLOCK * 32, 0, s>>>(
      rawCmds, rawOffsets, tus, tuOffsets, params, nSub, coeffs, ctxIn, sc.cmdBuf, sc.cmdOffsets, sc.cmdCounts,
      tuFlags, ctxOut, status ? status : sc.statusScratch, sc.wpp.sync ? nullptr : sc.order, sc.cursor, nullptr, nullptr, nullptr, nullptr, sc.wpp, sc.coeff16 ? 1 : 0);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const size_t perBlock = RC_WARPS * 32;
  range_code_kernel<<<unsigned((nSub + perBlock - 1) / perBlock), RC_WARPS * 32, 0, s>>>(
      sc.cmdBuf, sc.cmdOffsets, sc.cmdCounts, params, nSub, out, outOffsets, outSizes, numBins,
      status ? status : sc.statusScratch, sc.order);
  return cudaGetLastError();
}

cudaError_t launchDecodeResidual(cudaStream_t s, int device, const Sched &sc, const uint8_t *in,
                                 const uint64_t *inOffsets, const cabac_tu_desc *tus, const uint64_t *tuOffsets,
                                 const cabac_substream_params *params, size_t nSub, const cabac_ctx_model *ctxIn,
                                 int32_t *coeffs, uint32_t *numBins, uint32_t *bitsRead, uint8_t *tuFlags,
                                 cabac_ctx_model *ctxOut, uint32_t *status) {
  if (nSub == 0) return cudaSuccess;
  // LPT cost of a substream to decode: 8 x bytes + 8 per block + 1/16 per coefficient position (sweep in
  // DESIGN.md section 4; bytes alone leave 8 % on the table: sparse substreams still pay for their blocks)
  static const int byteWeight = getenv("CABAC_B200_DEC_BYTE_WEIGHT") ? atoi(getenv("CABAC_B200_DEC_BYTE_WEIGHT")) : 8;
  static const int tuConst = getenv("CABAC_B200_DEC_TU_CONST") ? atoi(getenv("CABAC_B200_DEC_TU_CONST")) : 8;
  static const int areaShift = getenv("CABAC_B200_DEC_AREA_SHIFT") ? atoi(getenv("CABAC_B200_DEC_AREA_SHIFT")) : 4;
  const int costThreads = nSub >= 16384 ? 32 : 256;  // many substreams are short substreams
  if (byteWeight > 0 && tus)
    cost_decode_kernel<<<costGrid(nSub, 1, costThreads == 32 ? 32 : 16), costThreads, 0, s>>>(inOffsets, tus, tuOffsets, nSub, uint32_t(byteWeight),
                                                        uint32_t(tuConst), uint32_t(areaShift), sc.cost);
  else
    cost_bytes_kernel<<<costGrid(nSub, 256), 256, 0, s>>>(inOffsets, nSub, sc.cost);
  cudaError_t e = launchOrder(s, sc, nSub);
  if (e != cudaSuccess) return e;
  if (sc.threadDecode && !sc.wpp.sync) {  // one thread per substream (cabac_decode_thread.cuh)
    const unsigned zb = unsigned(std::min<size_t>(nSub * 4 + 1, size_t(148) * 32));
    // One warp per block: a block keeps its slot until its slowest substream ends, so small blocks retire sooner
    // (100 000 substreams: 139 ms at 128 threads, 116 at 64, 111 at 32; profiles/r2_decoder_modes.md). Dev knobs.
    static const int tdBlock = getenv("CABAC_B200_TD_BLOCK") ? atoi(getenv("CABAC_B200_TD_BLOCK")) : 32;
    static const int tdDense = getenv("CABAC_B200_TD_DENSE") ? atoi(getenv("CABAC_B200_TD_DENSE")) : 0;
    static const long tdSpreadMax = getenv("CABAC_B200_TD_SPREAD_MAX") ? atol(getenv("CABAC_B200_TD_SPREAD_MAX")) : 2048;
    const int spread = nSub <= size_t(tdSpreadMax) ? 1 : 0;
    const unsigned bs = spread ? 32u : unsigned(tdBlock >= 32 && tdBlock <= 128 ? tdBlock & ~31 : 32);
    const unsigned grid = spread ? unsigned(nSub) : unsigned((nSub + bs - 1) / bs);
#define TD_LAUNCH(T, MINB)                                                                                              \
  decode_residual_thread_kernel<T, MINB><<<grid, bs, 0, s>>>(in, inOffsets, tus, tuOffsets, params, nSub, ctxIn,        \
                                                             reinterpret_cast<T *>(coeffs), numBins, bitsRead, tuFlags, \
                                                             ctxOut, status, sc.order, sc.threadCtx, spread)
    if (sc.coeff16) {
      td_zero_blocks_kernel<int16_t><<<zb, 256, 0, s>>>(tus, tuOffsets, nSub, reinterpret_cast<int16_t *>(coeffs));
      if (tdDense) TD_LAUNCH(int16_t, 8); else TD_LAUNCH(int16_t, 1);
    } else {
      td_zero_blocks_kernel<int32_t><<<zb, 256, 0, s>>>(tus, tuOffsets, nSub, coeffs);
      if (tdDense) TD_LAUNCH(int32_t, 8); else TD_LAUNCH(int32_t, 1);
    }
#undef TD_LAUNCH
    return cudaGetLastError();
  }
#define LS_LAUNCH(W, C)                                                                                            \
  decode_residual_kernel<W, C><<<persistentGrid(decode_residual_kernel<W, C>, nSub, device), WARPS_PER_BLOCK * 32, 0, s>>>( \
      in, inOffsets, tus, tuOffsets, params, nSub, ctxIn, coeffs, numBins, bitsRead, tuFlags, ctxOut, status,        \
      (sc.wpp.sync && !(sc.wpp.debug & 4u)) ? nullptr : sc.order, sc.cursor, sc.wpp)
  if (sc.wpp.sync) {
    if (sc.coeff16) LS_LAUNCH(true, true); else LS_LAUNCH(true, false);
  } else {
    if (sc.coeff16) LS_LAUNCH(false, true); else LS_LAUNCH(false, false);
  }
#undef LS_LAUNCH
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// Allocation-free residual decode (cabac_decode_residual_static_dev; no counterpart in the reference).
// The thread-per-substream decoder with its context rows in a file-scope device array instead of a cudaMalloc'd
// buffer: nothing is allocated per call, so the call is legal where cudaMalloc is not. Every GPU that loads the
// library holds the array (kStaticMaxSubstreams x 230 words = 14.4 MiB) until exit. All calls share it: the kernel
// rewrites a substream's whole row before the first bin, so stale contents never matter, but two calls that overlap
// on the GPU (different streams) race on the rows.
// ---------------------------------------------------------------------------
constexpr size_t kStaticMaxSubstreams = CABAC_STATIC_MAX_SUBSTREAMS;
static __device__ __align__(256) uint32_t g_staticCtxRows[kStaticMaxSubstreams * RCTX];

// In host code `g_staticCtxRows` is only a handle for the symbol, not a device address: the address is looked up once
// per device (with that device current) and kept here.
constexpr int kMaxDevices = 64;
static uint32_t *s_staticCtxRows[kMaxDevices] = {};

cudaError_t setupStaticScratch(int device) {
  if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
  if (s_staticCtxRows[device]) return cudaSuccess;
  void *p = nullptr;
  cudaError_t e = cudaGetSymbolAddress(&p, g_staticCtxRows);
  if (e != cudaSuccess) return e;
  if (!p) return cudaErrorInvalidSymbol;
  // Load the two kernels now (lazy module loading would otherwise do it, and allocate, inside the first call).
  cudaFuncAttributes fa;
  e = cudaFuncGetAttributes(&fa, td_zero_blocks_kernel<int32_t>);
  if (e != cudaSuccess) return e;
  e = cudaFuncGetAttributes(&fa, decode_residual_thread_kernel<int32_t, 1>);
  if (e != cudaSuccess) return e;
  s_staticCtxRows[device] = static_cast<uint32_t *>(p);
  return cudaSuccess;
}

cudaError_t launchDecodeResidualStatic(cudaStream_t s, int device, const uint8_t *in, const uint64_t *inOffsets,
                                       const cabac_tu_desc *tus, const uint64_t *tuOffsets,
                                       const cabac_substream_params *params, size_t nSub, const cabac_ctx_model *ctxIn,
                                       int32_t *coeffs, uint32_t *numBins, uint32_t *bitsRead, uint8_t *tuFlags,
                                       cabac_ctx_model *ctxOut, uint32_t *status) {
  if (nSub == 0) return cudaSuccess;
  if (nSub > kStaticMaxSubstreams || device < 0 || device >= kMaxDevices) return cudaErrorInvalidValue;
  uint32_t *rows = s_staticCtxRows[device];
  if (!rows) return cudaErrorNotReady;  // setupStaticScratch did not run on this device
  const unsigned zb = unsigned(std::min<size_t>(nSub * 4 + 1, size_t(148) * 32));
  td_zero_blocks_kernel<int32_t><<<zb, 256, 0, s>>>(tus, tuOffsets, nSub, coeffs);
  decode_residual_thread_kernel<int32_t, 1><<<unsigned((nSub + 31) / 32), 32, 0, s>>>(
      in, inOffsets, tus, tuOffsets, params, nSub, ctxIn, coeffs, numBins, bitsRead, tuFlags, ctxOut, status, nullptr,
      rows, 0);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// Coefficient range [lo, hi) (TCoeff units) the TUs of each chunk of substreams touch:
// lets the host-pointer entry points stream the slab over PCIe chunk by chunk while earlier
// chunks are already being coded.  ranges[2k] = lo, ranges[2k+1] = hi; pre-set to ~0 / 0.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) coeff_range_kernel(const cabac_tu_desc *__restrict__ tus,
                                                          const uint64_t *__restrict__ tuOffsets,
                                                          const uint32_t *__restrict__ chunkStart,
                                                          unsigned long long *__restrict__ ranges,
                                                          unsigned long long *__restrict__ areaSums) {
  const int k = blockIdx.y;
  const uint64_t t0 = tuOffsets[chunkStart[k]], t1 = tuOffsets[chunkStart[k + 1]];
  unsigned long long lo = ~0ull, hi = 0, sum = 0;
  for (uint64_t t = t0 + blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; t < t1; t += uint64_t(gridDim.x) * blockDim.x) {
    const cabac_tu_desc d = loadDesc(tus + t);
    const unsigned long long a = d.coeff_offset, b = a + (1ull << (d.log2_w + d.log2_h));
    lo = a < lo ? a : lo;
    hi = b > hi ? b : hi;
    sum += b - a;
  }
  if (areaSums) {
    sum = __reduce_add_sync(FULL, uint32_t(sum)) + (uint64_t(__reduce_add_sync(FULL, uint32_t(sum >> 32))) << 32);
    if ((threadIdx.x & 31) == 0 && sum) atomicAdd(areaSums + k, sum);
  }
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long l2 = __shfl_xor_sync(FULL, lo, o), h2 = __shfl_xor_sync(FULL, hi, o);
    lo = l2 < lo ? l2 : lo;
    hi = h2 > hi ? h2 : hi;
  }
  if ((threadIdx.x & 31) == 0 && hi) {
    atomicMin(ranges + 2 * k, lo);
    atomicMax(ranges + 2 * k + 1, hi);
  }
}

cudaError_t launchCoeffRanges(cudaStream_t s, const cabac_tu_desc *tus, const uint64_t *tuOffsets,
                              const uint32_t *chunkStart, int nChunks, uint64_t *ranges, uint64_t *areaSums) {
  if (nChunks <= 0) return cudaSuccess;
  coeff_range_kernel<<<dim3(64, nChunks), 256, 0, s>>>(tus, tuOffsets, chunkStart,
                                                        reinterpret_cast<unsigned long long *>(ranges),
                                                        reinterpret_cast<unsigned long long *>(areaSums));
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// K7: start-code emulation (the step right after the path).
//   count : OutputBitstream::countStartCodeEmulations per substream (bit_stream.cpp:157-181) -
//           the term VTM adds to every entry-point offset (EncGOP.cpp:3811-3830)
//   insert: writeNaluContent (VTM Lib/EncoderLib/NALwrite.cpp:67-126): an emulation_prevention_
//           three_byte before every byte <= 3 that follows two zero bytes, the zero count
//           restarting after each insertion; a trailing zero byte gets a final 0x03.
// Both are per-byte decisions once the zero run in front of a byte is known: in a run of L zeros
// the sequential rule inserts before the run's bytes 2, 4, 6 ... and before a following byte <= 3
// iff L >= 2 and L is even.  Arithmetic-coded bytes rarely hold zero runs, so each candidate byte
// simply walks back over the zeros in front of it.
// ---------------------------------------------------------------------------
__device__ __forceinline__ bool epbBefore(const uint8_t *__restrict__ data, uint64_t s0, uint64_t i) {
  if (data[i] > 3) return false;
  uint64_t j = i;
  while (j > s0 && data[j - 1] == 0) j--;
  const uint64_t run = i - j;
  return run >= 2 && (run & 1) == 0;
}

// one warp per segment
__global__ void __launch_bounds__(256) emulation_count_kernel(const uint8_t *__restrict__ data,
                                                              const uint64_t *__restrict__ offsets, size_t nSeg,
                                                              uint32_t *__restrict__ counts) {
  const int lane = threadIdx.x & 31;
  const size_t warp = (size_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const size_t nWarps = (size_t(gridDim.x) * blockDim.x) >> 5;
  for (size_t sgm = warp; sgm < nSeg; sgm += nWarps) {
    const uint64_t a = offsets[sgm], b = offsets[sgm + 1];
    uint32_t c = 0;
    // "end() - 1 prevents finding a trailing two byte sequence" is automatic: a flagged byte exists
    for (uint64_t i = a + lane; i < b; i += 32) c += epbBefore(data, a, i) ? 1u : 0u;
    c = __reduce_add_sync(FULL, c);
    if (lane == 0) counts[sgm] = c;
  }
}

constexpr int EPB_PER_THREAD = 16;
constexpr int EPB_TILE = 256 * EPB_PER_THREAD;

// pass A: insertions per 4 KiB tile
__global__ void __launch_bounds__(256) emulation_tile_count_kernel(const uint8_t *__restrict__ data, uint64_t n,
                                                                   uint32_t *__restrict__ tileCounts) {
  const uint64_t base = uint64_t(blockIdx.x) * EPB_TILE + uint64_t(threadIdx.x) * EPB_PER_THREAD;
  uint32_t c = 0;
  for (int k = 0; k < EPB_PER_THREAD; k++)
    if (base + k < n) c += epbBefore(data, 0, base + k) ? 1u : 0u;
  __shared__ uint32_t ws[8];
  c = __reduce_add_sync(FULL, c);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < 8; w++) t += ws[w];
    tileCounts[blockIdx.x] = t;
  }
}

// pass B: copy with insertions; tileBase = exclusive scan of the tile counts
__global__ void __launch_bounds__(256) emulation_insert_kernel(const uint8_t *__restrict__ data, uint64_t n,
                                                               const uint64_t *__restrict__ tileBase,
                                                               uint8_t *__restrict__ out, uint64_t outCap,
                                                               uint64_t *__restrict__ outSize) {
  const uint64_t base = uint64_t(blockIdx.x) * EPB_TILE + uint64_t(threadIdx.x) * EPB_PER_THREAD;
  uint32_t mask = 0;
  for (int k = 0; k < EPB_PER_THREAD; k++)
    if (base + k < n && epbBefore(data, 0, base + k)) mask |= 1u << k;
  const uint32_t c = __popc(mask);
  // block-wide exclusive scan of c
  __shared__ uint32_t ws[8];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint32_t inc = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(FULL, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) ws[w] = inc;
  __syncthreads();
  uint32_t before = inc - c;
  for (int k = 0; k < w; k++) before += ws[k];
  uint64_t o = base + tileBase[blockIdx.x] + before;
  for (int k = 0; k < EPB_PER_THREAD && base + k < n; k++) {
    if ((mask >> k) & 1u) {
      if (o < outCap) out[o] = 3;
      o++;
    }
    if (o < outCap) out[o] = data[base + k];
    o++;
  }
  if (n > 0 && base <= n - 1 && n - 1 < base + EPB_PER_THREAD) {  // the thread that owns the last byte
    if (data[n - 1] == 0) {  // 7.4.1.1: a final 0x03 after a trailing zero byte
      if (o < outCap) out[o] = 3;
      o++;
    }
    *outSize = o;
  }
}

cudaError_t launchEmulationCount(cudaStream_t s, const uint8_t *data, const uint64_t *offsets, size_t nSeg,
                                 uint32_t *counts) {
  if (nSeg == 0) return cudaSuccess;
  emulation_count_kernel<<<costGrid(nSeg, 8), 256, 0, s>>>(data, offsets, nSeg, counts);
  return cudaGetLastError();
}

size_t emulationTiles(size_t n) { return (n + EPB_TILE - 1) / EPB_TILE; }

// tileCounts: emulationTiles(n) u32, tileBase: emulationTiles(n) + 1 u64 (device scratch)
cudaError_t launchEmulationInsert(cudaStream_t s, const uint8_t *data, size_t n, uint32_t *tileCounts,
                                  uint64_t *tileBase, uint8_t *out, size_t outCap, uint64_t *outSize) {
  if (n == 0) return cudaMemsetAsync(outSize, 0, sizeof(uint64_t), s);
  const size_t tiles = emulationTiles(n);
  emulation_tile_count_kernel<<<unsigned(tiles), 256, 0, s>>>(data, n, tileCounts);
  entry_points_kernel<<<1, 1024, 0, s>>>(tileCounts, nullptr, tiles, tileBase);
  emulation_insert_kernel<<<unsigned(tiles), 256, 0, s>>>(data, n, tileBase, out, outCap, outSize);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// K8: the decode side of substream assembly.
//   remove : convertPayloadToRBSP (VTM Lib/DecoderLib/NALread.cpp:59-106): drop every
//            emulation_prevention_three_byte, record its position in the escaped buffer.  In a payload
//            the CHECKs of that function accept, a byte is an EPB iff it is 0x03 and follows two zero
//            bytes (the zero count restarts after an EPB, and a third zero / a 01 / a 02 after two zeros
//            is one of the refused cases), so the decision is local; refused patterns are counted.
//   syntax : sh_entry_offset_len_minus1 / sh_entry_point_offset_minus1 (VLCWriter.cpp:2833-2866) from the
//            substream sizes + emulation counts (EncSlice.cpp:2118-2125)
//   ranges : the substreams' RBSP byte ranges from the signalled offsets and the EPB locations
//            (VLCReader.cpp:4346-4381, DecSlice.cpp:106-114, bit_stream.cpp:382-415)
// ---------------------------------------------------------------------------
__device__ __forceinline__ bool isEpb(const uint8_t *__restrict__ d, uint64_t i) {
  return i >= 2 && d[i] == 3 && d[i - 1] == 0 && d[i - 2] == 0;
}
__device__ __forceinline__ bool epbViolation(const uint8_t *__restrict__ d, uint64_t n, uint64_t i) {
  if (i >= 2 && d[i] < 3 && d[i - 1] == 0 && d[i - 2] == 0) return true;  // 00 00 00 / 01 / 02
  if (isEpb(d, i) && i + 1 < n && d[i + 1] > 3) return true;              // "Read a value bigger than '3'"
  return i + 1 == n && d[i] == 0;                                          // "Zero count not '0'"
}

__global__ void __launch_bounds__(256) epb_remove_count_kernel(const uint8_t *__restrict__ data, uint64_t n,
                                                               uint32_t *__restrict__ tileCounts,
                                                               uint32_t *__restrict__ violations) {
  const uint64_t base = uint64_t(blockIdx.x) * EPB_TILE + uint64_t(threadIdx.x) * EPB_PER_THREAD;
  uint32_t c = 0, v = 0;
  for (int k = 0; k < EPB_PER_THREAD; k++)
    if (base + k < n) {
      c += isEpb(data, base + k) ? 1u : 0u;
      v += epbViolation(data, n, base + k) ? 1u : 0u;
    }
  __shared__ uint32_t ws[8];
  c = __reduce_add_sync(FULL, c);
  v = __reduce_add_sync(FULL, v);
  if ((threadIdx.x & 31) == 0) {
    ws[threadIdx.x >> 5] = c;
    if (v) atomicAdd(violations, v);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < 8; w++) t += ws[w];
    tileCounts[blockIdx.x] = t;
  }
}

// tileBase = exclusive scan of the tile counts ([tiles] = total).  sizes[0] = RBSP length, sizes[1] = #EPB.
__global__ void __launch_bounds__(256) epb_remove_kernel(const uint8_t *__restrict__ data, uint64_t n,
                                                         const uint64_t *__restrict__ tileBase, uint64_t tiles,
                                                         int stripTrailingZeros, uint8_t *__restrict__ out,
                                                         uint32_t *__restrict__ epbLocations,
                                                         uint64_t *__restrict__ sizes) {
  const uint64_t base = uint64_t(blockIdx.x) * EPB_TILE + uint64_t(threadIdx.x) * EPB_PER_THREAD;
  uint32_t mask = 0;
  for (int k = 0; k < EPB_PER_THREAD; k++)
    if (base + k < n && isEpb(data, base + k)) mask |= 1u << k;
  const uint32_t c = __popc(mask);
  __shared__ uint32_t ws[8];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint32_t inc = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(FULL, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) ws[w] = inc;
  __syncthreads();
  uint64_t before = tileBase[blockIdx.x] + (inc - c);
  for (int k = 0; k < w; k++) before += ws[k];
  for (int k = 0; k < EPB_PER_THREAD && base + k < n; k++) {
    if ((mask >> k) & 1u) {
      if (epbLocations) epbLocations[before] = uint32_t(base + k);
      before++;
    } else {
      out[base + k - before] = data[base + k];
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    uint64_t len = n - tileBase[tiles];
    if (stripTrailingZeros) {  // cabac_zero_words at the end of a VCL NAL unit, NALread.cpp:92-103
      uint64_t i = n;          // walk back over trailing zero bytes of the ESCAPED buffer (EPBs are not zero,
      while (len > 0) {        // so map the last RBSP byte back: skip EPBs from the end)
        while (i > 0 && isEpb(data, i - 1)) i--;
        if (i == 0 || data[i - 1] != 0) break;
        i--;
        len--;
      }
    }
    sizes[0] = len;
    sizes[1] = tileBase[tiles];
  }
}

cudaError_t launchEmulationRemove(cudaStream_t s, const uint8_t *data, size_t n, int stripTrailingZeros,
                                  uint32_t *tileCounts, uint64_t *tileBase, uint8_t *out, uint32_t *epbLocations,
                                  uint64_t *sizes, uint32_t *violations) {
  cudaError_t e = cudaMemsetAsync(violations, 0, sizeof(uint32_t), s);
  if (e != cudaSuccess) return e;
  if (n == 0) return cudaMemsetAsync(sizes, 0, 2 * sizeof(uint64_t), s);
  const size_t tiles = emulationTiles(n);
  epb_remove_count_kernel<<<unsigned(tiles), 256, 0, s>>>(data, n, tileCounts, violations);
  entry_points_kernel<<<1, 1024, 0, s>>>(tileCounts, nullptr, tiles, tileBase);
  epb_remove_kernel<<<unsigned(tiles), 256, 0, s>>>(data, n, tileBase, tiles, stripTrailingZeros, out, epbLocations, sizes);
  return cudaGetLastError();
}

// one CTA: offsets[i] = sizes[i] + emulations[i] - 1 (i < n - 1), *lenMinus1 = bits of the largest size - 1
__global__ void __launch_bounds__(256) entry_point_syntax_kernel(const uint32_t *__restrict__ sizes,
                                                                 const uint32_t *__restrict__ emulations, size_t n,
                                                                 uint32_t *__restrict__ offsetMinus1,
                                                                 uint32_t *__restrict__ lenMinus1) {
  __shared__ uint32_t sMax;
  if (threadIdx.x == 0) sMax = 0;
  __syncthreads();
  uint32_t m = 0;
  for (size_t i = threadIdx.x; i + 1 < n; i += blockDim.x) {
    const uint32_t v = sizes[i] + (emulations ? emulations[i] : 0u);
    offsetMinus1[i] = v - 1u;
    m = max(m, v);
  }
  m = __reduce_max_sync(FULL, m);
  if ((threadIdx.x & 31) == 0) atomicMax(&sMax, m);
  __syncthreads();
  if (threadIdx.x == 0) *lenMinus1 = sMax < 2u ? 0u : uint32_t(31 - __clz(sMax));  // smallest len with max < 2^(len+1)
}

// one CTA: the loops of parseSliceHeader as lower-bound searches in the (ascending) EPB locations
__global__ void __launch_bounds__(256) substream_ranges_kernel(const uint32_t *__restrict__ offsetMinus1, size_t nEntry,
                                                               const uint32_t *__restrict__ epb, size_t nEpb,
                                                               uint64_t hdrEnd, uint64_t rbspSize,
                                                               uint64_t *__restrict__ inOffsets) {
  __shared__ uint64_t sHdrEsc;
  __shared__ uint64_t part[256];
  auto lowerBound = [&](uint64_t v) {  // EPBs with location < v
    size_t lo = 0, hi = nEpb;
    while (lo < hi) {
      const size_t mid = (lo + hi) >> 1;
      if (uint64_t(epb[mid]) < v) lo = mid + 1; else hi = mid;
    }
    return uint64_t(lo);
  };
  if (threadIdx.x == 0) {  // endOfSliceHeaderLocation in escaped coordinates, VLCReader.cpp:4350-4360
    uint64_t h = hdrEnd;
    for (size_t k = 0; k < nEpb; k++)
      if (uint64_t(epb[k]) < h) h++;
    sHdrEsc = h;
  }
  __syncthreads();
  // sequential over entries per thread block-chunk: prefix sums of the signalled and the RBSP sizes
  const size_t per = (nEntry + blockDim.x - 1) / blockDim.x;
  const size_t a = min(nEntry, threadIdx.x * per), b = min(nEntry, a + per);
  uint64_t sumSig = 0;
  for (size_t i = a; i < b; i++) sumSig += uint64_t(offsetMinus1[i]) + 1u;
  part[threadIdx.x] = sumSig;
  __syncthreads();
  uint64_t cur = 0;
  for (int t = 0; t < int(threadIdx.x); t++) cur += part[t];
  // RBSP position of entry i's start = hdrEnd + signalled bytes before it - EPBs inside [hdr, its start)
  for (size_t i = a; i < b; i++) {
    cur += uint64_t(offsetMinus1[i]) + 1u;
    const uint64_t removed = lowerBound(cur + sHdrEsc) - lowerBound(sHdrEsc);
    inOffsets[i + 1] = hdrEnd + cur - removed;
  }
  if (threadIdx.x == 0) {
    inOffsets[0] = hdrEnd;
    inOffsets[nEntry + 1] = rbspSize;
  }
}

cudaError_t launchEntryPointSyntax(cudaStream_t s, const uint32_t *sizes, const uint32_t *emulations, size_t n,
                                   uint32_t *offsetMinus1, uint32_t *lenMinus1) {
  entry_point_syntax_kernel<<<1, 256, 0, s>>>(sizes, emulations, n, offsetMinus1, lenMinus1);
  return cudaGetLastError();
}
cudaError_t launchSubstreamRanges(cudaStream_t s, const uint32_t *offsetMinus1, size_t nEntry, const uint32_t *epb,
                                  size_t nEpb, uint64_t hdrEnd, uint64_t rbspSize, uint64_t *inOffsets) {
  substream_ranges_kernel<<<1, 256, 0, s>>>(offsetMinus1, nEntry, epb, nEpb, hdrEnd, rbspSize, inOffsets);
  return cudaGetLastError();
}

cudaError_t launchPack(cudaStream_t s, int device, const uint8_t *out, const uint64_t *outOffsets,
                       const uint32_t *outSizes, size_t nSub, uint8_t *packed, uint64_t *entryPoints) {
  entry_points_kernel<<<1, 1024, 0, s>>>(outSizes, outOffsets, nSub, entryPoints);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess || nSub == 0) return e;
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  const size_t wantBlocks = (nSub + 7) / 8;
  const size_t blocks = wantBlocks < size_t(sms) * 8 ? wantBlocks : size_t(sms) * 8;
  pack_kernel<<<int(blocks), 256, 0, s>>>(out, outOffsets, outSizes, nSub, packed, entryPoints);
  return cudaGetLastError();
}

}  // namespace cabac

// K9: whole-substream decoder (CTU-level syntax on the device)
#include "cabac_slice.cuh"
