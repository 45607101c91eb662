// cabac_sliced.cuh - residual encoder for FEW, LONG substreams (slices, frames): every stage parallel
// inside a substream.  Included by cabac_kernels.cu (shares its producers and helpers).
//
// One warp per substream (the fused kernels) turns a 10^7-bin slice into one serial chain of ~10^8
// dependent instructions.  Almost nothing in the ENCODER needs that chain:
//
//   A  binarisation is block-local (residual_coding, cabac_writer.cpp:2424-3062: the bins of a TU depend
//      on its coefficients only)            -> all TUs of all substreams in parallel, unresolved words
//   B  the probability of a context-coded bin depends on the earlier bins of the SAME context
//      (BinProbModel_Std::update, contexts.cpp:903-913): 230 independent chains per substream
//                                           -> 8 warps per substream, each owning 1/8 of the contexts
//   C  of the interval coder (arith_codec.cpp:553-582) only the RANGE recurrence is serial, nine bits of
//      state; a segment of words that begins at an LPS word depends on the range before it only through
//      range >> 5 (getLPS, contexts.cpp:945-950)
//                                           -> per segment the 8 class trajectories in parallel (C1), one
//                                              short serial composition per substream (C2)
//   D  `low` is the affine recurrence low' = (low << s) + a, i.e. the code value is
//      X = sum_i a_i << (T - S_i); every segment is coded on its own from low = 0 (D1) and the pieces are
//      placed at their bit offsets (D2); neighbouring pieces overlap in the 9-bit window of the coder,
//      which is added with carry propagation by one thread per substream (D3).
//   The byte stream the reference writes (writeOut / finish, arith_codec.cpp:339-357, 524-546) is the bit
//   string of X >> 8 on T + 1 bits, then the stop bit (bit_stream.cpp:152-155).  tests/tools/segment_model.py
//   checks both statements against the CPU checker in exact integer arithmetic.
#pragma once

namespace cabac {

constexpr int SL_SEG = 2048;      // nominal words per segment

// ---------------------------------------------------------------------------
// multi-block exclusive scan u32 -> u64 (out[n] = total)
// ---------------------------------------------------------------------------
constexpr int SCAN_TILE = 4096;  // elements per CTA (256 threads x 16)
__global__ void __launch_bounds__(256) scan_tile_sums_kernel(const uint32_t *__restrict__ in, size_t n,
                                                             uint32_t *__restrict__ tileSums) {
  const size_t base = size_t(blockIdx.x) * SCAN_TILE;
  uint64_t s = 0;
  for (int k = 0; k < 16; k++) {
    const size_t i = base + size_t(k) * 256 + threadIdx.x;
    if (i < n) s += in[i];
  }
  __shared__ uint64_t ws[8];
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(FULL, s, o);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint64_t t = 0;
    for (int w = 0; w < 8; w++) t += ws[w];
    tileSums[blockIdx.x] = uint32_t(t);  // a tile's sum fits 32 bits (<= 4096 x 2^20)
  }
}
__global__ void __launch_bounds__(256) scan_apply_kernel(const uint32_t *__restrict__ in, size_t n,
                                                         const uint64_t *__restrict__ tileBase,
                                                         uint64_t *__restrict__ out) {
  const size_t base = size_t(blockIdx.x) * SCAN_TILE + size_t(threadIdx.x) * 16;
  uint32_t v[16];
  uint64_t s = 0;
  for (int k = 0; k < 16; k++) {
    v[k] = base + k < n ? in[base + k] : 0u;
    s += v[k];
  }
  // block-wide exclusive scan of the per-thread sums
  __shared__ uint64_t ws[8];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint64_t inc = s;
  for (int o = 1; o < 32; o <<= 1) {
    const uint64_t t = __shfl_up_sync(FULL, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) ws[w] = inc;
  __syncthreads();
  uint64_t before = tileBase[blockIdx.x] + inc - s;
  for (int k = 0; k < w; k++) before += ws[k];
  for (int k = 0; k < 16; k++) {
    if (base + k < n) out[base + k] = before;
    before += v[k];
  }
  if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 255) out[n] = tileBase[gridDim.x];
}
// tileSums: ceil(n / 4096) u32, tileBase: that + 1 u64
static cudaError_t launchScan(cudaStream_t s, const uint32_t *in, size_t n, uint64_t *out, uint32_t *tileSums,
                              uint64_t *tileBase) {
  if (n == 0) return cudaMemsetAsync(out, 0, sizeof(uint64_t), s);
  const size_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  scan_tile_sums_kernel<<<unsigned(tiles), 256, 0, s>>>(in, n, tileSums);
  entry_points_kernel<<<1, 1024, 0, s>>>(tileSums, nullptr, tiles, tileBase);
  scan_apply_kernel<<<unsigned(tiles), 256, 0, s>>>(in, n, tileBase, out);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// A0: command-word capacity of every TU
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) sl_capacity_kernel(const cabac_tu_desc *__restrict__ tus, size_t nTu,
                                                          uint32_t *__restrict__ cap) {
  for (size_t t = blockIdx.x * size_t(blockDim.x) + threadIdx.x; t < nTu; t += size_t(gridDim.x) * blockDim.x)
    cap[t] = tuCmdCapacity(loadDesc(tus + t));
}

// ---------------------------------------------------------------------------
// A: binarisation of every TU into UNRESOLVED command words (compact residual context ids).
// Persistent warps pull TU indices from a cursor; the substream of a TU is found by bisection.
// ---------------------------------------------------------------------------
struct __align__(16) SlProduceSmem {
  int16_t tile[32 * 32];
  uint32_t queue[QUEUE_WORDS];
  uint16_t cgMask[64];
};

__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
sl_produce_kernel(const cabac_tu_desc *__restrict__ tus, const uint64_t *__restrict__ tuOffsets,
                  const cabac_substream_params *__restrict__ params, size_t nSub, size_t nTu,
                  const int32_t *__restrict__ coeffs, const uint64_t *__restrict__ wordOff,
                  uint32_t *__restrict__ words, uint32_t *__restrict__ wordCnt, uint8_t *__restrict__ tuFlags,
                  uint32_t *__restrict__ status, uint32_t *cursor, const int c16) {
  __shared__ SlProduceSmem smem[WARPS_PER_BLOCK];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  SlProduceSmem &sm = smem[wib];
  for (;;) {
    uint32_t ti = 0;
    if (lane == 0) ti = atomicAdd(cursor, 1u);
    ti = __shfl_sync(FULL, ti, 0);
    if (ti >= nTu) break;
    const uint64_t t = ti;
    size_t lo = 0, hi = nSub;  // substream with tuOffsets[s] <= t < tuOffsets[s + 1]
    while (hi - lo > 1) {
      const size_t mid = (lo + hi) >> 1;
      if (tuOffsets[mid] <= t) lo = mid; else hi = mid;
    }
    const size_t sub = lo;
    const cabac_substream_params p = params[sub];
    const bool depQuant = (p.flags & CABAC_SUB_DEP_QUANT) != 0;
    const bool signHiding = (p.flags & CABAC_SUB_SIGN_HIDING) != 0;
    const cabac_tu_desc d = loadDesc(tus + t);
    const TuGeom g = makeGeom(d);
    const bool ts = (d.flags & CABAC_TU_TS) != 0;
    const bool tsPath = ts && !(p.flags & CABAC_SUB_TSRC_DISABLED);
    const bool bdpcm = (d.flags & CABAC_TU_BDPCM) != 0;
    uint32_t res = ts ? CABAC_TUR_TS : 0, st = (d.flags & DESC_CLAMPED) ? uint32_t(CABAC_SS_BAD_DESC) : 0u;
    uint32_t *dst = words + wordOff[t];
    uint32_t cnt = 0;
    if (!loadTile(sm.tile, c16 ? reinterpret_cast<const int32_t *>(reinterpret_cast<const int16_t *>(coeffs) + d.coeff_offset)
                               : coeffs + d.coeff_offset, g, lane, c16 != 0))
      st |= CABAC_SS_COEFF_RANGE;
    // ---- pre-scan: non-zero mask per coefficient group, last position ----
    int lastCG = -1;
    const int cgSize = 1 << g.log2cg;
    const int perIter = 32 >> g.log2cg;
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
    } else {
      int nq = 0, s, scanPosLast = 0, state = 0, remRegBins = 0, remBinsTS = 0;
      uint64_t sigGroups = 0;
      if (d.flags & CABAC_TU_TS_ALLOWED) sm.queue[nq++] = CABAC_MK_CTX(R_TS_FLAG + (g.comp ? 1 : 0), ts);
      if (tsPath) {
        s = 0;
        remBinsTS = ((g.w * g.h) * 7) >> 2;
      } else {
        scanPosLast = (lastCG << g.log2cg) + (31 - __clz(uint32_t(sm.cgMask[lastCG])));
        if (!ts && g.w >= 4 && g.h >= 4) {
          const int maxLfnstPos = ((g.h == 4 && g.w == 4) || (g.h == 8 && g.w == 8)) ? 7 : 15;
          if (scanPosLast > maxLfnstPos) res |= CABAC_TUR_VIOLATES_LFNST;
          if (scanPosLast >= 1) res |= CABAC_TUR_LFNST_LAST_POS;
        }
        if (g.comp == 0 && !ts && scanPosLast >= 1) res |= CABAC_TUR_MTS_LAST_POS;
        int x, y;
        scanXY(g, scanPosLast, x, y);
        nq += pushLastSigCoeff(sm.queue + nq, g, x, y, lane);
        remRegBins = regBinLimit(g);
        s = lastCG;
      }
      __syncwarp();
      bool endTu = false;
      while (!endTu) {
        if (tsPath) {
          nq += produceTsCG(sm.queue + nq, sm.tile, g, bdpcm, sm.cgMask, s, remBinsTS, sigGroups, lane);
          s++;
          endTu = s == g.numCG;
        } else {
          const uint32_t c = __ldg(&g_rom.cgScan[g.lwig][g.lhig][s]);
          const int cgx = c & 15, cgy = c >> 4, cgPos = cgy * g.wig + cgx;
          const bool sig = sm.cgMask[s] != 0;
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
        __syncwarp();
        for (int i = lane; i < nq; i += 32) dst[cnt + i] = sm.queue[i];
        cnt += uint32_t(nq);
        nq = 0;
        __syncwarp();
      }
    }
    if (lane == 0) {
      wordCnt[t] = cnt;
      if (tuFlags) tuFlags[t] = uint8_t(res);
      if (st) atomicOr(status + sub, st);
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------
// A2: compaction.  One warp per TU: its words move from the capacity slab to the substream's compact,
// still unresolved, word array (a contiguous stream is what stage B wants to prefetch).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) sl_compact_kernel(const uint64_t *__restrict__ wordOff,
                                                         const uint32_t *__restrict__ words,
                                                         const uint32_t *__restrict__ wordCnt,
                                                         const uint64_t *__restrict__ wordPos, size_t nTu,
                                                         uint32_t *__restrict__ uwords) {
  const int lane = threadIdx.x & 31;
  const size_t warp = (size_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const size_t nWarps = (size_t(gridDim.x) * blockDim.x) >> 5;
  for (size_t t = warp; t < nTu; t += nWarps) {
    const uint32_t n = wordCnt[t];
    const uint32_t *in = words + wordOff[t];
    uint32_t *out = uwords + wordPos[t];
    for (uint32_t i = lane; i < n; i += 32) out[i] = in[i];
  }
}

// ---------------------------------------------------------------------------
// B: probability resolution as 230 independent chains per substream.
//   B1  count the words of every context per chunk of the substream's word stream
//   B2  per substream: bucket starts (exclusive scan over contexts) and the chunks' offsets inside them
//   B3  stable scatter: context-coded word i goes to its context's bucket as (bin << 31 | i - w0);
//       bypass words are final already and go straight to the resolved array
//   B4  one THREAD per context walks its bucket with the model in registers (BinProbModel_Std::update,
//       contexts.cpp:903-913; the LPS row of getLPS, :945-950) and writes the resolved word back to
//       position i.  The longest chain - a few percent of the substream - is the critical path.
// ---------------------------------------------------------------------------
constexpr int SL_CHUNK = 4096;  // words per counting chunk (one warp, 128 batches)

struct SlSubWords {       // per substream
  uint64_t w0, w1;        // compact word range
  uint32_t chunk0, nChunk;
};
__global__ void sl_words_plan_kernel(const uint64_t *__restrict__ tuOffsets, const uint64_t *__restrict__ wordPos,
                                     size_t nSub, SlSubWords *__restrict__ sw) {
  if (blockIdx.x || threadIdx.x) return;
  uint32_t c = 0;
  for (size_t s = 0; s < nSub; s++) {
    SlSubWords r;
    r.w0 = wordPos[tuOffsets[s]];
    r.w1 = wordPos[tuOffsets[s + 1]];
    r.chunk0 = c;
    r.nChunk = uint32_t((r.w1 - r.w0 + SL_CHUNK - 1) / SL_CHUNK);
    c += r.nChunk;
    sw[s] = r;
  }
}

__global__ void __launch_bounds__(128) sl_ctxcount_kernel(const SlSubWords *__restrict__ sw,
                                                          const uint32_t *__restrict__ uwords,
                                                          uint32_t *__restrict__ chunkCnt, uint32_t *__restrict__ binsAcc) {
  __shared__ uint32_t cnt[4][RCTX];
  const size_t sub = blockIdx.y;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const SlSubWords r = sw[sub];
  uint32_t bins = 0;
  for (uint32_t c = blockIdx.x * 4 + wib; c < r.nChunk; c += gridDim.x * 4) {
    for (int k = lane; k < RCTX; k += 32) cnt[wib][k] = 0;
    __syncwarp();
    const uint64_t a = r.w0 + uint64_t(c) * SL_CHUNK, b = min(r.w1, a + SL_CHUNK);
    for (uint64_t i = a + lane; i < b; i += 32) {
      const uint32_t w = __ldg(uwords + i);
      if ((w >> 29) == CABAC_CMD_CTX) {
        atomicAdd(&cnt[wib][min(w & 511u, uint32_t(RCTX - 1))], 1u);
        bins++;
      } else {
        bins += (w >> 24) & 31u;
      }
    }
    __syncwarp();
    uint32_t *dst = chunkCnt + size_t(r.chunk0 + c) * RCTX;
    for (int k = lane; k < RCTX; k += 32) dst[k] = cnt[wib][k];
    __syncwarp();
  }
  bins = __reduce_add_sync(FULL, bins);
  if (lane == 0 && bins) atomicAdd(binsAcc + sub, bins);
}

// one CTA per substream, thread k = context k: chunkCnt becomes the chunk's offset inside bucket k,
// bucketStart[sub][k] the bucket's start (relative to w0), [RCTX] the number of context-coded words
__global__ void __launch_bounds__(256) sl_ctxscan_kernel(const SlSubWords *__restrict__ sw,
                                                         uint32_t *__restrict__ chunkCnt,
                                                         uint32_t *__restrict__ bucketStart) {
  __shared__ uint32_t tot[256];
  const size_t sub = blockIdx.x;
  const int k = threadIdx.x;
  const SlSubWords r = sw[sub];
  uint32_t t = 0;
  if (k < RCTX)
    for (uint32_t c = 0; c < r.nChunk; c++) t += chunkCnt[size_t(r.chunk0 + c) * RCTX + k];
  tot[k] = k < RCTX ? t : 0u;
  __syncthreads();
  if (k == 0) {  // 230 values: a serial exclusive scan is a microsecond
    uint32_t run = 0;
    for (int j = 0; j < RCTX; j++) {
      const uint32_t v = tot[j];
      tot[j] = run;
      run += v;
    }
    bucketStart[sub * (RCTX + 1) + RCTX] = run;
  }
  __syncthreads();
  if (k < RCTX) {
    uint32_t run = tot[k];
    bucketStart[sub * (RCTX + 1) + k] = run;
    for (uint32_t c = 0; c < r.nChunk; c++) {
      uint32_t *p = chunkCnt + size_t(r.chunk0 + c) * RCTX + k;
      const uint32_t v = *p;
      *p = run;
      run += v;
    }
  }
}

__global__ void __launch_bounds__(128) sl_ctxscatter_kernel(const SlSubWords *__restrict__ sw,
                                                            const uint32_t *__restrict__ uwords,
                                                            const uint32_t *__restrict__ chunkOff,
                                                            uint32_t *__restrict__ items, uint32_t *__restrict__ cwords) {
  __shared__ uint32_t run[4][RCTX];
  const size_t sub = blockIdx.y;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const SlSubWords r = sw[sub];
  const uint32_t below = (1u << lane) - 1u;
  for (uint32_t c = blockIdx.x * 4 + wib; c < r.nChunk; c += gridDim.x * 4) {
    const uint32_t *src = chunkOff + size_t(r.chunk0 + c) * RCTX;
    for (int k = lane; k < RCTX; k += 32) run[wib][k] = src[k];
    __syncwarp();
    const uint64_t a = r.w0 + uint64_t(c) * SL_CHUNK, b = min(r.w1, a + SL_CHUNK);
    for (uint64_t base = a; base < b; base += 32) {
      const uint64_t i = base + lane;
      const uint32_t w = i < b ? __ldg(uwords + i) : (uint32_t(CABAC_CMD_EP) << 29);
      const bool isCtx = i < b && (w >> 29) == CABAC_CMD_CTX;
      const uint32_t idx = min(w & 511u, uint32_t(RCTX - 1));
      const uint32_t peers = __match_any_sync(FULL, isCtx ? idx : (0x8000u | uint32_t(lane)));
      if (isCtx) {
        const uint32_t rank = __popc(peers & below);
        const uint32_t pos = run[wib][idx] + rank;
        items[r.w0 + pos] = (((w >> 9) & 1u) << 31) | uint32_t(i - r.w0);
      } else if (i < b) {
        cwords[i] = w;  // bypass words need no model
      }
      __syncwarp();
      if (isCtx && (peers & below) == 0) run[wib][idx] += uint32_t(__popc(peers));
      __syncwarp();
    }
  }
}

__global__ void __launch_bounds__(32) sl_ctxresolve_kernel(const SlSubWords *__restrict__ sw,
                                                            const cabac_substream_params *__restrict__ params,
                                                            const cabac_ctx_model *__restrict__ ctxIn,
                                                            const uint32_t *__restrict__ bucketStart,
                                                            const uint32_t *__restrict__ items,
                                                            uint32_t *__restrict__ cwords, const uint32_t *__restrict__ binsAcc,
                                                            uint32_t *__restrict__ numBins, cabac_ctx_model *__restrict__ ctxOut) {
  // grid (8, nSub) x 32 threads: the 230 chains of a substream spread over eight SMs (their scattered
  // stores are what limits one SM)
  const size_t sub = blockIdx.y;
  const int k = int(blockIdx.x) * 32 + int(threadIdx.x);
  const cabac_substream_params p = params[sub];
  const cabac_ctx_model *src = ((p.flags & CABAC_SUB_CTX_IN) && ctxIn) ? ctxIn + sub * CABAC_NUM_CTX : nullptr;
  const SlSubWords r = sw[sub];
  const int qp = clipQp(p.qp);
  if (k == 0 && numBins) numBins[sub] = binsAcc[sub] + ((p.flags & CABAC_SUB_TERMINATE) ? 1u : 0u);
  if (k < RCTX) {
    const int id = rctx_inv(k);
    const uint32_t win = CABAC_CTX_INIT_TABLE[3][id];
    const uint4 e = src ? ctx_entry(src[id].state0, src[id].state1, win)
                        : ctx_entry_init(qp, CABAC_CTX_INIT_TABLE[p.init_id][id], win);
    uint32_t x = e.x;
    const uint32_t j0 = bucketStart[sub * (RCTX + 1) + k], j1 = bucketStart[sub * (RCTX + 1) + k + 1];
    const uint32_t *it = items + r.w0;
    uint32_t *out = cwords + r.w0;
    // the chain: ~10 dependent operations per word; the items of the next 16 words are in flight meanwhile
    constexpr int PF = 16;
    uint32_t nxt[PF];
#pragma unroll
    for (int u = 0; u < PF; u++) nxt[u] = j0 + u < j1 ? __ldg(it + j0 + u) : 0u;
    for (uint32_t j = j0; j < j1; j += PF) {
      uint32_t cur[PF];
#pragma unroll
      for (int u = 0; u < PF; u++) cur[u] = nxt[u];
#pragma unroll
      for (int u = 0; u < PF; u++) nxt[u] = j + PF + u < j1 ? __ldg(it + j + PF + u) : 0u;
#pragma unroll
      for (int u = 0; u < PF; u++) {
        if (j + u < j1) {
          const uint32_t v = cur[u];
          const uint32_t bin = v >> 31;
          const uint32_t q = ((x >> 16) + (x & 0xFFFFu)) >> 8;
          const uint32_t mps = q >> 7;
          const uint32_t qq = q ^ (0xFFu & (0u - mps));
          const uint32_t d0 = (x >> e.z) & (MASK_0 << 16);
          const uint32_t d1 = ((x & 0xFFFFu) >> e.w) & MASK_1;
          x = x - d0 - d1 + (bin ? e.y : 0u);
          out[v & 0x7FFFFFFFu] = (qq >> 2) | ((bin ^ mps) << 6);
        }
      }
    }
    if (ctxOut) {
      cabac_ctx_model m;
      m.state0 = uint16_t(x >> 16);
      m.state1 = uint16_t(x & 0xFFFFu);
      ctxOut[sub * CABAC_NUM_CTX + id] = m;
    }
  }
  if (ctxOut) {  // contexts the residual syntax does not touch pass through
    for (int id = k; id < CABAC_NUM_CTX; id += 256) {
      if (is_residual_ctx(id)) continue;
      cabac_ctx_model m;
      if (src) {
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

// ---------------------------------------------------------------------------
// C0: segments.  Substream s owns compact words [wordPos[tuOffsets[s]], wordPos[tuOffsets[s + 1]]) and
// nSeg = max(1, ceil(words / SL_SEG)) segments; segment j >= 1 begins at the first LPS word at or after
// its nominal start (possibly at the end: an empty segment).
// ---------------------------------------------------------------------------
struct SlSub {          // per substream
  uint64_t w0, w1;      // compact word range
  uint32_t seg0, nSeg;  // segment range
};
__global__ void sl_plan_kernel(const uint64_t *__restrict__ tuOffsets, const uint64_t *__restrict__ wordPos,
                               size_t nSub, SlSub *__restrict__ subs, uint32_t *__restrict__ nSegTotal) {
  if (blockIdx.x || threadIdx.x) return;
  uint32_t seg = 0;
  for (size_t s = 0; s < nSub; s++) {
    SlSub r;
    r.w0 = wordPos[tuOffsets[s]];
    r.w1 = wordPos[tuOffsets[s + 1]];
    const uint64_t n = r.w1 - r.w0;
    r.seg0 = seg;
    r.nSeg = uint32_t(n ? (n + SL_SEG - 1) / SL_SEG : 1);
    seg += r.nSeg;
    subs[s] = r;
  }
  *nSegTotal = seg;
}
__global__ void __launch_bounds__(256) sl_bounds_kernel(const SlSub *__restrict__ subs, size_t nSub,
                                                        const uint32_t *__restrict__ cwords,
                                                        uint64_t *__restrict__ segBound, uint32_t *__restrict__ segSub) {
  // one CTA per substream slice of segments: grid.y = substream, threads stride over its segments
  const size_t s = blockIdx.y;
  if (s >= nSub) return;
  const SlSub r = subs[s];
  for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j <= r.nSeg; j += gridDim.x * blockDim.x) {
    uint64_t b;
    if (j == 0) b = r.w0;
    else if (j == r.nSeg) b = r.w1;
    else {
      b = r.w0 + uint64_t(j) * SL_SEG;
      while (b < r.w1) {
        const uint32_t w = cwords[b];
        if ((w >> 29) == CABAC_CMD_CTX && (w & 64u)) break;
        b++;
      }
    }
    if (j < r.nSeg) {
      segBound[r.seg0 + j] = b;
      segSub[r.seg0 + j] = uint32_t(s);
    } else if (s + 1 == nSub) {
      segBound[r.seg0 + j] = b;  // the closing bound of the last substream
    }
  }
}

// the range recurrence of one word; returns the renormalisation shift
__device__ __forceinline__ uint32_t slRangeStep(uint32_t w, uint32_t &range) {
  const bool isCtx = (w >> 29) == CABAC_CMD_CTX;
  const bool isLps = (w >> 6) & 1u;
  const uint32_t lps = (((w & 63u) * (range >> 5)) >> 1) + 4;
  const uint32_t rmps = range - lps;
  const uint32_t nCtx = isLps ? uint32_t(__clz(int(lps | 7u)) - 23) : 1u - (rmps >> 8);
  range = isCtx ? (isLps ? lps : rmps) << nCtx : range;
  return isCtx ? nCtx : (w >> 24) & 31u;
}

// ---------------------------------------------------------------------------
// C1: class maps.  Thread (segment, class c): the range after the segment and its total shift when the
// range before it lies in class 8 + c (256 + 32 c .. 287 + 32 c).  A substream's first segment starts at
// range 510 whatever the class.  Segment ends of a substream's last segment include encodeBinTrm(1).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) sl_classmap_kernel(const SlSub *__restrict__ subs,
                                                          const cabac_substream_params *__restrict__ params,
                                                          const uint64_t *__restrict__ segBound,
                                                          const uint32_t *__restrict__ segSub,
                                                          const uint32_t *__restrict__ nSegPtr,
                                                          const uint32_t *__restrict__ cwords,
                                                          uint16_t *__restrict__ mapR, uint32_t *__restrict__ mapT) {
  const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t j = gid >> 3, c = gid & 7u;
  if (j >= *nSegPtr) return;
  const uint32_t s = segSub[j];
  const SlSub r = subs[s];
  const bool first = j == r.seg0, last = j + 1 == r.seg0 + r.nSeg;
  const uint64_t a = segBound[j], b = last ? r.w1 : segBound[j + 1];
  uint32_t range = first ? 510u : ((8u + c) << 5);
  uint32_t T = 0;
  for (uint64_t i = a; i < b; i++) T += slRangeStep(__ldg(cwords + i), range);
  const bool trm = last && (params[s].flags & CABAC_SUB_TERMINATE);
  if (trm) {
    T += 7;
    range = 256;
  }
  if (a == b && !trm && !first) range = 0;  // an empty segment: the range passes through unchanged (see sl_compose_kernel)
  mapR[size_t(j) * 8 + c] = uint16_t(range);
  mapT[size_t(j) * 8 + c] = T;
}

// ---------------------------------------------------------------------------
// C2: serial composition per substream: the range and the bit offset every segment starts at; the
// substream's total shift; local byte scratch of every segment.
// ---------------------------------------------------------------------------
struct SlSeg {
  uint64_t tStart;   // shift total before the segment (stream bit offset of its 9-bit window)
  uint64_t scratch;  // byte offset of its local bit string
  uint32_t tLen;     // its own shift total
  uint16_t rStart;
  uint16_t window;   // D1: top bits (window + carry) of its local value
};
__global__ void __launch_bounds__(128) sl_compose_kernel(const SlSub *__restrict__ subs, size_t nSub,
                                                         const uint16_t *__restrict__ mapR,
                                                         const uint32_t *__restrict__ mapT, SlSeg *__restrict__ segs,
                                                         uint64_t *__restrict__ subBits) {
  // one warp per substream: the maps of four segments (4 x 8 classes) arrive with one load per lane, the
  // serial walk then runs on registers
  const int lane = threadIdx.x & 31;
  const size_t s = (blockIdx.x * size_t(blockDim.x) + threadIdx.x) >> 5;
  if (s >= nSub) return;
  const SlSub r = subs[s];
  uint32_t range = 510;
  uint64_t T = 0;
  for (uint32_t j0 = r.seg0; j0 < r.seg0 + r.nSeg; j0 += 4) {
    const uint32_t j = j0 + (lane >> 3);
    const bool have = j < r.seg0 + r.nSeg;
    const uint32_t mr = have ? mapR[size_t(j) * 8 + (lane & 7)] : 0u;
    const uint32_t mt = have ? mapT[size_t(j) * 8 + (lane & 7)] : 0u;
    for (int k = 0; k < 4 && j0 + k < r.seg0 + r.nSeg; k++) {
      const uint32_t c = (range >> 5) - 8u;
      const uint32_t after = __shfl_sync(FULL, mr, k * 8 + int(c));
      const uint32_t tl = __shfl_sync(FULL, mt, k * 8 + int(c));
      if (lane == 0) {
        SlSeg g;
        g.tStart = T;
        g.tLen = tl;
        g.rStart = uint16_t(range);
        g.window = 0;
        g.scratch = 0;
        segs[j0 + k] = g;
      }
      if (after) range = after;
      T += tl;
    }
  }
  if (lane == 0) subBits[s] = T;
}
// local scratch offsets: segment j needs ceil((9 + tLen) / 8) + 12 bytes, 4-byte aligned
__global__ void __launch_bounds__(256) sl_scratch_need_kernel(const SlSeg *__restrict__ segs,
                                                              const uint32_t *__restrict__ nSegPtr, uint32_t nSegMax,
                                                              uint32_t *__restrict__ need) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < nSegMax) need[j] = j < *nSegPtr ? ((9u + segs[j].tLen + 7u) / 8u + 12u + 3u) & ~3u : 0u;
}

// ---------------------------------------------------------------------------
// D1: every segment coded on its own from low = 0 and its true start range: the local value
// V = sum a_i << (tLen - S_i) as a big-endian bit string of 9 + tLen bits (the coder's 9-bit window first),
// written to the segment's scratch.  One THREAD per segment; a carry out of the window is kept in `window`.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(128) sl_code_kernel(const SlSub *__restrict__ subs,
                                                      const cabac_substream_params *__restrict__ params,
                                                      const uint64_t *__restrict__ segBound,
                                                      const uint32_t *__restrict__ segSub,
                                                      const uint32_t *__restrict__ nSegPtr,
                                                      const uint32_t *__restrict__ cwords,
                                                      const uint64_t *__restrict__ scratchOff, SlSeg *__restrict__ segs,
                                                      uint8_t *__restrict__ scratch, uint64_t scratchBytes) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= *nSegPtr) return;
  if (scratchOff[j + 1] > scratchBytes) return;  // more bits than the output capacity allows: reported by sl_fix_kernel
  const uint32_t s = segSub[j];
  const SlSub r = subs[s];
  const bool last = j + 1 == r.seg0 + r.nSeg;
  const uint64_t a = segBound[j], b = last ? r.w1 : segBound[j + 1];
  SlSeg g = segs[j];
  uint8_t *dst = scratch + scratchOff[j];
  const uint32_t cap = uint32_t(scratchOff[j + 1] - scratchOff[j]);
  // the coder: `low` keeps the not yet emitted bits, bytes leave it as in writeOut (one at a time when
  // bitsLeft < 12); a carry is an add into the bytes already stored (they are private to this thread)
  uint64_t low = 0;
  uint32_t range = g.rStart, n = 0, carryOut = 0;
  int bitsLeft = 23;
  auto emit = [&]() {
    while (bitsLeft < 12) {
      const int sh = 24 - bitsLeft;
      const uint32_t lead = uint32_t(low >> sh);  // 8 bits + carry
      low -= uint64_t(lead) << sh;
      bitsLeft += 8;
      if (lead >> 8) {  // carry into the stored bytes
        uint32_t k = n;
        for (;;) {
          if (k == 0) {
            carryOut++;
            break;
          }
          k--;
          const uint8_t v = uint8_t(dst[k] + 1);
          dst[k] = v;
          if (v) break;
        }
      }
      if (n < cap) dst[n] = uint8_t(lead);
      n++;
    }
  };
  for (uint64_t i = a; i < b; i++) {
    const uint32_t w = __ldg(cwords + i);
    const bool isCtx = (w >> 29) == CABAC_CMD_CTX;
    const bool isLps = (w >> 6) & 1u;
    const uint32_t epN = (w >> 24) & 31u;
    const uint32_t lps = (((w & 63u) * (range >> 5)) >> 1) + 4;
    const uint32_t rmps = range - lps;
    const uint32_t nCtx = isLps ? uint32_t(__clz(int(lps | 7u)) - 23) : 1u - (rmps >> 8);
    const uint32_t sh = isCtx ? nCtx : epN;
    const uint64_t add = isCtx ? uint64_t(isLps ? rmps << nCtx : 0u) : uint64_t(range) * (w & 0xFFFFFFu);
    range = isCtx ? (isLps ? lps : rmps) << nCtx : range;
    low = (low << sh) + add;
    bitsLeft -= int(sh);
    emit();
  }
  if (last && (params[s].flags & CABAC_SUB_TERMINATE)) {  // encodeBinTrm(1), arith_codec.cpp:460-478
    range -= 2;
    low = (low + range) << 7;
    bitsLeft -= 7;
    emit();
  }
  // what is left in `low`: 32 - bitsLeft bits (+ a carry bit above them)
  const int rest = 32 - bitsLeft;
  const uint32_t carry = uint32_t(low >> rest);
  if (carry) {
    low -= uint64_t(carry) << rest;
    uint32_t k = n;
    for (;;) {
      if (k == 0) {
        carryOut++;
        break;
      }
      k--;
      const uint8_t v = uint8_t(dst[k] + 1);
      dst[k] = v;
      if (v) break;
    }
  }
  // append the rest left-aligned (rest <= 20 bits)
  const uint32_t tailBits = uint32_t(low) << (32 - rest);
  for (int k = 0; k < 4; k++)
    if (n + k < cap) dst[n + k] = uint8_t(tailBits >> (24 - 8 * k));
  // the 9-bit window (first 9 bits of the string) plus the carry out of it
  const uint32_t b0 = cap > 0 ? dst[0] : 0u, b1 = cap > 1 ? dst[1] : 0u;
  g.window = uint16_t(((b0 << 1) | (b1 >> 7)) + (carryOut << 9));
  segs[j] = g;
}

// ---------------------------------------------------------------------------
// D2: placement.  One warp per segment: bits [9, 9 + tLen) of its local string go to stream bits
// [tStart + 9, tStart + 9 + tLen) of the substream's (pre-zeroed) output slab.  Disjoint bit ranges:
// whole words are stored, the two boundary words are OR-ed atomically.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t slLoadBits32(const uint8_t *p, uint64_t bit) {  // 32 bits from bit offset (big-endian)
  const uint8_t *q = p + (bit >> 3);
  const uint64_t v = (uint64_t(q[0]) << 32) | (uint64_t(q[1]) << 24) | (uint64_t(q[2]) << 16) | (uint64_t(q[3]) << 8) |
                     uint64_t(q[4]);
  return uint32_t(v >> (8 - (bit & 7)));
}
__global__ void __launch_bounds__(256) sl_place_kernel(const SlSeg *__restrict__ segs,
                                                       const uint32_t *__restrict__ segSub,
                                                       const uint32_t *__restrict__ nSegPtr,
                                                       const uint64_t *__restrict__ scratchOff,
                                                       const uint8_t *__restrict__ scratch, uint64_t scratchBytes,
                                                       uint8_t *__restrict__ out,
                                                       const uint64_t *__restrict__ outOffsets) {
  const int lane = threadIdx.x & 31;
  const uint32_t j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (j >= *nSegPtr) return;
  const SlSeg g = segs[j];
  if (g.tLen == 0 || scratchOff[j + 1] > scratchBytes) return;
  const uint32_t s = segSub[j];
  const uint8_t *src = scratch + scratchOff[j];
  const uint64_t capBits = (outOffsets[s + 1] - outOffsets[s]) * 8;
  uint32_t *dst = reinterpret_cast<uint32_t *>(out + outOffsets[s]);  // slabs are 16-byte aligned
  const uint64_t d0 = g.tStart + 9, d1 = d0 + g.tLen;                  // destination bit range
  const uint64_t w0 = d0 >> 5, w1 = (d1 - 1) >> 5;
  for (uint64_t w = w0 + lane; w <= w1; w += 32) {
    const uint64_t wb = w << 5;  // first stream bit of this word
    // source bit of stream bit wb: 9 + (wb - d0); may be negative for the first word
    uint32_t v;
    uint32_t mask = 0xFFFFFFFFu;
    if (wb >= d0) {
      v = slLoadBits32(src, 9 + (wb - d0));
    } else {
      const uint32_t skip = uint32_t(d0 - wb);  // 1..31 leading bits belong to the neighbour
      v = slLoadBits32(src, 9) >> skip;
      mask = 0xFFFFFFFFu >> skip;
    }
    if (wb + 32 > d1) mask &= ~((1u << uint32_t(wb + 32 - d1)) - 1u);
    v &= mask;
    if (wb + 32 > capBits) continue;  // output capacity exhausted: reported by the fix kernel
    const uint32_t le = __byte_perm(v, 0, 0x0123);  // big-endian bit order in memory
    if (mask == 0xFFFFFFFFu) dst[w] = le;
    else if (v) atomicOr(dst + w, le);
  }
}

// ---------------------------------------------------------------------------
// D3: windows + finish.  One thread per substream adds every segment's window (its value's top bits,
// which overlap the previous segment's last 9 stream positions) with carry propagation, sets the stop bit
// and reports the size: stream = X >> 8 on T + 1 bits, then '1', zero padded (finish + writeByteAlignment).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(128) sl_fix_kernel(const SlSub *__restrict__ subs, const cabac_substream_params *__restrict__ params,
                              size_t nSub, const SlSeg *__restrict__ segs, const uint64_t *__restrict__ subBits,
                              uint8_t *__restrict__ out, const uint64_t *__restrict__ outOffsets,
                              uint32_t *__restrict__ outSizes, uint32_t *__restrict__ status) {
  const int lane = threadIdx.x & 31;
  const size_t s = (blockIdx.x * size_t(blockDim.x) + threadIdx.x) >> 5;
  if (s >= nSub) return;
  const SlSub r = subs[s];
  uint8_t *dst = out + outOffsets[s];
  const uint64_t cap = outOffsets[s + 1] - outOffsets[s];
  const uint64_t T = subBits[s];
  const bool align = (params[s].flags & CABAC_SUB_ALIGN) != 0;
  const uint64_t totalBits = T + 1 + (align ? 1 : 0);
  const uint64_t totalBytes = (totalBits + 7) >> 3;
  if (totalBytes + 4 > cap) {
    if (lane == 0) {
      outSizes[s] = uint32_t(min(totalBytes, uint64_t(0xFFFFFFFFu)));
      atomicOr(status + s, uint32_t(CABAC_SS_OUT_OVERFLOW));
    }
    return;
  }
  // windows from the last segment to the first; 32 records per load, the adds themselves stay serial
  for (uint32_t hi = r.seg0 + r.nSeg; hi > r.seg0;) {
    const uint32_t lo = hi - r.seg0 >= 32 ? hi - 32 : r.seg0;
    const uint32_t jl = lo + uint32_t(lane);
    uint64_t myT = 0;
    uint32_t myW = 0;
    if (jl < hi) {
      myT = segs[jl].tStart;
      myW = segs[jl].window;
    }
    for (int k = int(hi - lo) - 1; k >= 0; k--) {
      const uint32_t win = __shfl_sync(FULL, myW, k);
      const uint64_t ts = __shfl_sync(FULL, myT, k);
      if (!win || lane) continue;
      const uint64_t p = ts + 8;  // the window's LSB sits at stream bit tStart + 8
      uint64_t byte = p >> 3;
      uint32_t v = win << (7 - uint32_t(p & 7));  // <= 10 + 7 bits, aligned to the end of `byte`
      for (;;) {
        v += dst[byte];
        dst[byte] = uint8_t(v);
        v >>= 8;
        if (!v || byte == 0) break;
        byte--;
      }
    }
    hi = lo;
  }
  if (lane) return;
  if (align) dst[(T + 1) >> 3] |= uint8_t(0x80u >> ((T + 1) & 7));  // stop bit at stream bit T + 1
  outSizes[s] = uint32_t(totalBytes);
}

// sizes the host needs before it can size the scratch: {#TU, output capacity, "every substream is terminated"}
__global__ void sl_probe_kernel(const uint64_t *__restrict__ tuOffsets, const uint64_t *__restrict__ outOffsets,
                                const cabac_substream_params *__restrict__ params, size_t nSub,
                                uint64_t *__restrict__ probe) {
  __shared__ int ok;
  if (threadIdx.x == 0) ok = 1;
  __syncthreads();
  for (size_t i = threadIdx.x; i < nSub; i += blockDim.x)
    if (!(params[i].flags & CABAC_SUB_TERMINATE)) ok = 0;
  __syncthreads();
  if (threadIdx.x == 0) {
    probe[0] = tuOffsets[nSub] - tuOffsets[0];
    probe[1] = outOffsets[nSub];
    probe[2] = uint64_t(ok) | (tuOffsets[0] ? 2u : 0u);
  }
}

// Grow-only device buffer of the sliced encoder (owned by the handle through SlicedWork).
static cudaError_t slGrow(void **p, size_t *have, size_t need) {
  if (need <= *have) return cudaSuccess;
  if (*p) cudaFree(*p);
  *p = nullptr;
  *have = 0;
  const size_t sz = need + (need >> 3) + 4096;
  const cudaError_t e = cudaMalloc(p, sz);
  if (e == cudaSuccess) *have = sz;
  return e;
}

// Returns cudaSuccess and *used = false when the batch is not one for this path (a substream that is not
// terminated: the fused kernels report the reference's interim byte counts for those).
cudaError_t launchEncodeSliced(cudaStream_t st, int device, SlicedWork &W, const cabac_tu_desc *tus,
                               const uint64_t *tuOffsets, const cabac_substream_params *params, size_t nSub,
                               const int32_t *coeffs, bool coeff16, const cabac_ctx_model *ctxIn, uint8_t *out,
                               const uint64_t *outOffsets, uint32_t *outSizes, uint32_t *numBins, uint8_t *tuFlags,
                               cabac_ctx_model *ctxOut, uint32_t *status, size_t minTusPerSub, size_t minTus,
                               bool *used, int *launches) {
  *used = false;
  cudaError_t e;
  auto buf = [&](int k, size_t bytes) { return slGrow(&W.buf[k], &W.bytes[k], bytes); };
  if (!W.pinned && (e = cudaHostAlloc(&W.pinned, 8 * sizeof(uint64_t), cudaHostAllocDefault)) != cudaSuccess) return e;
  enum { B_SMALL, B_CAP, B_WOFF, B_WCNT, B_WPOS, B_TILES, B_WORDS, B_CWORDS, B_SEG, B_SCRATCH, B_UWORDS, B_ITEMS, B_CHUNK };
  // ---- sizes, round 1 ----
  if ((e = buf(B_SMALL, 4096 + nSub * (sizeof(SlSub) + sizeof(SlSubWords) + 8 + (RCTX + 4) * 4) + 64)) != cudaSuccess) return e;
  uint64_t *dProbe = static_cast<uint64_t *>(W.buf[B_SMALL]);        // 8 x u64
  uint32_t *dCursor = reinterpret_cast<uint32_t *>(dProbe + 8);      // cursor, nSegTotal
  uint32_t *dNSeg = dCursor + 1;
  SlSub *dSubs = reinterpret_cast<SlSub *>(static_cast<char *>(W.buf[B_SMALL]) + 4096);
  uint64_t *dSubBits = reinterpret_cast<uint64_t *>(dSubs + nSub);
  sl_probe_kernel<<<1, 256, 0, st>>>(tuOffsets, outOffsets, params, nSub, dProbe);
  if ((e = cudaMemcpyAsync(W.pinned, dProbe, 3 * sizeof(uint64_t), cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
  if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return e;
  *launches += 1;
  const size_t nTu = size_t(W.pinned[0]), outCap = size_t(W.pinned[1]);
  if (!(W.pinned[2] & 1u) || (W.pinned[2] & 2u) || nTu == 0 || nTu >= (size_t(1) << 32) || nTu < minTusPerSub * nSub || nTu < minTus)
    return cudaSuccess;
  // ---- A: capacity, offsets, produce ----
  if ((e = buf(B_CAP, nTu * 4)) != cudaSuccess || (e = buf(B_WOFF, (nTu + 1) * 8)) != cudaSuccess ||
      (e = buf(B_WCNT, nTu * 4)) != cudaSuccess || (e = buf(B_WPOS, (nTu + 1) * 8)) != cudaSuccess)
    return e;
  uint32_t *dCap = static_cast<uint32_t *>(W.buf[B_CAP]), *dWCnt = static_cast<uint32_t *>(W.buf[B_WCNT]);
  uint64_t *dWOff = static_cast<uint64_t *>(W.buf[B_WOFF]), *dWPos = static_cast<uint64_t *>(W.buf[B_WPOS]);
  sl_capacity_kernel<<<costGrid(nTu, 256), 256, 0, st>>>(tus, nTu, dCap);
  // tile scratch of the scans
  uint32_t *dTileSums = nullptr;
  uint64_t *dTileBase = nullptr;
  auto tileScratch = [&](size_t nElems) {
    const size_t tl = (nElems + SCAN_TILE - 1) / SCAN_TILE + 64;
    const cudaError_t ee = buf(B_TILES, tl * 12 + 4096);
    dTileSums = static_cast<uint32_t *>(W.buf[B_TILES]);
    dTileBase = reinterpret_cast<uint64_t *>(static_cast<char *>(W.buf[B_TILES]) + ((tl * 4 + 255) & ~size_t(255)));
    return ee;
  };
  if ((e = tileScratch(nTu)) != cudaSuccess) return e;
  if ((e = launchScan(st, dCap, nTu, dWOff, dTileSums, dTileBase)) != cudaSuccess) return e;
  if ((e = cudaMemcpyAsync(W.pinned, dWOff + nTu, sizeof(uint64_t), cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
  if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return e;
  const size_t capWords = size_t(W.pinned[0]);
  if ((e = buf(B_WORDS, capWords * 4 + 256)) != cudaSuccess) return e;
  uint32_t *dWords = static_cast<uint32_t *>(W.buf[B_WORDS]);
  if ((e = cudaMemsetAsync(dCursor, 0, 8, st)) != cudaSuccess) return e;
  if ((e = cudaMemsetAsync(status, 0, nSub * 4, st)) != cudaSuccess) return e;  // the kernels OR their bits in
  sl_produce_kernel<<<persistentGrid(sl_produce_kernel, nTu, device), WARPS_PER_BLOCK * 32, 0, st>>>(
      tus, tuOffsets, params, nSub, nTu, coeffs, dWOff, dWords, dWCnt, tuFlags, status, dCursor, coeff16 ? 1 : 0);
  if ((e = launchScan(st, dWCnt, nTu, dWPos, dTileSums, dTileBase)) != cudaSuccess) return e;
  if ((e = cudaMemcpyAsync(W.pinned, dWPos + nTu, sizeof(uint64_t), cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
  if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return e;
  *launches += 8;
  const size_t nWords = size_t(W.pinned[0]);
  // ---- B: resolve into the compact array ----
  if (nWords >= (size_t(1) << 31)) return cudaSuccess;  // bucket items carry a 31-bit position
  const size_t nChunkMax = nWords / SL_CHUNK + nSub + 1;
  if ((e = buf(B_CWORDS, nWords * 4 + 256)) != cudaSuccess || (e = buf(B_UWORDS, nWords * 4 + 256)) != cudaSuccess ||
      (e = buf(B_ITEMS, nWords * 4 + 256)) != cudaSuccess || (e = buf(B_CHUNK, nChunkMax * RCTX * 4 + 256)) != cudaSuccess)
    return e;
  uint32_t *dCWords = static_cast<uint32_t *>(W.buf[B_CWORDS]), *dUWords = static_cast<uint32_t *>(W.buf[B_UWORDS]);
  uint32_t *dItems = static_cast<uint32_t *>(W.buf[B_ITEMS]), *dChunk = static_cast<uint32_t *>(W.buf[B_CHUNK]);
  SlSubWords *dSW = reinterpret_cast<SlSubWords *>(dSubBits + nSub);          // 8-byte aligned records first
  uint32_t *dBucket = reinterpret_cast<uint32_t *>(dSW + nSub);                // [nSub][RCTX + 1]
  uint32_t *dBinsAcc = dBucket + nSub * (RCTX + 1);                            // [nSub]
  sl_compact_kernel<<<costGrid(nTu, 8), 256, 0, st>>>(dWOff, dWords, dWCnt, dWPos, nTu, dUWords);
  if ((e = cudaMemsetAsync(dBinsAcc, 0, nSub * 4, st)) != cudaSuccess) return e;
  sl_words_plan_kernel<<<1, 1, 0, st>>>(tuOffsets, dWPos, nSub, dSW);
  const unsigned chunkBlocks = unsigned(std::min<size_t>(nWords / (size_t(SL_CHUNK) * 4 * nSub) + 1, 1024));
  sl_ctxcount_kernel<<<dim3(chunkBlocks, unsigned(nSub)), 128, 0, st>>>(dSW, dUWords, dChunk, dBinsAcc);
  sl_ctxscan_kernel<<<unsigned(nSub), 256, 0, st>>>(dSW, dChunk, dBucket);
  sl_ctxscatter_kernel<<<dim3(chunkBlocks, unsigned(nSub)), 128, 0, st>>>(dSW, dUWords, dChunk, dItems, dCWords);
  sl_ctxresolve_kernel<<<dim3(8, unsigned(nSub)), 32, 0, st>>>(dSW, params, ctxIn, dBucket, dItems, dCWords, dBinsAcc, numBins, ctxOut);
  // ---- C: segments, class maps, composition ----
  const size_t nSegMax = nWords / SL_SEG + nSub + 1;
  if ((e = tileScratch(nSegMax > nTu ? nSegMax : nTu)) != cudaSuccess) return e;
  size_t off = 0;
  auto carve = [&](size_t bytes) {
    const size_t at = off;
    off += (bytes + 255) & ~size_t(255);
    return at;
  };
  const size_t oBound = carve((nSegMax + 1) * 8), oSub = carve(nSegMax * 4), oMapR = carve(nSegMax * 16),
               oMapT = carve(nSegMax * 32), oSegs = carve(nSegMax * sizeof(SlSeg)), oNeed = carve(nSegMax * 4),
               oScrOff = carve((nSegMax + 1) * 8);
  if ((e = buf(B_SEG, off)) != cudaSuccess) return e;
  char *sb = static_cast<char *>(W.buf[B_SEG]);
  uint64_t *dBound = reinterpret_cast<uint64_t *>(sb + oBound), *dScrOff = reinterpret_cast<uint64_t *>(sb + oScrOff);
  uint32_t *dSegSub = reinterpret_cast<uint32_t *>(sb + oSub), *dMapT = reinterpret_cast<uint32_t *>(sb + oMapT),
           *dNeed = reinterpret_cast<uint32_t *>(sb + oNeed);
  uint16_t *dMapR = reinterpret_cast<uint16_t *>(sb + oMapR);
  SlSeg *dSegs = reinterpret_cast<SlSeg *>(sb + oSegs);
  sl_plan_kernel<<<1, 1, 0, st>>>(tuOffsets, dWPos, nSub, dSubs, dNSeg);
  sl_bounds_kernel<<<dim3(64, unsigned(nSub)), 256, 0, st>>>(dSubs, nSub, dCWords, dBound, dSegSub);
  sl_classmap_kernel<<<unsigned((nSegMax * 8 + 255) / 256), 256, 0, st>>>(dSubs, params, dBound, dSegSub, dNSeg, dCWords, dMapR, dMapT);
  sl_compose_kernel<<<unsigned((nSub + 3) / 4), 128, 0, st>>>(dSubs, nSub, dMapR, dMapT, dSegs, dSubBits);
  sl_scratch_need_kernel<<<unsigned((nSegMax + 255) / 256), 256, 0, st>>>(dSegs, dNSeg, uint32_t(nSegMax), dNeed);
  if ((e = launchScan(st, dNeed, nSegMax, dScrOff, dTileSums, dTileBase)) != cudaSuccess) return e;
  // ---- D: code every segment, place the pieces, add the windows ----
  const size_t scratchBytes = outCap + 24 * nSegMax + 4096;
  if ((e = buf(B_SCRATCH, scratchBytes)) != cudaSuccess) return e;
  uint8_t *dScratch = static_cast<uint8_t *>(W.buf[B_SCRATCH]);
  if ((e = cudaMemsetAsync(out, 0, outCap, st)) != cudaSuccess) return e;
  sl_code_kernel<<<unsigned((nSegMax + 127) / 128), 128, 0, st>>>(dSubs, params, dBound, dSegSub, dNSeg, dCWords, dScrOff, dSegs,
                                                                   dScratch, scratchBytes);
  sl_place_kernel<<<unsigned((nSegMax * 32 + 255) / 256), 256, 0, st>>>(dSegs, dSegSub, dNSeg, dScrOff, dScratch, scratchBytes, out,
                                                                        outOffsets);
  sl_fix_kernel<<<unsigned((nSub + 3) / 4), 128, 0, st>>>(dSubs, params, nSub, dSegs, dSubBits, out, outOffsets, outSizes, status);
  *launches += 19;
  *used = true;
  return cudaGetLastError();
}

void freeSlicedWork(SlicedWork &W) {
  for (int k = 0; k < SlicedWork::kBufs; k++)
    if (W.buf[k]) cudaFree(W.buf[k]);
  if (W.pinned) cudaFreeHost(W.pinned);
  W = SlicedWork();
}

}  // namespace cabac
