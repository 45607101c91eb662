// cabac_slice.cuh - K9: whole-substream decoder (cabac_decode_slices).
// Included by cabac_kernels.cu after the lockstep residual decoder.
//
// One warp per substream (a slice or a tile: an independently initialised and terminated substream,
// EncSlice.cpp:2046-2054).  The warp runs the CTU-level parser of cabac_syntax.cuh in lockstep: all 32
// lanes hold the same arithmetic-decoder state and take the same branches, exactly as in
// decode_residual_kernel; inside residual_coding the lanes split the coefficient group's template work
// (lsResidualRegular / lsResidualTS).  Per warp in shared memory: all 379 contexts as expanded entries
// (6 KB; the 230 residual ones first, in the dense order the residual decoder indexes), the 32x32
// coefficient tile (4 KB), and the neighbour lines of the partitioner contexts (one 16-bit word per 4
// luma samples of the substream's width, per tree: <= 4 KB for 4096 samples).  Parsing a substream
// is a dependent chain from the first bin to the last - every bin picks the context of the next - so one
// substream cannot be spread over more than one warp; the launch is parallel over substreams only and a
// batch of few, long slices is latency-bound.
#pragma once
#include <new>

#include "cabac_syntax.cuh"

namespace cabac {

constexpr int SLICE_WARPS_PER_BLOCK = 2;
constexpr int SLICE_MAX_LINE = 1024;  // 4096 luma samples
constexpr int SLICE_MAX_CTU_COLS = 128;

// Neighbour state of the walk (cabac_syntax.cuh), per warp: used by the decoder and by the writer-direction emitter.
struct SliceLines {
  uint16_t above[2][SLICE_MAX_LINE];    // per tree: what the CU above left behind, one word per 4 luma samples
  uint16_t left[2][32];
  uint16_t edgeLeft[32];                // luma: right edge (x + w) of the CU last written at a row
  uint8_t modeAbove[SLICE_MAX_LINE];    // luma intra direction neighbours see (MPM derivation)
  uint8_t modeLeft[32];
  uint8_t lumaMap[256];                 // luma directions of the current 64x64 node, 4x4 granularity (chroma candidates)
  uint8_t ctuLine[SLICE_MAX_CTU_COLS];
};

struct __align__(16) SliceWarpSmem {
  LsDecoder dec;  // the arithmetic decoder's state: identical on all lanes, kept here between calls (see sliceBin)
  int32_t tile[32 * 32];
  uint4 ctx[CABAC_NUM_CTX];  // [0, RCTX): residual contexts in rctx() order; [RCTX, 379): the others
  uint32_t rowCtx[CABAC_NUM_CTX];  // WPP: the states (entry .x) after the first CTU of the current row
  SliceLines lines;
};

// reference context id -> slot in SliceWarpSmem::ctx
__host__ __device__ constexpr int sliceCtxSlot(int id) {
  return id < 86 ? RCTX + id
                 : id < 292 ? id - 86
                            : id < 310 ? RCTX + 86 + (id - 292)
                                       : id < 312 ? 206 + (id - 310) : id < 357 ? RCTX + 104 + (id - 312) : 208 + (id - 357);
}
static_assert(sliceCtxSlot(86) == rctx(86) && sliceCtxSlot(291) == rctx(291) && sliceCtxSlot(310) == rctx(310) &&
                  sliceCtxSlot(357) == rctx(357) && sliceCtxSlot(378) == rctx(378) && sliceCtxSlot(356) == CABAC_NUM_CTX - 1 &&
                  sliceCtxSlot(0) == RCTX && sliceCtxSlot(292) == RCTX + 86 && sliceCtxSlot(312) == RCTX + 104,
              "context slot map");

// residual_coding of one block (cabac_reader.cpp:2647-2735 with ts_flag :2737-2752): the body of
// decode_residual_kernel's TU loop.  Out of line - the CU parser calls it from four places.
__device__ __noinline__ uint32_t sliceResidual(SliceWarpSmem *sm, const cabac_tu_desc d, uint32_t subFlags,
                                               int32_t *__restrict__ coeffs, int lane) {
  __syncwarp();
  LsDecoder dec = sm->dec;
  uint4 *ctx = sm->ctx;
  int32_t *tile = sm->tile;
  const TuGeom g = makeGeom(d);
  const int n = g.tw * g.th;
  for (int i = lane; i < n; i += 32) tile[i] = 0;
  __syncwarp();
  bool ts = (d.flags & CABAC_TU_BDPCM) != 0;
  if (d.flags & CABAC_TU_TS_ALLOWED) ts = lsDecodeBin(dec, ctx + R_TS_FLAG + (g.comp ? 1 : 0)) != 0;
  uint32_t res = ts ? CABAC_TUR_TS : 0;
  if (ts && !(subFlags & CABAC_SUB_TSRC_DISABLED))
    lsResidualTS(dec, ctx, tile, g, (d.flags & CABAC_TU_BDPCM) != 0, lane);
  else
    res |= lsResidualRegular(dec, ctx, tile, g, subFlags, ts, lane);
  __syncwarp();
  const int total = g.w * g.h;
  int32_t *dst = coeffs + d.coeff_offset;
  for (int i = lane; i < total; i += 32) {
    const int y = i >> g.log2w, x = i & (g.w - 1);
    dst[i] = (x < g.tw && y < g.th) ? tile[(y << g.log2tw) + x] : 0;
  }
  __syncwarp();
  sm->dec = dec;
  __syncwarp();
  return res;
}

// decodeBin / decodeBinsEP / decodeBinTrm out of line for the CU parser: it has some sixty + seventy call sites.
// Inlined (about a hundred instructions each) the kernel was 485 KB of code, and with the decoder state passed by
// value every call site still marshalled twelve words each way (468 KB).  The state therefore lives in shared
// memory between calls: every lane loads it, steps it identically and stores the same values back (the contexts
// are shared the same way).  The __syncwarp keeps the lanes together across the store / load pair.
__device__ __noinline__ uint32_t sliceBin(SliceWarpSmem *sm, int slot) {
  __syncwarp();
  LsDecoder dec = sm->dec;
  const uint32_t b = dec.decodeBin(sm->ctx + slot);
  sm->dec = dec;
  return b;
}
__device__ __noinline__ uint32_t sliceEp(SliceWarpSmem *sm, int n) {
  __syncwarp();
  LsDecoder dec = sm->dec;
  const uint32_t v = n <= 16 ? dec.ep(n) : dec.epLong(uint32_t(n));
  sm->dec = dec;
  return v;
}
__device__ __noinline__ uint32_t sliceTrm(SliceWarpSmem *sm) {
  __syncwarp();
  LsDecoder dec = sm->dec;
  const uint32_t b = dec.decodeBinTrm();
  sm->dec = dec;
  return b;
}

#ifdef CABAC_SLICE_TRACE  // debug builds only (tests/tools/slice_trace.py): the first substream's calls as command words,
                          // recorded by lane 0 and by lane 5 (lockstep: the two must agree)
__device__ uint32_t g_sliceTrace[2][1 << 17];
__device__ uint32_t g_sliceTraceN[2];
#define SLICE_TRACE(w)                                            \
  do {                                                            \
    if ((lane == 0 || lane == 5) && traceOn) {                    \
      const int t_ = lane ? 1 : 0;                                \
      const uint32_t i_ = g_sliceTraceN[t_]++;                    \
      if (i_ < (1u << 17)) g_sliceTrace[t_][i_] = (w);            \
    }                                                             \
    __syncwarp();                                                 \
  } while (0)
#else
#define SLICE_TRACE(w) ((void)0)
#endif

struct SliceEngine {
  static constexpr bool kEncode = false;  // decoder direction: the `want` arguments of the walk are ignored
  __device__ __forceinline__ const cabac_cu_desc &getCu(uint32_t) const { return *cus; }
  __device__ __forceinline__ const cabac_ctu_desc &getCtu(uint32_t) const { return *ctus; }
  bool traceOn = false;
  SliceWarpSmem &sm;
  const uint32_t subFlags;
  cabac_cu_desc *cus;
  cabac_tu_desc *tus;
  uint8_t *tuFlags;
  cabac_ctu_desc *ctus;
  int32_t *coeffs;
  bool dead = false;

  // The engine object itself sits in shared memory, one per warp, written identically by all lanes - so nothing in
  // it may depend on the lane.
  __device__ __forceinline__ SliceEngine(SliceWarpSmem &s, uint32_t f, cabac_cu_desc *c, cabac_tu_desc *t, uint8_t *tf,
                                         cabac_ctu_desc *ct, int32_t *co)
      : sm(s), subFlags(f), cus(c), tus(t), tuFlags(tf), ctus(ct), coeffs(co) {}
  __device__ __forceinline__ static int laneId() { return int(threadIdx.x & 31u); }

  __device__ __forceinline__ uint32_t bin(int id, uint32_t = 0) {
    const int lane = laneId();
    (void)lane;
    const uint32_t b = sliceBin(&sm, sliceCtxSlot(id));
    SLICE_TRACE(CABAC_MK_CTX(id, b));
    return b;
  }
  __device__ __forceinline__ uint32_t ep(int n, uint32_t = 0) {
    const int lane = laneId();
    (void)lane;
    const uint32_t v = sliceEp(&sm, n);
    SLICE_TRACE(CABAC_MK_EP(v, n));
    return v;
  }
  __device__ __forceinline__ uint32_t trm(uint32_t = 1) {
    const int lane = laneId();
    (void)lane;
    const uint32_t b = sliceTrm(&sm);
    SLICE_TRACE(CABAC_MK_TRM(b));
    return b;
  }
  __device__ __forceinline__ uint32_t residual(uint32_t idx, const cabac_tu_desc &d) {
    const int lane = laneId();
    (void)lane;
    const uint32_t res = sliceResidual(&sm, d, subFlags, coeffs, lane);
    SLICE_TRACE(CABAC_MK_TU() | sm.dec.nbins);
    if (lane == 0) {
      cabac_tu_desc o = d;
      if (res & CABAC_TUR_TS) o.flags |= CABAC_TU_TS;
      tus[idx] = o;
      if (tuFlags) tuFlags[idx] = uint8_t(res);
    }
    __syncwarp();  // lockstep: lane 0 must not fall behind - the next bin reads and rewrites a shared context
    return res;
  }
  __device__ __forceinline__ uint16_t *aboveLine(int ch) { return sm.lines.above[ch]; }
  __device__ __forceinline__ uint16_t *leftCol(int ch) { return sm.lines.left[ch]; }
  __device__ __forceinline__ uint8_t *ctuLine() { return sm.lines.ctuLine; }
  __device__ __forceinline__ uint8_t *modeAbove() { return sm.lines.modeAbove; }
  __device__ __forceinline__ uint8_t *modeLeft() { return sm.lines.modeLeft; }
  __device__ __forceinline__ uint16_t *edgeLeft() { return sm.lines.edgeLeft; }
  __device__ __forceinline__ uint8_t *lumaMap() { return sm.lines.lumaMap; }
  __device__ __forceinline__ void fill8(uint8_t *p, int n, uint8_t v) {
    const int lane = laneId();
    for (int i = lane; i < n; i += 32) p[i] = v;
    __syncwarp();
  }
  __device__ __forceinline__ void fillMap(int bx, int by, int bw, int bh, uint8_t v) {
    const int lane = laneId();
    for (int i = lane; i < bw * bh; i += 32) sm.lines.lumaMap[((by + i / bw) << 4) + bx + i % bw] = v;
    __syncwarp();
  }
  __device__ __forceinline__ void ctuLineStore(int col, uint8_t v) {
    const int lane = laneId();
    (void)lane;
    if (lane == 0 && col < SLICE_MAX_CTU_COLS) sm.lines.ctuLine[col] = v;
    __syncwarp();
  }
  __device__ __forceinline__ void fill(uint16_t *p, int n, uint16_t v) {
    const int lane = laneId();
    (void)lane;
    for (int i = lane; i < n; i += 32) p[i] = v;
    __syncwarp();
  }
  __device__ __forceinline__ void putCu(uint32_t i, const cabac_cu_desc &c) {
    const int lane = laneId();
    (void)lane;
    if (lane == 0 && cus) cus[i] = c;
    __syncwarp();
  }
  __device__ __forceinline__ void putCtu(uint32_t i, const cabac_ctu_desc &c) {
    const int lane = laneId();
    (void)lane;
    if (lane == 0 && ctus) ctus[i] = c;
    __syncwarp();
  }
  // WPP row hand-over (see SliceParser::run)
  __device__ __forceinline__ bool rows() const { return false; }
  __device__ __forceinline__ void saveRowContexts() {
    const int lane = laneId();
    (void)lane;
    __syncwarp();
    for (int k = lane; k < CABAC_NUM_CTX; k += 32) sm.rowCtx[k] = sm.ctx[k].x;
    __syncwarp();
  }
  __device__ __forceinline__ void nextRow() {
    const int lane = laneId();
    (void)lane;
    __syncwarp();
    const LsDecoder old = sm.dec;
    const uint32_t used = rowStart + (old.bitsRead() >> 3) + 1u;  // the row's bytes: floor(bits / 8) + 1
    for (int k = lane; k < CABAC_NUM_CTX; k += 32) sm.ctx[k].x = sm.rowCtx[k];
    __syncwarp();
    LsDecoder d;
    if (used >= sliceBytes) {
      d = old;
      d.status |= CABAC_SS_IN_OVERRUN;
      dead = true;
    } else {
      d.start(sliceIn + used, sliceBytes - used);
      d.status = old.status;
      d.nbins = old.nbins;
      rowStart = used;
    }
    sm.dec = d;
    __syncwarp();
  }
  const uint8_t *sliceIn = nullptr;
  uint32_t sliceBytes = 0, rowStart = 0;
  __device__ __forceinline__ void abort() { dead = true; }
  __device__ __forceinline__ bool aborted() const { return dead; }
};

__global__ void __launch_bounds__(SLICE_WARPS_PER_BLOCK * 32)
decode_slice_kernel(const uint8_t *__restrict__ in, const uint64_t *__restrict__ inOffsets,
                    const cabac_substream_params *__restrict__ params,
                    const cabac_slice_params *__restrict__ sliceParams, size_t nSub,
                    const cabac_ctx_model *__restrict__ ctxIn, const uint64_t *__restrict__ cuOffsets,
                    const uint64_t *__restrict__ tuOffsets, const uint64_t *__restrict__ ctuOffsets,
                    const uint64_t *__restrict__ coeffOffsets, cabac_cu_desc *__restrict__ cus,
                    cabac_tu_desc *__restrict__ tus, uint8_t *__restrict__ tuFlags, cabac_ctu_desc *__restrict__ ctus,
                    int32_t *__restrict__ coeffs, uint32_t *__restrict__ counts, uint32_t *__restrict__ numBins,
                    uint32_t *__restrict__ bitsRead, cabac_ctx_model *__restrict__ ctxOut, uint32_t *__restrict__ status,
                    const uint32_t *__restrict__ order, uint32_t *cursor) {
  using Parser = cabac_syntax::SliceParser<SliceEngine>;
  __shared__ SliceWarpSmem smem[SLICE_WARPS_PER_BLOCK];
  // engine + parser objects: one copy per warp in shared memory (all lanes write the same values), instead of 32
  // per-lane copies in local memory - 12 resident warps x 32 lanes of stack no longer compete for the L1
  __shared__ __align__(16) unsigned char engBuf[SLICE_WARPS_PER_BLOCK][(sizeof(SliceEngine) + 15) & ~size_t(15)];
  __shared__ __align__(16) unsigned char psBuf[SLICE_WARPS_PER_BLOCK][(sizeof(Parser) + 15) & ~size_t(15)];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  SliceWarpSmem &sm = smem[wib];
  size_t sub;
  while (nextSubstream(order, cursor, nSub, lane, sub)) {
    const cabac_substream_params p = params[sub];
    const cabac_slice_params sp = sliceParams[sub];
    const cabac_ctx_model *src = ((p.flags & CABAC_SUB_CTX_IN) && ctxIn) ? ctxIn + sub * CABAC_NUM_CTX : nullptr;
    const int qp = clipQp(p.qp);
    for (int id = lane; id < CABAC_NUM_CTX; id += 32) {
      const uint32_t win = CABAC_CTX_INIT_TABLE[3][id];
      sm.ctx[sliceCtxSlot(id)] = src ? ctx_entry(src[id].state0, src[id].state1, win)
                                     : ctx_entry_init(qp, CABAC_CTX_INIT_TABLE[p.init_id][id], win);
    }
    __syncwarp();
    const uint64_t cu0 = cuOffsets[sub], tu0 = tuOffsets[sub], ctu0 = ctuOffsets[sub], co0 = coeffOffsets[sub];
    __syncwarp();
    SliceEngine &eng = *new (engBuf[wib]) SliceEngine(sm, p.flags, cus ? cus + cu0 : nullptr, tus + tu0,
                                                      tuFlags ? tuFlags + tu0 : nullptr, ctus ? ctus + ctu0 : nullptr, coeffs);
    __syncwarp();
    {
      LsDecoder d0;
      eng.sliceIn = in + inOffsets[sub];
      eng.sliceBytes = uint32_t(inOffsets[sub + 1] - inOffsets[sub]);
      d0.start(eng.sliceIn, eng.sliceBytes);
      sm.dec = d0;
      __syncwarp();
    }
#ifdef CABAC_SLICE_TRACE
    eng.traceOn = sub == 0;
    if (sub == 0 && lane == 0) g_sliceTraceN[0] = g_sliceTraceN[1] = 0;
#endif
    Parser &ps = *new (psBuf[wib]) Parser(eng, sp, uint32_t(cuOffsets[sub + 1] - cu0), uint32_t(tuOffsets[sub + 1] - tu0),
                                          uint32_t(ctuOffsets[sub + 1] - ctu0), coeffOffsets[sub + 1] - co0, co0);
    __syncwarp();
    if (((sp.x1 - sp.x0 + 3) >> 2) > uint32_t(SLICE_MAX_LINE) ||
        ((sp.x1 - sp.x0) >> sp.log2_ctu) >= uint32_t(SLICE_MAX_CTU_COLS))
      ps.status |= CABAC_SS_UNSUPPORTED;
    else
      ps.run((p.flags & CABAC_SUB_TERMINATE) != 0);
    __syncwarp();
    LsDecoder dEnd = sm.dec;
    dEnd.checkOverrun();
    if (lane == 0) {
      if (counts) {
        counts[3 * sub + 0] = ps.nCu;
        counts[3 * sub + 1] = ps.nTu;
        counts[3 * sub + 2] = uint32_t(ps.nCoeff);
      }
      if (numBins) numBins[sub] = dEnd.nbins;
      if (bitsRead) bitsRead[sub] = 8u * eng.rowStart + dEnd.bitsRead();
      if (status) status[sub] = dEnd.status | ps.status;
    }
    if (ctxOut) {
      __syncwarp();
      cabac_ctx_model *dst = ctxOut + sub * CABAC_NUM_CTX;
      for (int id = lane; id < CABAC_NUM_CTX; id += 32) {
        const uint32_t x = sm.ctx[sliceCtxSlot(id)].x;
        cabac_ctx_model m;
        m.state0 = uint16_t(x >> 16);
        m.state1 = uint16_t(x & 0xFFFFu);
        dst[id] = m;
      }
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------
// K10: writer direction (cabac_encode_slices).  The same walk on an engine that turns the caller's CU / TU / CTU
// records into the BinEncIf call sequence of the slice - command words with CABAC_CMD_TU markers - which the
// stream encoder (cabac_encode_stream: all encoder kernels) then codes.  No arithmetic coding here: one warp per
// slice writes into a staging region bounded by the slice's record counts (emitBound), emit_compact_kernel packs
// the slices at the scan of their word counts.
// ---------------------------------------------------------------------------
// Row substreams of the slice being written (WPP: one per CTU row; otherwise the slice itself): what the stream
// encoder's wavefront mode takes per substream.  cmdOff is relative to the slice's first word until
// emit_compact_kernel has placed the slice.
struct SliceRowOut {
  uint64_t *cmdOff;      // [row]: first command word (absolute in the command buffer)
  uint64_t *tuOff;       // [row]: first TU (absolute in tus)
  cabac_wpp_sync *sync;  // [row]: predecessor and the hand-over point (after the row's first CTU)
  uint32_t base, count;  // the slice's rows are [base, base + count)
  uint64_t tuBase;
  bool allow;            // the caller described the rows (cabac_encode_slices_wpp): WPP slices can be written
};

struct SliceEmitEngine {
  static constexpr bool kEncode = true;
  SliceLines &ln;
  const cabac_cu_desc *cusIn;
  const cabac_tu_desc *tusIn;
  const uint8_t *tuFlagsIn;
  const cabac_ctu_desc *ctusIn;
  uint32_t nCuIn, nTuIn, nCtuIn;
  uint32_t *cmds;  // the slice's staging region (capWords words: the bound of emitBound())
  uint32_t capWords;
  SliceRowOut ro;
  uint32_t n = 0, row = 0, rowStartN = 0, rowStartTu = 0, tuSeen = 0;
  bool dead = false, bad = false;

  __device__ __forceinline__ SliceEmitEngine(SliceLines &l, const cabac_cu_desc *c, uint32_t nc, const cabac_tu_desc *t,
                                             const uint8_t *tf, uint32_t nt, const cabac_ctu_desc *ct, uint32_t nct,
                                             uint32_t *out, uint32_t cap, const SliceRowOut &r)
      : ln(l), cusIn(c), tusIn(t), tuFlagsIn(tf), ctusIn(ct), nCuIn(nc), nTuIn(nt), nCtuIn(nct), cmds(out), capWords(cap),
        ro(r) {}
  __device__ __forceinline__ bool rows() const { return ro.allow; }
  __device__ __forceinline__ static int laneId() { return int(threadIdx.x & 31u); }
  __device__ __forceinline__ void push(uint32_t w) {
    if (laneId() == 0 && n < capWords) cmds[n] = w;
    n++;
  }
  __device__ __forceinline__ cabac_cu_desc getCu(uint32_t i) {
    if (i >= nCuIn) {
      bad = dead = true;
      return cabac_cu_desc{};
    }
    return cusIn[i];
  }
  __device__ __forceinline__ cabac_ctu_desc getCtu(uint32_t i) {
    if (i >= nCtuIn) {
      bad = dead = true;
      return cabac_ctu_desc{};
    }
    return ctusIn[i];
  }
  __device__ __forceinline__ uint32_t bin(int ctx, uint32_t want) {
    push(CABAC_MK_CTX(ctx, want & 1u));
    return want & 1u;
  }
  __device__ __forceinline__ uint32_t ep(int nb, uint32_t want) {
    const uint32_t v = nb < 32 ? want & ((1u << nb) - 1u) : want;
    int left_ = nb;
    while (left_ > 16) {
      push(CABAC_MK_EP((v >> (left_ - 16)) & 0xFFFFu, 16));
      left_ -= 16;
    }
    push(CABAC_MK_EP(v & ((1u << left_) - 1u), left_));
    return v;
  }
  __device__ __forceinline__ uint32_t trm(uint32_t want) {
    push(CABAC_MK_TRM(want & 1u));
    return want & 1u;
  }
  __device__ __forceinline__ uint32_t residual(uint32_t idx, const cabac_tu_desc &d) {
    if (idx >= nTuIn) {
      bad = dead = true;
      return 0;
    }
    const cabac_tu_desc t = tusIn[idx];
    // the shape and the flags are the walk's; transform_skip_flag is the record's, where the syntax has one (a BDPCM
    // block is transform-skipped by inference)
    const bool tsOk = !(t.flags & CABAC_TU_TS) || (d.flags & (CABAC_TU_TS_ALLOWED | CABAC_TU_BDPCM));
    if (t.log2_w != d.log2_w || t.log2_h != d.log2_h || t.comp_id != d.comp_id || ((t.flags ^ d.flags) & uint8_t(~CABAC_TU_TS)) ||
        !tsOk) {
      bad = dead = true;
      return 0;
    }
    push(CABAC_MK_TU());
    tuSeen = idx + 1;
    return tuFlagsIn[idx];
  }
  __device__ __forceinline__ uint16_t *aboveLine(int ch) { return ln.above[ch]; }
  __device__ __forceinline__ uint16_t *leftCol(int ch) { return ln.left[ch]; }
  __device__ __forceinline__ uint8_t *ctuLine() { return ln.ctuLine; }
  __device__ __forceinline__ uint8_t *modeAbove() { return ln.modeAbove; }
  __device__ __forceinline__ uint8_t *modeLeft() { return ln.modeLeft; }
  __device__ __forceinline__ uint16_t *edgeLeft() { return ln.edgeLeft; }
  __device__ __forceinline__ uint8_t *lumaMap() { return ln.lumaMap; }
  __device__ __forceinline__ void ctuLineStore(int col, uint8_t v) {
    if (laneId() == 0 && col < SLICE_MAX_CTU_COLS) ln.ctuLine[col] = v;
    __syncwarp();
  }
  __device__ __forceinline__ void fill(uint16_t *p, int cnt, uint16_t v) {
    for (int i = laneId(); i < cnt; i += 32) p[i] = v;
    __syncwarp();
  }
  __device__ __forceinline__ void fill8(uint8_t *p, int cnt, uint8_t v) {
    for (int i = laneId(); i < cnt; i += 32) p[i] = v;
    __syncwarp();
  }
  __device__ __forceinline__ void fillMap(int bx, int by, int bw, int bh, uint8_t v) {
    for (int i = laneId(); i < bw * bh; i += 32) ln.lumaMap[((by + i / bw) << 4) + bx + i % bw] = v;
    __syncwarp();
  }
  __device__ __forceinline__ void putCu(uint32_t, const cabac_cu_desc &) {}
  __device__ __forceinline__ void putCtu(uint32_t, const cabac_ctu_desc &) {}
  // WPP: the contexts are handed to the next row after the row's first CTU (EncSlice.cpp:2085-2089) ...
  __device__ __forceinline__ void saveRowContexts() {
    if (laneId() == 0 && row < ro.count) {
      ro.sync[ro.base + row].sync_cmds = n - rowStartN;
      ro.sync[ro.base + row].sync_tus = tuSeen - rowStartTu;
    }
  }
  // ... and the row ends with end_of_subset_one_bit, finish() and the alignment: the stream encoder's
  // (CABAC_SUB_TERMINATE | CABAC_SUB_ALIGN on every row), so the terminating bin just pushed is taken back
  __device__ __forceinline__ void nextRow() {
    n--;
    row++;
    rowStartN = n;
    rowStartTu = tuSeen;
    if (row >= ro.count) {
      bad = dead = true;
      return;
    }
    if (laneId() == 0) {
      ro.cmdOff[ro.base + row] = n;
      ro.tuOff[ro.base + row] = ro.tuBase + tuSeen;
    }
  }
  __device__ __forceinline__ void abort() { dead = true; }
  __device__ __forceinline__ bool aborted() const { return dead; }
};

constexpr int EMIT_WARPS_PER_BLOCK = 4;

// One walk per slice into its staging region (stageOffsets: exclusive scan of emitBound(), the most words a slice of
// that many CUs / TUs / CTUs can need), with the row-level arrays the stream encoder takes (one row per slice; one
// per CTU row of a WPP slice); emit_compact_kernel then packs the slices at the scanned word counts.
struct EmitRowArrays {
  const uint64_t *rowOffsets;  // [nSub + 1] or nullptr: every slice is one row (WPP slices refused)
  uint64_t *cmdOff, *tuOff;    // [nRows + 1]
  cabac_wpp_sync *sync;        // [nRows]
  cabac_substream_params *params;
  uint32_t *slice;             // [nRows]: the slice a row belongs to
  cabac_ctx_model *ctxIn;      // [nRows][379] or nullptr
};

// Words per CU: <= 12 for the split flags of its share of the tree, 14 luma / 7 chroma prediction syntax, 16
// cu_qp_delta, 6 chroma QP offset, 13 cbf / joint flags, 12 TU markers, 6 LFNST / MTS; per CTU: SAO 3 x (4 x 31
// offset bins + 7) + ALF / CC-ALF 35.
__host__ __device__ inline uint64_t emitBound(uint64_t nCu, uint64_t nTu, uint64_t nCtu) {
  return 80 * nCu + nTu + 450 * nCtu + 64;
}

__global__ void emit_bound_kernel(const uint64_t *__restrict__ cuOffsets, const uint64_t *__restrict__ tuOffsets,
                                  const uint64_t *__restrict__ ctuOffsets, size_t n, uint32_t *__restrict__ bound) {
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) {
    const uint64_t b = emitBound(cuOffsets[i + 1] - cuOffsets[i], tuOffsets[i + 1] - tuOffsets[i], ctuOffsets[i + 1] - ctuOffsets[i]);
    bound[i] = b > 0xFFFFFFFFull ? 0xFFFFFFFFu : uint32_t(b);
  }
}

__global__ void __launch_bounds__(EMIT_WARPS_PER_BLOCK * 32)
emit_slice_kernel(const cabac_substream_params *__restrict__ params, const cabac_slice_params *__restrict__ sliceParams,
                  size_t nSub, const cabac_cu_desc *__restrict__ cus, const uint64_t *__restrict__ cuOffsets,
                  const cabac_tu_desc *__restrict__ tus, const uint8_t *__restrict__ tuFlags,
                  const uint64_t *__restrict__ tuOffsets, const cabac_ctu_desc *__restrict__ ctus,
                  const uint64_t *__restrict__ ctuOffsets, const cabac_ctx_model *__restrict__ ctxIn,
                  uint32_t *__restrict__ stage, const uint64_t *__restrict__ stageOffsets, uint32_t *__restrict__ counts,
                  uint32_t *__restrict__ emitStatus, EmitRowArrays ra, const uint32_t *__restrict__ order,
                  uint32_t *cursor) {
  using Emitter = cabac_syntax::SliceParser<SliceEmitEngine>;
  __shared__ SliceLines lines[EMIT_WARPS_PER_BLOCK];
  __shared__ __align__(16) unsigned char engBuf[EMIT_WARPS_PER_BLOCK][(sizeof(SliceEmitEngine) + 15) & ~size_t(15)];
  __shared__ __align__(16) unsigned char psBuf[EMIT_WARPS_PER_BLOCK][(sizeof(Emitter) + 15) & ~size_t(15)];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  size_t sub = 0;
  while (nextSubstream(order, cursor, nSub, lane, sub)) {  // longest slice first (by CU count), warps pull from a cursor
    const cabac_slice_params sp = sliceParams[sub];
    const uint64_t cu0 = cuOffsets[sub], tu0 = tuOffsets[sub], ctu0 = ctuOffsets[sub];
    SliceRowOut ro{};
    ro.allow = ra.rowOffsets != nullptr;
    ro.base = ra.rowOffsets ? uint32_t(ra.rowOffsets[sub]) : uint32_t(sub);
    ro.count = ra.rowOffsets ? uint32_t(ra.rowOffsets[sub + 1] - ra.rowOffsets[sub]) : 1u;
    const uint32_t ctuRows = (sp.y1 - sp.y0 + (1u << sp.log2_ctu) - 1) >> sp.log2_ctu;
    const bool rowsOk = ro.count == ((sp.tools & CABAC_SL_WPP) ? ctuRows : 1u);
    ro.cmdOff = ra.cmdOff;
    ro.tuOff = ra.tuOff;
    ro.sync = ra.sync;
    ro.tuBase = tu0;
    const cabac_substream_params p = params[sub];
    for (uint32_t r = lane; r < ro.count; r += 32) {  // defaults; the walk overwrites offsets and hand-over points
      cabac_substream_params pr = p;
      if (r) pr.flags = uint8_t(pr.flags & ~CABAC_SUB_CTX_IN);  // rows below the first start from their predecessor
      ra.params[ro.base + r] = pr;
      ra.slice[ro.base + r] = uint32_t(sub);
      ra.sync[ro.base + r] = cabac_wpp_sync{r ? ro.base + r - 1 : CABAC_WPP_NONE, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u};
      ra.cmdOff[ro.base + r] = 0;
      ra.tuOff[ro.base + r] = tu0;
    }
    if (ra.ctxIn && ctxIn && (p.flags & CABAC_SUB_CTX_IN))
      for (int k = lane; k < CABAC_NUM_CTX; k += 32) ra.ctxIn[size_t(ro.base) * CABAC_NUM_CTX + k] = ctxIn[sub * CABAC_NUM_CTX + k];
    __syncwarp();
    const uint64_t room = stageOffsets[sub + 1] - stageOffsets[sub];
    SliceEmitEngine &eng = *new (engBuf[wib]) SliceEmitEngine(
        lines[wib], cus + cu0, uint32_t(cuOffsets[sub + 1] - cu0), tus + tu0, tuFlags + tu0, uint32_t(tuOffsets[sub + 1] - tu0),
        ctus + ctu0, uint32_t(ctuOffsets[sub + 1] - ctu0), stage + stageOffsets[sub], uint32_t(room), ro);
    __syncwarp();
    Emitter &ps = *new (psBuf[wib]) Emitter(eng, sp, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, ~uint64_t(0), 0);
    __syncwarp();
    uint32_t st = 0;
    if (((sp.x1 - sp.x0 + 3) >> 2) > uint32_t(SLICE_MAX_LINE) || ((sp.x1 - sp.x0) >> sp.log2_ctu) >= uint32_t(SLICE_MAX_CTU_COLS))
      st |= CABAC_SS_UNSUPPORTED;
    else if (!rowsOk)
      st |= ro.allow ? CABAC_SS_BAD_DESC : CABAC_SS_UNSUPPORTED;  // rows not described: a WPP slice cannot be written
    else
      ps.run(false);  // the terminating bin and finish() are the stream encoder's (CABAC_SUB_TERMINATE)
    st |= ps.status | (eng.bad ? uint32_t(CABAC_SS_BAD_DESC) : 0u) | (eng.n > eng.capWords ? uint32_t(CABAC_SS_CAPACITY) : 0u);
    if (!st && (ps.nCu != eng.nCuIn || ps.nTu != eng.nTuIn || eng.row + 1 != ro.count)) st |= CABAC_SS_BAD_DESC;  // left-overs
    __syncwarp();
    if (st)  // a refused slice codes nothing (its status says why): empty rows
      for (uint32_t r = lane; r < ro.count; r += 32) ra.cmdOff[ro.base + r] = 0, ra.tuOff[ro.base + r] = tu0;
    if (lane == 0) {
      counts[sub] = st ? 0u : eng.n;
      emitStatus[sub] = st;
    }
    __syncwarp();
  }
}

// Packs the staged slices at cmdOffsets (the scan of their word counts), makes the rows' command offsets absolute
// and closes the row arrays.  A slice that does not fit the caller's scratch is flagged and not copied.
__global__ void __launch_bounds__(256)
emit_compact_kernel(const uint32_t *__restrict__ stage, const uint64_t *__restrict__ stageOffsets,
                    const uint64_t *__restrict__ cmdOffsets, size_t nSub, uint64_t capWords, uint32_t *__restrict__ cmds,
                    uint32_t *__restrict__ emitStatus, EmitRowArrays ra, const uint64_t *__restrict__ tuOffsets) {
  for (size_t sub = blockIdx.x; sub < nSub; sub += gridDim.x) {
    const uint64_t c0 = cmdOffsets[sub], c1 = cmdOffsets[sub + 1];
    const uint32_t base = ra.rowOffsets ? uint32_t(ra.rowOffsets[sub]) : uint32_t(sub);
    const uint32_t count = ra.rowOffsets ? uint32_t(ra.rowOffsets[sub + 1] - ra.rowOffsets[sub]) : 1u;
    for (uint32_t r = threadIdx.x; r < count; r += blockDim.x) ra.cmdOff[base + r] += c0;
    if (threadIdx.x == 0 && sub + 1 == nSub) {
      ra.cmdOff[base + count] = c1;
      ra.tuOff[base + count] = tuOffsets[nSub];
    }
    if (c1 > capWords) {
      if (threadIdx.x == 0) emitStatus[sub] |= CABAC_SS_CAPACITY;
      continue;
    }
    const uint32_t *src = stage + stageOffsets[sub];
    for (uint64_t i = threadIdx.x; i < c1 - c0; i += blockDim.x) cmds[c0 + i] = src[i];
  }
}

__global__ void or_status_kernel(uint32_t *__restrict__ status, const uint32_t *__restrict__ extra,
                                 const uint32_t *__restrict__ rowSlice, size_t nRows) {
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < nRows; i += size_t(gridDim.x) * blockDim.x) {
    const uint32_t e = extra[rowSlice[i]];
    if (e) status[i] |= e;
  }
}

cudaError_t launchEmitSlices(cudaStream_t s, int device, const Sched &sc, const cabac_substream_params *params,
                             const cabac_slice_params *sliceParams, size_t nSub, const cabac_cu_desc *cus,
                             const uint64_t *cuOffsets, const cabac_tu_desc *tus, const uint8_t *tuFlags,
                             const uint64_t *tuOffsets, const cabac_ctu_desc *ctus, const uint64_t *ctuOffsets,
                             const cabac_ctx_model *ctxIn, uint32_t *stage, uint64_t *stageOffsets, uint32_t *cmds,
                             uint64_t capWords, uint32_t *counts, uint32_t *emitStatus, uint64_t *cmdOffsets,
                             const uint64_t *rowOffsets, uint64_t *rowCmdOff, uint64_t *rowTuOff, cabac_wpp_sync *rowSync,
                             cabac_substream_params *rowParams, uint32_t *rowSlice, cabac_ctx_model *rowCtxIn) {
  if (nSub == 0) return cudaSuccess;
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  const size_t want = (nSub + EMIT_WARPS_PER_BLOCK - 1) / EMIT_WARPS_PER_BLOCK, cap = size_t(sms) * 8;
  const int blocks = int(want < cap ? want : cap);
  const EmitRowArrays ra{rowOffsets, rowCmdOff, rowTuOff, rowSync, rowParams, rowSlice, rowCtxIn};
  emit_bound_kernel<<<costGrid(nSub, 256), 256, 0, s>>>(cuOffsets, tuOffsets, ctuOffsets, nSub, counts);
  entry_points_kernel<<<1, 1024, 0, s>>>(counts, nullptr, nSub, stageOffsets);
  cost_bytes_kernel<<<costGrid(nSub, 256), 256, 0, s>>>(cuOffsets, nSub, sc.cost);  // CUs per slice: the walk's length
  cudaError_t e = launchOrder(s, sc, nSub);
  if (e != cudaSuccess) return e;
  emit_slice_kernel<<<blocks, EMIT_WARPS_PER_BLOCK * 32, 0, s>>>(params, sliceParams, nSub, cus, cuOffsets, tus, tuFlags,
                                                                tuOffsets, ctus, ctuOffsets, ctxIn, stage, stageOffsets, counts,
                                                                emitStatus, ra, sc.order, sc.cursor);
  entry_points_kernel<<<1, 1024, 0, s>>>(counts, nullptr, nSub, cmdOffsets);
  const size_t cb = nSub < size_t(sms) * 8 ? nSub : size_t(sms) * 8;
  emit_compact_kernel<<<unsigned(cb), 256, 0, s>>>(stage, stageOffsets, cmdOffsets, nSub, capWords, cmds, emitStatus, ra,
                                                  tuOffsets);
  return cudaGetLastError();
}

uint64_t emitStageWords(uint64_t nCu, uint64_t nTu, uint64_t nCtu, uint64_t nSub) {
  return emitBound(nCu, nTu, nCtu) + 64 * nSub;
}

cudaError_t launchOrStatus(cudaStream_t s, uint32_t *status, const uint32_t *extra, const uint32_t *rowSlice, size_t nRows) {
  if (!nRows || !status) return cudaSuccess;
  or_status_kernel<<<unsigned((nRows + 255) / 256 < 1024 ? (nRows + 255) / 256 : 1024), 256, 0, s>>>(status, extra, rowSlice,
                                                                                                   nRows);
  return cudaGetLastError();
}

cudaError_t launchDecodeSlices(cudaStream_t s, int device, const Sched &sc, const uint8_t *in, const uint64_t *inOffsets,
                               const cabac_substream_params *params, const cabac_slice_params *sliceParams, size_t nSub,
                               const cabac_ctx_model *ctxIn, const uint64_t *cuOffsets, const uint64_t *tuOffsets,
                               const uint64_t *ctuOffsets, const uint64_t *coeffOffsets, cabac_cu_desc *cus,
                               cabac_tu_desc *tus, uint8_t *tuFlags, cabac_ctu_desc *ctus, int32_t *coeffs,
                               uint32_t *counts, uint32_t *numBins, uint32_t *bitsRead, cabac_ctx_model *ctxOut,
                               uint32_t *status) {
  if (nSub == 0) return cudaSuccess;
  cost_bytes_kernel<<<costGrid(nSub, 256), 256, 0, s>>>(inOffsets, nSub, sc.cost);
  cudaError_t e = launchOrder(s, sc, nSub);
  if (e != cudaSuccess) return e;
  int sms = 148, perSm = 1;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  cudaFuncSetAttribute(decode_slice_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, decode_slice_kernel, SLICE_WARPS_PER_BLOCK * 32, 0) !=
          cudaSuccess ||
      perSm < 1)
    perSm = 1;
  // few substreams: one warp per CTA spreads them over the SMs instead of pairing them
  const size_t blocksWanted = nSub <= size_t(sms) ? nSub : (nSub + SLICE_WARPS_PER_BLOCK - 1) / SLICE_WARPS_PER_BLOCK;
  const size_t cap = size_t(sms) * perSm;
  const int blocks = int(blocksWanted < 1 ? 1 : (blocksWanted > cap ? cap : blocksWanted));
  decode_slice_kernel<<<blocks, SLICE_WARPS_PER_BLOCK * 32, 0, s>>>(
      in, inOffsets, params, sliceParams, nSub, ctxIn, cuOffsets, tuOffsets, ctuOffsets, coeffOffsets, cus, tus, tuFlags,
      ctus, coeffs, counts, numBins, bitsRead, ctxOut, status, sc.order, sc.cursor);
  return cudaGetLastError();
}

}  // namespace cabac

#ifdef CABAC_SLICE_TRACE
extern "C" int cabac_debug_slice_trace(int which, uint32_t *out, size_t cap, uint32_t *n) {
  cudaDeviceSynchronize();
  uint32_t cnt[2];
  if (cudaMemcpyFromSymbol(cnt, cabac::g_sliceTraceN, 8) != cudaSuccess) return -1;
  *n = cnt[which & 1];
  const size_t m = *n < cap ? *n : cap;
  return cudaMemcpyFromSymbol(out, cabac::g_sliceTrace, m * 4, size_t(which & 1) * (size_t(1) << 17) * 4) == cudaSuccess ? 0 : -1;
}
#endif
