// oracle/syntax_host.cpp - TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Runs the product's CTU-level parser (entropy_coding_b200/csrc/cabac_syntax.cuh, control flow only)
// on the CPU, on top of the oracle's scalar arithmetic decoder / residual decoder
// (cabac_oracle.c:orc_dec_*), and records every call it makes in the command-word format of
// include/cabac_b200.h.  tests/test_syntax.py compares that trace with the BinEncIf call sequence the
// reference's own CABACWriter recorded for the same substream (tests/golden/vtm_*.npz: cmds with
// CABAC_CMD_TU markers), so the walk is pinned to the reference without a GPU; the `-m gpu` tests then
// compare the device kernel's records with the fixtures and with this host run.
#include <cstdint>
#include <cstring>
#include <vector>

#include "cabac_syntax.cuh"

extern "C" {
#include "cabac_oracle.h"
}

namespace {
struct HostEngine {
  static constexpr bool kEncode = false;
  const cabac_cu_desc &getCu(uint32_t) const { return *cusOut; }      // writer direction only
  const cabac_ctu_desc &getCtu(uint32_t) const { return *ctusOut; }
  void *dec = nullptr;
  std::vector<uint32_t> trace;
  std::vector<uint16_t> above[2], left[2];
  std::vector<uint8_t> ctus, mAbove, mLeft, lmap;
  std::vector<uint16_t> eLeft;
  cabac_cu_desc *cusOut;
  cabac_tu_desc *tusOut;
  uint8_t *tuFlagsOut;
  cabac_ctu_desc *ctusOut;
  int32_t *coeffs;
  bool dead = false;

  uint32_t bin(int ctx, uint32_t = 0) {
    const uint32_t b = orc_dec_bin(dec, unsigned(ctx));
    trace.push_back(CABAC_MK_CTX(ctx, b));
    return b;
  }
  uint32_t ep(int n, uint32_t = 0) {
    const uint32_t v = orc_dec_ep(dec, unsigned(n));
    trace.push_back(CABAC_MK_EP(v, n));
    return v;
  }
  uint32_t trm(uint32_t = 1) {
    const uint32_t b = orc_dec_trm(dec);
    trace.push_back(CABAC_MK_TRM(b));
    return b;
  }
  uint32_t residual(uint32_t idx, const cabac_tu_desc &d) {
    uint8_t fl = 0;
    orc_dec_tu(dec, &d, coeffs, &fl);
    trace.push_back(CABAC_MK_TU());
    cabac_tu_desc o = d;
    if (fl & CABAC_TUR_TS) o.flags |= CABAC_TU_TS;
    tusOut[idx] = o;
    tuFlagsOut[idx] = fl;
    return fl;
  }
  uint16_t *aboveLine(int ch) { return above[ch].data(); }
  uint16_t *leftCol(int ch) { return left[ch].data(); }
  uint8_t *ctuLine() { return ctus.data(); }
  uint8_t *modeAbove() { return mAbove.data(); }
  uint8_t *modeLeft() { return mLeft.data(); }
  uint16_t *edgeLeft() { return eLeft.data(); }
  uint8_t *lumaMap() { return lmap.data(); }
  void fill8(uint8_t *p, int n, uint8_t v) {
    for (int i = 0; i < n; i++) p[i] = v;
  }
  void fillMap(int bx, int by, int bw, int bh, uint8_t v) {
    for (int y = 0; y < bh; y++)
      for (int x = 0; x < bw; x++) lmap[size_t(((by + y) << 4) + bx + x)] = v;
  }
  void ctuLineStore(int col, uint8_t v) { ctus[size_t(col)] = v; }
  void fill(uint16_t *p, int n, uint16_t v) {
    for (int i = 0; i < n; i++) p[i] = v;
  }
  void putCu(uint32_t i, const cabac_cu_desc &c) { cusOut[i] = c; }
  void putCtu(uint32_t i, const cabac_ctu_desc &c) { ctusOut[i] = c; }
  // WPP rows (SliceParser::run)
  const uint8_t *sliceIn = nullptr;
  long sliceBytes = 0, rowStart = 0;
  cabac_ctx_model rowCtx[CABAC_NUM_CTX];
  bool rows() const { return false; }
  void saveRowContexts() { orc_dec_ctx_get(dec, rowCtx); }
  void nextRow() {
    const long used = rowStart + (orc_dec_bits(dec) >> 3) + 1;
    orc_dec_ctx_set(dec, rowCtx);
    if (used >= sliceBytes) {
      dead = true;
      return;
    }
    orc_dec_restart(dec, sliceIn + used, sliceBytes - used);
    rowStart = used;
  }
  void abort() { dead = true; }
  bool aborted() const { return dead; }
};
}  // namespace

extern "C" {

void syn_ctx_offsets(uint32_t *o) { cabac_syntax::syntaxCtxOffsets(o); }

// One substream.  Capacities are element counts; counts[0..2] = CUs, TUs, coefficients; trace holds up to
// traceCap words and *nTrace the number produced.  Returns the bits read.
long syn_decode_slice(const uint8_t *in, long nbytes, const cabac_substream_params *p, const cabac_slice_params *sp,
                      const cabac_ctx_model *ctxIn, cabac_cu_desc *cus, long capCu, cabac_tu_desc *tus,
                      uint8_t *tuFlags, long capTu, cabac_ctu_desc *ctus, long capCtu, int32_t *coeffs, long capCoeff,
                      uint32_t *counts, uint32_t *numBins, cabac_ctx_model *ctxOut, uint32_t *status, uint32_t *trace,
                      long traceCap, long *nTrace) {
  HostEngine e;
  e.dec = orc_dec_open(in, nbytes, p, ctxIn);
  e.sliceIn = in;
  e.sliceBytes = nbytes;
  for (int ch = 0; ch < 2; ch++) {
    e.above[ch].assign(cabac_syntax::SX_MAX_LINE, 0);
    e.left[ch].assign(32, 0);
  }
  e.ctus.assign(512, 0);
  e.mAbove.assign(cabac_syntax::SX_MAX_LINE, 0);
  e.mLeft.assign(32, 0);
  e.eLeft.assign(32, 0);
  e.lmap.assign(256, 0);
  e.cusOut = cus;
  e.tusOut = tus;
  e.tuFlagsOut = tuFlags;
  e.ctusOut = ctus;
  e.coeffs = coeffs;
  cabac_syntax::SliceParser<HostEngine> ps(e, *sp, uint32_t(capCu), uint32_t(capTu), uint32_t(capCtu),
                                           uint64_t(capCoeff), 0);
  ps.run((p->flags & CABAC_SUB_TERMINATE) != 0);
  counts[0] = ps.nCu;
  counts[1] = ps.nTu;
  counts[2] = uint32_t(ps.nCoeff);
  uint32_t st = 0;
  const long bits = 8 * e.rowStart + orc_dec_close(e.dec, ctxOut, &st, numBins);
  *status = st | ps.status;
  *nTrace = long(e.trace.size());
  for (long i = 0; i < *nTrace && i < traceCap; i++) trace[i] = e.trace[size_t(i)];
  return bits;
}
}

// ---- generator: the same parser on an engine that CHOOSES every bin and codes it -------------------------------
// A legal CTU-level stream for any tool configuration, with the records the walk produced as ground truth.  Not a
// conformance check (parser and generator share the walk; conformance is pinned by the VTM fixtures): it exercises
// what no fixture holds - large pictures, every CTU size, BDPCM, cu_qp_delta, chroma QP offsets, ALF chroma / CC-ALF
// CTU syntax, 32-sample maximum transform size, capacity limits - and gives the device a dense, realistic-size input.
namespace {
struct GenEngine {
  static constexpr bool kEncode = false;  // it chooses values itself: the walk records them like a decoder
  const cabac_cu_desc &getCu(uint32_t) const { return *cusOut; }
  const cabac_ctu_desc &getCtu(uint32_t) const { return *ctusOut; }
  void *enc = nullptr;
  uint64_t rng;
  double splitP, density;
  std::vector<uint16_t> above[2], left[2], eLeft;
  std::vector<uint8_t> ctus, mAbove, mLeft, lmap;
  cabac_cu_desc *cusOut;
  cabac_tu_desc *tusOut;
  uint8_t *tuFlagsOut;
  cabac_ctu_desc *ctusOut;
  int32_t *coeffs;
  cabac_ctx_model rowCtx[CABAC_NUM_CTX];
  bool dead = false;
  uint32_t subFlags = 0;

  uint64_t next() {
    rng ^= rng << 13;
    rng ^= rng >> 7;
    rng ^= rng << 17;
    return rng;
  }
  double uni() { return double(next() >> 11) * (1.0 / 9007199254740992.0); }
  uint32_t bin(int ctx, uint32_t = 0) {
    using namespace cabac_syntax;
    double p1 = 0.5;
    if (ctx >= SX_SPLIT_FLAG && ctx < SX_SPLIT_QT_FLAG) p1 = splitP;
    else if (ctx >= SX_QT_CBF_Y && ctx < SX_QT_CBF_CR + 3) p1 = 0.7;
    else if (ctx >= SX_MIP_FLAG && ctx < SX_MIP_FLAG + 4) p1 = 0.2;
    else if (ctx >= SX_BDPCM_MODE && ctx < SX_BDPCM_MODE + 4) p1 = 0.25;
    else if (ctx >= SX_ISP_MODE && ctx < SX_ISP_MODE + 2) p1 = 0.3;
    else if (ctx >= SX_MULTI_REF_LINE_IDX && ctx < SX_MULTI_REF_LINE_IDX + 2) p1 = 0.25;
    else if (ctx == SX_INTRA_LUMA_MPM_FLAG) p1 = 0.7;
    else if (ctx >= SX_LFNST_IDX && ctx < SX_LFNST_IDX + 3) p1 = 0.4;
    const uint32_t b = uni() < p1 ? 1u : 0u;
    orc_enc_bin(enc, b, unsigned(ctx));
    return b;
  }
  uint32_t ep(int n, uint32_t = 0) {
    uint32_t v = uint32_t(next() >> 20);
    if (n < 32) v &= (1u << n) - 1u;
    int left_ = n;  // encodeBinsEP takes at most 32 bins; feed 16 at a time, MSB first
    while (left_ > 16) {
      orc_enc_ep(enc, (v >> (left_ - 16)) & 0xFFFFu, 16);
      left_ -= 16;
    }
    orc_enc_ep(enc, v & ((1u << left_) - 1u), unsigned(left_));
    return v;
  }
  uint32_t trm(uint32_t = 1) {
    orc_enc_trm(enc, 1);
    return 1;
  }
  uint32_t residual(uint32_t idx, const cabac_tu_desc &d) {
    cabac_tu_desc o = d;
    const int w = 1 << d.log2_w, h = 1 << d.log2_h;
    const bool bdpcm = (d.flags & CABAC_TU_BDPCM) != 0;
    bool ts = bdpcm;
    if (d.flags & CABAC_TU_TS_ALLOWED) ts = uni() < 0.25;
    if (ts) o.flags |= CABAC_TU_TS;
    const int tw = ts ? w : (w > 32 ? 32 : w), th = ts ? h : (h > 32 ? 32 : h);
    int32_t *blk = coeffs + d.coeff_offset;
    memset(blk, 0, sizeof(int32_t) * size_t(w) * size_t(h));
    int nz = 0;
    for (int y = 0; y < th; y++)
      for (int x = 0; x < tw; x++) {
        // energy compaction: fewer and smaller levels away from the top-left corner (not for transform skip)
        const double fall = ts ? 1.0 : 1.0 / (1.0 + 0.35 * double(x + y));
        if (uni() < density * fall) {
          int mag = 1;
          while (mag < 400 && uni() < 0.55 * fall + 0.1) mag += 1 + int(next() & 3);
          blk[y * w + x] = (next() & 1) ? mag : -mag;
          nz++;
        }
      }
    if (!nz) blk[0] = (next() & 1) ? 1 : -1;
    uint8_t fl = 0;
    const unsigned st = orc_enc_tu(enc, &o, coeffs, &fl);
    if (st) dead = true;
    tusOut[idx] = o;
    tuFlagsOut[idx] = fl;
    return fl;
  }
  uint16_t *aboveLine(int ch) { return above[ch].data(); }
  uint16_t *leftCol(int ch) { return left[ch].data(); }
  uint8_t *ctuLine() { return ctus.data(); }
  void ctuLineStore(int col, uint8_t v) { ctus[size_t(col)] = v; }
  uint8_t *modeAbove() { return mAbove.data(); }
  uint8_t *modeLeft() { return mLeft.data(); }
  uint16_t *edgeLeft() { return eLeft.data(); }
  uint8_t *lumaMap() { return lmap.data(); }
  void fill(uint16_t *p, int n, uint16_t v) {
    for (int i = 0; i < n; i++) p[i] = v;
  }
  void fill8(uint8_t *p, int n, uint8_t v) {
    for (int i = 0; i < n; i++) p[i] = v;
  }
  void fillMap(int bx, int by, int bw, int bh, uint8_t v) {
    for (int y = 0; y < bh; y++)
      for (int x = 0; x < bw; x++) lmap[size_t(((by + y) << 4) + bx + x)] = v;
  }
  void putCu(uint32_t i, const cabac_cu_desc &c) { cusOut[i] = c; }
  void putCtu(uint32_t i, const cabac_ctu_desc &c) { ctusOut[i] = c; }
  bool rows() const { return false; }
  void saveRowContexts() { orc_enc_ctx_get(enc, rowCtx); }
  void nextRow() {  // the row's terminating bin is coded: finish + alignment, fresh coder, contexts of the row above
    orc_enc_flush(enc, 1);
    orc_enc_ctx_set(enc, rowCtx);
  }
  void abort() { dead = true; }
  bool aborted() const { return dead; }
};
}  // namespace

extern "C" long syn_generate_slice(uint64_t seed, double splitP, double density, const cabac_substream_params *p,
                                   const cabac_slice_params *sp, uint8_t *out, long outCap, cabac_cu_desc *cus, long capCu,
                                   cabac_tu_desc *tus, uint8_t *tuFlags, long capTu, cabac_ctu_desc *ctus, long capCtu,
                                   int32_t *coeffs, long capCoeff, uint32_t *counts, uint32_t *numBins,
                                   cabac_ctx_model *ctxOut, uint32_t *status) {
  GenEngine e;
  e.rng = seed * 0x9E3779B97F4A7C15ull + 0x1234567ull;
  e.splitP = splitP;
  e.density = density;
  e.enc = orc_enc_open(out, outCap, p, nullptr);
  e.subFlags = p->flags;
  for (int ch = 0; ch < 2; ch++) {
    e.above[ch].assign(cabac_syntax::SX_MAX_LINE, 0);
    e.left[ch].assign(32, 0);
  }
  e.ctus.assign(512, 0);
  e.mAbove.assign(cabac_syntax::SX_MAX_LINE, 0);
  e.mLeft.assign(32, 0);
  e.eLeft.assign(32, 0);
  e.lmap.assign(256, 0);
  e.cusOut = cus;
  e.tusOut = tus;
  e.tuFlagsOut = tuFlags;
  e.ctusOut = ctus;
  e.coeffs = coeffs;
  cabac_syntax::SliceParser<GenEngine> ps(e, *sp, uint32_t(capCu), uint32_t(capTu), uint32_t(capCtu), uint64_t(capCoeff), 0);
  ps.run(true);
  counts[0] = ps.nCu;
  counts[1] = ps.nTu;
  counts[2] = uint32_t(ps.nCoeff);
  *status = ps.status | (e.dead ? uint32_t(CABAC_SS_CAPACITY) : 0u);
  orc_enc_flush(e.enc, 0);
  return orc_enc_close(e.enc, ctxOut, numBins);
}

// ---- writer direction on the host: the walk on an engine that turns records into the call sequence -------------
namespace {
struct EmitEngine {
  static constexpr bool kEncode = true;
  const cabac_cu_desc *cusIn;
  const cabac_ctu_desc *ctusIn;
  const cabac_tu_desc *tusIn;
  const uint8_t *tuFlagsIn;
  uint32_t nCuIn, nTuIn, nCtuIn;
  std::vector<uint32_t> cmds;
  std::vector<uint16_t> above[2], left[2], eLeft;
  std::vector<uint8_t> ctus, mAbove, mLeft, lmap;
  bool dead = false, bad = false;
  cabac_cu_desc dummyCu{};
  cabac_ctu_desc dummyCtu{};

  const cabac_cu_desc &getCu(uint32_t i) {
    if (i >= nCuIn) {
      bad = dead = true;
      return dummyCu;
    }
    return cusIn[i];
  }
  const cabac_ctu_desc &getCtu(uint32_t i) {
    if (i >= nCtuIn) {
      bad = dead = true;
      return dummyCtu;
    }
    return ctusIn[i];
  }
  uint32_t bin(int ctx, uint32_t want) {
    cmds.push_back(CABAC_MK_CTX(ctx, want & 1u));
    return want & 1u;
  }
  uint32_t ep(int n, uint32_t want) {
    const uint32_t v = n < 32 ? want & ((1u << n) - 1u) : want;
    int left_ = n;
    while (left_ > 16) {
      cmds.push_back(CABAC_MK_EP((v >> (left_ - 16)) & 0xFFFFu, 16));
      left_ -= 16;
    }
    cmds.push_back(CABAC_MK_EP(v & ((1u << left_) - 1u), left_));
    return v;
  }
  uint32_t trm(uint32_t want) {
    cmds.push_back(CABAC_MK_TRM(want & 1u));
    return want & 1u;
  }
  uint32_t residual(uint32_t idx, const cabac_tu_desc &d) {
    if (idx >= nTuIn || tusIn[idx].log2_w != d.log2_w || tusIn[idx].log2_h != d.log2_h || tusIn[idx].comp_id != d.comp_id ||
        ((tusIn[idx].flags ^ d.flags) & uint8_t(~CABAC_TU_TS)) ||
        ((tusIn[idx].flags & CABAC_TU_TS) && !(d.flags & (CABAC_TU_TS_ALLOWED | CABAC_TU_BDPCM)))) {
      bad = dead = true;
      return 0;
    }
    cmds.push_back(CABAC_MK_TU());
    return tuFlagsIn[idx];
  }
  uint16_t *aboveLine(int ch) { return above[ch].data(); }
  uint16_t *leftCol(int ch) { return left[ch].data(); }
  uint8_t *ctuLine() { return ctus.data(); }
  void ctuLineStore(int col, uint8_t v) { ctus[size_t(col)] = v; }
  uint8_t *modeAbove() { return mAbove.data(); }
  uint8_t *modeLeft() { return mLeft.data(); }
  uint16_t *edgeLeft() { return eLeft.data(); }
  uint8_t *lumaMap() { return lmap.data(); }
  void fill(uint16_t *p, int n, uint16_t v) {
    for (int i = 0; i < n; i++) p[i] = v;
  }
  void fill8(uint8_t *p, int n, uint8_t v) {
    for (int i = 0; i < n; i++) p[i] = v;
  }
  void fillMap(int bx, int by, int bw, int bh, uint8_t v) {
    for (int y = 0; y < bh; y++)
      for (int x = 0; x < bw; x++) lmap[size_t(((by + y) << 4) + bx + x)] = v;
  }
  void putCu(uint32_t, const cabac_cu_desc &) {}
  void putCtu(uint32_t, const cabac_ctu_desc &) {}
  bool rows() const { return true; }  // the row ends stay in the sequence as terminating bins, as the writer made them
  void saveRowContexts() {}
  void nextRow() {}
  void abort() { dead = true; }
  bool aborted() const { return dead; }
};
}  // namespace

// Records of one slice (as cabac_decode_slices returns them) -> the BinEncIf call sequence of its writer, as
// command words with CABAC_CMD_TU markers (the input of cabac_encode_stream).  Returns the number of words.
extern "C" long syn_emit_slice(const cabac_slice_params *sp, const cabac_cu_desc *cus, long nCu, const cabac_tu_desc *tus,
                               const uint8_t *tuFlags, long nTu, const cabac_ctu_desc *ctus, long nCtu, uint32_t *cmds,
                               long cap, uint32_t *status) {
  EmitEngine e;
  e.cusIn = cus;
  e.ctusIn = ctus;
  e.tusIn = tus;
  e.tuFlagsIn = tuFlags;
  e.nCuIn = uint32_t(nCu);
  e.nTuIn = uint32_t(nTu);
  e.nCtuIn = uint32_t(nCtu);
  for (int ch = 0; ch < 2; ch++) {
    e.above[ch].assign(cabac_syntax::SX_MAX_LINE, 0);
    e.left[ch].assign(32, 0);
  }
  e.ctus.assign(512, 0);
  e.mAbove.assign(cabac_syntax::SX_MAX_LINE, 0);
  e.mLeft.assign(32, 0);
  e.eLeft.assign(32, 0);
  e.lmap.assign(256, 0);
  cabac_syntax::SliceParser<EmitEngine> ps(e, *sp, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, ~uint64_t(0), 0);
  ps.run(true);
  *status = ps.status | (e.bad ? uint32_t(CABAC_SS_BAD_DESC) : 0u);
  if (ps.nCu != e.nCuIn || ps.nTu != e.nTuIn) *status |= CABAC_SS_BAD_DESC;  // records left over
  for (long i = 0; i < long(e.cmds.size()) && i < cap; i++) cmds[i] = e.cmds[size_t(i)];
  return long(e.cmds.size());
}
