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

  uint32_t bin(int ctx) {
    const uint32_t b = orc_dec_bin(dec, unsigned(ctx));
    trace.push_back(CABAC_MK_CTX(ctx, b));
    return b;
  }
  uint32_t ep(int n) {
    const uint32_t v = orc_dec_ep(dec, unsigned(n));
    trace.push_back(CABAC_MK_EP(v, n));
    return v;
  }
  uint32_t trm() {
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
