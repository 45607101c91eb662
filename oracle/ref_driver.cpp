// oracle/ref_driver.cpp - TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// C-ABI driver around the UNMODIFIED reference sources (/root/reference/src),
// compiled where they lie by oracle/Makefile into oracle/_ref/libcabac_ref.so.
// It (i) dumps the context-init tables, (ii) runs recorded bin-level command
// streams through BinEncoder_Std / BinDecoder_Std, (iii) drives
// CABACWriter::residual_coding / CABACReader::residual_coding from the flat TU
// descriptors of include/cabac_b200.h, and (iv) times (ii)/(iii) on host
// threads for bench.py's cpu_baseline / --impl reference arm.
//
// Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may load it.
#include <atomic>
#include <chrono>
#include <cstring>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include "arith_codec.hpp"
#include "cabac_reader.hpp"
#include "cabac_writer.hpp"
#include "coding_structure.hpp"
#include "context_modelling.hpp"
#include "rom.hpp"
#include "slice.hpp"
#include "unit.hpp"

#include "cabac_b200.h"
#include "ref_harness.hpp"

using namespace Common;
using namespace EntropyCoding;

using namespace ref_harness;

namespace {

uint8_t tuResultFlags(const CUCtx &cuCtx, const TransformUnit &tu, int comp) {
  uint8_t f = 0;
  if (tu.mtsIdx[comp] == MTS_SKIP) f |= CABAC_TUR_TS;
  if (cuCtx.violatesLfnstConstrained[toChannelType(ComponentID(comp))])
    f |= CABAC_TUR_VIOLATES_LFNST;
  if (cuCtx.lfnstLastScanPos) f |= CABAC_TUR_LFNST_LAST_POS;
  if (cuCtx.mtsLastScanPos) f |= CABAC_TUR_MTS_LAST_POS;
  if (cuCtx.violatesMtsCoeffConstraint) f |= CABAC_TUR_VIOLATES_MTS;
  return f;
}

void importCtx(Ctx &ctx, const cabac_ctx_model *in) {
  CtxStore<BinProbModel_Std> &st =
      static_cast<CtxStore<BinProbModel_Std> &>(ctx);
  for (unsigned k = 0; k < Ctx::NumberOfContexts; k++) {
    const uint8_t rate = st[k].getRate();
    st[k] = BinProbModel_Std(in[k].state0, in[k].state1, rate);
  }
}

void exportCtx(const Ctx &ctx, cabac_ctx_model *out) {
  const CtxStore<BinProbModel_Std> &st =
      static_cast<const CtxStore<BinProbModel_Std> &>(ctx);
  for (unsigned k = 0; k < Ctx::NumberOfContexts; k++) {
    out[k].state0 = st[k].getState0();
    out[k].state1 = st[k].getState1();
  }
}

long copyOut(OutputBitstream &bs, uint8_t *out, long cap) {
  const std::vector<uint8_t> &f = bs.getFIFO();
  long n = (long)f.size();
  if (out) memcpy(out, f.data(), (size_t)std::min<long>(n, cap));
  return n;
}

}  // namespace

extern "C" {

int ref_num_contexts() { return (int)Ctx::NumberOfContexts; }

// Raw init-table rows (0=B,1=P,2=I,3=window sizes).
void ref_init_table(int row, uint8_t *out) {
  const std::vector<uint8_t> &t = ContextSetCfg::getInitTable((unsigned)row);
  memcpy(out, t.data(), t.size());
}

// Ctx::init(qp, initId): states and m_rate per context.
void ref_ctx_init(int qp, int initId, cabac_ctx_model *out, uint8_t *rates) {
  BinEncoder_Std enc;
  enc.Ctx::init(qp, initId);
  exportCtx(enc, out);
  if (rates) {
    const CtxStore<BinProbModel_Std> &st =
        static_cast<const CtxStore<BinProbModel_Std> &>(
            static_cast<const Ctx &>(enc));
    for (unsigned k = 0; k < Ctx::NumberOfContexts; k++)
      rates[k] = st[k].getRate();
  }
}

// Scan tables as CoeffCodingContext sees them (g_scanOrder[GROUPED_4x4][DIAG]).
// out: w*h entries of blkPos.  Returns log2CGw | log2CGh << 8.
int ref_scan_table(int log2w, int log2h, uint32_t *out) {
  ensureRom();
  const unsigned w = 1u << log2w, h = 1u << log2h;
  const ScanElement *scan =
      g_scanOrder[SCAN_GROUPED_4x4][SCAN_DIAG][gp_sizeIdxInfo->idxFrom(w)]
                 [gp_sizeIdxInfo->idxFrom(h)];
  for (unsigned i = 0; i < w * h; i++) out[i] = scan[i].idx;
  return (int)(g_log2SbbSize[log2w][log2h][0] |
               (g_log2SbbSize[log2w][log2h][1] << 8));
}

// Context-set offsets in the order oracle/cabac_oracle.c:orc_ctx_offsets lists them.
void ref_ctx_offsets(uint32_t *o) {
  int k = 0;
  for (int i = 0; i < 2; i++) o[k++] = Ctx::SigCoeffGroup[i]();
  for (int i = 0; i < 6; i++) o[k++] = Ctx::SigFlag[i]();
  for (int i = 0; i < 2; i++) o[k++] = Ctx::ParFlag[i]();
  for (int i = 0; i < 4; i++) o[k++] = Ctx::GtxFlag[i]();
  for (int i = 0; i < 2; i++) o[k++] = Ctx::LastX[i]();
  for (int i = 0; i < 2; i++) o[k++] = Ctx::LastY[i]();
  o[k++] = Ctx::TransformSkipFlag();
  o[k++] = Ctx::TsSigCoeffGroup();
  o[k++] = Ctx::TsSigFlag();
  o[k++] = Ctx::TsParFlag();
  o[k++] = Ctx::TsGtxFlag();
  o[k++] = Ctx::TsLrg1Flag();
  o[k++] = Ctx::TsResidualSign();
}


// Offsets of the context sets the CTU-level syntax uses (cabac_reader.cpp:103-2645), in the order
// entropy_coding_b200/csrc/cabac_syntax.cuh lists them (SX_* constants); tests/test_syntax.py compares.
int ref_syntax_ctx_offsets(uint32_t *o) {
  int k = 0;
  o[k++] = Ctx::SplitFlag();
  o[k++] = Ctx::SplitQtFlag();
  o[k++] = Ctx::SplitHvFlag();
  o[k++] = Ctx::Split12Flag();
  o[k++] = Ctx::ModeConsFlag();
  o[k++] = Ctx::MipFlag();
  o[k++] = Ctx::IntraLumaMpmFlag();
  o[k++] = Ctx::IntraLumaPlanarFlag();
  o[k++] = Ctx::CclmModeFlag();
  o[k++] = Ctx::CclmModeIdx();
  o[k++] = Ctx::IntraChromaPredMode();
  o[k++] = Ctx::MultiRefLineIdx();
  o[k++] = Ctx::ISPMode();
  o[k++] = Ctx::BDPCMMode();
  o[k++] = Ctx::QtCbf[0]();
  o[k++] = Ctx::QtCbf[1]();
  o[k++] = Ctx::QtCbf[2]();
  o[k++] = Ctx::DeltaQP();
  o[k++] = Ctx::ChromaQpAdjFlag();
  o[k++] = Ctx::ChromaQpAdjIdc();
  o[k++] = Ctx::JointCbCrFlag();
  o[k++] = Ctx::MTSIdx();
  o[k++] = Ctx::LFNSTIdx();
  o[k++] = Ctx::SaoMergeFlag();
  o[k++] = Ctx::SaoTypeIdx();
  o[k++] = Ctx::ctbAlfFlag();
  o[k++] = Ctx::ctbAlfAlternative();
  o[k++] = Ctx::AlfUseTemporalFilt();
  o[k++] = Ctx::CcAlfFilterControlFlag();
  return k;
}

// Runs one recorded command stream through BinEncoder_Std.  Returns the byte
// count (may exceed cap; bytes are truncated to cap).
long ref_encode_bins(const uint32_t *cmds, long ncmd,
                     const cabac_substream_params *p,
                     const cabac_ctx_model *ctxIn, uint8_t *out, long cap,
                     uint32_t *numBins, cabac_ctx_model *ctxOut) {
  BinEncoder_Std enc;
  OutputBitstream bs;
  enc.init(&bs);
  enc.reset(p->qp, p->init_id);
  if ((p->flags & CABAC_SUB_CTX_IN) && ctxIn) importCtx(enc, ctxIn);
  for (long i = 0; i < ncmd; i++) {
    const uint32_t w = cmds[i];
    switch (CABAC_CMD_KIND(w)) {
      case CABAC_CMD_CTX: enc.encodeBin((w >> 9) & 1, w & 511); break;
      case CABAC_CMD_EP: enc.encodeBinsEP(w & 0xFFFFFF, (w >> 24) & 31); break;
      case CABAC_CMD_REM:
        enc.encodeRemAbsEP(w & 0xFFFFFF, (w >> 24) & 31, 5, 15);
        break;
      case CABAC_CMD_TRM: enc.encodeBinTrm(w & 1); break;
      case CABAC_CMD_ALIGN: enc.align(); break;
      default: return -1;
    }
  }
  if (p->flags & CABAC_SUB_TERMINATE) {
    enc.encodeBinTrm(1);
    enc.finish();
  }
  if (p->flags & CABAC_SUB_ALIGN) bs.writeByteAlignment();
  if (numBins) *numBins = enc.getNumBins();
  if (ctxOut) exportCtx(enc, ctxOut);
  return copyOut(bs, out, cap);
}

// Decodes with a known call skeleton.  Returns getNumBitsRead().
long ref_decode_bins(const uint8_t *in, long nbytes, const uint32_t *cmds,
                     long ncmd, const cabac_substream_params *p,
                     const cabac_ctx_model *ctxIn, uint32_t *values,
                     cabac_ctx_model *ctxOut, uint32_t *status) {
  BinDecoder_Std dec;
  InputBitstream bs;
  bs.getFifo().assign(in, in + nbytes);
  // guard bytes: the reference CHECKs on overrun; the flat API reports it
  bs.getFifo().resize((size_t)nbytes + 8, 0);
  dec.init(&bs);
  dec.reset(p->qp, p->init_id);
  if ((p->flags & CABAC_SUB_CTX_IN) && ctxIn) importCtx(dec, ctxIn);
  uint32_t st = 0;
  try {
    for (long i = 0; i < ncmd; i++) {
      const uint32_t w = cmds[i];
      uint32_t v = 0;
      switch (CABAC_CMD_KIND(w)) {
        case CABAC_CMD_CTX: v = dec.decodeBin(w & 511); break;
        case CABAC_CMD_EP: v = dec.decodeBinsEP((w >> 24) & 31); break;
        case CABAC_CMD_REM: v = dec.decodeRemAbsEP((w >> 24) & 31, 5, 15); break;
        case CABAC_CMD_TRM: v = dec.decodeBinTrm(); break;
        case CABAC_CMD_ALIGN: dec.align(); break;
        default: st |= CABAC_SS_BAD_CMD; break;
      }
      if (values) values[i] = v;
    }
    if (p->flags & CABAC_SUB_TERMINATE) {
      if (!dec.decodeBinTrm()) st |= CABAC_SS_NO_TERMINATE;
    }
  } catch (const Exception &) {
    st |= CABAC_SS_IN_OVERRUN;
  }
  if (ctxOut) exportCtx(dec, ctxOut);
  if (status) *status = st;
  // InputBitstream::m_numBitsRead only advances under ENABLE_TRACING
  // (bit_stream.cpp:268-274), so rebuild the traced value from the byte index.
  return 8L * (long)bs.getByteLocation() + (long)(int32_t)dec.getNumBitsRead();
}

// reset + residual_coding over a TU list (+ terminate / align).
long ref_encode_residual(const cabac_tu_desc *tus, long ntu,
                         const cabac_substream_params *p, const int32_t *coeffs,
                         const cabac_ctx_model *ctxIn, uint8_t *out, long cap,
                         uint32_t *numBins, uint8_t *tuFlags,
                         cabac_ctx_model *ctxOut, int useMTS) {
  Harness hs(*p, useMTS != 0);
  BinEncoder_Std enc;
  CABACWriter wr(enc);
  OutputBitstream bs;
  wr.initBitstream(&bs);
  enc.reset(p->qp, p->init_id);
  enc.riceStatReset(10);
  if ((p->flags & CABAC_SUB_CTX_IN) && ctxIn) importCtx(enc, ctxIn);
  for (long t = 0; t < ntu; t++) {
    TuPair tp;
    makeTu(tp, hs, tus[t], const_cast<TCoeff *>(coeffs));
    applyTsAllowed(tp, tus[t], useMTS != 0);
    CUCtx cuCtx(p->qp);
    wr.residual_coding(tp.tu, ComponentID(tus[t].comp_id), &cuCtx);
    if (tuFlags) tuFlags[t] = tuResultFlags(cuCtx, tp.tu, tus[t].comp_id);
  }
  if (p->flags & CABAC_SUB_TERMINATE) wr.end_of_slice();
  if (p->flags & CABAC_SUB_ALIGN) bs.writeByteAlignment();
  if (numBins) *numBins = enc.getNumBins();
  if (ctxOut) exportCtx(enc, ctxOut);
  return copyOut(bs, out, cap);
}

// Recorded command stream with CABAC_CMD_TU markers: BinEncIf calls and
// CABACWriter::residual_coding calls interleaved on one coder, as the
// reference's slice writer issues them.
long ref_encode_stream(const uint32_t *cmds, long ncmd, const cabac_tu_desc *tus,
                       long ntu, const cabac_substream_params *p,
                       const int32_t *coeffs, const cabac_ctx_model *ctxIn,
                       uint8_t *out, long cap, uint32_t *numBins,
                       uint8_t *tuFlags, cabac_ctx_model *ctxOut) {
  Harness hs(*p, false);
  BinEncoder_Std enc;
  CABACWriter wr(enc);
  OutputBitstream bs;
  wr.initBitstream(&bs);
  enc.reset(p->qp, p->init_id);
  enc.riceStatReset(10);
  if ((p->flags & CABAC_SUB_CTX_IN) && ctxIn) importCtx(enc, ctxIn);
  long t = 0;
  for (long i = 0; i < ncmd; i++) {
    const uint32_t w = cmds[i];
    switch (CABAC_CMD_KIND(w)) {
      case CABAC_CMD_CTX: enc.encodeBin((w >> 9) & 1, w & 511); break;
      case CABAC_CMD_EP: enc.encodeBinsEP(w & 0xFFFFFF, (w >> 24) & 31); break;
      case CABAC_CMD_REM:
        enc.encodeRemAbsEP(w & 0xFFFFFF, (w >> 24) & 31, 5, 15);
        break;
      case CABAC_CMD_TRM: enc.encodeBinTrm(w & 1); break;
      case CABAC_CMD_ALIGN: enc.align(); break;
      case CABAC_CMD_TU: {
        if (t >= ntu) return -1;
        TuPair tp;
        makeTu(tp, hs, tus[t], const_cast<TCoeff *>(coeffs));
        applyTsAllowed(tp, tus[t], false);
        CUCtx cuCtx(p->qp);
        wr.residual_coding(tp.tu, ComponentID(tus[t].comp_id), &cuCtx);
        if (tuFlags) tuFlags[t] = tuResultFlags(cuCtx, tp.tu, tus[t].comp_id);
        t++;
        break;
      }
      default: return -1;
    }
  }
  if (p->flags & CABAC_SUB_TERMINATE) wr.end_of_slice();
  if (p->flags & CABAC_SUB_ALIGN) bs.writeByteAlignment();
  if (numBins) *numBins = enc.getNumBins();
  if (ctxOut) exportCtx(enc, ctxOut);
  return copyOut(bs, out, cap);
}

long ref_decode_residual(const uint8_t *in, long nbytes,
                         const cabac_tu_desc *tus, long ntu,
                         const cabac_substream_params *p,
                         const cabac_ctx_model *ctxIn, int32_t *coeffs,
                         uint8_t *tuFlags, cabac_ctx_model *ctxOut,
                         uint32_t *status, int useMTS) {
  Harness hs(*p, useMTS != 0);
  BinDecoder_Std dec;
  CABACReader rd(dec);
  InputBitstream bs;
  bs.getFifo().assign(in, in + nbytes);
  bs.getFifo().resize((size_t)nbytes + 8, 0);
  rd.initBitstream(&bs);
  dec.reset(p->qp, p->init_id);
  dec.riceStatReset(10);
  if ((p->flags & CABAC_SUB_CTX_IN) && ctxIn) importCtx(dec, ctxIn);
  uint32_t st = 0;
  try {
    for (long t = 0; t < ntu; t++) {
      TuPair tp;
      makeTu(tp, hs, tus[t], coeffs);
      applyTsAllowed(tp, tus[t], useMTS != 0);
      const unsigned n = 1u << (tus[t].log2_w + tus[t].log2_h);
      memset(coeffs + tus[t].coeff_offset, 0, n * sizeof(int32_t));
      CUCtx cuCtx(p->qp);
      rd.residual_coding(tp.tu, ComponentID(tus[t].comp_id), cuCtx);
      if (tuFlags) tuFlags[t] = tuResultFlags(cuCtx, tp.tu, tus[t].comp_id);
    }
    if (p->flags & CABAC_SUB_TERMINATE) {
      if (!dec.decodeBinTrm()) st |= CABAC_SS_NO_TERMINATE;
    }
  } catch (const Exception &) {
    st |= CABAC_SS_COEFF_RANGE;
  }
  if (ctxOut) exportCtx(dec, ctxOut);
  if (status) *status = st;
  // InputBitstream::m_numBitsRead only advances under ENABLE_TRACING
  // (bit_stream.cpp:268-274), so rebuild the traced value from the byte index.
  return 8L * (long)bs.getByteLocation() + (long)(int32_t)dec.getNumBitsRead();
}

// ---- timed batch runs for the CPU baseline --------------------------------
// Substreams are handed out dynamically (atomic counter) to nthreads host
// threads, one coder object per call.  Returns wall seconds; totals via the out pointers.
double ref_time_encode_residual(const cabac_tu_desc *tus,
                                const uint64_t *tuOffsets,
                                const cabac_substream_params *params,
                                long nSub, const int32_t *coeffs,
                                int nthreads, uint64_t *totalBins,
                                uint64_t *totalBytes, uint8_t *out,
                                const uint64_t *outOffsets,
                                uint32_t *outSizes) {
  ensureRom();
  std::atomic<uint64_t> bins{0}, bytes{0};
  std::atomic<long> next{0};
  auto t0 = std::chrono::steady_clock::now();
  std::vector<std::thread> th;
  for (int k = 0; k < nthreads; k++) {
    th.emplace_back([&, k] {
      uint64_t b = 0, by = 0;
      (void)k;
      for (long i = next++; i < nSub; i = next++) {
        uint32_t nb = 0;
        uint8_t *o = out ? out + outOffsets[i] : nullptr;
        long cap = out ? (long)(outOffsets[i + 1] - outOffsets[i]) : 0;
        long n = ref_encode_residual(tus + tuOffsets[i],
                                     (long)(tuOffsets[i + 1] - tuOffsets[i]),
                                     &params[i], coeffs, nullptr, o, cap, &nb,
                                     nullptr, nullptr, 0);
        if (outSizes) outSizes[i] = (uint32_t)n;
        b += nb;
        by += (uint64_t)n;
      }
      bins += b;
      bytes += by;
    });
  }
  for (auto &t : th) t.join();
  auto t1 = std::chrono::steady_clock::now();
  if (totalBins) *totalBins = bins.load();
  if (totalBytes) *totalBytes = bytes.load();
  return std::chrono::duration<double>(t1 - t0).count();
}

double ref_time_decode_residual(const uint8_t *in, const uint64_t *inOffsets,
                                const uint32_t *inSizes,
                                const cabac_tu_desc *tus,
                                const uint64_t *tuOffsets,
                                const cabac_substream_params *params,
                                long nSub, int32_t *coeffs, int nthreads) {
  ensureRom();
  std::atomic<long> next{0};
  auto t0 = std::chrono::steady_clock::now();
  std::vector<std::thread> th;
  for (int k = 0; k < nthreads; k++) {
    th.emplace_back([&, k] {
      (void)k;
      for (long i = next++; i < nSub; i = next++) {
        ref_decode_residual(in + inOffsets[i], (long)inSizes[i],
                            tus + tuOffsets[i],
                            (long)(tuOffsets[i + 1] - tuOffsets[i]), &params[i],
                            nullptr, coeffs, nullptr, nullptr, nullptr, 0);
      }
    });
  }
  for (auto &t : th) t.join();
  auto t1 = std::chrono::steady_clock::now();
  return std::chrono::duration<double>(t1 - t0).count();
}

// The reference's own OutputBitstream::countStartCodeEmulations (bit_stream.cpp:157-181).
long ref_count_emulations(const uint8_t *d, long n) {
  OutputBitstream bs;
  bs.getFIFO().assign(d, d + n);
  return bs.countStartCodeEmulations();
}

// ProbModelTables::m_binFracBits (contexts.cpp:791-878) through the public accessor: out[2 * q + bin].
void ref_frac_bits(uint32_t *out) {
  for (unsigned q = 0; q < 256; q++) {
    BinProbModel_Std m(uint16_t(q << 7), uint16_t(q << 7), 0x58);  // state() = (s0 + s1) >> 8 = q
    const BinFracBits f = m.getFracBitsArray();
    out[2 * q] = f.intBits[0];
    out[2 * q + 1] = f.intBits[1];
  }
}

// RDO estimation of a TU list: CABACWriter::residual_coding on the reference's BitEstimator_Std
// (arith_codec.cpp:603-711).  Returns getEstFracBits(); estPerTu[t] = the estimate after TU t.
uint64_t ref_estimate_residual(const cabac_tu_desc *tus, long ntu, const cabac_substream_params *p,
                               const int32_t *coeffs, const cabac_ctx_model *ctxIn, cabac_ctx_model *ctxOut,
                               uint64_t *estPerTu, int useMTS) {
  Harness hs(*p, useMTS != 0);
  BitEstimator_Std est;
  CABACWriter wr(est);
  est.reset(p->qp, p->init_id);
  if ((p->flags & CABAC_SUB_CTX_IN) && ctxIn) importCtx(est, ctxIn);
  for (long t = 0; t < ntu; t++) {
    TuPair tp;
    makeTu(tp, hs, tus[t], const_cast<TCoeff *>(coeffs));
    applyTsAllowed(tp, tus[t], useMTS != 0);
    CUCtx cuCtx(p->qp);
    wr.residual_coding(tp.tu, ComponentID(tus[t].comp_id), &cuCtx);
    if (estPerTu) estPerTu[t] = est.getEstFracBits();
  }
  if (ctxOut) exportCtx(est, ctxOut);
  return est.getEstFracBits();
}

// BinCounter view of a TU list on the real encoder: ctxBins[k] = getNumBins(k)
// (arith_codec.cpp:281-316), and *choice = CABACWriter::getCtxInitId(slice)
// (cabac_writer.cpp:41-81; SliceType B = 0, P = 1, I = 2) evaluated on that coder.
long ref_residual_ctx_bins(const cabac_tu_desc *tus, long ntu, const cabac_substream_params *p,
                           const int32_t *coeffs, const cabac_ctx_model *ctxIn, cabac_ctx_model *ctxOut,
                           uint32_t *ctxBins, int *choice) {
  Harness hs(*p, false);
  BinEncoder_Std enc;
  CABACWriter wr(enc);
  OutputBitstream bs;
  wr.initBitstream(&bs);
  enc.reset(p->qp, p->init_id);
  enc.riceStatReset(10);
  if ((p->flags & CABAC_SUB_CTX_IN) && ctxIn) importCtx(enc, ctxIn);
  for (long t = 0; t < ntu; t++) {
    TuPair tp;
    makeTu(tp, hs, tus[t], const_cast<TCoeff *>(coeffs));
    applyTsAllowed(tp, tus[t], false);
    CUCtx cuCtx(p->qp);
    wr.residual_coding(tp.tu, ComponentID(tus[t].comp_id), &cuCtx);
  }
  for (unsigned k = 0; k < Ctx::NumberOfContexts; k++) ctxBins[k] = enc.getNumBins(k);
  if (choice) *choice = int(wr.getCtxInitId(*hs.slice));
  if (ctxOut) exportCtx(enc, ctxOut);
  return long(enc.getNumBins());
}

}  // extern "C"
