// gpu_slice_reader.hpp - decode-side shim for WHOLE substreams: the batch
// counterpart of the decoder's per-CTU loop
//     for (ctu in slice) cabacReader.coding_tree_unit(cs, ctuArea, qps, ctuRsAddr);   DecSlice.cpp:150-262
//     cabacReader.terminating_bit()
// for all-intra slices / tiles with separate luma / chroma trees (cabac_decode_slices).
// A slice is queued as its payload bytes plus the slice / SPS / PPS facts the CTU syntax depends on
// (sliceParams(): the getters CABACReader::coding_tree_unit and its callees read, cabac_reader.cpp:103-2863);
// flush() parses every queued slice on the device and hands back, per slice, flat records in coding order:
// one cabac_cu_desc per CodingUnit (what coding_unit() leaves in CodingUnit / PredictionUnit as syntax element
// values), one cabac_tu_desc + TCoeff block per coded transform block, one cabac_ctu_desc per CTU (SAO / ALF),
// the final contexts (WPP / inter-slice hand-over) and the bits read (CABACReader::getNumBitsRead).
// Rebuilding the reference's CodingStructure from the records is the caller's loop over cus[] (addCU / addPU /
// addTU with the recorded areas): the records carry everything the reader sets.
#pragma once
#include <algorithm>
#include <vector>

#include "cabac_host.hpp"
#include "coding_structure.hpp"
#include "slice.hpp"
#include "unit.hpp"
#include "unit_tools.hpp"

namespace cabac_b200 {
using Common::Exception;  // CHECK / THROW (type_def.hpp:295-329) name it unqualified

// The slice-level inputs of the CTU syntax.  [x0, x1) x [y0, y1) is the rectangle of CTUs the substream covers
// (a slice of whole CTU rows, or a tile), in luma samples, clipped to the picture.
inline cabac_slice_params sliceParams(const Common::CodingStructure &cs, unsigned x0, unsigned y0, unsigned x1,
                                      unsigned y1) {
  namespace C = Common;
  const C::SPS &sps = *cs.sps;
  const C::PPS &pps = *cs.pps;
  const C::Slice &sl = *cs.slice;
  const C::PreCalcValues &pcv = *cs.pcv;
  auto l2 = [](unsigned v) { return uint32_t(C::floorLog2(v)); };
  cabac_slice_params p{};
  p.pic_w = pps.getPicWidthInLumaSamples();
  p.pic_h = pps.getPicHeightInLumaSamples();
  p.x0 = x0;
  p.y0 = y0;
  p.x1 = std::min<unsigned>(x1, p.pic_w);
  p.y1 = std::min<unsigned>(y1, p.pic_h);
  p.log2_ctu = pcv.maxCUWidthLog2;
  for (int ch = 0; ch < 2; ch++) {
    const C::ChannelType ct = C::ChannelType(ch);
    p.log2_min_qt[ch] = l2(pcv.getMinQtSize(sl, ct));
    p.log2_max_bt[ch] = l2(pcv.getMaxBtSize(sl, ct));
    p.log2_max_tt[ch] = l2(pcv.getMaxTtSize(sl, ct));
    p.max_mtt_depth[ch] = pcv.getMaxBtDepth(sl, ct);
    p.log2_min_bt[ch] = l2(pcv.getMinBtSize(sl, ct));
    p.log2_min_tt[ch] = l2(pcv.getMinTtSize(sl, ct));
  }
  p.log2_max_tb = l2(sps.getMaxTbSize());
  p.log2_max_ts = sps.getLog2MaxTransformSkipBlockSize();
  p.bit_depth = uint32_t(sps.getBitDepth(C::CHANNEL_TYPE_LUMA));
  p.alf_num_aps_luma = uint32_t(sl.getNumAlfApsIdsLuma());
  p.alf_chroma_num_alts = 1;
  const bool alf = sps.getALFEnabledFlag() && sl.getAlfEnabledFlag(C::COMPONENT_Y);
  if (sps.getALFEnabledFlag() && (sl.getAlfEnabledFlag(C::COMPONENT_Cb) || sl.getAlfEnabledFlag(C::COMPONENT_Cr))) {
    const C::APS *aps = sl.getAlfAPSs()[sl.getAlfApsIdChroma()];
    if (aps) p.alf_chroma_num_alts = uint32_t(aps->getAlfAPSParam().numAlternativesChroma);
  }
  const bool ccalf = sps.getCCALFEnabledFlag();
  p.ccalf_filter_count[0] = ccalf ? uint32_t(sl.m_ccAlfFilterParam.ccAlfFilterCount[0]) : 0;
  p.ccalf_filter_count[1] = ccalf ? uint32_t(sl.m_ccAlfFilterParam.ccAlfFilterCount[1]) : 0;
  p.cu_qp_delta_subdiv = sl.getCuQpDeltaSubdiv();
  p.cu_chroma_qp_offset_subdiv = sl.getCuChromaQpOffsetSubdiv();
  p.chroma_qp_offset_list_len = uint32_t(pps.getChromaQpOffsetListLen());
  p.chroma_format = uint32_t(sps.getChromaFormatIdc());
  uint32_t t = 0;
  t |= C::CS::isDualITree(cs) ? CABAC_SL_DUAL_TREE : 0;
  t |= (sps.getSAOEnabledFlag() && sl.getSaoEnabledFlag(C::CHANNEL_TYPE_LUMA)) ? CABAC_SL_SAO_LUMA : 0;
  t |= (sps.getSAOEnabledFlag() && sl.getSaoEnabledFlag(C::CHANNEL_TYPE_CHROMA)) ? CABAC_SL_SAO_CHROMA : 0;
  t |= alf ? CABAC_SL_ALF_Y : 0;
  t |= (alf && sl.getAlfEnabledFlag(C::COMPONENT_Cb)) ? CABAC_SL_ALF_CB : 0;
  t |= (alf && sl.getAlfEnabledFlag(C::COMPONENT_Cr)) ? CABAC_SL_ALF_CR : 0;
  t |= (ccalf && sl.m_ccAlfFilterParam.ccAlfFilterEnabled[0]) ? CABAC_SL_CCALF_CB : 0;
  t |= (ccalf && sl.m_ccAlfFilterParam.ccAlfFilterEnabled[1]) ? CABAC_SL_CCALF_CR : 0;
  t |= sps.getUseIntraMTS() ? CABAC_SL_MTS_INTRA : 0;
  t |= sps.getUseLFNST() ? CABAC_SL_LFNST : 0;
  t |= sps.getUseISP() ? CABAC_SL_ISP : 0;
  t |= sps.getUseMIP() ? CABAC_SL_MIP : 0;
  t |= sps.getUseMRL() ? CABAC_SL_MRL : 0;
  t |= sps.getTransformSkipEnabledFlag() ? CABAC_SL_TS : 0;
  t |= sps.getBDPCMEnabledFlag() ? CABAC_SL_BDPCM : 0;
  t |= sps.getJointCbCrEnabledFlag() ? CABAC_SL_JCCR : 0;
  t |= sps.getUseLMChroma() ? CABAC_SL_CCLM : 0;
  t |= pps.getUseDQP() ? CABAC_SL_DQP : 0;
  t |= sl.getUseChromaQpAdj() ? CABAC_SL_CHROMA_QP_ADJ : 0;
  t |= sps.getEntropyCodingSyncEnabledFlag() ? CABAC_SL_WPP : 0;
  t |= sps.getIBCFlag() ? CABAC_SL_IBC : 0;
  t |= sps.getPLTMode() ? CABAC_SL_PLT : 0;
  t |= sps.getUseColorTrans() ? CABAC_SL_ACT : 0;
  t |= sl.isIntra() ? CABAC_SL_INTRA : 0;
  t |= sps.getUseMTS() ? CABAC_SL_MTS : 0;
  p.tools = t;
  return p;
}

class GpuSliceReader {
 public:
  struct Result {  // one decoded slice; pointers into the reader's buffers, valid until the next flush()
    const cabac_cu_desc *cus;
    size_t nCus;
    const cabac_tu_desc *tus;     // coeff_offset indexes coeffs()
    const uint8_t *tuFlags;       // CABAC_TUR_*
    size_t nTus;
    const cabac_ctu_desc *ctus;
    size_t nCtus;
    const cabac_ctx_model *ctx;   // Ctx after the slice (CABACReader::getCtx())
    uint32_t numBins, bitsRead, status;
  };

  explicit GpuSliceReader(Engine &eng) : eng_(eng) {}

  // Queues one slice / tile: its payload as InputBitstream::getFifo holds it, the slice QP / init table
  // (CABACReader::initCtxModels, cabac_reader.cpp:34-60) and the slice-level facts (sliceParams()).
  size_t add(const uint8_t *bytes, size_t nbytes, const Common::CodingStructure &cs, int qp, int initId,
             const cabac_slice_params &sp, const cabac_ctx_model *ctxIn = nullptr) {
    {  // VVC v1 residual coding only (as GpuResidualReader)
      const auto &rext = cs.sps->getSpsRangeExtension();
      CHECK(rext.getPersistentRiceAdaptationEnabledFlag() || rext.getRrcRiceExtensionEnableFlag() ||
                rext.getExtendedPrecisionProcessingFlag() ||
                (rext.getTSRCRicePresentFlag() && cs.slice->get_tsrc_index() != 0),
            "B200 CABAC engine: range-extension residual coding tools are not supported");
    }
    const Common::Slice &sl = *cs.slice;
    cabac_substream_params p{};
    p.qp = int8_t(qp);
    p.init_id = uint8_t(initId);
    p.flags = uint8_t((sl.getDepQuantEnabledFlag() ? CABAC_SUB_DEP_QUANT : 0) |
                      (sl.getSignDataHidingEnabledFlag() ? CABAC_SUB_SIGN_HIDING : 0) |
                      (sl.getTSResidualCodingDisabledFlag() ? CABAC_SUB_TSRC_DISABLED : 0));
    return addRaw(bytes, nbytes, p, sp, ctxIn);
  }
  // The same with the substream parameters already flat (qp, init table, CABAC_SUB_DEP_QUANT / SIGN_HIDING /
  // TSRC_DISABLED); terminate and ctx-in are set here.
  size_t addRaw(const uint8_t *bytes, size_t nbytes, cabac_substream_params p, const cabac_slice_params &sp,
                const cabac_ctx_model *ctxIn = nullptr) {
    p.flags = uint8_t((p.flags & (CABAC_SUB_DEP_QUANT | CABAC_SUB_SIGN_HIDING | CABAC_SUB_TSRC_DISABLED)) |
                      CABAC_SUB_TERMINATE | (ctxIn ? CABAC_SUB_CTX_IN : 0));
    params_.push_back(p);
    sps_.push_back(sp);
    in_.insert(in_.end(), bytes, bytes + nbytes);
    inOff_.push_back(in_.size());
    ctxIn_.resize(params_.size() * CABAC_NUM_CTX);
    if (ctxIn) std::copy(ctxIn, ctxIn + CABAC_NUM_CTX, ctxIn_.begin() + (params_.size() - 1) * CABAC_NUM_CTX);
    anyCtxIn_ |= ctxIn != nullptr;
    return params_.size() - 1;
  }

  // Parses everything queued in one launch.  Throws Common::Exception like the reference's CHECKs when a slice is
  // malformed (overrun, missing terminating bit, level out of range) or uses a tool outside the built scope.
  void flush() {
    const size_t n = params_.size();
    cuOff_.assign(1, 0);
    tuOff_.assign(1, 0);
    ctuOff_.assign(1, 0);
    coOff_.assign(1, 0);
    for (size_t i = 0; i < n; i++) {
      const cabac_slice_params &s = sps_[i];
      const uint64_t w = s.x1 - s.x0, h = s.y1 - s.y0, area = w * h, ctu = uint64_t(1) << s.log2_ctu;
      cuOff_.push_back(cuOff_.back() + area / 16 * 2 + 8);
      tuOff_.push_back(tuOff_.back() + area / 16 * 2 + 8);
      ctuOff_.push_back(ctuOff_.back() + ((w + ctu - 1) / ctu) * ((h + ctu - 1) / ctu));
      coOff_.push_back(coOff_.back() + area * 3 / 2 + 64);
    }
    cus_.resize(cuOff_.back());
    tus_.resize(tuOff_.back());
    tuFlags_.resize(tuOff_.back());
    ctus_.resize(ctuOff_.back());
    coeffs_.resize(coOff_.back());
    counts_.resize(3 * n);
    bins_.resize(n);
    bits_.resize(n);
    status_.resize(n);
    ctxOut_.resize(n * CABAC_NUM_CTX);
    std::vector<uint64_t> inOff(1, 0);
    inOff.insert(inOff.end(), inOff_.begin(), inOff_.end());
    const int rc = cabac_decode_slices(eng_.get(), in_.data(), inOff.data(), params_.data(), sps_.data(), n,
                                       anyCtxIn_ ? ctxIn_.data() : nullptr, cuOff_.data(), tuOff_.data(), ctuOff_.data(),
                                       coOff_.data(), cus_.data(), tus_.data(), tuFlags_.data(), ctus_.data(),
                                       coeffs_.data(), counts_.data(), bins_.data(), bits_.data(), ctxOut_.data(),
                                       status_.data());
    eng_.check(rc, "cabac_decode_slices");
    for (size_t i = 0; i < n; i++) {
      const uint32_t st = status_[i];
      CHECK(st & CABAC_SS_UNSUPPORTED, "B200 CABAC engine: slice uses a tool outside the device parser's scope");
      CHECK(st & CABAC_SS_IN_OVERRUN, "FIFO exceeded");                                  // bit_stream.cpp:269
      CHECK(st & CABAC_SS_COEFF_RANGE, "TransCoeffLevel outside allowable range");        // cabac_reader.cpp:2942
      CHECK(st & CABAC_SS_NO_TERMINATE, "Expecting a terminating bit");                   // DecSlice.cpp:252
      CHECK(st & CABAC_SS_CAPACITY, "B200 CABAC engine: record capacity exceeded");
    }
    done_ = n;
    params_.clear();
    sps_.clear();
    in_.clear();
    inOff_.clear();
    ctxIn_.clear();
    anyCtxIn_ = false;
  }

  size_t size() const { return done_; }
  const int32_t *coeffs() const { return coeffs_.data(); }
  Result result(size_t i) const {
    Result r;
    r.cus = cus_.data() + cuOff_[i];
    r.nCus = counts_[3 * i];
    r.tus = tus_.data() + tuOff_[i];
    r.tuFlags = tuFlags_.data() + tuOff_[i];
    r.nTus = counts_[3 * i + 1];
    r.ctus = ctus_.data() + ctuOff_[i];
    r.nCtus = size_t(ctuOff_[i + 1] - ctuOff_[i]);
    r.ctx = ctxOut_.data() + i * CABAC_NUM_CTX;
    r.numBins = bins_[i];
    r.bitsRead = bits_[i];
    r.status = status_[i];
    return r;
  }

 private:
  Engine &eng_;
  std::vector<cabac_substream_params> params_;
  std::vector<cabac_slice_params> sps_;
  std::vector<uint8_t> in_, tuFlags_;
  std::vector<uint64_t> inOff_, cuOff_, tuOff_, ctuOff_, coOff_;
  std::vector<cabac_ctx_model> ctxIn_, ctxOut_;
  std::vector<cabac_cu_desc> cus_;
  std::vector<cabac_tu_desc> tus_;
  std::vector<cabac_ctu_desc> ctus_;
  std::vector<int32_t> coeffs_;
  std::vector<uint32_t> counts_, bins_, bits_, status_;
  bool anyCtxIn_ = false;
  size_t done_ = 0;
};

}  // namespace cabac_b200
