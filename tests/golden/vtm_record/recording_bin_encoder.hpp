// recording_bin_encoder.hpp - fixture recorder for the reference's final bitstream pass.
//
// TEST INFRASTRUCTURE.  Compiled only into a /tmp copy of p-sawicki/entropy_coding by
// tests/golden/make_vtm_streams.py (never into the product, never into /root/reference).
// It replaces the type of CABACEncoder::m_entropyCodingBinEncoderStd
// (VTM/Lib/EncoderLib/CABACWriter.h:306) by a subclass of the library's own
// EntropyCoding::BinEncoder_Std that forwards every BinEncIf call (arith_codec.hpp:40-69) to
// the base coder - so str.bin stays what the unpatched build writes - and records it with
// the repo's reference-independent cabac_b200::StreamRecorder (host/cabac_host.hpp).
// EntropyCoding::CABACWriter::residual_coding (src/entropy_codec/cabac_writer.cpp:2424) is
// bracketed by a TuScope: the bins of the TU are replaced by one CABAC_CMD_TU marker plus the
// flat descriptor and the TCoeff block, exactly what GpuBinEncoder::residual_coding records.
//
// One record per substream (start() .. finish()) is appended to $VTM_RECORD_OUT:
//   header  u32 magic 'CBRC', i32 qp, u32 init_id, u32 flags, u32 n_cmds, u32 n_tus,
//           u32 n_coeffs, u32 n_payload, u32 num_bins, u32 n_ctu, u32 tools, u32 n_full
//   u32 cmds[n_cmds]; cabac_tu_desc tus[n_tus]; i32 coeffs[n_coeffs];
//   u8 cuctx_before[n_tus], cuctx_after[n_tus]   (CABAC_TUR_* bits of the CUCtx around the call)
//   u32 ctu_marks[2 * n_ctu]                     (cmd index, tu index at every coding_tree_unit)
//   u16 ctx_in[2 * 379], ctx_out[2 * 379]        (state0, state1 at the first bin / after finish)
//   u32 ctx_bins[379]                            (BinCounter::getCtx(k) at finish)
//   u8 payload[n_payload]                        (FIFO after finish() + writeByteAlignment())
//   u32 full_cmds[n_full]                        (every call incl. the residual bins: the decoder's skeleton)
// Records with magic 'CBR2' carry two more header words (n_slice, n_cus) and, after full_cmds,
//   u32 slice[n_slice]                           (the cabac_slice_params fields the CTU parser needs, see sliceWords())
//   CuRec cus[n_cus]                             (one per CABACWriter::coding_unit call, 16 bytes each)
#pragma once
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "arith_codec.hpp"
#include "cabac_host.hpp"
#include "coding_structure.hpp"
#include "context_modelling.hpp"
#include "slice.hpp"
#include "unit.hpp"
#include "unit_tools.hpp"

namespace vtm_record {

class RecordingBinEncoder : public EntropyCoding::BinEncoder_Std {
  using Base = EntropyCoding::BinEncoder_Std;

 public:
  void start() override {
    Base::start();
    rec_.clear();
    full_.clear();
    rec_.params.flags = 0;
    coeffs_.clear();
    before_.clear();
    after_.clear();
    ctuMarks_.clear();
    slice_.clear();
    cus_.clear();
    haveCtxIn_ = false;
    inTu_ = 0;
    tools_ = 0;
  }
  void reset(int qp, int initId) override {
    rec_.params.qp = int8_t(qp);
    rec_.params.init_id = uint8_t(initId);
    Base::reset(qp, initId);  // -> start()
    rec_.params.qp = int8_t(qp);
    rec_.params.init_id = uint8_t(initId);
  }
  void encodeBin(unsigned bin, unsigned ctxId) override {
    first();
    if (!inTu_) rec_.encodeBin(bin, ctxId);
    full_.encodeBin(bin, ctxId);
    Base::encodeBin(bin, ctxId);
  }
  void encodeBinEP(unsigned bin) override {
    first();
    if (!inTu_) rec_.encodeBinEP(bin);
    full_.encodeBinEP(bin);
    Base::encodeBinEP(bin);
  }
  void encodeBinsEP(unsigned bins, unsigned numBins) override {
    first();
    if (!nested_) {  // BinEncoderBase::encodeRemAbsEP emits its two words through this virtual (arith_codec.cpp:426-458)
      if (!inTu_) rec_.encodeBinsEP(bins, numBins);
      full_.encodeBinsEP(bins, numBins);
    }
    Base::encodeBinsEP(bins, numBins);
  }
  void encodeRemAbsEP(unsigned bins, unsigned goRicePar, unsigned cutoff, int maxLog2TrDynamicRange) override {
    first();
    if (!inTu_) rec_.encodeRemAbsEP(bins, goRicePar, cutoff, maxLog2TrDynamicRange);
    full_.encodeRemAbsEP(bins, goRicePar, cutoff, maxLog2TrDynamicRange);
    nested_ = true;
    Base::encodeRemAbsEP(bins, goRicePar, cutoff, maxLog2TrDynamicRange);
    nested_ = false;
  }
  void encodeBinTrm(unsigned bin) override {
    first();
    if (!inTu_) rec_.encodeBinTrm(bin);
    full_.encodeBinTrm(bin);
    Base::encodeBinTrm(bin);
  }
  void align() override {
    first();
    if (!inTu_) rec_.align();
    full_.align();
    Base::align();
  }
  void finish() override {
    first();
    Base::finish();
    dump();
  }

  // ---- hooks called from the patched reference writer ----
  void markCtu(const Common::CodingStructure &cs, const Common::UnitArea &area) {
    first();
    ctuMarks_.push_back(uint32_t(rec_.cmds.size()));
    ctuMarks_.push_back(uint32_t(rec_.tus.size()));
    if (slice_.empty()) sliceWords(cs);
    // the substream's CTU rectangle (luma samples, exclusive end, clipped to the picture)
    const uint32_t x0 = area.lumaPos().x, y0 = area.lumaPos().y;
    const uint32_t x1 = std::min<uint32_t>(x0 + cs.pcv->maxCUWidth, cs.pps->getPicWidthInLumaSamples());
    const uint32_t y1 = std::min<uint32_t>(y0 + cs.pcv->maxCUHeight, cs.pps->getPicHeightInLumaSamples());
    slice_[2] = std::min(slice_[2], x0);
    slice_[3] = std::min(slice_[3], y0);
    slice_[4] = std::max(slice_[4], x1);
    slice_[5] = std::max(slice_[5], y1);
  }
  struct CuRec {
    uint16_t x, y;  // luma samples
    uint8_t log2w, log2h, chType, qtDepth, mtDepth, flags, bdpcm, mrl, isp, lfnst, mts, dir;
  };
  void markCu(const Common::CodingUnit &cu) {
    namespace C = Common;
    CuRec r{};
    const C::CompArea &b = cu.blocks[cu.chType];
    const int sh = cu.chType == C::CHANNEL_TYPE_CHROMA ? 1 : 0;  // 4:2:0
    r.x = uint16_t(b.x << sh);
    r.y = uint16_t(b.y << sh);
    r.log2w = uint8_t(C::floorLog2(b.width << sh));
    r.log2h = uint8_t(C::floorLog2(b.height << sh));
    r.chType = uint8_t(cu.chType);
    r.qtDepth = cu.qtDepth;
    r.mtDepth = cu.mtDepth;
    const C::PredictionUnit *pu = cu.firstPU;
    r.flags = uint8_t((cu.mipFlag ? 1 : 0) | ((pu && pu->mipTransposedFlag) ? 2 : 0) | (cu.predMode == C::MODE_INTRA ? 4 : 0));
    r.bdpcm = uint8_t(cu.chType == C::CHANNEL_TYPE_CHROMA ? cu.bdpcmModeChroma : cu.bdpcmMode);
    r.mrl = uint8_t(pu ? pu->multiRefIdx : 0);
    r.isp = cu.ispMode;
    r.lfnst = uint8_t(cu.lfnstIdx);
    r.mts = uint8_t(cu.firstTU ? cu.firstTU->mtsIdx[C::COMPONENT_Y] : 0);
    r.dir = uint8_t(pu ? pu->intraDir[cu.chType == C::CHANNEL_TYPE_CHROMA ? 1 : 0] : 0);
    cus_.push_back(r);
  }
  // The slice / C::SPS / C::PPS facts CABACReader::coding_tree_unit reads (cabac_reader.cpp:103-209 and callees);
  // the word order is cabac_slice_params' field order (include/cabac_b200.h).
  void sliceWords(const Common::CodingStructure &cs) {
    namespace C = Common;
    const C::SPS &sps = *cs.sps;
    const C::PPS &pps = *cs.pps;
    const C::Slice &sl = *cs.slice;
    const C::PreCalcValues &pcv = *cs.pcv;
    auto l2 = [](unsigned v) { return uint32_t(C::floorLog2(v)); };
    slice_.assign(32, 0);
    slice_[0] = pps.getPicWidthInLumaSamples();
    slice_[1] = pps.getPicHeightInLumaSamples();
    slice_[2] = slice_[3] = 0xFFFFFFFFu;  // x0, y0 (min over the CTUs)
    slice_[4] = slice_[5] = 0;            // x1, y1
    slice_[6] = pcv.maxCUWidthLog2;
    for (int ch = 0; ch < 2; ch++) {
      const C::ChannelType ct = C::ChannelType(ch);
      slice_[7 + ch] = l2(pcv.getMinQtSize(sl, ct));
      slice_[9 + ch] = l2(pcv.getMaxBtSize(sl, ct));
      slice_[11 + ch] = l2(pcv.getMaxTtSize(sl, ct));
      slice_[13 + ch] = pcv.getMaxBtDepth(sl, ct);
      slice_[15 + ch] = l2(pcv.getMinBtSize(sl, ct));
      slice_[17 + ch] = l2(pcv.getMinTtSize(sl, ct));
    }
    slice_[19] = l2(sps.getMaxTbSize());
    slice_[20] = sps.getLog2MaxTransformSkipBlockSize();
    slice_[21] = uint32_t(sps.getBitDepth(C::CHANNEL_TYPE_LUMA));
    slice_[22] = uint32_t(sl.getNumAlfApsIdsLuma());
    uint32_t numAlts = 1;
    if (sps.getALFEnabledFlag() && (sl.getAlfEnabledFlag(C::COMPONENT_Cb) || sl.getAlfEnabledFlag(C::COMPONENT_Cr))) {
      const C::APS *aps = sl.getAlfAPSs()[sl.getAlfApsIdChroma()];
      if (aps) numAlts = uint32_t(aps->getAlfAPSParam().numAlternativesChroma);
    }
    slice_[23] = numAlts;
    slice_[24] = sps.getCCALFEnabledFlag() ? uint32_t(sl.m_ccAlfFilterParam.ccAlfFilterCount[0]) : 0;
    slice_[25] = sps.getCCALFEnabledFlag() ? uint32_t(sl.m_ccAlfFilterParam.ccAlfFilterCount[1]) : 0;
    slice_[26] = sl.getCuQpDeltaSubdiv();
    slice_[27] = sl.getCuChromaQpOffsetSubdiv();
    slice_[28] = uint32_t(pps.getChromaQpOffsetListLen());
    slice_[29] = uint32_t(sps.getChromaFormatIdc());
    const bool alf = sps.getALFEnabledFlag() && sl.getAlfEnabledFlag(C::COMPONENT_Y);
    const bool ccalf = sps.getCCALFEnabledFlag();
    uint32_t t = 0;
    t |= Common::CS::isDualITree(cs) ? 1u << 0 : 0;
    t |= (sps.getSAOEnabledFlag() && sl.getSaoEnabledFlag(C::CHANNEL_TYPE_LUMA)) ? 1u << 1 : 0;
    t |= (sps.getSAOEnabledFlag() && sl.getSaoEnabledFlag(C::CHANNEL_TYPE_CHROMA)) ? 1u << 2 : 0;
    t |= alf ? 1u << 3 : 0;
    t |= (alf && sl.getAlfEnabledFlag(C::COMPONENT_Cb)) ? 1u << 4 : 0;
    t |= (alf && sl.getAlfEnabledFlag(C::COMPONENT_Cr)) ? 1u << 5 : 0;
    t |= (ccalf && sl.m_ccAlfFilterParam.ccAlfFilterEnabled[0]) ? 1u << 6 : 0;
    t |= (ccalf && sl.m_ccAlfFilterParam.ccAlfFilterEnabled[1]) ? 1u << 7 : 0;
    t |= sps.getUseIntraMTS() ? 1u << 8 : 0;
    t |= sps.getUseLFNST() ? 1u << 9 : 0;
    t |= sps.getUseISP() ? 1u << 10 : 0;
    t |= sps.getUseMIP() ? 1u << 11 : 0;
    t |= sps.getUseMRL() ? 1u << 12 : 0;
    t |= sps.getTransformSkipEnabledFlag() ? 1u << 13 : 0;
    t |= sps.getBDPCMEnabledFlag() ? 1u << 14 : 0;
    t |= sps.getJointCbCrEnabledFlag() ? 1u << 15 : 0;
    t |= sps.getUseLMChroma() ? 1u << 16 : 0;
    t |= pps.getUseDQP() ? 1u << 17 : 0;
    t |= sl.getUseChromaQpAdj() ? 1u << 18 : 0;
    t |= sps.getEntropyCodingSyncEnabledFlag() ? 1u << 19 : 0;
    t |= sps.getIBCFlag() ? 1u << 20 : 0;
    t |= sps.getPLTMode() ? 1u << 21 : 0;
    t |= sps.getUseColorTrans() ? 1u << 22 : 0;
    t |= sl.isIntra() ? 1u << 23 : 0;
    t |= sps.getUseMTS() ? 1u << 24 : 0;
    slice_[30] = t;
  }
  static uint8_t cuCtxBits(const Common::CUCtx *c, Common::ComponentID compID) {
    if (!c) return 0;
    const Common::ChannelType ch = Common::toChannelType(compID);
    return uint8_t((c->violatesLfnstConstrained[ch] ? CABAC_TUR_VIOLATES_LFNST : 0) |
                   (c->lfnstLastScanPos ? CABAC_TUR_LFNST_LAST_POS : 0) |
                   (c->mtsLastScanPos ? CABAC_TUR_MTS_LAST_POS : 0) |
                   (c->violatesMtsCoeffConstraint ? CABAC_TUR_VIOLATES_MTS : 0));
  }
  // Mirrors GpuBinEncoder::residual_coding (host/gpu_bin_encoder.hpp): same descriptor, same slab layout.
  void beginTu(const Common::TransformUnit &tu, Common::ComponentID compID, const Common::CUCtx *cuCtx) {
    first();
    if (compID == Common::COMPONENT_Cr && tu.jointCbCr == 3) {  // cabac_writer.cpp:2428-2430: nothing is coded
      inTu_++;
      skipped_ = true;
      return;
    }
    skipped_ = false;
    const Common::Slice &sl = *tu.cs->slice;
    rec_.params.flags = uint8_t((sl.getDepQuantEnabledFlag() ? CABAC_SUB_DEP_QUANT : 0) |
                                (sl.getSignDataHidingEnabledFlag() ? CABAC_SUB_SIGN_HIDING : 0) |
                                (sl.getTSResidualCodingDisabledFlag() ? CABAC_SUB_TSRC_DISABLED : 0));
    const auto &rext = sl.getSPS()->getSpsRangeExtension();
    tools_ = (rext.getPersistentRiceAdaptationEnabledFlag() ? 1u : 0u) | (rext.getExtendedPrecisionProcessingFlag() ? 2u : 0u) |
             (tu.cs->sps->getUseMTS() ? 4u : 0u) | (tu.cs->sps->getUseLFNST() ? 8u : 0u);
    const Common::CompArea &blk = tu.blocks[compID];
    cabac_tu_desc d{};
    d.log2_w = uint8_t(Common::floorLog2(blk.width));
    d.log2_h = uint8_t(Common::floorLog2(blk.height));
    d.comp_id = uint8_t(compID);
    const Common::CodingUnit &cu = *tu.cu;
    const bool bdpcm = Common::isLuma(compID) ? cu.bdpcmMode != 0 : cu.bdpcmModeChroma != 0;
    d.flags = uint8_t((tu.mtsIdx[compID] == Common::MTS_SKIP ? CABAC_TU_TS : 0) |
                      (Common::TU::isTSAllowed(tu, compID) ? CABAC_TU_TS_ALLOWED : 0) | (bdpcm ? CABAC_TU_BDPCM : 0) |
                      ((tu.cs->sps->getUseMTS() && cu.sbtInfo != 0 && compID == Common::COMPONENT_Y)
                           ? CABAC_TU_SBT_ZEROOUT : 0));
    d.coeff_offset = coeffs_.size();
    const Common::TCoeff *src = tu.getCoeffs(compID).buf;
    coeffs_.insert(coeffs_.end(), src, src + size_t(blk.width) * blk.height);
    rec_.residual(d);
    before_.push_back(cuCtxBits(cuCtx, compID));
    inTu_++;
  }
  void endTu(Common::ComponentID compID, const Common::CUCtx *cuCtx) {
    inTu_--;
    if (!skipped_) after_.push_back(cuCtxBits(cuCtx, compID));
    skipped_ = false;
  }

 private:
  void snapshot(std::vector<uint16_t> &dst) {
    const auto &st = static_cast<const Common::CtxStore<Common::BinProbModel_Std> &>(static_cast<const Common::Ctx &>(*this));
    dst.resize(2 * CABAC_NUM_CTX);
    for (unsigned k = 0; k < CABAC_NUM_CTX; k++) {
      dst[2 * k] = st[k].getState0();
      dst[2 * k + 1] = st[k].getState1();
    }
  }
  void first() {
    if (haveCtxIn_) return;
    haveCtxIn_ = true;
    snapshot(ctxIn_);
  }
  template <class T>
  static void put(FILE *f, const std::vector<T> &v) {
    if (!v.empty()) fwrite(v.data(), sizeof(T), v.size(), f);
  }
  void dump() {
    const char *path = std::getenv("VTM_RECORD_OUT");
    if (!path) return;
    Common::OutputBitstream copy = *m_Bitstream;  // the caller's next step: EncSlice.cpp:2114
    copy.writeByteAlignment();
    const std::vector<uint8_t> &payload = copy.getFIFO();
    std::vector<uint16_t> ctxOut;
    snapshot(ctxOut);
    std::vector<uint32_t> ctxBins(CABAC_NUM_CTX);
    for (unsigned k = 0; k < CABAC_NUM_CTX; k++) ctxBins[k] = Base::getNumBins(k);
    FILE *f = fopen(path, "ab");
    if (!f) return;
    const uint32_t hdr[14] = {0x32524243u, uint32_t(int32_t(rec_.params.qp)), rec_.params.init_id, rec_.params.flags,
                              uint32_t(rec_.cmds.size()), uint32_t(rec_.tus.size()), uint32_t(coeffs_.size()),
                              uint32_t(payload.size()), Base::getNumBins(), uint32_t(ctuMarks_.size() / 2), tools_, uint32_t(full_.cmds.size()),
                              uint32_t(slice_.size()), uint32_t(cus_.size())};
    fwrite(hdr, 4, 14, f);
    put(f, rec_.cmds);
    put(f, rec_.tus);
    put(f, coeffs_);
    put(f, before_);
    put(f, after_);
    put(f, ctuMarks_);
    put(f, ctxIn_);
    put(f, ctxOut);
    put(f, ctxBins);
    put(f, payload);
    put(f, full_.cmds);
    put(f, slice_);
    put(f, cus_);
    fclose(f);
  }

  cabac_b200::StreamRecorder rec_, full_;
  std::vector<int32_t> coeffs_;
  std::vector<uint8_t> before_, after_;
  std::vector<uint32_t> ctuMarks_;
  std::vector<uint16_t> ctxIn_;
  std::vector<uint32_t> slice_;
  std::vector<CuRec> cus_;
  bool haveCtxIn_ = false, skipped_ = false, nested_ = false;
  int inTu_ = 0;
  uint32_t tools_ = 0;
};

// RAII bracket for EntropyCoding::CABACWriter::residual_coding; a no-op on the bit estimator.
struct TuScope {
  RecordingBinEncoder *r;
  Common::ComponentID comp;
  const Common::CUCtx *cu;
  TuScope(EntropyCoding::BinEncIf &enc, const Common::TransformUnit &tu, Common::ComponentID compID, const Common::CUCtx *cuCtx)
      : r(dynamic_cast<RecordingBinEncoder *>(&enc)), comp(compID), cu(cuCtx) {
    if (r) r->beginTu(tu, compID, cuCtx);
  }
  ~TuScope() {
    if (r) r->endTu(comp, cu);
  }
};
inline void markCtu(EntropyCoding::BinEncIf &enc, const Common::CodingStructure &cs, const Common::UnitArea &area) {
  if (auto *r = dynamic_cast<RecordingBinEncoder *>(&enc)) r->markCtu(cs, area);
}
inline void markCu(EntropyCoding::BinEncIf &enc, const Common::CodingUnit &cu) {
  if (auto *r = dynamic_cast<RecordingBinEncoder *>(&enc)) r->markCu(cu);
}

}  // namespace vtm_record
