// gpu_slice_writer.hpp - encode-side shim for WHOLE substreams: the batch counterpart
// of the encoder's per-CTU loop
//     for (ctu in slice) cabacWriter.coding_tree_unit(cs, ctuArea, qps, ctuRsAddr);   cabac_writer.cpp:115-194
//     cabacWriter.end_of_slice()
// for all-intra slices / tiles with separate luma / chroma trees (cabac_encode_slices, SURVEY.md section 8 f2, the
// writer direction).  Nothing is binarised on the host: a slice is queued as the encoder's DECISIONS - one
// cabac_cu_desc per CodingUnit in coding order (partitioning path, prediction mode, coded block flags, ...), one
// cabac_tu_desc + TCoeff block per coded transform block, one cabac_ctu_desc per CTU (SAO / ALF) - and flush()
// walks the syntax on the device (split_cu_mode with the partitioner's legality rules, MPM derivation, cbf / LFNST /
// MTS conditions) and codes it; result(i).bytes is what CABACWriter + BinEncoder would have left in the slice's
// OutputBitstream (terminating bin, finish() and the rbsp trailing bits included) - for a slice with
// entropy_coding_sync_enabled its CTU-row substreams back to back, with the row sizes for the entry points
// (cabac_encode_slices_wpp: the rows are coded in one launch, contexts handed over on the device).
#pragma once
#include <algorithm>
#include <vector>

#include "gpu_bin_encoder.hpp"   // GpuBinEncoder::sideFlags: the CUCtx side effects of residual_coding per block
#include "gpu_slice_reader.hpp"  // sliceParams()
#include "picture.hpp"
#include "unit_partitioner.hpp"

namespace cabac_b200 {

// SAO / ALF / CC-ALF CTU syntax of one CTU, from where CABACWriter::coding_tree_unit reads it
// (sao :196-232 + sao_offset_pars :262-352; codeAlfCtuEnableFlag / FilterIndex / Alternative, codeCcAlfFilterControlIdc).
inline cabac_ctu_desc ctuDesc(const Common::CodingStructure &cs, unsigned ctuRsAddr) {
  namespace C = Common;
  cabac_ctu_desc c{};
  const C::Slice &sl = *cs.slice;
  C::Picture &pic = *cs.picture;
  if (cs.sps->getSAOEnabledFlag()) {
    const C::SAOBlkParam &sao = pic.getSAO()[ctuRsAddr];
    for (int comp = 0; comp < 3; comp++) {
      const C::SAOOffset &o = sao[comp];
      c.sao_mode[comp] = uint8_t(o.modeIdc);  // SAO_MODE_OFF / NEW / MERGE = 0 / 1 / 2
      if (o.modeIdc == C::SAO_MODE_MERGE) {
        c.sao_type[comp] = uint8_t(o.typeIdc);  // SAO_MERGE_LEFT / ABOVE
      } else if (o.modeIdc == C::SAO_MODE_NEW) {
        const bool band = o.typeIdc == C::SAO_TYPE_BO;
        c.sao_type[comp] = band ? 0 : uint8_t(1 + o.typeIdc - C::SAO_TYPE_START_EO);
        c.sao_band[comp] = band ? uint8_t(o.typeAuxInfo) : 0;
        static const int eoClass[4] = {0, 1, 3, 4};  // SAO_CLASS_EO_PLAIN (2) carries no offset
        for (int k = 0; k < 4; k++)
          c.sao_offset[comp][k] = int8_t(band ? o.offset[(o.typeAuxInfo + k) % 32] : o.offset[eoClass[k]]);
      }
    }
  }
  if (cs.sps->getALFEnabledFlag()) {
    for (int comp = 0; comp < 3; comp++)
      c.alf_flag[comp] = sl.getAlfEnabledFlag(C::ComponentID(comp)) ? pic.getAlfCtuEnableFlag(comp)[ctuRsAddr] : 0;
    c.alf_filter_idx = uint8_t(pic.getAlfCtbFilterIndex()[ctuRsAddr]);
    for (int comp = 1; comp < 3; comp++) c.alf_alt[comp - 1] = pic.getAlfCtuAlternativeData(comp)[ctuRsAddr];
  }
  if (cs.sps->getCCALFEnabledFlag())
    for (int k = 0; k < 2; k++)
      if (sl.m_ccAlfFilterParam.ccAlfFilterEnabled[k] && sl.m_ccAlfFilterControl[k])
        c.ccalf_idc[k] = sl.m_ccAlfFilterControl[k][ctuRsAddr];
  return c;
}

class GpuSliceWriter {
 public:
  struct Result {
    const uint8_t *bytes;         // the slice data: a WPP slice's CTU-row substreams back to back
    size_t nbytes;
    const uint32_t *rowBytes;     // bytes per row substream (the entry points of a WPP slice); nRows == 1 otherwise
    size_t nRows;
    const cabac_ctx_model *ctx;   // Ctx after the slice
    uint32_t numBins, status;
  };

  explicit GpuSliceWriter(Engine &eng) : eng_(eng) {}

  // Opens the next slice / tile: slice QP / init table as CABACWriter::initCtxModels takes them
  // (cabac_writer.cpp:60-85), the slice-level facts (sliceParams(), gpu_slice_reader.hpp).
  size_t beginSlice(const Common::CodingStructure &cs, int qp, int initId, const cabac_slice_params &sp,
                    const cabac_ctx_model *ctxIn = nullptr) {
    const Common::Slice &sl = *cs.slice;
    cabac_substream_params p{};
    p.qp = int8_t(qp);
    p.init_id = uint8_t(initId);
    p.flags = uint8_t((sl.getDepQuantEnabledFlag() ? CABAC_SUB_DEP_QUANT : 0) |
                      (sl.getSignDataHidingEnabledFlag() ? CABAC_SUB_SIGN_HIDING : 0) |
                      (sl.getTSResidualCodingDisabledFlag() ? CABAC_SUB_TSRC_DISABLED : 0));
    return beginSliceRaw(p, sp, ctxIn);
  }
  size_t beginSliceRaw(cabac_substream_params p, const cabac_slice_params &sp, const cabac_ctx_model *ctxIn = nullptr) {
    closeSlice();
    p.flags = uint8_t((p.flags & (CABAC_SUB_DEP_QUANT | CABAC_SUB_SIGN_HIDING | CABAC_SUB_TSRC_DISABLED)) |
                      CABAC_SUB_TERMINATE | CABAC_SUB_ALIGN | (ctxIn ? CABAC_SUB_CTX_IN : 0));
    params_.push_back(p);
    sps_.push_back(sp);
    ctxIn_.resize(params_.size() * CABAC_NUM_CTX);
    if (ctxIn) std::copy(ctxIn, ctxIn + CABAC_NUM_CTX, ctxIn_.begin() + (params_.size() - 1) * CABAC_NUM_CTX);
    anyCtxIn_ |= ctxIn != nullptr;
    const uint32_t ctu = 1u << sp.log2_ctu;
    rowOff_.push_back(rowOff_.back() + ((sp.tools & CABAC_SL_WPP) ? (sp.y1 - sp.y0 + ctu - 1) / ctu : 1u));
    rowCoeffs_.resize(rowOff_.back(), 0);
    curRow_ = 0;
    open_ = true;
    return params_.size() - 1;
  }

  // One CTU of the open slice, in coding order: its SAO / ALF record.  Its CUs follow through addCu() in the order
  // the syntax holds them - the order in which CABACWriter::coding_tree (cabac_writer.cpp:355-512) reaches
  // coding_unit(): per CTU the luma tree then the chroma tree, per 64x64 node when the CTU is 128x128.
  void addCtu(const Common::CodingStructure &cs, unsigned ctuRsAddr) { addCtuRaw(ctuDesc(cs, ctuRsAddr)); }
  void addCtuRaw(const cabac_ctu_desc &c) { ctus_.push_back(c); }

  // One CodingUnit of the open slice (intra, dual tree): what CABACWriter::coding_unit :586-700, cu_pred_data,
  // cu_residual / transform_tree :2059-2217 and transform_unit :2219-2354 read from it.
  // qpDelta: cu.qp - CU::predictQP(cu, prevQp), what cu_qp_delta would code if this CU opens the coded part of a
  // quantisation group (:2356-2395; ignored by the device where the syntax has no cu_qp_delta).
  void addCu(const Common::CodingUnit &cu, int qpDelta = 0) {
    namespace C = Common;
    CHECK(cu.predMode != C::MODE_INTRA, "B200 CABAC engine: the device writer codes intra CUs");
    const bool chroma = cu.chType == C::CHANNEL_TYPE_CHROMA;
    const C::CompArea &b = cu.blocks[chroma ? C::COMPONENT_Cb : C::COMPONENT_Y];
    const int sh = chroma ? 1 : 0;  // 4:2:0
    const C::PredictionUnit &pu = *cu.firstPU;
    cabac_cu_desc d{};
    d.x = uint16_t(b.x << sh);
    d.y = uint16_t(b.y << sh);
    setRow(d.y);
    d.log2_w = uint8_t(C::floorLog2(b.width << sh));
    d.log2_h = uint8_t(C::floorLog2(b.height << sh));
    d.ch_type = uint8_t(cu.chType);
    d.qt_depth = cu.qtDepth;
    d.mt_depth = cu.mtDepth;
    d.flags = uint8_t(CABAC_CU_FROM_DIR | (cu.mipFlag && !chroma ? CABAC_CU_MIP : 0) |
                      (cu.mipFlag && !chroma && pu.mipTransposedFlag ? CABAC_CU_MIP_TRANSPOSED : 0));
    d.bdpcm = uint8_t(chroma ? cu.bdpcmModeChroma : cu.bdpcmMode);
    d.mrl_idx = chroma ? 0 : uint8_t(pu.multiRefIdx);
    d.isp = chroma ? 0 : cu.ispMode;
    d.lfnst_idx = uint8_t(cu.lfnstIdx);
    d.intra_dir = uint8_t(pu.intraDir[chroma ? 1 : 0]);
    d.qp_delta = int8_t(qpDelta);
    d.chroma_qp_adj = uint8_t(cu.chromaQpAdj);
    // CU::getSplitAtDepth (unit_tools.cpp:115-140): SPLIT_DMULT bits per depth, CU_QUAD_SPLIT .. CU_TRIV_SPLIT = 1 .. 5.
    // The record's path starts at the root of the walk: the 64x64 node below the shared implicit quad split of a
    // 128x128 dual-tree CTU, else the CTU.
    const unsigned skip = cu.cs->pcv->maxCUWidth > 64 ? 1 : 0;
    for (unsigned k = skip; k < cu.depth && k - skip < 10; k++)
      d.split_series |= uint32_t((cu.splitSeries >> (k * C::SPLIT_DMULT)) & 7u) << (3 * (k - skip));
    d.first_tu = uint32_t(tus_.size() - tuOff_.back());
    unsigned part = 0;  // ISP sub-partition / maximum-TB tile index within the CU
    for (const C::TransformUnit *tu = cu.firstTU; tu; tu = tu == cu.lastTU ? nullptr : tu->next, part++) {
      CHECK(part > 3, "B200 CABAC engine: more than four transform units in a CU");
      if (!chroma) {
        if (C::TU::getCbf(*tu, C::COMPONENT_Y)) {
          d.cbf_mask |= uint16_t(1u << part);
          addTu(*tu, C::COMPONENT_Y);
        }
        if (part == 0) d.mts_idx = tu->mtsIdx[C::COMPONENT_Y];
      } else {
        const bool cb = C::TU::getCbf(*tu, C::COMPONENT_Cb), cr = C::TU::getCbf(*tu, C::COMPONENT_Cr);
        d.cbf_mask |= uint16_t((cb ? 1u << (4 + part) : 0u) | (cr ? 1u << (8 + part) : 0u));
        if (tu->jointCbCr) {
          d.cbf_mask |= uint16_t(1u << (12 + part));
          d.flags |= CABAC_CU_JCCR;
        }
        if (cb) addTu(*tu, C::COMPONENT_Cb);
        if (cr && !(tu->jointCbCr == 3)) addTu(*tu, C::COMPONENT_Cr);  // cabac_writer.cpp:2428-2430
      }
    }
    d.n_tu = uint16_t(tus_.size() - tuOff_.back() - d.first_tu);
    addCuRaw(d);
  }
  // Raw records: a CU, then its coded transform blocks (so that the output capacity can be sized per CTU row).
  void addCuRaw(const cabac_cu_desc &d) {
    setRow(d.y);
    cus_.push_back(d);
  }
  // A coded transform block with its coefficients (row-major, stride == width) and CABAC_TUR_* side flags.
  void addTuRaw(cabac_tu_desc d, const int32_t *coeff, uint8_t tuFlags) {
    d.coeff_offset = coeffs_.size();
    coeffs_.insert(coeffs_.end(), coeff, coeff + (size_t(1) << (d.log2_w + d.log2_h)));
    rowCoeffs_[rowOff_[params_.size() - 1] + curRow_] += uint64_t(1) << (d.log2_w + d.log2_h);
    tus_.push_back(d);
    tuFlags_.push_back(tuFlags);
  }

  // Codes everything queued in one call.  Throws Common::Exception like the reference's CHECKs when the records do
  // not describe a legal partitioning or use a tool outside the built scope.
  void flush() {
    closeSlice();
    const size_t n = params_.size();
    const size_t nRows = rowOff_.back();
    std::vector<uint64_t> rowCap(1, 0);
    size_t words = 0;
    for (size_t i = 0; i < n; i++) {
      const cabac_slice_params &s = sps_[i];
      const uint64_t area = uint64_t(s.x1 - s.x0) * (s.y1 - s.y0), rows = rowOff_[i + 1] - rowOff_[i];
      for (uint64_t r = rowOff_[i]; r < rowOff_[i + 1]; r++)  // <= 20 bits per coefficient + syntax
        rowCap.push_back(rowCap.back() + area / rows / 4 + rowCoeffs_[r] * 3 + 256);
    }
    words = 80 * cus_.size() + tus_.size() + 450 * ctus_.size() + 64 * n + 64;  // the library's own staging bound
    rows_.resize(rowCap.back());
    sizes_.resize(nRows);
    rowBins_.resize(nRows);
    rowStatus_.resize(nRows);
    rowCtx_.resize(nRows * CABAC_NUM_CTX);
    const int32_t none = 0;
    const int rc = cabac_encode_slices_wpp(eng_.get(), params_.data(), sps_.data(), n, cus_.data(), cuOff_.data(), tus_.data(),
                                           tuFlags_.data(), tuOff_.data(), ctus_.data(), ctuOff_.data(),
                                           coeffs_.empty() ? &none : coeffs_.data(), coeffs_.size(),
                                           anyCtxIn_ ? ctxIn_.data() : nullptr, words, rowOff_.data(), nRows, rows_.data(),
                                           rowCap.data(), sizes_.data(), rowBins_.data(), rowCtx_.data(), rowStatus_.data());
    eng_.check(rc, "cabac_encode_slices_wpp");
    // per slice: the rows back to back, bins summed, the last row's contexts
    out_.clear();
    outOff_.assign(1, 0);
    bins_.assign(n, 0);
    status_.assign(n, 0);
    ctxOut_.resize(n * CABAC_NUM_CTX);
    for (size_t i = 0; i < n; i++) {
      for (uint64_t r = rowOff_[i]; r < rowOff_[i + 1]; r++) {
        out_.insert(out_.end(), rows_.begin() + rowCap[r], rows_.begin() + rowCap[r] + sizes_[r]);
        bins_[i] += rowBins_[r];
        status_[i] |= rowStatus_[r];
      }
      outOff_.push_back(out_.size());
      std::copy_n(rowCtx_.begin() + (rowOff_[i + 1] - 1) * CABAC_NUM_CTX, CABAC_NUM_CTX, ctxOut_.begin() + i * CABAC_NUM_CTX);
      const uint32_t st = status_[i];
      CHECK(st & CABAC_SS_UNSUPPORTED, "B200 CABAC engine: slice uses a tool outside the device writer's scope");
      CHECK(st & CABAC_SS_BAD_DESC, "B200 CABAC engine: the CU / TU records do not describe a legal partitioning");
      CHECK(st & CABAC_SS_COEFF_RANGE, "TransCoeffLevel outside allowable range");
      CHECK(st & (CABAC_SS_CAPACITY | CABAC_SS_OUT_OVERFLOW), "B200 CABAC engine: output capacity exceeded");
    }
    doneRowOff_ = rowOff_;
    rowOff_.assign(1, 0);
    rowCoeffs_.clear();
    done_ = n;
    params_.clear();
    sps_.clear();
    cus_.clear();
    tus_.clear();
    tuFlags_.clear();
    ctus_.clear();
    coeffs_.clear();
    ctxIn_.clear();
    cuOff_.assign(1, 0);
    tuOff_.assign(1, 0);
    ctuOff_.assign(1, 0);
    coOff_.assign(1, 0);
    anyCtxIn_ = false;
  }

  size_t size() const { return done_; }
  Result result(size_t i) const {
    return Result{out_.data() + outOff_[i], size_t(outOff_[i + 1] - outOff_[i]), sizes_.data() + doneRowOff_[i],
                  size_t(doneRowOff_[i + 1] - doneRowOff_[i]), ctxOut_.data() + i * CABAC_NUM_CTX, bins_[i], status_[i]};
  }

 private:
  void addTu(const Common::TransformUnit &tu, Common::ComponentID compID) {
    namespace C = Common;
    GpuBinEncoder::refuseUnsupportedTools(tu);
    const C::CompArea &blk = tu.blocks[compID];
    const C::CodingUnit &cu = *tu.cu;
    const bool bdpcm = C::isLuma(compID) ? cu.bdpcmMode != 0 : cu.bdpcmModeChroma != 0;
    const bool ts = tu.mtsIdx[compID] == C::MTS_SKIP;
    cabac_tu_desc d{};
    d.log2_w = uint8_t(C::floorLog2(blk.width));
    d.log2_h = uint8_t(C::floorLog2(blk.height));
    d.comp_id = uint8_t(compID);
    d.flags = uint8_t((ts ? CABAC_TU_TS : 0) | (C::TU::isTSAllowed(tu, compID) ? CABAC_TU_TS_ALLOWED : 0) |
                      (bdpcm ? CABAC_TU_BDPCM : 0));
    addTuRaw(d, tu.getCoeffs(compID).buf, uint8_t(GpuBinEncoder::sideFlags(tu, compID) | (ts ? CABAC_TUR_TS : 0)));
  }
  void setRow(unsigned y) {
    const cabac_slice_params &sp = sps_.back();
    curRow_ = (sp.tools & CABAC_SL_WPP) ? (y - sp.y0) >> sp.log2_ctu : 0;
    CHECK(rowOff_[params_.size() - 1] + curRow_ >= rowOff_.back(), "B200 CABAC engine: CU outside the slice's rectangle");
  }
  void closeSlice() {
    if (!open_) return;
    cuOff_.push_back(cus_.size());
    tuOff_.push_back(tus_.size());
    ctuOff_.push_back(ctus_.size());
    coOff_.push_back(coeffs_.size());
    open_ = false;
  }

  Engine &eng_;
  std::vector<cabac_substream_params> params_;
  std::vector<cabac_slice_params> sps_;
  std::vector<cabac_cu_desc> cus_;
  std::vector<cabac_tu_desc> tus_;
  std::vector<uint8_t> tuFlags_, out_;
  std::vector<cabac_ctu_desc> ctus_;
  std::vector<int32_t> coeffs_;
  std::vector<uint8_t> rows_;
  std::vector<uint64_t> cuOff_{0}, tuOff_{0}, ctuOff_{0}, coOff_{0}, outOff_, rowOff_{0}, doneRowOff_, rowCoeffs_;
  std::vector<cabac_ctx_model> ctxIn_, ctxOut_, rowCtx_;
  std::vector<uint32_t> sizes_, bins_, status_, rowBins_, rowStatus_;
  unsigned curRow_ = 0;
  bool anyCtxIn_ = false, open_ = false;
  size_t done_ = 0;
};

}  // namespace cabac_b200
