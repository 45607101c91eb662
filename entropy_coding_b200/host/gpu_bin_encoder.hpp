// gpu_bin_encoder.hpp - the shim a maintainer of p-sawicki/entropy_coding adds to route
// the final-pass bitstream writer through the B200 engine.
//
// Compiled against the REFERENCE's headers (arith_codec.hpp, cabac_writer.hpp, unit.hpp):
//   GpuBinEncoder : EntropyCoding::BinEncIf   records the 21 virtuals of arith_codec.hpp:40-69;
//       CABACWriter(BinEncIf&) (cabac_writer.hpp:16-18) works on it unchanged
//   GpuEncodeSession                          batches the substreams of many GpuBinEncoders
//       into one cabac_encode_stream call and delivers bytes to their OutputBitstreams
//   GpuBinEncoder::residual_coding(tu, compID)  replaces CABACWriter::residual_coding
//       (cabac_writer.hpp:128-129) by one TU marker: binarisation runs on the device
#pragma once
#include <memory>

#include "arith_codec.hpp"
#include "cabac_host.hpp"
#include "unit.hpp"
#include "coding_structure.hpp"
#include "context_modelling.hpp"
#include "slice.hpp"
#include "unit_tools.hpp"

namespace cabac_b200 {
using Common::Exception;  // CHECK / THROW (type_def.hpp:295-329) name it unqualified

class GpuBinEncoder;

class GpuEncodeSession {
 public:
  explicit GpuEncodeSession(Engine &eng) : eng_(eng) {}
  Engine &engine() { return eng_; }
  // TCoeff staging slab the TU markers index (H2D source)
  std::vector<int32_t> coeffs;
  // Inter slices: also count the device-coded bins per context, so that getNumBins(ctxId) and
  // getCtxInitId() see the residual syntax too (one more estimator-kernel pass per flush).
  bool countCtxBins = false;
  void submit(GpuBinEncoder *enc, const StreamRecorder &rec, bool ctxIn);
  // Runs every submitted substream in one launch sequence and completes the encoders:
  // bytes into their OutputBitstream, contexts into their Ctx, CUCtx side flags.
  void flush();
  size_t pending() const { return owners_.size(); }

 private:
  Engine &eng_;
  EncodeBatch batch_;
  std::vector<GpuBinEncoder *> owners_;
};

class GpuBinEncoder : public EntropyCoding::BinEncIf {
 public:
  // immediate mode: finish() launches this one substream and returns with the bytes written
  explicit GpuBinEncoder(Engine &eng)
      : BinEncIf(static_cast<const Common::BinProbModel_Std *>(nullptr)), own_(new GpuEncodeSession(eng)),
        session_(own_.get()) {}
  // batched mode: finish() queues the substream; GpuEncodeSession::flush() completes it
  explicit GpuBinEncoder(GpuEncodeSession &s)
      : BinEncIf(static_cast<const Common::BinProbModel_Std *>(nullptr)), session_(&s) {}

  // ---- BinEncIf (arith_codec.hpp:40-69) ----
  void init(Common::OutputBitstream *bitstream) override { bs_ = bitstream; }
  void uninit() override { bs_ = nullptr; }
  void start() override {  // arith_codec.cpp:329-337
    rec_.clear();
    tuSinks_.clear();
    numCtx_.assign(Common::Ctx::NumberOfContexts, 0);
    devCtx_.assign(Common::Ctx::NumberOfContexts, 0);
    numEP_ = numTrm_ = 0;
    writtenBits_ = 0;
    deviceBins_ = 0;
    completed_ = false;
  }
  void restart() override { start(); }  // arith_codec.cpp:359-365: same coder state as start()
  void reset(int qp, int initId) override {  // arith_codec.cpp:367-370
    Common::Ctx::init(qp, initId);
    rec_.params.qp = int8_t(qp);
    rec_.params.init_id = uint8_t(initId);
    ctxDirty_ = false;
    start();
  }
  void resetBits() override { start(); }
  uint64_t getEstFracBits() const override { THROW("not supported"); }  // as BinEncoderBase, arith_codec.cpp:380-383
  // BinCounter::getNumBins(ctxId), arith_codec.hpp:72-93.  Bins coded on the device (TU markers) are
  // included once the batch completed with GpuEncodeSession::countCtxBins set.
  unsigned getNumBins(unsigned ctxId) const override { return numCtx_[ctxId] + devCtx_[ctxId]; }
  // CABACWriter::getCtxInitId(slice) (cabac_writer.cpp:41-81) on this coder: the init table the next
  // inter slice should start from, evaluated by cabac_ctx_init_id from the final contexts and
  // getNumBins(k).  Returns the SliceType value (B_SLICE 0, P_SLICE 1, I_SLICE 2).
  int getCtxInitId(const Common::Slice &slice) {
    if (slice.isIntra()) return int(Common::I_SLICE);
    const auto &st = static_cast<const Common::CtxStore<Common::BinProbModel_Std> &>(
        static_cast<const Common::Ctx &>(*this));
    std::vector<cabac_ctx_model> ref(CABAC_NUM_CTX);
    std::vector<uint32_t> bins(CABAC_NUM_CTX);
    for (unsigned k = 0; k < CABAC_NUM_CTX; k++) {
      ref[k] = {st[k].getState0(), st[k].getState1()};
      bins[k] = getNumBins(k);
    }
    cabac_substream_params p{};
    p.qp = int8_t(slice.getSliceQp());
    p.init_id = 0;
    uint8_t choice = 0;
    Engine &eng = session_->engine();
    eng.check(cabac_ctx_init_id(eng.get(), &p, ref.data(), bins.data(), 1, &choice, nullptr), "cabac_ctx_init_id");
    return int(choice);
  }

  void encodeBin(unsigned bin, unsigned ctxId) override {
    rec_.encodeBin(bin, ctxId);
    numCtx_[ctxId]++;
  }
  void encodeBinEP(unsigned bin) override {
    rec_.encodeBinEP(bin);
    numEP_++;
  }
  void encodeBinsEP(unsigned bins, unsigned numBins) override {
    rec_.encodeBinsEP(bins, numBins);
    numEP_ += numBins;
  }
  void encodeRemAbsEP(unsigned bins, unsigned goRicePar, unsigned cutoff, int maxLog2TrDynamicRange) override {
    rec_.encodeRemAbsEP(bins, goRicePar, cutoff, maxLog2TrDynamicRange);
    numEP_ += remAbsLength(bins, goRicePar, cutoff, maxLog2TrDynamicRange);
  }
  void encodeBinTrm(unsigned bin) override {
    rec_.encodeBinTrm(bin);
    numTrm_++;
  }
  void align() override { rec_.align(); }
  // terminating bin + flush, i.e. the end_of_slice pair (cabac_writer.cpp:104-107)
  void finish() override {
    CHECK(!bs_, "GpuBinEncoder::finish without a bitstream");
    CHECK(rec_.cmds.empty() || rec_.cmds.back() != CABAC_MK_TRM(1),
          "GpuBinEncoder::finish() must follow encodeBinTrm(1) (end_of_slice)");
    rec_.cmds.pop_back();
    rec_.params.flags |= CABAC_SUB_TERMINATE | CABAC_SUB_ALIGN;  // the stop bit is peeled off again in complete()
    session_->submit(this, rec_, ctxDirty_);
    if (own_) own_->flush();
  }

  uint32_t getNumBins() override {  // BinCounter::getAll, arith_codec.cpp:302-308
    uint32_t n = numEP_ + numTrm_;
    for (uint32_t c : numCtx_) n += c;
    return n + deviceBins_;
  }
  bool isEncoding() override { return true; }
  // Exact only once the substream has been coded: the interval state lives on the device.
  unsigned getNumWrittenBits() override {
    CHECK(!completed_, "GpuBinEncoder::getNumWrittenBits before the batch was flushed");
    return writtenBits_;
  }
  void setBinStorage(bool) override {}
  const EntropyCoding::BinStore *getBinStore() const override { return nullptr; }
  EntropyCoding::BinEncIf *getTestBinEncoder() const override { return nullptr; }

  // ---- per-slice switches CABACWriter reads from the Slice (cabac_writer.cpp:2441, 2484) ----
  void setSliceTools(bool depQuant, bool signHiding, bool tsrcDisabled) {
    rec_.params.flags = uint8_t((depQuant ? CABAC_SUB_DEP_QUANT : 0) | (signHiding ? CABAC_SUB_SIGN_HIDING : 0) |
                                (tsrcDisabled ? CABAC_SUB_TSRC_DISABLED : 0));
  }
  // Start from handed-over contexts (WPP sync / Ctx::operator=): call after assigning to getCtx().
  void contextsAssigned() { ctxDirty_ = true; }

  // ---- CABACWriter::residual_coding(tu, compID, cuCtx) on the device ----
  // cuCtx (may be null) receives the LFNST / MTS side flags BEFORE this returns: the reference's
  // writer reads them right after the TUs of a CU (mts_idx, cabac_writer.cpp:2541-2543;
  // residual_lfnst_mode, :2601-2618), long before the batch runs.  They follow from the last
  // scan position and the significant coefficient groups alone, so the shim derives them from the
  // block it copies into the slab anyway, with the reference's own CoeffCodingContext for the scan;
  // the device reports the same bits per TU (tu_flags) and complete() cross-checks them.
  void residual_coding(const Common::TransformUnit &tu, Common::ComponentID compID, Common::CUCtx *cuCtx = nullptr) {
    if (compID == Common::COMPONENT_Cr && tu.jointCbCr == 3) return;  // cabac_writer.cpp:2428-2430
    refuseUnsupportedTools(tu);
    const Common::CompArea &blk = tu.blocks[compID];
    const uint8_t hostFlags = sideFlags(tu, compID);
    if (cuCtx) {  // cabac_writer.cpp:2460-2478, 2520-2523
      cuCtx->violatesLfnstConstrained[Common::toChannelType(compID)] |= (hostFlags & CABAC_TUR_VIOLATES_LFNST) != 0;
      cuCtx->lfnstLastScanPos |= (hostFlags & CABAC_TUR_LFNST_LAST_POS) != 0;
      cuCtx->mtsLastScanPos |= (hostFlags & CABAC_TUR_MTS_LAST_POS) != 0;
      cuCtx->violatesMtsCoeffConstraint |= (hostFlags & CABAC_TUR_VIOLATES_MTS) != 0;
    }
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
    std::vector<int32_t> &slab = session_->coeffs;
    d.coeff_offset = slab.size();
    const Common::TCoeff *src = tu.getCoeffs(compID).buf;  // row-major, stride == width (unit.hpp:551)
    slab.insert(slab.end(), src, src + size_t(blk.width) * blk.height);
    rec_.residual(d);
    tuSinks_.push_back({hostFlags});
  }

  // The CUCtx side effects of CABACWriter::residual_coding for one block, as CABAC_TUR_* bits.
  static uint8_t sideFlags(const Common::TransformUnit &tu, Common::ComponentID compID) {
    const bool ts = tu.mtsIdx[compID] == Common::MTS_SKIP;
    if (ts && !tu.cs->slice->getTSResidualCodingDisabledFlag()) return 0;  // residual_codingTS touches no CUCtx
    Common::CoeffCodingContext cctx(tu, compID, tu.cs->slice->getSignDataHidingEnabledFlag());
    const Common::TCoeff *coeff = tu.getCoeffs(compID).buf;
    const Common::CompArea &blk = tu.blocks[compID];
    const bool luma = Common::isLuma(compID);
    const bool sbtZero = tu.cs->sps->getUseMTS() && tu.cu->sbtInfo != 0 && blk.height <= 32 && blk.width <= 32 &&
                         compID == Common::COMPONENT_Y;
    int last = -1;
    bool farGroup = false;  // a significant, coded group beyond the 16 x 16 low-frequency corner
    for (int scanPos = 0; scanPos < int(cctx.maxNumCoeff()); scanPos++) {
      if (!coeff[cctx.blockPos(scanPos)]) continue;
      last = scanPos;
      const int cg = scanPos >> cctx.log2CGSize();
      if (luma && !farGroup) {
        cctx.initSubblock(cg, true);
        const bool zeroed = sbtZero && ((blk.height == 32 && cctx.cgPosY() >= (16 >> cctx.log2CGHeight())) ||
                                        (blk.width == 32 && cctx.cgPosX() >= (16 >> cctx.log2CGWidth())));
        farGroup = !zeroed && (cctx.cgPosY() > 3 || cctx.cgPosX() > 3);
      }
    }
    CHECK(last < 0, "Coefficient coding called for empty TU");
    uint8_t f = 0;
    if (!ts && blk.height >= 4 && blk.width >= 4) {
      const int maxLfnstPos = ((blk.height == 4 && blk.width == 4) || (blk.height == 8 && blk.width == 8)) ? 7 : 15;
      if (last > maxLfnstPos) f |= CABAC_TUR_VIOLATES_LFNST;
      if (last >= (luma ? Common::LFNST_LAST_SIG_LUMA : Common::LFNST_LAST_SIG_CHROMA)) f |= CABAC_TUR_LFNST_LAST_POS;
    }
    if (luma && !ts && last >= 1) f |= CABAC_TUR_MTS_LAST_POS;
    if (farGroup) f |= CABAC_TUR_VIOLATES_MTS;
    return f;
  }
  // Range-extension tools change the Rice derivation (persistent Rice adaptation, cabac_writer.cpp:2495-2503,
  // 2822-2850; RRC Rice extension, context_modelling.cpp:74-78; TSRC Rice, cabac_writer.cpp:2898-2920; extended
  // precision: maxLog2TrDynamicRange != 15).  The device path implements VVC v1: refuse instead of mis-coding.
  static void refuseUnsupportedTools(const Common::TransformUnit &tu) {
    const auto &rext = tu.cs->sps->getSpsRangeExtension();
    CHECK(rext.getPersistentRiceAdaptationEnabledFlag(), "B200 CABAC engine: persistent Rice adaptation is not supported");
    CHECK(rext.getRrcRiceExtensionEnableFlag(), "B200 CABAC engine: RRC Rice extension is not supported");
    CHECK(rext.getExtendedPrecisionProcessingFlag(), "B200 CABAC engine: extended precision processing is not supported");
    CHECK(rext.getTSRCRicePresentFlag() && tu.cs->slice->get_tsrc_index() != 0,
          "B200 CABAC engine: TSRC Rice index != 0 is not supported");
  }

 private:
  friend class GpuEncodeSession;
  static unsigned remAbsLength(unsigned bins, unsigned rice, unsigned cutoff, int maxLog2) {
    if (bins < (cutoff << rice)) return (bins >> rice) + 1 + rice;
    const unsigned maxPrefix = 32 - cutoff - unsigned(maxLog2);
    const unsigned code = (bins >> rice) - cutoff;
    unsigned prefix = 0;
    if (code >= ((1u << maxPrefix) - 1)) return maxPrefix + cutoff + unsigned(maxLog2);
    while (code > ((2u << prefix) - 2u)) prefix++;
    return prefix + cutoff + prefix + rice + 1;
  }
  void complete(const SubstreamResult &r, const uint8_t *tuFlags);

  struct TuSink {
    uint8_t hostFlags;  // sideFlags() at record time; the device must report the same
  };
  std::unique_ptr<GpuEncodeSession> own_;
  GpuEncodeSession *session_;
  Common::OutputBitstream *bs_ = nullptr;
  StreamRecorder rec_;
  std::vector<TuSink> tuSinks_;
  std::vector<uint32_t> numCtx_ = std::vector<uint32_t>(Common::Ctx::NumberOfContexts, 0);
  std::vector<uint32_t> devCtx_ = std::vector<uint32_t>(Common::Ctx::NumberOfContexts, 0);
  uint32_t numEP_ = 0, numTrm_ = 0, deviceBins_ = 0;
  unsigned writtenBits_ = 0;
  bool ctxDirty_ = false, completed_ = false;
};

inline void GpuEncodeSession::submit(GpuBinEncoder *enc, const StreamRecorder &rec, bool ctxIn) {
  std::vector<cabac_ctx_model> models;
  if (ctxIn) {  // export the encoder's Ctx (contexts.hpp:184-186 exist for exactly this bridging)
    const auto &st = static_cast<const Common::CtxStore<Common::BinProbModel_Std> &>(
        static_cast<const Common::Ctx &>(*enc));
    models.resize(CABAC_NUM_CTX);
    for (unsigned k = 0; k < CABAC_NUM_CTX; k++) models[k] = {st[k].getState0(), st[k].getState1()};
  }
  batch_.add(rec, ctxIn ? models.data() : nullptr);
  owners_.push_back(enc);
}

inline void GpuEncodeSession::flush() {
  if (owners_.empty()) return;
  static const int32_t none = 0;
  batch_.flush(eng_, coeffs.empty() ? &none : coeffs.data(), coeffs.size(), /*wantCtx=*/true, countCtxBins);
  const uint8_t *tf = batch_.tuFlags().data();
  for (size_t i = 0; i < owners_.size(); i++) {
    owners_[i]->complete(batch_.result(i), tf);
    tf += owners_[i]->tuSinks_.size();
  }
  owners_.clear();
  batch_.clear();
  coeffs.clear();
}

inline void GpuBinEncoder::complete(const SubstreamResult &r, const uint8_t *tuFlags) {
  // status bits -> the exceptions the reference throws at the same points (type_def.hpp:295-329)
  CHECK(r.status & CABAC_SS_EMPTY_TU, "Coefficient coding called for empty TU");
  CHECK(r.status != 0, "B200 CABAC engine reported substream status " + std::to_string(r.status));
  // The device appended writeByteAlignment(); finish() itself must leave the bits after the
  // terminating bin un-aligned in the OutputBitstream (arith_codec.cpp:356), so peel the stop
  // bit off again: the caller adds it with writeByteAlignment() (EncSlice.cpp:2114).
  CHECK(r.size == 0, "empty substream");
  for (uint32_t k = 0; k + 1 < r.size; k++) bs_->write(r.bytes[k], 8);
  const unsigned last = r.bytes[r.size - 1];
  const unsigned zeros = unsigned(__builtin_ctz(last | 0x100u));
  if (zeros < 7) bs_->write(last >> (zeros + 1), 7 - zeros);
  writtenBits_ = 8 * (r.size - 1) + (7 - zeros);
  uint32_t host = numEP_ + numTrm_;
  for (uint32_t c : numCtx_) host += c;
  deviceBins_ = r.numBins - host;
  if (r.ctxBins) devCtx_.assign(r.ctxBins, r.ctxBins + CABAC_NUM_CTX);
  if (r.ctx) {
    auto &st = static_cast<Common::CtxStore<Common::BinProbModel_Std> &>(static_cast<Common::Ctx &>(*this));
    for (unsigned k = 0; k < CABAC_NUM_CTX; k++)
      st[k] = Common::BinProbModel_Std(r.ctx[k].state0, r.ctx[k].state1, st[k].getRate());
  }
  constexpr uint8_t side = CABAC_TUR_VIOLATES_LFNST | CABAC_TUR_LFNST_LAST_POS | CABAC_TUR_MTS_LAST_POS | CABAC_TUR_VIOLATES_MTS;
  for (size_t t = 0; t < tuSinks_.size(); t++)  // the CUCtx was filled at record time; the device agrees or we stop
    CHECK((tuFlags[t] & side) != tuSinks_[t].hostFlags,
          "CUCtx side flags of TU " + std::to_string(t) + ": device " + std::to_string(tuFlags[t] & side) + " vs host " +
              std::to_string(tuSinks_[t].hostFlags));
  completed_ = true;
}

}  // namespace cabac_b200
