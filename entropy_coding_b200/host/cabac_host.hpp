// cabac_host.hpp - C++17 host layer above the C ABI (include/cabac_b200.h).
//
// Independent of the reference's headers: it records the calls a BinEncIf user
// makes (arith_codec.hpp:40-69) as command words, collects many substreams into one
// batch and hands the batch to the GPU.  The reference-facing adapters that derive
// from EntropyCoding::BinEncIf live in gpu_bin_encoder.hpp.
//
// No CPU fallback: every coding call goes to libcabac_b200.so; a failure throws.
#pragma once
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "cabac_b200.h"

namespace cabac_b200 {

struct Error : std::runtime_error {
  int code;
  Error(int c, const std::string &what) : std::runtime_error(what), code(c) {}
};

// RAII handle; one CUDA stream per engine (thread-compatible like the reference's coder objects).
class Engine {
 public:
  explicit Engine(int device = -1) : h_(cabac_create(device)) {
    if (!h_) throw Error(CABAC_ERR_CUDA, "cabac_create failed: no CUDA device (there is no CPU fallback)");
  }
  ~Engine() { cabac_destroy(h_); }
  Engine(const Engine &) = delete;
  Engine &operator=(const Engine &) = delete;
  cabac_handle *get() const { return h_; }
  void check(int rc, const char *what) const {
    if (rc != CABAC_OK && rc != CABAC_ERR_SUBSTREAM) throw Error(rc, std::string(what) + ": " + cabac_last_error(h_));
  }

 private:
  cabac_handle *h_;
};

// One substream as its syntax writer produced it: BinEncIf calls as command words, with a
// CABAC_CMD_TU marker wherever residual_coding() is delegated to the device.
class StreamRecorder {
 public:
  std::vector<uint32_t> cmds;
  std::vector<cabac_tu_desc> tus;
  cabac_substream_params params{32, 2, 0, 0};

  void clear() {
    cmds.clear();
    tus.clear();
  }
  void encodeBin(unsigned bin, unsigned ctxId) { cmds.push_back(CABAC_MK_CTX(ctxId, bin)); }
  void encodeBinEP(unsigned bin) { cmds.push_back(CABAC_MK_EP(bin & 1u, 1)); }
  // encodeBinsEP splits into <= 8-bin steps itself (arith_codec.cpp:401-424); any split is
  // arithmetically identical, so runs longer than a command word's 24 bins are cut MSB first.
  void encodeBinsEP(unsigned bins, unsigned numBins) {
    while (numBins > 24) {
      numBins -= 24;
      cmds.push_back(CABAC_MK_EP((bins >> numBins) & 0xFFFFFFu, 24));
    }
    if (numBins) cmds.push_back(CABAC_MK_EP(bins & ((1u << numBins) - 1u), numBins));
  }
  // encodeRemAbsEP(.., cutoff, maxLog2TrDynamicRange): the device command implements the
  // (5, 15) case every VVC v1 call site uses; other parameters are expanded here into
  // EP runs exactly as arith_codec.cpp:426-458 does.
  void encodeRemAbsEP(unsigned bins, unsigned rice, unsigned cutoff, int maxLog2TrDR) {
    if (cutoff == 5 && maxLog2TrDR == 15 && bins < (1u << 24) && rice < 32) {
      cmds.push_back(CABAC_MK_REM(bins, rice));
      return;
    }
    const unsigned threshold = cutoff << rice;
    if (bins < threshold) {
      const unsigned len = (bins >> rice) + 1;
      encodeBinsEP((1u << len) - 2, len);
      encodeBinsEP(bins & ((1u << rice) - 1), rice);
      return;
    }
    const unsigned maxPrefix = 32 - cutoff - unsigned(maxLog2TrDR);
    unsigned prefix = 0, code = (bins >> rice) - cutoff, suffix;
    if (code >= ((1u << maxPrefix) - 1)) {
      prefix = maxPrefix;
      suffix = unsigned(maxLog2TrDR);
    } else {
      while (code > ((2u << prefix) - 2)) prefix++;
      suffix = prefix + rice + 1;
    }
    const unsigned total = prefix + cutoff;
    const unsigned bitMask = (1u << rice) - 1;
    encodeBinsEP((1u << total) - 1, total);
    encodeBinsEP(((code - ((1u << prefix) - 1)) << rice) | (bins & bitMask), suffix);
  }
  void encodeBinTrm(unsigned bin) { cmds.push_back(CABAC_MK_TRM(bin)); }
  void align() { cmds.push_back(CABAC_MK_ALIGN()); }
  // residual_coding(tu, compID) delegated to the device; `coeff_offset` indexes the batch slab.
  void residual(const cabac_tu_desc &d) {
    cmds.push_back(CABAC_MK_TU());
    tus.push_back(d);
  }
};

struct SubstreamResult {
  const uint8_t *bytes = nullptr;
  uint32_t size = 0, numBins = 0, status = 0;
  const cabac_ctx_model *ctx = nullptr;  // CABAC_NUM_CTX models after the run
  const uint32_t *ctxBins = nullptr;     // CABAC_NUM_CTX context-coded bin counts of the TU markers (on request)
};

// A batch of recorded substreams -> one cabac_encode_stream call.  The coefficient slab is
// caller-owned (TransformUnit::m_coeffs style, unit.hpp:551); descriptors index into it.
class EncodeBatch {
 public:
  // Returns the substream's index in the batch.  ctxIn (CABAC_NUM_CTX models) is copied when
  // the substream starts from handed-over contexts (WPP sync, CABAC_SUB_CTX_IN).
  size_t add(const StreamRecorder &s, const cabac_ctx_model *ctxIn = nullptr) {
    const size_t i = params_.size();
    cmds_.insert(cmds_.end(), s.cmds.begin(), s.cmds.end());
    tus_.insert(tus_.end(), s.tus.begin(), s.tus.end());
    cmdOff_.push_back(cmds_.size());
    tuOff_.push_back(tus_.size());
    cabac_substream_params p = s.params;
    if (ctxIn) p.flags |= CABAC_SUB_CTX_IN;
    params_.push_back(p);
    if (ctxIn) {
      ctxIn_.resize((i + 1) * CABAC_NUM_CTX);
      std::memcpy(&ctxIn_[i * CABAC_NUM_CTX], ctxIn, CABAC_NUM_CTX * sizeof(cabac_ctx_model));
      haveCtxIn_ = true;
    }
    // capacity bound: an EP bin costs exactly one bit; a context-coded bin at most
    // log2(510/4) < 7 bits; a TU at most ~5 bytes per coefficient position
    size_t bits = 64;
    for (uint32_t w : s.cmds) {
      const uint32_t k = CABAC_CMD_KIND(w);
      bits += k == CABAC_CMD_EP ? ((w >> 24) & 31u) : (k == CABAC_CMD_REM ? 32u : 8u);
    }
    size_t cap = bits / 8 + 16;
    for (const cabac_tu_desc &d : s.tus) cap += (size_t(5) << (d.log2_w + d.log2_h)) + 16;
    outOff_.push_back(outOff_.back() + ((cap + 15) & ~size_t(15)));
    return i;
  }
  size_t size() const { return params_.size(); }

  // Runs the batch.  Throws Error on a CUDA / argument failure; per-substream problems are
  // reported in SubstreamResult::status (CABAC_SS_*), as the C ABI does.
  // wantCtxBins: also fetch the per-context bin counts of the TU markers (BinCounter view of the
  // device-side residual syntax, cabac_residual_ctx_bins) - what getNumBins(ctxId) needs for
  // CABACWriter::getCtxInitId (cabac_writer.cpp:41-81).  The counts follow from the coefficients alone.
  void flush(Engine &eng, const int32_t *coeffs, size_t nCoeffs, bool wantCtx = false, bool wantCtxBins = false) {
    const size_t n = params_.size();
    ctxBins_.clear();
    out_.assign(outOff_.back(), 0);
    sizes_.assign(n, 0);
    bins_.assign(n, 0);
    status_.assign(n, 0);
    tuFlags_.assign(tus_.size(), 0);
    if (wantCtx) ctxOut_.assign(n * CABAC_NUM_CTX, cabac_ctx_model{0, 0});
    if (haveCtxIn_) ctxIn_.resize(n * CABAC_NUM_CTX);
    if (!n) return;
    const int rc = cabac_encode_stream(eng.get(), cmds_.data(), cmdOff_.data(), tus_.empty() ? nullptr : tus_.data(),
                                       tus_.empty() ? nullptr : tuOff_.data(), params_.data(), n, coeffs, nCoeffs,
                                       haveCtxIn_ ? ctxIn_.data() : nullptr, out_.data(), outOff_.data(),
                                       sizes_.data(), bins_.data(), tus_.empty() ? nullptr : tuFlags_.data(),
                                       wantCtx ? ctxOut_.data() : nullptr, status_.data());
    eng.check(rc, "cabac_encode_stream");
    if (wantCtxBins && !tus_.empty()) {
      ctxBins_.assign(n * CABAC_NUM_CTX, 0);
      const int rb = cabac_residual_ctx_bins(eng.get(), tus_.data(), tuOff_.data(), params_.data(), n, coeffs, nCoeffs,
                                             nullptr, ctxBins_.data(), nullptr, nullptr, nullptr);
      eng.check(rb, "cabac_residual_ctx_bins");
    }
  }
  SubstreamResult result(size_t i) const {
    SubstreamResult r;
    r.bytes = out_.data() + outOff_[i];
    r.size = sizes_[i];
    r.numBins = bins_[i];
    r.status = status_[i];
    r.ctx = ctxOut_.empty() ? nullptr : &ctxOut_[i * CABAC_NUM_CTX];
    r.ctxBins = ctxBins_.empty() ? nullptr : &ctxBins_[i * CABAC_NUM_CTX];
    return r;
  }
  const std::vector<uint8_t> &tuFlags() const { return tuFlags_; }
  void clear() { *this = EncodeBatch(); }

 private:
  std::vector<uint32_t> cmds_;
  std::vector<cabac_tu_desc> tus_;
  std::vector<uint64_t> cmdOff_{0}, tuOff_{0}, outOff_{0};
  std::vector<cabac_substream_params> params_;
  std::vector<cabac_ctx_model> ctxIn_, ctxOut_;
  bool haveCtxIn_ = false;
  std::vector<uint8_t> out_, tuFlags_;
  std::vector<uint32_t> sizes_, bins_, status_, ctxBins_;
};

// Residual decode of a batch of substreams: bytes + TU lists in, coefficients out
// (CABACReader::residual_coding, cabac_reader.hpp:111, at batch granularity).
class DecodeBatch {
 public:
  size_t add(const uint8_t *bytes, size_t nbytes, const cabac_substream_params &p, const cabac_tu_desc *tus,
             size_t ntu) {
    in_.insert(in_.end(), bytes, bytes + nbytes);
    inOff_.push_back(in_.size());
    tus_.insert(tus_.end(), tus, tus + ntu);
    tuOff_.push_back(tus_.size());
    params_.push_back(p);
    return params_.size() - 1;
  }
  size_t size() const { return params_.size(); }
  // coeffs: caller-owned slab of nCoeffs TCoeff the descriptors index; fully overwritten.
  void flush(Engine &eng, int32_t *coeffs, size_t nCoeffs, bool wantCtx = false) {
    const size_t n = params_.size();
    bins_.assign(n, 0);
    bitsRead_.assign(n, 0);
    status_.assign(n, 0);
    tuFlags_.assign(tus_.size(), 0);
    if (wantCtx) ctxOut_.assign(n * CABAC_NUM_CTX, cabac_ctx_model{0, 0});
    if (!n) return;
    if (in_.empty()) in_.push_back(0);
    const int rc = cabac_decode_residual(eng.get(), in_.data(), inOff_.data(), tus_.data(), tuOff_.data(),
                                         params_.data(), n, nullptr, coeffs, nCoeffs, bins_.data(), bitsRead_.data(),
                                         tuFlags_.data(), wantCtx ? ctxOut_.data() : nullptr, status_.data());
    eng.check(rc, "cabac_decode_residual");
  }
  uint32_t numBins(size_t i) const { return bins_[i]; }
  uint32_t bitsRead(size_t i) const { return bitsRead_[i]; }
  uint32_t status(size_t i) const { return status_[i]; }
  const cabac_ctx_model *ctx(size_t i) const { return ctxOut_.empty() ? nullptr : &ctxOut_[i * CABAC_NUM_CTX]; }
  const std::vector<uint8_t> &tuFlags() const { return tuFlags_; }

 private:
  std::vector<uint8_t> in_, tuFlags_;
  std::vector<uint64_t> inOff_{0}, tuOff_{0};
  std::vector<cabac_tu_desc> tus_;
  std::vector<cabac_substream_params> params_;
  std::vector<cabac_ctx_model> ctxOut_;
  std::vector<uint32_t> bins_, bitsRead_, status_;
};

// RDO bit estimation of a batch of candidates (CABACWriter::residual_coding on a BitEstimator_Std,
// arith_codec.cpp:603-711): every candidate is a TU list priced from its own starting contexts
// (`ctxIn`: the Ctx snapshot the encoder's RDO loop restores between candidates, or null for a fresh
// Ctx::init(qp, initId)).  One launch prices them all.
class EstimateBatch {
 public:
  size_t add(const cabac_substream_params &p, const cabac_tu_desc *tus, size_t ntu,
             const cabac_ctx_model *ctxIn = nullptr) {
    const size_t i = params_.size();
    tus_.insert(tus_.end(), tus, tus + ntu);
    tuOff_.push_back(tus_.size());
    params_.push_back(p);
    if (ctxIn) {
      haveCtxIn_ = true;
      ctxIn_.resize((i + 1) * CABAC_NUM_CTX, cabac_ctx_model{0, 0});
      std::copy(ctxIn, ctxIn + CABAC_NUM_CTX, ctxIn_.begin() + i * CABAC_NUM_CTX);
      params_.back().flags |= CABAC_SUB_CTX_IN;
    }
    return i;
  }
  size_t size() const { return params_.size(); }
  void flush(Engine &eng, const int32_t *coeffs, size_t nCoeffs, bool wantCtx = false) {
    const size_t n = params_.size();
    est_.assign(n, 0);
    perTu_.assign(tus_.size(), 0);
    status_.assign(n, 0);
    if (wantCtx) ctxOut_.assign(n * CABAC_NUM_CTX, cabac_ctx_model{0, 0});
    if (!n) return;
    if (haveCtxIn_) ctxIn_.resize(n * CABAC_NUM_CTX, cabac_ctx_model{0, 0});
    const int rc = cabac_estimate_residual_bits(eng.get(), tus_.data(), tuOff_.data(), params_.data(), n, coeffs, nCoeffs,
                                                haveCtxIn_ ? ctxIn_.data() : nullptr, est_.data(), perTu_.data(),
                                                nullptr, nullptr, wantCtx ? ctxOut_.data() : nullptr, status_.data());
    eng.check(rc, "cabac_estimate_residual_bits");
  }
  // BitEstimatorBase::getEstFracBits() after the candidate's last TU / after its k-th TU (1/32768 bit)
  uint64_t fracBits(size_t i) const { return est_[i]; }
  uint64_t fracBitsAfterTu(size_t i, size_t k) const { return perTu_[tuOff_[i] + k]; }
  uint32_t status(size_t i) const { return status_[i]; }
  const cabac_ctx_model *ctx(size_t i) const { return ctxOut_.empty() ? nullptr : &ctxOut_[i * CABAC_NUM_CTX]; }
  void clear() { *this = EstimateBatch(); }

 private:
  std::vector<uint64_t> tuOff_{0}, est_, perTu_;
  std::vector<cabac_tu_desc> tus_;
  std::vector<cabac_substream_params> params_;
  std::vector<cabac_ctx_model> ctxIn_, ctxOut_;
  std::vector<uint32_t> status_;
  bool haveCtxIn_ = false;
};

}  // namespace cabac_b200
