// gpu_residual_reader.hpp - decode-side shim: the batch counterpart of
// CABACReader::residual_coding(TransformUnit&, ComponentID, CUCtx&) (cabac_reader.hpp:111).
//
// Every decoded bin decides the next context, so the decode seam is the residual level
// (SURVEY.md section 8b): a substream = its bytes + the list of TUs to parse, in order.
// The caller queues TUs per substream; flush() parses all substreams on the device and
// writes the coefficients into the TUs' own buffers (TransformUnit::getCoeffs, unit.hpp:551),
// sets tu.mtsIdx from the parsed transform_skip_flag and the CUCtx side flags.
#pragma once
#include "cabac_host.hpp"
#include "coding_structure.hpp"
#include "context_modelling.hpp"
#include "slice.hpp"
#include "unit.hpp"
#include "unit_tools.hpp"

namespace cabac_b200 {
using Common::Exception;  // CHECK / THROW (type_def.hpp:295-329) name it unqualified

class GpuResidualReader {
 public:
  explicit GpuResidualReader(Engine &eng) : eng_(eng) {}

  // Opens a substream: its payload (as InputBitstream::getFifo holds it, bit_stream.hpp), the slice
  // QP / type (CABACReader::initCtxModels, cabac_reader.cpp:34-60) and the slice-level tools.
  size_t beginSubstream(const uint8_t *bytes, size_t nbytes, int qp, int initId, bool depQuant, bool signHiding,
                        bool tsrcDisabled, bool expectTerminate) {
    Sub s;
    s.bytes.assign(bytes, bytes + nbytes);
    s.p.qp = int8_t(qp);
    s.p.init_id = uint8_t(initId);
    s.p.flags = uint8_t((depQuant ? CABAC_SUB_DEP_QUANT : 0) | (signHiding ? CABAC_SUB_SIGN_HIDING : 0) |
                        (tsrcDisabled ? CABAC_SUB_TSRC_DISABLED : 0) | (expectTerminate ? CABAC_SUB_TERMINATE : 0));
    subs_.push_back(std::move(s));
    return subs_.size() - 1;
  }
  // Queues residual_coding(tu, compID, cuCtx) on the current substream.
  void residual_coding(Common::TransformUnit &tu, Common::ComponentID compID, Common::CUCtx *cuCtx = nullptr) {
    if (compID == Common::COMPONENT_Cr && tu.jointCbCr == 3) return;  // cabac_reader.cpp:2651-2653
    {  // VVC v1 residual coding only: refuse the range-extension Rice tools instead of mis-parsing
      const auto &rext = tu.cs->sps->getSpsRangeExtension();
      CHECK(rext.getPersistentRiceAdaptationEnabledFlag() || rext.getRrcRiceExtensionEnableFlag() ||
                rext.getExtendedPrecisionProcessingFlag() ||
                (rext.getTSRCRicePresentFlag() && tu.cs->slice->get_tsrc_index() != 0),
            "B200 CABAC engine: range-extension residual coding tools are not supported");
    }
    const Common::CompArea &blk = tu.blocks[compID];
    const Common::CodingUnit &cu = *tu.cu;
    cabac_tu_desc d{};
    d.log2_w = uint8_t(Common::floorLog2(blk.width));
    d.log2_h = uint8_t(Common::floorLog2(blk.height));
    d.comp_id = uint8_t(compID);
    const bool bdpcm = Common::isLuma(compID) ? cu.bdpcmMode != 0 : cu.bdpcmModeChroma != 0;
    // ts_flag (cabac_reader.cpp:2737-2752): parsed when allowed, else inferred (BDPCM implies skip)
    d.flags = uint8_t((Common::TU::isTSAllowed(tu, compID) ? CABAC_TU_TS_ALLOWED : 0) | (bdpcm ? CABAC_TU_BDPCM : 0) |
                      ((bdpcm || tu.mtsIdx[compID] == Common::MTS_SKIP) ? CABAC_TU_TS : 0) |
                      ((tu.cs->sps->getUseMTS() && cu.sbtInfo != 0 && compID == Common::COMPONENT_Y)
                           ? CABAC_TU_SBT_ZEROOUT : 0));
    d.coeff_offset = nCoeffs_;
    nCoeffs_ += size_t(blk.width) * blk.height;
    Sub &s = subs_.back();
    s.tus.push_back(d);
    s.sinks.push_back({&tu, cuCtx, int(compID)});
  }
  // Parses everything queued.  Throws Common::Exception where the reference would
  // (bit_stream.cpp:269, cabac_reader.cpp:2942-2946, arith_codec.cpp:68-73).
  void flush() {
    DecodeBatch b;
    for (Sub &s : subs_) b.add(s.bytes.data(), s.bytes.size(), s.p, s.tus.data(), s.tus.size());
    std::vector<int32_t> coeffs(nCoeffs_ ? nCoeffs_ : 1);
    b.flush(eng_, coeffs.data(), nCoeffs_);
    const uint8_t *tf = b.tuFlags().data();
    for (size_t i = 0; i < subs_.size(); i++) {
      const uint32_t st = b.status(i);
      CHECK(st & CABAC_SS_IN_OVERRUN, "Exceeded FIFO size");
      CHECK(st & CABAC_SS_COEFF_RANGE, "TransCoeffLevel outside allowable range");
      CHECK(st & CABAC_SS_NO_TERMINATE, "Expecting a terminating bit");
      CHECK(st != 0, "B200 CABAC engine reported substream status " + std::to_string(st));
      Sub &s = subs_[i];
      for (size_t t = 0; t < s.sinks.size(); t++, tf++) {
        const Sink &k = s.sinks[t];
        const Common::ComponentID comp = Common::ComponentID(k.comp);
        const Common::CompArea &blk = k.tu->blocks[comp];
        Common::TCoeff *dst = k.tu->getCoeffs(comp).buf;
        std::memcpy(dst, &coeffs[s.tus[t].coeff_offset], size_t(blk.width) * blk.height * sizeof(int32_t));
        k.tu->mtsIdx[comp] = (*tf & CABAC_TUR_TS) ? Common::MTS_SKIP : Common::MTS_DCT2_DCT2;
        if (Common::CUCtx *c = k.cuCtx) {  // cabac_reader.cpp:2672-2690, 2730-2733
          const Common::ChannelType ch = Common::toChannelType(comp);
          c->violatesLfnstConstrained[ch] |= (*tf & CABAC_TUR_VIOLATES_LFNST) != 0;
          c->lfnstLastScanPos |= (*tf & CABAC_TUR_LFNST_LAST_POS) != 0;
          if (k.comp == 0) c->mtsLastScanPos |= (*tf & CABAC_TUR_MTS_LAST_POS) != 0;
          c->violatesMtsCoeffConstraint |= (*tf & CABAC_TUR_VIOLATES_MTS) != 0;
        }
      }
      s.numBins = b.numBins(i);
      s.bitsRead = b.bitsRead(i);
    }
    done_ = std::move(subs_);
    subs_.clear();
    nCoeffs_ = 0;
  }
  uint32_t numBins(size_t i) const { return done_[i].numBins; }
  uint32_t bitsRead(size_t i) const { return done_[i].bitsRead; }  // BinDecoderBase::getNumBitsRead, arith_codec.cpp:201-203

 private:
  struct Sink {
    Common::TransformUnit *tu;
    Common::CUCtx *cuCtx;
    int comp;
  };
  struct Sub {
    std::vector<uint8_t> bytes;
    cabac_substream_params p{};
    std::vector<cabac_tu_desc> tus;
    std::vector<Sink> sinks;
    uint32_t numBins = 0, bitsRead = 0;
  };
  Engine &eng_;
  std::vector<Sub> subs_, done_;
  size_t nCoeffs_ = 0;
};

}  // namespace cabac_b200
