// oracle/ref_harness.hpp - TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// The objects CABACWriter::residual_coding / CABACReader::residual_coding read besides the
// TU itself (SURVEY.md section 8c "Harness objects"), built from the flat descriptors of
// include/cabac_b200.h.  Shared by ref_driver.cpp (the compiled-reference checker) and
// tests/dropin/dropin_main.cpp (the reference's own writer / reader running on the shims).
#pragma once
#include <array>
#include <memory>
#include <mutex>

#include "arith_codec.hpp"
#include "cabac_reader.hpp"
#include "cabac_writer.hpp"
#include "coding_structure.hpp"
#include "context_modelling.hpp"
#include "rom.hpp"
#include "slice.hpp"
#include "unit.hpp"

#include "cabac_b200.h"

namespace ref_harness {
using namespace Common;
using namespace EntropyCoding;

inline std::once_flag g_romOnce;
inline void ensureRom() { std::call_once(g_romOnce, [] { initROM(); }); }

// Everything residual_coding reads besides the TU itself
// (SURVEY.md section 8c "Harness objects").
struct Harness {
  std::shared_ptr<SPS> sps;
  std::shared_ptr<Slice> slice;
  std::shared_ptr<PreCalcValues> pcv;
  std::unique_ptr<CodingStructure> cs;

  Harness(const cabac_substream_params &p, bool useMTS) {
    ensureRom();
    SPSRExt rext(false, false, false, false);
    BitDepths bd;
    bd.recon[0] = bd.recon[1] = 10;
    int qpBdOffset[2] = {12, 12};
    sps = std::make_shared<SPS>(
        false, false, /*SBT*/ true, false, CHROMA_420, 2, 128u, 128u,
        /*transformSkipEnabled*/ true, /*log2MaxTSSize*/ 5, true, true, bd,
        false, qpBdOffset, 6u, false, rext, false, false, 0u, false, 0u, false,
        false, /*MTS*/ useMTS, useMTS, useMTS, false, false, false, false,
        false, false, false, false, 6u, 6u, 5u);
    bool sao[2] = {false, false};
    int numRef[2] = {0, 0};
    int symRef[2] = {0, 0};
    WeightPredTable wpt{};
    std::array<APS *, ALF_CTB_MAX_NUM_APS> apss;
    apss.fill(nullptr);
    bool alf[3] = {false, false, false};
    unsigned riceBit[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    CcAlfFilterParam ccalf{};
    uint8_t *ccctl[2] = {nullptr, nullptr};
    SliceType st = p.init_id == 2 ? I_SLICE : (p.init_id == 1 ? P_SLICE : B_SLICE);
    slice = std::make_shared<Slice>(
        sao, st, int(p.qp), false, (p.flags & CABAC_SUB_DEP_QUANT) != 0, 4,
        (p.flags & CABAC_SUB_SIGN_HIDING) != 0,
        (p.flags & CABAC_SUB_TSRC_DISABLED) != 0, numRef, false, false, symRef,
        0u, wpt, false, st, apss, alf, 0, 0, 0, riceBit, ccalf, ccctl);
    slice->setSPS(sps.get());
    unsigned z3[3] = {0, 0, 0};
    pcv = std::make_shared<PreCalcValues>(CHROMA_420, false, 128u, 128u, 127u,
                                          127u, 7u, 7u, 1u, 1u, false, false,
                                          z3, z3, z3, z3, z3, z3);
    std::array<UnitScale, MAX_NUM_COMPONENT> us;
    PLTBuf plt{};
    unsigned *idx2[2] = {nullptr, nullptr};
    TCoeff *c3[3] = {nullptr, nullptr, nullptr};
    Pel *p3[3] = {nullptr, nullptr, nullptr};
    bool *r2[2] = {nullptr, nullptr};
    int offs[3] = {0, 0, 0};
    cs.reset(new CodingStructure(UnitArea(CHROMA_420, Area(0, 0, 128, 128)), us,
                                 0, TREE_D, MODE_TYPE_ALL, plt, false, idx2,
                                 idx2, idx2, 0u, 0u, 0u, c3, p3, r2, offs));
    cs->sps = sps;
    cs->slice = slice;
    cs->pcv = pcv;
  }
};

struct TuPair {
  CodingUnit cu;
  TransformUnit tu;
};

// Builds the (CodingUnit, TransformUnit) pair a descriptor stands for.
inline void makeTu(TuPair &tp, Harness &hs, const cabac_tu_desc &d, TCoeff *coeffBase) {
  const unsigned w = 1u << d.log2_w, h = 1u << d.log2_h;
  const ComponentID comp = ComponentID(d.comp_id);
  CodingUnit &cu = tp.cu;
  cu.cs = hs.cs.get();
  cu.slice = hs.slice.get();
  cu.chType = toChannelType(comp);
  cu.predMode = MODE_INTRA;
  cu.sbtInfo = (d.flags & CABAC_TU_SBT_ZEROOUT) ? 1 : 0;
  cu.ispMode = 0;
  cu.bdpcmMode = (comp == COMPONENT_Y && (d.flags & CABAC_TU_BDPCM)) ? 1 : 0;
  cu.bdpcmModeChroma = (comp != COMPONENT_Y && (d.flags & CABAC_TU_BDPCM)) ? 1 : 0;
  cu.colorTransform = false;
  cu.lfnstIdx = 0;
  TransformUnit &tu = tp.tu;
  // every block gets the coded block's size; only blocks[comp] is read
  UnitBlocksType blocks;
  for (int c = 0; c < 3; c++)
    blocks.push_back(CompArea(ComponentID(c), CHROMA_420, 0, 0, w, h));
  static_cast<UnitArea &>(tu) = UnitArea(CHROMA_420, blocks);
  tu.cu = &cu;
  tu.cs = hs.cs.get();
  tu.chType = toChannelType(comp);
  tu.initData();
  tu.mtsIdx[comp] = (d.flags & CABAC_TU_TS) ? MTS_SKIP : MTS_DCT2_DCT2;
  tu.jointCbCr = 0;
  tu.next = tu.prev = nullptr;
  TCoeff *c3[3] = {nullptr, nullptr, nullptr};
  Pel *p3[3] = {nullptr, nullptr, nullptr};
  bool *r2[2] = {nullptr, nullptr};
  c3[comp] = coeffBase + d.coeff_offset;
  tu.init(c3, p3, r2);
}

// TU::isTSAllowed is evaluated by the reference itself; a descriptor that says
// "not allowed" is realised through the one input that only disables the flag:
// ispMode for luma, sbtInfo otherwise (both make isTSAllowed() false without
// touching residual coding when SBT zero-out / MTS are off).
inline void applyTsAllowed(TuPair &tp, const cabac_tu_desc &d, bool useMTS) {
  const bool want = (d.flags & CABAC_TU_TS_ALLOWED) != 0;
  if (want) return;
  if (d.comp_id == 0) {
    tp.cu.ispMode = 1;  // HOR_INTRA_SUBPARTITIONS: only read by isTSAllowed here
  } else if (!useMTS) {
    tp.cu.sbtInfo = 1;  // without MTS sbtInfo has no other effect on this path
  } else {
    tp.cu.sbtInfo = 1;  // chroma: SBT zero-out applies to COMPONENT_Y only
  }
}

}  // namespace ref_harness
