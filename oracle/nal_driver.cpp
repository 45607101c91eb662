// oracle/nal_driver.cpp - TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// C entry points over VTM's own NAL byte-stream code, compiled UNMODIFIED from where it lies in the
// reference tree (oracle/Makefile target `nalref`, output oracle/_ref/libnal_ref.so):
//   VVCSoftware_VTM/source/Lib/EncoderLib/NALwrite.cpp   writeNaluContent   (emulation prevention insertion)
//   VVCSoftware_VTM/source/Lib/DecoderLib/NALread.cpp    read / convertPayloadToRBSP (removal + EPB locations)
//   VVCSoftware_VTM/source/Lib/CommonLib/BitStream.cpp
// Pins the restatements orc_insert_emulation_prevention / orc_remove_emulation_prevention
// (oracle/cabac_oracle.c) and through them the device kernels.
#include <cstring>
#include <sstream>

#include "CommonLib/CommonDef.h"
#include "DecoderLib/NALread.h"
#include "EncoderLib/NALwrite.h"

void writeNaluContent(std::ostream &out, OutputNALUnit &nalu);  // NALwrite.cpp:67 (not declared in the header)

extern "C" {

// RBSP -> NAL payload bytes.  Returns the output length.
long nalref_insert(const uint8_t *rbsp, long n, uint8_t *out) {
  OutputNALUnit nalu(NAL_UNIT_CODED_SLICE_IDR_N_LP);
  nalu.m_Bitstream.getFIFO().assign(rbsp, rbsp + n);
  std::ostringstream os;
  writeNaluContent(os, nalu);
  const std::string s = os.str();
  memcpy(out, s.data(), s.size());
  return long(s.size());
}

// NAL unit (2 header bytes + escaped payload) -> RBSP (header bytes included, as VTM keeps them in the
// FIFO) and the EPB locations.  Returns the RBSP length, or -1 when VTM throws (malformed payload).
long nalref_remove(const uint8_t *nal, long n, uint8_t *rbsp, uint32_t *epbLocations, long *nEpb) {
  try {
    InputNALUnit nalu;
    nalu.getBitstream().getFifo().assign(nal, nal + n);
    read(nalu);  // NALread.cpp:157-166: convertPayloadToRBSP + header parse
    const std::vector<uint8_t> &f = nalu.getBitstream().getFifo();
    memcpy(rbsp, f.data(), f.size());
    const uint32_t k = nalu.getBitstream().numEmulationPreventionBytesRead();
    for (uint32_t i = 0; i < k; i++) epbLocations[i] = nalu.getBitstream().getEmulationPreventionByteLocation(i);
    *nEpb = long(k);
    return long(f.size());
  } catch (...) {
    return -1;
  }
}

}  // extern "C"

// msg() of CommonDef.h reads it
MsgLevel g_verbosity = SILENT;
