// cabac_device.cuh - device-side CABAC engine for sm_100a.
//
// Probability-model arithmetic shared by all kernels.  The coders themselves live in
// cabac_lockstep.cuh (decoder) and cabac_lockstep_enc.cuh (encoders).
//
// Reference behaviour restated here (paths relative to p-sawicki/entropy_coding):
//   src/common/contexts.cpp:893-956      BinProbModel_Std
//   src/entropy_codec/arith_codec.cpp    BinEncoderBase / TBinEncoder / BinDecoderBase / TBinDecoder
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "cabac_b200.h"

namespace cabac {

constexpr unsigned FULL = 0xFFFFFFFFu;

// ---------------------------------------------------------------------------
// Probability model, packed into one 32-bit shared-memory word:
//   bits  1..14  state1 (MASK_1 = 0x7FFE)          contexts.hpp:20-21
//   bits 16..17  rate0 - 2                          contexts.cpp:915-920
//   bits 18..19  rate1 - rate0 - 3
//   bits 21..30  state0 (MASK_0 = 0x7FE0 << 16)     contexts.hpp:18-19
// The rate is a per-context-index constant, so carrying it in the free bits of
// the word makes one LDS fetch everything a bin needs.
// ---------------------------------------------------------------------------
constexpr uint32_t MASK_0 = 0x7FE0u;
constexpr uint32_t MASK_1 = 0x7FFEu;

__host__ __device__ inline uint32_t model_pack(uint32_t s0, uint32_t s1, uint32_t win) {
  return (s1 & MASK_1) | ((s0 & MASK_0) << 16) | (((win >> 2) & 3u) << 16) | ((win & 3u) << 18);
}
__host__ __device__ inline uint32_t model_state0(uint32_t w) { return (w >> 16) & MASK_0; }
__host__ __device__ inline uint32_t model_state1(uint32_t w) { return w & MASK_1; }

// BinProbModel_Std::init, contexts.cpp:893-901 (qp already clipped to 0..63, :1010)
__host__ __device__ inline uint32_t model_init(int qp, uint32_t initVal, uint32_t win) {
  const int slope = int(initVal >> 3) - 4;
  const int offset = int(initVal & 7) * 18 + 1;
  int s = ((slope * (qp - 16)) >> 1) + offset;
  s = s < 1 ? 1 : (s > 127 ? 127 : s);
  const uint32_t p1 = uint32_t(s) << 8;
  return model_pack(p1, p1, win);
}

// state() >> and getLPS(), contexts.cpp:939-948.  q is the 8-bit state.
__device__ __forceinline__ uint32_t model_q(uint32_t w) {
  return ((model_state0(w) + model_state1(w)) >> 8) & 0xFFu;
}
__device__ __forceinline__ uint32_t lps_from_q(uint32_t q, uint32_t range) {
  const uint32_t qq = (q & 0x80u) ? (q ^ 0xFFu) : q;
  return ((((qq >> 2) * (range >> 5)) >> 1) + 4) & 0xFFu;
}
// m_RenormTable_32[lps >> 3], contexts.cpp:787-789: 6,5,4,4,3,3,3,3,2x8,1x16
__device__ __forceinline__ int renorm_lps(uint32_t lps) { return __clz(lps | 7u) - 24 + 1; }

// update(), contexts.cpp:903-913
__device__ __forceinline__ uint32_t model_update(uint32_t w, uint32_t bin) {
  const uint32_t r0 = 2u + ((w >> 16) & 3u);
  const uint32_t r1 = 3u + r0 + ((w >> 18) & 3u);
  uint32_t s0 = model_state0(w), s1 = model_state1(w);
  s0 -= (s0 >> r0) & MASK_0;
  s1 -= (s1 >> r1) & MASK_1;
  if (bin) {
    s0 += (0x7FFFu >> r0) & MASK_0;
    s1 += (0x7FFFu >> r1) & MASK_1;
  }
  return (w & 0x000F0000u) | s1 | (s0 << 16);
}

constexpr int STAGE_BYTES = 256;  // per-warp output ring of the fused lockstep encoder

}  // namespace cabac
