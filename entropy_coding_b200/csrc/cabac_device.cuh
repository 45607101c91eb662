// cabac_device.cuh - device-side CABAC engine for sm_100a.
//
// One warp owns one substream.  The arithmetic coder itself is a strictly
// serial recurrence (range -> LPS -> renorm -> model update), so it runs on the
// warp's lane 0 with its state in registers and the substream's probability
// models in the warp's private shared-memory slice; the other 31 lanes do the
// data-parallel work around it (coefficient staging, binarisation, context
// selection, coalesced output flushes).
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

// ---------------------------------------------------------------------------
// Output staging: bytes are produced by lane 0 into a per-warp shared-memory
// ring and leave for HBM as coalesced 16-byte stores issued by all lanes.
// ---------------------------------------------------------------------------
constexpr int STAGE_BYTES = 256;

struct OutStage {
  uint8_t *ring;    // shared, STAGE_BYTES, 16-byte aligned
  uint8_t *dst;     // global base of this substream's slab
  uint32_t cap;     // capacity in bytes
  uint32_t n;       // bytes produced (lane 0's copy is authoritative)
  uint32_t flushed; // bytes already in global memory (multiple of 16)
};

// Lane 0 only.  When the ring is full (an outstanding-0xFF run longer than the
// ring was released between two warp-wide flushes) lane 0 spills the whole ring
// itself; rare, so the serial stores do not matter.
__device__ __noinline__ void stage_spill(OutStage &o) {
  for (uint32_t b = o.flushed; b < o.flushed + uint32_t(STAGE_BYTES); b++)
    if (b < o.cap) o.dst[b] = o.ring[b & (STAGE_BYTES - 1)];
  o.flushed += STAGE_BYTES;
}
__device__ __forceinline__ void stage_put(OutStage &o, uint32_t byte) {
  if (o.n - o.flushed >= uint32_t(STAGE_BYTES)) stage_spill(o);
  o.ring[o.n & (STAGE_BYTES - 1)] = uint8_t(byte);
  o.n++;
}

// Warp-converged.  Moves every complete 16-byte chunk (all bytes when `final`)
// of [flushed, n) to global memory.
__device__ inline void stage_flush(OutStage &o, int lane, bool final) {
  const uint32_t n = __shfl_sync(FULL, o.n, 0);
  const uint32_t f = __shfl_sync(FULL, o.flushed, 0);  // lane 0 may have spilled
  const uint32_t limChunks = n & ~15u;
  __syncwarp();
  if ((reinterpret_cast<uintptr_t>(o.dst) & 15u) == 0) {
    for (uint32_t b = f + 16u * lane; b < limChunks; b += 16u * 32u) {
      if (b + 16u <= o.cap) {
        const uint4 v = *reinterpret_cast<const uint4 *>(o.ring + (b & (STAGE_BYTES - 1)));
        *reinterpret_cast<uint4 *>(o.dst + b) = v;
      } else {
        for (uint32_t k = b; k < b + 16u && k < o.cap; k++) o.dst[k] = o.ring[k & (STAGE_BYTES - 1)];
      }
    }
  } else {
    for (uint32_t b = f + lane; b < limChunks; b += 32u)
      if (b < o.cap) o.dst[b] = o.ring[b & (STAGE_BYTES - 1)];
  }
  uint32_t nf = limChunks;
  if (final) {
    for (uint32_t b = nf + lane; b < n; b += 32u)
      if (b < o.cap) o.dst[b] = o.ring[b & (STAGE_BYTES - 1)];
    nf = n;
  }
  __syncwarp();
  o.flushed = nf;
  o.n = n;
}

// ---------------------------------------------------------------------------
// Arithmetic encoder (lane 0).  arith_codec.cpp:329-582.
// ---------------------------------------------------------------------------
struct Encoder {
  uint32_t low, range, bufferedByte;
  int32_t numBuffered, bitsLeft;
  uint32_t nbins;
  OutStage out;

  // start(), :329-337
  __device__ __forceinline__ void start() {
    low = 0;
    range = 510;
    bufferedByte = 0xFF;
    numBuffered = 0;
    bitsLeft = 23;
    nbins = 0;
  }

  // writeOut(), :524-546
  __device__ __noinline__ void writeOut() {
    const uint32_t lead = low >> (24 - bitsLeft);
    bitsLeft += 8;
    low &= 0xFFFFFFFFu >> bitsLeft;
    if (lead == 0xFF) {
      numBuffered++;
    } else if (numBuffered > 0) {
      const uint32_t carry = lead >> 8;
      stage_put(out, (bufferedByte + carry) & 0xFF);
      bufferedByte = lead & 0xFF;
      const uint32_t fill = (0xFF + carry) & 0xFF;
      while (numBuffered > 1) {
        stage_put(out, fill);
        numBuffered--;
      }
    } else {
      numBuffered = 1;
      bufferedByte = lead;
    }
  }

  // TBinEncoder::encodeBin, :553-582.  `m` is the context's shared-memory word.
  __device__ __forceinline__ void encodeBin(uint32_t bin, uint32_t *m) {
    const uint32_t w = *m;
    const uint32_t q = model_q(w);
    const uint32_t lps = lps_from_q(q, range);
    nbins++;
    range -= lps;
    if (bin != (q >> 7)) {
      const int n = renorm_lps(lps);
      bitsLeft -= n;
      low = (low + range) << n;
      range = lps << n;
    } else if (range < 256) {
      bitsLeft -= 1;
      low <<= 1;
      range <<= 1;
    }
    *m = model_update(w, bin);
    if (bitsLeft < 12) writeOut();
  }

  // encodeBinsEP, :401-424 (the aligned variant :491-522 is the same
  // arithmetic and the same chunking for in-contract input, bins < 2^n)
  __device__ __forceinline__ void encodeBinsEP(uint32_t bins, uint32_t n) {
    nbins += n;
    while (n > 8) {
      n -= 8;
      const uint32_t pat = bins >> n;
      low = (low << 8) + range * pat;
      bins -= pat << n;
      bitsLeft -= 8;
      if (bitsLeft < 12) writeOut();
    }
    low = (low << n) + range * bins;
    bitsLeft -= int(n);
    if (bitsLeft < 12) writeOut();
  }

  // encodeRemAbsEP(v, rice, cutoff = 5, maxLog2TrDynamicRange = 15), :426-458
  __device__ __forceinline__ void encodeRemAbsEP(uint32_t v, uint32_t rice) {
    const uint32_t cutoff = 5, maxLog2 = 15;
    if (v < (cutoff << rice)) {
      const uint32_t len = (v >> rice) + 1;
      encodeBinsEP((1u << len) - 2u, len);
      encodeBinsEP(v & ((1u << rice) - 1u), rice);
    } else {
      const uint32_t maxPrefix = 32 - cutoff - maxLog2;
      const uint32_t code = (v >> rice) - cutoff;
      uint32_t prefix, suffixLen;
      if (code >= ((1u << maxPrefix) - 1u)) {
        prefix = maxPrefix;
        suffixLen = maxLog2;
      } else {
        prefix = 31 - __clz(code + 1);  // smallest p with code <= 2^(p+1) - 2
        suffixLen = prefix + rice + 1;
      }
      const uint32_t total = prefix + cutoff;
      const uint32_t suffix = ((code - ((1u << prefix) - 1u)) << rice) | (v & ((1u << rice) - 1u));
      encodeBinsEP((1u << total) - 1u, total);
      encodeBinsEP(suffix, suffixLen);
    }
  }

  // encodeBinTrm, :460-478
  __device__ __forceinline__ void encodeBinTrm(uint32_t bin) {
    nbins++;
    range -= 2;
    if (bin) {
      low = (low + range) << 7;
      range = 256;
      bitsLeft -= 7;
    } else if (range >= 256) {
      return;
    } else {
      low <<= 1;
      range <<= 1;
      bitsLeft--;
    }
    if (bitsLeft < 12) writeOut();
  }

  // finish(), :339-357, followed (optionally) by writeByteAlignment(),
  // bit_stream.cpp:152-155.  Without finish() the tail bits stay in `low`.
  __device__ void finish(bool terminate, bool align) {
    uint32_t tail = 0, ntail = 0;  // bits that OutputBitstream would still hold
    if (terminate) {
      encodeBinTrm(1);
      if (low >> (32 - bitsLeft)) {
        stage_put(out, (bufferedByte + 1) & 0xFF);
        while (numBuffered > 1) {
          stage_put(out, 0x00);
          numBuffered--;
        }
        low -= 1u << (32 - bitsLeft);
      } else {
        if (numBuffered > 0) stage_put(out, bufferedByte & 0xFF);
        while (numBuffered > 1) {
          stage_put(out, 0xFF);
          numBuffered--;
        }
      }
      tail = low >> 8;
      ntail = uint32_t(24 - bitsLeft);
    }
    if (align) {
      tail = (tail << 1) | 1u;  // stop bit
      ntail += 1;
    }
    // emit whole bytes MSB-first; a trailing partial byte is zero padded only
    // when aligning (otherwise the held bits are dropped, as m_held_bits would
    // stay inside the OutputBitstream)
    while (ntail >= 8) {
      stage_put(out, (tail >> (ntail - 8)) & 0xFF);
      ntail -= 8;
    }
    if (align && ntail) stage_put(out, (tail << (8 - ntail)) & 0xFF);
  }
};

// ---------------------------------------------------------------------------
// Arithmetic decoder (lane 0).  arith_codec.cpp:60-277.
// ---------------------------------------------------------------------------
struct Decoder {
  const uint8_t *in;
  uint32_t nbytes, pos;
  uint32_t range, value;
  int32_t bitsNeeded;
  uint32_t status, nbins;

  // InputBitstream::readByte, bit_stream.cpp:268-274
  __device__ __forceinline__ uint32_t readByte() {
    uint32_t b = 0;
    if (pos < nbytes)
      b = __ldg(in + pos);
    else
      status |= CABAC_SS_IN_OVERRUN;
    pos++;
    return b;
  }

  // start(), :60-66
  __device__ __forceinline__ void start(const uint8_t *p, uint32_t n) {
    in = p;
    nbytes = n;
    pos = 0;
    status = 0;
    nbins = 0;
    range = 510;
    value = readByte() << 8;
    value += readByte();
    bitsNeeded = -8;
  }

  // TBinDecoder::decodeBin, :242-277
  __device__ __forceinline__ uint32_t decodeBin(uint32_t *m) {
    const uint32_t w = *m;
    const uint32_t q = model_q(w);
    const uint32_t lps = lps_from_q(q, range);
    uint32_t bin = q >> 7;
    nbins++;
    range -= lps;
    const uint32_t sr = range << 7;
    if (value < sr) {
      if (range < 256) {
        range <<= 1;
        value <<= 1;
        if (++bitsNeeded >= 0) {
          value += readByte() << bitsNeeded;
          bitsNeeded -= 8;
        }
      }
    } else {
      bin ^= 1u;
      const int n = renorm_lps(lps);
      value = (value - sr) << n;
      range = lps << n;
      bitsNeeded += n;
      if (bitsNeeded >= 0) {
        value += readByte() << bitsNeeded;
        bitsNeeded -= 8;
      }
    }
    *m = model_update(w, bin);
    return bin;
  }

  // decodeBinEP, :100-114
  __device__ __forceinline__ uint32_t decodeBinEP() {
    nbins++;
    value += value;
    if (++bitsNeeded >= 0) {
      value += readByte();
      bitsNeeded = -8;
    }
    const uint32_t sr = range << 7;
    if (value >= sr) {
      value -= sr;
      return 1;
    }
    return 0;
  }

  // decodeAlignedBinsEP, :205-235
  __device__ uint32_t decodeAlignedBinsEP(uint32_t n) {
    uint32_t rem = n, bins = 0;
    while (rem > 0) {
      const uint32_t k = rem < 8 ? rem : 8;
      const uint32_t nb = (value >> (15 - k)) & ((1u << k) - 1u);
      bins = (bins << k) | nb;
      value = (value << k) & 0x7FFF;
      rem -= k;
      bitsNeeded += int(k);
      if (bitsNeeded >= 0) {
        value |= readByte() << bitsNeeded;
        bitsNeeded -= 8;
      }
    }
    return bins;
  }

  // decodeBinsEP, :116-151
  __device__ uint32_t decodeBinsEP(uint32_t n) {
    nbins += n;
    if (range == 256) return decodeAlignedBinsEP(n);
    uint32_t rem = n, bins = 0;
    while (rem > 8) {
      value = (value << 8) + (readByte() << (8 + bitsNeeded));
      uint32_t sr = range << 15;
#pragma unroll
      for (int i = 0; i < 8; i++) {
        bins += bins;
        sr >>= 1;
        if (value >= sr) {
          bins++;
          value -= sr;
        }
      }
      rem -= 8;
    }
    bitsNeeded += int(rem);
    value <<= rem;
    if (bitsNeeded >= 0) {
      value += readByte() << bitsNeeded;
      bitsNeeded -= 8;
    }
    uint32_t sr = range << (rem + 7);
    for (uint32_t i = 0; i < rem; i++) {
      bins += bins;
      sr >>= 1;
      if (value >= sr) {
        bins++;
        value -= sr;
      }
    }
    return bins;
  }

  // decodeRemAbsEP(rice, cutoff = 5, maxLog2TrDynamicRange = 15), :153-179
  __device__ uint32_t decodeRemAbsEP(uint32_t rice) {
    const uint32_t cutoff = 5, maxLog2 = 15, maxPrefix = 32 - maxLog2;
    uint32_t prefix = 0, cw;
    do {
      prefix++;
      cw = decodeBinEP();
    } while (cw && prefix < maxPrefix);
    prefix -= 1 - cw;
    uint32_t len = rice, offset;
    if (prefix < cutoff) {
      offset = prefix << rice;
    } else {
      offset = ((1u << (prefix - cutoff)) + cutoff - 1) << rice;
      len += (prefix == maxPrefix) ? maxLog2 - rice : prefix - cutoff;
    }
    return offset + decodeBinsEP(len);
  }

  // decodeBinTrm, :181-197
  __device__ __forceinline__ uint32_t decodeBinTrm() {
    nbins++;
    range -= 2;
    const uint32_t sr = range << 7;
    if (value >= sr) return 1;
    if (range < 256) {
      range += range;
      value += value;
      if (++bitsNeeded == 0) {
        value += readByte();
        bitsNeeded = -8;
      }
    }
    return 0;
  }

  // getNumBitsRead as traced builds count it: 8 * bytes read + m_bitsNeeded, :201-203
  __device__ __forceinline__ uint32_t bitsRead() const { return 8u * pos + uint32_t(bitsNeeded); }
};

}  // namespace cabac
