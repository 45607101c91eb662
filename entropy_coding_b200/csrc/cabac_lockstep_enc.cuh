// cabac_lockstep_enc.cuh - warp-lockstep CABAC encoder for sm_100a.
//
// The encoder's bins do not depend on the coder state, so a substream is coded in
// two alternating steps per batch (a coefficient group, a last-position prefix, ...):
//   produce : the warp binarises in parallel (one lane per scan position: template
//             sums, context ids, Rice parameters, Golomb-Rice words already packed as
//             EP bin strings) into a shared-memory queue of 32-bit command words
//   drain   : ONE instance of the arithmetic coder in the whole kernel walks the
//             queue, executed uniformly by all lanes (state in registers, the next
//             command and its context entry are fetched one step ahead)
// Exactly one encodeBin site keeps the hot loop inside the instruction cache, which
// was the limiter of the first version (profiles/r1_b_ncu_details.txt).
//
// Arithmetic (arith_codec.cpp:329-582) is restated on a 64-bit `low`: up to 24 EP
// bins enter in one multiply-add and bytes leave in a loop of writeOut steps; the
// byte stream is identical because it only depends on the exact sum.
#pragma once
#include "cabac_lockstep.cuh"

namespace cabac {

// Output staging, uniform on all lanes: bytes go to a shared-memory ring and leave
// for HBM as coalesced 16-byte stores.
// All arguments by value and out of line: one copy of the flush code in the kernel, and the
// encoder state never has its address taken (it stays in registers).  Returns the new `flushed`.
__device__ __noinline__ uint32_t lsFlushRing(const uint8_t *ring, uint8_t *dst, uint32_t cap, uint32_t n,
                                             uint32_t flushed, int lane, bool final) {
  const uint32_t lim = n & ~15u;
  __syncwarp();
  if ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0) {
    for (uint32_t b = flushed + 16u * lane; b < lim; b += 512u) {
      if (b + 16u <= cap) {
        *reinterpret_cast<uint4 *>(dst + b) = *reinterpret_cast<const uint4 *>(ring + (b & (STAGE_BYTES - 1)));
      } else {
        for (uint32_t k = b; k < b + 16u && k < cap; k++) dst[k] = ring[k & (STAGE_BYTES - 1)];
      }
    }
  } else {
    for (uint32_t b = flushed + lane; b < lim; b += 32u)
      if (b < cap) dst[b] = ring[b & (STAGE_BYTES - 1)];
  }
  uint32_t done = lim;
  if (final) {
    for (uint32_t b = lim + lane; b < n; b += 32u)
      if (b < cap) dst[b] = ring[b & (STAGE_BYTES - 1)];
    done = n;
  }
  __syncwarp();
  return done;
}

struct LsOut {  // output staging of the fused lockstep encoder, uniform on all lanes
  uint8_t *ring;  // shared, STAGE_BYTES, 16-byte aligned
  uint8_t *dst;   // this substream's slab
  uint32_t cap, n, flushed;
  int lane;

  __device__ __forceinline__ void flush(bool final) { flushed = lsFlushRing(ring, dst, cap, n, flushed, lane, final); }
  __device__ __forceinline__ void put(uint32_t byte) {
    if (n - flushed >= uint32_t(STAGE_BYTES)) flush(false);  // long outstanding-0xFF run released at once
    ring[n & (STAGE_BYTES - 1)] = uint8_t(byte);
    n++;
  }
};

template <typename Out>
struct ArithEncoder {
  uint64_t low;
  uint32_t range, bufferedByte;
  int numBuffered, bitsLeft;
  uint32_t nbins;
  Out out;

  // start(), arith_codec.cpp:329-337
  __device__ __forceinline__ void start() {
    low = 0;
    range = 510;
    bufferedByte = 0xFF;
    numBuffered = 0;
    bitsLeft = 23;
    nbins = 0;
  }
  // one writeOut() step, arith_codec.cpp:524-546
  __device__ __forceinline__ void emitByte() {
    const uint32_t lead = uint32_t(low >> (24 - bitsLeft));
    bitsLeft += 8;
    low &= (1ull << (32 - bitsLeft)) - 1ull;
    if (lead == 0xFF) {
      numBuffered++;
    } else if (numBuffered > 0) {
      const uint32_t carry = lead >> 8;
      out.put((bufferedByte + carry) & 0xFF);
      bufferedByte = lead & 0xFF;
      const uint32_t fill = (0xFF + carry) & 0xFF;
      while (numBuffered > 1) {
        out.put(fill);
        numBuffered--;
      }
    } else {
      numBuffered = 1;
      bufferedByte = lead;
    }
  }
  __device__ __forceinline__ void flushBits() {
    while (bitsLeft < 12) emitByte();
  }
  // TBinEncoder::encodeBin, arith_codec.cpp:553-582; returns the updated state word
  __device__ __forceinline__ uint32_t encodeCtx(uint32_t bin, const uint4 e) {
    const uint32_t q = ((e.x >> 16) + (e.x & 0xFFFFu)) >> 8;
    const uint32_t mps = q >> 7;
    const uint32_t qq = q ^ (0xFFu & (0u - mps));
    const uint32_t lps = (((qq >> 2) * (range >> 5)) >> 1) + 4;
    nbins++;
    const uint32_t rmps = range - lps;
    const bool isLps = bin != mps;
    const int n = isLps ? __clz(int(lps | 7u)) - 23 : 1 - int(rmps >> 8);
    low = (low + (isLps ? rmps : 0u)) << n;
    range = (isLps ? lps : rmps) << n;
    bitsLeft -= n;
    const uint32_t d0 = (e.x >> e.z) & (MASK_0 << 16);
    const uint32_t d1 = ((e.x & 0xFFFFu) >> e.w) & MASK_1;
    return e.x - d0 - d1 + (bin ? e.y : 0u);
  }
  // the same with the probability already resolved by stage P: k = LPS table row, isLps
  __device__ __forceinline__ void encodeResolved(uint32_t k, bool isLps) {
    const uint32_t lps = ((k * (range >> 5)) >> 1) + 4;
    nbins++;
    const uint32_t rmps = range - lps;
    const int n = isLps ? __clz(int(lps | 7u)) - 23 : 1 - int(rmps >> 8);
    low = (low + (isLps ? rmps : 0u)) << n;
    range = (isLps ? lps : rmps) << n;
    bitsLeft -= n;
  }
  // One resolved CTX word or one EP word, without a branch between the two:
  //   (low + x) << n == (low << n) + (x << n), so both kinds are low = (low << n) + addend.
  __device__ __forceinline__ void step(uint32_t w) {
    const bool isCtx = (w >> 29) == CABAC_CMD_CTX;
    const bool isLps = (w >> 6) & 1u;
    const uint32_t epN = (w >> 24) & 31u;
    const uint32_t lps = (((w & 63u) * (range >> 5)) >> 1) + 4;
    const uint32_t rmps = range - lps;
    const uint32_t nCtx = isLps ? uint32_t(__clz(int(lps | 7u)) - 23) : 1u - (rmps >> 8);
    const uint32_t n = isCtx ? nCtx : epN;
    const uint64_t add = isCtx ? uint64_t(isLps ? rmps << nCtx : 0u) : uint64_t(range) * (w & 0xFFFFFFu);
    range = isCtx ? (isLps ? lps : rmps) << nCtx : range;
    low = (low << n) + add;
    bitsLeft -= int(n);
    nbins += isCtx ? 1u : epN;
  }
  // encodeBinsEP, arith_codec.cpp:401-424, n <= 24 bins at once
  __device__ __forceinline__ void ep(uint32_t bins, int n) {
    low = (low << n) + uint64_t(range) * bins;
    bitsLeft -= n;
    nbins += uint32_t(n);
  }
  // encodeBinTrm, arith_codec.cpp:460-478
  __device__ __forceinline__ void trm(uint32_t bin) {
    nbins++;
    range -= 2;
    if (bin) {
      low = (low + range) << 7;
      range = 256;
      bitsLeft -= 7;
    } else if (range < 256) {
      low <<= 1;
      range <<= 1;
      bitsLeft--;
    }
    flushBits();
  }
  // finish(), arith_codec.cpp:339-357, then optionally writeByteAlignment(), bit_stream.cpp:152-155.
  // Without `terminate` the tail bits stay in `low` (nothing is emitted), as the reference
  // leaves them in the coder until finish().
  __device__ __forceinline__ void finish(bool terminate, bool align) {
    uint32_t tail = 0, ntail = 0;  // bits an OutputBitstream would still hold
    if (terminate) {
      trm(1);
      if (low >> (32 - bitsLeft)) {
        out.put((bufferedByte + 1) & 0xFF);
        while (numBuffered > 1) {
          out.put(0x00);
          numBuffered--;
        }
        low -= 1ull << (32 - bitsLeft);
      } else {
        if (numBuffered > 0) out.put(bufferedByte & 0xFF);
        while (numBuffered > 1) {
          out.put(0xFF);
          numBuffered--;
        }
      }
      tail = uint32_t(low >> 8);
      ntail = uint32_t(24 - bitsLeft);
    }
    if (align) {
      tail = (tail << 1) | 1u;  // stop bit
      ntail += 1;
    }
    while (ntail >= 8) {
      out.put((tail >> (ntail - 8)) & 0xFF);
      ntail -= 8;
    }
    if (align && ntail) out.put((tail << (8 - ntail)) & 0xFF);
  }
};

typedef ArithEncoder<LsOut> LsEncoder;

// The single arithmetic-coder site: walks n command words (CTX / EP kinds).
__device__ __forceinline__ void lsDrain(LsEncoder &enc, uint4 *ctx, const uint32_t *queue, int n) {
  if (n <= 0) return;
  uint32_t w = queue[0];
  uint4 e = ctx[(w >> 29) ? 0u : (w & 511u)];
#pragma unroll 1
  for (int i = 0; i < n; i++) {
    // one step ahead: the next command and (speculatively) its context entry
    const uint32_t wn = queue[min(i + 1, n - 1)];
    const uint32_t idxn = (wn >> 29) ? 0u : (wn & 511u);
    uint4 en = ctx[idxn];
    if ((w >> 29) == CABAC_CMD_CTX) {
      const uint32_t idx = w & 511u;
      const uint32_t nx = enc.encodeCtx((w >> 9) & 1u, e);
      ctx[idx].x = nx;
      if (idxn == idx) en.x = nx;  // the entry fetched ahead was the one just updated
    } else {
      enc.ep(w & 0xFFFFFFu, int((w >> 24) & 31u));
    }
    enc.flushBits();
    w = wn;
    e = en;
  }
}

// The same walk over words whose probabilities were already resolved in parallel
// (lsResolveQueue below): only the interval recurrence is left on the serial path.
__device__ __forceinline__ void lsDrainResolved(LsEncoder &enc, const uint32_t *queue, int n) {
  if (n <= 0) return;
  uint32_t w = queue[0];
#pragma unroll 1
  for (int i = 0; i < n; i++) {
    const uint32_t wn = queue[min(i + 1, n - 1)];
    enc.step(w);
    enc.flushBits();
    w = wn;
  }
}

// ---------------------------------------------------------------------------
// LsCoder: the interval coder of the fused kernel's resolved walk, uniform on all lanes.
//
// Same arithmetic as ArithEncoder (arith_codec.cpp:329-582 on a 64-bit `low`), organised for the
// fewest instructions per word on the serial path:
//   * finished bytes leave `low` three to five at a time (when bitsLeft < -4) instead of one per
//     writeOut: until then a carry is an ordinary add inside the 64-bit register;
//   * emitted bytes wait big-endian in `acc` (4..7 most recent bytes), where a late carry is again
//     a plain integer add; only a carry through all pending bytes patches stored bytes (carrySlow);
//   * groups of 4 bytes go to the shared-memory ring and from there to HBM as 16-byte stores;
//   * bins are counted by the parallel resolve step, not here.
// `vtmPending` mirrors m_numBufferedBytes so that a substream that is NOT terminated reports
// exactly the bytes the reference coder would already have handed to its OutputBitstream.
// ---------------------------------------------------------------------------
__device__ __noinline__ void lsCarrySlow(uint8_t *ring, uint8_t *dst, uint32_t cap, uint32_t n, uint32_t flushed,
                                         int lane) {
  __syncwarp();
  if (lane == 0) {
    while (n > 0) {
      n--;
      uint8_t b;
      if (n >= flushed) {
        b = uint8_t(ring[n & (STAGE_BYTES - 1)] + 1);
        ring[n & (STAGE_BYTES - 1)] = b;
      } else if (n < cap) {
        b = uint8_t(dst[n] + 1);
        dst[n] = b;
      } else {
        continue;
      }
      if (b != 0) break;
    }
  }
  __syncwarp();
}

struct LsCoder {
  uint64_t low, acc;
  uint32_t range;
  int bitsLeft;
  uint32_t cnt;         // bytes pending in acc (0..7)
  uint32_t n;           // bytes handed to the ring so far (multiple of 4 until finish)
  uint32_t flushed;     // bytes of the ring already stored to HBM
  uint32_t vtmPending;  // m_numBufferedBytes of the reference coder
  uint32_t nbins;
  int lazyBelow;        // bytes leave `low` when bitsLeft drops below this
  bool trackPending;    // keep vtmPending (substream not terminated)
  uint8_t *ring, *dst;
  uint32_t cap;
  int lane;

  // `terminated`: the substream ends with finish(), so nobody observes the reference's outstanding-
  // byte count and finished bits may wait in `low` (lazy emission).  Otherwise bytes leave one per
  // writeOut() exactly as in the reference, which keeps `vtmPending` exact.
  __device__ __forceinline__ void start(uint8_t *ring_, uint8_t *dst_, uint32_t cap_, int lane_, bool terminated) {
    lazyBelow = terminated ? -4 : 12;
    trackPending = !terminated;
    low = 0;
    acc = 0;
    range = 510;
    bitsLeft = 23;
    cnt = 0;
    n = 0;
    flushed = 0;
    vtmPending = 0;
    nbins = 0;
    ring = ring_;
    dst = dst_;
    cap = cap_;
    lane = lane_;
  }
  __device__ __forceinline__ void flushRing(bool final) { flushed = lsFlushRing(ring, dst, cap, n, flushed, lane, final); }
  __device__ __forceinline__ void spill4() {  // oldest 4 of the pending bytes -> ring
    const uint32_t be = uint32_t(acc >> (8 * (cnt - 4)));
    *reinterpret_cast<uint32_t *>(ring + (n & (STAGE_BYTES - 1))) = __byte_perm(be, 0, 0x0123);
    n += 4;
    cnt -= 4;
    acc &= (1ull << (8 * cnt)) - 1ull;
    if (n - flushed >= uint32_t(STAGE_BYTES / 2)) flushRing(false);
  }
  __device__ __forceinline__ void addCarry(uint32_t carry) {
    acc += carry;
    if (cnt < 8 && (acc >> (8 * cnt))) {  // rippled through every pending byte
      acc &= (1ull << (8 * cnt)) - 1ull;
      lsCarrySlow(ring, dst, cap, n, flushed, lane);
    }
  }
  // k writeOut() steps at once (arith_codec.cpp:524-546), 1 <= k <= 5, cnt <= 3 on entry
  __device__ __forceinline__ void emitBytes(uint32_t k) {
    const int s = 24 - bitsLeft - 8 * int(k - 1);
    const uint64_t chunk = low >> s;  // carry bit + k bytes
    low -= chunk << s;
    bitsLeft += 8 * int(k);
    if (cnt + k > 8) spill4();  // cnt <= 7, k <= 5: one spill always makes room
    const uint32_t carry = uint32_t(chunk >> (8 * k));
    addCarry(carry);
    const uint64_t bytes = chunk & ((1ull << (8 * k)) - 1ull);
    acc = (acc << (8 * k)) | bytes;
    cnt += k;
    if (trackPending) {
      // m_numBufferedBytes, only observable when the substream is not terminated: a lead byte of
      // exactly 0xFF (no carry) extends the outstanding run, anything else restarts it at 1.  Exact
      // whenever bytes leave one at a time; a packed EP word can release up to 3 at once, where the
      // reference's 8-bin steps (arith_codec.cpp:401-424) may see a carry arrive one byte later -
      // the byte stream is unaffected, only this interim count can then differ.
      const uint32_t t = min(uint32_t(__ffsll(~(long long)bytes) - 1) >> 3, k);
      vtmPending = t == k ? (carry ? k : vtmPending + k) : t + 1u;
    }
    if (cnt == 8) spill4();
  }
  // everything the reference has written by now: back to bitsLeft >= 12
  __device__ __forceinline__ void flushBits() {
    if (bitsLeft < 12) emitBytes(uint32_t(19 - bitsLeft) >> 3);
  }
  // lazy variant used inside the walk: lets up to 16 + 24 finished bits wait in `low`
  __device__ __forceinline__ void flushLazy() {
    if (bitsLeft < lazyBelow) emitBytes(uint32_t(19 - bitsLeft) >> 3);
  }
  // resolved context-coded word: bits 0..5 LPS table row, bit 6 "bin is the LPS"
  __device__ __forceinline__ void ctxWord(uint32_t w) {
    const uint32_t lps = (((w & 63u) * (range >> 5)) >> 1) + 4;
    const uint32_t rmps = range - lps;
    const bool isLps = (w & 64u) != 0;
    const uint32_t nr = isLps ? lps : rmps;  // rmps >= 128, so one normalisation rule serves both
    const int sh = __clz(int(nr | 7u)) - 23;
    low = (low + (isLps ? rmps : 0u)) << sh;
    range = nr << sh;
    bitsLeft -= sh;
  }
  __device__ __forceinline__ void epWord(uint32_t w) {  // encodeBinsEP, <= 24 bins
    const int k = int((w >> 24) & 31u);
    low = (low << k) + uint64_t(range) * (w & 0xFFFFFFu);
    bitsLeft -= k;
  }
  // encodeBinTrm, arith_codec.cpp:460-478
  __device__ __forceinline__ void trm(uint32_t bin) {
    nbins++;
    range -= 2;
    if (bin) {
      low = (low + range) << 7;
      range = 256;
      bitsLeft -= 7;
    } else if (range < 256) {
      low <<= 1;
      range <<= 1;
      bitsLeft--;
    }
    flushBits();
  }
  __device__ __forceinline__ void putAt(uint32_t at, uint32_t b) {
    if (at >= flushed) ring[at & (STAGE_BYTES - 1)] = uint8_t(b);
    else if (at < cap && lane == 0) dst[at] = uint8_t(b);
  }
  // finish() + writeByteAlignment() (arith_codec.cpp:339-357, bit_stream.cpp:152-155); returns the byte
  // count and stores everything.  Without `terminate` only the bytes the reference coder would already
  // have written are reported (its buffered byte and 0xFF run stay back).
  __device__ __forceinline__ uint32_t finish(bool terminate, bool align) {
    flushBits();
    uint32_t tail = 0, ntail = 0;
    uint32_t keepBack = vtmPending;
    if (terminate) {
      trm(1);
      const uint32_t carry = uint32_t(low >> (32 - bitsLeft));
      addCarry(carry);
      low -= uint64_t(carry) << (32 - bitsLeft);
      tail = uint32_t(low >> 8);
      ntail = uint32_t(24 - bitsLeft);
      keepBack = 0;
    }
    if (align) {
      tail = (tail << 1) | 1u;  // stop bit
      ntail += 1;
    }
    if (n + cnt - flushed > uint32_t(STAGE_BYTES) - 8u) flushRing(false);
    for (uint32_t k = cnt; k > 0; k--) {  // pending bytes, oldest first
      ring[n & (STAGE_BYTES - 1)] = uint8_t(acc >> (8 * (k - 1)));
      n++;
    }
    cnt = 0;
    uint32_t total = n - keepBack;
    while (ntail >= 8) {
      putAt(total++, (tail >> (ntail - 8)) & 0xFF);
      ntail -= 8;
    }
    if (align && ntail) putAt(total++, (tail << (8 - ntail)) & 0xFF);
    if (total < flushed) flushed = total;  // nothing beyond `total` is reported
    n = total;
    flushRing(true);
    return total;
  }
};

// The walk over resolved words on an LsCoder: a uniform branch per word (all lanes hold the same
// word), one lazy byte-emission test; the queue is read one word ahead.
__device__ __forceinline__ void lsWalkResolved(LsCoder &enc, const uint32_t *queue, int n) {
  if (n <= 0) return;
  const uint32_t *q = queue, *end = queue + n;
  uint32_t w = *q;
#pragma unroll 1
  do {
    q++;
    const uint32_t wn = *q;  // queue[n] is readable (QUEUE_WORDS leaves a spare slot)
    if ((w >> 29) == CABAC_CMD_CTX) enc.ctxWord(w); else enc.epWord(w);
    enc.flushLazy();
    w = wn;
  } while (q != end);
}

// The walk for recorded streams: besides resolved CTX / EP words they carry the rare kinds of
// BinEncIf (terminating bins, align()).
__device__ __forceinline__ void lsWalkResolvedAny(LsCoder &enc, const uint32_t *queue, int n) {
  if (n <= 0) return;
  const uint32_t *q = queue, *end = queue + n;
  uint32_t w = *q;
#pragma unroll 1
  do {
    q++;
    const uint32_t wn = *q;
    const uint32_t kind = w >> 29;
    if (kind == CABAC_CMD_CTX) {
      enc.ctxWord(w);
    } else if (kind == CABAC_CMD_EP) {
      enc.epWord(w);
    } else if (kind == CABAC_CMD_TRM) {
      enc.flushBits();  // encodeBinTrm works on the eager state (arith_codec.cpp:460-478)
      enc.trm(w & 1u);
    } else {  // CABAC_CMD_ALIGN: align(), arith_codec.cpp:480
      enc.range = 256;
    }
    enc.flushLazy();
    w = wn;
  } while (q != end);
}

// encodeRemAbsEP(v, rice, cutoff = 5, maxLog2TrDynamicRange = 15), arith_codec.cpp:426-458, as one or
// two EP command words (prefix and suffix concatenated when they fit 24 bins).  Returns the count.
__device__ __forceinline__ int lsRemWords(uint32_t v, uint32_t rice, uint32_t &w0, uint32_t &w1) {
  const uint32_t mask = (1u << rice) - 1u;
  if (v < (5u << rice)) {
    const uint32_t len = (v >> rice) + 1u;
    w0 = CABAC_MK_EP((((1u << len) - 2u) << rice) | (v & mask), len + rice);
    return 1;
  }
  const uint32_t code = (v >> rice) - 5u;
  uint32_t prefix, suffixLen;
  if (code >= ((1u << 12) - 1u)) {
    prefix = 12;
    suffixLen = 15;
  } else {
    prefix = 31 - __clz(int(code + 1u));  // smallest p with code <= 2^(p+1) - 2
    suffixLen = prefix + rice + 1u;
  }
  const uint32_t total = prefix + 5u;
  const uint32_t suffix = ((code - ((1u << prefix) - 1u)) << rice) | (v & mask);
  if (total + suffixLen <= 24u) {
    w0 = CABAC_MK_EP((((1u << total) - 1u) << suffixLen) | suffix, total + suffixLen);
    return 1;
  }
  w0 = CABAC_MK_EP((1u << total) - 1u, total);
  w1 = CABAC_MK_EP(suffix, suffixLen);
  return 2;
}


// ===========================================================================
// Two-stage encoder.  The bins of an encoder do not depend on the coder state, and
// the probability a context-coded bin is coded with depends only on the earlier
// bins of the SAME context.  So everything except the interval recurrence is done
// by the substream's warp in parallel (stage P), which leaves "resolved" command
// words in HBM; stage D is a pure interval coder with one THREAD per substream:
// 32 substreams per warp-instruction, no context table, no model update.
//
// Resolved word: kind in bits 31..29 as in cabac_b200.h; a CTX word carries
//   bits 0..5 = (q ^ mps-mask) >> 2  (the LPS table row, BinProbModel_Std::getLPS, contexts.cpp:945-950)
//   bit  6    = 1 when the bin is the LPS
// ===========================================================================

// Stage P, probability modelling of one queue (in place).  Lane i owns command i of each
// 32-command slice; commands that share a context are serialised by their rank among
// the slice's users of that context (match_any), everything else runs in parallel.
// kTranslate: the words carry the residual kernels' compact context ids while `ctx` is the full
// 379-entry table.
template <bool kTranslate>
__device__ __forceinline__ uint32_t lsResolveQueue(uint4 *ctx, uint32_t *queue, int n, int lane) {
  uint32_t bins = 0;  // BinCounter::getAll() of the CTX / EP words (uniform)
  for (int base = 0; base < n; base += 32) {
    const int i = base + lane;
    uint32_t w = i < n ? queue[i] : (uint32_t(CABAC_CMD_EP) << 29);
    const bool isCtx = (w >> 29) == CABAC_CMD_CTX;
    const uint32_t idx = kTranslate ? uint32_t(rctx_inv(int(w & 511u))) : (w & 511u), bin = (w >> 9) & 1u;
    const uint32_t peers = __match_any_sync(FULL, isCtx ? idx : (0x8000u | uint32_t(lane)));
    const int rank = __popc(peers & ((1u << lane) - 1u));
    const int rounds = __reduce_max_sync(FULL, isCtx ? rank : 0);
    for (int r = 0; r <= rounds; r++) {
      if (isCtx && rank == r) {
        const uint4 e = ctx[idx];
        const uint32_t q = ((e.x >> 16) + (e.x & 0xFFFFu)) >> 8;
        const uint32_t mps = q >> 7;
        const uint32_t qq = q ^ (0xFFu & (0u - mps));
        const uint32_t d0 = (e.x >> e.z) & (MASK_0 << 16);
        const uint32_t d1 = ((e.x & 0xFFFFu) >> e.w) & MASK_1;
        ctx[idx].x = e.x - d0 - d1 + (bin ? e.y : 0u);
        w = (qq >> 2) | ((bin ^ mps) << 6);
      }
      __syncwarp();
    }
    if (i < n) queue[i] = w;
    bins += __reduce_add_sync(FULL, i < n ? (isCtx ? 1u : ((w >> 29) == CABAC_CMD_EP ? (w >> 24) & 31u : 0u)) : 0u);
  }
  __syncwarp();
  return bins;
}


// lsResolveQueue for the warp-specialised producers: 4-byte context states, the per-index constants
// (y, z, w of ctx_entry) in one table shared by the CTA.
__device__ __forceinline__ uint32_t lsResolveQueueStates(uint32_t *ctx, const uint4 *rates, uint32_t *queue, int n,
                                                         int lane) {
  uint32_t bins = 0;
  for (int base = 0; base < n; base += 32) {
    const int i = base + lane;
    uint32_t w = i < n ? queue[i] : (uint32_t(CABAC_CMD_EP) << 29);
    const bool isCtx = (w >> 29) == CABAC_CMD_CTX;
    const uint32_t idx = w & 511u, bin = (w >> 9) & 1u;
    const uint32_t peers = __match_any_sync(FULL, isCtx ? idx : (0x8000u | uint32_t(lane)));
    const int rank = __popc(peers & ((1u << lane) - 1u));
    const int rounds = __reduce_max_sync(FULL, isCtx ? rank : 0);
    const uint4 k = rates[isCtx ? idx : 0u];
    for (int r = 0; r <= rounds; r++) {
      if (isCtx && rank == r) {
        const uint32_t x = ctx[idx];
        const uint32_t q = ((x >> 16) + (x & 0xFFFFu)) >> 8;
        const uint32_t mps = q >> 7;
        const uint32_t qq = q ^ (0xFFu & (0u - mps));
        const uint32_t d0 = (x >> k.z) & (MASK_0 << 16);
        const uint32_t d1 = ((x & 0xFFFFu) >> k.w) & MASK_1;
        ctx[idx] = x - d0 - d1 + (bin ? k.y : 0u);
        w = (qq >> 2) | ((bin ^ mps) << 6);
      }
      __syncwarp();
    }
    if (i < n) queue[i] = w;
    bins += __reduce_add_sync(FULL, i < n ? (isCtx ? 1u : ((w >> 29) == CABAC_CMD_EP ? (w >> 24) & 31u : 0u)) : 0u);
  }
  __syncwarp();
  return bins;
}

// The RDO bit estimator (BitEstimator_Std, arith_codec.cpp:603-711) over one queue: the same
// rank-serialised model walk as lsResolveQueue, but what leaves is the 15-bit fixed-point cost of the
// words - m_binFracBits[state()][bin] for a context-coded bin BEFORE its update (estFracBitsUpdate,
// contexts.cpp:922-925), numBins << SCALE_BITS for bypass bins (estFracBitsEP, contexts.cpp:882-884).
// There is no interval and no serial walk.  Returns the (uniform) cost; `bins` as lsResolveQueue.
// ctxBins (nullable, this substream's CABAC_NUM_CTX counters, zeroed by the caller) is the BinCounter
// view (BinCounter::addCtx, arith_codec.hpp:72-93): the first lane of each same-context group adds the
// group's size, the warp owns the row.
__device__ __forceinline__ uint64_t lsEstimateQueue(uint4 *ctx, const uint32_t *queue, int n, int lane,
                                                    const uint32_t *__restrict__ fracBits, uint32_t &bins,
                                                    uint32_t *ctxBins = nullptr) {
  uint32_t cost = 0;  // this lane's share; <= 2 words of <= 2^20 per slice, queues are << 2^10 slices
  uint32_t nb = 0;
  for (int base = 0; base < n; base += 32) {
    const int i = base + lane;
    const uint32_t w = i < n ? queue[i] : (uint32_t(CABAC_CMD_EP) << 29);
    const bool isCtx = (w >> 29) == CABAC_CMD_CTX;
    const uint32_t idx = w & 511u, bin = (w >> 9) & 1u;
    const uint32_t peers = __match_any_sync(FULL, isCtx ? idx : (0x8000u | uint32_t(lane)));
    const int rank = __popc(peers & ((1u << lane) - 1u));
    const int rounds = __reduce_max_sync(FULL, isCtx ? rank : 0);
    if (ctxBins && isCtx && rank == 0) atomicAdd(ctxBins + rctx_inv(int(idx)), uint32_t(__popc(peers)));
    for (int r = 0; r <= rounds; r++) {
      if (isCtx && rank == r) {
        const uint4 e = ctx[idx];
        const uint32_t q = ((e.x >> 16) + (e.x & 0xFFFFu)) >> 8;
        const uint32_t d0 = (e.x >> e.z) & (MASK_0 << 16);
        const uint32_t d1 = ((e.x & 0xFFFFu) >> e.w) & MASK_1;
        ctx[idx].x = e.x - d0 - d1 + (bin ? e.y : 0u);
        cost += __ldg(&fracBits[2u * q + bin]);
        nb++;
      }
      __syncwarp();
    }
    if (!isCtx && (w >> 29) == CABAC_CMD_EP) {
      const uint32_t k = (w >> 24) & 31u;
      cost += k << 15;
      nb += k;
    }
  }
  __syncwarp();
  bins += __reduce_add_sync(FULL, nb);
  // 64-bit sum of the lanes' 32-bit shares
  return uint64_t(__reduce_add_sync(FULL, cost & 0xFFFFu)) + (uint64_t(__reduce_add_sync(FULL, cost >> 16)) << 16);
}

// ---------------------------------------------------------------------------
// Stage D coder: interval recurrence + byte output for ONE substream per lane.
//
// Output bytes are assembled big-endian in a 64-bit register that always keeps the most
// recent 4..7 bytes pending, so a carry out of `low` is a plain integer add on the pending
// bytes (the reference holds one byte + a count of 0xFF bytes for the same purpose,
// arith_codec.cpp:524-546).  Only a carry through >= 4 pending 0xFF bytes has to patch bytes
// already stored in the slab (carrySlow).  `vtmPending` tracks m_numBufferedBytes so that a
// substream that is NOT terminated reports exactly the bytes the reference would have written.
// ---------------------------------------------------------------------------
struct RangeCoder {
  uint64_t low, acc;
  uint32_t range, nbins;
  int bitsLeft;
  uint32_t cnt;         // bytes pending in acc (0..7)
  uint32_t n;           // bytes stored to the slab
  uint32_t vtmPending;  // m_numBufferedBytes of the reference coder
  uint8_t *dst;
  uint32_t cap;

  __device__ __forceinline__ void start(uint8_t *p, uint32_t c) {
    low = 0;
    acc = 0;
    range = 510;
    nbins = 0;
    bitsLeft = 23;
    cnt = 0;
    n = 0;
    vtmPending = 0;
    dst = p;
    cap = c;
  }
  __device__ __forceinline__ void storeByte(uint32_t at, uint32_t b) {
    if (at < cap) dst[at] = uint8_t(b);
  }
  // carry out of all pending bytes: increment the stored bytes backwards (essentially never)
  __device__ __noinline__ static void carrySlow(uint8_t *dst, uint32_t cap, uint32_t n) {
    while (n > 0) {
      n--;
      if (n >= cap) continue;
      const uint8_t b = uint8_t(dst[n] + 1);
      dst[n] = b;
      if (b != 0) break;
    }
  }
  __device__ __forceinline__ void spill4() {  // oldest 4 of 8 pending bytes -> slab
    const uint32_t be = uint32_t(acc >> 32);
    if ((reinterpret_cast<uintptr_t>(dst) & 3u) == 0 && n + 4u <= cap) {
      *reinterpret_cast<uint32_t *>(dst + n) = __byte_perm(be, 0, 0x0123);
    } else {
      storeByte(n, be >> 24);
      storeByte(n + 1, (be >> 16) & 0xFF);
      storeByte(n + 2, (be >> 8) & 0xFF);
      storeByte(n + 3, be & 0xFF);
    }
    n += 4;
    acc &= 0xFFFFFFFFull;
    cnt = 4;
  }
  __device__ __forceinline__ void addCarry(uint32_t carry) {
    acc += carry;
    if (cnt < 8 && (acc >> (8 * cnt))) {  // rippled through every pending byte
      acc &= (1ull << (8 * cnt)) - 1ull;
      carrySlow(dst, cap, n);
    }
  }
  // one writeOut() step, arith_codec.cpp:524-546
  __device__ __forceinline__ void emitByte() {
    const int s = 24 - bitsLeft;
    const uint32_t lead = uint32_t(low >> s);
    low -= uint64_t(lead) << s;
    bitsLeft += 8;
    addCarry(lead >> 8);
    acc = (acc << 8) | (lead & 0xFFu);
    cnt++;
    vtmPending = lead == 0xFFu ? vtmPending + 1u : 1u;
    if (cnt == 8) spill4();
  }
  __device__ __forceinline__ void flushBits() {
    while (bitsLeft < 12) emitByte();
  }
  // One resolved CTX word or one EP word, without a branch between the two:
  //   (low + x) << n == (low << n) + (x << n), so both kinds are low = (low << n) + addend.
  __device__ __forceinline__ void step(uint32_t w) {
    const bool isCtx = (w >> 29) == CABAC_CMD_CTX;
    const bool isLps = (w >> 6) & 1u;
    const uint32_t epN = (w >> 24) & 31u;
    const uint32_t lps = (((w & 63u) * (range >> 5)) >> 1) + 4;
    const uint32_t rmps = range - lps;
    const uint32_t nCtx = isLps ? uint32_t(__clz(int(lps | 7u)) - 23) : 1u - (rmps >> 8);
    const uint32_t n_ = isCtx ? nCtx : epN;
    const uint64_t add = isCtx ? uint64_t(isLps ? rmps << nCtx : 0u) : uint64_t(range) * (w & 0xFFFFFFu);
    range = isCtx ? (isLps ? lps : rmps) << nCtx : range;
    low = (low << n_) + add;
    bitsLeft -= int(n_);
    nbins += isCtx ? 1u : epN;
  }
  // encodeBinTrm, arith_codec.cpp:460-478
  __device__ __forceinline__ void trm(uint32_t bin) {
    nbins++;
    range -= 2;
    if (bin) {
      low = (low + range) << 7;
      range = 256;
      bitsLeft -= 7;
    } else if (range < 256) {
      low <<= 1;
      range <<= 1;
      bitsLeft--;
    }
    flushBits();
  }
  // finish() + writeByteAlignment() (arith_codec.cpp:339-357, bit_stream.cpp:152-155); returns the
  // substream's byte count.  Without `terminate` only the bytes the reference coder would already
  // have handed to its OutputBitstream are reported (the buffered byte and its 0xFF run stay back).
  __device__ __forceinline__ uint32_t finish(bool terminate, bool align) {
    uint32_t tail = 0, ntail = 0;
    uint32_t keepBack = vtmPending;
    if (terminate) {
      trm(1);
      const uint32_t carry = uint32_t(low >> (32 - bitsLeft));
      addCarry(carry);
      low -= uint64_t(carry) << (32 - bitsLeft);
      tail = uint32_t(low >> 8);
      ntail = uint32_t(24 - bitsLeft);
      keepBack = 0;
    }
    if (align) {
      tail = (tail << 1) | 1u;  // stop bit
      ntail += 1;
    }
    // pending bytes, oldest first
    for (uint32_t k = cnt; k > 0; k--) storeByte(n + (cnt - k), uint32_t(acc >> (8 * (k - 1))) & 0xFFu);
    uint32_t total = n + cnt - keepBack;
    while (ntail >= 8) {
      storeByte(total++, (tail >> (ntail - 8)) & 0xFF);
      ntail -= 8;
    }
    if (align && ntail) storeByte(total++, (tail << (8 - ntail)) & 0xFF);
    return total;
  }
};

// ===========================================================================
// Warp-specialised fused encoder.
//
// Of the interval coder only the RANGE recurrence is inherently serial: range' = f(range, word), nine
// bits of state, ~10 instructions.  `low` is a sum of shifted addends - associative - so it can be
// accumulated in groups by the lanes of a warp once the range each word saw is known.
//
//   producer warp (one substream): binarise -> resolve probabilities -> push resolved words into
//       its ring in shared memory                                    (all lane-parallel)
//   walker warp (one per CTA): lane j runs the range recurrence of producer j's ring and leaves
//       the range every word saw next to it            (serial, but WS_PRODUCERS chains per warp-
//       instruction instead of one)
//   producer warp again: 32 words at a time, lanes compute (shift, addend) from (word, range), a prefix
//       sum groups consecutive words into <= 39-bit pieces ("low = (low << K) + C"), and the warp's
//       LsCoder takes one group - two to five output bytes - per serial step.
// ===========================================================================
#ifndef CABAC_WS_PRODUCERS
#define CABAC_WS_PRODUCERS 19
#endif
#ifndef CABAC_WS_WINDOW_BITS
#define CABAC_WS_WINDOW_BITS 4
#endif
constexpr int WS_PRODUCERS = CABAC_WS_PRODUCERS;
#ifndef CABAC_WS_RING
#define CABAC_WS_RING 256
#endif
constexpr uint32_t WS_RING = CABAC_WS_RING;  // words per channel (a power of two > QUEUE_WORDS + one drain slice)
constexpr uint32_t WS_KIND_CTL = 6;  // control word: bits 28..27 = op
constexpr uint32_t WS_OP_START = 0;  // new substream: range = 510
constexpr uint32_t WS_OP_EXIT = 2;
constexpr uint32_t WS_SPIN_LIMIT = 1u << 24;  // x >= 32 ns: a stuck hand-off traps instead of hanging the GPU
#define WS_MK_CTL(op) ((WS_KIND_CTL << 29) | (uint32_t(op) << 27))

struct __align__(16) WsChannel {
  uint32_t ring[WS_RING];  // resolved words
  uint16_t rng[WS_RING];   // m_Range each word saw (written by the walker)
  uint2 pieces[32];        // (K, C) pieces of the slice being drained (producer only)
  uint32_t head;           // words published by the producer (free-running)
  uint32_t tail;           // words the walker has passed (free-running)
  uint32_t want;           // the producer is waiting for the walker
  uint32_t pad;
};

// Head / tail hand-off inside one CTA: the data words (ring[], rng[]) are plain shared-memory stores, the
// counters volatile ones.  A __threadfence_block() sits between the data and the counter on both sides
// (writer: data, fence, counter; reader: counter, fence, data) so the order is what the PTX memory model
// guarantees, not just what current hardware happens to do; one fence per push / walker round is free at
// this rate.
__device__ __forceinline__ uint32_t wsLoad(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(v) : "r"(uint32_t(__cvta_generic_to_shared(p))) : "memory");
  return v;
}
__device__ __forceinline__ void wsStore(uint32_t *p, uint32_t v) {
  asm volatile("st.volatile.shared.u32 [%0], %1;" ::"r"(uint32_t(__cvta_generic_to_shared(p))), "r"(v) : "memory");
}

// producer side; the caller has made room (wsDrain in the kernel)
__device__ __forceinline__ void wsPush(WsChannel &ch, uint32_t &head, const uint32_t *q, int n, int lane) {
  for (int i = lane; i < n; i += 32) ch.ring[(head + uint32_t(i)) & (WS_RING - 1u)] = q[i];
  head += uint32_t(n);
  __threadfence_block();
  __syncwarp();
  if (lane == 0) wsStore(&ch.head, head);
}
__device__ __forceinline__ void wsPushWord(WsChannel &ch, uint32_t &head, uint32_t w, int lane) {
  if (lane == 0) {
    ch.ring[head & (WS_RING - 1u)] = w;
    __threadfence_block();
    wsStore(&ch.head, head + 1u);
  }
  head += 1u;
  __syncwarp();
}

// The walker warp: lane j < WS_PRODUCERS runs the range recurrence over channel j until EXIT.
//  * Words arrive in bursts (one coefficient group at a time), out of step between channels.  The
//    walker is one instruction stream for all its lanes, so it lets work pile up until several
//    channels have a backlog (or a producer is waiting, or a ring is half full) and then takes up to
//    32 words from every channel in one go: few issue slots are spent on mostly idle lanes.
//  * WS_WALK_UNROLL words per lane and round: their loads are issued together, the critical path
//    is one shared-memory round trip plus the dependent range steps (~9 instructions each).
#ifndef CABAC_WS_UNROLL
#define CABAC_WS_UNROLL 4
#endif
#ifndef CABAC_WS_ROUNDS
#define CABAC_WS_ROUNDS 8
#endif
constexpr int WS_WALK_UNROLL = CABAC_WS_UNROLL;
constexpr int WS_WALK_ROUNDS = CABAC_WS_ROUNDS;
#ifndef CABAC_WS_BACKLOG
#define CABAC_WS_BACKLOG 16
#endif
#ifndef CABAC_WS_MIN_READY
#define CABAC_WS_MIN_READY (CABAC_WS_PRODUCERS / 2)
#endif
constexpr uint32_t WS_WALK_BACKLOG = CABAC_WS_BACKLOG;  // words that make a channel "ready"
constexpr int WS_WALK_MIN_READY = CABAC_WS_MIN_READY;
__device__ __forceinline__ void wsWalk(WsChannel *chans, int lane) {
  bool alive = lane < WS_PRODUCERS;
  WsChannel &ch = chans[alive ? lane : 0];
  uint32_t tail = 0, idle = 0, range = 510;
#pragma unroll 1
  while (__any_sync(FULL, alive)) {
    uint32_t avail = 0;
    bool urgent = false;
    if (alive) {
      avail = wsLoad(&ch.head) - tail;
      urgent = avail != 0 && (wsLoad(&ch.want) != 0 || avail >= WS_RING / 2);
    }
    const int ready = __popc(__ballot_sync(FULL, avail >= WS_WALK_BACKLOG));
    if (ready < WS_WALK_MIN_READY && !__any_sync(FULL, urgent)) {
      __nanosleep(128);
      if (++idle > WS_SPIN_LIMIT) __trap();  // a stuck hand-off traps instead of hanging the GPU
      continue;
    }
    idle = 0;
    __threadfence_block();  // head observed above; the ring words below
#pragma unroll 1
    for (int round = 0; round < WS_WALK_ROUNDS && __any_sync(FULL, avail != 0); round++) {
      uint32_t w[WS_WALK_UNROLL];
#pragma unroll
      for (int i = 0; i < WS_WALK_UNROLL; i++) w[i] = ch.ring[(tail + uint32_t(i)) & (WS_RING - 1u)];  // may be stale past head
      const uint32_t n = min(avail, uint32_t(WS_WALK_UNROLL));
#pragma unroll
      for (int i = 0; i < WS_WALK_UNROLL; i++) {
        if (uint32_t(i) < n) {
          ch.rng[(tail + uint32_t(i)) & (WS_RING - 1u)] = uint16_t(range);
          const uint32_t kind = w[i] >> 29;
          // BinProbModel_Std::getLPS + the renormalisation of encodeBin (arith_codec.cpp:359-385), range only
          const uint32_t lps = (((w[i] & 63u) * (range >> 5)) >> 1) + 4;
          const uint32_t nr = (w[i] & 64u) ? lps : range - lps;
          const uint32_t next = nr << (__clz(int(nr | 7u)) - 23);
          range = kind == CABAC_CMD_CTX ? next : range;
          if (kind == WS_KIND_CTL) {  // rare
            if (((w[i] >> 27) & 3u) == WS_OP_START) range = 510;
            else alive = false;  // EXIT is the last word a producer pushes
          }
        }
      }
      tail += n;
      avail -= n;
    }
    __threadfence_block();  // rng[] before the tail that releases it
    if (lane < WS_PRODUCERS) wsStore(&ch.tail, tail);
  }
}

// Producer side: words [d, d + m) of the ring (m <= 32, all passed by the walker) go through the
// warp's coder.  Lane l owns word d + l: shift k and addend c of "low = (low << k) + c" from the word
// and the range it saw; consecutive words whose bit positions fall into the same 16-bit window are
// summed into one (K, C) piece (K <= 39, C < 2^49), and the serial coder takes one piece per step.
// `single`: one piece per word (a substream that is not terminated reports the reference's
// outstanding-byte count, which is only exact when bytes leave as they do there).
__device__ __forceinline__ void wsDrainSlice(LsCoder &rc, WsChannel &ch, uint32_t d, int m, int lane, bool single) {
  uint2 *const kc = ch.pieces;
  const uint32_t slot = (d + uint32_t(lane)) & (WS_RING - 1u);
  const bool active = lane < m;
  const uint32_t w = active ? ch.ring[slot] : (uint32_t(CABAC_CMD_EP) << 29);
  const uint32_t r = ch.rng[slot];
  const bool isCtx = (w >> 29) == CABAC_CMD_CTX;
  const uint32_t lps = (((w & 63u) * (r >> 5)) >> 1) + 4;
  const uint32_t rmps = r - lps;
  const bool isLps = (w & 64u) != 0;
  const uint32_t nr = isLps ? lps : rmps;
  const uint32_t sh = uint32_t(__clz(int(nr | 7u)) - 23);
  const uint32_t k = isCtx ? sh : (w >> 24) & 31u;
  const uint64_t c = isCtx ? (isLps ? uint64_t(rmps) << sh : 0ull) : uint64_t(r) * (w & 0xFFFFFFu);
  // inclusive prefix sum of the shifts
  uint32_t P = k;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(FULL, P, o);
    if (lane >= o) P += t;
  }
  const uint32_t E = P - k;
  const uint32_t g = single ? uint32_t(lane) : E >> CABAC_WS_WINDOW_BITS;
  const uint32_t gPrev = __shfl_up_sync(FULL, g, 1);
  const bool head = active && (lane == 0 || g != gPrev);
  const uint32_t heads = __ballot_sync(FULL, head);
  // last lane of this lane's piece: the one before the next head (or the last active lane)
  const uint32_t above = heads & ~((2u << lane) - 1u);
  const int last = (above ? __ffs(int(above)) - 1 : m) - 1;
  const uint32_t Pend = __shfl_sync(FULL, P, last & 31);
  const uint64_t v = active ? c << (Pend - P) : 0ull;
  // inclusive 64-bit prefix sum of the aligned addends (< 32 * 2^49)
  uint64_t X = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint64_t t = __shfl_up_sync(FULL, X, o);
    if (lane >= o) X += t;
  }
  const uint64_t Xprev = __shfl_up_sync(FULL, X, 1), Xlast = __shfl_sync(FULL, X, last & 31);
  const uint32_t Pprev = __shfl_up_sync(FULL, P, 1);
  if (head) {
    const uint64_t C = Xlast - (lane ? Xprev : 0ull);
    const uint32_t K = Pend - (lane ? Pprev : 0u);
    kc[__popc(heads & ((1u << lane) - 1u))] = make_uint2(uint32_t(C), (uint32_t(C >> 32) << 8) | K);
  }
  __syncwarp();
  const int pieces = __popc(heads);
#pragma unroll 1
  for (int i = 0; i < pieces; i++) {
    const uint2 e = kc[i];
    const uint32_t K = e.y & 0xFFu;
    rc.low = (rc.low << K) + ((uint64_t(e.y >> 8) << 32) | e.x);
    rc.bitsLeft -= int(K);
    rc.flushBits();
  }
  __syncwarp();
}

}  // namespace cabac
