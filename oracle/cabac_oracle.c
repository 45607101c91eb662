/* oracle/cabac_oracle.c - TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C, single-threaded CPU restatement of the reference's CABAC hot path
 * (p-sawicki/entropy_coding, VTM-14-derived).  It exists only as a checker:
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg load it; the
 * product (entropy_coding_b200/) never links, imports or calls it.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this file against
 *   - the compiled reference itself (oracle/_ref, built from /root/reference/src
 *     by oracle/Makefile) on seeded random inputs, when that library is present;
 *   - tests/golden/ (vectors generated from the compiled reference by
 *     tests/golden/make_golden.py, committed), always;
 *   - the known-answer vectors KAT0-KAT3 of SURVEY.md section 8c.
 * The reference's own golden data (test/hashes.txt) needs three QCIF clips
 * that are not in the checkout (.MISSING_LARGE_BLOBS) and cannot be reproduced.
 *
 * Every function cites the reference file:line it restates (paths relative to
 * the reference root).
 */
#include "cabac_oracle.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define CABAC_CTX_TABLE_QUALIFIER static const
#include "ctx_tables.inc"
#include "frac_bits.inc"

/* ------------------------------------------------------------------------- */
/* context-set offsets: positional in src/common/contexts.cpp:77-770          */
/* (checked against the compiled reference through orc_ctx_offsets)           */
/* ------------------------------------------------------------------------- */
static const uint16_t OFS_SIG_GROUP[2] = {86, 88};
static const uint16_t OFS_SIG_FLAG[6] = {90, 102, 110, 122, 130, 142};
static const uint16_t OFS_PAR_FLAG[2] = {150, 171};
static const uint16_t OFS_GTX_FLAG[4] = {182, 203, 214, 235};
static const uint16_t OFS_LAST_X[2] = {246, 266};
static const uint16_t OFS_LAST_Y[2] = {269, 289};
enum {
  OFS_TS_FLAG = 310,
  OFS_TS_SIG_GROUP = 357,
  OFS_TS_SIG_FLAG = 360,
  OFS_TS_PAR_FLAG = 363,
  OFS_TS_GTX_FLAG = 364,
  OFS_TS_LRG1_FLAG = 369,
  OFS_TS_SIGN = 373
};

void orc_ctx_offsets(uint32_t *o) {
  int k = 0;
  for (int i = 0; i < 2; i++) o[k++] = OFS_SIG_GROUP[i];
  for (int i = 0; i < 6; i++) o[k++] = OFS_SIG_FLAG[i];
  for (int i = 0; i < 2; i++) o[k++] = OFS_PAR_FLAG[i];
  for (int i = 0; i < 4; i++) o[k++] = OFS_GTX_FLAG[i];
  for (int i = 0; i < 2; i++) o[k++] = OFS_LAST_X[i];
  for (int i = 0; i < 2; i++) o[k++] = OFS_LAST_Y[i];
  o[k++] = OFS_TS_FLAG;
  o[k++] = OFS_TS_SIG_GROUP;
  o[k++] = OFS_TS_SIG_FLAG;
  o[k++] = OFS_TS_PAR_FLAG;
  o[k++] = OFS_TS_GTX_FLAG;
  o[k++] = OFS_TS_LRG1_FLAG;
  o[k++] = OFS_TS_SIGN;
}

/* ------------------------------------------------------------------------- */
/* probability model: src/common/contexts.{hpp,cpp}                            */
/* ------------------------------------------------------------------------- */
#define MASK_0 0x7FE0u /* contexts.hpp:18-19 */
#define MASK_1 0x7FFEu /* contexts.hpp:20-21 */

typedef struct {
  cabac_ctx_model m[CABAC_NUM_CTX];
  uint8_t rate[CABAC_NUM_CTX];
} CtxTab;

int orc_num_contexts(void) { return CABAC_NUM_CTX; }

void orc_init_table(int row, uint8_t *out) { memcpy(out, CABAC_CTX_INIT_TABLE[row], CABAC_NUM_CTX); }

/* BinProbModel_Std::init, contexts.cpp:893-901; setLog2WindowSize :915-920;
 * CtxStore::init :996-1015 (qp clipped to 0..63). */
static void ctxtab_init(CtxTab *c, int qp, int initId) {
  if (qp < 0) qp = 0;
  if (qp > 63) qp = 63;
  for (int k = 0; k < CABAC_NUM_CTX; k++) {
    const int iv = CABAC_CTX_INIT_TABLE[initId][k];
    const int slope = (iv >> 3) - 4;
    const int offset = ((iv & 7) * 18) + 1;
    int s = ((slope * (qp - 16)) >> 1) + offset;
    s = s < 1 ? 1 : (s > 127 ? 127 : s);
    const unsigned p1 = (unsigned)s << 8;
    c->m[k].state0 = (uint16_t)(p1 & MASK_0);
    c->m[k].state1 = (uint16_t)(p1 & MASK_1);
    const int win = CABAC_CTX_INIT_TABLE[3][k];
    const int r0 = 2 + ((win >> 2) & 3);
    const int r1 = 3 + r0 + (win & 3);
    c->rate[k] = (uint8_t)(16 * r0 + r1);
  }
}

void orc_ctx_init(int qp, int initId, cabac_ctx_model *out, uint8_t *rates) {
  CtxTab c;
  ctxtab_init(&c, qp, initId);
  memcpy(out, c.m, sizeof(c.m));
  if (rates) memcpy(rates, c.rate, sizeof(c.rate));
}

/* state(), mps(), contexts.cpp:939-941 */
static inline unsigned mdl_state(const cabac_ctx_model *m) { return ((unsigned)m->state0 + m->state1) >> 8 & 0xFF; }
static inline unsigned mdl_mps(const cabac_ctx_model *m) { return mdl_state(m) >> 7; }

/* getLPS, contexts.cpp:943-948 (returns uint8_t) */
static inline unsigned mdl_lps(const cabac_ctx_model *m, unsigned range) {
  unsigned q = mdl_state(m);
  if (q & 0x80) q ^= 0xFF;
  return ((((q >> 2) * (range >> 5)) >> 1) + 4) & 0xFF;
}

/* update, contexts.cpp:903-913 */
static inline void mdl_update(cabac_ctx_model *m, unsigned rate, unsigned bin) {
  const unsigned r0 = rate >> 4, r1 = rate & 15;
  unsigned s0 = m->state0, s1 = m->state1;
  s0 -= (s0 >> r0) & MASK_0;
  s1 -= (s1 >> r1) & MASK_1;
  if (bin) {
    s0 += (0x7FFFu >> r0) & MASK_0;
    s1 += (0x7FFFu >> r1) & MASK_1;
  }
  m->state0 = (uint16_t)s0;
  m->state1 = (uint16_t)s1;
}

/* m_RenormTable_32[LPS >> 3], contexts.cpp:787-789, as a function */
static inline int renorm_bits_lps(unsigned lps) {
  int n = 0;
  unsigned v = lps >> 3; /* 0..31 */
  if (v == 0) return 6;
  while ((v << n) < 16) n++;
  return n + 1;
}

/* ------------------------------------------------------------------------- */
/* output bit writer: OutputBitstream::write, src/common/bit_stream.cpp:70-117 */
/* ------------------------------------------------------------------------- */
typedef struct {
  uint8_t *buf;
  long cap, n; /* n may exceed cap: bytes beyond cap are counted, not stored */
  unsigned held, nheld;
} BitW;

static inline void bw_byte(BitW *b, unsigned v) {
  if (b->n < b->cap) b->buf[b->n] = (uint8_t)v;
  b->n++;
}

static void bw_write(BitW *b, uint32_t bits, unsigned nbits) {
  for (int i = (int)nbits - 1; i >= 0; i--) { /* MSB first */
    b->held = (b->held << 1) | ((bits >> i) & 1);
    if (++b->nheld == 8) {
      bw_byte(b, b->held & 0xFF);
      b->held = 0;
      b->nheld = 0;
    }
  }
}

/* writeByteAlignment, bit_stream.cpp:152-155 (+ writeAlignZero :125-132) */
static void bw_byte_alignment(BitW *b) {
  bw_write(b, 1, 1);
  if (b->nheld) bw_write(b, 0, 8 - b->nheld);
}

/* ------------------------------------------------------------------------- */
/* arithmetic encoder: src/entropy_codec/arith_codec.cpp:329-582               */
/* ------------------------------------------------------------------------- */
typedef struct {
  BitW bw;
  uint32_t low, range, bufferedByte;
  int32_t numBuffered, bitsLeft;
  uint32_t nbins;
  /* estimate mode = the reference's BitEstimator_Std behind the same BinEncIf calls
   * (arith_codec.cpp:603-711): no interval, no bytes; 15-bit fixed-point bits are summed */
  int est;
  uint64_t estBits;
  /* BinCounter::addCtx (arith_codec.hpp:72-93) when non-NULL: context-coded bins per context id */
  uint32_t *ctxBins;
} Enc;

/* start, arith_codec.cpp:329-337 */
static void enc_start(Enc *e, uint8_t *out, long cap) {
  memset(e, 0, sizeof(*e));
  e->bw.buf = out;
  e->bw.cap = cap;
  e->low = 0;
  e->range = 510;
  e->bufferedByte = 0xFF;
  e->numBuffered = 0;
  e->bitsLeft = 23;
}

/* writeOut, arith_codec.cpp:524-546 */
static void enc_write_out(Enc *e) {
  const unsigned lead = e->low >> (24 - e->bitsLeft);
  e->bitsLeft += 8;
  e->low &= 0xFFFFFFFFu >> e->bitsLeft;
  if (lead == 0xFF) {
    e->numBuffered++;
  } else if (e->numBuffered > 0) {
    const unsigned carry = lead >> 8;
    bw_byte(&e->bw, (e->bufferedByte + carry) & 0xFF);
    e->bufferedByte = lead & 0xFF;
    const unsigned fill = (0xFF + carry) & 0xFF;
    while (e->numBuffered > 1) {
      bw_byte(&e->bw, fill);
      e->numBuffered--;
    }
  } else {
    e->numBuffered = 1;
    e->bufferedByte = lead;
  }
}

/* TBinEncoder::encodeBin, arith_codec.cpp:553-582 */
static void enc_bin(Enc *e, CtxTab *c, unsigned bin, unsigned ctxId) {
  cabac_ctx_model *m = &c->m[ctxId];
  e->nbins++;
  if (e->ctxBins) e->ctxBins[ctxId]++;
  if (e->est) { /* estFracBitsUpdate, contexts.cpp:922-929: table[state()][bin], then update */
    const unsigned q = ((unsigned)m->state0 + m->state1) >> 8;
    e->estBits += CABAC_FRAC_BITS[2 * q + (bin & 1)];
    mdl_update(m, c->rate[ctxId], bin);
    return;
  }
  const unsigned lps = mdl_lps(m, e->range);
  e->range -= lps;
  if (bin != mdl_mps(m)) {
    const int n = renorm_bits_lps(lps);
    e->bitsLeft -= n;
    e->low += e->range;
    e->low <<= n;
    e->range = lps << n;
    if (e->bitsLeft < 12) enc_write_out(e);
  } else if (e->range < 256) {
    e->bitsLeft -= 1;
    e->low <<= 1;
    e->range <<= 1;
    if (e->bitsLeft < 12) enc_write_out(e);
  }
  mdl_update(m, c->rate[ctxId], bin);
}

/* encodeBinsEP, arith_codec.cpp:401-424 (encodeAlignedBinsEP :491-522 is the
 * same arithmetic with range == 256 and the same 8-bin chunking) */
static void enc_bins_ep(Enc *e, unsigned bins, unsigned n) {
  e->nbins += n;
  if (e->est) { /* estFracBitsEP(numBins) = numBins << SCALE_BITS, contexts.cpp:880-884 */
    e->estBits += (uint64_t)n << 15;
    return;
  }
  /* the aligned variant masks every chunk to its width (:513-514); only visible
   * for out-of-contract input (bins >= 2^n) */
  if (e->range == 256 && n < 32) bins &= (1u << n) - 1;
  while (n > 8) {
    n -= 8;
    const unsigned pat = bins >> n;
    e->low <<= 8;
    e->low += e->range * pat;
    bins -= pat << n;
    e->bitsLeft -= 8;
    if (e->bitsLeft < 12) enc_write_out(e);
  }
  e->low <<= n;
  e->low += e->range * bins;
  e->bitsLeft -= (int)n;
  if (e->bitsLeft < 12) enc_write_out(e);
}

/* encodeBinEP, arith_codec.cpp:389-399 */
static void enc_bin_ep(Enc *e, unsigned bin) { enc_bins_ep(e, bin & 1, 1); }

/* encodeRemAbsEP, arith_codec.cpp:426-458 */
static void enc_rem_abs(Enc *e, unsigned v, unsigned rice, unsigned cutoff, int maxLog2) {
  const unsigned threshold = cutoff << rice;
  if (v < threshold) {
    const unsigned len = (v >> rice) + 1;
    enc_bins_ep(e, (1u << len) - 2, len);
    enc_bins_ep(e, v & ((1u << rice) - 1), rice);
  } else {
    const unsigned maxPrefix = 32 - cutoff - (unsigned)maxLog2;
    unsigned prefix = 0, suffixLen;
    const unsigned code = (v >> rice) - cutoff;
    if (code >= ((1u << maxPrefix) - 1)) {
      prefix = maxPrefix;
      suffixLen = (unsigned)maxLog2;
    } else {
      while (code > ((2u << prefix) - 2u)) prefix++;
      suffixLen = prefix + rice + 1;
    }
    const unsigned total = prefix + cutoff;
    const unsigned suffix = ((code - ((1u << prefix) - 1)) << rice) | (v & ((1u << rice) - 1));
    enc_bins_ep(e, (1u << total) - 1, total);
    enc_bins_ep(e, suffix, suffixLen);
  }
}

/* encodeBinTrm, arith_codec.cpp:460-478 */
static void enc_trm(Enc *e, unsigned bin) {
  e->nbins++;
  if (e->est) { /* estFracBitsTrm, contexts.cpp:931-933 */
    e->estBits += bin ? 0x3bfbb : 0x0010c;
    return;
  }
  e->range -= 2;
  if (bin) {
    e->low += e->range;
    e->low <<= 7;
    e->range = 2 << 7;
    e->bitsLeft -= 7;
  } else if (e->range >= 256) {
    return;
  } else {
    e->low <<= 1;
    e->range <<= 1;
    e->bitsLeft--;
  }
  if (e->bitsLeft < 12) enc_write_out(e);
}

/* finish, arith_codec.cpp:339-357 */
static void enc_finish(Enc *e) {
  if (e->low >> (32 - e->bitsLeft)) {
    bw_byte(&e->bw, (e->bufferedByte + 1) & 0xFF);
    while (e->numBuffered > 1) {
      bw_byte(&e->bw, 0x00);
      e->numBuffered--;
    }
    e->low -= 1u << (32 - e->bitsLeft);
  } else {
    if (e->numBuffered > 0) bw_byte(&e->bw, e->bufferedByte & 0xFF);
    while (e->numBuffered > 1) {
      bw_byte(&e->bw, 0xFF);
      e->numBuffered--;
    }
  }
  bw_write(&e->bw, e->low >> 8, (unsigned)(24 - e->bitsLeft));
}

/* ------------------------------------------------------------------------- */
/* arithmetic decoder: src/entropy_codec/arith_codec.cpp:60-277                */
/* ------------------------------------------------------------------------- */
typedef struct {
  const uint8_t *in;
  long n, pos;
  uint32_t range, value;
  int32_t bitsNeeded;
  uint32_t status;
  uint32_t nbins;
} Dec;

/* InputBitstream::readByte, bit_stream.cpp:268-274 ("FIFO exceeded" -> status) */
static inline unsigned dec_byte(Dec *d) {
  if (d->pos < d->n) return d->in[d->pos++];
  d->pos++;
  d->status |= CABAC_SS_IN_OVERRUN;
  return 0;
}

/* start, arith_codec.cpp:60-66 */
static void dec_start(Dec *d, const uint8_t *in, long n) {
  memset(d, 0, sizeof(*d));
  d->in = in;
  d->n = n;
  d->range = 510;
  d->value = (dec_byte(d) << 8);
  d->value += dec_byte(d);
  d->bitsNeeded = -8;
}

/* TBinDecoder::decodeBin, arith_codec.cpp:242-277 */
static unsigned dec_bin(Dec *d, CtxTab *c, unsigned ctxId) {
  cabac_ctx_model *m = &c->m[ctxId];
  d->nbins++;
  unsigned bin = mdl_mps(m);
  const unsigned lps = mdl_lps(m, d->range);
  d->range -= lps;
  const uint32_t sr = d->range << 7;
  if (d->value < sr) {
    if (d->range < 256) {
      d->range <<= 1;
      d->value <<= 1;
      if (++d->bitsNeeded >= 0) {
        d->value += dec_byte(d) << d->bitsNeeded;
        d->bitsNeeded -= 8;
      }
    }
  } else {
    bin = 1 - bin;
    const int n = renorm_bits_lps(lps);
    d->value -= sr;
    d->value <<= n;
    d->range = lps << n;
    d->bitsNeeded += n;
    if (d->bitsNeeded >= 0) {
      d->value += dec_byte(d) << d->bitsNeeded;
      d->bitsNeeded -= 8;
    }
  }
  mdl_update(m, c->rate[ctxId], bin);
  return bin;
}

/* decodeBinEP, arith_codec.cpp:100-114 */
static unsigned dec_bin_ep(Dec *d) {
  d->nbins++;
  d->value += d->value;
  if (++d->bitsNeeded >= 0) {
    d->value += dec_byte(d);
    d->bitsNeeded = -8;
  }
  const uint32_t sr = d->range << 7;
  if (d->value >= sr) {
    d->value -= sr;
    return 1;
  }
  return 0;
}

/* decodeAlignedBinsEP, arith_codec.cpp:205-235 */
static unsigned dec_aligned_bins_ep(Dec *d, unsigned n) {
  unsigned rem = n, bins = 0;
  while (rem > 0) {
    const unsigned k = rem < 8 ? rem : 8;
    const unsigned mask = (1u << k) - 1;
    const unsigned nb = (d->value >> (15 - k)) & mask;
    bins = (bins << k) | nb;
    d->value = (d->value << k) & 0x7FFF;
    rem -= k;
    d->bitsNeeded += (int)k;
    if (d->bitsNeeded >= 0) {
      d->value |= dec_byte(d) << d->bitsNeeded;
      d->bitsNeeded -= 8;
    }
  }
  return bins;
}

/* decodeBinsEP, arith_codec.cpp:116-151 */
static unsigned dec_bins_ep(Dec *d, unsigned n) {
  d->nbins += n;
  if (d->range == 256) return dec_aligned_bins_ep(d, n);
  unsigned rem = n, bins = 0;
  while (rem > 8) {
    d->value = (d->value << 8) + (dec_byte(d) << (8 + d->bitsNeeded));
    uint32_t sr = d->range << 15;
    for (int i = 0; i < 8; i++) {
      bins += bins;
      sr >>= 1;
      if (d->value >= sr) {
        bins++;
        d->value -= sr;
      }
    }
    rem -= 8;
  }
  d->bitsNeeded += (int)rem;
  d->value <<= rem;
  if (d->bitsNeeded >= 0) {
    d->value += dec_byte(d) << d->bitsNeeded;
    d->bitsNeeded -= 8;
  }
  uint32_t sr = d->range << (rem + 7);
  for (unsigned i = 0; i < rem; i++) {
    bins += bins;
    sr >>= 1;
    if (d->value >= sr) {
      bins++;
      d->value -= sr;
    }
  }
  return bins;
}

/* decodeRemAbsEP, arith_codec.cpp:153-179 */
static unsigned dec_rem_abs(Dec *d, unsigned rice, unsigned cutoff, int maxLog2) {
  unsigned prefix = 0;
  const unsigned maxPrefix = 32 - (unsigned)maxLog2;
  unsigned cw;
  do {
    prefix++;
    cw = dec_bin_ep(d);
  } while (cw && prefix < maxPrefix);
  prefix -= 1 - cw;
  unsigned len = rice, offset;
  if (prefix < cutoff) {
    offset = prefix << rice;
  } else {
    offset = ((1u << (prefix - cutoff)) + cutoff - 1) << rice;
    len += (prefix == maxPrefix) ? (unsigned)maxLog2 - rice : prefix - cutoff;
  }
  return offset + dec_bins_ep(d, len);
}

/* decodeBinTrm, arith_codec.cpp:181-197 */
static unsigned dec_trm(Dec *d) {
  d->nbins++;
  d->range -= 2;
  const uint32_t sr = d->range << 7;
  if (d->value >= sr) return 1;
  if (d->range < 256) {
    d->range += d->range;
    d->value += d->value;
    if (++d->bitsNeeded == 0) {
      d->value += dec_byte(d);
      d->bitsNeeded = -8;
    }
  }
  return 0;
}

/* getNumBitsRead as it counts under ENABLE_TRACING: 8 * bytes read + m_bitsNeeded
 * (arith_codec.cpp:201-203, bit_stream.cpp:268-274) */
static long dec_bits_read(const Dec *d) { return 8 * d->pos + d->bitsNeeded; }

/* ------------------------------------------------------------------------- */
/* ROM: src/common/rom.{hpp,cpp}                                               */
/* ------------------------------------------------------------------------- */
static const uint8_t GROUP_IDX[64] = {/* g_groupIdx, rom.cpp:21-25 */
    0,  1,  2,  3,  4,  4,  5,  5,  6,  6,  6,  6,  7,  7,  7,  7,  8,  8,  8,  8,  8,  8,
    8,  8,  9,  9,  9,  9,  9,  9,  9,  9,  10, 10, 10, 10, 10, 10, 10, 10, 10, 10, 10, 10,
    10, 10, 10, 10, 11, 11, 11, 11, 11, 11, 11, 11, 11, 11, 11, 11, 11, 11, 11, 11};
static const uint8_t MIN_IN_GROUP[14] = {0, 1, 2, 3, 4, 6, 8, 12, 16, 24, 32, 48, 64, 96}; /* rom.cpp:18-19 */
static const uint8_t GO_RICE_PARS[32] = {/* g_goRiceParsCoeff, rom.cpp:27-29 */
    0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 3, 3, 3, 3};
/* g_log2SbbSize[log2w][log2h][0..1], rom.cpp:41-50 (rows 4..7 are equal) */
static const uint8_t LOG2_SBB[5][5][2] = {
    {{0, 0}, {0, 1}, {0, 2}, {0, 3}, {0, 4}},
    {{1, 0}, {1, 1}, {1, 1}, {1, 3}, {1, 3}},
    {{2, 0}, {1, 1}, {2, 2}, {2, 2}, {2, 2}},
    {{3, 0}, {3, 1}, {2, 2}, {2, 2}, {2, 2}},
    {{4, 0}, {3, 1}, {2, 2}, {2, 2}, {2, 2}}};

static void log2_sbb(int log2w, int log2h, int *cw, int *ch) {
  const int a = log2w > 4 ? 4 : log2w, b = log2h > 4 ? 4 : log2h;
  *cw = LOG2_SBB[a][b][0];
  *ch = LOG2_SBB[a][b][1];
}

typedef struct {
  uint16_t idx, x, y;
} ScanEl;

/* Diagonal scan of a bw x bh block: ScanGenerator SCAN_DIAG, rom.cpp:75-93.
 * Every anti-diagonal runs from its bottom-left end to its top-right end. */
static void diag_scan(int bw, int bh, uint8_t *xs, uint8_t *ys) {
  int k = 0;
  for (int d = 0; d <= bw + bh - 2; d++) {
    int y = d < bh ? d : bh - 1;
    int x = d - y;
    while (y >= 0 && x < bw) {
      xs[k] = (uint8_t)x;
      ys[k] = (uint8_t)y;
      k++;
      y--;
      x++;
    }
  }
}

#define MAX_LOG2 6
static ScanEl *g_scan[MAX_LOG2 + 1][MAX_LOG2 + 1];    /* SCAN_GROUPED_4x4 / DIAG */
static uint8_t *g_scanCG[MAX_LOG2 + 1][MAX_LOG2 + 1]; /* SCAN_UNGROUPED / DIAG over CGs: CG raster idx */
static pthread_once_t g_romOnce = PTHREAD_ONCE_INIT;

/* initROM grouped scans, rom.cpp:205-262 (incl. the zero-out fill for 64-sized blocks) */
static void rom_init(void) {
  uint8_t gx[1024], gy[1024], cx[16], cy[16];
  for (int lw = 0; lw <= MAX_LOG2; lw++)
    for (int lh = 0; lh <= MAX_LOG2; lh++) {
      const int w = 1 << lw, h = 1 << lh, total = w * h;
      int lcw, lch;
      log2_sbb(lw, lh, &lcw, &lch);
      const int gw = 1 << lcw, gh = 1 << lch;
      const int wig = (w < 32 ? w : 32) >> lcw, hig = (h < 32 ? h : 32) >> lch;
      ScanEl *s = (ScanEl *)malloc(sizeof(ScanEl) * (size_t)total);
      for (int i = 0; i < total; i++) {
        s[i].idx = (uint16_t)(total - 1);
        s[i].x = (uint16_t)(w - 1);
        s[i].y = (uint16_t)(h - 1);
      }
      diag_scan(wig, hig, gx, gy);
      diag_scan(gw, gh, cx, cy);
      uint8_t *cg = (uint8_t *)malloc((size_t)(wig * hig));
      for (int g = 0; g < wig * hig; g++) {
        cg[g] = (uint8_t)(gy[g] * wig + gx[g]);
        for (int i = 0; i < gw * gh; i++) {
          const int x = gx[g] * gw + cx[i], y = gy[g] * gh + cy[i];
          ScanEl *e = &s[g * gw * gh + i];
          e->x = (uint16_t)x;
          e->y = (uint16_t)y;
          e->idx = (uint16_t)(y * w + x);
        }
      }
      g_scan[lw][lh] = s;
      g_scanCG[lw][lh] = cg;
    }
}

int orc_scan_table(int log2w, int log2h, uint32_t *out) {
  pthread_once(&g_romOnce, rom_init);
  const int total = 1 << (log2w + log2h);
  for (int i = 0; i < total; i++) out[i] = g_scan[log2w][log2h][i].idx;
  int cw, ch;
  log2_sbb(log2w, log2h, &cw, &ch);
  return cw | (ch << 8);
}

/* ------------------------------------------------------------------------- */
/* CoeffCodingContext: src/common/context_modelling.{hpp,cpp}                  */
/* ------------------------------------------------------------------------- */
typedef struct {
  /* constants, context_modelling.cpp:7-79 */
  int comp, chType, w, h, log2w, log2h, log2cgw, log2cgh, log2cg, wig, hig, maxNumCoeff;
  int signHiding, bdpcm, maxLog2TrDR;
  const ScanEl *scan;
  const uint8_t *scanCG;
  int maxLastPosX, maxLastPosY, lastOffsetX, lastOffsetY, lastShiftX, lastShiftY;
  /* state */
  int scanPosLast, subSetId, subSetPos, subSetPosX, subSetPosY, minSubPos, maxSubPos;
  unsigned sigGroupCtxId, sigGroupCtxIdTS;
  int tmplCpSum1, tmplCpDiag;
  int remBins; /* regBinLimit (regular) / m_remainingContextBins (TS) */
  uint64_t sigGroupFlags;
} CCtx;

static int iabs(int v) { return v < 0 ? -v : v; }
static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

static void cctx_init(CCtx *c, const cabac_tu_desc *d, int signHide, int bdpcm) {
  static const int prefixCtx[8] = {0, 0, 0, 3, 6, 10, 15, 21};
  pthread_once(&g_romOnce, rom_init);
  memset(c, 0, sizeof(*c));
  c->comp = d->comp_id;
  c->chType = d->comp_id ? 1 : 0;
  c->log2w = d->log2_w;
  c->log2h = d->log2_h;
  c->w = 1 << c->log2w;
  c->h = 1 << c->log2h;
  log2_sbb(c->log2w, c->log2h, &c->log2cgw, &c->log2cgh);
  c->log2cg = c->log2cgw + c->log2cgh;
  c->wig = imin(32, c->w) >> c->log2cgw;
  c->hig = imin(32, c->h) >> c->log2cgh;
  c->maxNumCoeff = c->w * c->h;
  c->signHiding = signHide;
  c->bdpcm = bdpcm;
  c->maxLog2TrDR = 15; /* no extended precision, slice.hpp:186-191 */
  c->scan = g_scan[c->log2w][c->log2h];
  c->scanCG = g_scanCG[c->log2w][c->log2h];
  c->maxLastPosX = GROUP_IDX[imin(32, c->w) - 1];
  c->maxLastPosY = GROUP_IDX[imin(32, c->h) - 1];
  if (c->chType) {
    c->lastShiftX = imax(0, imin(2, c->w >> 3));
    c->lastShiftY = imax(0, imin(2, c->h >> 3));
  } else {
    c->lastOffsetX = prefixCtx[c->log2w];
    c->lastOffsetY = prefixCtx[c->log2h];
    c->lastShiftX = (c->log2w + 1) >> 2;
    c->lastShiftY = (c->log2h + 1) >> 2;
  }
  c->scanPosLast = -1;
  c->tmplCpSum1 = -1;
  c->tmplCpDiag = -1;
}

/* initSubblock, context_modelling.cpp:81-106 */
static void cctx_init_subblock(CCtx *c, int subSetId, int sigGroupFlag) {
  c->subSetId = subSetId;
  c->subSetPos = c->scanCG[subSetId];
  c->subSetPosY = c->subSetPos / c->wig;
  c->subSetPosX = c->subSetPos - c->subSetPosY * c->wig;
  c->minSubPos = subSetId << c->log2cg;
  c->maxSubPos = c->minSubPos + (1 << c->log2cg) - 1;
  if (sigGroupFlag) c->sigGroupFlags |= 1ull << c->subSetPos;
  const unsigned right = (c->subSetPosX + 1 < c->wig) ? (unsigned)(c->sigGroupFlags >> (c->subSetPos + 1)) & 1 : 0;
  const unsigned lower =
      (c->subSetPosY + 1 < c->hig) ? (unsigned)(c->sigGroupFlags >> (c->subSetPos + c->wig)) & 1 : 0;
  c->sigGroupCtxId = OFS_SIG_GROUP[c->chType] + (right | lower);
  const unsigned left = c->subSetPosX > 0 ? (unsigned)(c->sigGroupFlags >> (c->subSetPos - 1)) & 1 : 0;
  const unsigned above = c->subSetPosY > 0 ? (unsigned)(c->sigGroupFlags >> (c->subSetPos - c->wig)) & 1 : 0;
  c->sigGroupCtxIdTS = OFS_TS_SIG_GROUP + left + above;
}

static int cctx_is_sig_group(const CCtx *c) { return (int)((c->sigGroupFlags >> c->subSetPos) & 1); }

/* sigCtxIdAbs, context_modelling.hpp:71-111 */
static unsigned cctx_sig_ctx(CCtx *c, int scanPos, const int32_t *coeff, int state) {
  const int x = c->scan[scanPos].x, y = c->scan[scanPos].y;
  const int32_t *p = coeff + x + y * c->w;
  int numPos = 0, sumAbs = 0;
#define UPD(v)                        \
  {                                   \
    const int a = iabs(v);            \
    sumAbs += imin(4 + (a & 1), a);   \
    numPos += a != 0;                 \
  }
  if (x < c->w - 1) {
    UPD(p[1]);
    if (x < c->w - 2) UPD(p[2]);
    if (y < c->h - 1) UPD(p[c->w + 1]);
  }
  if (y < c->h - 1) {
    UPD(p[c->w]);
    if (y < c->h - 2) UPD(p[c->w << 1]);
  }
#undef UPD
  const int diag = x + y;
  int ofs = imin((sumAbs + 1) >> 1, 3) + (diag < 2 ? 4 : 0);
  if (c->chType == 0) ofs += diag < 5 ? 4 : 0;
  c->tmplCpDiag = diag;
  c->tmplCpSum1 = sumAbs - numPos;
  return OFS_SIG_FLAG[c->chType + 2 * imax(0, state - 1)] + ofs;
}

/* ctxOffsetAbs, context_modelling.hpp:113-124 */
static unsigned cctx_offset_abs(const CCtx *c) {
  int ofs = 0;
  if (c->tmplCpDiag != -1) {
    ofs = imin(c->tmplCpSum1, 4) + 1;
    if (c->tmplCpDiag == 0)
      ofs += c->chType == 0 ? 15 : 5;
    else if (c->chType == 0)
      ofs += c->tmplCpDiag < 3 ? 10 : (c->tmplCpDiag < 10 ? 5 : 0);
  }
  return (unsigned)ofs;
}

/* templateAbsSum + deriveRice, context_modelling.hpp:135-157, 235-239 */
static unsigned cctx_derive_rice(const CCtx *c, int scanPos, const int32_t *coeff, int baseLevel) {
  const int x = c->scan[scanPos].x, y = c->scan[scanPos].y;
  const int32_t *p = coeff + x + y * c->w;
  int sum = 0;
  if (x < c->w - 1) {
    sum += iabs(p[1]);
    if (x < c->w - 2) sum += iabs(p[2]);
    if (y < c->h - 1) sum += iabs(p[c->w + 1]);
  }
  if (y < c->h - 1) {
    sum += iabs(p[c->w]);
    if (y < c->h - 2) sum += iabs(p[c->w << 1]);
  }
  return GO_RICE_PARS[imax(imin(sum - 5 * baseLevel, 31), 0)];
}

/* TS helpers, context_modelling.hpp:246-385 */
static void cctx_neigh_ts(const CCtx *c, int scanPos, const int32_t *coeff, int *left, int *above) {
  const int x = c->scan[scanPos].x, y = c->scan[scanPos].y;
  const int32_t *p = coeff + x + y * c->w;
  *left = x > 0 ? p[-1] : 0;
  *above = y > 0 ? p[-c->w] : 0;
}
static unsigned cctx_sig_ctx_ts(const CCtx *c, int scanPos, const int32_t *coeff) {
  int l, a;
  cctx_neigh_ts(c, scanPos, coeff, &l, &a);
  return OFS_TS_SIG_FLAG + (l != 0) + (a != 0);
}
static unsigned cctx_lrg1_ctx_ts(const CCtx *c, int scanPos, const int32_t *coeff) {
  int l, a;
  if (c->bdpcm) return OFS_TS_LRG1_FLAG + 3;
  cctx_neigh_ts(c, scanPos, coeff, &l, &a);
  return OFS_TS_LRG1_FLAG + (l != 0) + (a != 0);
}
static unsigned cctx_sign_ctx_ts(const CCtx *c, int scanPos, const int32_t *coeff) {
  int l, a;
  cctx_neigh_ts(c, scanPos, coeff, &l, &a);
  const int sl = (l > 0) - (l < 0), sa = (a > 0) - (a < 0);
  unsigned ctx;
  if ((sl == 0 && sa == 0) || sl * sa < 0)
    ctx = 0;
  else if (sl >= 0 && sa >= 0)
    ctx = 1;
  else
    ctx = 2;
  if (c->bdpcm) ctx += 3;
  return OFS_TS_SIGN + ctx;
}
/* deriveModCoeff :345-366 */
static int cctx_mod_coeff(int left, int above, int absCoeff, int bdpcm) {
  if (absCoeff == 0) return 0;
  int v = absCoeff;
  if (!bdpcm) {
    const int pred = imax(iabs(above), iabs(left));
    if (v == pred)
      v = 1;
    else if (v < pred)
      v = v + 1;
  }
  return v;
}
/* decDeriveModCoeff :368-385 */
static int cctx_dec_mod_coeff(int left, int above, int absCoeff) {
  if (absCoeff == 0) return 0;
  const int pred = imax(iabs(above), iabs(left));
  if (absCoeff == 1 && pred > 0) return pred;
  return absCoeff - (absCoeff <= pred);
}

/* ------------------------------------------------------------------------- */
/* residual coding, encoder: src/entropy_codec/cabac_writer.cpp:2424-3062      */
/* ------------------------------------------------------------------------- */
#define STATE_NEXT(tab, st, par) (((tab) >> (((st) << 2) + (((par)&1) << 1))) & 3) /* cabac_writer.cpp:2808 */

/* last_sig_coeff, cabac_writer.cpp:2639-2722 */
static void w_last_sig_coeff(Enc *e, CtxTab *ct, const CCtx *c, int sbtZero) {
  const unsigned blk = c->scan[c->scanPosLast].idx;
  unsigned posY = blk / (unsigned)c->w, posX = blk - posY * (unsigned)c->w;
  const unsigned gx = GROUP_IDX[posX], gy = GROUP_IDX[posY];
  unsigned maxX = (unsigned)c->maxLastPosX, maxY = (unsigned)c->maxLastPosY;
  if (sbtZero && c->w <= 32 && c->h <= 32 && c->comp == 0) {
    if (c->w == 32) maxX = GROUP_IDX[15];
    if (c->h == 32) maxY = GROUP_IDX[15];
  }
  unsigned k;
  for (k = 0; k < gx; k++) enc_bin(e, ct, 1, OFS_LAST_X[c->chType] + c->lastOffsetX + (k >> c->lastShiftX));
  if (gx < maxX) enc_bin(e, ct, 0, OFS_LAST_X[c->chType] + c->lastOffsetX + (k >> c->lastShiftX));
  for (k = 0; k < gy; k++) enc_bin(e, ct, 1, OFS_LAST_Y[c->chType] + c->lastOffsetY + (k >> c->lastShiftY));
  if (gy < maxY) enc_bin(e, ct, 0, OFS_LAST_Y[c->chType] + c->lastOffsetY + (k >> c->lastShiftY));
  if (gx > 3) {
    posX -= MIN_IN_GROUP[gx];
    for (int i = (int)((gx - 2) >> 1) - 1; i >= 0; i--) enc_bin_ep(e, (posX >> i) & 1);
  }
  if (gy > 3) {
    posY -= MIN_IN_GROUP[gy];
    for (int i = (int)((gy - 2) >> 1) - 1; i >= 0; i--) enc_bin_ep(e, (posY >> i) & 1);
  }
}

/* residual_coding_subblock, cabac_writer.cpp:2724-2872 */
static void w_subblock(Enc *e, CtxTab *ct, CCtx *c, const int32_t *coeff, int stateTab, int *state) {
  const int minSubPos = c->minSubPos;
  const int isLast = (c->scanPosLast >> c->log2cg) == c->subSetId;
  const int firstSigPos = isLast ? c->scanPosLast : c->maxSubPos;
  int nextSigPos = firstSigPos;
  if (!isLast && c->subSetId != 0) {
    if (cctx_is_sig_group(c)) {
      enc_bin(e, ct, 1, c->sigGroupCtxId);
    } else {
      enc_bin(e, ct, 0, c->sigGroupCtxId);
      return;
    }
  }
  unsigned ctxOffset[16];
  const int inferSigPos = nextSigPos != c->scanPosLast ? (c->subSetId != 0 ? minSubPos : -1) : nextSigPos;
  int firstNZPos = nextSigPos, lastNZPos = -1, numNonZero = 0;
  unsigned signPattern = 0;
  int remRegBins = c->remBins;

  for (; nextSigPos >= minSubPos && remRegBins >= 4; nextSigPos--) {
    const int v = coeff[c->scan[nextSigPos].idx];
    const unsigned sig = v != 0;
    if (numNonZero || nextSigPos != inferSigPos) {
      const unsigned ctxId = cctx_sig_ctx(c, nextSigPos, coeff, *state);
      enc_bin(e, ct, sig, ctxId);
      remRegBins--;
    } else if (nextSigPos != c->scanPosLast) {
      cctx_sig_ctx(c, nextSigPos, coeff, *state);
    }
    if (sig) {
      const unsigned ofs = ctxOffset[nextSigPos - minSubPos] = cctx_offset_abs(c);
      numNonZero++;
      firstNZPos = nextSigPos;
      lastNZPos = imax(lastNZPos, nextSigPos);
      int rem = iabs(v) - 1;
      if (nextSigPos != c->scanPosLast) signPattern <<= 1;
      if (v < 0) signPattern++;
      const unsigned gt1 = rem != 0;
      enc_bin(e, ct, gt1, OFS_GTX_FLAG[c->chType + 2] + ofs);
      remRegBins--;
      if (gt1) {
        rem -= 1;
        enc_bin(e, ct, (unsigned)rem & 1, OFS_PAR_FLAG[c->chType] + ofs);
        rem >>= 1;
        remRegBins--;
        enc_bin(e, ct, rem != 0, OFS_GTX_FLAG[c->chType] + ofs);
        remRegBins--;
      }
    }
    *state = STATE_NEXT(stateTab, *state, v);
  }
  const int firstPosMode2 = nextSigPos;
  c->remBins = remRegBins;

  /* 2nd pass: Golomb-Rice remainders of the context-coded positions */
  for (int sp = firstSigPos; sp > firstPosMode2; sp--) {
    const unsigned rice = cctx_derive_rice(c, sp, coeff, 4);
    const unsigned a = (unsigned)iabs(coeff[c->scan[sp].idx]);
    if (a >= 4) enc_rem_abs(e, (a - 4) >> 1, rice, 5, c->maxLog2TrDR);
  }
  /* bypass-only positions */
  for (int sp = firstPosMode2; sp >= minSubPos; sp--) {
    const int v = coeff[c->scan[sp].idx];
    const unsigned a = (unsigned)iabs(v);
    const unsigned rice = cctx_derive_rice(c, sp, coeff, 0);
    const unsigned pos0 = (*state < 2 ? 1u : 2u) << rice; /* g_goRicePosCoeff0, rom.hpp:25-27 */
    const unsigned rem = a == 0 ? pos0 : (a <= pos0 ? a - 1 : a);
    enc_rem_abs(e, rem, rice, 5, c->maxLog2TrDR);
    *state = STATE_NEXT(stateTab, *state, a);
    if (a) {
      numNonZero++;
      firstNZPos = sp;
      lastNZPos = imax(lastNZPos, sp);
      signPattern <<= 1;
      if (v < 0) signPattern++;
    }
  }
  unsigned numSigns = (unsigned)numNonZero;
  if (c->signHiding && lastNZPos - firstNZPos >= 4) { /* SBH_THRESHOLD, common_def.hpp:96 */
    numSigns--;
    signPattern >>= 1;
  }
  enc_bins_ep(e, signPattern, numSigns);
}

/* residual_coding_subblockTS, cabac_writer.cpp:2925-3062 */
static void w_subblock_ts(Enc *e, CtxTab *ct, CCtx *c, const int32_t *coeff) {
  const int lastPos = c->maxSubPos; /* the reference names this minSubPos */
  const int firstSigPos = c->minSubPos;
  int nextSigPos = firstSigPos;
  const int lastSubSet = (c->maxNumCoeff - 1) >> c->log2cg;
  const int isLastSubSet = lastSubSet == c->subSetId;
  /* only1stSigGroup(): no CG flagged except possibly the last one, context_modelling.hpp:24-26 */
  const uint64_t others = c->sigGroupFlags & ~(1ull << lastSubSet);
  if (!isLastSubSet || others != 0) {
    if (cctx_is_sig_group(c)) {
      enc_bin(e, ct, 1, c->sigGroupCtxIdTS);
    } else {
      enc_bin(e, ct, 0, c->sigGroupCtxIdTS);
      return;
    }
  }
  const int inferSigPos = lastPos;
  int numNonZero = 0, l, a;
  int lastScanPosPass1 = -1, lastScanPosPass2 = -1;
  for (; nextSigPos <= lastPos && c->remBins >= 4; nextSigPos++) {
    const int v = coeff[c->scan[nextSigPos].idx];
    const unsigned sig = v != 0;
    if (numNonZero || nextSigPos != inferSigPos) {
      enc_bin(e, ct, sig, cctx_sig_ctx_ts(c, nextSigPos, coeff));
      c->remBins--;
    }
    if (sig) {
      enc_bin(e, ct, v < 0, cctx_sign_ctx_ts(c, nextSigPos, coeff));
      c->remBins--;
      numNonZero++;
      cctx_neigh_ts(c, nextSigPos, coeff, &l, &a);
      int rem = cctx_mod_coeff(l, a, iabs(v), c->bdpcm) - 1;
      const unsigned gt1 = rem != 0;
      enc_bin(e, ct, gt1, cctx_lrg1_ctx_ts(c, nextSigPos, coeff));
      c->remBins--;
      if (gt1) {
        rem -= 1;
        enc_bin(e, ct, (unsigned)rem & 1, OFS_TS_PAR_FLAG);
        c->remBins--;
      }
    }
    lastScanPosPass1 = nextSigPos;
  }
  for (int sp = firstSigPos; sp <= lastPos && c->remBins >= 4; sp++) {
    cctx_neigh_ts(c, sp, coeff, &l, &a);
    const int absLevel = cctx_mod_coeff(l, a, iabs(coeff[c->scan[sp].idx]), c->bdpcm);
    int cutoff = 2;
    for (int i = 0; i < 4; i++) {
      if (absLevel >= cutoff) {
        enc_bin(e, ct, absLevel >= cutoff + 2, OFS_TS_GTX_FLAG + (cutoff >> 1));
        c->remBins--;
      }
      cutoff += 2;
    }
    lastScanPosPass2 = sp;
  }
  for (int sp = firstSigPos; sp <= lastPos; sp++) {
    cctx_neigh_ts(c, sp, coeff, &l, &a);
    const int cutoff = sp <= lastScanPosPass2 ? 10 : (sp <= lastScanPosPass1 ? 2 : 0);
    const int v = coeff[c->scan[sp].idx];
    const int absLevel = cctx_mod_coeff(l, a, iabs(v), c->bdpcm || !cutoff);
    if (absLevel >= cutoff) {
      const unsigned rem = sp <= lastScanPosPass1 ? (unsigned)(absLevel - cutoff) >> 1 : (unsigned)absLevel;
      enc_rem_abs(e, rem, 1, 5, c->maxLog2TrDR);
      if (absLevel && sp > lastScanPosPass1) enc_bin_ep(e, v < 0);
    }
  }
}

/* residual_coding (+ ts_flag, residual_codingTS), cabac_writer.cpp:2424-2535, 2874-2923.
 * Returns CABAC_SS_* status bits. */
static unsigned w_residual_coding(Enc *e, CtxTab *ct, const cabac_tu_desc *d, const int32_t *coeffBase,
                                  unsigned subFlags, uint8_t *tuFlagOut) {
  const int32_t *coeff = coeffBase + d->coeff_offset;
  const int ts = (d->flags & CABAC_TU_TS) != 0;
  const int isLuma = d->comp_id == 0;
  unsigned res = ts ? CABAC_TUR_TS : 0;
  if (d->flags & CABAC_TU_TS_ALLOWED) enc_bin(e, ct, (unsigned)ts, OFS_TS_FLAG + (isLuma ? 0 : 1));
  CCtx c;
  if (ts && !(subFlags & CABAC_SUB_TSRC_DISABLED)) {
    cctx_init(&c, d, 0, (d->flags & CABAC_TU_BDPCM) != 0);
    c.remBins = (c.maxNumCoeff * 7) >> 2;
    uint64_t sig = 0;
    for (int sp = 0; sp < c.maxNumCoeff; sp++)
      if (coeff[c.scan[sp].idx]) sig |= 1ull << (sp >> c.log2cg);
    for (int s = 0; s <= (c.maxNumCoeff - 1) >> c.log2cg; s++) {
      cctx_init_subblock(&c, s, (int)((sig >> s) & 1));
      w_subblock_ts(e, ct, &c, coeff);
    }
    if (tuFlagOut) *tuFlagOut = (uint8_t)res;
    return 0;
  }
  cctx_init(&c, d, (subFlags & CABAC_SUB_SIGN_HIDING) != 0, 0);
  int scanPosLast = -1;
  uint64_t sig = 0;
  for (int sp = 0; sp < c.maxNumCoeff; sp++)
    if (coeff[c.scan[sp].idx]) {
      scanPosLast = sp;
      if ((sp >> c.log2cg) < 64) sig |= 1ull << (sp >> c.log2cg);
    }
  if (scanPosLast < 0) return CABAC_SS_EMPTY_TU;
  if ((scanPosLast >> c.log2cg) >= c.wig * c.hig) return CABAC_SS_COEFF_RANGE; /* zero-out region not zero */
  c.scanPosLast = scanPosLast;
  /* CUCtx side effects, cabac_writer.cpp:2460-2478 */
  if (!ts && c.w >= 4 && c.h >= 4) {
    const int maxLfnstPos = ((c.h == 4 && c.w == 4) || (c.h == 8 && c.w == 8)) ? 7 : 15;
    if (scanPosLast > maxLfnstPos) res |= CABAC_TUR_VIOLATES_LFNST;
    if (scanPosLast >= 1) res |= CABAC_TUR_LFNST_LAST_POS;
  }
  if (isLuma && !ts && scanPosLast >= 1) res |= CABAC_TUR_MTS_LAST_POS;
  const int sbtZero = (d->flags & CABAC_TU_SBT_ZEROOUT) && c.w <= 32 && c.h <= 32 && isLuma;
  w_last_sig_coeff(e, ct, &c, sbtZero);
  const int stateTab = (subFlags & CABAC_SUB_DEP_QUANT) ? 32040 : 0;
  int state = 0;
  {
    /* getTbAreaAfterCoefZeroOut, unit.cpp:465-478; x28 >> 4, common_def.hpp:281-282 */
    int zw = c.w, zh = c.h;
    if (sbtZero) {
      if (c.w == 32) zw = 16;
      if (c.h == 32) zh = 16;
    }
    c.remBins = (imin(32, zw) * imin(32, zh) * 28) >> 4;
  }
  for (int s = scanPosLast >> c.log2cg; s >= 0; s--) {
    cctx_init_subblock(&c, s, (int)((sig >> s) & 1));
    if (sbtZero) {
      if ((c.h == 32 && c.subSetPosY >= (16 >> c.log2cgh)) || (c.w == 32 && c.subSetPosX >= (16 >> c.log2cgw)))
        continue;
    }
    w_subblock(e, ct, &c, coeff, stateTab, &state);
    if (isLuma && cctx_is_sig_group(&c) && (c.subSetPosY > 3 || c.subSetPosX > 3)) res |= CABAC_TUR_VIOLATES_MTS;
  }
  if (tuFlagOut) *tuFlagOut = (uint8_t)res;
  return 0;
}

/* ------------------------------------------------------------------------- */
/* residual coding, decoder: src/entropy_codec/cabac_reader.cpp:2647-3339      */
/* ------------------------------------------------------------------------- */
/* last_sig_coeff, cabac_reader.cpp:2865-2940 */
static int r_last_sig_coeff(Dec *d, CtxTab *ct, const CCtx *c, int sbtZero) {
  unsigned posX = 0, posY = 0;
  unsigned maxX = (unsigned)c->maxLastPosX, maxY = (unsigned)c->maxLastPosY;
  if (sbtZero && c->w <= 32 && c->h <= 32 && c->comp == 0) {
    if (c->w == 32) maxX = GROUP_IDX[15];
    if (c->h == 32) maxY = GROUP_IDX[15];
  }
  for (; posX < maxX; posX++)
    if (!dec_bin(d, ct, OFS_LAST_X[c->chType] + c->lastOffsetX + (posX >> c->lastShiftX))) break;
  for (; posY < maxY; posY++)
    if (!dec_bin(d, ct, OFS_LAST_Y[c->chType] + c->lastOffsetY + (posY >> c->lastShiftY))) break;
  if (posX > 3) {
    unsigned t = 0;
    const int cnt = (int)((posX - 2) >> 1);
    for (int i = cnt - 1; i >= 0; i--) t += dec_bin_ep(d) << i;
    posX = MIN_IN_GROUP[posX] + t;
  }
  if (posY > 3) {
    unsigned t = 0;
    const int cnt = (int)((posY - 2) >> 1);
    for (int i = cnt - 1; i >= 0; i--) t += dec_bin_ep(d) << i;
    posY = MIN_IN_GROUP[posY] + t;
  }
  const int blk = (int)(posX + posY * (unsigned)c->w);
  int sp = 0;
  for (; sp < c->maxNumCoeff - 1; sp++)
    if (blk == c->scan[sp].idx) break;
  return sp;
}

static void check_conformance(Dec *d, int v) { /* cabac_reader.cpp:2942-2946 */
  if (v < -32768 || v > 32767) d->status |= CABAC_SS_COEFF_RANGE;
}

/* residual_coding_subblock, cabac_reader.cpp:2948-3128 */
static void r_subblock(Dec *d, CtxTab *ct, CCtx *c, int32_t *coeff, int stateTab, int *state) {
  const int minSubPos = c->minSubPos;
  const int isLast = (c->scanPosLast >> c->log2cg) == c->subSetId;
  const int firstSigPos = isLast ? c->scanPosLast : c->maxSubPos;
  int nextSigPos = firstSigPos;
  int sigGroup = isLast || !minSubPos;
  if (!sigGroup) sigGroup = (int)dec_bin(d, ct, c->sigGroupCtxId);
  if (sigGroup)
    c->sigGroupFlags |= 1ull << c->subSetPos;
  else
    return;
  unsigned ctxOffset[16];
  const int inferSigPos = nextSigPos != c->scanPosLast ? (c->subSetId != 0 ? minSubPos : -1) : nextSigPos;
  int firstNZPos = nextSigPos, lastNZPos = -1, numNonZero = 0;
  int remRegBins = c->remBins;
  int sigBlkPos[16];
  for (; nextSigPos >= minSubPos && remRegBins >= 4; nextSigPos--) {
    const int blk = c->scan[nextSigPos].idx;
    unsigned sig = (!numNonZero && nextSigPos == inferSigPos);
    if (!sig) {
      sig = dec_bin(d, ct, cctx_sig_ctx(c, nextSigPos, coeff, *state));
      remRegBins--;
    } else if (nextSigPos != c->scanPosLast) {
      cctx_sig_ctx(c, nextSigPos, coeff, *state);
    }
    if (sig) {
      const unsigned ofs = ctxOffset[nextSigPos - minSubPos] = cctx_offset_abs(c);
      sigBlkPos[numNonZero++] = blk;
      firstNZPos = nextSigPos;
      lastNZPos = imax(lastNZPos, nextSigPos);
      const unsigned gt1 = dec_bin(d, ct, OFS_GTX_FLAG[c->chType + 2] + ofs);
      remRegBins--;
      unsigned par = 0, gt2 = 0;
      if (gt1) {
        par = dec_bin(d, ct, OFS_PAR_FLAG[c->chType] + ofs);
        remRegBins--;
        gt2 = dec_bin(d, ct, OFS_GTX_FLAG[c->chType] + ofs);
        remRegBins--;
      }
      coeff[blk] += (int)(1 + par + gt1 + (gt2 << 1));
    }
    *state = STATE_NEXT(stateTab, *state, coeff[blk]);
  }
  const int firstPosMode2 = nextSigPos;
  c->remBins = remRegBins;
  for (int sp = firstSigPos; sp > firstPosMode2; sp--) {
    const unsigned rice = cctx_derive_rice(c, sp, coeff, 4);
    int32_t *t = &coeff[c->scan[sp].idx];
    if (*t >= 4) *t += (int)(dec_rem_abs(d, rice, 5, c->maxLog2TrDR) << 1);
  }
  for (int sp = firstPosMode2; sp >= minSubPos; sp--) {
    const unsigned rice = cctx_derive_rice(c, sp, coeff, 0);
    const unsigned pos0 = (*state < 2 ? 1u : 2u) << rice;
    const unsigned rem = dec_rem_abs(d, rice, 5, c->maxLog2TrDR);
    const int t = (int)(rem == pos0 ? 0 : (rem < pos0 ? rem + 1 : rem));
    *state = STATE_NEXT(stateTab, *state, t);
    if (t) {
      const int blk = c->scan[sp].idx;
      sigBlkPos[numNonZero++] = blk;
      firstNZPos = sp;
      lastNZPos = imax(lastNZPos, sp);
      coeff[blk] = t;
    }
  }
  const unsigned numSigns =
      (c->signHiding && lastNZPos - firstNZPos >= 4) ? (unsigned)numNonZero - 1 : (unsigned)numNonZero;
  /* "<< (32 - numSigns)" with numSigns == 0 is UB in the reference (cabac_reader.cpp:3109);
   * decodeBinsEP(0) is 0 so the pattern is defined as 0. */
  unsigned signPattern = numSigns ? dec_bins_ep(d, numSigns) << (32 - numSigns) : dec_bins_ep(d, 0);
  int sumAbs = 0;
  for (unsigned k = 0; k < numSigns; k++) {
    const int a = coeff[sigBlkPos[k]];
    sumAbs += a;
    coeff[sigBlkPos[k]] = (signPattern & (1u << 31)) ? -a : a;
    signPattern <<= 1;
    check_conformance(d, coeff[sigBlkPos[k]]);
  }
  if ((unsigned)numNonZero > numSigns) {
    const int a = coeff[sigBlkPos[numSigns]];
    sumAbs += a;
    coeff[sigBlkPos[numSigns]] = (sumAbs & 1) ? -a : a;
    check_conformance(d, coeff[sigBlkPos[numSigns]]);
  }
}

/* residual_coding_subblockTS, cabac_reader.cpp:3153-3339 */
static void r_subblock_ts(Dec *d, CtxTab *ct, CCtx *c, int32_t *coeff) {
  const int lastPos = c->maxSubPos;
  const int firstSigPos = c->minSubPos;
  int nextSigPos = firstSigPos;
  unsigned signPattern = 0;
  const int lastSubSet = (c->maxNumCoeff - 1) >> c->log2cg;
  int sigGroup = (lastSubSet == c->subSetId) && c->sigGroupFlags == 0;
  if (!sigGroup) sigGroup = (int)dec_bin(d, ct, c->sigGroupCtxIdTS);
  if (sigGroup)
    c->sigGroupFlags |= 1ull << c->subSetPos;
  else
    return;
  const int inferSigPos = lastPos;
  int numNonZero = 0;
  int sigBlkPos[16];
  int lastScanPosPass1 = -1, lastScanPosPass2 = -1;
  for (; nextSigPos <= lastPos && c->remBins >= 4; nextSigPos++) {
    const int blk = c->scan[nextSigPos].idx;
    unsigned sig = (!numNonZero && nextSigPos == inferSigPos);
    if (!sig) {
      sig = dec_bin(d, ct, cctx_sig_ctx_ts(c, nextSigPos, coeff));
      c->remBins--;
    }
    if (sig) {
      const unsigned sign = dec_bin(d, ct, cctx_sign_ctx_ts(c, nextSigPos, coeff));
      c->remBins--;
      signPattern += sign << numNonZero;
      sigBlkPos[numNonZero++] = blk;
      const unsigned gt1 = dec_bin(d, ct, cctx_lrg1_ctx_ts(c, nextSigPos, coeff));
      c->remBins--;
      unsigned par = 0;
      if (gt1) {
        par = dec_bin(d, ct, OFS_TS_PAR_FLAG);
        c->remBins--;
      }
      coeff[blk] = (sign ? -1 : 1) * (int)(1 + par + gt1);
    }
    lastScanPosPass1 = nextSigPos;
  }
  for (int sp = firstSigPos; sp <= lastPos && c->remBins >= 4; sp++) {
    int32_t *t = &coeff[c->scan[sp].idx];
    int cutoff = 2;
    for (int i = 0; i < 4; i++) {
      if (*t < 0) *t = -*t;
      if (*t >= cutoff) {
        const unsigned gt2 = dec_bin(d, ct, OFS_TS_GTX_FLAG + (cutoff >> 1));
        *t += (int)(gt2 << 1);
        c->remBins--;
      }
      cutoff += 2;
    }
    lastScanPosPass2 = sp;
  }
  for (int sp = firstSigPos; sp <= lastPos; sp++) {
    int32_t *t = &coeff[c->scan[sp].idx];
    const int cutoff = sp <= lastScanPosPass2 ? 10 : (sp <= lastScanPosPass1 ? 2 : 0);
    if (*t < 0) *t = -*t;
    if (*t >= cutoff) {
      const int rem = (int)dec_rem_abs(d, 1, 5, c->maxLog2TrDR);
      *t += (sp <= lastScanPosPass1) ? (rem << 1) : rem;
      if (*t && sp > lastScanPosPass1) {
        const unsigned sign = dec_bin_ep(d);
        signPattern += sign << numNonZero;
        sigBlkPos[numNonZero++] = c->scan[sp].idx;
      }
    }
    if (!c->bdpcm && cutoff) {
      if (*t > 0) {
        int l, a;
        cctx_neigh_ts(c, sp, coeff, &l, &a);
        *t = cctx_dec_mod_coeff(l, a, *t);
      }
    }
  }
  for (int k = 0; k < numNonZero; k++) {
    const int a = coeff[sigBlkPos[k]];
    coeff[sigBlkPos[k]] = (signPattern & 1) ? -a : a;
    signPattern >>= 1;
    check_conformance(d, coeff[sigBlkPos[k]]);
  }
}

/* residual_coding (+ ts_flag, residual_codingTS), cabac_reader.cpp:2647-2752, 3130-3151 */
static void r_residual_coding(Dec *dc, CtxTab *ct, const cabac_tu_desc *d, int32_t *coeffBase, unsigned subFlags,
                              uint8_t *tuFlagOut) {
  int32_t *coeff = coeffBase + d->coeff_offset;
  const int isLuma = d->comp_id == 0;
  int ts = (d->flags & CABAC_TU_BDPCM) ? 1 : ((d->flags & CABAC_TU_TS) != 0);
  if (d->flags & CABAC_TU_TS_ALLOWED) ts = (int)dec_bin(dc, ct, OFS_TS_FLAG + (isLuma ? 0 : 1));
  unsigned res = ts ? CABAC_TUR_TS : 0;
  CCtx c;
  if (ts && !(subFlags & CABAC_SUB_TSRC_DISABLED)) {
    cctx_init(&c, d, 0, (d->flags & CABAC_TU_BDPCM) != 0);
    c.remBins = (c.maxNumCoeff * 7) >> 2;
    for (int s = 0; s <= (c.maxNumCoeff - 1) >> c.log2cg; s++) {
      cctx_init_subblock(&c, s, 0);
      r_subblock_ts(dc, ct, &c, coeff);
    }
    if (tuFlagOut) *tuFlagOut = (uint8_t)res;
    return;
  }
  cctx_init(&c, d, (subFlags & CABAC_SUB_SIGN_HIDING) != 0, 0);
  const int sbtZero = (d->flags & CABAC_TU_SBT_ZEROOUT) && c.w <= 32 && c.h <= 32 && isLuma;
  c.scanPosLast = r_last_sig_coeff(dc, ct, &c, sbtZero);
  if ((c.scanPosLast >> c.log2cg) >= c.wig * c.hig) { /* position inside the zero-out region */
    dc->status |= CABAC_SS_COEFF_RANGE;
    if (tuFlagOut) *tuFlagOut = (uint8_t)res;
    return;
  }
  if (!ts && c.w >= 4 && c.h >= 4) {
    const int maxLfnstPos = ((c.h == 4 && c.w == 4) || (c.h == 8 && c.w == 8)) ? 7 : 15;
    if (c.scanPosLast > maxLfnstPos) res |= CABAC_TUR_VIOLATES_LFNST;
    if (c.scanPosLast >= 1) res |= CABAC_TUR_LFNST_LAST_POS;
  }
  if (isLuma && !ts && c.scanPosLast >= 1) res |= CABAC_TUR_MTS_LAST_POS;
  const int stateTab = (subFlags & CABAC_SUB_DEP_QUANT) ? 32040 : 0;
  int state = 0;
  {
    int zw = c.w, zh = c.h;
    if (sbtZero) {
      if (c.w == 32) zw = 16;
      if (c.h == 32) zh = 16;
    }
    c.remBins = (imin(32, zw) * imin(32, zh) * 28) >> 4;
  }
  for (int s = c.scanPosLast >> c.log2cg; s >= 0; s--) {
    cctx_init_subblock(&c, s, 0);
    if (sbtZero) {
      if ((c.h == 32 && c.subSetPosY >= (16 >> c.log2cgh)) || (c.w == 32 && c.subSetPosX >= (16 >> c.log2cgw)))
        continue;
    }
    r_subblock(dc, ct, &c, coeff, stateTab, &state);
    if (isLuma && cctx_is_sig_group(&c) && (c.subSetPosY > 3 || c.subSetPosX > 3)) res |= CABAC_TUR_VIOLATES_MTS;
  }
  if (tuFlagOut) *tuFlagOut = (uint8_t)res;
}

/* ------------------------------------------------------------------------- */
/* flat entry points                                                           */
/* ------------------------------------------------------------------------- */
static void ctx_begin(CtxTab *ct, const cabac_substream_params *p, const cabac_ctx_model *ctxIn) {
  ctxtab_init(ct, p->qp, p->init_id); /* reset(qp, initId), arith_codec.cpp:367-370 / 75-78 */
  if ((p->flags & CABAC_SUB_CTX_IN) && ctxIn) memcpy(ct->m, ctxIn, sizeof(ct->m));
}

static long enc_end(Enc *e, const cabac_substream_params *p, const CtxTab *ct, uint32_t *numBins,
                    cabac_ctx_model *ctxOut) {
  if (p->flags & CABAC_SUB_TERMINATE) { /* end_of_slice, cabac_writer.cpp:104-107 */
    enc_trm(e, 1);
    enc_finish(e);
  }
  if (p->flags & CABAC_SUB_ALIGN) bw_byte_alignment(&e->bw);
  if (numBins) *numBins = e->nbins;
  if (ctxOut) memcpy(ctxOut, ct->m, sizeof(ct->m));
  return e->bw.n;
}

long orc_encode_bins(const uint32_t *cmds, long ncmd, const cabac_substream_params *p,
                     const cabac_ctx_model *ctxIn, uint8_t *out, long cap, uint32_t *numBins,
                     cabac_ctx_model *ctxOut) {
  CtxTab ct;
  Enc e;
  ctx_begin(&ct, p, ctxIn);
  enc_start(&e, out, cap);
  for (long i = 0; i < ncmd; i++) {
    const uint32_t w = cmds[i];
    switch (CABAC_CMD_KIND(w)) {
      case CABAC_CMD_CTX: enc_bin(&e, &ct, (w >> 9) & 1, w & 511); break;
      case CABAC_CMD_EP: enc_bins_ep(&e, w & 0xFFFFFF, (w >> 24) & 31); break;
      case CABAC_CMD_REM: enc_rem_abs(&e, w & 0xFFFFFF, (w >> 24) & 31, 5, 15); break;
      case CABAC_CMD_TRM: enc_trm(&e, w & 1); break;
      case CABAC_CMD_ALIGN: e.range = 256; break; /* align(), arith_codec.cpp:480 */
      default: return -1;
    }
  }
  return enc_end(&e, p, &ct, numBins, ctxOut);
}

long orc_decode_bins(const uint8_t *in, long nbytes, const uint32_t *cmds, long ncmd,
                     const cabac_substream_params *p, const cabac_ctx_model *ctxIn, uint32_t *values,
                     cabac_ctx_model *ctxOut, uint32_t *status) {
  CtxTab ct;
  Dec d;
  ctx_begin(&ct, p, ctxIn);
  dec_start(&d, in, nbytes);
  for (long i = 0; i < ncmd; i++) {
    const uint32_t w = cmds[i];
    uint32_t v = 0;
    switch (CABAC_CMD_KIND(w)) {
      case CABAC_CMD_CTX: v = dec_bin(&d, &ct, w & 511); break;
      case CABAC_CMD_EP: v = dec_bins_ep(&d, (w >> 24) & 31); break;
      case CABAC_CMD_REM: v = dec_rem_abs(&d, (w >> 24) & 31, 5, 15); break;
      case CABAC_CMD_TRM: v = dec_trm(&d); break;
      case CABAC_CMD_ALIGN: d.range = 256; break;
      default: d.status |= CABAC_SS_BAD_CMD; break;
    }
    if (values) values[i] = v;
  }
  if (p->flags & CABAC_SUB_TERMINATE)
    if (!dec_trm(&d)) d.status |= CABAC_SS_NO_TERMINATE;
  if (ctxOut) memcpy(ctxOut, ct.m, sizeof(ct.m));
  if (status) *status = d.status;
  return dec_bits_read(&d);
}

long orc_encode_residual(const cabac_tu_desc *tus, long ntu, const cabac_substream_params *p,
                         const int32_t *coeffs, const cabac_ctx_model *ctxIn, uint8_t *out, long cap,
                         uint32_t *numBins, uint8_t *tuFlags, cabac_ctx_model *ctxOut, int useMTS) {
  (void)useMTS; /* SBT zero-out is carried by the descriptor flag */
  CtxTab ct;
  Enc e;
  ctx_begin(&ct, p, ctxIn);
  enc_start(&e, out, cap);
  for (long t = 0; t < ntu; t++) {
    const unsigned st = w_residual_coding(&e, &ct, &tus[t], coeffs, p->flags, tuFlags ? &tuFlags[t] : NULL);
    if (st) return -(long)st;
  }
  return enc_end(&e, p, &ct, numBins, ctxOut);
}

/* m_binFracBits[256][2] flattened (contexts.cpp:791-878), generated by gen_tables.py */
void orc_frac_bits(uint32_t *out) { memcpy(out, CABAC_FRAC_BITS, sizeof(CABAC_FRAC_BITS)); }

/* RDO bit estimate of a TU list: CABACWriter::residual_coding on BitEstimator_Std.  Returns
 * getEstFracBits(); estPerTu[t] = the running estimate after TU t. */
uint64_t orc_estimate_residual(const cabac_tu_desc *tus, long ntu, const cabac_substream_params *p,
                               const int32_t *coeffs, const cabac_ctx_model *ctxIn, cabac_ctx_model *ctxOut,
                               uint64_t *estPerTu, int useMTS) {
  CtxTab ct;
  (void)useMTS; /* the reference harness needs it to set up its SPS; the TU descriptors carry the effect */
  Enc e;
  uint8_t dummy[8];
  ctx_begin(&ct, p, ctxIn);
  enc_start(&e, dummy, 0);
  e.est = 1;
  for (long t = 0; t < ntu; t++) {
    if (w_residual_coding(&e, &ct, &tus[t], coeffs, p->flags, NULL)) return ~0ull;
    if (estPerTu) estPerTu[t] = e.estBits;
  }
  if (ctxOut) memcpy(ctxOut, ct.m, sizeof(ct.m));
  return e.estBits;
}

/* BinCounter view of a TU list: ctxBins[k] = getNumBins(k) (arith_codec.cpp:281-316) after
 * residual_coding over the list; the contexts evolve as in orc_encode_residual. */
long orc_residual_ctx_bins(const cabac_tu_desc *tus, long ntu, const cabac_substream_params *p,
                           const int32_t *coeffs, const cabac_ctx_model *ctxIn, cabac_ctx_model *ctxOut,
                           uint32_t *ctxBins) {
  CtxTab ct;
  Enc e;
  uint8_t dummy[8];
  ctx_begin(&ct, p, ctxIn);
  enc_start(&e, dummy, 0);
  e.est = 1;
  memset(ctxBins, 0, sizeof(uint32_t) * CABAC_NUM_CTX);
  e.ctxBins = ctxBins;
  for (long t = 0; t < ntu; t++)
    if (w_residual_coding(&e, &ct, &tus[t], coeffs, p->flags, NULL)) return -1;
  if (ctxOut) memcpy(ctxOut, ct.m, sizeof(ct.m));
  return (long)e.nbins;
}

/* xGetCtxInitId, cabac_writer.cpp:41-73: which init table (0 = B, 1 = P) the next inter slice should
 * use - the one whose initial models cost the fewest excess bits against the contexts this slice
 * ended with, weighted by its per-context bin counts.  estFracExcessBits, contexts.cpp:965-968.
 * costs[0..1] = the two sums; returns the choice (ties keep B, the first candidate). */
int orc_ctx_init_id(int qp, const cabac_ctx_model *ctxRef, const uint32_t *ctxBins, uint64_t *costs) {
  uint64_t c[2];
  for (int t = 0; t < 2; t++) {
    cabac_ctx_model test[CABAC_NUM_CTX];
    uint8_t rates[CABAC_NUM_CTX];
    orc_ctx_init(qp, t, test, rates);
    c[t] = 0;
    for (int k = 0; k < CABAC_NUM_CTX; k++) {
      if (!ctxBins[k]) continue;
      const unsigned qs = ((unsigned)ctxRef[k].state0 + ctxRef[k].state1) >> 8; /* state() of the final model */
      const unsigned qt = ((unsigned)test[k].state0 + test[k].state1) >> 8;
      const uint64_t n = 2 * qs + 1;
      const uint64_t x = ((512 - n) * CABAC_FRAC_BITS[2 * qt] + n * CABAC_FRAC_BITS[2 * qt + 1] + 256) >> 9;
      c[t] += (uint64_t)ctxBins[k] * x;
    }
  }
  if (costs) costs[0] = c[0], costs[1] = c[1];
  return c[1] < c[0] ? 1 : 0;
}

/* Recorded command stream with CABAC_CMD_TU markers (cabac_encode_stream): the BinEncIf calls of a
 * CABACWriter interleaved with its residual_coding() calls, on one coder object. */
long orc_encode_stream(const uint32_t *cmds, long ncmd, const cabac_tu_desc *tus, long ntu,
                       const cabac_substream_params *p, const int32_t *coeffs, const cabac_ctx_model *ctxIn,
                       uint8_t *out, long cap, uint32_t *numBins, uint8_t *tuFlags, cabac_ctx_model *ctxOut) {
  CtxTab ct;
  Enc e;
  long t = 0;
  ctx_begin(&ct, p, ctxIn);
  enc_start(&e, out, cap);
  for (long i = 0; i < ncmd; i++) {
    const uint32_t w = cmds[i];
    switch (CABAC_CMD_KIND(w)) {
      case CABAC_CMD_CTX: enc_bin(&e, &ct, (w >> 9) & 1, w & 511); break;
      case CABAC_CMD_EP: enc_bins_ep(&e, w & 0xFFFFFF, (w >> 24) & 31); break;
      case CABAC_CMD_REM: enc_rem_abs(&e, w & 0xFFFFFF, (w >> 24) & 31, 5, 15); break;
      case CABAC_CMD_TRM: enc_trm(&e, w & 1); break;
      case CABAC_CMD_ALIGN: e.range = 256; break;
      case CABAC_CMD_TU: {
        if (t >= ntu) return -1;
        const unsigned st = w_residual_coding(&e, &ct, &tus[t], coeffs, p->flags, tuFlags ? &tuFlags[t] : NULL);
        if (st) return -(long)st;
        t++;
        break;
      }
      default: return -1;
    }
  }
  return enc_end(&e, p, &ct, numBins, ctxOut);
}

long orc_decode_residual(const uint8_t *in, long nbytes, const cabac_tu_desc *tus, long ntu,
                         const cabac_substream_params *p, const cabac_ctx_model *ctxIn, int32_t *coeffs,
                         uint8_t *tuFlags, cabac_ctx_model *ctxOut, uint32_t *status, int useMTS) {
  (void)useMTS;
  CtxTab ct;
  Dec d;
  ctx_begin(&ct, p, ctxIn);
  dec_start(&d, in, nbytes);
  for (long t = 0; t < ntu; t++) {
    const size_t n = (size_t)1 << (tus[t].log2_w + tus[t].log2_h);
    memset(coeffs + tus[t].coeff_offset, 0, n * sizeof(int32_t));
    r_residual_coding(&d, &ct, &tus[t], coeffs, p->flags, tuFlags ? &tuFlags[t] : NULL);
  }
  if (p->flags & CABAC_SUB_TERMINATE)
    if (!dec_trm(&d)) d.status |= CABAC_SS_NO_TERMINATE;
  if (ctxOut) memcpy(ctxOut, ct.m, sizeof(ct.m));
  if (status) *status = d.status;
  return dec_bits_read(&d);
}


/* ---- stepwise decoder handle ----------------------------------------------
 * One BinDecoder_Std + Ctx object driven call by call (decodeBin / decodeBinsEP / decodeBinTrm /
 * CABACReader::residual_coding).  oracle/syntax_host.cpp runs the CTU-level parser
 * (entropy_coding_b200/csrc/cabac_syntax.cuh compiled for the host) on top of it, so the parser's
 * control flow can be checked on the CPU against the reference writer's recorded call sequence. */
typedef struct {
  Dec d;
  CtxTab ct;
  cabac_substream_params p;
} OrcDecHandle;

void *orc_dec_open(const uint8_t *in, long n, const cabac_substream_params *p, const cabac_ctx_model *ctxIn) {
  OrcDecHandle *h = (OrcDecHandle *)malloc(sizeof(OrcDecHandle));
  if (!h) return NULL;
  h->p = *p;
  ctx_begin(&h->ct, p, ctxIn);
  dec_start(&h->d, in, n);
  return h;
}
unsigned orc_dec_bin(void *hv, unsigned ctxId) {
  OrcDecHandle *h = (OrcDecHandle *)hv;
  return dec_bin(&h->d, &h->ct, ctxId);
}
unsigned orc_dec_ep(void *hv, unsigned n) {
  OrcDecHandle *h = (OrcDecHandle *)hv;
  return n == 1 ? dec_bin_ep(&h->d) : dec_bins_ep(&h->d, n);
}
unsigned orc_dec_trm(void *hv) { return dec_trm(&((OrcDecHandle *)hv)->d); }
void orc_dec_tu(void *hv, const cabac_tu_desc *tu, int32_t *coeffs, uint8_t *tuFlags) {
  OrcDecHandle *h = (OrcDecHandle *)hv;
  memset(coeffs + tu->coeff_offset, 0, sizeof(int32_t) << (tu->log2_w + tu->log2_h));
  r_residual_coding(&h->d, &h->ct, tu, coeffs, h->p.flags, tuFlags);
}
/* WPP rows inside one slice: restart the arithmetic decoder on the next row's bytes (contexts stay), and
 * snapshot / restore the contexts (Ctx::operator=, contexts.cpp:1096-1108). */
long orc_dec_bits(void *hv) { return dec_bits_read(&((OrcDecHandle *)hv)->d); }
void orc_dec_restart(void *hv, const uint8_t *in, long n) {
  OrcDecHandle *h = (OrcDecHandle *)hv;
  const uint32_t st = h->d.status, nb = h->d.nbins;
  dec_start(&h->d, in, n);
  h->d.status = st;
  h->d.nbins = nb;
}
void orc_dec_ctx_get(void *hv, cabac_ctx_model *out) { memcpy(out, ((OrcDecHandle *)hv)->ct.m, sizeof(((OrcDecHandle *)hv)->ct.m)); }
void orc_dec_ctx_set(void *hv, const cabac_ctx_model *in) { memcpy(((OrcDecHandle *)hv)->ct.m, in, sizeof(((OrcDecHandle *)hv)->ct.m)); }
long orc_dec_close(void *hv, cabac_ctx_model *ctxOut, uint32_t *status, uint32_t *numBins) {
  OrcDecHandle *h = (OrcDecHandle *)hv;
  if (ctxOut) memcpy(ctxOut, h->ct.m, sizeof(h->ct.m));
  if (status) *status = h->d.status;
  if (numBins) *numBins = h->d.nbins;
  const long bits = dec_bits_read(&h->d);
  free(h);
  return bits;
}


/* ---- stepwise encoder handle ------------------------------------------------
 * One BinEncoder_Std + Ctx object driven call by call.  oracle/syntax_host.cpp uses it to GENERATE legal
 * CTU-level streams: the parser header runs on an engine that picks every bin value itself and codes it. */
typedef struct {
  Enc e;
  CtxTab ct;
  cabac_substream_params p;
  uint8_t *out;
  long cap, done; /* bytes of finished rows in front of the running coder */
  uint32_t bins;
} OrcEncHandle;

void *orc_enc_open(uint8_t *out, long cap, const cabac_substream_params *p, const cabac_ctx_model *ctxIn) {
  OrcEncHandle *h = (OrcEncHandle *)malloc(sizeof(OrcEncHandle));
  if (!h) return NULL;
  h->p = *p;
  h->out = out;
  h->cap = cap;
  h->done = 0;
  h->bins = 0;
  ctx_begin(&h->ct, p, ctxIn);
  enc_start(&h->e, out, cap);
  return h;
}
void orc_enc_bin(void *hv, unsigned bin, unsigned ctxId) { enc_bin(&((OrcEncHandle *)hv)->e, &((OrcEncHandle *)hv)->ct, bin, ctxId); }
void orc_enc_ep(void *hv, unsigned bins, unsigned n) {
  if (n) enc_bins_ep(&((OrcEncHandle *)hv)->e, bins, n);
}
void orc_enc_trm(void *hv, unsigned bin) { enc_trm(&((OrcEncHandle *)hv)->e, bin); }
unsigned orc_enc_tu(void *hv, const cabac_tu_desc *tu, const int32_t *coeffs, uint8_t *tuFlags) {
  OrcEncHandle *h = (OrcEncHandle *)hv;
  return w_residual_coding(&h->e, &h->ct, tu, coeffs, h->p.flags, tuFlags);
}
void orc_enc_ctx_get(void *hv, cabac_ctx_model *out) { memcpy(out, ((OrcEncHandle *)hv)->ct.m, sizeof(((OrcEncHandle *)hv)->ct.m)); }
void orc_enc_ctx_set(void *hv, const cabac_ctx_model *in) { memcpy(((OrcEncHandle *)hv)->ct.m, in, sizeof(((OrcEncHandle *)hv)->ct.m)); }
/* end of a WPP row / of the slice: the terminating bin was coded by the caller; finish(), byte alignment, and - for
 * a row - a fresh coder behind the bytes written so far.  Returns the total bytes. */
long orc_enc_flush(void *hv, int restart) {
  OrcEncHandle *h = (OrcEncHandle *)hv;
  enc_finish(&h->e);
  bw_byte_alignment(&h->e.bw);
  h->done += h->e.bw.n;
  h->bins += h->e.nbins;
  if (restart) enc_start(&h->e, h->out + (h->done < h->cap ? h->done : h->cap), h->cap > h->done ? h->cap - h->done : 0);
  return h->done;
}
long orc_enc_close(void *hv, cabac_ctx_model *ctxOut, uint32_t *numBins) {
  OrcEncHandle *h = (OrcEncHandle *)hv;
  if (ctxOut) memcpy(ctxOut, h->ct.m, sizeof(h->ct.m));
  if (numBins) *numBins = h->bins;
  const long n = h->done;
  free(h);
  return n;
}

/* ---- timed batches (bench.py cpu_baseline, kind = "port") ---------------- */
typedef struct {
  int k, nthreads, decode;
  long nSub;
  const cabac_tu_desc *tus;
  const uint64_t *tuOffsets;
  const cabac_substream_params *params;
  const int32_t *coeffs;
  int32_t *coeffsOut;
  uint8_t *out;
  const uint64_t *outOffsets;
  uint32_t *outSizes;
  const uint8_t *in;
  const uint64_t *inOffsets;
  const uint32_t *inSizes;
  uint64_t bins, bytes;
  long *next; /* shared work counter: substreams are handed out dynamically */
} Job;

static void *job_run(void *arg) {
  Job *j = (Job *)arg;
  for (long i = __atomic_fetch_add(j->next, 1, __ATOMIC_RELAXED); i < j->nSub;
       i = __atomic_fetch_add(j->next, 1, __ATOMIC_RELAXED)) {
    const cabac_tu_desc *t = j->tus + j->tuOffsets[i];
    const long nt = (long)(j->tuOffsets[i + 1] - j->tuOffsets[i]);
    if (j->decode) {
      orc_decode_residual(j->in + j->inOffsets[i], (long)j->inSizes[i], t, nt, &j->params[i], NULL, j->coeffsOut,
                          NULL, NULL, NULL, 0);
    } else {
      uint32_t nb = 0;
      uint8_t *o = j->out ? j->out + j->outOffsets[i] : NULL;
      const long cap = j->out ? (long)(j->outOffsets[i + 1] - j->outOffsets[i]) : 0;
      const long n = orc_encode_residual(t, nt, &j->params[i], j->coeffs, NULL, o, cap, &nb, NULL, NULL, 0);
      if (j->outSizes) j->outSizes[i] = (uint32_t)n;
      j->bins += nb;
      j->bytes += (uint64_t)(n > 0 ? n : 0);
    }
  }
  return NULL;
}

static double run_jobs(Job *proto, int nthreads, uint64_t *bins, uint64_t *bytes) {
  pthread_once(&g_romOnce, rom_init);
  if (nthreads < 1) nthreads = 1;
  Job *jobs = (Job *)calloc((size_t)nthreads, sizeof(Job));
  pthread_t *th = (pthread_t *)calloc((size_t)nthreads, sizeof(pthread_t));
  struct timespec t0, t1;
  long next = 0;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (int k = 0; k < nthreads; k++) {
    jobs[k] = *proto;
    jobs[k].next = &next;
    jobs[k].k = k;
    jobs[k].nthreads = nthreads;
    pthread_create(&th[k], NULL, job_run, &jobs[k]);
  }
  uint64_t b = 0, by = 0;
  for (int k = 0; k < nthreads; k++) {
    pthread_join(th[k], NULL);
    b += jobs[k].bins;
    by += jobs[k].bytes;
  }
  clock_gettime(CLOCK_MONOTONIC, &t1);
  if (bins) *bins = b;
  if (bytes) *bytes = by;
  free(jobs);
  free(th);
  return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}

double orc_time_encode_residual(const cabac_tu_desc *tus, const uint64_t *tuOffsets,
                                const cabac_substream_params *params, long nSub, const int32_t *coeffs,
                                int nthreads, uint64_t *totalBins, uint64_t *totalBytes, uint8_t *out,
                                const uint64_t *outOffsets, uint32_t *outSizes) {
  Job j;
  memset(&j, 0, sizeof(j));
  j.nSub = nSub;
  j.tus = tus;
  j.tuOffsets = tuOffsets;
  j.params = params;
  j.coeffs = coeffs;
  j.out = out;
  j.outOffsets = outOffsets;
  j.outSizes = outSizes;
  return run_jobs(&j, nthreads, totalBins, totalBytes);
}

double orc_time_decode_residual(const uint8_t *in, const uint64_t *inOffsets, const uint32_t *inSizes,
                                const cabac_tu_desc *tus, const uint64_t *tuOffsets,
                                const cabac_substream_params *params, long nSub, int32_t *coeffs,
                                int nthreads) {
  Job j;
  memset(&j, 0, sizeof(j));
  j.decode = 1;
  j.nSub = nSub;
  j.tus = tus;
  j.tuOffsets = tuOffsets;
  j.params = params;
  j.coeffsOut = coeffs;
  j.in = in;
  j.inOffsets = inOffsets;
  j.inSizes = inSizes;
  return run_jobs(&j, nthreads, NULL, NULL);
}

/* ---- start-code emulation ----------------------------------------- */
/* OutputBitstream::countStartCodeEmulations, src/common/bit_stream.cpp:157-181: repeatedly find the
 * next pair of zero bytes that is followed by another byte (the search stops one byte short of the
 * end); when that byte is <= 3 count one emulation and continue the search AT that byte. */
long orc_count_emulations(const uint8_t *d, long n) {
  long cnt = 0, it = 0;
  while (it < n) {
    long found = it;
    for (;;) {
      long p = found; /* search_n(found, end - 1, 2, 0): both zeros inside [found, n - 1) */
      while (p + 1 <= n - 2 && !(d[p] == 0 && d[p + 1] == 0)) p++;
      if (p + 1 > n - 2) p = n - 1; /* not found: search_n returns its `last` = end - 1 */
      found = p + 1;
      if (found >= n) { found = n; break; }
      found++;
      if (d[found] <= 3) break;
    }
    it = found;
    if (found != n) cnt++;
  }
  return cnt;
}

/* writeNaluContent, VVCSoftware_VTM/source/Lib/EncoderLib/NALwrite.cpp:67-126 (the VTM submodule;
 * not part of the reference's src/): RBSP -> NAL payload.  Returns the output length. */
long orc_insert_emulation_prevention(const uint8_t *in, long n, uint8_t *out) {
  long o = 0;
  int zeroCount = 0;
  for (long i = 0; i < n; i++) {
    const uint8_t v = in[i];
    if (zeroCount == 2 && v <= 3) {
      out[o++] = 3;
      zeroCount = 0;
    }
    if (v == 0) zeroCount++; else zeroCount = 0;
    out[o++] = v;
  }
  if (zeroCount > 0) out[o++] = 3;
  return o;
}

/* convertPayloadToRBSP, VVCSoftware_VTM/source/Lib/DecoderLib/NALread.cpp:59-106: NAL unit bytes -> RBSP.
 * An emulation_prevention_three_byte is a 0x03 that follows two zero bytes of the output; its position
 * in the ESCAPED buffer is recorded (pushEmulationPreventionByteLocation).  vcl: trailing zero bytes
 * (cabac_zero_words) are dropped.  Returns the RBSP length, or -1 where VTM's CHECKs throw. */
long orc_remove_emulation_prevention(const uint8_t *nal, long n, int vcl, uint8_t *rbsp, uint32_t *epb,
                                     long *n_epb) {
  long w = 0, k = 0;
  unsigned zeroCount = 0;
  for (long r = 0; r < n; r++) {
    if (zeroCount >= 2 && nal[r] < 3) return -1; /* "Zero count is '2' and read value is small than '3'" */
    if (zeroCount == 2 && nal[r] == 3) {
      if (epb) epb[k] = (uint32_t)r;
      k++;
      r++;
      zeroCount = 0;
      if (r == n) break;
      if (nal[r] > 3) return -1; /* "Read a value bigger than '3'" */
    }
    zeroCount = nal[r] == 0 ? zeroCount + 1 : 0;
    rbsp[w++] = nal[r];
  }
  if (zeroCount != 0) return -1; /* "Zero count not '0'" */
  if (vcl)
    while (w > 0 && rbsp[w - 1] == 0) w--;
  if (n_epb) *n_epb = k;
  return w;
}

/* Slice-header form of the entry points.  EncSlice.cpp:2118-2125: the signalled size of substream i
 * (all but the last) is its byte count plus its start-code emulations; HLSWriter::codeTilesWPPEntryPoint,
 * VLCWriter.cpp:2833-2866: sh_entry_offset_len_minus1 is the smallest length that holds the largest
 * size, each sh_entry_point_offset_minus1 is size - 1.  Returns sh_entry_offset_len_minus1. */
int orc_entry_point_syntax(const uint32_t *sizes, const uint32_t *emulations, long n_sub, uint32_t *offset_minus1) {
  uint32_t maxOffset = 0;
  for (long i = 0; i + 1 < n_sub; i++) {
    const uint32_t s = sizes[i] + emulations[i];
    if (s > maxOffset) maxOffset = s;
    offset_minus1[i] = s - 1;
  }
  int len = 0;
  while (maxOffset >= (1u << (len + 1))) len++;
  return len;
}

/* Decoder side: HLSyntaxReader::parseSliceHeader, VLCReader.cpp:4346-4381 (the signalled sizes minus the
 * emulation prevention bytes inside each substream) and DecSlice.cpp:106-114 + InputBitstream::
 * extractSubstream, bit_stream.cpp:382-415 (consecutive byte ranges, the last substream takes the rest).
 * hdr_end: RBSP byte position of the first slice-data byte; epb: locations in the escaped buffer.
 * in_offsets[n_entry + 2]: RBSP byte range of every substream. */
void orc_substream_ranges(const uint32_t *offset_minus1, long n_entry, const uint32_t *epb, long n_epb,
                          uint64_t hdr_end, uint64_t rbsp_size, uint64_t *in_offsets) {
  uint64_t endOfSliceHeaderLocation = hdr_end;
  for (long k = 0; k < n_epb; k++)
    if (epb[k] < endOfSliceHeaderLocation) endOfSliceHeaderLocation++;
  uint64_t cur = 0, prev = 0, pos = hdr_end;
  in_offsets[0] = pos;
  for (long i = 0; i < n_entry; i++) {
    cur += (uint64_t)offset_minus1[i] + 1;
    uint64_t cnt = 0;
    for (long k = 0; k < n_epb; k++)
      if (epb[k] >= prev + endOfSliceHeaderLocation && epb[k] < cur + endOfSliceHeaderLocation) cnt++;
    pos += (uint64_t)offset_minus1[i] + 1 - cnt;
    in_offsets[i + 1] = pos;
    prev = cur;
  }
  in_offsets[n_entry + 1] = rbsp_size;
}
