/*
 * cabac_b200.h - C ABI of the B200-native VVC CABAC engine.
 *
 * This is the drop-in boundary for the entropy-coding hot path of
 * p-sawicki/entropy_coding (reference paths below are relative to that repo):
 *
 *   (1) the arithmetic coder    src/entropy_codec/arith_codec.{hpp,cpp}
 *   (2) the probability models  src/common/contexts.{hpp,cpp}
 *   (3) residual binarisation   src/entropy_codec/cabac_{writer,reader}.cpp
 *                               src/common/context_modelling.{hpp,cpp}
 *
 * Plain C: pointers and sizes only, no C++/torch types.  Every entry point
 * exists in two flavours:
 *   - `*_dev`  : all buffer arguments are DEVICE pointers, the launch is
 *                asynchronous on the handle's CUDA stream (call cabac_sync);
 *   - no suffix: all buffer arguments are HOST pointers; the call copies
 *                inputs H2D, launches, copies results D2H and synchronises.
 * There is no CPU fallback: without a CUDA device every call fails with
 * CABAC_ERR_CUDA.
 *
 * One independent substream (slice, tile, WPP CTU row, all-intra frame) is
 * coded by one warp.  Batches of 10^4..10^6 substreams per launch are the
 * design point.
 */
#ifndef CABAC_B200_H
#define CABAC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Ctx::NumberOfContexts, src/common/contexts.cpp:773-774 */
#define CABAC_NUM_CTX 379

/* ---- status codes (returned by every call) ------------------------------ */
enum {
  CABAC_OK = 0,
  CABAC_ERR_CUDA = -1,     /* CUDA runtime failure; see cabac_last_error() */
  CABAC_ERR_ARG = -2,      /* bad argument                                  */
  CABAC_ERR_SUBSTREAM = -3 /* >=1 substream failed; see the status[] array  */
};

/* ---- per-substream status bits (status[] output arrays) ----------------- */
/* Mirror the reference's CHECK/THROW sites (src/common/type_def.hpp:295-329). */
enum {
  CABAC_SS_OK = 0,
  CABAC_SS_OUT_OVERFLOW = 1, /* output capacity exhausted (bytes truncated)  */
  CABAC_SS_EMPTY_TU = 2,     /* "Coefficient coding called for empty TU",
                                cabac_writer.cpp:2457                        */
  CABAC_SS_BAD_CMD = 4,      /* malformed command word                       */
  CABAC_SS_IN_OVERRUN = 8,   /* decoder read past the substream's bytes,
                                bit_stream.cpp:269                           */
  CABAC_SS_COEFF_RANGE = 16, /* "TransCoeffLevel outside allowable range",
                                cabac_reader.cpp:2942-2946                   */
  CABAC_SS_NO_TERMINATE = 32, /* terminating bin / stop pattern missing,
                                arith_codec.cpp:68-73                        */
  CABAC_SS_WPP_TIMEOUT = 64, /* cabac_*_wpp: the predecessor never published
                                its contexts (malformed dep chain)           */
  CABAC_SS_BAD_DESC = 128    /* a cabac_tu_desc with log2_w / log2_h > 6 or
                                comp_id > 2 (the reference CHECKs the block
                                size, context_modelling.cpp:7-79); the block
                                was coded with the fields clamped into range,
                                so the substream's output is not meaningful   */
};

/* ---- probability model as exchanged over the boundary ------------------- */
/* BinProbModel_Std::m_state[0..1], contexts.hpp:184-190.  The rate
 * (m_rate) is a per-context-index constant (row [3] of the init tables,
 * contexts.cpp:1012) and is exported once by cabac_ctx_rates().            */
typedef struct cabac_ctx_model {
  uint16_t state0; /* 10 significant bits, MASK_0 = 0x7FE0 */
  uint16_t state1; /* 14 significant bits, MASK_1 = 0x7FFE */
} cabac_ctx_model;

/* ---- per-substream parameters ------------------------------------------- */
enum {
  CABAC_SUB_DEP_QUANT = 1,   /* Slice::getDepQuantEnabledFlag()              */
  CABAC_SUB_SIGN_HIDING = 2, /* Slice::getSignDataHidingEnabledFlag()        */
  CABAC_SUB_TSRC_DISABLED = 4, /* Slice::getTSResidualCodingDisabledFlag()   */
  CABAC_SUB_TERMINATE = 8,   /* after the last unit: encodeBinTrm(1) +
                                finish() (end_of_slice, cabac_writer.cpp:104)
                                / decode: terminating_bit + finish check     */
  CABAC_SUB_ALIGN = 16,      /* encode: append writeByteAlignment()
                                (bit_stream.cpp:152-155, EncSlice.cpp:2114)  */
  CABAC_SUB_CTX_IN = 32      /* start from ctx_in[] instead of Ctx::init     */
};

typedef struct cabac_substream_params {
  int8_t qp;       /* slice QP, clipped to 0..63 as CtxStore::init does      */
  uint8_t init_id; /* 0 = B, 1 = P, 2 = I table (type_def.hpp:22-25)         */
  uint8_t flags;   /* CABAC_SUB_*                                            */
  uint8_t reserved;
} cabac_substream_params;

/* ---- bin-level command words (BinEncIf / BinDecoderBase calls) ---------- */
/* One 32-bit word per call recorded from arith_codec.hpp:53-59 / 238-246.
 *   bits 31..29 kind
 *   CTX   : bit 9 = bin, bits 8..0 = ctxId        encodeBin / decodeBin
 *   EP    : bits 28..24 = n (0..24), 23..0 = bins encodeBinsEP / decodeBinsEP
 *           (encodeBinEP(b) == EP with n = 1; longer runs are split by the
 *            recorder, which is arithmetically identical)
 *   REM   : bits 28..24 = rice (0..12), 23..0 = value
 *                                                 encodeRemAbsEP(v, rice, 5, 15)
 *   TRM   : bit 0 = bin                           encodeBinTrm / decodeBinTrm
 *   ALIGN :                                       align()  (m_Range = 256)
 */
enum {
  CABAC_CMD_CTX = 0,
  CABAC_CMD_EP = 1,
  CABAC_CMD_REM = 2,
  CABAC_CMD_TRM = 3,
  CABAC_CMD_ALIGN = 4,
  CABAC_CMD_TU = 5 /* cabac_encode_stream only: residual_coding() of the
                      substream's next TU descriptor                        */
};
#define CABAC_CMD_KIND(w) ((uint32_t)(w) >> 29)
#define CABAC_MK_CTX(ctx, bin) \
  (((uint32_t)CABAC_CMD_CTX << 29) | (((uint32_t)(bin) & 1u) << 9) | ((uint32_t)(ctx) & 511u))
#define CABAC_MK_EP(bins, n) \
  (((uint32_t)CABAC_CMD_EP << 29) | (((uint32_t)(n) & 31u) << 24) | ((uint32_t)(bins) & 0xFFFFFFu))
#define CABAC_MK_REM(value, rice) \
  (((uint32_t)CABAC_CMD_REM << 29) | (((uint32_t)(rice) & 31u) << 24) | ((uint32_t)(value) & 0xFFFFFFu))
#define CABAC_MK_TRM(bin) (((uint32_t)CABAC_CMD_TRM << 29) | ((uint32_t)(bin) & 1u))
#define CABAC_MK_ALIGN() ((uint32_t)CABAC_CMD_ALIGN << 29)
#define CABAC_MK_TU() ((uint32_t)CABAC_CMD_TU << 29)

/* ---- flat transform-unit descriptor (residual level) --------------------- */
/* Replaces the (TransformUnit&, ComponentID) argument pair of
 * CABACWriter::residual_coding (cabac_writer.hpp:128-129) and
 * CABACReader::residual_coding (cabac_reader.hpp:111).                      */
enum {
  CABAC_TU_TS = 1,          /* tu.mtsIdx[comp] == MTS_SKIP  (encode input)   */
  CABAC_TU_TS_ALLOWED = 2,  /* TU::isTSAllowed() -> transform_skip_flag is
                               coded (unit_tools.cpp:651-664)                */
  CABAC_TU_BDPCM = 4,       /* cu.bdpcmMode / bdpcmModeChroma for this comp  */
  CABAC_TU_SBT_ZEROOUT = 8  /* sps.getUseMTS() && cu.sbtInfo != 0 (luma):
                               32-sized dims are zeroed beyond 16
                               (cabac_writer.cpp:2508-2517, 2661-2667)       */
};

typedef struct cabac_tu_desc {
  uint64_t coeff_offset; /* first TCoeff (int32) of the block in the slab;
                            row-major, stride == width (unit.hpp:551)        */
  uint8_t log2_w;        /* width  = 1 << log2_w, 1..64 (1 and 2: ISP parts) */
  uint8_t log2_h;        /* height = 1 << log2_h, 1..64                      */
  uint8_t comp_id;       /* 0 = Y, 1 = Cb, 2 = Cr                            */
  uint8_t flags;         /* CABAC_TU_*                                       */
  uint32_t reserved;
} cabac_tu_desc;

/* Per-TU result flags (side effects of residual_coding on CUCtx / tu,
 * cabac_writer.cpp:2460-2478, 2520-2523; cabac_reader.cpp:2672-2690, 2751). */
enum {
  CABAC_TUR_TS = 1,               /* decoded transform_skip_flag             */
  CABAC_TUR_VIOLATES_LFNST = 2,   /* violatesLfnstConstrained[chType]        */
  CABAC_TUR_LFNST_LAST_POS = 4,   /* lfnstLastScanPos                        */
  CABAC_TUR_MTS_LAST_POS = 8,     /* mtsLastScanPos                          */
  CABAC_TUR_VIOLATES_MTS = 16     /* violatesMtsCoeffConstraint              */
};

typedef struct cabac_handle cabac_handle;

/* ---- lifetime ------------------------------------------------------------ */
/* device < 0 selects the current CUDA device.  Returns NULL on failure.     */
cabac_handle *cabac_create(int device);
void cabac_destroy(cabac_handle *h);
/* Use an externally owned CUDA stream (cudaStream_t as void*); NULL restores
 * the handle's own stream.                                                  */
int cabac_set_stream(cabac_handle *h, void *cuda_stream);
int cabac_sync(cabac_handle *h);
/* Residual encoder variant.  AUTO picks by batch shape: FUSED = one warp per
 * substream does binarisation and arithmetic coding (few, long substreams);
 * TWO_STAGE = warps binarise and resolve probabilities into command words,
 * then one thread per substream runs the interval coder (many substreams).
 * WARP_SPECIALISED (residual TU lists only; AUTO picks it between 5 624 and 32 768
 * substreams, where every producer slot of the GPU is filled at least twice) =
 * producer warps binarise and resolve, one walker warp per CTA runs the range
 * recurrence of all of them lane-parallel, the producers fold their words into the
 * code value in lane-parallel groups.
 * All are byte-identical.  Also settable with
 * CABAC_B200_ENCODER=fused|two_stage|warp_specialised. */
enum {
  CABAC_ENCODER_AUTO = 0,
  CABAC_ENCODER_FUSED = 1,
  CABAC_ENCODER_TWO_STAGE = 2,
  CABAC_ENCODER_WARP_SPECIALISED = 3,
  CABAC_ENCODER_SLICED = 4 /* few, LONG substreams (slices, frames), terminated: every stage parallel INSIDE
                              a substream - all TUs binarised at once, the contexts resolved as independent
                              chains, the interval coder split into segments whose pieces are summed
                              (DESIGN.md section 4).  AUTO picks it below 2 048 substreams of >= 256 TUs. */
};
int cabac_set_encoder_mode(cabac_handle *h, int mode);
/* Residual decoder variant.  LOCKSTEP = one warp per substream, all lanes run the same arithmetic decoder
 * and share the parallel parts (context templates, signs).  THREAD = one thread per substream with its
 * context models in a global-memory row.  Identical results.  AUTO = LOCKSTEP at every batch size: on
 * B200 the thread kernel issues 4.8x fewer warp instructions but diverges to 7.4 active lanes and is
 * 1.2x slower at 10^5 substreams (profiles/r2_decoder_modes.md).  CABAC_B200_DECODER=lockstep|thread. */
enum { CABAC_DECODER_AUTO = 0, CABAC_DECODER_LOCKSTEP = 1, CABAC_DECODER_THREAD = 2 };
int cabac_set_decoder_mode(cabac_handle *h, int mode);
const char *cabac_last_error(cabac_handle *h);
const char *cabac_version(void);
/* Number of kernel launches issued through this handle so far.              */
uint64_t cabac_launch_count(cabac_handle *h);

/* Page-locked host memory for the host-pointer entry points (any host pointer
 * is accepted; pinned buffers let the H2D / D2H copies run at link speed).  */
void *cabac_host_alloc(size_t bytes);
void cabac_host_free(void *p);

/* ---- batch descriptors resident on the device ----------------------------- */
/* The host-pointer residual calls upload tus / tu_offsets / params on every call.
 * A caller that makes several calls over the SAME descriptor arrays (encode and
 * decode of one batch, an estimate followed by the final pass, repeated passes of
 * an RDO loop) uploads them once: after cabac_pin_descriptors the host-pointer
 * residual entry points that are passed exactly these three pointers and n_sub
 * use the device copy instead.  The arrays must stay unchanged until
 * cabac_unpin_descriptors (or the next cabac_pin_descriptors / cabac_destroy). */
int cabac_pin_descriptors(cabac_handle *h, const cabac_tu_desc *tus,
                          const uint64_t *tu_offsets,
                          const cabac_substream_params *params, size_t n_sub);
int cabac_unpin_descriptors(cabac_handle *h);

/* ---- (2) context initialisation ----------------------------------------- */
/* Ctx::init(qp, initId) for n (qp, init_id) pairs, contexts.cpp:996-1015,
 * 893-901.  out: n * CABAC_NUM_CTX models.                                   */
int cabac_ctx_init(cabac_handle *h, const cabac_substream_params *params,
                   size_t n, cabac_ctx_model *out);
int cabac_ctx_init_dev(cabac_handle *h, const cabac_substream_params *params,
                       size_t n, cabac_ctx_model *out);
/* m_rate of every context index (host memory, CABAC_NUM_CTX bytes);
 * BinProbModel_Std::setLog2WindowSize, contexts.cpp:915-920.                 */
int cabac_ctx_rates(uint8_t *rates_out);
/* Raw init table row (0=B,1=P,2=I,3=window sizes), CABAC_NUM_CTX bytes.     */
int cabac_ctx_init_table(int row, uint8_t *table_out);

/* ---- (1) bin-level arithmetic coder -------------------------------------- */
/* Encodes n_sub recorded command streams.  Substream i owns command words
 * cmds[cmd_offsets[i] .. cmd_offsets[i+1]) and output bytes
 * out[out_offsets[i] .. out_offsets[i+1]) (its capacity).
 *   out_sizes[i]  bytes produced (incl. finish/alignment when flagged)
 *   num_bins[i]   BinCounter::getAll(), arith_codec.cpp:302-308   (optional)
 *   ctx_in        n_sub*CABAC_NUM_CTX models, used when CABAC_SUB_CTX_IN
 *   ctx_out       n_sub*CABAC_NUM_CTX models after the run       (optional)
 *   status[i]     CABAC_SS_* bits                                 (optional)
 */
int cabac_encode_bins(cabac_handle *h, const uint32_t *cmds,
                      const uint64_t *cmd_offsets,
                      const cabac_substream_params *params, size_t n_sub,
                      const cabac_ctx_model *ctx_in, uint8_t *out,
                      const uint64_t *out_offsets, uint32_t *out_sizes,
                      uint32_t *num_bins, cabac_ctx_model *ctx_out,
                      uint32_t *status);
int cabac_encode_bins_dev(cabac_handle *h, const uint32_t *cmds,
                          const uint64_t *cmd_offsets,
                          const cabac_substream_params *params, size_t n_sub,
                          const cabac_ctx_model *ctx_in, uint8_t *out,
                          const uint64_t *out_offsets, uint32_t *out_sizes,
                          uint32_t *num_bins, cabac_ctx_model *ctx_out,
                          uint32_t *status);

/* Decodes with a known call skeleton (the context of every bin is given by
 * the command words; bin/bins/value payload fields are ignored on input).
 * values[k] receives the decoded bin / bins / remainder of command k.
 * Substream i reads bytes in[in_offsets[i] .. in_offsets[i+1]).
 *   bits_read[i]  BinDecoderBase::getNumBitsRead(), arith_codec.cpp:201-203
 */
int cabac_decode_bins(cabac_handle *h, const uint8_t *in,
                      const uint64_t *in_offsets, const uint32_t *cmds,
                      const uint64_t *cmd_offsets,
                      const cabac_substream_params *params, size_t n_sub,
                      const cabac_ctx_model *ctx_in, uint32_t *values,
                      uint32_t *bits_read, cabac_ctx_model *ctx_out,
                      uint32_t *status);
int cabac_decode_bins_dev(cabac_handle *h, const uint8_t *in,
                          const uint64_t *in_offsets, const uint32_t *cmds,
                          const uint64_t *cmd_offsets,
                          const cabac_substream_params *params, size_t n_sub,
                          const cabac_ctx_model *ctx_in, uint32_t *values,
                          uint32_t *bits_read, cabac_ctx_model *ctx_out,
                          uint32_t *status);

/* ---- (3) residual-level coder ------------------------------------------- */
/* For each substream i: reset(qp, init_id) [or ctx_in], then
 * residual_coding() for TUs tus[tu_offsets[i] .. tu_offsets[i+1]) in order,
 * then (CABAC_SUB_TERMINATE) encodeBinTrm(1) + finish() and
 * (CABAC_SUB_ALIGN) writeByteAlignment().
 *   coeffs       the TCoeff (int32) slab the descriptors point into
 *   n_coeffs     slab length in TCoeff units (bounds the H2D copy)
 *   tu_flags[t]  CABAC_TUR_* per TU                               (optional)
 */
int cabac_encode_residual(cabac_handle *h, const cabac_tu_desc *tus,
                          const uint64_t *tu_offsets,
                          const cabac_substream_params *params, size_t n_sub,
                          const int32_t *coeffs, size_t n_coeffs,
                          const cabac_ctx_model *ctx_in, uint8_t *out,
                          const uint64_t *out_offsets, uint32_t *out_sizes,
                          uint32_t *num_bins, uint8_t *tu_flags,
                          cabac_ctx_model *ctx_out, uint32_t *status);
int cabac_encode_residual_dev(cabac_handle *h, const cabac_tu_desc *tus,
                              const uint64_t *tu_offsets,
                              const cabac_substream_params *params,
                              size_t n_sub, const int32_t *coeffs,
                              const cabac_ctx_model *ctx_in, uint8_t *out,
                              const uint64_t *out_offsets, uint32_t *out_sizes,
                              uint32_t *num_bins, uint8_t *tu_flags,
                              cabac_ctx_model *ctx_out, uint32_t *status);

/* ---- (1)+(3) recorded command streams ------------------------------------- */
/* The full CABACWriter seam: a substream is the recorded sequence of BinEncIf
 * calls its syntax writer made (command words as above, raw context ids) with
 * a CABAC_CMD_TU marker wherever the writer called residual_coding(); marker k
 * of substream i codes TU tus[tu_offsets[i] + k].  This is how coding_tree /
 * coding_unit / cbf / intra-mode bins recorded on the host interleave with
 * device-side residual binarisation.  Arguments as in
 * cabac_encode_bins + cabac_encode_residual; tus / tu_offsets / coeffs may be
 * NULL when no substream contains a marker.                                  */
int cabac_encode_stream(cabac_handle *h, const uint32_t *cmds,
                        const uint64_t *cmd_offsets, const cabac_tu_desc *tus,
                        const uint64_t *tu_offsets,
                        const cabac_substream_params *params, size_t n_sub,
                        const int32_t *coeffs, size_t n_coeffs,
                        const cabac_ctx_model *ctx_in, uint8_t *out,
                        const uint64_t *out_offsets, uint32_t *out_sizes,
                        uint32_t *num_bins, uint8_t *tu_flags,
                        cabac_ctx_model *ctx_out, uint32_t *status);
int cabac_encode_stream_dev(cabac_handle *h, const uint32_t *cmds,
                            const uint64_t *cmd_offsets,
                            const cabac_tu_desc *tus,
                            const uint64_t *tu_offsets,
                            const cabac_substream_params *params, size_t n_sub,
                            const int32_t *coeffs,
                            const cabac_ctx_model *ctx_in, uint8_t *out,
                            const uint64_t *out_offsets, uint32_t *out_sizes,
                            uint32_t *num_bins, uint8_t *tu_flags,
                            cabac_ctx_model *ctx_out, uint32_t *status);

/* Host-pointer encode that returns what the slice writer consumes
 * (EncSlice.cpp:2114-2125, EncGOP.cpp:3811-3824): the substreams compacted in
 * order into `packed` plus the entry-point table, i.e. cabac_encode_residual
 * followed by cabac_pack with a single D2H copy of the produced bytes.
 * out_offsets describes the per-substream capacity of the DEVICE-side slab
 * only; entry_points has n_sub + 1 entries; packed_cap bounds `packed`.     */
int cabac_encode_residual_packed(cabac_handle *h, const cabac_tu_desc *tus,
                                 const uint64_t *tu_offsets,
                                 const cabac_substream_params *params,
                                 size_t n_sub, const int32_t *coeffs,
                                 size_t n_coeffs, const cabac_ctx_model *ctx_in,
                                 const uint64_t *out_offsets, uint8_t *packed,
                                 size_t packed_cap, uint64_t *entry_points,
                                 uint32_t *out_sizes, uint32_t *num_bins,
                                 uint8_t *tu_flags, cabac_ctx_model *ctx_out,
                                 uint32_t *status);

/* Mirror of the above.  The kernel zeroes each TU's coefficient block before
 * parsing (the reference requires the caller to pre-zero,
 * cabac_reader.cpp:2952-2953).  descriptor flag CABAC_TU_TS is ignored when
 * CABAC_TU_TS_ALLOWED is set (the flag is parsed) and otherwise supplies the
 * inferred value (cabac_reader.cpp:2738-2741).  The host-pointer flavour
 * returns the whole slab: positions no descriptor covers come back as zero.
 */
int cabac_decode_residual(cabac_handle *h, const uint8_t *in,
                          const uint64_t *in_offsets, const cabac_tu_desc *tus,
                          const uint64_t *tu_offsets,
                          const cabac_substream_params *params, size_t n_sub,
                          const cabac_ctx_model *ctx_in, int32_t *coeffs,
                          size_t n_coeffs, uint32_t *num_bins,
                          uint32_t *bits_read, uint8_t *tu_flags,
                          cabac_ctx_model *ctx_out, uint32_t *status);
int cabac_decode_residual_dev(cabac_handle *h, const uint8_t *in,
                              const uint64_t *in_offsets,
                              const cabac_tu_desc *tus,
                              const uint64_t *tu_offsets,
                              const cabac_substream_params *params,
                              size_t n_sub, const cabac_ctx_model *ctx_in,
                              int32_t *coeffs, uint32_t *num_bins,
                              uint32_t *bits_read, uint8_t *tu_flags,
                              cabac_ctx_model *ctx_out, uint32_t *status);

/* Allocation-free residual decode.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc
 * (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, meaning and results as cabac_decode_residual_dev (all
 * pointers are device pointers); runs the thread-per-substream kernel with its
 * per-substream context rows in a static device array of
 * CABAC_STATIC_MAX_SUBSTREAMS x 230 words (14.4 MiB, held by every GPU that
 * loads the library until the process exits) and calls no cudaMalloc /
 * cudaFree / allocating routine.  n_sub > CABAC_STATIC_MAX_SUBSTREAMS is
 * refused with CABAC_ERR_ARG (cabac_last_error says so).  The array is shared
 * by every call and every handle on the device: a call rewrites the rows it
 * uses before reading them, so nothing has to be cleared between calls, but
 * two calls whose kernels overlap on the GPU (different streams) race.       */
#define CABAC_STATIC_MAX_SUBSTREAMS 16384
int cabac_decode_residual_static_dev(cabac_handle *h, const uint8_t *in,
                                     const uint64_t *in_offsets,
                                     const cabac_tu_desc *tus,
                                     const uint64_t *tu_offsets,
                                     const cabac_substream_params *params,
                                     size_t n_sub,
                                     const cabac_ctx_model *ctx_in,
                                     int32_t *coeffs, uint32_t *num_bins,
                                     uint32_t *bits_read, uint8_t *tu_flags,
                                     cabac_ctx_model *ctx_out,
                                     uint32_t *status);

/* ---- int16 coefficient slabs ----------------------------------------------- */
/* The residual calls with the coefficient slab as int16 instead of the reference's
 * int32 TCoeff (unit.hpp:551).  VVC v1 levels fit 16 bits (maxLog2TrDynamicRange =
 * 15, context_modelling.cpp:66-72; cabac_reader.cpp:2942-2946 refuses anything else),
 * and the shim copies every block out of TransformUnit::m_coeffs into the batch slab
 * anyway (host/gpu_bin_encoder.hpp), so it can narrow there: half the bytes cross
 * PCIe.  Same descriptors (coeff_offset in coefficients), same results.        */
int cabac_encode_residual_i16_dev(cabac_handle *h, const cabac_tu_desc *tus,
                                  const uint64_t *tu_offsets,
                                  const cabac_substream_params *params, size_t n_sub,
                                  const int16_t *coeffs, const cabac_ctx_model *ctx_in,
                                  uint8_t *out, const uint64_t *out_offsets,
                                  uint32_t *out_sizes, uint32_t *num_bins,
                                  uint8_t *tu_flags, cabac_ctx_model *ctx_out,
                                  uint32_t *status);
int cabac_decode_residual_i16_dev(cabac_handle *h, const uint8_t *in,
                                  const uint64_t *in_offsets, const cabac_tu_desc *tus,
                                  const uint64_t *tu_offsets,
                                  const cabac_substream_params *params, size_t n_sub,
                                  const cabac_ctx_model *ctx_in, int16_t *coeffs,
                                  uint32_t *num_bins, uint32_t *bits_read,
                                  uint8_t *tu_flags, cabac_ctx_model *ctx_out,
                                  uint32_t *status);
int cabac_encode_residual_packed_i16(cabac_handle *h, const cabac_tu_desc *tus,
                                     const uint64_t *tu_offsets,
                                     const cabac_substream_params *params, size_t n_sub,
                                     const int16_t *coeffs, size_t n_coeffs,
                                     const cabac_ctx_model *ctx_in,
                                     const uint64_t *out_offsets, uint8_t *packed,
                                     size_t packed_cap, uint64_t *entry_points,
                                     uint32_t *out_sizes, uint32_t *num_bins,
                                     uint8_t *tu_flags, cabac_ctx_model *ctx_out,
                                     uint32_t *status);
int cabac_decode_residual_i16(cabac_handle *h, const uint8_t *in,
                              const uint64_t *in_offsets, const cabac_tu_desc *tus,
                              const uint64_t *tu_offsets,
                              const cabac_substream_params *params, size_t n_sub,
                              const cabac_ctx_model *ctx_in, int16_t *coeffs,
                              size_t n_coeffs, uint32_t *num_bins, uint32_t *bits_read,
                              uint8_t *tu_flags, cabac_ctx_model *ctx_out,
                              uint32_t *status);

/* ---- WPP: CTU rows with context hand-over, one launch ---------------------- */
/* With entropy_coding_sync_enabled a slice is one substream per CTU row and row r
 * starts from the contexts row r-1 held after its FIRST CTU
 * (VTM Lib/EncoderLib/EncSlice.cpp:2057-2072 load, :2085-2089 store;
 * Lib/DecoderLib/DecSlice.cpp:194-210, 241-245; the copy is Ctx::operator=,
 * src/common/contexts.cpp:1096-1108).  The *_wpp entry points take one
 * cabac_wpp_sync per substream and resolve the hand-over on the device: the
 * predecessor's warp publishes its contexts at its sync point, the dependant's
 * warp waits for them - every CTU is coded once, in one launch.
 *   dep        index of the substream this one starts from, CABAC_WPP_NONE for
 *              a row that starts from Ctx::init / its ctx_in row (first row of
 *              a slice or tile).  Must be < the substream's own index.
 *   sync_tus   residual entry points: the substream's contexts are published
 *              after this many of its TUs (its first CTU)
 *   sync_cmds  stream entry point: ... after this many of its command words
 *              (TU markers count as one word)
 * A substream shorter than its sync point publishes at its end.             */
#define CABAC_WPP_NONE 0xFFFFFFFFu
typedef struct cabac_wpp_sync {
  uint32_t dep;
  uint32_t sync_tus;
  uint32_t sync_cmds;
  uint32_t reserved;
} cabac_wpp_sync;

int cabac_encode_residual_wpp_dev(cabac_handle *h, const cabac_tu_desc *tus,
                                  const uint64_t *tu_offsets,
                                  const cabac_substream_params *params,
                                  const cabac_wpp_sync *wpp, size_t n_sub,
                                  const int32_t *coeffs,
                                  const cabac_ctx_model *ctx_in, uint8_t *out,
                                  const uint64_t *out_offsets, uint32_t *out_sizes,
                                  uint32_t *num_bins, uint8_t *tu_flags,
                                  cabac_ctx_model *ctx_out, uint32_t *status);
/* Host pointers; returns the packed payload + entry points like
 * cabac_encode_residual_packed.                                             */
int cabac_encode_residual_wpp(cabac_handle *h, const cabac_tu_desc *tus,
                              const uint64_t *tu_offsets,
                              const cabac_substream_params *params,
                              const cabac_wpp_sync *wpp, size_t n_sub,
                              const int32_t *coeffs, size_t n_coeffs,
                              const cabac_ctx_model *ctx_in,
                              const uint64_t *out_offsets, uint8_t *packed,
                              size_t packed_cap, uint64_t *entry_points,
                              uint32_t *out_sizes, uint32_t *num_bins,
                              uint8_t *tu_flags, cabac_ctx_model *ctx_out,
                              uint32_t *status);
int cabac_decode_residual_wpp_dev(cabac_handle *h, const uint8_t *in,
                                  const uint64_t *in_offsets,
                                  const cabac_tu_desc *tus,
                                  const uint64_t *tu_offsets,
                                  const cabac_substream_params *params,
                                  const cabac_wpp_sync *wpp, size_t n_sub,
                                  const cabac_ctx_model *ctx_in, int32_t *coeffs,
                                  uint32_t *num_bins, uint32_t *bits_read,
                                  uint8_t *tu_flags, cabac_ctx_model *ctx_out,
                                  uint32_t *status);
int cabac_decode_residual_wpp(cabac_handle *h, const uint8_t *in,
                              const uint64_t *in_offsets, const cabac_tu_desc *tus,
                              const uint64_t *tu_offsets,
                              const cabac_substream_params *params,
                              const cabac_wpp_sync *wpp, size_t n_sub,
                              const cabac_ctx_model *ctx_in, int32_t *coeffs,
                              size_t n_coeffs, uint32_t *num_bins,
                              uint32_t *bits_read, uint8_t *tu_flags,
                              cabac_ctx_model *ctx_out, uint32_t *status);
/* Recorded command streams (the CABACWriter seam) with WPP hand-over.       */
int cabac_encode_stream_wpp_dev(cabac_handle *h, const uint32_t *cmds,
                                const uint64_t *cmd_offsets,
                                const cabac_tu_desc *tus,
                                const uint64_t *tu_offsets,
                                const cabac_substream_params *params,
                                const cabac_wpp_sync *wpp, size_t n_sub,
                                const int32_t *coeffs,
                                const cabac_ctx_model *ctx_in, uint8_t *out,
                                const uint64_t *out_offsets, uint32_t *out_sizes,
                                uint32_t *num_bins, uint8_t *tu_flags,
                                cabac_ctx_model *ctx_out, uint32_t *status);
int cabac_encode_stream_wpp(cabac_handle *h, const uint32_t *cmds,
                            const uint64_t *cmd_offsets, const cabac_tu_desc *tus,
                            const uint64_t *tu_offsets,
                            const cabac_substream_params *params,
                            const cabac_wpp_sync *wpp, size_t n_sub,
                            const int32_t *coeffs, size_t n_coeffs,
                            const cabac_ctx_model *ctx_in, uint8_t *out,
                            const uint64_t *out_offsets, uint32_t *out_sizes,
                            uint32_t *num_bins, uint8_t *tu_flags,
                            cabac_ctx_model *ctx_out, uint32_t *status);

/* ---- substream packing + entry points ------------------------------------ */
/* Compacts the per-substream output slabs of an encode call into one
 * contiguous payload in substream order (OutputBitstream::addSubstream for
 * byte-aligned substreams, bit_stream.cpp:139-150; EncGOP.cpp:3811-3824) and
 * returns the entry-point table: entry_points[i] = byte offset of substream i
 * in `packed`, entry_points[n_sub] = total bytes.  packed must hold
 * sum(out_sizes).  Device pointers; asynchronous.
 */
int cabac_pack_dev(cabac_handle *h, const uint8_t *out,
                   const uint64_t *out_offsets, const uint32_t *out_sizes,
                   size_t n_sub, uint8_t *packed, uint64_t *entry_points);
int cabac_pack(cabac_handle *h, const uint8_t *out, const uint64_t *out_offsets,
               const uint32_t *out_sizes, size_t n_sub, uint8_t *packed,
               size_t packed_cap, uint64_t *entry_points);

/* ---- RDO bit estimation: the encoder-side caller of the path ---------------- */
/* CABACWriter::residual_coding() (cabac_writer.cpp:2438-2585) over a TU list with the
 * writer bound to a BitEstimator_Std (arith_codec.hpp:189-259, arith_codec.cpp:603-711)
 * instead of the arithmetic coder - what an encoder's rate-distortion search calls
 * to price a candidate.  est_bits[i] = getEstFracBits() of substream i after its last
 * TU, in 1/32768 bit (SCALE_BITS = 15): m_binFracBits[state()][bin] per context-coded
 * bin (contexts.cpp:791-878, 922-929), 1 << 15 per bypass bin (contexts.cpp:880-884).
 * est_per_tu (nullable) [n_tu] = the running total after each TU, i.e. the value an
 * encoder reads between two residual_coding() calls.  Contexts evolve exactly as in
 * cabac_encode_residual; CABAC_SUB_TERMINATE / CABAC_SUB_ALIGN are ignored (the
 * estimator has neither; num_bins therefore counts the residual syntax only, one less
 * than cabac_encode_residual's for a terminated substream).  num_bins / tu_flags /
 * ctx_out / status nullable.
 * One launch prices every candidate list of a batch: n_sub is the number of
 * candidates, ctx_in their (shared or individual) starting contexts. */
int cabac_estimate_residual_bits_dev(cabac_handle *h, const cabac_tu_desc *tus,
                                     const uint64_t *tu_offsets,
                                     const cabac_substream_params *params,
                                     size_t n_sub, const int32_t *coeffs,
                                     const cabac_ctx_model *ctx_in,
                                     uint64_t *est_bits, uint64_t *est_per_tu,
                                     uint32_t *num_bins, uint8_t *tu_flags,
                                     cabac_ctx_model *ctx_out, uint32_t *status);
int cabac_estimate_residual_bits(cabac_handle *h, const cabac_tu_desc *tus,
                                 const uint64_t *tu_offsets,
                                 const cabac_substream_params *params,
                                 size_t n_sub, const int32_t *coeffs,
                                 size_t n_coeffs, const cabac_ctx_model *ctx_in,
                                 uint64_t *est_bits, uint64_t *est_per_tu,
                                 uint32_t *num_bins, uint8_t *tu_flags,
                                 cabac_ctx_model *ctx_out, uint32_t *status);
/* ---- inter-slice init-table choice: CABACWriter::getCtxInitId ----------------
 * Replaces the two inputs and the decision of CABACWriter::getCtxInitId /
 * xGetCtxInitId (cabac_writer.cpp:41-81) for the residual syntax coded on the device.
 *
 * cabac_residual_ctx_bins: ctx_bins[i * CABAC_NUM_CTX + k] = BinEncIf::getNumBins(k)
 * (BinCounter::addCtx, arith_codec.hpp:72-93, arith_codec.cpp:281-316) after
 * residual_coding over TU list i - the context-coded bins per context id; bypass and
 * terminate bins are not part of it, as in the reference.  Contexts evolve exactly as
 * in cabac_encode_residual (ctx_out = getCtx() after the list).  A host shim adds the
 * rows to the counts of the bins it recorded itself (GpuBinEncoder::getNumBins(k)).
 * num_bins / ctx_out / status nullable.
 *
 * cabac_ctx_init_id: for slice i with QP params[i].qp, final contexts
 * ctx_ref[i * CABAC_NUM_CTX ..] and bin counts ctx_bins[i * CABAC_NUM_CTX ..]:
 * costs[2i + t] = sum_k ctx_bins[k] * estFracExcessBits(ctx_ref[k], model k of init
 * table t at that QP) (contexts.cpp:965-968), t = 0 (B_SLICE), 1 (P_SLICE);
 * choice[i] = the cheaper t (ties keep B), or 2 (I_SLICE) when params[i].init_id == 2
 * as the reference returns for an intra slice.  costs nullable. */
int cabac_residual_ctx_bins_dev(cabac_handle *h, const cabac_tu_desc *tus,
                                const uint64_t *tu_offsets,
                                const cabac_substream_params *params, size_t n_sub,
                                const int32_t *coeffs, const cabac_ctx_model *ctx_in,
                                uint32_t *ctx_bins, uint32_t *num_bins,
                                cabac_ctx_model *ctx_out, uint32_t *status);
int cabac_residual_ctx_bins(cabac_handle *h, const cabac_tu_desc *tus,
                            const uint64_t *tu_offsets,
                            const cabac_substream_params *params, size_t n_sub,
                            const int32_t *coeffs, size_t n_coeffs,
                            const cabac_ctx_model *ctx_in, uint32_t *ctx_bins,
                            uint32_t *num_bins, cabac_ctx_model *ctx_out,
                            uint32_t *status);
int cabac_ctx_init_id_dev(cabac_handle *h, const cabac_substream_params *params,
                          const cabac_ctx_model *ctx_ref, const uint32_t *ctx_bins,
                          size_t n, uint8_t *choice, uint64_t *costs);
int cabac_ctx_init_id(cabac_handle *h, const cabac_substream_params *params,
                      const cabac_ctx_model *ctx_ref, const uint32_t *ctx_bins,
                      size_t n, uint8_t *choice, uint64_t *costs);
/* m_binFracBits flattened as table_out[2 * state + bin], 512 entries. */
int cabac_frac_bits_table(uint32_t *table_out);

/* ---- start-code emulation: the step right after the path ------------------- */
/* counts[i] = OutputBitstream::countStartCodeEmulations() of segment
 * data[offsets[i] .. offsets[i+1]) (bit_stream.cpp:157-181): the number of
 * emulation_prevention_three_bytes the NAL writer will insert into substream i,
 * which VTM adds to the substream's entry-point offset (EncGOP.cpp:3811-3830). */
int cabac_count_emulations_dev(cabac_handle *h, const uint8_t *data,
                               const uint64_t *offsets, size_t n_seg,
                               uint32_t *counts);
int cabac_count_emulations(cabac_handle *h, const uint8_t *data,
                           const uint64_t *offsets, size_t n_seg,
                           uint32_t *counts);
/* RBSP -> NAL payload: inserts 0x03 before every byte <= 3 that follows two
 * zero bytes (zero count restarting after each insertion) and after a trailing
 * zero byte - writeNaluContent, VTM Lib/EncoderLib/NALwrite.cpp:67-126.
 * out must hold up to n + n/2 + 1 bytes; *out_size receives the length
 * (device memory for the _dev flavour).                                      */
int cabac_insert_emulation_prevention_dev(cabac_handle *h, const uint8_t *rbsp,
                                          size_t n, uint8_t *out,
                                          size_t out_cap, uint64_t *out_size);
int cabac_insert_emulation_prevention(cabac_handle *h, const uint8_t *rbsp,
                                      size_t n, uint8_t *out, size_t out_cap,
                                      uint64_t *out_size);

/* ---- the decode side of substream assembly -------------------------------- */
/* NAL unit bytes -> RBSP: convertPayloadToRBSP, VTM Lib/DecoderLib/NALread.cpp:59-106.
 * Every emulation_prevention_three_byte (a 0x03 after two zero bytes) is dropped and
 * its position in the ESCAPED buffer recorded (InputBitstream::
 * pushEmulationPreventionByteLocation) - the slice-header parser needs the positions
 * to turn signalled entry points into RBSP ranges.
 *   strip_trailing_zeros  VCL NAL units: drop trailing cabac_zero_words (NALread.cpp:92-103)
 *   rbsp                  n bytes of room
 *   sizes[2]              [0] RBSP length, [1] number of EPBs
 *   epb_locations         nullable; n / 3 + 1 entries of room, ascending
 *   violations            number of positions where VTM's CHECKs throw (00 00 0x with
 *                         x < 3, a byte > 3 after an EPB, a trailing zero byte); the
 *                         host flavour returns CABAC_ERR_ARG when it is non-zero       */
int cabac_remove_emulation_prevention_dev(cabac_handle *h, const uint8_t *nal, size_t n,
                                          int strip_trailing_zeros, uint8_t *rbsp,
                                          uint64_t *sizes, uint32_t *epb_locations,
                                          uint32_t *violations);
int cabac_remove_emulation_prevention(cabac_handle *h, const uint8_t *nal, size_t n,
                                      int strip_trailing_zeros, uint8_t *rbsp,
                                      uint64_t *sizes, uint32_t *epb_locations,
                                      uint32_t *violations);
/* Slice-header form of the entry points.  offset_minus1[i] (i < n_sub - 1) =
 * sh_entry_point_offset_minus1 = sizes[i] + emulations[i] - 1 (EncSlice.cpp:2118-2125:
 * the signalled size counts the substream's emulation prevention bytes; emulations from
 * cabac_count_emulations, nullable = none); *offset_len_minus1 =
 * sh_entry_offset_len_minus1 (HLSWriter::codeTilesWPPEntryPoint, VLCWriter.cpp:2833-2866). */
int cabac_entry_point_syntax_dev(cabac_handle *h, const uint32_t *sizes,
                                 const uint32_t *emulations, size_t n_sub,
                                 uint32_t *offset_minus1, uint32_t *offset_len_minus1);
int cabac_entry_point_syntax(cabac_handle *h, const uint32_t *sizes,
                             const uint32_t *emulations, size_t n_sub,
                             uint32_t *offset_minus1, uint32_t *offset_len_minus1);
/* The inverse on the decoder: RBSP byte ranges of the n_entry + 1 substreams of a
 * slice from the signalled offsets and the EPB locations (HLSyntaxReader::
 * parseSliceHeader, VLCReader.cpp:4346-4381: each signalled size minus the EPBs inside
 * it; DecSlice.cpp:106-114 + InputBitstream::extractSubstream, bit_stream.cpp:382-415:
 * consecutive ranges, the last substream takes the rest).
 *   slice_data_pos  RBSP byte position of the first slice-data byte
 *   in_offsets      n_entry + 2 entries: substream i is rbsp[in_offsets[i] ..
 *                   in_offsets[i + 1]) - the in / in_offsets arguments of cabac_decode_*   */
int cabac_substream_ranges_dev(cabac_handle *h, const uint32_t *offset_minus1,
                               size_t n_entry, const uint32_t *epb_locations,
                               size_t n_epb, uint64_t slice_data_pos,
                               uint64_t rbsp_size, uint64_t *in_offsets);
int cabac_substream_ranges(cabac_handle *h, const uint32_t *offset_minus1,
                           size_t n_entry, const uint32_t *epb_locations, size_t n_epb,
                           uint64_t slice_data_pos, uint64_t rbsp_size,
                           uint64_t *in_offsets);


/* ---- whole substreams: the CTU-level syntax on the device ---------------
 * Replaces, for all-intra substreams with separate luma / chroma trees, the decoder's per-CTU call
 *   CABACReader::coding_tree_unit(cs, area, qps, ctuRsAddr)          cabac_reader.cpp:103-209
 * and everything it calls (sao :284, readAlfCtuFilterIndex :211, ccAlfFilterControlIdc :237,
 * coding_tree :429, split_cu_mode :677, coding_unit :764, bdpcm_mode :1063, intra_luma_pred_modes
 * :1231, extend_ref_line :1188, isp_mode :2785, mip_flag / mip_pred_mode :3381-3408,
 * intra_chroma_pred_mode :1353, cu_residual :1396, transform_tree :2297, transform_unit :2456,
 * cbf_comp :2353, cu_qp_delta :2578, cu_chroma_qp_offset :2602, joint_cb_cr :2632, residual_coding
 * :2647, mts_idx :2754, residual_lfnst_mode :2808), the QTBTPartitioner's legality rules
 * (unit_partitioner.cpp:303-470) and DeriveCtx::CtxSplit / CtxMipFlag / CtxQtCbf
 * (context_modelling.cpp:108-348), followed by the slice's terminating bit (DecSlice.cpp:241-262).
 * One call parses n_sub independent substreams (slices / tiles) from their bytes alone: no TU list
 * is needed, the kernel emits it.  What the reference keeps in CodingStructure / CodingUnit /
 * TransformUnit comes back as flat records.                                                       */
enum { /* cabac_slice_params.tools */
  CABAC_SL_DUAL_TREE = 1 << 0,   /* CS::isDualITree: I slice, sps qtbtt_dual_tree_intra_flag    */
  CABAC_SL_SAO_LUMA = 1 << 1,    /* sps SAO && slice.getSaoEnabledFlag(CHANNEL_TYPE_LUMA)       */
  CABAC_SL_SAO_CHROMA = 1 << 2,
  CABAC_SL_ALF_Y = 1 << 3,       /* sps ALF && slice.getAlfEnabledFlag(COMPONENT_Y)             */
  CABAC_SL_ALF_CB = 1 << 4,
  CABAC_SL_ALF_CR = 1 << 5,
  CABAC_SL_CCALF_CB = 1 << 6,    /* slice.m_ccAlfFilterParam.ccAlfFilterEnabled[0]              */
  CABAC_SL_CCALF_CR = 1 << 7,
  CABAC_SL_MTS_INTRA = 1 << 8,   /* sps.getUseIntraMTS()                                        */
  CABAC_SL_LFNST = 1 << 9,
  CABAC_SL_ISP = 1 << 10,
  CABAC_SL_MIP = 1 << 11,
  CABAC_SL_MRL = 1 << 12,
  CABAC_SL_TS = 1 << 13,         /* sps.getTransformSkipEnabledFlag()                           */
  CABAC_SL_BDPCM = 1 << 14,
  CABAC_SL_JCCR = 1 << 15,
  CABAC_SL_CCLM = 1 << 16,
  CABAC_SL_DQP = 1 << 17,        /* pps.getUseDQP()                                             */
  CABAC_SL_CHROMA_QP_ADJ = 1 << 18,
  CABAC_SL_WPP = 1 << 19,        /* entropy coding sync: the substream's bytes are the slice's
                                    CTU-row substreams back to back; rows are parsed one after
                                    the other with the context hand-over of DecSlice.cpp:194-245 */
  CABAC_SL_IBC = 1 << 20,        /* refused                                                     */
  CABAC_SL_PLT = 1 << 21,        /* refused                                                     */
  CABAC_SL_ACT = 1 << 22,        /* refused                                                     */
  CABAC_SL_INTRA = 1 << 23,      /* slice.isIntra(); P / B slices are refused                   */
  CABAC_SL_MTS = 1 << 24         /* sps.getUseMTS()                                             */
};
/* Field order == word order of the fixture records (tests/golden/vtm_record). */
typedef struct cabac_slice_params {
  uint32_t pic_w, pic_h;        /* pps.getPicWidth/HeightInLumaSamples                          */
  uint32_t x0, y0, x1, y1;      /* the CTU rectangle this substream covers, luma samples,
                                   x1 / y1 exclusive and clipped to the picture; neighbours
                                   outside it are unavailable (other slice / tile)               */
  uint32_t log2_ctu;            /* 5..7                                                          */
  uint32_t log2_min_qt[2];      /* [luma tree, chroma tree], luma samples (PreCalcValues)        */
  uint32_t log2_max_bt[2];
  uint32_t log2_max_tt[2];
  uint32_t max_mtt_depth[2];
  uint32_t log2_min_bt[2];
  uint32_t log2_min_tt[2];
  uint32_t log2_max_tb;         /* sps.getMaxTbSize()                                            */
  uint32_t log2_max_ts;         /* sps.getLog2MaxTransformSkipBlockSize()                        */
  uint32_t bit_depth;           /* luma; SAO offset range                                        */
  uint32_t alf_num_aps_luma;    /* slice.getNumAlfApsIdsLuma()                                   */
  uint32_t alf_chroma_num_alts; /* AlfParam::numAlternativesChroma of the chroma APS             */
  uint32_t ccalf_filter_count[2];
  uint32_t cu_qp_delta_subdiv;  /* slice.getCuQpDeltaSubdiv()                                    */
  uint32_t cu_chroma_qp_offset_subdiv;
  uint32_t chroma_qp_offset_list_len;
  uint32_t chroma_format;       /* 1 = 4:2:0 (anything else is refused)                          */
  uint32_t tools;               /* CABAC_SL_*                                                    */
  uint32_t reserved;
} cabac_slice_params;

enum { /* cabac_cu_desc.flags */
  CABAC_CU_MIP = 1,            /* cu.mipFlag                                                    */
  CABAC_CU_MIP_TRANSPOSED = 2, /* pu.mipTransposedFlag                                          */
  CABAC_CU_MPM = 4,            /* intra_luma_mpm_flag (inferred 1 with a reference line != 0)    */
  CABAC_CU_CCLM = 8,           /* chroma tree: cclm_mode_flag                                    */
  CABAC_CU_JCCR = 16,          /* some TU of the CU has tu.jointCbCr != 0                        */
  CABAC_CU_DQP = 32,           /* cu_qp_delta_abs was coded in this CU (value in qp_delta)       */
  CABAC_CU_FROM_DIR = 64       /* cabac_encode_slices only: take the prediction mode from intra_dir
                                  (pu.intraDir[chType]) and derive CABAC_CU_MPM / CABAC_CU_CCLM and
                                  pred_syntax from it (the MPM list and the chroma candidates are
                                  derived on the device as the writer derives them); without it
                                  pred_syntax is coded as given                                   */
};
/* One record per CodingUnit in decoding order: what coding_unit() leaves in CodingUnit /
 * PredictionUnit (unit.hpp:283-341), the syntax element values and the intra direction derived
 * from them. */
typedef struct cabac_cu_desc {
  uint16_t x, y;          /* top-left in luma samples                                           */
  uint8_t log2_w, log2_h; /* luma samples (a chroma-tree CU covers half of that per dimension)  */
  uint8_t ch_type;        /* 0 = luma tree, 1 = chroma tree                                     */
  uint8_t qt_depth, mt_depth;
  uint8_t flags;          /* CABAC_CU_*                                                         */
  uint8_t bdpcm;          /* cu.bdpcmMode (luma tree) / cu.bdpcmModeChroma: 0, 1 = hor, 2 = ver */
  uint8_t mrl_idx;        /* pu.multiRefIdx                                                     */
  uint8_t isp;            /* cu.ispMode: 0, 1 = horizontal, 2 = vertical                        */
  uint8_t lfnst_idx;      /* cu.lfnstIdx                                                        */
  uint8_t mts_idx;        /* firstTU mtsIdx[Y]: 0 DCT2, 1 transform skip, 2..5 explicit MTS      */
  uint8_t pred_syntax;    /* luma: MIP mode | mpm_idx (0 = planar) | rem_intra_luma_pred_mode;
                             chroma: 0 = DM, 1..4 = intra_chroma_pred_mode candidate + 1,
                             with CABAC_CU_CCLM: cclm_mode_idx 0..2                              */
  uint32_t first_tu;      /* index of the CU's first record in this substream's TU list          */
  uint16_t n_tu;          /* coded blocks (cbf = 1, residual_coding called)                      */
  int8_t qp_delta;        /* CuQpDeltaVal when CABAC_CU_DQP                                      */
  uint8_t chroma_qp_adj;  /* cu_chroma_qp_offset_flag + idx                                      */
  uint8_t intra_dir;      /* pu.intraDir[chType] as the reader leaves it: luma 0..66 from the MPM
                             list (PU::getIntraMPMs, unit_tools.cpp:485-577) or the MIP mode;
                             chroma 0..66, 67..69 = LM / MDLM_L / MDLM_T, 70 = DM_CHROMA_IDX      */
  uint8_t reserved;
  uint16_t cbf_mask;      /* coded block flags as parsed: bits 0-3 luma (ISP sub-partition / transform tile i),
                             4-7 Cb, 8-11 Cr (per tile), 12-15 tu_joint_cbcr_residual_flag (per tile)     */
  uint32_t split_series;  /* the path from the tree's root (the CTU; the 64x64 node of a 128x128 dual-tree
                             CTU) to this CU: 3 bits per depth, 1 QT, 2 BT hor, 3 BT ver, 4 TT hor, 5 TT ver
                             (CU::getSplitAtDepth, unit_tools.cpp); 0 from the CU's own depth on          */
} cabac_cu_desc;

/* Per-CTU in-loop filter syntax (SAOBlkParam, sample_adaptive_offset.hpp; the ALF CTU maps of
 * Picture, picture.hpp), one record per CTU in decoding order. */
typedef struct cabac_ctu_desc {
  uint8_t sao_mode[3];      /* 0 off, 1 new, 2 merge                                            */
  uint8_t sao_type[3];      /* new: 0 band, 1..4 = EO class + 1; merge: 0 left, 1 above          */
  uint8_t sao_band[3];      /* sao_band_position                                                 */
  int8_t sao_offset[3][4];  /* signed as parsed (band) / magnitudes (edge)                       */
  uint8_t alf_flag[3];      /* alf_ctb_flag per component                                        */
  uint8_t alf_filter_idx;   /* AlfCtbFilterSetIndex (luma)                                       */
  uint8_t alf_alt[2];       /* alf_ctb_filter_alt_idx Cb, Cr                                     */
  uint8_t ccalf_idc[2];     /* filterControlIdc Cb, Cr                                           */
  uint8_t reserved[3];
} cabac_ctu_desc;

#define CABAC_SS_UNSUPPORTED 256 /* cabac_decode_slices: a tool outside the built scope (P / B
                                    slice, single tree, IBC, palette, ACT, non-4:2:0)           */
#define CABAC_SS_CAPACITY 512    /* cabac_decode_slices: a record or coefficient region is full */

/* Capacities: substream i may write cus[cu_offsets[i] .. cu_offsets[i+1]), likewise tus / ctus /
 * coeffs (TCoeff units; a region of 3/2 x the rectangle's luma area always suffices, and
 * area / 16 x 3/2 TU records, area / 16 x 2 CU records).  TU records come back in the layout
 * cabac_decode_residual takes (coeff_offset absolute in `coeffs`; flags = CABAC_TU_* as the
 * encoder-side recorder would have set them, CABAC_TU_TS holding the decoded flag), tu_flags in
 * CABAC_TUR_*.  counts[3*i + 0..2] = CUs, TUs, coefficients written by substream i.
 * All pointers are device pointers in the _dev form and host pointers otherwise.                 */
int cabac_decode_slices_dev(cabac_handle *h, const uint8_t *in, const uint64_t *in_offsets,
                            const cabac_substream_params *params,
                            const cabac_slice_params *slice_params, size_t n_sub,
                            const cabac_ctx_model *ctx_in, const uint64_t *cu_offsets,
                            const uint64_t *tu_offsets, const uint64_t *ctu_offsets,
                            const uint64_t *coeff_offsets, cabac_cu_desc *cus, cabac_tu_desc *tus,
                            uint8_t *tu_flags, cabac_ctu_desc *ctus, int32_t *coeffs,
                            uint32_t *counts, uint32_t *num_bins, uint32_t *bits_read,
                            cabac_ctx_model *ctx_out, uint32_t *status);
int cabac_decode_slices(cabac_handle *h, const uint8_t *in, const uint64_t *in_offsets,
                        const cabac_substream_params *params,
                        const cabac_slice_params *slice_params, size_t n_sub,
                        const cabac_ctx_model *ctx_in, const uint64_t *cu_offsets,
                        const uint64_t *tu_offsets, const uint64_t *ctu_offsets,
                        const uint64_t *coeff_offsets, cabac_cu_desc *cus, cabac_tu_desc *tus,
                        uint8_t *tu_flags, cabac_ctu_desc *ctus, int32_t *coeffs,
                        uint32_t *counts, uint32_t *num_bins, uint32_t *bits_read,
                        cabac_ctx_model *ctx_out, uint32_t *status);

/* Writer direction of the same syntax: CABACWriter::coding_tree_unit (cabac_writer.cpp:115) and what it
 * calls - sao :196, coding_tree :355, split_cu_mode :514, coding_unit :586, intra_luma_pred_modes :949,
 * intra_chroma_pred_mode :1167, transform_tree :2059, transform_unit :2219, cu_qp_delta :2356,
 * residual_coding :2424 - as one call:
 * records in, slice payloads out, byte-identical with the reference writer's.  The records are exactly the
 * ones cabac_decode_slices produces (an encoder fills them from its CodingStructure the way
 * host/gpu_slice_reader.hpp reads them back): CUs in coding order with split_series, cbf_mask and the
 * prediction syntax, TUs in coding order with coeff_offset absolute in `coeffs` and tu_flags (CABAC_TUR_*,
 * from cabac_decode_slices or from cabac_estimate_residual on the same TUs), one CTU record per CTU.
 * Here the offsets are exact: slice i owns cus[cu_offsets[i] .. cu_offsets[i+1]) etc.
 * One warp per slice walks the partitioning and writes the slice's BinEncIf call sequence as command words
 * (into handle-owned staging sized by the record counts: 80 words per CU + 1 per TU + 450 per CTU), a copy
 * packs the slices into cmd_scratch (cmd_scratch_words uint32 available; the packed stream needs one word per
 * bin outside the residuals plus one per TU - num_bins of the slices always suffices); the stream encoder
 * (cabac_encode_stream_dev) then codes them, terminating bin and finish() included
 * (params[i].flags: CABAC_SUB_TERMINATE | CABAC_SUB_ALIGN as for every slice payload).
 * status[i]: CABAC_SS_* of the stream encoder, plus CABAC_SS_UNSUPPORTED (see cabac_decode_slices; also
 * CABAC_SL_WPP slices, whose rows are separate substreams: cabac_encode_slices_wpp), CABAC_SS_BAD_DESC (records that do
 * not describe a legal partitioning of the rectangle, or left-over records) and CABAC_SS_CAPACITY
 * (cmd_scratch too small: nothing is encoded and the call returns CABAC_ERR_SUBSTREAM).
 * The _dev form synchronises the stream twice (it reads the record totals and the total command count back).
 * The host-pointer form returns CABAC_ERR_ARG when a TU record's coeff_offset + area exceeds n_coeffs.     */
int cabac_encode_slices_dev(cabac_handle *h, const cabac_substream_params *params,
                            const cabac_slice_params *slice_params, size_t n_sub,
                            const cabac_cu_desc *cus, const uint64_t *cu_offsets,
                            const cabac_tu_desc *tus, const uint8_t *tu_flags,
                            const uint64_t *tu_offsets, const cabac_ctu_desc *ctus,
                            const uint64_t *ctu_offsets, const int32_t *coeffs,
                            const cabac_ctx_model *ctx_in, uint32_t *cmd_scratch,
                            size_t cmd_scratch_words, uint8_t *out, const uint64_t *out_offsets,
                            uint32_t *out_sizes, uint32_t *num_bins, cabac_ctx_model *ctx_out,
                            uint32_t *status);
int cabac_encode_slices(cabac_handle *h, const cabac_substream_params *params,
                        const cabac_slice_params *slice_params, size_t n_sub,
                        const cabac_cu_desc *cus, const uint64_t *cu_offsets,
                        const cabac_tu_desc *tus, const uint8_t *tu_flags,
                        const uint64_t *tu_offsets, const cabac_ctu_desc *ctus,
                        const uint64_t *ctu_offsets, const int32_t *coeffs, size_t n_coeffs,
                        const cabac_ctx_model *ctx_in, size_t cmd_words, uint8_t *out,
                        const uint64_t *out_offsets, uint32_t *out_sizes, uint32_t *num_bins,
                        cabac_ctx_model *ctx_out, uint32_t *status);

/* The same for slices with entropy_coding_sync_enabled (CABAC_SL_WPP) - and any mix of WPP and other slices.
 * The output is per ROW substream: a WPP slice has one per CTU row (EncSlice.cpp:2057-2089: the row's
 * end_of_subset_one_bit, finish() and alignment close it, the next row starts from the contexts held after the
 * first CTU of the row above), any other slice has one.  row_offsets[n_sub + 1] is the exclusive scan of the
 * rows per slice (a WPP slice: ceil((y1 - y0) / CTU size); otherwise 1), n_rows = row_offsets[n_sub];
 * out_offsets[n_rows + 1], out_sizes, num_bins, ctx_out ([n_rows][CABAC_NUM_CTX]) and status are indexed by row
 * substream (slice i's rows are row_offsets[i] .. row_offsets[i + 1]; their payloads back to back are the slice
 * data, their sizes the entry points: cabac_pack_dev).  The records are one
 * list per slice as above - the walk cuts its command stream at the row ends and notes the hand-over point of
 * every row; the stream encoder's wavefront mode (cabac_encode_stream_wpp_dev) codes all rows in one launch,
 * every CTU once.  A slice whose row count does not match its rectangle gets CABAC_SS_BAD_DESC.            */
int cabac_encode_slices_wpp_dev(cabac_handle *h, const cabac_substream_params *params,
                                const cabac_slice_params *slice_params, size_t n_sub,
                                const cabac_cu_desc *cus, const uint64_t *cu_offsets,
                                const cabac_tu_desc *tus, const uint8_t *tu_flags,
                                const uint64_t *tu_offsets, const cabac_ctu_desc *ctus,
                                const uint64_t *ctu_offsets, const int32_t *coeffs,
                                const cabac_ctx_model *ctx_in, uint32_t *cmd_scratch,
                                size_t cmd_scratch_words, const uint64_t *row_offsets,
                                size_t n_rows, uint8_t *out, const uint64_t *out_offsets,
                                uint32_t *out_sizes, uint32_t *num_bins,
                                cabac_ctx_model *ctx_out, uint32_t *status);
int cabac_encode_slices_wpp(cabac_handle *h, const cabac_substream_params *params,
                            const cabac_slice_params *slice_params, size_t n_sub,
                            const cabac_cu_desc *cus, const uint64_t *cu_offsets,
                            const cabac_tu_desc *tus, const uint8_t *tu_flags,
                            const uint64_t *tu_offsets, const cabac_ctu_desc *ctus,
                            const uint64_t *ctu_offsets, const int32_t *coeffs, size_t n_coeffs,
                            const cabac_ctx_model *ctx_in, size_t cmd_words,
                            const uint64_t *row_offsets, size_t n_rows, uint8_t *out,
                            const uint64_t *out_offsets, uint32_t *out_sizes, uint32_t *num_bins,
                            cabac_ctx_model *ctx_out, uint32_t *status);

#ifdef __cplusplus
}
#endif
#endif /* CABAC_B200_H */
