/* oracle/cabac_oracle.h - TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Entry points of the plain-C restatement (cabac_oracle.c).  Signatures are
 * identical to oracle/ref_driver.cpp's `ref_*` functions so oracle/checker.py
 * drives both through the same code.
 */
#ifndef CABAC_ORACLE_H
#define CABAC_ORACLE_H
#include <stdint.h>

#include "cabac_b200.h"

int orc_num_contexts(void);
void orc_init_table(int row, uint8_t *out);
void orc_ctx_init(int qp, int initId, cabac_ctx_model *out, uint8_t *rates);
int orc_scan_table(int log2w, int log2h, uint32_t *out);
void orc_ctx_offsets(uint32_t *out);

long orc_encode_bins(const uint32_t *cmds, long ncmd, const cabac_substream_params *p,
                     const cabac_ctx_model *ctxIn, uint8_t *out, long cap, uint32_t *numBins,
                     cabac_ctx_model *ctxOut);
long orc_decode_bins(const uint8_t *in, long nbytes, const uint32_t *cmds, long ncmd,
                     const cabac_substream_params *p, const cabac_ctx_model *ctxIn, uint32_t *values,
                     cabac_ctx_model *ctxOut, uint32_t *status);
long orc_encode_residual(const cabac_tu_desc *tus, long ntu, const cabac_substream_params *p,
                         const int32_t *coeffs, const cabac_ctx_model *ctxIn, uint8_t *out, long cap,
                         uint32_t *numBins, uint8_t *tuFlags, cabac_ctx_model *ctxOut, int useMTS);
long orc_encode_stream(const uint32_t *cmds, long ncmd, const cabac_tu_desc *tus, long ntu,
                       const cabac_substream_params *p, const int32_t *coeffs, const cabac_ctx_model *ctxIn,
                       uint8_t *out, long cap, uint32_t *numBins, uint8_t *tuFlags, cabac_ctx_model *ctxOut);
long orc_decode_residual(const uint8_t *in, long nbytes, const cabac_tu_desc *tus, long ntu,
                         const cabac_substream_params *p, const cabac_ctx_model *ctxIn, int32_t *coeffs,
                         uint8_t *tuFlags, cabac_ctx_model *ctxOut, uint32_t *status, int useMTS);
double orc_time_encode_residual(const cabac_tu_desc *tus, const uint64_t *tuOffsets,
                                const cabac_substream_params *params, long nSub, const int32_t *coeffs,
                                int nthreads, uint64_t *totalBins, uint64_t *totalBytes, uint8_t *out,
                                const uint64_t *outOffsets, uint32_t *outSizes);
double orc_time_decode_residual(const uint8_t *in, const uint64_t *inOffsets, const uint32_t *inSizes,
                                const cabac_tu_desc *tus, const uint64_t *tuOffsets,
                                const cabac_substream_params *params, long nSub, int32_t *coeffs,
                                int nthreads);
void orc_frac_bits(uint32_t *out);
uint64_t orc_estimate_residual(const cabac_tu_desc *tus, long ntu, const cabac_substream_params *p,
                               const int32_t *coeffs, const cabac_ctx_model *ctxIn, cabac_ctx_model *ctxOut,
                               uint64_t *estPerTu, int useMTS);
long orc_residual_ctx_bins(const cabac_tu_desc *tus, long ntu, const cabac_substream_params *p,
                           const int32_t *coeffs, const cabac_ctx_model *ctxIn, cabac_ctx_model *ctxOut,
                           uint32_t *ctxBins);
int orc_ctx_init_id(int qp, const cabac_ctx_model *ctxRef, const uint32_t *ctxBins, uint64_t *costs);
long orc_count_emulations(const uint8_t *d, long n);
long orc_insert_emulation_prevention(const uint8_t *in, long n, uint8_t *out);
long orc_remove_emulation_prevention(const uint8_t *nal, long n, int vcl, uint8_t *rbsp, uint32_t *epb, long *n_epb);
int orc_entry_point_syntax(const uint32_t *sizes, const uint32_t *emulations, long n_sub, uint32_t *offset_minus1);
void orc_substream_ranges(const uint32_t *offset_minus1, long n_entry, const uint32_t *epb, long n_epb,
                          uint64_t hdr_end, uint64_t rbsp_size, uint64_t *in_offsets);
/* stepwise decoder handle (for oracle/syntax_host.cpp) */
void *orc_dec_open(const uint8_t *in, long n, const cabac_substream_params *p, const cabac_ctx_model *ctxIn);
unsigned orc_dec_bin(void *h, unsigned ctxId);
unsigned orc_dec_ep(void *h, unsigned n);
unsigned orc_dec_trm(void *h);
void orc_dec_tu(void *h, const cabac_tu_desc *tu, int32_t *coeffs, uint8_t *tuFlags);
long orc_dec_bits(void *h);
void orc_dec_restart(void *h, const uint8_t *in, long n);
void orc_dec_ctx_get(void *h, cabac_ctx_model *out);
void orc_dec_ctx_set(void *h, const cabac_ctx_model *in);
long orc_dec_close(void *h, cabac_ctx_model *ctxOut, uint32_t *status, uint32_t *numBins);
/* stepwise encoder handle (oracle/syntax_host.cpp: generator of legal CTU-level streams) */
void *orc_enc_open(uint8_t *out, long cap, const cabac_substream_params *p, const cabac_ctx_model *ctxIn);
void orc_enc_bin(void *h, unsigned bin, unsigned ctxId);
void orc_enc_ep(void *h, unsigned bins, unsigned n);
void orc_enc_trm(void *h, unsigned bin);
unsigned orc_enc_tu(void *h, const cabac_tu_desc *tu, const int32_t *coeffs, uint8_t *tuFlags);
void orc_enc_ctx_get(void *h, cabac_ctx_model *out);
void orc_enc_ctx_set(void *h, const cabac_ctx_model *in);
long orc_enc_flush(void *h, int restart);
long orc_enc_close(void *h, cabac_ctx_model *ctxOut, uint32_t *numBins);
#endif
