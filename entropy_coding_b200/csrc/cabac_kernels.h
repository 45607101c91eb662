// cabac_kernels.h - launchers exported by cabac_kernels.cu to the C-ABI layer.
#pragma once
#include <cuda_runtime.h>

#include "cabac_b200.h"

namespace cabac {

// WPP hand-over inside one launch (cabac_wpp_sync, include/cabac_b200.h): per-substream sync records, the
// snapshot rows (CABAC_NUM_CTX models per substream) and the "published" flags.  sync == nullptr: off.
struct WppArgs {
  const cabac_wpp_sync *sync = nullptr;
  cabac_ctx_model *snap = nullptr;
  uint32_t *ready = nullptr;
  uint32_t spinLimit = 1u << 24;  // polls of >= 100 ns before a wait gives up with CABAC_SS_WPP_TIMEOUT
  uint32_t debug = 0;             // dev: bit 0 skips the waits, bit 1 the snapshots, bit 2 keeps the LPT order
};

// Device scratch of the scheduler (owned by the handle): per-substream cost, the
// longest-first order and the work cursor.  Kernels launched per call: 3.
struct Sched {
  uint32_t *cost;
  uint32_t *order;
  uint32_t *cursor;
  // residual encode (two-stage): command-slab bookkeeping
  uint32_t *capWords;       // [n]   upper bound of command words per substream
  uint32_t *cmdCounts;      // [n]   words stage P produced
  uint32_t *statusScratch;  // [n]   when the caller passes no status array
  uint64_t *cmdOffsets;     // [n+1] slab offsets (words); [n] = total
  uint32_t *cmdBuf;         // the slab
  bool warpSpecialised = false;  // fused kernel: producer warps + one walker warp per CTA (FUSED_WS)
  bool fusedResolve = false;  // fused kernel: resolve probabilities lane-parallel before the serial walk
  bool twoStage;            // residual encode: stage P + stage D instead of the fused lockstep kernel
  bool threadDecode = false;  // residual decode: one thread per substream instead of one warp (>= 10^5 substreams)
  uint32_t *threadCtx = nullptr;  // [n][230] packed models, the thread decoder's context rows
  bool coeff16 = false;     // the coefficient slab holds int16 levels (cabac_*_i16) instead of int32 TCoeff
  WppArgs wpp;              // WPP launches: substreams are claimed in index order instead of longest-first
};
constexpr int kLaunchesPerCodingCall = 3;

cudaError_t uploadRom();
const unsigned char *hostInitTable(int row);
const unsigned int *hostFracBits();

cudaError_t launchCtxInit(cudaStream_t s, int device, const cabac_substream_params *params, size_t n,
                          cabac_ctx_model *out);
cudaError_t launchDecodeBins(cudaStream_t s, int device, const Sched &sc, const uint8_t *in, const uint64_t *inOffsets,
                             const uint32_t *cmds, const uint64_t *cmdOffsets,
                             const cabac_substream_params *params, size_t nSub, const cabac_ctx_model *ctxIn,
                             uint32_t *values, uint32_t *bitsRead, cabac_ctx_model *ctxOut, uint32_t *status);
cudaError_t launchDecodeSlices(cudaStream_t s, int device, const Sched &sc, const uint8_t *in, const uint64_t *inOffsets,
                               const cabac_substream_params *params, const cabac_slice_params *sliceParams, size_t nSub,
                               const cabac_ctx_model *ctxIn, const uint64_t *cuOffsets, const uint64_t *tuOffsets,
                               const uint64_t *ctuOffsets, const uint64_t *coeffOffsets, cabac_cu_desc *cus,
                               cabac_tu_desc *tus, uint8_t *tuFlags, cabac_ctu_desc *ctus, int32_t *coeffs,
                               uint32_t *counts, uint32_t *numBins, uint32_t *bitsRead, cabac_ctx_model *ctxOut,
                               uint32_t *status);
cudaError_t launchEmitSlices(cudaStream_t s, int device, const Sched &sc, const cabac_substream_params *params,
                             const cabac_slice_params *sliceParams, size_t nSub, const cabac_cu_desc *cus,
                             const uint64_t *cuOffsets, const cabac_tu_desc *tus, const uint8_t *tuFlags,
                             const uint64_t *tuOffsets, const cabac_ctu_desc *ctus, const uint64_t *ctuOffsets,
                             const cabac_ctx_model *ctxIn, uint32_t *stage, uint64_t *stageOffsets, uint32_t *cmds,
                             uint64_t capWords, uint32_t *counts, uint32_t *emitStatus, uint64_t *cmdOffsets,
                             const uint64_t *rowOffsets, uint64_t *rowCmdOff, uint64_t *rowTuOff, cabac_wpp_sync *rowSync,
                             cabac_substream_params *rowParams, uint32_t *rowSlice, cabac_ctx_model *rowCtxIn);
uint64_t emitStageWords(uint64_t nCu, uint64_t nTu, uint64_t nCtu, uint64_t nSub);
cudaError_t launchOrStatus(cudaStream_t s, uint32_t *status, const uint32_t *extra, const uint32_t *rowSlice, size_t nRows);
cudaError_t launchEncodePrepare(cudaStream_t s, int device, const Sched &sc, const cabac_tu_desc *tus,
                                const uint64_t *tuOffsets, size_t nSub, const int32_t *coeffs,
                                const uint64_t *rawOffsets, const cabac_substream_params *params = nullptr,
                                int sampleMask = -1);
cudaError_t launchEncodeStream(cudaStream_t s, int device, const Sched &sc, const uint32_t *rawCmds,
                               const uint64_t *rawOffsets, const cabac_tu_desc *tus, const uint64_t *tuOffsets,
                               const cabac_substream_params *params, size_t nSub, const int32_t *coeffs,
                               const cabac_ctx_model *ctxIn, uint8_t *out, const uint64_t *outOffsets,
                               uint32_t *outSizes, uint32_t *numBins, uint8_t *tuFlags, cabac_ctx_model *ctxOut,
                               uint32_t *status);
cudaError_t launchEncodeResidual(cudaStream_t s, int device, const Sched &sc, const cabac_tu_desc *tus, const uint64_t *tuOffsets,
                                 const cabac_substream_params *params, size_t nSub, const int32_t *coeffs,
                                 const cabac_ctx_model *ctxIn, uint8_t *out, const uint64_t *outOffsets,
                                 uint32_t *outSizes, uint32_t *numBins, uint8_t *tuFlags,
                                 cabac_ctx_model *ctxOut, uint32_t *status);
// Scratch of the sliced encoder (few, long substreams; cabac_sliced.cuh), owned by the handle, grow-only.
struct SlicedWork {
  static constexpr int kBufs = 13;
  void *buf[kBufs] = {};
  size_t bytes[kBufs] = {};
  uint64_t *pinned = nullptr;
};
cudaError_t launchEncodeSliced(cudaStream_t s, int device, SlicedWork &W, const cabac_tu_desc *tus,
                               const uint64_t *tuOffsets, const cabac_substream_params *params, size_t nSub,
                               const int32_t *coeffs, bool coeff16, const cabac_ctx_model *ctxIn, uint8_t *out,
                               const uint64_t *outOffsets, uint32_t *outSizes, uint32_t *numBins, uint8_t *tuFlags,
                               cabac_ctx_model *ctxOut, uint32_t *status, size_t minTusPerSub, size_t minTus,
                               bool *used, int *launches);
void freeSlicedWork(SlicedWork &W);
cudaError_t launchEstimateResidual(cudaStream_t s, int device, const Sched &sc, const cabac_tu_desc *tus,
                                   const uint64_t *tuOffsets, const cabac_substream_params *params, size_t nSub,
                                   const int32_t *coeffs, const cabac_ctx_model *ctxIn, uint64_t *estBits,
                                   uint64_t *estPerTu, uint32_t *numBins, uint8_t *tuFlags, cabac_ctx_model *ctxOut,
                                   uint32_t *status, uint32_t *ctxBins = nullptr);
cudaError_t launchCtxInitId(cudaStream_t s, const cabac_substream_params *params, const cabac_ctx_model *ctxRef,
                            const uint32_t *ctxBins, size_t n, uint8_t *choice, uint64_t *costs);
cudaError_t launchDecodeResidual(cudaStream_t s, int device, const Sched &sc, const uint8_t *in, const uint64_t *inOffsets,
                                 const cabac_tu_desc *tus, const uint64_t *tuOffsets,
                                 const cabac_substream_params *params, size_t nSub, const cabac_ctx_model *ctxIn,
                                 int32_t *coeffs, uint32_t *numBins, uint32_t *bitsRead, uint8_t *tuFlags,
                                 cabac_ctx_model *ctxOut, uint32_t *status);
// Allocation-free variant (context rows in a file-scope device array; cabac_kernels.cu).
cudaError_t setupStaticScratch(int device);  // once per device, with that device current
cudaError_t launchDecodeResidualStatic(cudaStream_t s, int device, const uint8_t *in, const uint64_t *inOffsets,
                                       const cabac_tu_desc *tus, const uint64_t *tuOffsets,
                                       const cabac_substream_params *params, size_t nSub, const cabac_ctx_model *ctxIn,
                                       int32_t *coeffs, uint32_t *numBins, uint32_t *bitsRead, uint8_t *tuFlags,
                                       cabac_ctx_model *ctxOut, uint32_t *status);
// per-chunk coefficient ranges: chunk k = substreams [chunkStart[k], chunkStart[k+1]); device pointers
cudaError_t launchCoeffRanges(cudaStream_t s, const cabac_tu_desc *tus, const uint64_t *tuOffsets,
                              const uint32_t *chunkStart, int nChunks, uint64_t *ranges,
                              uint64_t *areaSums = nullptr);
cudaError_t launchEmulationCount(cudaStream_t s, const uint8_t *data, const uint64_t *offsets, size_t nSeg,
                                 uint32_t *counts);
size_t emulationTiles(size_t n);
cudaError_t launchEmulationInsert(cudaStream_t s, const uint8_t *data, size_t n, uint32_t *tileCounts,
                                  uint64_t *tileBase, uint8_t *out, size_t outCap, uint64_t *outSize);
cudaError_t launchEmulationRemove(cudaStream_t s, const uint8_t *data, size_t n, int stripTrailingZeros,
                                  uint32_t *tileCounts, uint64_t *tileBase, uint8_t *out, uint32_t *epbLocations,
                                  uint64_t *sizes, uint32_t *violations);
cudaError_t launchEntryPointSyntax(cudaStream_t s, const uint32_t *sizes, const uint32_t *emulations, size_t n,
                                   uint32_t *offsetMinus1, uint32_t *lenMinus1);
cudaError_t launchSubstreamRanges(cudaStream_t s, const uint32_t *offsetMinus1, size_t nEntry, const uint32_t *epb,
                                  size_t nEpb, uint64_t hdrEnd, uint64_t rbspSize, uint64_t *inOffsets);
cudaError_t launchPack(cudaStream_t s, int device, const uint8_t *out, const uint64_t *outOffsets,
                       const uint32_t *outSizes, size_t nSub, uint8_t *packed, uint64_t *entryPoints);

}  // namespace cabac
