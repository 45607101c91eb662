// cabac_api.cu - the C ABI of include/cabac_b200.h.
//
// `*_dev` entry points launch asynchronously on the handle's stream with
// device pointers.  The host-pointer entry points wrap them: pinned staging,
// H2D copies, launch, D2H copies, one synchronise.  There is no CPU fallback:
// every failure to reach the GPU is reported as CABAC_ERR_CUDA.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "cabac_b200.h"
#include "cabac_kernels.h"

struct cabac_handle {
  int device = 0;
  cudaStream_t ownStream = nullptr;
  cudaStream_t stream = nullptr;
  std::string err;
  uint64_t launches = 0;
  // grow-only device scratch for the host-pointer entry points
  void *scratch = nullptr;
  size_t scratchBytes = 0;
  // scheduler scratch (cost / order / cursor), grow-only
  uint32_t *sched = nullptr;
  size_t schedSubs = 0;
  // resolved command words of the two-stage residual encoder, grow-only
  uint32_t *cmdBuf = nullptr;
  size_t cmdCapWords = 0;
  uint64_t *pinnedTotal = nullptr;  // page-locked landing spot for the slab size read-back
  int encoderMode = CABAC_ENCODER_AUTO;
  int decoderMode = CABAC_DECODER_AUTO;
  uint32_t *threadCtx = nullptr;  // context rows of the thread-per-substream decoder, grow-only
  size_t threadCtxSubs = 0;
  bool zeroCopy = true;      // CABAC_B200_ZERO_COPY=0: always stage pinned host slabs through device memory
  bool areaCostNext = false; // next residual encode: schedule from areas, do not read the slab (zero-copy)
  bool fusedResolve = true;  // CABAC_B200_FUSED_RESOLVE=0 selects the classic single-site drain
  // chunked pipeline of the host-pointer residual entry points: PCIe copies of the coefficient
  // slab overlap coding (one copy stream; one compute stream per chunk, so a chunk's long substreams never
  // hold back the chunks behind it)
  static constexpr int kMaxChunks = 16;
  int pipeChunks = 8;
  size_t pipeMinCoeffs = size_t(8) << 20, pipeMinSubstreams = 512;  // below: one copy, one launch
  cudaStream_t copyStream = nullptr, compStream[kMaxChunks] = {};
  cudaEvent_t evCopy[kMaxChunks] = {}, evDone[kMaxChunks] = {}, evJoin = nullptr;
  void *epbScratch = nullptr;  // tile counts / bases of the emulation-prevention pass, grow-only
  size_t epbTiles = 0;
  void *emitScratch = nullptr;  // cabac_encode_slices_dev: per-slice word counts / statuses, row-level arrays; grow-only
  size_t emitBytes = 0;
  uint32_t *emitStage = nullptr;  // the walk's staging words (one bounded region per slice), grow-only
  size_t emitStageWords = 0;
  bool trace = false;  // CABAC_B200_TRACE=1: per-chunk copy / compute completion times on stderr
  cudaEvent_t evT0 = nullptr;
  uint64_t *pinnedRanges = nullptr;  // [2 * kMaxChunks] ranges + [kMaxChunks + 1] u32 chunk starts
  // WPP launches (cabac_*_wpp*): the sync records of the NEXT coding call (device pointer, consumed by it),
  // snapshot rows + flags (grow-only), and a device copy of host-side sync records
  bool coeff16Next = false;  // cabac_*_i16: the coefficient slab of the next residual call holds int16 levels
  const cabac_wpp_sync *wppNext = nullptr;
  void *wppScratch = nullptr;
  size_t wppSubs = 0;
  cabac_wpp_sync *wppDev = nullptr;
  size_t wppDevSubs = 0;
  cabac::SlicedWork sliced;  // scratch of the sliced encoder
  // cabac_pin_descriptors: device copies of a batch description the caller promised not to change
  const void *pinTus = nullptr, *pinTuOff = nullptr, *pinParams = nullptr;
  size_t pinSubs = 0, pinTuCount = 0;
  void *pinDev = nullptr;  // [tus | tu_offsets | params], grow-only
  size_t pinDevBytes = 0;
  cabac_tu_desc *pinDevTus = nullptr;
  uint64_t *pinDevTuOff = nullptr;
  cabac_substream_params *pinDevParams = nullptr;
  bool pinned(const void *t, const void *o, const void *p, size_t n) const {
    return pinDev && pinTuOff && t == pinTus && o == pinTuOff && p == pinParams && n == pinSubs;
  }
};

namespace {

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
  }
  ~DeviceGuard() {
    int cur = -1;
    cudaGetDevice(&cur);
    if (prev >= 0 && cur != prev) cudaSetDevice(prev);
  }
};

int fail(cabac_handle *h, cudaError_t e, const char *what) {
  if (h) h->err = std::string(what) + ": " + cudaGetErrorString(e);
  return CABAC_ERR_CUDA;
}
int badArg(cabac_handle *h, const char *what) {
  if (h) h->err = std::string("bad argument: ") + what;
  return CABAC_ERR_ARG;
}

// Bump allocator over the handle's scratch buffer (256-byte aligned pieces).
struct Arena {
  cabac_handle *h;
  size_t need = 0;
  std::vector<std::pair<void **, size_t>> slots;
  explicit Arena(cabac_handle *hh) : h(hh) {}
  void want(void **p, size_t bytes) {
    slots.emplace_back(p, need);
    need += (bytes + 255) & ~size_t(255);
  }
  cudaError_t commit() {
    if (need > h->scratchBytes) {
      if (h->scratch) cudaFree(h->scratch);
      h->scratch = nullptr;
      h->scratchBytes = 0;
      const size_t sz = need + (need >> 2) + (1 << 20);
      cudaError_t e = cudaMalloc(&h->scratch, sz);
      if (e != cudaSuccess) return e;
      h->scratchBytes = sz;
    }
    for (auto &s : slots) *s.first = static_cast<char *>(h->scratch) + s.second;
    return cudaSuccess;
  }
};

// Scheduler scratch for n substreams.  Stream-ordered use only, so growing it
// (free + malloc, both synchronising) is safe.
cudaError_t schedFor(cabac_handle *h, size_t n, cabac::Sched &sc) {
  if (n > h->schedSubs) {
    if (h->sched) cudaFree(h->sched);
    h->sched = nullptr;
    h->schedSubs = 0;
    const size_t cap = n + (n >> 2) + 1024;
    cudaError_t e = cudaMalloc(&h->sched, (7 * cap + 64 + 2) * sizeof(uint32_t));
    if (e != cudaSuccess) return e;
    h->schedSubs = cap;
  }
  const size_t cap = h->schedSubs;
  sc.cursor = h->sched;
  sc.cmdOffsets = reinterpret_cast<uint64_t *>(h->sched + 64);  // 2 * (cap + 1) words, 8-byte aligned
  uint32_t *p = h->sched + 64 + 2 * (cap + 1);
  sc.cost = p;
  sc.order = p + cap;
  sc.capWords = p + 2 * cap;
  sc.cmdCounts = p + 3 * cap;
  sc.statusScratch = p + 4 * cap;
  sc.cmdBuf = h->cmdBuf;
  sc.fusedResolve = h->fusedResolve;
  sc.warpSpecialised = h->encoderMode == CABAC_ENCODER_WARP_SPECIALISED;
  sc.coeff16 = h->coeff16Next;
  sc.wpp = cabac::WppArgs();
  if (h->wppNext) {  // a WPP launch: snapshot rows + "published" flags for n substreams, flags cleared on the stream
    const cabac_wpp_sync *sync = h->wppNext;
    h->wppNext = nullptr;
    if (n > h->wppSubs) {
      if (h->wppScratch) cudaFree(h->wppScratch);
      h->wppScratch = nullptr;
      h->wppSubs = 0;
      const size_t cap2 = n + (n >> 2) + 64;
      cudaError_t e = cudaMalloc(&h->wppScratch, cap2 * (CABAC_NUM_CTX * sizeof(cabac_ctx_model) + sizeof(uint32_t)));
      if (e != cudaSuccess) return e;
      h->wppSubs = cap2;
    }
    static const uint32_t spinLimit = getenv("CABAC_B200_WPP_SPIN") ? uint32_t(atoll(getenv("CABAC_B200_WPP_SPIN"))) : (1u << 24);
    sc.wpp.spinLimit = spinLimit;
    static const uint32_t dbg = getenv("CABAC_B200_WPP_DEBUG") ? uint32_t(atoi(getenv("CABAC_B200_WPP_DEBUG"))) : 0u;
    sc.wpp.debug = dbg;
    sc.wpp.sync = sync;
    sc.wpp.snap = static_cast<cabac_ctx_model *>(h->wppScratch);
    sc.wpp.ready = reinterpret_cast<uint32_t *>(sc.wpp.snap + h->wppSubs * CABAC_NUM_CTX);
    cudaError_t e = cudaMemsetAsync(sc.wpp.ready, 0, n * sizeof(uint32_t), h->stream);
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

// AUTO keeps the warp-lockstep decoder at every batch size: measured at 10^5 substreams the thread-per-substream
// kernel issues 4.8x fewer warp instructions but runs them 7.4 lanes wide at 15 % occupancy and ends up 1.2x slower
// (profiles/r2_decoder_modes.md). CABAC_DECODER_THREAD selects it explicitly.
cudaError_t threadDecodeFor(cabac_handle *h, size_t n, cabac::Sched &sc) {
  sc.threadDecode = !sc.wpp.sync && n > 0 && h->decoderMode == CABAC_DECODER_THREAD;
  if (!sc.threadDecode) return cudaSuccess;
  if (n > h->threadCtxSubs) {
    if (h->threadCtx) cudaFree(h->threadCtx);
    h->threadCtx = nullptr;
    h->threadCtxSubs = 0;
    const size_t cap = n + (n >> 3) + 1024;
    cudaError_t e = cudaMalloc(&h->threadCtx, cap * 230 * sizeof(uint32_t));
    if (e != cudaSuccess) return e;
    h->threadCtxSubs = cap;
  }
  sc.threadCtx = h->threadCtx;
  return cudaSuccess;
}

// Host-side sync records -> the handle's device copy (stream-ordered); marks the next coding call as WPP.
cudaError_t stageWpp(cabac_handle *h, const cabac_wpp_sync *wpp, size_t n) {
  if (n > h->wppDevSubs) {
    if (h->wppDev) cudaFree(h->wppDev);
    h->wppDev = nullptr;
    h->wppDevSubs = 0;
    const size_t cap = n + (n >> 2) + 64;
    cudaError_t e = cudaMalloc(&h->wppDev, cap * sizeof(cabac_wpp_sync));
    if (e != cudaSuccess) return e;
    h->wppDevSubs = cap;
  }
  cudaError_t e = cudaMemcpyAsync(h->wppDev, wpp, n * sizeof(cabac_wpp_sync), cudaMemcpyHostToDevice, h->stream);
  if (e != cudaSuccess) return e;
  h->wppNext = h->wppDev;
  return cudaSuccess;
}
// dep must precede the substream (claim order = index order: the deadlock-freedom argument of the kernels)
bool wppValid(const cabac_wpp_sync *wpp, size_t n) {
  for (size_t i = 0; i < n; i++)
    if (wpp[i].dep != CABAC_WPP_NONE && wpp[i].dep >= i) return false;
  return true;
}

// Reads the command-slab size the prepare kernels computed (8 bytes; the one synchronisation
// point of a two-stage encode call) and grows the slab if needed.
cudaError_t ensureCmdSlab(cabac_handle *h, cabac::Sched &sc, size_t n_sub) {
  cudaError_t e;
  if (!h->pinnedTotal && (e = cudaHostAlloc(&h->pinnedTotal, sizeof(uint64_t), cudaHostAllocDefault)) != cudaSuccess)
    return e;
  if ((e = cudaMemcpyAsync(h->pinnedTotal, sc.cmdOffsets + n_sub, sizeof(uint64_t), cudaMemcpyDeviceToHost,
                           h->stream)) != cudaSuccess)
    return e;
  if ((e = cudaStreamSynchronize(h->stream)) != cudaSuccess) return e;
  const size_t needWords = size_t(*h->pinnedTotal);
  if (needWords > h->cmdCapWords) {
    if (h->cmdBuf) cudaFree(h->cmdBuf);
    h->cmdBuf = nullptr;
    h->cmdCapWords = 0;
    const size_t words = needWords + (needWords >> 3) + 1024;
    if ((e = cudaMalloc(&h->cmdBuf, words * sizeof(uint32_t))) != cudaSuccess) return e;
    h->cmdCapWords = words;
  }
  sc.cmdBuf = h->cmdBuf;
  return cudaSuccess;
}

// Page-locked, device-mapped host memory (cudaHostAlloc / cudaHostRegister under UVA): kernels can
// read and write it directly over PCIe.  Returns the device alias, or nullptr for any other pointer.
void *mappedHostAlias(const void *p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return (a.type == cudaMemoryTypeHost && a.devicePointer) ? a.devicePointer : nullptr;
}
// Zero-copy pays when the slab is large against the per-TU PCIe round trip.
constexpr size_t kZeroCopyMinCoeffs = size_t(4) << 20;

// Streams / events / pinned landing area of the chunked pipeline (created on first use).
cudaError_t ensurePipe(cabac_handle *h) {
  if (h->copyStream) return cudaSuccess;
  cudaError_t e;
  if ((e = cudaStreamCreateWithFlags(&h->copyStream, cudaStreamNonBlocking)) != cudaSuccess) return e;
  for (int i = 0; i < cabac_handle::kMaxChunks; i++)
    if ((e = cudaStreamCreateWithFlags(&h->compStream[i], cudaStreamNonBlocking)) != cudaSuccess) return e;
  for (int i = 0; i < cabac_handle::kMaxChunks; i++) {
    const unsigned fl = h->trace ? cudaEventDefault : cudaEventDisableTiming;
    if ((e = cudaEventCreateWithFlags(&h->evCopy[i], fl)) != cudaSuccess) return e;
    if ((e = cudaEventCreateWithFlags(&h->evDone[i], fl)) != cudaSuccess) return e;
  }
  if (h->trace && (e = cudaEventCreate(&h->evT0)) != cudaSuccess) return e;
  if ((e = cudaEventCreateWithFlags(&h->evJoin, h->trace ? cudaEventDefault : cudaEventDisableTiming)) != cudaSuccess) return e;
  return cudaHostAlloc(&h->pinnedRanges, (3 * cabac_handle::kMaxChunks + 2) * sizeof(uint64_t), cudaHostAllocDefault);
}

void traceChunks(cabac_handle *h, const char *what, int K, bool copyFirst) {
  if (!h->trace) return;
  fprintf(stderr, "[cabac_b200 trace] %s: %d chunks, ms since pipeline start (%s):", what, K,
          copyFirst ? "H2D done / coded" : "decoded");
  for (int k = 0; k < K; k++) {
    float a = 0, b = 0;
    if (copyFirst) cudaEventElapsedTime(&a, h->evT0, h->evCopy[k]);
    cudaEventElapsedTime(&b, h->evT0, h->evDone[k]);
    if (copyFirst) fprintf(stderr, " %.1f/%.1f", a, b); else fprintf(stderr, " %.1f", b);
  }
  float j = 0;
  cudaEventElapsedTime(&j, h->evT0, h->evJoin);
  fprintf(stderr, " | joined %.1f\n", j);
}

// Splits n_sub substreams into <= want chunks of about equal TU count (tu_offsets is a host array).
int planChunks(const uint64_t *tu_offsets, size_t n_sub, int want, uint32_t *start) {
  const uint64_t nTu = tu_offsets[n_sub] - tu_offsets[0];
  int k = 0;
  start[0] = 0;
  for (int c = 1; c < want; c++) {
    const uint64_t target = tu_offsets[0] + nTu * uint64_t(c) / uint64_t(want);
    size_t lo = start[k], hi = n_sub;  // first substream whose TUs begin at or after the target
    while (lo < hi) {
      const size_t mid = (lo + hi) / 2;
      if (tu_offsets[mid] < target) lo = mid + 1; else hi = mid;
    }
    if (lo > start[k] && lo < n_sub) start[++k] = uint32_t(lo);
  }
  start[++k] = uint32_t(n_sub);
  return k;
}

cabac::Sched chunkSched(const cabac::Sched &base, size_t s0, int k) {
  cabac::Sched c = base;
  c.cost += s0;
  c.order += s0;
  c.capWords += s0;
  c.cmdCounts += s0;
  c.statusScratch += s0;
  c.cmdOffsets += s0 + size_t(k);
  c.cursor = base.cursor + k;
  c.twoStage = false;
  return c;
}

// AUTO takes the sliced encoder (parallel inside every substream) for batches of fewer substreams than
// fill the warp slots of the fused kernels, when they are long: measured crossover on 1080p frames cut
// into n substreams: ~2 500 (the fused kernels need one warp per substream, 2 960 resident).
constexpr size_t kSlicedMaxSubstreams = 2048, kSlicedMinTusPerSubstream = 256, kSlicedMinTus = 65536;
// From this many substreams on, stage D (32 substreams per warp) fills the machine.
constexpr size_t kTwoStageMinSubstreams = 32768;
// AUTO: the warp-specialised fused kernel from here up to the two-stage threshold (148 SMs x 19 producers x 2)
constexpr size_t kWarpSpecialisedMinSubstreams = 5624;

#define CK(call, what)                                   \
  do {                                                   \
    cudaError_t e__ = (call);                            \
    if (e__ != cudaSuccess) return fail(h, e__, what);   \
  } while (0)

int anySubstreamFailed(const uint32_t *status, size_t n) {
  if (!status) return CABAC_OK;
  for (size_t i = 0; i < n; i++)
    if (status[i]) return CABAC_ERR_SUBSTREAM;
  return CABAC_OK;
}

}  // namespace

extern "C" {

const char *cabac_version(void) { return "cabac_b200 0.1 (sm_100a)"; }

cabac_handle *cabac_create(int device) {
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return nullptr;
  if (device < 0) {
    if (cudaGetDevice(&device) != cudaSuccess) return nullptr;
  }
  if (device >= count) return nullptr;
  cabac_handle *h = new cabac_handle();
  h->device = device;
  DeviceGuard g(device);
  if (cudaStreamCreateWithFlags(&h->ownStream, cudaStreamNonBlocking) != cudaSuccess ||
      cabac::uploadRom() != cudaSuccess || cabac::setupStaticScratch(device) != cudaSuccess) {
    delete h;
    return nullptr;
  }
  h->err.reserve(256);  // error texts of the allocation-free entry point fit without a heap call
  h->stream = h->ownStream;
  if (const char *c = getenv("CABAC_B200_PIPE_CHUNKS")) {  // 1 disables the chunked copy/compute pipeline
    const int v = atoi(c);
    if (v >= 1 && v <= cabac_handle::kMaxChunks) h->pipeChunks = v;
  }
  if (const char *c = getenv("CABAC_B200_ZERO_COPY")) h->zeroCopy = atoi(c) != 0;
  if (const char *c = getenv("CABAC_B200_FUSED_RESOLVE")) h->fusedResolve = atoi(c) != 0;
  if (const char *c = getenv("CABAC_B200_TRACE")) h->trace = atoi(c) != 0;
  if (const char *c = getenv("CABAC_B200_PIPE_MIN_COEFFS")) h->pipeMinCoeffs = size_t(atoll(c));
  if (const char *c = getenv("CABAC_B200_PIPE_MIN_SUBSTREAMS")) h->pipeMinSubstreams = size_t(atoll(c));
  if (const char *m = getenv("CABAC_B200_ENCODER")) {
    if (!strcmp(m, "fused")) h->encoderMode = CABAC_ENCODER_FUSED;
    if (!strcmp(m, "two_stage")) h->encoderMode = CABAC_ENCODER_TWO_STAGE;
    if (!strcmp(m, "warp_specialised")) h->encoderMode = CABAC_ENCODER_WARP_SPECIALISED;
    if (!strcmp(m, "sliced")) h->encoderMode = CABAC_ENCODER_SLICED;
  }
  if (const char *m = getenv("CABAC_B200_DECODER")) {
    if (!strcmp(m, "lockstep")) h->decoderMode = CABAC_DECODER_LOCKSTEP;
    if (!strcmp(m, "thread")) h->decoderMode = CABAC_DECODER_THREAD;
  }
  return h;
}

void cabac_destroy(cabac_handle *h) {
  if (!h) return;
  DeviceGuard g(h->device);
  cudaStreamSynchronize(h->stream);
  if (h->scratch) cudaFree(h->scratch);
  if (h->sched) cudaFree(h->sched);
  if (h->cmdBuf) cudaFree(h->cmdBuf);
  if (h->pinnedTotal) cudaFreeHost(h->pinnedTotal);
  if (h->pinnedRanges) cudaFreeHost(h->pinnedRanges);
  if (h->epbScratch) cudaFree(h->epbScratch);
  if (h->emitScratch) cudaFree(h->emitScratch);
  if (h->emitStage) cudaFree(h->emitStage);
  if (h->pinDev) cudaFree(h->pinDev);
  cabac::freeSlicedWork(h->sliced);
  if (h->threadCtx) cudaFree(h->threadCtx);
  if (h->wppScratch) cudaFree(h->wppScratch);
  if (h->wppDev) cudaFree(h->wppDev);
  for (int i = 0; i < cabac_handle::kMaxChunks; i++) {
    if (h->evCopy[i]) cudaEventDestroy(h->evCopy[i]);
    if (h->evDone[i]) cudaEventDestroy(h->evDone[i]);
  }
  if (h->evJoin) cudaEventDestroy(h->evJoin);
  if (h->copyStream) cudaStreamDestroy(h->copyStream);
  for (int i = 0; i < cabac_handle::kMaxChunks; i++)
    if (h->compStream[i]) cudaStreamDestroy(h->compStream[i]);
  if (h->ownStream) cudaStreamDestroy(h->ownStream);
  delete h;
}

int cabac_set_stream(cabac_handle *h, void *cuda_stream) {
  if (!h) return CABAC_ERR_ARG;
  h->stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : h->ownStream;
  return CABAC_OK;
}

int cabac_set_encoder_mode(cabac_handle *h, int mode) {
  if (!h || mode < CABAC_ENCODER_AUTO || mode > CABAC_ENCODER_SLICED) return CABAC_ERR_ARG;
  h->encoderMode = mode;
  return CABAC_OK;
}

int cabac_set_decoder_mode(cabac_handle *h, int mode) {
  if (!h || mode < CABAC_DECODER_AUTO || mode > CABAC_DECODER_THREAD) return CABAC_ERR_ARG;
  h->decoderMode = mode;
  return CABAC_OK;
}

int cabac_sync(cabac_handle *h) {
  if (!h) return CABAC_ERR_ARG;
  DeviceGuard g(h->device);
  CK(cudaStreamSynchronize(h->stream), "cabac_sync");
  return CABAC_OK;
}

const char *cabac_last_error(cabac_handle *h) { return h ? h->err.c_str() : "null handle"; }
uint64_t cabac_launch_count(cabac_handle *h) { return h ? h->launches : 0; }

void *cabac_host_alloc(size_t bytes) {
  void *p = nullptr;
  if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) return nullptr;
  return p;
}
void cabac_host_free(void *p) {
  if (p) cudaFreeHost(p);
}

int cabac_ctx_rates(uint8_t *rates_out) {
  if (!rates_out) return CABAC_ERR_ARG;
  const unsigned char *win = cabac::hostInitTable(3);
  for (int k = 0; k < CABAC_NUM_CTX; k++) {
    const int r0 = 2 + ((win[k] >> 2) & 3), r1 = 3 + r0 + (win[k] & 3);
    rates_out[k] = uint8_t(16 * r0 + r1);
  }
  return CABAC_OK;
}

int cabac_ctx_init_table(int row, uint8_t *table_out) {
  if (row < 0 || row > 3 || !table_out) return CABAC_ERR_ARG;
  memcpy(table_out, cabac::hostInitTable(row), CABAC_NUM_CTX);
  return CABAC_OK;
}

int cabac_frac_bits_table(uint32_t *table_out) {
  if (!table_out) return CABAC_ERR_ARG;
  memcpy(table_out, cabac::hostFracBits(), 512 * sizeof(uint32_t));
  return CABAC_OK;
}

// ---- device-pointer entry points -----------------------------------------
int cabac_ctx_init_dev(cabac_handle *h, const cabac_substream_params *params, size_t n, cabac_ctx_model *out) {
  if (!h || (n && (!params || !out))) return badArg(h, "cabac_ctx_init_dev");
  DeviceGuard g(h->device);
  CK(cabac::launchCtxInit(h->stream, h->device, params, n, out), "ctx_init_kernel");
  h->launches += n ? 1 : 0;
  return CABAC_OK;
}

int cabac_encode_stream_dev(cabac_handle *h, const uint32_t *cmds, const uint64_t *cmd_offsets,
                            const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                            const cabac_substream_params *params, size_t n_sub, const int32_t *coeffs,
                            const cabac_ctx_model *ctx_in, uint8_t *out, const uint64_t *out_offsets,
                            uint32_t *out_sizes, uint32_t *num_bins, uint8_t *tu_flags, cabac_ctx_model *ctx_out,
                            uint32_t *status);

// A recorded bin stream is a command stream without TU markers.
int cabac_encode_bins_dev(cabac_handle *h, const uint32_t *cmds, const uint64_t *cmd_offsets,
                          const cabac_substream_params *params, size_t n_sub, const cabac_ctx_model *ctx_in,
                          uint8_t *out, const uint64_t *out_offsets, uint32_t *out_sizes, uint32_t *num_bins,
                          cabac_ctx_model *ctx_out, uint32_t *status) {
  return cabac_encode_stream_dev(h, cmds, cmd_offsets, nullptr, nullptr, params, n_sub, nullptr, ctx_in, out,
                                 out_offsets, out_sizes, num_bins, nullptr, ctx_out, status);
}

int cabac_decode_bins_dev(cabac_handle *h, const uint8_t *in, const uint64_t *in_offsets, const uint32_t *cmds,
                          const uint64_t *cmd_offsets, const cabac_substream_params *params, size_t n_sub,
                          const cabac_ctx_model *ctx_in, uint32_t *values, uint32_t *bits_read,
                          cabac_ctx_model *ctx_out, uint32_t *status) {
  if (!h || (n_sub && (!in || !in_offsets || !cmd_offsets || !params))) return badArg(h, "cabac_decode_bins_dev");
  DeviceGuard g(h->device);
  cabac::Sched sc;
  CK(schedFor(h, n_sub, sc), "scheduler scratch");
  CK(cabac::launchDecodeBins(h->stream, h->device, sc, in, in_offsets, cmds, cmd_offsets, params, n_sub, ctx_in,
                             values, bits_read, ctx_out, status),
     "decode_bins_kernel");
  h->launches += n_sub ? cabac::kLaunchesPerCodingCall : 0;
  return CABAC_OK;
}

int cabac_encode_residual_dev(cabac_handle *h, const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                              const cabac_substream_params *params, size_t n_sub, const int32_t *coeffs,
                              const cabac_ctx_model *ctx_in, uint8_t *out, const uint64_t *out_offsets,
                              uint32_t *out_sizes, uint32_t *num_bins, uint8_t *tu_flags,
                              cabac_ctx_model *ctx_out, uint32_t *status) {
  if (!h || (n_sub && (!tu_offsets || !params || !out || !out_offsets || !out_sizes)))
    return badArg(h, "cabac_encode_residual_dev");
  DeviceGuard g(h->device);
  if (!n_sub) return CABAC_OK;
  cabac::Sched sc;
  CK(schedFor(h, n_sub, sc), "scheduler scratch");
  // Few, long substreams (slices, frames): the sliced encoder, parallel inside every substream.  It needs the
  // slab on the device (several passes), terminated substreams and no WPP hand-over; otherwise it declines.
  if (!sc.wpp.sync && !h->areaCostNext &&
      (h->encoderMode == CABAC_ENCODER_SLICED || (h->encoderMode == CABAC_ENCODER_AUTO && n_sub < kSlicedMaxSubstreams))) {
    bool used = false;
    int launches = 0;
    const bool forced = h->encoderMode == CABAC_ENCODER_SLICED;
    CK(cabac::launchEncodeSliced(h->stream, h->device, h->sliced, tus, tu_offsets, params, n_sub, coeffs, sc.coeff16, ctx_in,
                                 out, out_offsets, out_sizes, num_bins, tu_flags, ctx_out, status ? status : sc.statusScratch,
                                 forced ? 0 : kSlicedMinTusPerSubstream, forced ? 0 : kSlicedMinTus, &used, &launches),
       "sliced encoder");
    h->launches += uint64_t(launches);
    if (used) return CABAC_OK;
  }
  static const bool forceAreaCost = getenv("CABAC_B200_AREA_COST") && atoi(getenv("CABAC_B200_AREA_COST"));  // experiment
  const bool areaCost = h->areaCostNext || forceAreaCost;
  h->areaCostNext = false;
  // Scheduling cost: the area / QP estimate of every block, calibrated on every 32nd block's coefficients
  // (3 % of the slab: cheap in HBM, affordable over PCIe for a zero-copy slab).  CABAC_B200_COST_SAMPLE_MASK=-1
  // restores the full pass over the slab (HBM) / the uncalibrated estimate (zero-copy).
  static const int sampleMask = getenv("CABAC_B200_COST_SAMPLE_MASK") ? atoi(getenv("CABAC_B200_COST_SAMPLE_MASK")) : 31;
  CK(cabac::launchEncodePrepare(h->stream, h->device, sc, tus, tu_offsets, n_sub,
                                areaCost && sampleMask < 0 ? nullptr : coeffs, nullptr, params, sampleMask),
     "encode prepare kernels");
  // Few long substreams: one warp per substream, fused.  Many substreams: the two-stage coder,
  // whose interval stage runs one substream per thread.
  sc.twoStage = !sc.wpp.sync &&  // WPP: the fused kernels (contexts are final at every TU boundary there)
                (h->encoderMode == CABAC_ENCODER_TWO_STAGE ||
                 (h->encoderMode == CABAC_ENCODER_AUTO && n_sub >= kTwoStageMinSubstreams));
  if (sc.twoStage) CK(ensureCmdSlab(h, sc, n_sub), "command slab");
  sc.cmdBuf = h->cmdBuf;
  // In between (every producer slot of the GPU filled at least twice): the warp-specialised kernel.
  sc.warpSpecialised = h->encoderMode == CABAC_ENCODER_WARP_SPECIALISED ||
                       (h->encoderMode == CABAC_ENCODER_AUTO && !sc.twoStage && n_sub >= kWarpSpecialisedMinSubstreams);
  CK(cabac::launchEncodeResidual(h->stream, h->device, sc, tus, tu_offsets, params, n_sub, coeffs, ctx_in, out, out_offsets,
                                 out_sizes, num_bins, tu_flags, ctx_out, status),
     "encode_residual kernels");
  h->launches += sc.twoStage ? 5 : 4;  // cost+capacity, order, slab scan, then fused | stage P + stage D
  return CABAC_OK;
}

int cabac_encode_stream_dev(cabac_handle *h, const uint32_t *cmds, const uint64_t *cmd_offsets,
                            const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                            const cabac_substream_params *params, size_t n_sub, const int32_t *coeffs,
                            const cabac_ctx_model *ctx_in, uint8_t *out, const uint64_t *out_offsets,
                            uint32_t *out_sizes, uint32_t *num_bins, uint8_t *tu_flags, cabac_ctx_model *ctx_out,
                            uint32_t *status) {
  if (!h || (n_sub && (!cmd_offsets || !params || !out || !out_offsets || !out_sizes)) || (tus && !tu_offsets))
    return badArg(h, "cabac_encode_stream_dev");
  DeviceGuard g(h->device);
  if (!n_sub) return CABAC_OK;
  cabac::Sched sc;
  CK(schedFor(h, n_sub, sc), "scheduler scratch");
  // Few long substreams (slices, frames): one warp each, fused.  Many: stage P + stage D (one thread each).
  sc.twoStage = h->encoderMode == CABAC_ENCODER_TWO_STAGE ||
                (h->encoderMode == CABAC_ENCODER_AUTO && n_sub >= kTwoStageMinSubstreams);
  CK(cabac::launchEncodePrepare(h->stream, h->device, sc, tus, tu_offsets, n_sub, coeffs, cmd_offsets),
     "encode prepare kernels");
  if (sc.twoStage) CK(ensureCmdSlab(h, sc, n_sub), "command slab");
  CK(cabac::launchEncodeStream(h->stream, h->device, sc, cmds, cmd_offsets, tus, tu_offsets, params, n_sub, coeffs,
                               ctx_in, out, out_offsets, out_sizes, num_bins, tu_flags, ctx_out, status),
     "encode_stream kernels");
  h->launches += sc.twoStage ? 5 : 4;
  return CABAC_OK;
}

static int estimateDev(cabac_handle *h, const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                       const cabac_substream_params *params, size_t n_sub, const int32_t *coeffs,
                       const cabac_ctx_model *ctx_in, uint64_t *est_bits, uint64_t *est_per_tu, uint32_t *num_bins,
                       uint8_t *tu_flags, cabac_ctx_model *ctx_out, uint32_t *status, uint32_t *ctx_bins) {
  DeviceGuard g(h->device);
  if (!n_sub) return CABAC_OK;
  cabac::Sched sc;
  CK(schedFor(h, n_sub, sc), "scheduler scratch");
  sc.twoStage = false;
  CK(cabac::launchEncodePrepare(h->stream, h->device, sc, tus, tu_offsets, n_sub, coeffs, nullptr, params),
     "estimate prepare kernels");
  CK(cabac::launchEstimateResidual(h->stream, h->device, sc, tus, tu_offsets, params, n_sub, coeffs, ctx_in, est_bits,
                                   est_per_tu, num_bins, tu_flags, ctx_out, status, ctx_bins),
     "estimate_residual kernel");
  h->launches += 4;  // cost, order, slab scan (shared prepare), estimate
  return CABAC_OK;
}

int cabac_estimate_residual_bits_dev(cabac_handle *h, const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                                     const cabac_substream_params *params, size_t n_sub, const int32_t *coeffs,
                                     const cabac_ctx_model *ctx_in, uint64_t *est_bits, uint64_t *est_per_tu,
                                     uint32_t *num_bins, uint8_t *tu_flags, cabac_ctx_model *ctx_out,
                                     uint32_t *status) {
  if (!h || (n_sub && (!tu_offsets || !params || !est_bits))) return badArg(h, "cabac_estimate_residual_bits_dev");
  return estimateDev(h, tus, tu_offsets, params, n_sub, coeffs, ctx_in, est_bits, est_per_tu, num_bins, tu_flags,
                     ctx_out, status, nullptr);
}

int cabac_residual_ctx_bins_dev(cabac_handle *h, const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                                const cabac_substream_params *params, size_t n_sub, const int32_t *coeffs,
                                const cabac_ctx_model *ctx_in, uint32_t *ctx_bins, uint32_t *num_bins,
                                cabac_ctx_model *ctx_out, uint32_t *status) {
  if (!h || (n_sub && (!tu_offsets || !params || !ctx_bins))) return badArg(h, "cabac_residual_ctx_bins_dev");
  return estimateDev(h, tus, tu_offsets, params, n_sub, coeffs, ctx_in, nullptr, nullptr, num_bins, nullptr, ctx_out,
                     status, ctx_bins);
}

int cabac_ctx_init_id_dev(cabac_handle *h, const cabac_substream_params *params, const cabac_ctx_model *ctx_ref,
                          const uint32_t *ctx_bins, size_t n, uint8_t *choice, uint64_t *costs) {
  if (!h || (n && (!params || !ctx_ref || !ctx_bins || !choice))) return badArg(h, "cabac_ctx_init_id_dev");
  DeviceGuard g(h->device);
  CK(cabac::launchCtxInitId(h->stream, params, ctx_ref, ctx_bins, n, choice, costs), "ctx_init_id_kernel");
  h->launches += n ? 1 : 0;
  return CABAC_OK;
}

int cabac_decode_residual_dev(cabac_handle *h, const uint8_t *in, const uint64_t *in_offsets,
                              const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                              const cabac_substream_params *params, size_t n_sub, const cabac_ctx_model *ctx_in,
                              int32_t *coeffs, uint32_t *num_bins, uint32_t *bits_read, uint8_t *tu_flags,
                              cabac_ctx_model *ctx_out, uint32_t *status) {
  if (!h || (n_sub && (!in || !in_offsets || !tu_offsets || !params || !coeffs)))
    return badArg(h, "cabac_decode_residual_dev");
  DeviceGuard g(h->device);
  cabac::Sched sc;
  CK(schedFor(h, n_sub, sc), "scheduler scratch");
  CK(threadDecodeFor(h, n_sub, sc), "thread decoder scratch");
  CK(cabac::launchDecodeResidual(h->stream, h->device, sc, in, in_offsets, tus, tu_offsets, params, n_sub, ctx_in,
                                 coeffs, num_bins, bits_read, tu_flags, ctx_out, status),
     "decode_residual_kernel");
  h->launches += n_sub ? cabac::kLaunchesPerCodingCall + (sc.threadDecode ? 1 : 0) : 0;
  return CABAC_OK;
}

// Writer direction of the slice syntax (cabac_slice.cuh K10 + the stream encoder).  Everything below the walk is
// row-level: one row substream per slice, or one per CTU row of a WPP slice (row_offsets given) with the hand-over
// resolved by the stream encoder's wavefront mode.  out_offsets / out_sizes / num_bins / ctx_out / status are per ROW.
static int encodeSliceRows(cabac_handle *h, const char *what, const cabac_substream_params *params,
                           const cabac_slice_params *slice_params, size_t n_sub, const cabac_cu_desc *cus,
                           const uint64_t *cu_offsets, const cabac_tu_desc *tus, const uint8_t *tu_flags,
                           const uint64_t *tu_offsets, const cabac_ctu_desc *ctus, const uint64_t *ctu_offsets,
                           const int32_t *coeffs, const cabac_ctx_model *ctx_in, uint32_t *cmd_scratch,
                           size_t cmd_scratch_words, const uint64_t *row_offsets, size_t n_rows, uint8_t *out,
                           const uint64_t *out_offsets, uint32_t *out_sizes, uint32_t *num_bins, cabac_ctx_model *ctx_out,
                           uint32_t *status) {
  if (!h || (n_sub && (!params || !slice_params || !cus || !cu_offsets || !tus || !tu_flags || !tu_offsets || !ctus ||
                       !ctu_offsets || !coeffs || !cmd_scratch || !out || !out_offsets || !out_sizes || !status)) ||
      n_rows < n_sub)
    return badArg(h, what);
  if (!n_sub) return CABAC_OK;
  DeviceGuard g(h->device);
  // staging for the walk: one region per slice, bounded by its record counts (the totals come back first)
  uint64_t tot[3] = {0, 0, 0};
  CK(cudaMemcpyAsync(&tot[0], cu_offsets + n_sub, 8, cudaMemcpyDeviceToHost, h->stream), "copy");
  CK(cudaMemcpyAsync(&tot[1], tu_offsets + n_sub, 8, cudaMemcpyDeviceToHost, h->stream), "copy");
  CK(cudaMemcpyAsync(&tot[2], ctu_offsets + n_sub, 8, cudaMemcpyDeviceToHost, h->stream), "copy");
  CK(cudaStreamSynchronize(h->stream), "sync");
  const uint64_t stageWords = cabac::emitStageWords(tot[0], tot[1], tot[2], n_sub);
  if (stageWords > h->emitStageWords) {
    if (h->emitStage) cudaFree(h->emitStage);
    h->emitStage = nullptr;
    h->emitStageWords = 0;
    CK(cudaMalloc(&h->emitStage, (stageWords + (stageWords >> 3)) * 4), "slice emit staging");
    h->emitStageWords = stageWords + (stageWords >> 3);
  }
  // scratch: u64 cmdOffsets[nSub+1], stageOffsets[nSub+1], rowCmdOff[nRows+1], rowTuOff[nRows+1] | sync[nRows] (16 B) |
  //          u32 counts[nSub], emitStatus[nSub], rowSlice[nRows], rowParams[nRows] | ctx rows[nRows][379] (if ctx_in)
  const size_t need = (2 * (n_sub + 1) + 2 * (n_rows + 1)) * 8 + n_rows * sizeof(cabac_wpp_sync) + (2 * n_sub + 2 * n_rows) * 4 +
                      (ctx_in ? n_rows * CABAC_NUM_CTX * sizeof(cabac_ctx_model) : 0) + 64;
  if (need > h->emitBytes) {
    CK(cudaStreamSynchronize(h->stream), "sync");
    if (h->emitScratch) cudaFree(h->emitScratch);
    h->emitScratch = nullptr;
    h->emitBytes = 0;
    CK(cudaMalloc(&h->emitScratch, need + (need >> 2)), "slice emit scratch");
    h->emitBytes = need + (need >> 2);
  }
  static_assert(sizeof(cabac_substream_params) == 4, "row parameter copies are packed as words");
  uint64_t *cmdOffsets = static_cast<uint64_t *>(h->emitScratch);
  uint64_t *stageOffsets = cmdOffsets + n_sub + 1;
  uint64_t *rowCmdOff = stageOffsets + n_sub + 1, *rowTuOff = rowCmdOff + n_rows + 1;
  cabac_wpp_sync *rowSync = reinterpret_cast<cabac_wpp_sync *>(rowTuOff + n_rows + 1);
  uint32_t *counts = reinterpret_cast<uint32_t *>(rowSync + n_rows), *emitStatus = counts + n_sub;
  uint32_t *rowSlice = emitStatus + n_sub;
  cabac_substream_params *rowParams = reinterpret_cast<cabac_substream_params *>(rowSlice + n_rows);
  cabac_ctx_model *rowCtxIn = ctx_in ? reinterpret_cast<cabac_ctx_model *>(rowParams + n_rows) : nullptr;
  cabac::Sched sc;
  CK(schedFor(h, n_sub, sc), "scheduler scratch");
  CK(cabac::launchEmitSlices(h->stream, h->device, sc, params, slice_params, n_sub, cus, cu_offsets, tus, tu_flags, tu_offsets,
                             ctus, ctu_offsets, ctx_in, h->emitStage, stageOffsets, cmd_scratch, cmd_scratch_words, counts,
                             emitStatus, cmdOffsets,
                             row_offsets, rowCmdOff, rowTuOff, rowSync, rowParams, rowSlice, rowCtxIn),
     "emit_slice kernels");
  h->launches += 8;
  // a slice whose commands did not fit (CABAC_SS_CAPACITY) has offsets past the scratch: the encoder must not read
  // them, so such a call stops here with every slice's status written
  CK(cudaMemsetAsync(out_sizes, 0, n_rows * 4, h->stream), "memset");
  uint64_t total = 0;
  CK(cudaMemcpyAsync(&total, cmdOffsets + n_sub, 8, cudaMemcpyDeviceToHost, h->stream), "copy");
  CK(cudaStreamSynchronize(h->stream), "sync");
  if (total > cmd_scratch_words) {
    CK(cudaMemsetAsync(status, 0, n_rows * 4, h->stream), "memset");
    CK(cabac::launchOrStatus(h->stream, status, emitStatus, rowSlice, n_rows), "or_status_kernel");
    h->err = std::string(what) + ": cmd_scratch too small (status[] says which slices: CABAC_SS_CAPACITY)";
    return CABAC_ERR_SUBSTREAM;
  }
  const int rc = row_offsets
                     ? cabac_encode_stream_wpp_dev(h, cmd_scratch, rowCmdOff, tus, rowTuOff, rowParams, rowSync, n_rows, coeffs,
                                                   rowCtxIn, out, out_offsets, out_sizes, num_bins, nullptr, ctx_out, status)
                     : cabac_encode_stream_dev(h, cmd_scratch, rowCmdOff, tus, rowTuOff, rowParams, n_rows, coeffs, rowCtxIn,
                                               out, out_offsets, out_sizes, num_bins, nullptr, ctx_out, status);
  if (rc) return rc;
  CK(cabac::launchOrStatus(h->stream, status, emitStatus, rowSlice, n_rows), "or_status_kernel");
  h->launches += 1;
  return CABAC_OK;
}

int cabac_encode_slices_dev(cabac_handle *h, const cabac_substream_params *params, const cabac_slice_params *slice_params,
                            size_t n_sub, const cabac_cu_desc *cus, const uint64_t *cu_offsets, const cabac_tu_desc *tus,
                            const uint8_t *tu_flags, const uint64_t *tu_offsets, const cabac_ctu_desc *ctus,
                            const uint64_t *ctu_offsets, const int32_t *coeffs, const cabac_ctx_model *ctx_in,
                            uint32_t *cmd_scratch, size_t cmd_scratch_words, uint8_t *out, const uint64_t *out_offsets,
                            uint32_t *out_sizes, uint32_t *num_bins, cabac_ctx_model *ctx_out, uint32_t *status) {
  return encodeSliceRows(h, "cabac_encode_slices_dev", params, slice_params, n_sub, cus, cu_offsets, tus, tu_flags, tu_offsets,
                         ctus, ctu_offsets, coeffs, ctx_in, cmd_scratch, cmd_scratch_words, nullptr, n_sub, out, out_offsets,
                         out_sizes, num_bins, ctx_out, status);
}

int cabac_encode_slices_wpp_dev(cabac_handle *h, const cabac_substream_params *params,
                                const cabac_slice_params *slice_params, size_t n_sub, const cabac_cu_desc *cus,
                                const uint64_t *cu_offsets, const cabac_tu_desc *tus, const uint8_t *tu_flags,
                                const uint64_t *tu_offsets, const cabac_ctu_desc *ctus, const uint64_t *ctu_offsets,
                                const int32_t *coeffs, const cabac_ctx_model *ctx_in, uint32_t *cmd_scratch,
                                size_t cmd_scratch_words, const uint64_t *row_offsets, size_t n_rows, uint8_t *out,
                                const uint64_t *out_offsets, uint32_t *out_sizes, uint32_t *num_bins,
                                cabac_ctx_model *ctx_out, uint32_t *status) {
  if (n_sub && !row_offsets) return badArg(h, "cabac_encode_slices_wpp_dev: row_offsets");
  return encodeSliceRows(h, "cabac_encode_slices_wpp_dev", params, slice_params, n_sub, cus, cu_offsets, tus, tu_flags,
                         tu_offsets, ctus, ctu_offsets, coeffs, ctx_in, cmd_scratch, cmd_scratch_words, row_offsets, n_rows,
                         out, out_offsets, out_sizes, num_bins, ctx_out, status);
}

int cabac_decode_slices_dev(cabac_handle *h, const uint8_t *in, const uint64_t *in_offsets,
                            const cabac_substream_params *params, const cabac_slice_params *slice_params, size_t n_sub,
                            const cabac_ctx_model *ctx_in, const uint64_t *cu_offsets, const uint64_t *tu_offsets,
                            const uint64_t *ctu_offsets, const uint64_t *coeff_offsets, cabac_cu_desc *cus,
                            cabac_tu_desc *tus, uint8_t *tu_flags, cabac_ctu_desc *ctus, int32_t *coeffs,
                            uint32_t *counts, uint32_t *num_bins, uint32_t *bits_read, cabac_ctx_model *ctx_out,
                            uint32_t *status) {
  if (!h || (n_sub && (!in || !in_offsets || !params || !slice_params || !cu_offsets || !tu_offsets || !ctu_offsets ||
                       !coeff_offsets || !tus || !coeffs)))
    return badArg(h, "cabac_decode_slices_dev");
  DeviceGuard g(h->device);
  cabac::Sched sc;
  CK(schedFor(h, n_sub, sc), "scheduler scratch");
  CK(cabac::launchDecodeSlices(h->stream, h->device, sc, in, in_offsets, params, slice_params, n_sub, ctx_in, cu_offsets,
                               tu_offsets, ctu_offsets, coeff_offsets, cus, tus, tu_flags, ctus, coeffs, counts, num_bins,
                               bits_read, ctx_out, status),
     "decode_slice_kernel");
  h->launches += n_sub ? cabac::kLaunchesPerCodingCall : 0;
  return CABAC_OK;
}

int cabac_decode_residual_static_dev(cabac_handle *h, const uint8_t *in, const uint64_t *in_offsets,
                                     const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                                     const cabac_substream_params *params, size_t n_sub, const cabac_ctx_model *ctx_in,
                                     int32_t *coeffs, uint32_t *num_bins, uint32_t *bits_read, uint8_t *tu_flags,
                                     cabac_ctx_model *ctx_out, uint32_t *status) {
  if (!h) return CABAC_ERR_ARG;
  if (n_sub && (!in || !in_offsets || !tu_offsets || !params || !coeffs)) {
    h->err.assign("bad argument: cabac_decode_residual_static_dev");
    return CABAC_ERR_ARG;
  }
  if (n_sub > size_t(CABAC_STATIC_MAX_SUBSTREAMS)) {
    char msg[160];
    snprintf(msg, sizeof(msg),
             "cabac_decode_residual_static_dev: n_sub %zu exceeds CABAC_STATIC_MAX_SUBSTREAMS (%d); split the batch or "
             "use cabac_decode_residual_dev",
             n_sub, CABAC_STATIC_MAX_SUBSTREAMS);
    h->err.assign(msg);
    return CABAC_ERR_ARG;
  }
  DeviceGuard g(h->device);
  cudaError_t e = cabac::launchDecodeResidualStatic(h->stream, h->device, in, in_offsets, tus, tu_offsets, params, n_sub,
                                                    ctx_in, coeffs, num_bins, bits_read, tu_flags, ctx_out, status);
  if (e != cudaSuccess) {
    h->err.assign("decode_residual_thread_kernel (static scratch): ");
    h->err.append(cudaGetErrorString(e));
    return CABAC_ERR_CUDA;
  }
  h->launches += n_sub ? 2 : 0;
  return CABAC_OK;
}

int cabac_pack_dev(cabac_handle *h, const uint8_t *out, const uint64_t *out_offsets, const uint32_t *out_sizes,
                   size_t n_sub, uint8_t *packed, uint64_t *entry_points) {
  if (!h || !entry_points || (n_sub && (!out || !out_offsets || !out_sizes || !packed)))
    return badArg(h, "cabac_pack_dev");
  DeviceGuard g(h->device);
  CK(cabac::launchPack(h->stream, h->device, out, out_offsets, out_sizes, n_sub, packed, entry_points),
     "pack kernels");
  h->launches += n_sub ? 2 : 1;
  return CABAC_OK;
}

int cabac_count_emulations_dev(cabac_handle *h, const uint8_t *data, const uint64_t *offsets, size_t n_seg,
                               uint32_t *counts) {
  if (!h || (n_seg && (!data || !offsets || !counts))) return badArg(h, "cabac_count_emulations_dev");
  DeviceGuard g(h->device);
  CK(cabac::launchEmulationCount(h->stream, data, offsets, n_seg, counts), "emulation_count_kernel");
  h->launches += n_seg ? 1 : 0;
  return CABAC_OK;
}

int cabac_insert_emulation_prevention_dev(cabac_handle *h, const uint8_t *rbsp, size_t n, uint8_t *out,
                                          size_t out_cap, uint64_t *out_size) {
  if (!h || !out_size || (n && (!rbsp || !out))) return badArg(h, "cabac_insert_emulation_prevention_dev");
  DeviceGuard g(h->device);
  const size_t tiles = cabac::emulationTiles(n);
  if (tiles > h->epbTiles) {
    if (h->epbScratch) cudaFree(h->epbScratch);
    h->epbScratch = nullptr;
    h->epbTiles = 0;
    const size_t cap = tiles + (tiles >> 2) + 256;
    CK(cudaMalloc(&h->epbScratch, cap * 4 + (cap + 1) * 8 + 16), "emulation scratch");
    h->epbTiles = cap;
  }
  uint64_t *tileBase = static_cast<uint64_t *>(h->epbScratch);
  uint32_t *tileCounts = reinterpret_cast<uint32_t *>(tileBase + h->epbTiles + 1);
  CK(cabac::launchEmulationInsert(h->stream, rbsp, n, tileCounts, tileBase, out, out_cap, out_size),
     "emulation insert kernels");
  h->launches += n ? 3 : 0;
  return CABAC_OK;
}

// ---- host-pointer entry points -------------------------------------------
#define H2D(dst, src, bytes)                                                                       \
  do {                                                                                             \
    if ((bytes) && (src))                                                                          \
      CK(cudaMemcpyAsync((dst), (src), (bytes), cudaMemcpyHostToDevice, h->stream), "H2D copy");   \
  } while (0)
#define D2H(dst, src, bytes)                                                                       \
  do {                                                                                             \
    if ((bytes) && (dst))                                                                          \
      CK(cudaMemcpyAsync((dst), (src), (bytes), cudaMemcpyDeviceToHost, h->stream), "D2H copy");   \
  } while (0)

int cabac_unpin_descriptors(cabac_handle *h) {
  if (!h) return CABAC_ERR_ARG;
  h->pinTus = h->pinTuOff = h->pinParams = nullptr;  // the device buffer stays for the next pin (grow-only)
  h->pinSubs = h->pinTuCount = 0;
  return CABAC_OK;
}

int cabac_pin_descriptors(cabac_handle *h, const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                          const cabac_substream_params *params, size_t n_sub) {
  if (!h || !tu_offsets || !params || !n_sub) return badArg(h, "cabac_pin_descriptors");
  cabac_unpin_descriptors(h);
  DeviceGuard g(h->device);
  const size_t nTu = tu_offsets[n_sub];
  if (nTu && !tus) return badArg(h, "cabac_pin_descriptors: tus");
  const size_t bTus = (nTu * sizeof(cabac_tu_desc) + 255) & ~size_t(255), bOff = ((n_sub + 1) * 8 + 255) & ~size_t(255);
  const size_t need = bTus + bOff + n_sub * sizeof(cabac_substream_params) + 256;
  if (need > h->pinDevBytes) {
    CK(cudaStreamSynchronize(h->stream), "sync");
    if (h->pinDev) cudaFree(h->pinDev);
    h->pinDev = nullptr;
    h->pinDevBytes = 0;
    CK(cudaMalloc(&h->pinDev, need + (need >> 3)), "descriptor copy");
    h->pinDevBytes = need + (need >> 3);
  }
  h->pinDevTus = static_cast<cabac_tu_desc *>(h->pinDev);
  h->pinDevTuOff = reinterpret_cast<uint64_t *>(static_cast<char *>(h->pinDev) + bTus);
  h->pinDevParams = reinterpret_cast<cabac_substream_params *>(static_cast<char *>(h->pinDev) + bTus + bOff);
  // stream-ordered: the calls that use the copy are queued behind these transfers on the same stream
  H2D(h->pinDevTus, tus, nTu * sizeof(cabac_tu_desc));
  H2D(h->pinDevTuOff, tu_offsets, (n_sub + 1) * 8);
  H2D(h->pinDevParams, params, n_sub * sizeof(cabac_substream_params));
  h->pinTus = tus;
  h->pinTuOff = tu_offsets;
  h->pinParams = params;
  h->pinSubs = n_sub;
  h->pinTuCount = nTu;
  return CABAC_OK;
}

int cabac_ctx_init(cabac_handle *h, const cabac_substream_params *params, size_t n, cabac_ctx_model *out) {
  if (!h || (n && (!params || !out))) return badArg(h, "cabac_ctx_init");
  if (!n) return CABAC_OK;
  DeviceGuard g(h->device);
  cabac_substream_params *dP;
  cabac_ctx_model *dO;
  Arena a(h);
  a.want((void **)&dP, n * sizeof(*params));
  a.want((void **)&dO, n * CABAC_NUM_CTX * sizeof(*out));
  CK(a.commit(), "scratch alloc");
  H2D(dP, params, n * sizeof(*params));
  int rc = cabac_ctx_init_dev(h, dP, n, dO);
  if (rc) return rc;
  D2H(out, dO, n * CABAC_NUM_CTX * sizeof(*out));
  CK(cudaStreamSynchronize(h->stream), "sync");
  return CABAC_OK;
}

int cabac_encode_stream(cabac_handle *h, const uint32_t *cmds, const uint64_t *cmd_offsets, const cabac_tu_desc *tus,
                        const uint64_t *tu_offsets, const cabac_substream_params *params, size_t n_sub,
                        const int32_t *coeffs, size_t n_coeffs, const cabac_ctx_model *ctx_in, uint8_t *out,
                        const uint64_t *out_offsets, uint32_t *out_sizes, uint32_t *num_bins, uint8_t *tu_flags,
                        cabac_ctx_model *ctx_out, uint32_t *status) {
  if (!h || (n_sub && (!cmd_offsets || !params || !out || !out_offsets || !out_sizes)))
    return badArg(h, "cabac_encode_stream");
  if (!n_sub) return CABAC_OK;
  DeviceGuard g(h->device);
  const size_t nCmd = cmd_offsets[n_sub], outCap = out_offsets[n_sub];
  const size_t nTu = tu_offsets ? tu_offsets[n_sub] : 0;
  if ((nCmd && !cmds) || (nTu && (!tus || !coeffs))) return badArg(h, "cabac_encode_stream: cmds/tus/coeffs");
  uint32_t *dCmds, *dSizes, *dBins, *dStatus;
  uint64_t *dCmdOff, *dOutOff, *dTuOff;
  cabac_tu_desc *dTus;
  int32_t *dCoeffs;
  cabac_substream_params *dP;
  cabac_ctx_model *dCin, *dCout;
  uint8_t *dOut, *dTuFlags;
  Arena a(h);
  a.want((void **)&dCmds, nCmd * 4);
  a.want((void **)&dCmdOff, (n_sub + 1) * 8);
  a.want((void **)&dTus, nTu * sizeof(*tus));
  a.want((void **)&dTuOff, (n_sub + 1) * 8);
  a.want((void **)&dCoeffs, nTu ? n_coeffs * 4 : 0);
  a.want((void **)&dOutOff, (n_sub + 1) * 8);
  a.want((void **)&dP, n_sub * sizeof(*params));
  a.want((void **)&dCin, ctx_in ? n_sub * CABAC_NUM_CTX * sizeof(*ctx_in) : 0);
  a.want((void **)&dCout, ctx_out ? n_sub * CABAC_NUM_CTX * sizeof(*ctx_out) : 0);
  a.want((void **)&dOut, outCap);
  a.want((void **)&dTuFlags, nTu);
  a.want((void **)&dSizes, n_sub * 4);
  a.want((void **)&dBins, n_sub * 4);
  a.want((void **)&dStatus, n_sub * 4);
  CK(a.commit(), "scratch alloc");
  H2D(dCmds, cmds, nCmd * 4);
  H2D(dCmdOff, cmd_offsets, (n_sub + 1) * 8);
  H2D(dTus, tus, nTu * sizeof(*tus));
  if (nTu) H2D(dTuOff, tu_offsets, (n_sub + 1) * 8);
  if (nTu) H2D(dCoeffs, coeffs, n_coeffs * 4);
  H2D(dOutOff, out_offsets, (n_sub + 1) * 8);
  H2D(dP, params, n_sub * sizeof(*params));
  H2D(dCin, ctx_in, n_sub * CABAC_NUM_CTX * sizeof(*ctx_in));
  int rc = cabac_encode_stream_dev(h, dCmds, dCmdOff, nTu ? dTus : nullptr, nTu ? dTuOff : nullptr, dP, n_sub,
                                   nTu ? dCoeffs : nullptr, ctx_in ? dCin : nullptr, dOut, dOutOff, dSizes, dBins,
                                   nTu ? dTuFlags : nullptr, ctx_out ? dCout : nullptr, dStatus);
  if (rc) return rc;
  D2H(out, dOut, outCap);
  D2H(out_sizes, dSizes, n_sub * 4);
  D2H(num_bins, dBins, n_sub * 4);
  D2H(tu_flags, dTuFlags, nTu);
  D2H(ctx_out, dCout, n_sub * CABAC_NUM_CTX * sizeof(*ctx_out));
  std::vector<uint32_t> st(status ? 0 : n_sub);
  uint32_t *hst = status ? status : st.data();
  D2H(hst, dStatus, n_sub * 4);
  CK(cudaStreamSynchronize(h->stream), "sync");
  return anySubstreamFailed(hst, n_sub);
}

int cabac_encode_bins(cabac_handle *h, const uint32_t *cmds, const uint64_t *cmd_offsets,
                      const cabac_substream_params *params, size_t n_sub, const cabac_ctx_model *ctx_in,
                      uint8_t *out, const uint64_t *out_offsets, uint32_t *out_sizes, uint32_t *num_bins,
                      cabac_ctx_model *ctx_out, uint32_t *status) {
  return cabac_encode_stream(h, cmds, cmd_offsets, nullptr, nullptr, params, n_sub, nullptr, 0, ctx_in, out,
                             out_offsets, out_sizes, num_bins, nullptr, ctx_out, status);
}

int cabac_decode_bins(cabac_handle *h, const uint8_t *in, const uint64_t *in_offsets, const uint32_t *cmds,
                      const uint64_t *cmd_offsets, const cabac_substream_params *params, size_t n_sub,
                      const cabac_ctx_model *ctx_in, uint32_t *values, uint32_t *bits_read,
                      cabac_ctx_model *ctx_out, uint32_t *status) {
  if (!h || (n_sub && (!in || !in_offsets || !cmd_offsets || !params))) return badArg(h, "cabac_decode_bins");
  if (!n_sub) return CABAC_OK;
  DeviceGuard g(h->device);
  const size_t nCmd = cmd_offsets[n_sub], nIn = in_offsets[n_sub];
  uint32_t *dCmds, *dVals, *dBits, *dStatus;
  uint64_t *dCmdOff, *dInOff;
  cabac_substream_params *dP;
  cabac_ctx_model *dCin, *dCout;
  uint8_t *dIn;
  Arena a(h);
  a.want((void **)&dCmds, nCmd * 4);
  a.want((void **)&dVals, nCmd * 4);
  a.want((void **)&dCmdOff, (n_sub + 1) * 8);
  a.want((void **)&dInOff, (n_sub + 1) * 8);
  a.want((void **)&dP, n_sub * sizeof(*params));
  a.want((void **)&dCin, ctx_in ? n_sub * CABAC_NUM_CTX * sizeof(*ctx_in) : 0);
  a.want((void **)&dCout, ctx_out ? n_sub * CABAC_NUM_CTX * sizeof(*ctx_out) : 0);
  a.want((void **)&dIn, nIn);
  a.want((void **)&dBits, n_sub * 4);
  a.want((void **)&dStatus, n_sub * 4);
  CK(a.commit(), "scratch alloc");
  H2D(dCmds, cmds, nCmd * 4);
  H2D(dCmdOff, cmd_offsets, (n_sub + 1) * 8);
  H2D(dInOff, in_offsets, (n_sub + 1) * 8);
  H2D(dP, params, n_sub * sizeof(*params));
  H2D(dCin, ctx_in, n_sub * CABAC_NUM_CTX * sizeof(*ctx_in));
  H2D(dIn, in, nIn);
  int rc = cabac_decode_bins_dev(h, dIn, dInOff, dCmds, dCmdOff, dP, n_sub, ctx_in ? dCin : nullptr, dVals, dBits,
                                 ctx_out ? dCout : nullptr, dStatus);
  if (rc) return rc;
  D2H(values, dVals, nCmd * 4);
  D2H(bits_read, dBits, n_sub * 4);
  D2H(ctx_out, dCout, n_sub * CABAC_NUM_CTX * sizeof(*ctx_out));
  std::vector<uint32_t> st(status ? 0 : n_sub);
  uint32_t *hst = status ? status : st.data();
  D2H(hst, dStatus, n_sub * 4);
  CK(cudaStreamSynchronize(h->stream), "sync");
  return anySubstreamFailed(hst, n_sub);
}

static int encodeSliceRowsHost(cabac_handle *h, const char *what, const cabac_substream_params *params,
                               const cabac_slice_params *slice_params, size_t n_sub, const cabac_cu_desc *cus,
                               const uint64_t *cu_offsets, const cabac_tu_desc *tus, const uint8_t *tu_flags,
                               const uint64_t *tu_offsets, const cabac_ctu_desc *ctus, const uint64_t *ctu_offsets,
                               const int32_t *coeffs, size_t n_coeffs, const cabac_ctx_model *ctx_in, size_t cmd_words,
                               const uint64_t *row_offsets, size_t n_rows, uint8_t *out, const uint64_t *out_offsets,
                               uint32_t *out_sizes, uint32_t *num_bins, cabac_ctx_model *ctx_out, uint32_t *status) {
  if (!h || (n_sub && (!params || !slice_params || !cus || !cu_offsets || !tus || !tu_flags || !tu_offsets || !ctus ||
                       !ctu_offsets || !coeffs || !out || !out_offsets || !out_sizes)))
    return badArg(h, what);
  if (!n_sub) return CABAC_OK;
  if (row_offsets && (row_offsets[0] != 0 || row_offsets[n_sub] != n_rows)) return badArg(h, what);
  DeviceGuard g(h->device);
  const size_t nCu = cu_offsets[n_sub], nTu = tu_offsets[n_sub], nCtu = ctu_offsets[n_sub], outCap = out_offsets[n_rows];
  // the kernels read coefficients where the TU records point: with host pointers the slab's size is known, so a record
  // pointing outside it is refused here (the _dev form trusts coeff_offset like every other _dev entry point)
  for (size_t t = 0; t < nTu; t++) {
    const uint64_t area = uint64_t(1) << ((tus[t].log2_w > 6 ? 6 : tus[t].log2_w) + (tus[t].log2_h > 6 ? 6 : tus[t].log2_h));
    if (tus[t].coeff_offset > n_coeffs || area > n_coeffs - tus[t].coeff_offset) {
      h->err = std::string(what) + ": TU record " + std::to_string(t) + " points outside the coefficient slab";
      return CABAC_ERR_ARG;
    }
  }
  cabac_substream_params *dP;
  cabac_slice_params *dSp;
  cabac_cu_desc *dCus;
  cabac_tu_desc *dTus;
  cabac_ctu_desc *dCtus;
  uint8_t *dTuFlags, *dOut;
  uint64_t *dCuOff, *dTuOff, *dCtuOff, *dOutOff, *dRowOff;
  int32_t *dCoeffs;
  cabac_ctx_model *dCin, *dCout;
  uint32_t *dCmds, *dSizes, *dBins, *dStatus;
  Arena a(h);
  a.want((void **)&dP, n_sub * sizeof(*params));
  a.want((void **)&dSp, n_sub * sizeof(*slice_params));
  a.want((void **)&dCus, nCu * sizeof(*cus));
  a.want((void **)&dTus, nTu * sizeof(*tus));
  a.want((void **)&dCtus, nCtu * sizeof(*ctus));
  a.want((void **)&dTuFlags, nTu);
  a.want((void **)&dCuOff, (n_sub + 1) * 8);
  a.want((void **)&dTuOff, (n_sub + 1) * 8);
  a.want((void **)&dCtuOff, (n_sub + 1) * 8);
  a.want((void **)&dRowOff, row_offsets ? (n_sub + 1) * 8 : 0);
  a.want((void **)&dOutOff, (n_rows + 1) * 8);
  a.want((void **)&dCoeffs, n_coeffs * 4);
  a.want((void **)&dCin, ctx_in ? n_sub * CABAC_NUM_CTX * sizeof(*ctx_in) : 0);
  a.want((void **)&dCout, ctx_out ? n_rows * CABAC_NUM_CTX * sizeof(*ctx_out) : 0);
  a.want((void **)&dCmds, cmd_words * 4);
  a.want((void **)&dOut, outCap);
  a.want((void **)&dSizes, n_rows * 4);
  a.want((void **)&dBins, n_rows * 4);
  a.want((void **)&dStatus, n_rows * 4);
  CK(a.commit(), "scratch alloc");
  H2D(dP, params, n_sub * sizeof(*params));
  H2D(dSp, slice_params, n_sub * sizeof(*slice_params));
  H2D(dCus, cus, nCu * sizeof(*cus));
  H2D(dTus, tus, nTu * sizeof(*tus));
  H2D(dCtus, ctus, nCtu * sizeof(*ctus));
  H2D(dTuFlags, tu_flags, nTu);
  H2D(dCuOff, cu_offsets, (n_sub + 1) * 8);
  H2D(dTuOff, tu_offsets, (n_sub + 1) * 8);
  H2D(dCtuOff, ctu_offsets, (n_sub + 1) * 8);
  H2D(dRowOff, row_offsets, (n_sub + 1) * 8);
  H2D(dOutOff, out_offsets, (n_rows + 1) * 8);
  H2D(dCoeffs, coeffs, n_coeffs * 4);
  H2D(dCin, ctx_in, n_sub * CABAC_NUM_CTX * sizeof(*ctx_in));
  const int rc = encodeSliceRows(h, what, dP, dSp, n_sub, dCus, dCuOff, dTus, dTuFlags, dTuOff, dCtus, dCtuOff, dCoeffs,
                                 ctx_in ? dCin : nullptr, dCmds, cmd_words, row_offsets ? dRowOff : nullptr, n_rows, dOut,
                                 dOutOff, dSizes, dBins, ctx_out ? dCout : nullptr, dStatus);
  std::vector<uint32_t> st(status ? 0 : n_rows);
  uint32_t *hst = status ? status : st.data();
  if (rc && rc != CABAC_ERR_SUBSTREAM) return rc;
  D2H(hst, dStatus, n_rows * 4);
  D2H(out_sizes, dSizes, n_rows * 4);
  if (!rc) {
    D2H(out, dOut, outCap);
    D2H(num_bins, dBins, n_rows * 4);
    D2H(ctx_out, dCout, n_rows * CABAC_NUM_CTX * sizeof(*ctx_out));
  }
  CK(cudaStreamSynchronize(h->stream), "sync");
  return rc ? rc : anySubstreamFailed(hst, n_rows);
}

int cabac_encode_slices(cabac_handle *h, const cabac_substream_params *params, const cabac_slice_params *slice_params,
                        size_t n_sub, const cabac_cu_desc *cus, const uint64_t *cu_offsets, const cabac_tu_desc *tus,
                        const uint8_t *tu_flags, const uint64_t *tu_offsets, const cabac_ctu_desc *ctus,
                        const uint64_t *ctu_offsets, const int32_t *coeffs, size_t n_coeffs, const cabac_ctx_model *ctx_in,
                        size_t cmd_words, uint8_t *out, const uint64_t *out_offsets, uint32_t *out_sizes,
                        uint32_t *num_bins, cabac_ctx_model *ctx_out, uint32_t *status) {
  return encodeSliceRowsHost(h, "cabac_encode_slices", params, slice_params, n_sub, cus, cu_offsets, tus, tu_flags, tu_offsets,
                             ctus, ctu_offsets, coeffs, n_coeffs, ctx_in, cmd_words, nullptr, n_sub, out, out_offsets,
                             out_sizes, num_bins, ctx_out, status);
}

int cabac_encode_slices_wpp(cabac_handle *h, const cabac_substream_params *params, const cabac_slice_params *slice_params,
                            size_t n_sub, const cabac_cu_desc *cus, const uint64_t *cu_offsets, const cabac_tu_desc *tus,
                            const uint8_t *tu_flags, const uint64_t *tu_offsets, const cabac_ctu_desc *ctus,
                            const uint64_t *ctu_offsets, const int32_t *coeffs, size_t n_coeffs,
                            const cabac_ctx_model *ctx_in, size_t cmd_words, const uint64_t *row_offsets, size_t n_rows,
                            uint8_t *out, const uint64_t *out_offsets, uint32_t *out_sizes, uint32_t *num_bins,
                            cabac_ctx_model *ctx_out, uint32_t *status) {
  if (n_sub && !row_offsets) return badArg(h, "cabac_encode_slices_wpp: row_offsets");
  return encodeSliceRowsHost(h, "cabac_encode_slices_wpp", params, slice_params, n_sub, cus, cu_offsets, tus, tu_flags,
                             tu_offsets, ctus, ctu_offsets, coeffs, n_coeffs, ctx_in, cmd_words, row_offsets, n_rows, out,
                             out_offsets, out_sizes, num_bins, ctx_out, status);
}

int cabac_decode_slices(cabac_handle *h, const uint8_t *in, const uint64_t *in_offsets,
                        const cabac_substream_params *params, const cabac_slice_params *slice_params, size_t n_sub,
                        const cabac_ctx_model *ctx_in, const uint64_t *cu_offsets, const uint64_t *tu_offsets,
                        const uint64_t *ctu_offsets, const uint64_t *coeff_offsets, cabac_cu_desc *cus,
                        cabac_tu_desc *tus, uint8_t *tu_flags, cabac_ctu_desc *ctus, int32_t *coeffs,
                        uint32_t *counts, uint32_t *num_bins, uint32_t *bits_read, cabac_ctx_model *ctx_out,
                        uint32_t *status) {
  if (!h || (n_sub && (!in || !in_offsets || !params || !slice_params || !cu_offsets || !tu_offsets || !ctu_offsets ||
                       !coeff_offsets || !tus || !coeffs || !counts)))
    return badArg(h, "cabac_decode_slices");
  if (!n_sub) return CABAC_OK;
  DeviceGuard g(h->device);
  const size_t nIn = in_offsets[n_sub], nCu = cu_offsets[n_sub], nTu = tu_offsets[n_sub], nCtu = ctu_offsets[n_sub],
               nCo = coeff_offsets[n_sub];
  uint8_t *dIn, *dTuFlags;
  uint64_t *dInOff, *dCuOff, *dTuOff, *dCtuOff, *dCoOff;
  cabac_substream_params *dP;
  cabac_slice_params *dSp;
  cabac_ctx_model *dCin, *dCout;
  cabac_cu_desc *dCus;
  cabac_tu_desc *dTus;
  cabac_ctu_desc *dCtus;
  int32_t *dCoeffs;
  uint32_t *dCounts, *dBins, *dBits, *dStatus;
  Arena a(h);
  a.want((void **)&dIn, nIn);
  a.want((void **)&dInOff, (n_sub + 1) * 8);
  a.want((void **)&dCuOff, (n_sub + 1) * 8);
  a.want((void **)&dTuOff, (n_sub + 1) * 8);
  a.want((void **)&dCtuOff, (n_sub + 1) * 8);
  a.want((void **)&dCoOff, (n_sub + 1) * 8);
  a.want((void **)&dP, n_sub * sizeof(*params));
  a.want((void **)&dSp, n_sub * sizeof(*slice_params));
  a.want((void **)&dCin, ctx_in ? n_sub * CABAC_NUM_CTX * sizeof(*ctx_in) : 0);
  a.want((void **)&dCout, ctx_out ? n_sub * CABAC_NUM_CTX * sizeof(*ctx_out) : 0);
  a.want((void **)&dCus, cus ? nCu * sizeof(*cus) : 0);
  a.want((void **)&dTus, nTu * sizeof(*tus));
  a.want((void **)&dTuFlags, nTu);
  a.want((void **)&dCtus, ctus ? nCtu * sizeof(*ctus) : 0);
  a.want((void **)&dCoeffs, nCo * 4);
  a.want((void **)&dCounts, n_sub * 12);
  a.want((void **)&dBins, n_sub * 4);
  a.want((void **)&dBits, n_sub * 4);
  a.want((void **)&dStatus, n_sub * 4);
  CK(a.commit(), "scratch alloc");
  H2D(dIn, in, nIn);
  H2D(dInOff, in_offsets, (n_sub + 1) * 8);
  H2D(dCuOff, cu_offsets, (n_sub + 1) * 8);
  H2D(dTuOff, tu_offsets, (n_sub + 1) * 8);
  H2D(dCtuOff, ctu_offsets, (n_sub + 1) * 8);
  H2D(dCoOff, coeff_offsets, (n_sub + 1) * 8);
  H2D(dP, params, n_sub * sizeof(*params));
  H2D(dSp, slice_params, n_sub * sizeof(*slice_params));
  H2D(dCin, ctx_in, n_sub * CABAC_NUM_CTX * sizeof(*ctx_in));
  int rc = cabac_decode_slices_dev(h, dIn, dInOff, dP, dSp, n_sub, ctx_in ? dCin : nullptr, dCuOff, dTuOff, dCtuOff, dCoOff,
                                   cus ? dCus : nullptr, dTus, dTuFlags, ctus ? dCtus : nullptr, dCoeffs, dCounts, dBins,
                                   dBits, ctx_out ? dCout : nullptr, dStatus);
  if (rc) return rc;
  // the counts decide how much of every region comes back
  D2H(counts, dCounts, n_sub * 12);
  std::vector<uint32_t> st(status ? 0 : n_sub);
  uint32_t *hst = status ? status : st.data();
  D2H(hst, dStatus, n_sub * 4);
  D2H(num_bins, dBins, n_sub * 4);
  D2H(bits_read, dBits, n_sub * 4);
  D2H(ctx_out, dCout, n_sub * CABAC_NUM_CTX * sizeof(*ctx_out));
  CK(cudaStreamSynchronize(h->stream), "sync");
  for (size_t i = 0; i < n_sub; i++) {
    const size_t cuCap = cu_offsets[i + 1] - cu_offsets[i], tuCap = tu_offsets[i + 1] - tu_offsets[i],
                 coCap = coeff_offsets[i + 1] - coeff_offsets[i];
    const size_t nc = std::min<size_t>(counts[3 * i], cuCap), nt = std::min<size_t>(counts[3 * i + 1], tuCap),
                 nk = std::min<size_t>(counts[3 * i + 2], coCap);
    if (cus) D2H(cus + cu_offsets[i], dCus + cu_offsets[i], nc * sizeof(*cus));
    D2H(tus + tu_offsets[i], dTus + tu_offsets[i], nt * sizeof(*tus));
    if (tu_flags) D2H(tu_flags + tu_offsets[i], dTuFlags + tu_offsets[i], nt);
    D2H(coeffs + coeff_offsets[i], dCoeffs + coeff_offsets[i], nk * 4);
  }
  if (ctus) D2H(ctus, dCtus, nCtu * sizeof(*ctus));
  CK(cudaStreamSynchronize(h->stream), "sync");
  return anySubstreamFailed(hst, n_sub);
}

int cabac_encode_residual(cabac_handle *h, const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                          const cabac_substream_params *params, size_t n_sub, const int32_t *coeffs,
                          size_t n_coeffs, const cabac_ctx_model *ctx_in, uint8_t *out,
                          const uint64_t *out_offsets, uint32_t *out_sizes, uint32_t *num_bins,
                          uint8_t *tu_flags, cabac_ctx_model *ctx_out, uint32_t *status) {
  if (!h || (n_sub && (!tu_offsets || !params || !out || !out_offsets || !out_sizes)))
    return badArg(h, "cabac_encode_residual");
  if (!n_sub) return CABAC_OK;
  DeviceGuard g(h->device);
  const size_t nTu = tu_offsets[n_sub], outCap = out_offsets[n_sub];
  if (nTu && (!tus || !coeffs)) return badArg(h, "cabac_encode_residual: tus/coeffs");
  cabac_tu_desc *dTus;
  uint64_t *dTuOff, *dOutOff;
  cabac_substream_params *dP;
  int32_t *dCoeffs;
  cabac_ctx_model *dCin, *dCout;
  uint8_t *dOut, *dTuFlags;
  uint32_t *dSizes, *dBins, *dStatus;
  Arena a(h);
  a.want((void **)&dTus, nTu * sizeof(*tus));
  a.want((void **)&dTuOff, (n_sub + 1) * 8);
  a.want((void **)&dOutOff, (n_sub + 1) * 8);
  a.want((void **)&dP, n_sub * sizeof(*params));
  a.want((void **)&dCoeffs, n_coeffs * 4);
  a.want((void **)&dCin, ctx_in ? n_sub * CABAC_NUM_CTX * sizeof(*ctx_in) : 0);
  a.want((void **)&dCout, ctx_out ? n_sub * CABAC_NUM_CTX * sizeof(*ctx_out) : 0);
  a.want((void **)&dOut, outCap);
  a.want((void **)&dTuFlags, nTu);
  a.want((void **)&dSizes, n_sub * 4);
  a.want((void **)&dBins, n_sub * 4);
  a.want((void **)&dStatus, n_sub * 4);
  CK(a.commit(), "scratch alloc");
  H2D(dTus, tus, nTu * sizeof(*tus));
  H2D(dTuOff, tu_offsets, (n_sub + 1) * 8);
  H2D(dOutOff, out_offsets, (n_sub + 1) * 8);
  H2D(dP, params, n_sub * sizeof(*params));
  H2D(dCoeffs, coeffs, n_coeffs * 4);
  H2D(dCin, ctx_in, n_sub * CABAC_NUM_CTX * sizeof(*ctx_in));
  int rc = cabac_encode_residual_dev(h, dTus, dTuOff, dP, n_sub, dCoeffs, ctx_in ? dCin : nullptr, dOut, dOutOff,
                                     dSizes, dBins, dTuFlags, ctx_out ? dCout : nullptr, dStatus);
  if (rc) return rc;
  D2H(out_sizes, dSizes, n_sub * 4);
  D2H(num_bins, dBins, n_sub * 4);
  D2H(tu_flags, dTuFlags, nTu);
  D2H(ctx_out, dCout, n_sub * CABAC_NUM_CTX * sizeof(*ctx_out));
  std::vector<uint32_t> st(status ? 0 : n_sub);
  uint32_t *hst = status ? status : st.data();
  D2H(hst, dStatus, n_sub * 4);
  CK(cudaStreamSynchronize(h->stream), "sync");
  // copy back only the bytes every substream produced
  for (size_t i = 0; i < n_sub; i++) {
    const size_t cap = out_offsets[i + 1] - out_offsets[i];
    const size_t nb = out_sizes[i] < cap ? out_sizes[i] : cap;
    if (nb) CK(cudaMemcpyAsync(out + out_offsets[i], dOut + out_offsets[i], nb, cudaMemcpyDeviceToHost, h->stream),
               "D2H copy");
  }
  CK(cudaStreamSynchronize(h->stream), "sync");
  return anySubstreamFailed(hst, n_sub);
}

static int estimateHost(cabac_handle *h, const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                        const cabac_substream_params *params, size_t n_sub, const int32_t *coeffs, size_t n_coeffs,
                        const cabac_ctx_model *ctx_in, uint64_t *est_bits, uint64_t *est_per_tu, uint32_t *num_bins,
                        uint8_t *tu_flags, cabac_ctx_model *ctx_out, uint32_t *status, uint32_t *ctx_bins) {
  if (!n_sub) return CABAC_OK;
  DeviceGuard g(h->device);
  const size_t nTu = tu_offsets[n_sub];
  if (nTu && (!tus || !coeffs)) return badArg(h, "cabac_estimate_residual_bits: tus/coeffs");
  cabac_tu_desc *dTus;
  uint64_t *dTuOff, *dEst, *dPerTu;
  cabac_substream_params *dP;
  int32_t *dCoeffs;
  cabac_ctx_model *dCin, *dCout;
  uint8_t *dTuFlags;
  uint32_t *dBins, *dStatus, *dCtxBins;
  Arena a(h);
  a.want((void **)&dTus, nTu * sizeof(*tus));
  a.want((void **)&dTuOff, (n_sub + 1) * 8);
  a.want((void **)&dCtxBins, ctx_bins ? n_sub * CABAC_NUM_CTX * 4 : 0);
  a.want((void **)&dEst, n_sub * 8);
  a.want((void **)&dPerTu, est_per_tu ? nTu * 8 : 0);
  a.want((void **)&dP, n_sub * sizeof(*params));
  a.want((void **)&dCoeffs, n_coeffs * 4);
  a.want((void **)&dCin, ctx_in ? n_sub * CABAC_NUM_CTX * sizeof(*ctx_in) : 0);
  a.want((void **)&dCout, ctx_out ? n_sub * CABAC_NUM_CTX * sizeof(*ctx_out) : 0);
  a.want((void **)&dTuFlags, nTu);
  a.want((void **)&dBins, n_sub * 4);
  a.want((void **)&dStatus, n_sub * 4);
  CK(a.commit(), "scratch alloc");
  H2D(dTus, tus, nTu * sizeof(*tus));
  H2D(dTuOff, tu_offsets, (n_sub + 1) * 8);
  H2D(dP, params, n_sub * sizeof(*params));
  H2D(dCoeffs, coeffs, n_coeffs * 4);
  H2D(dCin, ctx_in, n_sub * CABAC_NUM_CTX * sizeof(*ctx_in));
  int rc = estimateDev(h, dTus, dTuOff, dP, n_sub, dCoeffs, ctx_in ? dCin : nullptr, dEst,
                       est_per_tu ? dPerTu : nullptr, dBins, dTuFlags, ctx_out ? dCout : nullptr, dStatus,
                       ctx_bins ? dCtxBins : nullptr);
  if (rc) return rc;
  D2H(ctx_bins, dCtxBins, n_sub * CABAC_NUM_CTX * 4);
  D2H(est_bits, dEst, n_sub * 8);
  D2H(est_per_tu, dPerTu, nTu * 8);
  D2H(num_bins, dBins, n_sub * 4);
  D2H(tu_flags, dTuFlags, nTu);
  D2H(ctx_out, dCout, n_sub * CABAC_NUM_CTX * sizeof(*ctx_out));
  std::vector<uint32_t> st(status ? 0 : n_sub);
  uint32_t *hst = status ? status : st.data();
  D2H(hst, dStatus, n_sub * 4);
  CK(cudaStreamSynchronize(h->stream), "sync");
  return anySubstreamFailed(hst, n_sub);
}

int cabac_estimate_residual_bits(cabac_handle *h, const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                                 const cabac_substream_params *params, size_t n_sub, const int32_t *coeffs,
                                 size_t n_coeffs, const cabac_ctx_model *ctx_in, uint64_t *est_bits,
                                 uint64_t *est_per_tu, uint32_t *num_bins, uint8_t *tu_flags,
                                 cabac_ctx_model *ctx_out, uint32_t *status) {
  if (!h || (n_sub && (!tu_offsets || !params || !est_bits))) return badArg(h, "cabac_estimate_residual_bits");
  return estimateHost(h, tus, tu_offsets, params, n_sub, coeffs, n_coeffs, ctx_in, est_bits, est_per_tu, num_bins,
                      tu_flags, ctx_out, status, nullptr);
}

int cabac_residual_ctx_bins(cabac_handle *h, const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                            const cabac_substream_params *params, size_t n_sub, const int32_t *coeffs,
                            size_t n_coeffs, const cabac_ctx_model *ctx_in, uint32_t *ctx_bins, uint32_t *num_bins,
                            cabac_ctx_model *ctx_out, uint32_t *status) {
  if (!h || (n_sub && (!tu_offsets || !params || !ctx_bins))) return badArg(h, "cabac_residual_ctx_bins");
  return estimateHost(h, tus, tu_offsets, params, n_sub, coeffs, n_coeffs, ctx_in, nullptr, nullptr, num_bins, nullptr,
                      ctx_out, status, ctx_bins);
}

int cabac_ctx_init_id(cabac_handle *h, const cabac_substream_params *params, const cabac_ctx_model *ctx_ref,
                      const uint32_t *ctx_bins, size_t n, uint8_t *choice, uint64_t *costs) {
  if (!h || (n && (!params || !ctx_ref || !ctx_bins || !choice))) return badArg(h, "cabac_ctx_init_id");
  if (!n) return CABAC_OK;
  DeviceGuard g(h->device);
  cabac_substream_params *dP;
  cabac_ctx_model *dRef;
  uint32_t *dBins;
  uint8_t *dChoice;
  uint64_t *dCosts;
  Arena a(h);
  a.want((void **)&dP, n * sizeof(*params));
  a.want((void **)&dRef, n * CABAC_NUM_CTX * sizeof(*ctx_ref));
  a.want((void **)&dBins, n * CABAC_NUM_CTX * 4);
  a.want((void **)&dChoice, n);
  a.want((void **)&dCosts, n * 16);
  CK(a.commit(), "scratch alloc");
  H2D(dP, params, n * sizeof(*params));
  H2D(dRef, ctx_ref, n * CABAC_NUM_CTX * sizeof(*ctx_ref));
  H2D(dBins, ctx_bins, n * CABAC_NUM_CTX * 4);
  int rc = cabac_ctx_init_id_dev(h, dP, dRef, dBins, n, dChoice, dCosts);
  if (rc) return rc;
  D2H(choice, dChoice, n);
  D2H(costs, dCosts, n * 16);
  CK(cudaStreamSynchronize(h->stream), "sync");
  return CABAC_OK;
}

int cabac_encode_residual_packed(cabac_handle *h, const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                                 const cabac_substream_params *params, size_t n_sub, const int32_t *coeffs,
                                 size_t n_coeffs, const cabac_ctx_model *ctx_in, const uint64_t *out_offsets,
                                 uint8_t *packed, size_t packed_cap, uint64_t *entry_points, uint32_t *out_sizes,
                                 uint32_t *num_bins, uint8_t *tu_flags, cabac_ctx_model *ctx_out,
                                 uint32_t *status) {
  if (!h || !entry_points || (n_sub && (!tu_offsets || !params || !out_offsets || !packed)))
    return badArg(h, "cabac_encode_residual_packed");
  if (!n_sub) {
    entry_points[0] = 0;
    return CABAC_OK;
  }
  DeviceGuard g(h->device);
  const size_t nTu = tu_offsets[n_sub], outCap = out_offsets[n_sub];
  if (nTu && (!tus || !coeffs)) return badArg(h, "cabac_encode_residual_packed: tus/coeffs");
  cabac_tu_desc *dTus;
  uint64_t *dTuOff, *dOutOff, *dEntry, *dRanges;
  uint32_t *dChunkStart;
  cabac_substream_params *dP;
  int32_t *dCoeffs;
  cabac_ctx_model *dCin, *dCout;
  uint8_t *dOut, *dTuFlags, *dPacked;
  uint32_t *dSizes, *dBins, *dStatus;
  const bool pinnedDesc = h->pinned(tus, tu_offsets, params, n_sub);  // cabac_pin_descriptors: already on the device
  Arena a(h);
  a.want((void **)&dTus, pinnedDesc ? 0 : nTu * sizeof(*tus));
  a.want((void **)&dTuOff, pinnedDesc ? 0 : (n_sub + 1) * 8);
  a.want((void **)&dOutOff, (n_sub + 1) * 8);
  a.want((void **)&dEntry, (n_sub + 1) * 8);
  a.want((void **)&dRanges, 2 * cabac_handle::kMaxChunks * sizeof(uint64_t));
  a.want((void **)&dChunkStart, (cabac_handle::kMaxChunks + 1) * sizeof(uint32_t));
  a.want((void **)&dP, pinnedDesc ? 0 : n_sub * sizeof(*params));
  const bool twoStage = !h->wppNext && (h->encoderMode == CABAC_ENCODER_TWO_STAGE ||
                                        (h->encoderMode == CABAC_ENCODER_AUTO && n_sub >= kTwoStageMinSubstreams));
  const size_t esz = h->coeff16Next ? 2 : 4;  // bytes per coefficient of the caller's slab
  // few, long substreams go to the sliced encoder, which makes its passes over a slab in HBM
  const bool slicedLikely = !h->wppNext && (h->encoderMode == CABAC_ENCODER_SLICED ||
                                            (h->encoderMode == CABAC_ENCODER_AUTO && n_sub < kSlicedMaxSubstreams &&
                                             nTu >= kSlicedMinTus && nTu >= kSlicedMinTusPerSubstream * n_sub));
  const int32_t *zcCoeffs = (h->zeroCopy && !twoStage && !slicedLikely && n_coeffs >= kZeroCopyMinCoeffs)
                                ? static_cast<const int32_t *>(mappedHostAlias(coeffs)) : nullptr;
  a.want((void **)&dCoeffs, zcCoeffs ? 0 : n_coeffs * esz);
  a.want((void **)&dCin, ctx_in ? n_sub * CABAC_NUM_CTX * sizeof(*ctx_in) : 0);
  a.want((void **)&dCout, ctx_out ? n_sub * CABAC_NUM_CTX * sizeof(*ctx_out) : 0);
  a.want((void **)&dOut, outCap);
  a.want((void **)&dPacked, outCap);
  a.want((void **)&dTuFlags, nTu);
  a.want((void **)&dSizes, n_sub * 4);
  a.want((void **)&dBins, n_sub * 4);
  a.want((void **)&dStatus, n_sub * 4);
  CK(a.commit(), "scratch alloc");
  if (pinnedDesc) {
    dTus = h->pinDevTus;
    dTuOff = h->pinDevTuOff;
    dP = h->pinDevParams;
  } else {
    H2D(dTus, tus, nTu * sizeof(*tus));
    H2D(dTuOff, tu_offsets, (n_sub + 1) * 8);
    H2D(dP, params, n_sub * sizeof(*params));
  }
  H2D(dOutOff, out_offsets, (n_sub + 1) * 8);
  H2D(dCin, ctx_in, n_sub * CABAC_NUM_CTX * sizeof(*ctx_in));
  int rc;
  if (zcCoeffs) {
    // ---- zero-copy: the kernel reads every TU block straight from pinned host memory (each
    // coefficient crosses PCIe exactly once, under the coding of the other warps); the schedule comes
    // from block areas and QPs instead of a pass over the slab ----
    h->areaCostNext = true;
    rc = cabac_encode_residual_dev(h, dTus, dTuOff, dP, n_sub, zcCoeffs, ctx_in ? dCin : nullptr, dOut, dOutOff,
                                   dSizes, dBins, dTuFlags, ctx_out ? dCout : nullptr, dStatus);
    if (rc) return rc;
  } else if (h->pipeChunks > 1 && !twoStage && !slicedLikely && !h->wppNext && n_coeffs >= h->pipeMinCoeffs &&
             n_sub >= h->pipeMinSubstreams) {
    // ---- chunked pipeline: the slab crosses PCIe chunk by chunk while earlier chunks are coded ----
    CK(ensurePipe(h), "pipeline streams");
    uint64_t *hr = h->pinnedRanges;
    uint32_t *hStart = reinterpret_cast<uint32_t *>(hr + 2 * cabac_handle::kMaxChunks);
    const int K = planChunks(tu_offsets, n_sub, h->pipeChunks, hStart);
    for (int k = 0; k < K; k++) {
      hr[2 * k] = ~0ull;
      hr[2 * k + 1] = 0;
    }
    H2D(dRanges, hr, 2 * K * sizeof(uint64_t));
    H2D(dChunkStart, hStart, (K + 1) * sizeof(uint32_t));
    CK(cabac::launchCoeffRanges(h->stream, dTus, dTuOff, dChunkStart, K, dRanges), "coeff_range_kernel");
    h->launches += 1;
    D2H(hr, dRanges, 2 * K * sizeof(uint64_t));
    CK(cudaStreamSynchronize(h->stream), "sync");  // descriptors are resident from here on
    uint64_t cur = ~0ull;
    for (int k = 0; k < K; k++) {
      if (hr[2 * k + 1] > n_coeffs) return badArg(h, "cabac_encode_residual_packed: descriptor beyond n_coeffs");
      if (hr[2 * k] < cur) cur = hr[2 * k];
    }
    cabac::Sched base;
    CK(schedFor(h, n_sub, base), "scheduler scratch");
    if (h->trace) CK(cudaEventRecord(h->evT0, h->copyStream), "event");
    for (int k = 0; k < K; k++) {
      const uint64_t hi = hr[2 * k + 1];
      if (hi > cur) {  // everything below `cur` is already on its way: any slab layout stays correct
        CK(cudaMemcpyAsync(reinterpret_cast<char *>(dCoeffs) + cur * esz, reinterpret_cast<const char *>(coeffs) + cur * esz,
                           (hi - cur) * esz, cudaMemcpyHostToDevice, h->copyStream),
           "H2D copy");
        cur = hi;
      }
      CK(cudaEventRecord(h->evCopy[k], h->copyStream), "event");
      cudaStream_t cs = h->compStream[k];
      CK(cudaStreamWaitEvent(cs, h->evCopy[k], 0), "event wait");
      const size_t s0 = hStart[k], nk = hStart[k + 1] - s0;
      cabac::Sched sc = chunkSched(base, s0, k);
      CK(cabac::launchEncodePrepare(cs, h->device, sc, dTus, dTuOff + s0, nk, dCoeffs, nullptr), "encode prepare kernels");
      CK(cabac::launchEncodeResidual(cs, h->device, sc, dTus, dTuOff + s0, dP + s0, nk, dCoeffs,
                                     ctx_in ? dCin + s0 * CABAC_NUM_CTX : nullptr, dOut, dOutOff + s0, dSizes + s0,
                                     dBins + s0, dTuFlags, ctx_out ? dCout + s0 * CABAC_NUM_CTX : nullptr, dStatus + s0),
         "encode_residual kernels");
      h->launches += 4;
      CK(cudaEventRecord(h->evDone[k], cs), "event");
    }
    for (int k = 0; k < K; k++) CK(cudaStreamWaitEvent(h->stream, h->evDone[k], 0), "event wait");
    if (h->trace) {
      CK(cudaEventRecord(h->evJoin, h->stream), "event");
      CK(cudaStreamSynchronize(h->stream), "sync");
      traceChunks(h, "encode_residual_packed", K, true);
    }
  } else {
    H2D(dCoeffs, coeffs, n_coeffs * esz);
    rc = cabac_encode_residual_dev(h, dTus, dTuOff, dP, n_sub, dCoeffs, ctx_in ? dCin : nullptr, dOut, dOutOff,
                                   dSizes, dBins, dTuFlags, ctx_out ? dCout : nullptr, dStatus);
    if (rc) return rc;
  }
  rc = cabac_pack_dev(h, dOut, dOutOff, dSizes, n_sub, dPacked, dEntry);
  if (rc) return rc;
  D2H(entry_points, dEntry, (n_sub + 1) * 8);
  D2H(out_sizes, dSizes, n_sub * 4);
  D2H(num_bins, dBins, n_sub * 4);
  D2H(tu_flags, dTuFlags, nTu);
  D2H(ctx_out, dCout, n_sub * CABAC_NUM_CTX * sizeof(*ctx_out));
  std::vector<uint32_t> st(status ? 0 : n_sub);
  uint32_t *hst = status ? status : st.data();
  D2H(hst, dStatus, n_sub * 4);
  CK(cudaStreamSynchronize(h->stream), "sync");
  const size_t total = entry_points[n_sub];
  if (total > packed_cap) return badArg(h, "cabac_encode_residual_packed: packed_cap too small");
  D2H(packed, dPacked, total);
  CK(cudaStreamSynchronize(h->stream), "sync");
  return anySubstreamFailed(hst, n_sub);
}

int cabac_decode_residual(cabac_handle *h, const uint8_t *in, const uint64_t *in_offsets, const cabac_tu_desc *tus,
                          const uint64_t *tu_offsets, const cabac_substream_params *params, size_t n_sub,
                          const cabac_ctx_model *ctx_in, int32_t *coeffs, size_t n_coeffs, uint32_t *num_bins,
                          uint32_t *bits_read, uint8_t *tu_flags, cabac_ctx_model *ctx_out, uint32_t *status) {
  if (!h || (n_sub && (!in || !in_offsets || !tu_offsets || !params || !coeffs)))
    return badArg(h, "cabac_decode_residual");
  if (!n_sub) return CABAC_OK;
  DeviceGuard g(h->device);
  const size_t nTu = tu_offsets[n_sub], nIn = in_offsets[n_sub];
  cabac_tu_desc *dTus;
  uint64_t *dTuOff, *dInOff, *dRanges;
  uint32_t *dChunkStart;
  cabac_substream_params *dP;
  int32_t *dCoeffs;
  cabac_ctx_model *dCin, *dCout;
  uint8_t *dIn, *dTuFlags;
  uint32_t *dBins, *dBits, *dStatus;
  const bool pinnedDesc = h->pinned(tus, tu_offsets, params, n_sub);  // cabac_pin_descriptors: already on the device
  Arena a(h);
  a.want((void **)&dTus, pinnedDesc ? 0 : nTu * sizeof(*tus));
  a.want((void **)&dTuOff, pinnedDesc ? 0 : (n_sub + 1) * 8);
  a.want((void **)&dInOff, (n_sub + 1) * 8);
  a.want((void **)&dRanges, 2 * cabac_handle::kMaxChunks * sizeof(uint64_t));
  a.want((void **)&dChunkStart, (cabac_handle::kMaxChunks + 1) * sizeof(uint32_t));
  a.want((void **)&dP, pinnedDesc ? 0 : n_sub * sizeof(*params));
  const size_t esz = h->coeff16Next ? 2 : 4;  // bytes per coefficient of the caller's slab
  int32_t *zcCandidate = (h->zeroCopy && n_coeffs >= kZeroCopyMinCoeffs)
                             ? static_cast<int32_t *>(mappedHostAlias(coeffs)) : nullptr;
  a.want((void **)&dCoeffs, n_coeffs * esz);  // kept even for a zero-copy candidate: the coverage check may refuse
  a.want((void **)&dCin, ctx_in ? n_sub * CABAC_NUM_CTX * sizeof(*ctx_in) : 0);
  a.want((void **)&dCout, ctx_out ? n_sub * CABAC_NUM_CTX * sizeof(*ctx_out) : 0);
  a.want((void **)&dIn, nIn);
  a.want((void **)&dTuFlags, nTu);
  a.want((void **)&dBins, n_sub * 4);
  a.want((void **)&dBits, n_sub * 4);
  a.want((void **)&dStatus, n_sub * 4);
  CK(a.commit(), "scratch alloc");
  if (pinnedDesc) {
    dTus = h->pinDevTus;
    dTuOff = h->pinDevTuOff;
    dP = h->pinDevParams;
  } else {
    H2D(dTus, tus, nTu * sizeof(*tus));
    H2D(dTuOff, tu_offsets, (n_sub + 1) * 8);
    H2D(dP, params, n_sub * sizeof(*params));
  }
  H2D(dInOff, in_offsets, (n_sub + 1) * 8);
  H2D(dCin, ctx_in, n_sub * CABAC_NUM_CTX * sizeof(*ctx_in));
  H2D(dIn, in, nIn);
  // positions no descriptor covers come back as zero
  // ---- zero-copy: when the descriptors tile the whole slab (their areas sum to n_coeffs over the
  // range [0, n_coeffs)), every position is written by the kernel itself, so it can store straight into
  // the caller's pinned host slab: no device slab, no memset, no D2H copy ----
  int32_t *zcCoeffs = nullptr;
  if (zcCandidate) {
    CK(ensurePipe(h), "pipeline streams");  // owns the pinned landing area used below
    uint64_t *hr = h->pinnedRanges;
    uint32_t *hStart = reinterpret_cast<uint32_t *>(hr + 2 * cabac_handle::kMaxChunks);
    hStart[0] = 0;
    hStart[1] = uint32_t(n_sub);
    hr[0] = ~0ull;
    hr[1] = 0;
    hr[2] = 0;
    H2D(dRanges, hr, 3 * sizeof(uint64_t));
    H2D(dChunkStart, hStart, 2 * sizeof(uint32_t));
    CK(cabac::launchCoeffRanges(h->stream, dTus, dTuOff, dChunkStart, 1, dRanges, dRanges + 2), "coeff_range_kernel");
    h->launches += 1;
    D2H(hr, dRanges, 3 * sizeof(uint64_t));
    CK(cudaStreamSynchronize(h->stream), "sync");
    if (hr[0] == 0 && hr[1] == n_coeffs && hr[2] == n_coeffs) zcCoeffs = zcCandidate;
  }
  if (zcCoeffs) {
    int rc = cabac_decode_residual_dev(h, dIn, dInOff, dTus, dTuOff, dP, n_sub, ctx_in ? dCin : nullptr, zcCoeffs,
                                       dBins, dBits, dTuFlags, ctx_out ? dCout : nullptr, dStatus);
    if (rc) return rc;
  } else if (h->pipeChunks > 1 && !h->wppNext && n_coeffs >= h->pipeMinCoeffs && n_sub >= h->pipeMinSubstreams) {
    CK(cudaMemsetAsync(dCoeffs, 0, n_coeffs * esz, h->stream), "memset");
    // ---- chunked pipeline: finished coefficient ranges leave over PCIe while later chunks decode ----
    CK(ensurePipe(h), "pipeline streams");
    uint64_t *hr = h->pinnedRanges;
    uint32_t *hStart = reinterpret_cast<uint32_t *>(hr + 2 * cabac_handle::kMaxChunks);
    const int K = planChunks(tu_offsets, n_sub, h->pipeChunks, hStart);
    for (int k = 0; k < K; k++) {
      hr[2 * k] = ~0ull;
      hr[2 * k + 1] = 0;
    }
    H2D(dRanges, hr, 2 * K * sizeof(uint64_t));
    H2D(dChunkStart, hStart, (K + 1) * sizeof(uint32_t));
    CK(cabac::launchCoeffRanges(h->stream, dTus, dTuOff, dChunkStart, K, dRanges), "coeff_range_kernel");
    h->launches += 1;
    D2H(hr, dRanges, 2 * K * sizeof(uint64_t));
    CK(cudaStreamSynchronize(h->stream), "sync");  // inputs resident, slab zeroed
    uint64_t sufMin[cabac_handle::kMaxChunks + 1];
    sufMin[K] = n_coeffs;
    for (int k = K - 1; k >= 0; k--) {
      if (hr[2 * k + 1] > n_coeffs) return badArg(h, "cabac_decode_residual: descriptor beyond n_coeffs");
      sufMin[k] = hr[2 * k] < sufMin[k + 1] ? hr[2 * k] : sufMin[k + 1];
    }
    cabac::Sched base;
    CK(schedFor(h, n_sub, base), "scheduler scratch");
    uint64_t cur = 0;
    if (h->trace) CK(cudaEventRecord(h->evT0, h->compStream[0]), "event");
    for (int k = 0; k < K; k++) {
      cudaStream_t cs = h->compStream[k];
      const size_t s0 = hStart[k], nk = hStart[k + 1] - s0;
      cabac::Sched sc = chunkSched(base, s0, k);
      CK(cabac::launchDecodeResidual(cs, h->device, sc, dIn, dInOff + s0, dTus, dTuOff + s0, dP + s0, nk,
                                     ctx_in ? dCin + s0 * CABAC_NUM_CTX : nullptr, dCoeffs, dBins + s0, dBits + s0,
                                     dTuFlags, ctx_out ? dCout + s0 * CABAC_NUM_CTX : nullptr, dStatus + s0),
         "decode_residual_kernel");
      h->launches += cabac::kLaunchesPerCodingCall;
      CK(cudaEventRecord(h->evDone[k], cs), "event");
      // chunks 0..k are complete once the copy stream has passed evDone[0..k]; no later chunk
      // writes below sufMin[k + 1]
      CK(cudaStreamWaitEvent(h->copyStream, h->evDone[k], 0), "event wait");
      const uint64_t safe = sufMin[k + 1];
      if (safe > cur) {
        CK(cudaMemcpyAsync(reinterpret_cast<char *>(coeffs) + cur * esz, reinterpret_cast<char *>(dCoeffs) + cur * esz,
                           (safe - cur) * esz, cudaMemcpyDeviceToHost, h->copyStream),
           "D2H copy");
        cur = safe;
      }
    }
    CK(cudaEventRecord(h->evJoin, h->copyStream), "event");
    CK(cudaStreamWaitEvent(h->stream, h->evJoin, 0), "event wait");
    if (h->trace) {
      CK(cudaStreamSynchronize(h->stream), "sync");
      traceChunks(h, "decode_residual", K, false);
    }
  } else {
    CK(cudaMemsetAsync(dCoeffs, 0, n_coeffs * esz, h->stream), "memset");
    int rc = cabac_decode_residual_dev(h, dIn, dInOff, dTus, dTuOff, dP, n_sub, ctx_in ? dCin : nullptr, dCoeffs,
                                       dBins, dBits, dTuFlags, ctx_out ? dCout : nullptr, dStatus);
    if (rc) return rc;
    D2H(coeffs, dCoeffs, n_coeffs * esz);
  }
  D2H(num_bins, dBins, n_sub * 4);
  D2H(bits_read, dBits, n_sub * 4);
  D2H(tu_flags, dTuFlags, nTu);
  D2H(ctx_out, dCout, n_sub * CABAC_NUM_CTX * sizeof(*ctx_out));
  std::vector<uint32_t> st(status ? 0 : n_sub);
  uint32_t *hst = status ? status : st.data();
  D2H(hst, dStatus, n_sub * 4);
  CK(cudaStreamSynchronize(h->stream), "sync");
  return anySubstreamFailed(hst, n_sub);
}

int cabac_pack(cabac_handle *h, const uint8_t *out, const uint64_t *out_offsets, const uint32_t *out_sizes,
               size_t n_sub, uint8_t *packed, size_t packed_cap, uint64_t *entry_points) {
  if (!h || !entry_points || (n_sub && (!out || !out_offsets || !out_sizes || !packed)))
    return badArg(h, "cabac_pack");
  DeviceGuard g(h->device);
  size_t total = 0;
  for (size_t i = 0; i < n_sub; i++) total += out_sizes[i];
  if (total > packed_cap) return badArg(h, "cabac_pack: packed_cap too small");
  const size_t outCap = n_sub ? out_offsets[n_sub] : 0;
  uint8_t *dOut, *dPacked;
  uint64_t *dOutOff, *dEntry;
  uint32_t *dSizes;
  Arena a(h);
  a.want((void **)&dOut, outCap);
  a.want((void **)&dPacked, total);
  a.want((void **)&dOutOff, (n_sub + 1) * 8);
  a.want((void **)&dEntry, (n_sub + 1) * 8);
  a.want((void **)&dSizes, n_sub * 4);
  CK(a.commit(), "scratch alloc");
  H2D(dOut, out, outCap);
  H2D(dOutOff, out_offsets, (n_sub + 1) * 8);
  H2D(dSizes, out_sizes, n_sub * 4);
  int rc = cabac_pack_dev(h, dOut, dOutOff, dSizes, n_sub, dPacked, dEntry);
  if (rc) return rc;
  D2H(packed, dPacked, total);
  D2H(entry_points, dEntry, (n_sub + 1) * 8);
  CK(cudaStreamSynchronize(h->stream), "sync");
  return CABAC_OK;
}

int cabac_count_emulations(cabac_handle *h, const uint8_t *data, const uint64_t *offsets, size_t n_seg,
                           uint32_t *counts) {
  if (!h || (n_seg && (!data || !offsets || !counts))) return badArg(h, "cabac_count_emulations");
  if (!n_seg) return CABAC_OK;
  DeviceGuard g(h->device);
  const size_t nBytes = offsets[n_seg];
  uint8_t *dData;
  uint64_t *dOff;
  uint32_t *dCnt;
  Arena a(h);
  a.want((void **)&dData, nBytes);
  a.want((void **)&dOff, (n_seg + 1) * 8);
  a.want((void **)&dCnt, n_seg * 4);
  CK(a.commit(), "scratch alloc");
  H2D(dData, data, nBytes);
  H2D(dOff, offsets, (n_seg + 1) * 8);
  int rc = cabac_count_emulations_dev(h, dData, dOff, n_seg, dCnt);
  if (rc) return rc;
  D2H(counts, dCnt, n_seg * 4);
  CK(cudaStreamSynchronize(h->stream), "sync");
  return CABAC_OK;
}

int cabac_insert_emulation_prevention(cabac_handle *h, const uint8_t *rbsp, size_t n, uint8_t *out, size_t out_cap,
                                      uint64_t *out_size) {
  if (!h || !out_size || (n && (!rbsp || !out))) return badArg(h, "cabac_insert_emulation_prevention");
  DeviceGuard g(h->device);
  const size_t worst = n + n / 2 + 1;
  uint8_t *dIn, *dOut;
  uint64_t *dSize;
  Arena a(h);
  a.want((void **)&dIn, n);
  a.want((void **)&dOut, worst);
  a.want((void **)&dSize, 8);
  CK(a.commit(), "scratch alloc");
  H2D(dIn, rbsp, n);
  int rc = cabac_insert_emulation_prevention_dev(h, dIn, n, dOut, worst, dSize);
  if (rc) return rc;
  D2H(out_size, dSize, 8);
  CK(cudaStreamSynchronize(h->stream), "sync");
  if (*out_size > out_cap) return badArg(h, "cabac_insert_emulation_prevention: out_cap too small");
  D2H(out, dOut, size_t(*out_size));
  CK(cudaStreamSynchronize(h->stream), "sync");
  return CABAC_OK;
}

// ---- int16 coefficient slabs: the same calls, half the bytes across PCIe ----
// VVC levels fit 16 bits (maxLog2TrDynamicRange = 15, context_modelling.cpp:66-72); TCoeff is int32 in the
// reference (unit.hpp:551), so the shim narrows while it gathers the blocks into the batch slab anyway.
int cabac_encode_residual_i16_dev(cabac_handle *h, const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                                  const cabac_substream_params *params, size_t n_sub, const int16_t *coeffs,
                                  const cabac_ctx_model *ctx_in, uint8_t *out, const uint64_t *out_offsets,
                                  uint32_t *out_sizes, uint32_t *num_bins, uint8_t *tu_flags, cabac_ctx_model *ctx_out,
                                  uint32_t *status) {
  if (!h) return CABAC_ERR_ARG;
  h->coeff16Next = true;
  const int rc = cabac_encode_residual_dev(h, tus, tu_offsets, params, n_sub, reinterpret_cast<const int32_t *>(coeffs),
                                           ctx_in, out, out_offsets, out_sizes, num_bins, tu_flags, ctx_out, status);
  h->coeff16Next = false;
  return rc;
}

int cabac_decode_residual_i16_dev(cabac_handle *h, const uint8_t *in, const uint64_t *in_offsets, const cabac_tu_desc *tus,
                                  const uint64_t *tu_offsets, const cabac_substream_params *params, size_t n_sub,
                                  const cabac_ctx_model *ctx_in, int16_t *coeffs, uint32_t *num_bins, uint32_t *bits_read,
                                  uint8_t *tu_flags, cabac_ctx_model *ctx_out, uint32_t *status) {
  if (!h) return CABAC_ERR_ARG;
  h->coeff16Next = true;
  const int rc = cabac_decode_residual_dev(h, in, in_offsets, tus, tu_offsets, params, n_sub, ctx_in,
                                           reinterpret_cast<int32_t *>(coeffs), num_bins, bits_read, tu_flags, ctx_out,
                                           status);
  h->coeff16Next = false;
  return rc;
}

int cabac_encode_residual_packed_i16(cabac_handle *h, const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                                     const cabac_substream_params *params, size_t n_sub, const int16_t *coeffs,
                                     size_t n_coeffs, const cabac_ctx_model *ctx_in, const uint64_t *out_offsets,
                                     uint8_t *packed, size_t packed_cap, uint64_t *entry_points, uint32_t *out_sizes,
                                     uint32_t *num_bins, uint8_t *tu_flags, cabac_ctx_model *ctx_out, uint32_t *status) {
  if (!h) return CABAC_ERR_ARG;
  h->coeff16Next = true;
  const int rc = cabac_encode_residual_packed(h, tus, tu_offsets, params, n_sub, reinterpret_cast<const int32_t *>(coeffs),
                                              n_coeffs, ctx_in, out_offsets, packed, packed_cap, entry_points, out_sizes,
                                              num_bins, tu_flags, ctx_out, status);
  h->coeff16Next = false;
  return rc;
}

int cabac_decode_residual_i16(cabac_handle *h, const uint8_t *in, const uint64_t *in_offsets, const cabac_tu_desc *tus,
                              const uint64_t *tu_offsets, const cabac_substream_params *params, size_t n_sub,
                              const cabac_ctx_model *ctx_in, int16_t *coeffs, size_t n_coeffs, uint32_t *num_bins,
                              uint32_t *bits_read, uint8_t *tu_flags, cabac_ctx_model *ctx_out, uint32_t *status) {
  if (!h) return CABAC_ERR_ARG;
  h->coeff16Next = true;
  const int rc = cabac_decode_residual(h, in, in_offsets, tus, tu_offsets, params, n_sub, ctx_in,
                                       reinterpret_cast<int32_t *>(coeffs), n_coeffs, num_bins, bits_read, tu_flags,
                                       ctx_out, status);
  h->coeff16Next = false;
  return rc;
}

// ---- decode side of substream assembly ----
static int ensureEpbScratch(cabac_handle *h, size_t n) {
  const size_t tiles = cabac::emulationTiles(n);
  if (tiles > h->epbTiles) {
    if (h->epbScratch) cudaFree(h->epbScratch);
    h->epbScratch = nullptr;
    h->epbTiles = 0;
    const size_t cap = tiles + (tiles >> 2) + 256;
    CK(cudaMalloc(&h->epbScratch, cap * 4 + (cap + 1) * 8 + 16), "emulation scratch");
    h->epbTiles = cap;
  }
  return CABAC_OK;
}

int cabac_remove_emulation_prevention_dev(cabac_handle *h, const uint8_t *nal, size_t n, int strip_trailing_zeros,
                                          uint8_t *rbsp, uint64_t *sizes, uint32_t *epb_locations,
                                          uint32_t *violations) {
  if (!h || !sizes || !violations || (n && (!nal || !rbsp))) return badArg(h, "cabac_remove_emulation_prevention_dev");
  DeviceGuard g(h->device);
  if (int rc = ensureEpbScratch(h, n)) return rc;
  uint64_t *tileBase = static_cast<uint64_t *>(h->epbScratch);
  uint32_t *tileCounts = reinterpret_cast<uint32_t *>(tileBase + h->epbTiles + 1);
  CK(cabac::launchEmulationRemove(h->stream, nal, n, strip_trailing_zeros, tileCounts, tileBase, rbsp, epb_locations,
                                  sizes, violations),
     "emulation removal kernels");
  h->launches += n ? 3 : 0;
  return CABAC_OK;
}

int cabac_remove_emulation_prevention(cabac_handle *h, const uint8_t *nal, size_t n, int strip_trailing_zeros,
                                      uint8_t *rbsp, uint64_t *sizes, uint32_t *epb_locations, uint32_t *violations) {
  if (!h || !sizes || (n && (!nal || !rbsp))) return badArg(h, "cabac_remove_emulation_prevention");
  DeviceGuard g(h->device);
  uint8_t *dIn, *dOut;
  uint64_t *dSizes;
  uint32_t *dEpb, *dViol;
  Arena a(h);
  a.want((void **)&dIn, n);
  a.want((void **)&dOut, n);
  a.want((void **)&dSizes, 16);
  a.want((void **)&dViol, 4);
  a.want((void **)&dEpb, (n / 3 + 1) * 4);
  CK(a.commit(), "scratch alloc");
  H2D(dIn, nal, n);
  int rc = cabac_remove_emulation_prevention_dev(h, dIn, n, strip_trailing_zeros, dOut, dSizes, dEpb, dViol);
  if (rc) return rc;
  uint32_t viol = 0;
  D2H(sizes, dSizes, 16);
  D2H(&viol, dViol, 4);
  CK(cudaStreamSynchronize(h->stream), "sync");
  if (violations) *violations = viol;
  D2H(rbsp, dOut, size_t(sizes[0]));
  D2H(epb_locations, dEpb, size_t(sizes[1]) * 4);
  CK(cudaStreamSynchronize(h->stream), "sync");
  if (viol) return badArg(h, "cabac_remove_emulation_prevention: malformed NAL payload (NALread.cpp:66, 79, 84)");
  return CABAC_OK;
}

int cabac_entry_point_syntax_dev(cabac_handle *h, const uint32_t *sizes, const uint32_t *emulations, size_t n_sub,
                                 uint32_t *offset_minus1, uint32_t *offset_len_minus1) {
  if (!h || !offset_len_minus1 || (n_sub && !sizes) || (n_sub > 1 && !offset_minus1))
    return badArg(h, "cabac_entry_point_syntax_dev");
  DeviceGuard g(h->device);
  CK(cabac::launchEntryPointSyntax(h->stream, sizes, emulations, n_sub, offset_minus1, offset_len_minus1),
     "entry_point_syntax_kernel");
  h->launches += 1;
  return CABAC_OK;
}

int cabac_entry_point_syntax(cabac_handle *h, const uint32_t *sizes, const uint32_t *emulations, size_t n_sub,
                             uint32_t *offset_minus1, uint32_t *offset_len_minus1) {
  if (!h || !offset_len_minus1 || (n_sub && !sizes) || (n_sub > 1 && !offset_minus1))
    return badArg(h, "cabac_entry_point_syntax");
  DeviceGuard g(h->device);
  uint32_t *dS, *dE, *dO, *dL;
  Arena a(h);
  a.want((void **)&dS, n_sub * 4);
  a.want((void **)&dE, emulations ? n_sub * 4 : 0);
  a.want((void **)&dO, n_sub * 4);
  a.want((void **)&dL, 4);
  CK(a.commit(), "scratch alloc");
  H2D(dS, sizes, n_sub * 4);
  H2D(dE, emulations, n_sub * 4);
  int rc = cabac_entry_point_syntax_dev(h, dS, emulations ? dE : nullptr, n_sub, dO, dL);
  if (rc) return rc;
  if (n_sub > 1) D2H(offset_minus1, dO, (n_sub - 1) * 4);
  D2H(offset_len_minus1, dL, 4);
  CK(cudaStreamSynchronize(h->stream), "sync");
  return CABAC_OK;
}

int cabac_substream_ranges_dev(cabac_handle *h, const uint32_t *offset_minus1, size_t n_entry,
                               const uint32_t *epb_locations, size_t n_epb, uint64_t slice_data_pos,
                               uint64_t rbsp_size, uint64_t *in_offsets) {
  if (!h || !in_offsets || (n_entry && !offset_minus1) || (n_epb && !epb_locations))
    return badArg(h, "cabac_substream_ranges_dev");
  DeviceGuard g(h->device);
  CK(cabac::launchSubstreamRanges(h->stream, offset_minus1, n_entry, epb_locations, n_epb, slice_data_pos, rbsp_size,
                                  in_offsets),
     "substream_ranges_kernel");
  h->launches += 1;
  return CABAC_OK;
}

int cabac_substream_ranges(cabac_handle *h, const uint32_t *offset_minus1, size_t n_entry,
                           const uint32_t *epb_locations, size_t n_epb, uint64_t slice_data_pos, uint64_t rbsp_size,
                           uint64_t *in_offsets) {
  if (!h || !in_offsets || (n_entry && !offset_minus1) || (n_epb && !epb_locations))
    return badArg(h, "cabac_substream_ranges");
  DeviceGuard g(h->device);
  uint32_t *dO, *dE;
  uint64_t *dR;
  Arena a(h);
  a.want((void **)&dO, n_entry * 4);
  a.want((void **)&dE, n_epb * 4);
  a.want((void **)&dR, (n_entry + 2) * 8);
  CK(a.commit(), "scratch alloc");
  H2D(dO, offset_minus1, n_entry * 4);
  H2D(dE, epb_locations, n_epb * 4);
  int rc = cabac_substream_ranges_dev(h, dO, n_entry, dE, n_epb, slice_data_pos, rbsp_size, dR);
  if (rc) return rc;
  D2H(in_offsets, dR, (n_entry + 2) * 8);
  CK(cudaStreamSynchronize(h->stream), "sync");
  return CABAC_OK;
}

// ---- WPP entry points: the plain calls with the sync records attached to the launch ----
int cabac_encode_residual_wpp_dev(cabac_handle *h, const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                                  const cabac_substream_params *params, const cabac_wpp_sync *wpp, size_t n_sub,
                                  const int32_t *coeffs, const cabac_ctx_model *ctx_in, uint8_t *out,
                                  const uint64_t *out_offsets, uint32_t *out_sizes, uint32_t *num_bins,
                                  uint8_t *tu_flags, cabac_ctx_model *ctx_out, uint32_t *status) {
  if (!h || (n_sub && !wpp)) return badArg(h, "cabac_encode_residual_wpp_dev");
  h->wppNext = n_sub ? wpp : nullptr;
  const int rc = cabac_encode_residual_dev(h, tus, tu_offsets, params, n_sub, coeffs, ctx_in, out, out_offsets,
                                           out_sizes, num_bins, tu_flags, ctx_out, status);
  h->wppNext = nullptr;
  return rc;
}

int cabac_decode_residual_wpp_dev(cabac_handle *h, const uint8_t *in, const uint64_t *in_offsets,
                                  const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                                  const cabac_substream_params *params, const cabac_wpp_sync *wpp, size_t n_sub,
                                  const cabac_ctx_model *ctx_in, int32_t *coeffs, uint32_t *num_bins,
                                  uint32_t *bits_read, uint8_t *tu_flags, cabac_ctx_model *ctx_out,
                                  uint32_t *status) {
  if (!h || (n_sub && !wpp)) return badArg(h, "cabac_decode_residual_wpp_dev");
  h->wppNext = n_sub ? wpp : nullptr;
  const int rc = cabac_decode_residual_dev(h, in, in_offsets, tus, tu_offsets, params, n_sub, ctx_in, coeffs, num_bins,
                                           bits_read, tu_flags, ctx_out, status);
  h->wppNext = nullptr;
  return rc;
}

int cabac_encode_stream_wpp_dev(cabac_handle *h, const uint32_t *cmds, const uint64_t *cmd_offsets,
                                const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                                const cabac_substream_params *params, const cabac_wpp_sync *wpp, size_t n_sub,
                                const int32_t *coeffs, const cabac_ctx_model *ctx_in, uint8_t *out,
                                const uint64_t *out_offsets, uint32_t *out_sizes, uint32_t *num_bins,
                                uint8_t *tu_flags, cabac_ctx_model *ctx_out, uint32_t *status) {
  if (!h || (n_sub && !wpp)) return badArg(h, "cabac_encode_stream_wpp_dev");
  h->wppNext = n_sub ? wpp : nullptr;
  const int rc = cabac_encode_stream_dev(h, cmds, cmd_offsets, tus, tu_offsets, params, n_sub, coeffs, ctx_in, out,
                                         out_offsets, out_sizes, num_bins, tu_flags, ctx_out, status);
  h->wppNext = nullptr;
  return rc;
}

int cabac_encode_residual_wpp(cabac_handle *h, const cabac_tu_desc *tus, const uint64_t *tu_offsets,
                              const cabac_substream_params *params, const cabac_wpp_sync *wpp, size_t n_sub,
                              const int32_t *coeffs, size_t n_coeffs, const cabac_ctx_model *ctx_in,
                              const uint64_t *out_offsets, uint8_t *packed, size_t packed_cap,
                              uint64_t *entry_points, uint32_t *out_sizes, uint32_t *num_bins, uint8_t *tu_flags,
                              cabac_ctx_model *ctx_out, uint32_t *status) {
  if (!h || (n_sub && !wpp)) return badArg(h, "cabac_encode_residual_wpp");
  if (n_sub && !wppValid(wpp, n_sub)) return badArg(h, "cabac_encode_residual_wpp: dep must be < the substream's index");
  if (n_sub) {
    DeviceGuard g(h->device);
    CK(stageWpp(h, wpp, n_sub), "wpp sync records");
  }
  const int rc = cabac_encode_residual_packed(h, tus, tu_offsets, params, n_sub, coeffs, n_coeffs, ctx_in, out_offsets,
                                              packed, packed_cap, entry_points, out_sizes, num_bins, tu_flags, ctx_out,
                                              status);
  h->wppNext = nullptr;
  return rc;
}

int cabac_decode_residual_wpp(cabac_handle *h, const uint8_t *in, const uint64_t *in_offsets, const cabac_tu_desc *tus,
                              const uint64_t *tu_offsets, const cabac_substream_params *params,
                              const cabac_wpp_sync *wpp, size_t n_sub, const cabac_ctx_model *ctx_in, int32_t *coeffs,
                              size_t n_coeffs, uint32_t *num_bins, uint32_t *bits_read, uint8_t *tu_flags,
                              cabac_ctx_model *ctx_out, uint32_t *status) {
  if (!h || (n_sub && !wpp)) return badArg(h, "cabac_decode_residual_wpp");
  if (n_sub && !wppValid(wpp, n_sub)) return badArg(h, "cabac_decode_residual_wpp: dep must be < the substream's index");
  if (n_sub) {
    DeviceGuard g(h->device);
    CK(stageWpp(h, wpp, n_sub), "wpp sync records");
  }
  const int rc = cabac_decode_residual(h, in, in_offsets, tus, tu_offsets, params, n_sub, ctx_in, coeffs, n_coeffs,
                                       num_bins, bits_read, tu_flags, ctx_out, status);
  h->wppNext = nullptr;
  return rc;
}

int cabac_encode_stream_wpp(cabac_handle *h, const uint32_t *cmds, const uint64_t *cmd_offsets, const cabac_tu_desc *tus,
                            const uint64_t *tu_offsets, const cabac_substream_params *params,
                            const cabac_wpp_sync *wpp, size_t n_sub, const int32_t *coeffs, size_t n_coeffs,
                            const cabac_ctx_model *ctx_in, uint8_t *out, const uint64_t *out_offsets,
                            uint32_t *out_sizes, uint32_t *num_bins, uint8_t *tu_flags, cabac_ctx_model *ctx_out,
                            uint32_t *status) {
  if (!h || (n_sub && !wpp)) return badArg(h, "cabac_encode_stream_wpp");
  if (n_sub && !wppValid(wpp, n_sub)) return badArg(h, "cabac_encode_stream_wpp: dep must be < the substream's index");
  if (n_sub) {
    DeviceGuard g(h->device);
    CK(stageWpp(h, wpp, n_sub), "wpp sync records");
  }
  const int rc = cabac_encode_stream(h, cmds, cmd_offsets, tus, tu_offsets, params, n_sub, coeffs, n_coeffs, ctx_in, out,
                                     out_offsets, out_sizes, num_bins, tu_flags, ctx_out, status);
  h->wppNext = nullptr;
  return rc;
}

}  // extern "C"
