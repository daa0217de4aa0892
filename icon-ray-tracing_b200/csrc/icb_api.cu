// C ABI (include/icon_b200.h) on top of the host grid builder and the sm_100a kernels.
// There is no CPU fallback anywhere in this file: without a CUDA device icb_create fails.
#include <chrono>
#include <cstdlib>
#include <cmath>
#include <cstddef>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "icb_context.h"

static std::string g_createError;
#define SL(h) ((h)->slots[(h)->cur])

#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      h->err = std::string(#call) + ": " + cudaGetErrorString(e_);                                 \
      return ICB_ERR_CUDA;                                                                         \
    }                                                                                              \
  } while (0)

void icbFreeGrid(icb_context *h)
{
  cudaFree(h->dRec); cudaFree(h->dLines); cudaFree(h->dLinesAlt); cudaFree(h->dOrigToInt);
  cudaFree(h->dCol); cudaFree(h->dBinOff); cudaFree(h->dBinList); cudaFree(h->dStage); cudaFree(h->dGlobal); cudaFree(h->dSeed);
  cudaFree(h->dVtxOff); cudaFree(h->dVtxCols); cudaFree(h->dColLayers); cudaFree(h->dVtxScalars);
  h->dVtxOff = h->dVtxCols = h->dColLayers = nullptr; h->dVtxScalars = nullptr;
  cudaFree(h->dColTrig); h->dColTrig = nullptr;
  cudaFree(h->dPrismMask); h->dPrismMask = nullptr; h->grid.prismMask = nullptr; h->maskDirty = true;
  h->colTrigHost.clear(); h->colTrigHost.shrink_to_fit();
  h->majorDirty = h->shellDirty = 2;
  h->grid.vtxScalars = nullptr; h->grid.sampler = 0;
  h->dRec = nullptr; h->dLines = nullptr; h->dLinesAlt = nullptr; h->linesBuf = 0; h->dOrigToInt = nullptr;
  for (FrameSlot &sl : h->slots) sl.bufUsed[0] = sl.bufUsed[1] = false;
  h->dCol = nullptr; h->dBinOff = h->dBinList = nullptr; h->dStage = nullptr; h->dGlobal = nullptr; h->dSeed = nullptr;
  h->haveGrid = false;
}

template <typename T, typename A>
static cudaError_t upload(T *&dst, const std::vector<T, A> &src)
{
  cudaError_t e = cudaMalloc((void **)&dst, std::max<size_t>(1, src.size()) * sizeof(T));
  if (e != cudaSuccess) return e;
  return cudaMemcpy(dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice);
}

// Host arrays go up through two pinned 128 MB buffers: `fill(first, last, dst)` writes elements [first, last) of the array
// into the pinned buffer with all host threads while the previous buffer is on its way (a pageable cudaMemcpy stages through
// one thread).  The records' lines (256 bytes each, internal order) are gathered that way and never exist as a whole in
// host memory; the nodes (128 bytes each) are a parallel memcpy.
// behind a frame's (or a locate call's) kernel: the slot's end event, and the event of the lines copy the kernel read
static cudaError_t frameDone(icb_context *h, cudaStream_t stream)
{
  FrameSlot &sl = h->slots[h->cur];
  cudaError_t e = cudaEventRecord(sl.ev1, stream);
  if (e != cudaSuccess) return e;
  sl.bufUsed[h->linesBuf] = true;
  return cudaEventRecord(sl.evBuf[h->linesBuf], stream);
}

struct Stager {
  static constexpr size_t kBytes = (size_t)128 << 20;
  void *pin[2] = {nullptr, nullptr};
  cudaEvent_t done[2] = {nullptr, nullptr};
  int uses = 0;
  cudaError_t init()
  {
    cudaError_t e = cudaSuccess;
    for (int b = 0; b < 2 && e == cudaSuccess; ++b) {
      e = cudaMallocHost(&pin[b], kBytes);
      if (e == cudaSuccess) e = cudaEventCreateWithFlags(&done[b], cudaEventDisableTiming);
    }
    return e;
  }
  ~Stager() { for (int b = 0; b < 2; ++b) { if (pin[b]) cudaFreeHost(pin[b]); if (done[b]) cudaEventDestroy(done[b]); } }
};

template <typename F>
static int uploadStaged(icb_context *h, Stager &st, void *dst, size_t count, size_t elemBytes, F fill)
{
  if (count == 0) return ICB_OK;
  const size_t chunk = std::max<size_t>(1, Stager::kBytes / elemBytes);
  cudaError_t e = cudaSuccess;
  for (size_t a = 0; a < count && e == cudaSuccess; a += chunk) {
    const int b = st.uses & 1;
    const size_t end = std::min(count, a + chunk);
    if (st.uses >= 2) e = cudaEventSynchronize(st.done[b]);
    if (e != cudaSuccess) break;
    ++st.uses;
    fill(a, end, st.pin[b]);
    e = cudaMemcpyAsync((char *)dst + a * elemBytes, st.pin[b], (end - a) * elemBytes, cudaMemcpyHostToDevice, h->ownStream);
    if (e == cudaSuccess) e = cudaEventRecord(st.done[b], h->ownStream);
  }
  if (e != cudaSuccess) { h->err = cudaGetErrorString(e); return ICB_ERR_CUDA; }
  return ICB_OK;
}

static int uploadLines(icb_context *h, Stager &st, const void *records, const RawVec<int32_t> &origToInt)
{
  const size_t n = origToInt.size();
  CK(cudaMalloc((void **)&h->dLines, std::max<size_t>(1, n) * kLineStride * sizeof(float)));
  if (n == 0) return ICB_OK;
  RawVec<int32_t> intToOrig;
  invertPermutation(origToInt.data(), n, intToOrig);
  int rc = uploadStaged(h, st, h->dLines, n, kLineStride * sizeof(float),
                        [&](size_t a, size_t end, void *dst) { fillLines(records, intToOrig.data(), a, end, (float *)dst); });
  if (rc != ICB_OK) return rc;
  CK(cudaStreamSynchronize(h->ownStream));     // (intToOrig goes out of scope)
  return ICB_OK;
}

template <typename T, typename A>
static int uploadBig(icb_context *h, Stager &st, T *&dst, const std::vector<T, A> &src)
{
  CK(cudaMalloc((void **)&dst, std::max<size_t>(1, src.size()) * sizeof(T)));
  return uploadStaged(h, st, dst, src.size(), sizeof(T),
                      [&](size_t a, size_t end, void *p) { parallelCopy(p, src.data() + a, (end - a) * sizeof(T)); });
}

extern "C" {

const char *icb_version(void) { return "icon-b200 0.1 (sm_100a)"; }

const char *icb_last_error(icb_handle h) { return h ? h->err.c_str() : g_createError.c_str(); }

int icb_create(int device, icb_handle *out)
{
  if (!out) return ICB_ERR_INVALID;
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count <= 0) {
    g_createError = std::string("icb_create: no CUDA device (") + cudaGetErrorString(e) + "); this library has no CPU path";
    return ICB_ERR_NO_DEVICE;
  }
  if (device < 0 || device >= count) { g_createError = "icb_create: device index out of range"; return ICB_ERR_INVALID; }
  icb_context *h = new icb_context();
  h->device = device;
  auto fail = [&](const char *what, cudaError_t err) {
    g_createError = std::string("icb_create: ") + what + ": " + cudaGetErrorString(err);
    delete h;
    return ICB_ERR_CUDA;
  };
  if ((e = cudaSetDevice(device)) != cudaSuccess) return fail("cudaSetDevice", e);
  cudaDeviceProp prop;
  if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return fail("cudaGetDeviceProperties", e);
  h->numSMs = prop.multiProcessorCount;
  // experiment knob: how much L2 fetches from DRAM around a missing sector (32, 64 or 128 bytes; the driver's default is 64).
  // The record fetches are 80-byte and 256-byte reads at random addresses; on R2B10 (DRAM-bound) the granularity
  // decides how many bytes a fetch costs.
  if (const char *v = getenv("ICB_L2_FETCH")) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)std::max(32, std::min(128, atoi(v))));
  if (!getenv("ICB_NO_L2_PERSIST") && prop.persistingL2CacheMaxSize >= (64 << 20) && prop.accessPolicyMaxWindowSize >= (64 << 20)) {
    h->l2Persist = true;
    h->l2WindowMax = (size_t)prop.accessPolicyMaxWindowSize;
  }
  if ((e = configureKernels()) != cudaSuccess) return fail("kernel image for this device (built for sm_100a only)", e);
  h->blocksPerSM = renderBlocksPerSM(300);
  if ((e = cudaStreamCreateWithFlags(&h->ownStream, cudaStreamNonBlocking)) != cudaSuccess) return fail("cudaStreamCreate", e);
  for (FrameSlot &sl : h->slots) {
    if ((e = cudaEventCreate(&sl.ev0)) != cudaSuccess) return fail("cudaEventCreate", e);
    if ((e = cudaEventCreate(&sl.ev1)) != cudaSuccess) return fail("cudaEventCreate", e);
    for (int b = 0; b < 2; ++b)
      if ((e = cudaEventCreateWithFlags(&sl.evBuf[b], cudaEventDisableTiming)) != cudaSuccess) return fail("cudaEventCreate", e);
    if ((e = cudaStreamCreateWithFlags(&sl.stream, cudaStreamNonBlocking)) != cudaSuccess) return fail("cudaStreamCreate", e);
    if ((e = cudaMalloc((void **)&sl.dTileCounter, 2 * sizeof(int))) != cudaSuccess) return fail("cudaMalloc", e);
    if ((e = cudaMalloc((void **)&sl.dNHeavy, sizeof(int))) != cudaSuccess) return fail("cudaMalloc", e);
    cudaMemset(sl.dNHeavy, 0, sizeof(int));
    if ((e = cudaMalloc((void **)&sl.dCounters, sizeof(Counters))) != cudaSuccess) return fail("cudaMalloc", e);
    cudaMemset(sl.dCounters, 0, sizeof(Counters));
  }
  // defaults (hostCode.cu:924-925)
  h->rp.ambient[0] = h->rp.ambient[1] = h->rp.ambient[2] = 1.f;
  h->rp.ambientRadiance = 1.f;
  h->rp.termAlpha = 0.99f;
  h->rp.opacityScale = 1.f;
  // developer knobs (profiling sweeps; the defaults are the tuned values)
  if (const char *v = getenv("ICB_REFILL_MIN")) h->refillMin = std::max(1, std::min(32, atoi(v)));
  if (const char *v = getenv("ICB_SKIP_BACK")) h->skipBack = std::max(0, std::min(4, atoi(v)));
  if (const char *v = getenv("ICB_BLOCKS_PER_SM")) h->blocksOverride = std::max(0, std::min(8, atoi(v)));
  if (const char *v = getenv("ICB_MIN_BLOCKS")) h->minBlocks = std::max(0, std::min(4096, atoi(v)));
  if (const char *v = getenv("ICB_SPLIT_HEAVY")) h->splitMaxHeavy = std::max(0, std::min(4096, atoi(v)));
  if (const char *v = getenv("ICB_SPLIT_MODE")) h->splitMode = std::max(0, std::min(2, atoi(v)));
  if (const char *v = getenv("ICB_SPLIT_COST_LOG2")) h->splitMinCostLog2 = std::max(0, std::min(31, atoi(v)));
  if (const char *v = getenv("ICB_CHUNK_LOG2")) h->chunkLog2 = std::max(4, std::min(20, atoi(v)));
  *out = h;
  return ICB_OK;
}

int icb_destroy(icb_handle h)
{
  if (!h) return ICB_ERR_INVALID;
  cudaSetDevice(h->device);
  icbFreeGrid(h);
  for (FrameSlot &sl : h->slots) { cudaFree(sl.dFrameFb); cudaFree(sl.dFrameAccum); if (sl.stream) cudaStreamDestroy(sl.stream); }
  cudaFree(h->dLogTable); cudaFree(h->dNz);
  if (h->maskEvent) cudaEventDestroy(h->maskEvent);
  if (h->majorEvent) cudaEventDestroy(h->majorEvent);
  cudaFree(h->dRanges); cudaFree(h->dMaxOp);
  if (h->shellEvent) cudaEventDestroy(h->shellEvent);
  cudaFree(h->dShellRanges); cudaFree(h->dShellMaxOp);
  cudaFree(h->dLut);
  for (FrameSlot &sl : h->slots) {
    cudaFree(sl.dTileList); cudaFree(sl.dTileCounter); cudaFree(sl.dTileCost); cudaFree(sl.dOrder); cudaFree(sl.dCounters);
    cudaFree(sl.dNHeavy); cudaFree(sl.dCancel); cudaFree(sl.dChunks);
    if (sl.ev0) cudaEventDestroy(sl.ev0);
    if (sl.ev1) cudaEventDestroy(sl.ev1);
    for (int b = 0; b < 2; ++b) if (sl.evBuf[b]) cudaEventDestroy(sl.evBuf[b]);
  }
  if (h->ownStream) cudaStreamDestroy(h->ownStream);
  delete h;
  return ICB_OK;
}

int icb_set_frames_in_flight(icb_handle h, int n)
{
  if (!h || n < 1 || n > kFrameSlots) return ICB_ERR_INVALID;
  h->framesInFlight = n;
  return ICB_OK;
}

int icb_set_frame_slot(icb_handle h, int slot)
{
  if (!h || slot < 0 || slot >= kFrameSlots) return ICB_ERR_INVALID;
  h->cur = slot;
  return ICB_OK;
}

int icb_set_locator(icb_handle h, float marginMeters, int maxWalkSteps)
{
  if (!h) return ICB_ERR_INVALID;
  if (marginMeters >= 0.f) h->marginMeters = marginMeters;   // takes effect at the next icb_set_grid
  if (maxWalkSteps >= 0) { h->maxWalk = maxWalkSteps; h->grid.maxWalk = maxWalkSteps; }
  return ICB_OK;
}

int icb_set_skip_ahead(icb_handle h, int enabled)
{
  if (!h) return ICB_ERR_INVALID;
  h->skipAhead = enabled != 0;
  return ICB_OK;
}

int icb_set_ray_split(icb_handle h, int chunkSamplesLog2, int maxHeavyTiles, int minTileCostLog2, int mode)
{
  if (!h) return ICB_ERR_INVALID;
  if (mode >= 0) {
    if (mode > ICB_SPLIT_AUTO) { h->err = "icb_set_ray_split: unknown mode"; return ICB_ERR_INVALID; }
    h->splitMode = mode;
  }
  if (chunkSamplesLog2 >= 0) {
    if (chunkSamplesLog2 < 4 || chunkSamplesLog2 > 20) { h->err = "icb_set_ray_split: chunk size outside 2^4 .. 2^20 samples"; return ICB_ERR_INVALID; }
    h->chunkLog2 = chunkSamplesLog2;
  }
  if (maxHeavyTiles >= 0) h->splitMaxHeavy = std::min(maxHeavyTiles, 4096);
  if (minTileCostLog2 >= 0) h->splitMinCostLog2 = std::min(minTileCostLog2, 31);
  return ICB_OK;
}

int icb_set_integrator(icb_handle h, int integrator)
{
  if (!h) return ICB_ERR_INVALID;
  if (integrator != ICB_INTEGRATOR_MARCH && integrator != ICB_INTEGRATOR_WOODCOCK && integrator != ICB_INTEGRATOR_WOODCOCK_GRID &&
      integrator != ICB_INTEGRATOR_WOODCOCK_SPHERE) {
    h->err = "icb_set_integrator: unknown integrator";
    return ICB_ERR_INVALID;
  }
  // logf(1-u) of the free flights: computed on the device with glibc's operation sequence (glibc_logf.cuh, the FMA build's
  // form; dLogTable stays null) -- but only after that form has reproduced THIS host's libm for every argument a free flight
  // can take (2^24 of them: one host table, one check kernel, once per handle, ~0.2 s).  Where it does not (a libm that takes
  // the SSE2 ifunc path, another glibc), or with ICB_LOG_TABLE=1, the kernels read the host-built table as round 1 did:
  // "bit-identical to the reference's CPU path on the same host" then holds by construction, only slower.
  if (integrator != ICB_INTEGRATOR_MARCH && !h->dLogTable && (h->logfForm < 0 || getenv("ICB_LOG_TABLE"))) {
    CK(cudaSetDevice(h->device));
    std::vector<float> table;
    buildLogTable(table);
    float *dTable = nullptr;
    CK(upload(dTable, table));
    bool useTable = getenv("ICB_LOG_TABLE") != nullptr;
    if (!useTable) {
      unsigned long long *dBad = nullptr, bad[2] = {1, 1};
      cudaError_t e = cudaMalloc((void **)&dBad, sizeof(bad));
      if (e == cudaSuccess) e = cudaMemsetAsync(dBad, 0, sizeof(bad), h->ownStream);
      if (e == cudaSuccess) e = launchLogfCheck(dTable, dBad, h->ownStream);
      if (e == cudaSuccess) e = cudaMemcpyAsync(bad, dBad, sizeof(bad), cudaMemcpyDeviceToHost, h->ownStream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(h->ownStream);
      cudaFree(dBad);
      if (e != cudaSuccess) { cudaFree(dTable); h->err = std::string("icb_set_integrator: logf check: ") + cudaGetErrorString(e); return ICB_ERR_CUDA; }
      useTable = bad[0] != 0 || getenv("ICB_LOG_PRETEND_MISMATCH") != nullptr;     // (the second: a test hook for the fallback)
      h->logfForm = useTable ? 2 : 0;
    } else h->logfForm = 2;
    if (useTable) {
      h->dLogTable = dTable;
      if (h->l2Persist && cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)72 << 20) != cudaSuccess) { (void)cudaGetLastError(); h->l2Persist = false; }
    } else cudaFree(dTable);
  }
  h->integrator = integrator;
  return ICB_OK;
}

// (re)computes the per-vertex interface scalars from the values now in lines[]; allocates the table on first use
static int refreshVertexScalars(icb_context *h, cudaStream_t stream)
{
  if (!h->dVtxScalars) {
    const size_t bytes = h->nVert * (size_t)h->vtxStride * sizeof(float);
    if (bytes > ((size_t)64 << 30)) { h->err = "icb_set_sampler: vertex-scalar table above 64 GB (a column with very many layers?)"; return ICB_ERR_INVALID; }
    CK(cudaMalloc((void **)&h->dVtxScalars, std::max<size_t>(bytes, 4)));
    h->stats.deviceBytes += (int64_t)bytes;
  }
  h->grid.vtxScalars = h->dVtxScalars;
  h->grid.vtxStride = h->vtxStride;
  CK(launchVertexScalars(h->grid, h->dColLayers, h->dVtxOff, h->dVtxCols, h->nVert, h->dVtxScalars, stream));
  h->maskDirty = true;
  return ICB_OK;
}

// Woodcock ray-gens: the 64 MB logf table is gathered at random (one entry per tracking step); pin it in the 126 MB L2
// for the duration of the launch (access-policy window on the launching stream, best effort).
static void logTableWindow(icb_context *h, cudaStream_t stream, bool on)
{
  if (!h->l2Persist || !h->dLogTable) return;
  cudaStreamAttrValue attr;
  memset(&attr, 0, sizeof(attr));
  attr.accessPolicyWindow.base_ptr = h->dLogTable;
  attr.accessPolicyWindow.num_bytes = on ? std::min<size_t>((size_t)64 << 20, h->l2WindowMax) : 0;
  attr.accessPolicyWindow.hitRatio = 1.f;
  attr.accessPolicyWindow.hitProp = on ? cudaAccessPropertyPersisting : cudaAccessPropertyNormal;
  attr.accessPolicyWindow.missProp = on ? cudaAccessPropertyStreaming : cudaAccessPropertyNormal;
  if (cudaStreamSetAttribute(stream, cudaStreamAttributeAccessPolicyWindow, &attr) != cudaSuccess) (void)cudaGetLastError();
}

static MajorantGrid majorantGrid(const icb_context *h)
{
  MajorantGrid mg;
  mg.maxOpacities = h->dMaxOp;
  // (the occupancy bits live behind the majorants in the same allocation)
  mg.occupied = h->dMaxOp ? reinterpret_cast<const uint32_t *>(h->dMaxOp + (size_t)h->accelDims[0] * h->accelDims[1] * h->accelDims[2]) : nullptr;
  for (int k = 0; k < 3; ++k) { mg.dims[k] = h->accelDims[k]; mg.lo[k] = h->bounds.volLo[k]; mg.hi[k] = h->bounds.volHi[k]; }
  return mg;
}

// Tables shared by the frames in flight (majorant grid, prism masks) are rebuilt on the stream of the frame that finds
// them stale: that stream first waits for every frame still in flight on another slot's stream (it may be reading the
// old tables), and every later frame on another stream waits for the event recorded behind the rebuild.
static int waitFramesInFlight(icb_context *h, cudaStream_t stream)
{
  for (FrameSlot &sl : h->slots)
    if (sl.used) CK(cudaStreamWaitEvent(stream, sl.ev1, 0));
  return ICB_OK;
}

// brings the majorant grid up to date with the grid's scalars and the transfer function, on `stream`
static int ensureMajorants(icb_context *h, cudaStream_t stream)
{
  if (!h->majorEvent) CK(cudaEventCreateWithFlags(&h->majorEvent, cudaEventDisableTiming));
  const size_t numMC = (size_t)h->accelDims[0] * h->accelDims[1] * h->accelDims[2];
  if (!h->dColTrig) {
    if (h->colTrigHost.size() != h->nCol * kColTrigStride) { h->err = "majorant grid: column table missing"; return ICB_ERR_INVALID; }
    CK(upload(h->dColTrig, h->colTrigHost));
    h->colTrigHost.clear(); h->colTrigHost.shrink_to_fit();
  }
  if (!h->dRanges) { CK(cudaMalloc((void **)&h->dRanges, numMC * 2 * sizeof(float))); h->majorDirty = 2; }
  if (!h->dMaxOp) { CK(cudaMalloc((void **)&h->dMaxOp, (numMC + (numMC + 31) / 32) * sizeof(float))); h->majorDirty = std::max(h->majorDirty, 1); }
  const MajorantGrid mg = majorantGrid(h);
  if (h->majorDirty == 0) {
    CK(cudaStreamWaitEvent(stream, h->majorEvent, 0));
    return ICB_OK;
  }
  int rc = waitFramesInFlight(h, stream);
  if (rc != ICB_OK) return rc;
  if (h->majorDirty >= 2) CK(launchGridRanges(h->grid, h->dColTrig, mg, h->dRanges, stream));
  if (h->majorDirty >= 1) CK(launchMaxOpacities(h->dRanges, mg, h->dLut, h->lutSize, h->rp.tfLo, h->rp.tfHi, h->dMaxOp, const_cast<uint32_t *>(mg.occupied), stream));
  CK(cudaEventRecord(h->majorEvent, stream));
  h->majorDirty = 0;
  return ICB_OK;
}

int icb_set_accel_grid(icb_handle h, int nx, int ny, int nz)
{
  if (!h) return ICB_ERR_INVALID;
  if (nx < 1 || ny < 1 || nz < 1 || (int64_t)nx * ny * nz > ((int64_t)1 << 30)) { h->err = "icb_set_accel_grid: bad dimensions"; return ICB_ERR_INVALID; }
  CK(cudaSetDevice(h->device));
  CK(cudaDeviceSynchronize());
  cudaFree(h->dRanges); cudaFree(h->dMaxOp);
  h->dRanges = h->dMaxOp = nullptr;
  h->accelDims[0] = nx; h->accelDims[1] = ny; h->accelDims[2] = nz;
  h->majorDirty = 2;
  return ICB_OK;
}

int icb_get_majorants(icb_handle h, float *maxOpacities_host, float *valueRanges_host, size_t numMacrocells)
{
  if (!h || !maxOpacities_host) return ICB_ERR_INVALID;
  if (!h->haveGrid) { h->err = "icb_get_majorants: no grid"; return ICB_ERR_NO_GRID; }
  if (!h->dLut) { h->err = "icb_get_majorants: no transfer function"; return ICB_ERR_NO_TRANSFUNC; }
  const size_t numMC = (size_t)h->accelDims[0] * h->accelDims[1] * h->accelDims[2];
  if (numMacrocells != numMC) { h->err = "icb_get_majorants: macrocell count differs from the accel grid"; return ICB_ERR_INVALID; }
  CK(cudaSetDevice(h->device));
  int rc = ensureMajorants(h, h->ownStream);
  if (rc != ICB_OK) return rc;
  CK(cudaStreamSynchronize(h->ownStream));
  CK(cudaMemcpy(maxOpacities_host, h->dMaxOp, numMC * sizeof(float), cudaMemcpyDeviceToHost));
  if (valueRanges_host) CK(cudaMemcpy(valueRanges_host, h->dRanges, numMC * 2 * sizeof(float), cudaMemcpyDeviceToHost));
  return ICB_OK;
}

// ---- the (rad, lat, lon) shell accel of ICB_INTEGRATOR_WOODCOCK_SPHERE (hostCode.cu:652-666 buildShellAccel) ----
static MajorantGrid shellGrid(const icb_context *h)
{
  MajorantGrid sg;
  sg.maxOpacities = h->dShellMaxOp;
  sg.occupied = h->dShellMaxOp ? reinterpret_cast<const uint32_t *>(h->dShellMaxOp + (size_t)h->shellDims[0] * h->shellDims[1] * h->shellDims[2]) : nullptr;
  for (int k = 0; k < 3; ++k) { sg.dims[k] = h->shellDims[k]; sg.lo[k] = h->bounds.sphLo[k]; sg.hi[k] = h->bounds.sphHi[k]; }
  return sg;
}

static int ensureShellMajorants(icb_context *h, cudaStream_t stream)
{
  if (!h->shellEvent) CK(cudaEventCreateWithFlags(&h->shellEvent, cudaEventDisableTiming));
  const size_t numMC = (size_t)h->shellDims[0] * h->shellDims[1] * h->shellDims[2];
  if (!h->dColTrig) {
    if (h->colTrigHost.size() != h->nCol * kColTrigStride) { h->err = "shell accel: column table missing"; return ICB_ERR_INVALID; }
    CK(upload(h->dColTrig, h->colTrigHost));
    h->colTrigHost.clear(); h->colTrigHost.shrink_to_fit();
  }
  if (!h->dShellRanges) { CK(cudaMalloc((void **)&h->dShellRanges, numMC * 2 * sizeof(float))); h->shellDirty = 2; }
  if (!h->dShellMaxOp) { CK(cudaMalloc((void **)&h->dShellMaxOp, (numMC + (numMC + 31) / 32) * sizeof(float))); h->shellDirty = std::max(h->shellDirty, 1); }
  const MajorantGrid sg = shellGrid(h);
  if (h->shellDirty == 0) {
    CK(cudaStreamWaitEvent(stream, h->shellEvent, 0));
    return ICB_OK;
  }
  int rc = waitFramesInFlight(h, stream);
  if (rc != ICB_OK) return rc;
  if (h->shellDirty >= 2) CK(launchShellRanges(h->grid, h->dColTrig, sg, h->dShellRanges, stream));
  if (h->shellDirty >= 1) CK(launchMaxOpacities(h->dShellRanges, sg, h->dLut, h->lutSize, h->rp.tfLo, h->rp.tfHi, h->dShellMaxOp, const_cast<uint32_t *>(sg.occupied), stream));
  CK(cudaEventRecord(h->shellEvent, stream));
  h->shellDirty = 0;
  return ICB_OK;
}

int icb_set_accel_shell(icb_handle h, int nRad, int nLat, int nLon)
{
  if (!h) return ICB_ERR_INVALID;
  if (nRad < 1 || nLat < 1 || nLon < 1 || (int64_t)nRad * nLat * nLon > ((int64_t)1 << 28)) { h->err = "icb_set_accel_shell: bad dimensions"; return ICB_ERR_INVALID; }
  CK(cudaSetDevice(h->device));
  CK(cudaDeviceSynchronize());
  cudaFree(h->dShellRanges); cudaFree(h->dShellMaxOp);
  h->dShellRanges = h->dShellMaxOp = nullptr;
  h->shellDims[0] = nRad; h->shellDims[1] = nLat; h->shellDims[2] = nLon;
  h->shellDirty = 2;
  return ICB_OK;
}

int icb_get_shell_majorants(icb_handle h, float *maxOpacities_host, float *valueRanges_host, size_t numCells)
{
  if (!h || !maxOpacities_host) return ICB_ERR_INVALID;
  if (!h->haveGrid) { h->err = "icb_get_shell_majorants: no grid"; return ICB_ERR_NO_GRID; }
  if (!h->dLut) { h->err = "icb_get_shell_majorants: no transfer function"; return ICB_ERR_NO_TRANSFUNC; }
  const size_t numMC = (size_t)h->shellDims[0] * h->shellDims[1] * h->shellDims[2];
  if (numCells != numMC) { h->err = "icb_get_shell_majorants: cell count differs from the shell accel"; return ICB_ERR_INVALID; }
  CK(cudaSetDevice(h->device));
  int rc = ensureShellMajorants(h, h->ownStream);
  if (rc != ICB_OK) return rc;
  CK(cudaStreamSynchronize(h->ownStream));
  CK(cudaMemcpy(maxOpacities_host, h->dShellMaxOp, numMC * sizeof(float), cudaMemcpyDeviceToHost));
  if (valueRanges_host) CK(cudaMemcpy(valueRanges_host, h->dShellRanges, numMC * 2 * sizeof(float), cudaMemcpyDeviceToHost));
  return ICB_OK;
}

int icb_selftest_logf(icb_handle h, int64_t mismatches[2])
{
  if (!h || !mismatches) return ICB_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  std::vector<float> table;
  buildLogTable(table);
  float *dTable = nullptr;
  unsigned long long *dBad = nullptr;
  CK(upload(dTable, table));
  CK(cudaMalloc((void **)&dBad, 2 * sizeof(unsigned long long)));
  CK(cudaMemset(dBad, 0, 2 * sizeof(unsigned long long)));
  CK(launchLogfCheck(dTable, dBad, h->ownStream));
  unsigned long long bad[2] = {0, 0};
  CK(cudaMemcpyAsync(bad, dBad, sizeof(bad), cudaMemcpyDeviceToHost, h->ownStream));
  CK(cudaStreamSynchronize(h->ownStream));
  cudaFree(dTable); cudaFree(dBad);
  mismatches[0] = (int64_t)bad[0]; mismatches[1] = (int64_t)bad[1];
  return ICB_OK;
}

int icb_set_sampler(icb_handle h, int sampler)
{
  if (!h) return ICB_ERR_INVALID;
  if (sampler != ICB_SAMPLER_CELL && sampler != ICB_SAMPLER_PRISM) { h->err = "icb_set_sampler: unknown sampler"; return ICB_ERR_INVALID; }
  h->sampler = sampler;
  if (h->haveGrid) {
    CK(cudaSetDevice(h->device));
    if (sampler == ICB_SAMPLER_PRISM) {
      int rc = refreshVertexScalars(h, h->ownStream);
      if (rc != ICB_OK) { h->sampler = ICB_SAMPLER_CELL; return rc; }
      CK(cudaStreamSynchronize(h->ownStream));
    }
    h->grid.sampler = sampler;
    if (sampler == ICB_SAMPLER_CELL) h->grid.prismMask = nullptr;
    else h->maskDirty = true;
  }
  h->majorDirty = h->shellDirty = 2;       // the majorant grid's value ranges are those of the sampler (cell values | prism vertex scalars)
  return ICB_OK;
}

// (re)computes the per-vertex interface scalars -- defined below
static int refreshVertexScalars(icb_context *h, cudaStream_t stream);

extern "C++" int icbAdoptGrid(icb_context *h, const GridMeta &m, double setupSeconds)
{
  h->nRec = (size_t)m.nRec;
  h->nCol = (size_t)m.nCol;
  h->nVert = (size_t)m.nVert;
  h->vtxStride = m.maxColLayers + 2;      // interfaces 0..L of the tallest column, +1 so that reading (j, j+1) stays in the row
  DeviceGrid &g = h->grid;
  g.rec = h->dRec; g.lines = h->dLines; g.col = h->dCol;
  g.binOff = h->dBinOff; g.binList = h->dBinList; g.binN = m.binN; g.seed = h->dSeed;
  g.globalRecs = h->dGlobal; g.nGlobal = (int32_t)m.nGlobal; g.recOfOrig0 = m.recOfOrig0;
  g.nRec = (int32_t)h->nRec; g.nCol = (int32_t)h->nCol; g.maxWalk = h->maxWalk;
  g.cullLo = m.cullLo; g.cullHi = m.cullHi;
  h->bounds = m.bounds;
  h->marginUsed = m.marginUsed;
  for (int k = 0; k < 3; ++k) { h->rp.volLo[k] = m.bounds.volLo[k]; h->rp.volHi[k] = m.bounds.volHi[k]; }
  h->rp.rInner = m.bounds.sphLo[0];
  h->rp.rOuter = m.bounds.sphHi[0];
  if (!(h->rp.unitDistance > 0.f)) h->rp.unitDistance = m.bounds.unitDistance;
  if (!(h->rp.dt > 0.f)) h->rp.dt = h->rp.unitDistance / 2;
  icb_stats_t &st = h->stats;
  st = icb_stats_t{};
  st.numRecords = (int64_t)h->nRec; st.numColumns = (int64_t)h->nCol; st.numVertices = (int64_t)m.nVert;
  st.numBins = (int64_t)6 * m.binN * m.binN; st.numBinEntries = (int64_t)m.nBinList;
  st.numGenericColumns = m.numGeneric; st.numGlobalRecords = (int64_t)m.nGlobal;
  st.deviceBytes = (int64_t)(h->nRec * sizeof(RecBlock) + h->nRec * (size_t)kLineStride * 4 +
                             (m.nRec + m.nBinOff + m.nBinList + m.nSeed) * 4 +
                             m.nCol * sizeof(ColBlock) + (m.nVtxOff + m.nVtxCols + m.nColLayers) * 4);
  h->haveGrid = true;
  if (h->sampler == ICB_SAMPLER_PRISM) {
    int rc = refreshVertexScalars(h, h->ownStream);
    if (rc != ICB_OK) return rc;
    CK(cudaStreamSynchronize(h->ownStream));
    g.sampler = ICB_SAMPLER_PRISM;
  }
  st.setupSeconds = setupSeconds;
  return ICB_OK;
}

int icb_set_grid(icb_handle h, const void *records, size_t n)
{
  if (!h) return ICB_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  auto t0 = std::chrono::steady_clock::now();
  const bool timing = std::getenv("ICB_BUILD_TIMING") != nullptr;     // seconds per phase on stderr
  auto tPrev = t0;
  auto lap = [&](const char *what) {
    if (!timing) return;
    const auto t = std::chrono::steady_clock::now();
    std::fprintf(stderr, "icb_set_grid %-30s %.3f s\n", what, std::chrono::duration<double>(t - tPrev).count());
    tPrev = t;
  };
  HostGrid hg;
  int rc = buildGrid(records, n, h->marginMeters, hg, h->err, false);
  if (rc != ICB_OK) return rc;
  lap("buildGrid");
  CK(cudaDeviceSynchronize());
  icbFreeGrid(h);
  GridMeta m{};
  m.nRec = hg.rec.size(); m.nCol = hg.col.size(); m.nVert = (uint64_t)hg.numVertices;
  m.nBinOff = hg.binOff.size(); m.nBinList = hg.binList.size(); m.nGlobal = hg.globalRecs.size(); m.nSeed = hg.seed.size();
  m.nVtxOff = hg.vtxOff.size(); m.nVtxCols = hg.vtxCols.size(); m.nColLayers = hg.colLayers.size(); m.nColTrig = hg.colTrig.size();
  m.binN = hg.binN; m.recOfOrig0 = hg.origToInt[0]; m.maxColLayers = hg.maxColLayers;
  m.cullLo = hg.cullLo; m.cullHi = hg.cullHi; m.marginUsed = h->marginMeters;
  m.numGeneric = hg.numGeneric; m.bounds = hg.bounds;
  Stager st;
  CK(st.init());
  rc = uploadBig(h, st, h->dRec, hg.rec);
  if (rc != ICB_OK) return rc;
  CK(cudaStreamSynchronize(h->ownStream));
  RawVec<RecBlock>().swap(hg.rec);               // (32 GB at R2B10 x 90: not needed on the host any more)
  lap("nodes up");
  rc = uploadLines(h, st, records, hg.origToInt);
  if (rc != ICB_OK) return rc;
  lap("lines gathered and up");
  if ((rc = uploadBig(h, st, h->dOrigToInt, hg.origToInt)) != ICB_OK) return rc;
  if ((rc = uploadBig(h, st, h->dCol, hg.col)) != ICB_OK) return rc;
  if ((rc = uploadBig(h, st, h->dBinOff, hg.binOff)) != ICB_OK) return rc;
  if ((rc = uploadBig(h, st, h->dBinList, hg.binList)) != ICB_OK) return rc;
  if ((rc = uploadBig(h, st, h->dGlobal, hg.globalRecs)) != ICB_OK) return rc;
  if ((rc = uploadBig(h, st, h->dSeed, hg.seed)) != ICB_OK) return rc;
  if ((rc = uploadBig(h, st, h->dVtxOff, hg.vtxOff)) != ICB_OK) return rc;
  if ((rc = uploadBig(h, st, h->dVtxCols, hg.vtxCols)) != ICB_OK) return rc;
  if ((rc = uploadBig(h, st, h->dColLayers, hg.colLayers)) != ICB_OK) return rc;
  CK(cudaStreamSynchronize(h->ownStream));
  lap("tables up");
  h->colTrigHost = std::move(hg.colTrig);
  rc = icbAdoptGrid(h, m, std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
  lap("adopt");
  return rc;
}

int icb_set_scalars_dev(icb_handle h, const float *values_dev, size_t n, void *stream)
{
  if (!h || !values_dev) return ICB_ERR_INVALID;
  if (!h->haveGrid) { h->err = "icb_set_scalars: no grid"; return ICB_ERR_NO_GRID; }
  if (n != h->nRec) { h->err = "icb_set_scalars: record count differs from the grid"; return ICB_ERR_INVALID; }
  CK(cudaSetDevice(h->device));
  CK(launchScatterScalars(values_dev, h->dOrigToInt, h->linesBuf ? h->dLinesAlt : h->dLines, n, (cudaStream_t)stream));
  h->majorDirty = h->shellDirty = 2;
  if (h->sampler == ICB_SAMPLER_PRISM) return refreshVertexScalars(h, (cudaStream_t)stream);
  return ICB_OK;
}

int icb_set_scalars(icb_handle h, const float *values, size_t n)
{
  if (!h || !values) return ICB_ERR_INVALID;
  if (!h->haveGrid) { h->err = "icb_set_scalars: no grid"; return ICB_ERR_NO_GRID; }
  if (n != h->nRec) { h->err = "icb_set_scalars: record count differs from the grid"; return ICB_ERR_INVALID; }
  CK(cudaSetDevice(h->device));
  CK(cudaDeviceSynchronize());            // (frames in flight on other streams still read the old values)
  if (!h->dStage) CK(cudaMalloc((void **)&h->dStage, n * kHeightStride * sizeof(float)));
  CK(cudaMemcpyAsync(h->dStage, values, n * kHeightStride * sizeof(float), cudaMemcpyHostToDevice, h->ownStream));
  CK(launchScatterScalars(h->dStage, h->dOrigToInt, h->linesBuf ? h->dLinesAlt : h->dLines, n, h->ownStream));
  h->majorDirty = h->shellDirty = 2;
  if (h->sampler == ICB_SAMPLER_PRISM) {
    int rc = refreshVertexScalars(h, h->ownStream);
    if (rc != ICB_OK) return rc;
  }
  CK(cudaStreamSynchronize(h->ownStream));
  return ICB_OK;
}

int icb_get_bounds(icb_handle h, icb_bounds *out)
{
  if (!h || !out) return ICB_ERR_INVALID;
  if (!h->haveGrid) { h->err = "icb_get_bounds: no grid"; return ICB_ERR_NO_GRID; }
  *out = h->bounds;
  return ICB_OK;
}

int icb_set_transfunc(icb_handle h, const float *rgba, int size, float lo, float hi, float opacityScale)
{
  if (!h || !rgba || size < 1) return ICB_ERR_INVALID;
  if (size > kMaxLutEntries) { h->err = "icb_set_transfunc: more than 4096 entries"; return ICB_ERR_INVALID; }
  CK(cudaSetDevice(h->device));
  // the table is shared by every frame in flight (library-owned non-blocking slot streams included): a synchronous call
  CK(cudaDeviceSynchronize());
  if (size != h->lutSize) {
    cudaFree(h->dLut);
    h->dLut = nullptr;
    CK(cudaMalloc((void **)&h->dLut, (size_t)size * sizeof(float4)));
    h->lutSize = size;
  }
  CK(cudaMemcpy(h->dLut, rgba, (size_t)size * sizeof(float4), cudaMemcpyHostToDevice));
  {
    // prefix counts of the table entries that are NOT exactly clear (renderKernel builds the same table in shared memory)
    std::vector<int> nz((size_t)size + 1);
    const bool scaleOk = std::fabs(opacityScale) < INFINITY;
    int run = 0;
    for (int i = 0; i < size; ++i) {
      nz[i] = run;
      const float *c = rgba + 4 * (size_t)i;
      const bool clear = scaleOk && c[3] == 0.f && std::fabs(c[0]) < INFINITY && std::fabs(c[1]) < INFINITY && std::fabs(c[2]) < INFINITY;
      run += clear ? 0 : 1;
    }
    nz[size] = run;
    if (size + 1 > h->nzCap) {
      CK(cudaDeviceSynchronize());
      cudaFree(h->dNz);
      h->dNz = nullptr; h->nzCap = 0;
      CK(cudaMalloc((void **)&h->dNz, ((size_t)size + 1) * sizeof(int)));
      h->nzCap = size + 1;
    }
    CK(cudaMemcpy(h->dNz, nz.data(), nz.size() * sizeof(int), cudaMemcpyHostToDevice));
    h->maskDirty = true;
  }
  h->rp.tfLo = lo; h->rp.tfHi = hi; h->rp.opacityScale = opacityScale; h->rp.tfSize = size;
  h->majorDirty = std::max(h->majorDirty, 1); h->shellDirty = std::max(h->shellDirty, 1);
  return ICB_OK;
}

int icb_set_camera(icb_handle h, const float org[3], const float dir_00[3], const float dir_du[3], const float dir_dv[3])
{
  if (!h || !org || !dir_00 || !dir_du || !dir_dv) return ICB_ERR_INVALID;
  for (int k = 0; k < 3; ++k) { h->rp.org[k] = org[k]; h->rp.d00[k] = dir_00[k]; h->rp.du[k] = dir_du[k]; h->rp.dv[k] = dir_dv[k]; }
  h->haveCamera = true;
  return ICB_OK;
}

int icb_set_params(icb_handle h, float unitDistance, float dt, const float ambientColor[3], float ambientRadiance,
                   int accumID, unsigned flags, float termAlpha)
{
  if (!h) return ICB_ERR_INVALID;
  if (unitDistance > 0.f) h->rp.unitDistance = unitDistance;
  h->rp.dt = dt > 0.f ? dt : h->rp.unitDistance / 2;
  if (ambientColor) for (int k = 0; k < 3; ++k) h->rp.ambient[k] = ambientColor[k];
  h->rp.ambientRadiance = ambientRadiance;
  h->rp.accumID = accumID;
  h->rp.flags = flags;
  if (getenv("ICB_PRISM_NO_EMPTY_SKIP")) h->rp.flags |= 4u;      // developer knob: prism sampler without the transparent-prism shortcut
  h->rp.termAlpha = termAlpha > 0.f ? termAlpha : 0.99f;
  return ICB_OK;
}

int icb_set_tiles(icb_handle h, const int32_t *tileIds_host, int n)
{
  if (!h || n < 0 || (n > 0 && !tileIds_host)) return ICB_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  if (n > SL(h).tileListCap) {
    CK(cudaDeviceSynchronize());
    cudaFree(SL(h).dTileList);
    SL(h).dTileList = nullptr;
    SL(h).tileListCap = 0;
    CK(cudaMalloc((void **)&SL(h).dTileList, (size_t)n * sizeof(int32_t)));
    SL(h).tileListCap = n;
  }
  SL(h).nTileList = n;
  // synchronous copy; the caller must not change the list while a frame that uses it is still in flight
  if (n > 0) CK(cudaMemcpy(SL(h).dTileList, tileIds_host, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice));
  return ICB_OK;
}

static int checkReady(icb_context *h, int width, int height)
{
  if (!h->haveGrid) { h->err = "render: icb_set_grid has not been called"; return ICB_ERR_NO_GRID; }
  if (!h->dLut) { h->err = "render: icb_set_transfunc has not been called"; return ICB_ERR_NO_TRANSFUNC; }
  if (!h->haveCamera) { h->err = "render: icb_set_camera has not been called"; return ICB_ERR_INVALID; }
  if (width < 1 || height < 1 || (int64_t)width * height > (int64_t)1 << 30) { h->err = "render: bad frame size"; return ICB_ERR_INVALID; }
  return ICB_OK;
}

// the current slot's per-tile cost and order arrays, sized for a frame of `allTiles` tiles (grow-only; new arrays start at
// zero cost).  Shared by the render path and icb_set_tile_costs: a rank of a group that is dealt no tile in a frame
// (more ranks than blocks of tiles) still receives the all-reduced costs of that frame.
static int ensureTileCosts(icb_context *h, int allTiles, cudaStream_t stream)
{
  if (allTiles <= SL(h).tileCostCap) return ICB_OK;
  cudaFree(SL(h).dTileCost); cudaFree(SL(h).dOrder);
  SL(h).dTileCost = nullptr; SL(h).dOrder = nullptr;
  SL(h).tileCostCap = 0;
  CK(cudaMalloc((void **)&SL(h).dTileCost, (size_t)allTiles * sizeof(int)));
  CK(cudaMalloc((void **)&SL(h).dOrder, (size_t)allTiles * sizeof(int)));
  CK(cudaMemsetAsync(SL(h).dTileCost, 0, (size_t)allTiles * sizeof(int), stream));
  SL(h).tileCostCap = allTiles;
  return ICB_OK;
}

// prism sampler: bring the per-record transparency masks up to date (on `stream`); every render stream waits for the
// rebuild before its kernel reads them
static int ensurePrismMask(icb_context *h, cudaStream_t stream)
{
  if (!h->maskEvent) CK(cudaEventCreateWithFlags(&h->maskEvent, cudaEventDisableTiming));
  if (h->maskDirty) {
    if (!h->dPrismMask) CK(cudaMalloc((void **)&h->dPrismMask, std::max<size_t>(h->nRec, 1) * sizeof(uint32_t)));
    // (stream-ordered after whatever last changed the vertex scalars on this stream, and after every frame in flight that
    // still reads the old masks)
    int rcw = waitFramesInFlight(h, stream);
    if (rcw != ICB_OK) return rcw;
    CK(launchPrismMask(h->grid, h->dNz, h->rp.tfLo, h->rp.tfHi, h->rp.tfSize, h->dPrismMask, stream));
    CK(cudaEventRecord(h->maskEvent, stream));
    h->grid.prismMask = h->dPrismMask;
    h->maskDirty = false;
  } else {
    CK(cudaStreamWaitEvent(stream, h->maskEvent, 0));
  }
  return ICB_OK;
}

extern "C++" int icbRenderCommon(icb_context *h, uint32_t *fb, float *accum, float *depth, int32_t *perPixel, int width, int height,
                        bool compact, bool trace, cudaStream_t stream)
{
  int rc = checkReady(h, width, height);
  if (rc != ICB_OK) return rc;
  CK(cudaSetDevice(h->device));
  RenderParams p = h->rp;
  p.width = width; p.height = height;
  p.tilesX = (width + kTile - 1) / kTile;
  p.tilesY = (height + kTile - 1) / kTile;
  p.chunkLog2 = h->splitMode != ICB_SPLIT_NEVER ? h->chunkLog2 : 0;      // (0: the plain A.3 recurrence)
  // skip-ahead slack: 16 ulps of the largest coordinate magnitude in play (locator.cuh skipCount)
  {
    float M = h->rp.rOuter;
    for (int k = 0; k < 3; ++k) M = fmaxf(M, fabsf(p.org[k]));
    const float slack = 16.f * M * 1.1920929e-7f;
    p.skipRadSlack = slack;
    p.skipPlaneScale = h->marginUsed > 0.f ? 1.f + slack / h->marginUsed : 0.f;
    p.maxSkip = (h->skipAhead && h->marginUsed > 0.f && std::isfinite(M)) ? 1024 : 0;
    p.skipBack = h->skipBack;
  }
  if (h->integrator != ICB_INTEGRATOR_MARCH) {
    if (compact || trace) { h->err = "render: tile lists and traces apply to the marcher only"; return ICB_ERR_INVALID; }
    // the reference clears the frame when accumID == 0 (pipeline.cu:1058-1059) and leaves pixels whose ray misses
    // the volume's box untouched (deviceCode.cu:253-254)
    const size_t npx = (size_t)width * height;
    if (p.accumID == 0) {
      if (fb) CK(cudaMemsetAsync(fb, 0, npx * sizeof(uint32_t), stream));
      if (accum) CK(cudaMemsetAsync(accum, 0, npx * 4 * sizeof(float), stream));
      if (depth) CK(cudaMemsetAsync(depth, 0, npx * sizeof(float), stream));
    }
    CK(launchPrepareFrame(SL(h).dTileCounter, SL(h).dCounters, nullptr, 0, nullptr, 0, stream));
    RenderTargets wo{};
    wo.fb = fb; wo.accum = (float4 *)accum; wo.counters = SL(h).dCounters;
    if (h->integrator == ICB_INTEGRATOR_WOODCOCK_GRID) {
      const int builds = h->majorDirty >= 2 ? 3 : h->majorDirty;      // initRanges + gridRanges + maxOpacities | maxOpacities
      int rc2 = ensureMajorants(h, stream);
      if (rc2 != ICB_OK) return rc2;
      logTableWindow(h, stream, true);
      SL(h).used = true; CK(cudaEventRecord(SL(h).ev0, stream));
      CK(launchWoodcockGrid(h->grid, p, h->dLut, h->dLogTable, majorantGrid(h), wo, stream));
      CK(frameDone(h, stream));
      logTableWindow(h, stream, false);
      h->stats.kernelLaunches = 2 + builds;
      return ICB_OK;
    }
    if (h->integrator == ICB_INTEGRATOR_WOODCOCK_SPHERE) {
      const int builds = h->shellDirty >= 2 ? 3 : h->shellDirty;      // initRanges + shellRanges + maxOpacities | maxOpacities
      int rc2 = ensureShellMajorants(h, stream);
      if (rc2 != ICB_OK) return rc2;
      logTableWindow(h, stream, true);
      SL(h).used = true; CK(cudaEventRecord(SL(h).ev0, stream));
      CK(launchWoodcockSphere(h->grid, p, h->dLut, h->dLogTable, shellGrid(h), wo, stream));
      CK(frameDone(h, stream));
      logTableWindow(h, stream, false);
      h->stats.kernelLaunches = 2 + builds;
      return ICB_OK;
    }
    logTableWindow(h, stream, true);
    SL(h).used = true; CK(cudaEventRecord(SL(h).ev0, stream));
    CK(launchWoodcock(h->grid, p, h->dLut, h->dLogTable, wo, stream));
    CK(frameDone(h, stream));
    logTableWindow(h, stream, false);
    h->stats.kernelLaunches = 2;   // prepareFrameKernel + woodcockKernel
    return ICB_OK;
  }
  const int allTiles = p.tilesX * p.tilesY;
  const int *tileList = SL(h).nTileList > 0 ? SL(h).dTileList : nullptr;
  p.numTiles = tileList ? SL(h).nTileList : allTiles;
  if (compact && !tileList) { h->err = "icb_render_tiles: no tile list (icb_set_tiles)"; return ICB_ERR_INVALID; }
  { int rct = ensureTileCosts(h, allTiles, stream); if (rct != ICB_OK) return rct; }
  if (h->grid.sampler != 0) {
    int rcm = ensurePrismMask(h, stream);
    if (rcm != ICB_OK) return rcm;
  }
  // split rays (not for the trace kernels, whose id-sequence hash is a property of one lane marching the whole ray)
  // auto: a rank's share of a sort-first frame is short, so its longest rays ARE the frame time; on a full frame the
  // speculative chunks (those behind the sample at which a ray terminates) cost more than the tail they remove
  const bool splitWanted = h->splitMode == 1 || (h->splitMode == 2 && tileList && 2 * p.numTiles <= allTiles);
  const bool splitOn = !trace && splitWanted && h->splitMaxHeavy > 0 && h->grid.nGlobal == 0;
  if (splitOn && SL(h).splitCap < h->splitMaxHeavy) {
    CK(cudaStreamSynchronize(stream));
    cudaFree(SL(h).dCancel); cudaFree(SL(h).dChunks);
    SL(h).dCancel = nullptr; SL(h).dChunks = nullptr; SL(h).splitCap = 0;
    CK(cudaMalloc((void **)&SL(h).dCancel, (size_t)h->splitMaxHeavy * kTilePixels * sizeof(int32_t)));
    CK(cudaMalloc((void **)&SL(h).dChunks, (size_t)h->splitMaxHeavy * kTilePixels * kMaxChunks * sizeof(ChunkResult)));
    SL(h).splitCap = h->splitMaxHeavy;
  }
  // heaviest tiles first, judged by the previous frame's sample counts (all equal on a first frame)
  CK(launchOrderTiles(tileList, p.numTiles, SL(h).dTileCost, SL(h).dOrder, h->splitMinCostLog2, splitOn ? h->splitMaxHeavy : 0, SL(h).dNHeavy, stream));
  CK(launchPrepareFrame(SL(h).dTileCounter, SL(h).dCounters, SL(h).dTileCost, allTiles, splitOn ? SL(h).dCancel : nullptr,
                        splitOn ? h->splitMaxHeavy * kTilePixels : 0, stream));
  SplitArgs sp{};
  if (splitOn) { sp.nHeavy = SL(h).dNHeavy; sp.results = SL(h).dChunks; sp.cancelFrom = SL(h).dCancel; }
  RenderTargets out{};
  out.fb = fb; out.accum = (float4 *)accum; out.depth = depth; out.perPixel = perPixel;
  out.tileCost = SL(h).dTileCost; out.counters = SL(h).dCounters; out.compact = compact ? 1 : 0;
  if (h->lutSize != h->blocksForLut) { h->blocksPerSM = renderBlocksPerSM(h->lutSize); h->blocksForLut = h->lutSize; }
  if (h->grid.sampler != 0 && h->lutSize != h->blocksForLutPrism) { h->blocksPerSMPrism = renderBlocksPerSM(h->lutSize, true); h->blocksForLutPrism = h->lutSize; }
  const int perSM = h->grid.sampler != 0 ? h->blocksPerSMPrism : h->blocksPerSM;
  // frames in flight share the device: each frame's persistent kernel takes its share of the resident CTA slots, so
  // that the kernels of consecutive frames are co-resident instead of queueing for the slots of a kernel whose every
  // CTA holds one of the frame's long rays until the end
  int blocks = h->numSMs * (h->blocksOverride > 0 ? std::min(h->blocksOverride, perSM) : perSM);
  // (at least a quarter of the SMs' worth of CTAs: below that a kernel's own tail -- one long ray per warp -- outlasts its share)
  blocks = std::max(h->minBlocks > 0 ? h->minBlocks : h->numSMs / 2, (blocks + h->framesInFlight - 1) / h->framesInFlight);
  SL(h).used = true; CK(cudaEventRecord(SL(h).ev0, stream));
  CK(launchRender(h->grid, p, h->dLut, out, tileList, SL(h).dOrder, SL(h).dTileCounter, h->refillMin, blocks, trace, sp, stream));
  CK(frameDone(h, stream));
  h->stats.kernelLaunches = splitOn ? 4 : 3;   // orderTilesKernel + prepareFrameKernel (+ chunk pass) + renderKernel
  return ICB_OK;
}

int icb_render(icb_handle h, uint32_t *fbPointer_dev, float *accumBuffer_dev, float *fbDepth_dev, int width, int height, void *stream)
{
  if (!h) return ICB_ERR_INVALID;
  return icbRenderCommon(h, fbPointer_dev, accumBuffer_dev, fbDepth_dev, nullptr, width, height, false, false, (cudaStream_t)stream);
}

int icb_render_trace(icb_handle h, uint32_t *fbPointer_dev, float *accumBuffer_dev, int32_t *perPixel_dev, int width, int height, void *stream)
{
  if (!h || !perPixel_dev) return ICB_ERR_INVALID;
  return icbRenderCommon(h, fbPointer_dev, accumBuffer_dev, nullptr, perPixel_dev, width, height, false, true, (cudaStream_t)stream);
}

int icb_render_tiles(icb_handle h, uint32_t *rgba8_dev, float *accum_dev, int width, int height, void *stream)
{
  if (!h) return ICB_ERR_INVALID;
  return icbRenderCommon(h, rgba8_dev, accum_dev, nullptr, nullptr, width, height, true, false, (cudaStream_t)stream);
}

int icb_render_host(icb_handle h, uint32_t *fbPointer_host, float *accumBuffer_host, int width, int height)
{
  if (!h) return ICB_ERR_INVALID;
  int rc = checkReady(h, width, height);
  if (rc != ICB_OK) return rc;
  CK(cudaSetDevice(h->device));
  const size_t npx = (size_t)width * height;
  if (npx > SL(h).frameCap) {
    CK(cudaDeviceSynchronize());
    cudaFree(SL(h).dFrameFb); cudaFree(SL(h).dFrameAccum);
    SL(h).dFrameFb = nullptr; SL(h).dFrameAccum = nullptr; SL(h).frameCap = 0;
    CK(cudaMalloc((void **)&SL(h).dFrameFb, npx * sizeof(uint32_t)));
    CK(cudaMalloc((void **)&SL(h).dFrameAccum, npx * 4 * sizeof(float)));
    SL(h).frameCap = npx;
  }
  // accumBuffer_host == NULL keeps the running mean on the device between frames, as the reference does
  // (Frame::accumBuffer is device memory under RTCORE, fb.cu:22-46; present() copies fbPointer only)
  if (h->rp.accumID != 0 && accumBuffer_host)
    CK(cudaMemcpyAsync(SL(h).dFrameAccum, accumBuffer_host, npx * 4 * sizeof(float), cudaMemcpyHostToDevice, h->ownStream));
  rc = icbRenderCommon(h, SL(h).dFrameFb, SL(h).dFrameAccum, nullptr, nullptr, width, height, false, false, h->ownStream);
  if (rc != ICB_OK) return rc;
  if (fbPointer_host) CK(cudaMemcpyAsync(fbPointer_host, SL(h).dFrameFb, npx * sizeof(uint32_t), cudaMemcpyDeviceToHost, h->ownStream));
  if (accumBuffer_host)
    CK(cudaMemcpyAsync(accumBuffer_host, SL(h).dFrameAccum, npx * 4 * sizeof(float), cudaMemcpyDeviceToHost, h->ownStream));
  CK(cudaStreamSynchronize(h->ownStream));
  return ICB_OK;
}

// icb_render_host without the wait: renders on the current slot's own stream and enqueues the download of the RGBA8
// frame into page-locked host memory behind it; icb_frame_wait(slot) returns when that frame is in host memory.
int icb_render_host_async(icb_handle h, uint32_t *fbPointer_pinned, int width, int height)
{
  if (!h || !fbPointer_pinned) return ICB_ERR_INVALID;
  int rc = checkReady(h, width, height);
  if (rc != ICB_OK) return rc;
  // the running mean lives in the SLOT's accumulation buffer: frames of one progressive sequence (accumID > 0) must all
  // go through the same slot, one after the other
  if (h->rp.accumID > 0 && h->framesInFlight > 1) {
    h->err = "icb_render_host_async: accumID > 0 with several frames in flight (the running mean is per slot; use one slot)";
    return ICB_ERR_INVALID;
  }
  CK(cudaSetDevice(h->device));
  const size_t npx = (size_t)width * height;
  FrameSlot &sl = SL(h);
  if (npx > sl.frameCap) {
    CK(cudaStreamSynchronize(sl.stream));
    cudaFree(sl.dFrameFb); cudaFree(sl.dFrameAccum);
    sl.dFrameFb = nullptr; sl.dFrameAccum = nullptr; sl.frameCap = 0;
    CK(cudaMalloc((void **)&sl.dFrameFb, npx * sizeof(uint32_t)));
    CK(cudaMalloc((void **)&sl.dFrameAccum, npx * 4 * sizeof(float)));
    sl.frameCap = npx;
  }
  rc = icbRenderCommon(h, sl.dFrameFb, sl.dFrameAccum, nullptr, nullptr, width, height, false, false, sl.stream);
  if (rc != ICB_OK) return rc;
  CK(cudaMemcpyAsync(fbPointer_pinned, sl.dFrameFb, npx * sizeof(uint32_t), cudaMemcpyDeviceToHost, sl.stream));
  return ICB_OK;
}

int icb_frame_wait(icb_handle h, int slot)
{
  if (!h || slot < 0 || slot >= kFrameSlots) return ICB_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  CK(cudaStreamSynchronize(h->slots[slot].stream));
  return ICB_OK;
}

int icb_trace_pixel(icb_handle h, int x, int y, int width, int height, int32_t *ids_host, int maxIds, float rgba_host[4])
{
  if (!h || maxIds < 0 || (maxIds > 0 && !ids_host)) return ICB_ERR_INVALID;
  int rc = checkReady(h, width, height);
  if (rc != ICB_OK) return rc;
  if (x < 0 || y < 0 || x >= width || y >= height) return ICB_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  RenderParams p = h->rp;
  p.width = width; p.height = height;
  int *dIds = nullptr, *dN = nullptr;
  float4 *dRgba = nullptr;
  CK(cudaMalloc((void **)&dIds, std::max(1, maxIds) * sizeof(int)));
  CK(cudaMalloc((void **)&dN, sizeof(int)));
  CK(cudaMalloc((void **)&dRgba, sizeof(float4)));
  int n = 0;
  cudaError_t e = launchTracePixel(h->grid, p, h->dLut, x, y, dIds, maxIds, dRgba, dN, h->ownStream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(h->ownStream);
  if (e == cudaSuccess) e = cudaMemcpy(&n, dN, sizeof(int), cudaMemcpyDeviceToHost);
  if (e == cudaSuccess && maxIds > 0) e = cudaMemcpy(ids_host, dIds, sizeof(int) * std::min(n, maxIds), cudaMemcpyDeviceToHost);
  if (e == cudaSuccess && rgba_host) e = cudaMemcpy(rgba_host, dRgba, sizeof(float4), cudaMemcpyDeviceToHost);
  cudaFree(dIds); cudaFree(dN); cudaFree(dRgba);
  if (e != cudaSuccess) { h->err = std::string("icb_trace_pixel: ") + cudaGetErrorString(e); return ICB_ERR_CUDA; }
  return n;
}

int icb_locate(icb_handle h, const float *pos_dev, size_t n, int32_t *ids_dev, float *values_dev, unsigned flags, void *stream)
{
  if (!h || (n > 0 && (!pos_dev || !ids_dev || !values_dev))) return ICB_ERR_INVALID;
  if (!h->haveGrid) { h->err = "icb_locate: no grid"; return ICB_ERR_NO_GRID; }
  CK(cudaSetDevice(h->device));
  cudaStream_t s = (cudaStream_t)stream;
  CK(cudaMemsetAsync(SL(h).dCounters, 0, offsetof(Counters, totalSteps), s));
  SL(h).used = true; CK(cudaEventRecord(SL(h).ev0, s));
  CK(launchLocate(h->grid, pos_dev, n, ids_dev, values_dev, flags, SL(h).dCounters, s));
  CK(frameDone(h, s));
  h->stats.kernelLaunches = 1;
  return ICB_OK;
}

int icb_locate_host(icb_handle h, const float *pos_host, size_t n, int32_t *ids_host, float *values_host, unsigned flags)
{
  if (!h || (n > 0 && (!pos_host || !ids_host || !values_host))) return ICB_ERR_INVALID;
  if (!h->haveGrid) { h->err = "icb_locate: no grid"; return ICB_ERR_NO_GRID; }
  if (n == 0) return ICB_OK;
  CK(cudaSetDevice(h->device));
  float *dPos = nullptr, *dVal = nullptr;
  int *dIds = nullptr;
  CK(cudaMalloc((void **)&dPos, n * 3 * sizeof(float)));
  CK(cudaMalloc((void **)&dVal, n * sizeof(float)));
  CK(cudaMalloc((void **)&dIds, n * sizeof(int)));
  cudaError_t e = cudaMemcpyAsync(dPos, pos_host, n * 3 * sizeof(float), cudaMemcpyHostToDevice, h->ownStream);
  int rc = ICB_OK;
  if (e == cudaSuccess) rc = icb_locate(h, dPos, n, dIds, dVal, flags, h->ownStream);
  if (e == cudaSuccess && rc == ICB_OK) e = cudaMemcpyAsync(ids_host, dIds, n * sizeof(int), cudaMemcpyDeviceToHost, h->ownStream);
  if (e == cudaSuccess && rc == ICB_OK) e = cudaMemcpyAsync(values_host, dVal, n * sizeof(float), cudaMemcpyDeviceToHost, h->ownStream);
  if (e == cudaSuccess && rc == ICB_OK) e = cudaStreamSynchronize(h->ownStream);
  cudaFree(dPos); cudaFree(dVal); cudaFree(dIds);
  if (rc != ICB_OK) return rc;
  if (e != cudaSuccess) { h->err = std::string("icb_locate_host: ") + cudaGetErrorString(e); return ICB_ERR_CUDA; }
  return ICB_OK;
}

int icb_scatter_tiles(icb_handle h, const uint32_t *rgba8_dev, const float *accum_dev, const int32_t *tileIds_dev, int n,
                      uint32_t *fbPointer_dev, float *accumBuffer_dev, int width, int height, void *stream)
{
  if (!h || n < 0 || width < 1 || height < 1) return ICB_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  CK(launchScatterTiles(rgba8_dev, accum_dev, tileIds_dev, n, fbPointer_dev, accumBuffer_dev, width, height, (cudaStream_t)stream));
  return ICB_OK;
}

int icb_tile_costs(icb_handle h, int32_t *costs_dev, int numTiles, void *stream)
{
  if (!h || !costs_dev || numTiles < 0) return ICB_ERR_INVALID;
  if (numTiles > SL(h).tileCostCap) { h->err = "icb_tile_costs: more tiles than the last frame had"; return ICB_ERR_INVALID; }
  CK(cudaSetDevice(h->device));
  CK(cudaMemcpyAsync(costs_dev, SL(h).dTileCost, (size_t)numTiles * sizeof(int), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return ICB_OK;
}

int icb_set_tile_costs(icb_handle h, const int32_t *costs_dev, int numTiles, void *stream)
{
  if (!h || !costs_dev || numTiles < 0) return ICB_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  { int rct = ensureTileCosts(h, numTiles, (cudaStream_t)stream); if (rct != ICB_OK) return rct; }   // (a slot that has not rendered yet)
  CK(cudaMemcpyAsync(SL(h).dTileCost, costs_dev, (size_t)numTiles * sizeof(int), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return ICB_OK;
}

int icb_set_tiles_async(icb_handle h, const int32_t *tileIds_pinned, int n, void *stream)
{
  if (!h || n < 0 || (n > 0 && !tileIds_pinned)) return ICB_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  if (n > SL(h).tileListCap) {            // growing the list is the one synchronising case (first frame, larger frame)
    CK(cudaDeviceSynchronize());
    cudaFree(SL(h).dTileList);
    SL(h).dTileList = nullptr;
    SL(h).tileListCap = 0;
    const int cap = std::max(n, 1 << 16);
    CK(cudaMalloc((void **)&SL(h).dTileList, (size_t)cap * sizeof(int32_t)));
    SL(h).tileListCap = cap;
  }
  SL(h).nTileList = n;
  if (n > 0) CK(cudaMemcpyAsync(SL(h).dTileList, tileIds_pinned, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, (cudaStream_t)stream));
  return ICB_OK;
}

int icb_stats(icb_handle h, icb_stats_t *out)
{
  if (!h || !out) return ICB_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  if (h->stats.kernelLaunches > 0 && SL(h).used) {
    CK(cudaEventSynchronize(SL(h).ev1));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, SL(h).ev0, SL(h).ev1));
    h->stats.lastKernelMs = ms;
    Counters c;
    CK(cudaMemcpy(&c, SL(h).dCounters, sizeof(Counters), cudaMemcpyDeviceToHost));
    h->stats.steps = c.steps; h->stats.hits = c.hits; h->stats.slowSteps = c.slow; h->stats.exactQueries = c.exact;
    h->stats.raysMarched = c.rays;
    h->stats.totalSteps = c.totalSteps; h->stats.totalHits = c.totalHits;
    for (int k = 0; k < kFrameSlots; ++k)          // totals run over all frame slots
      if (k != h->cur && h->slots[k].used) {
        CK(cudaEventSynchronize(h->slots[k].ev1));
        Counters o;
        CK(cudaMemcpy(&o, h->slots[k].dCounters, sizeof(Counters), cudaMemcpyDeviceToHost));
        h->stats.totalSteps += o.totalSteps; h->stats.totalHits += o.totalHits;
      }
  }
  *out = h->stats;
  return ICB_OK;
}

int icb_debug_counters(icb_handle h, uint64_t out[8])
{
  if (!h || !out) return ICB_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  CK(cudaDeviceSynchronize());
  Counters c;
  CK(cudaMemcpy(&c, SL(h).dCounters, sizeof(Counters), cudaMemcpyDeviceToHost));
  for (int k = 0; k < 8; ++k) out[k] = c.dbg[k];
  return ICB_OK;
}

int icb_reset_totals(icb_handle h)
{
  if (!h) return ICB_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  CK(cudaDeviceSynchronize());
  for (FrameSlot &sl : h->slots) CK(cudaMemset(&sl.dCounters->totalSteps, 0, 2 * sizeof(unsigned long long)));
  return ICB_OK;
}

int icb_dev_alloc(icb_handle h, size_t bytes, void **out_dev)
{
  if (!h || !out_dev) return ICB_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  CK(cudaMalloc(out_dev, bytes > 0 ? bytes : 1));
  return ICB_OK;
}

int icb_dev_free(icb_handle h, void *ptr_dev)
{
  if (!h) return ICB_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  CK(cudaFree(ptr_dev));
  return ICB_OK;
}

int icb_dev_upload(icb_handle h, void *dst_dev, const void *src_host, size_t bytes)
{
  if (!h || (bytes > 0 && (!dst_dev || !src_host))) return ICB_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  CK(cudaMemcpy(dst_dev, src_host, bytes, cudaMemcpyHostToDevice));
  return ICB_OK;
}

int icb_dev_download(icb_handle h, void *dst_host, const void *src_dev, size_t bytes)
{
  if (!h || (bytes > 0 && (!dst_host || !src_dev))) return ICB_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  CK(cudaMemcpy(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost));
  return ICB_OK;
}

int icb_host_alloc(icb_handle h, size_t bytes, void **out_host)
{
  if (!h || !out_host) return ICB_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  CK(cudaHostAlloc(out_host, bytes > 0 ? bytes : 1, cudaHostAllocDefault));
  return ICB_OK;
}

int icb_host_free(icb_handle h, void *ptr_host)
{
  if (!h) return ICB_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  CK(cudaFreeHost(ptr_host));
  return ICB_OK;
}

int icb_dev_sync(icb_handle h)
{
  if (!h) return ICB_ERR_INVALID;
  CK(cudaSetDevice(h->device));
  CK(cudaDeviceSynchronize());
  return ICB_OK;
}

}  // extern "C"
