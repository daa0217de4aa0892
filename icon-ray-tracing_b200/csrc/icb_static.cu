// icb_trace_pixel_static: icb_trace_pixel for callers that must not allocate (include/icon_b200.h).
//
// icb_trace_pixel stages its result on the device -- the id sequence, the step count and the pixel's RGBA -- in three
// buffers it cudaMallocs and cudaFrees on every call.  That is fine for a test harness, and wrong for two kinds of caller:
//   * a present loop that keeps frames in flight (icb_render_host_async, icb_group_render) and wants to probe a pixel
//     while they run: cudaFree waits for ALL outstanding work on the device, so every probe drains the pipeline;
//   * a process whose device memory is spoken for -- an R2B10 replica is ~73 GB, a host application may have handed the
//     rest to its own pool (torch's caching allocator, a renderer's arena) -- where cudaMalloc can fail at any time, or
//     where the embedding forbids driver allocations after start-up altogether.
// For them the staging lives in the file-scope array below: part of the library's device image, present on every GPU
// that loads the library from module load until the process exits (64.25 KB per GPU), never allocated, never freed.
//
// There is ONE such array per GPU, shared by every handle and every call, and it keeps what the last call left in it:
//   * calls that can overlap on one GPU (two handles, two host threads) race on it -- serialise them;
//   * the header (step count, RGBA) is zeroed on the stream in front of every launch, so that a launch that did not run
//     cannot hand back the previous pixel's count; ids past the returned count are never copied out.
// The array's name is a device symbol: in host code it is only a stand-in, not an address the GPU can use.  The device
// address comes from cudaGetSymbolAddress, asked once per device with that device current, by the FIRST call on that
// device (scratchFor below).  That first call may make the runtime load this module on the device, so it should be made
// during start-up -- before any stream capture the caller intends to do, and before the caller's "no more allocations"
// point if it counts the driver's own module memory.
#include <algorithm>
#include <mutex>
#include <string>

#include "icb_context.h"

namespace {

constexpr int kHeaderBytes = 256;                      // RGBA (float4) at 0, step count (int) at 16; ids start 256-aligned
constexpr int kMaxIds = ICB_TRACE_STATIC_MAX_IDS;
constexpr int kMaxDevices = 64;

static __device__ __align__(256) unsigned char gTraceScratch[kHeaderBytes + (size_t)kMaxIds * sizeof(int)];

std::mutex gScratchMutex;
unsigned char *gScratchPtr[kMaxDevices] = {};          // device address of gTraceScratch on each device, once asked

// the device address of the array on `device`, which must be the current device
cudaError_t scratchFor(int device, unsigned char **out)
{
  std::lock_guard<std::mutex> lock(gScratchMutex);
  if (!gScratchPtr[device]) {
    void *p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, gTraceScratch);
    if (e != cudaSuccess) return e;
    gScratchPtr[device] = static_cast<unsigned char *>(p);
  }
  *out = gScratchPtr[device];
  return cudaSuccess;
}

}  // namespace

#define CKS(call)                                                                                  \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      h->err = std::string("icb_trace_pixel_static: " #call ": ") + cudaGetErrorString(e_);        \
      return ICB_ERR_CUDA;                                                                         \
    }                                                                                              \
  } while (0)

extern "C" int icb_trace_pixel_static(icb_handle h, int x, int y, int width, int height, int32_t *ids_host, int maxIds,
                                      float rgba_host[4])
{
  if (!h) return ICB_ERR_INVALID;
  if (maxIds < 0 || (maxIds > 0 && !ids_host)) { h->err = "icb_trace_pixel_static: bad ids buffer"; return ICB_ERR_INVALID; }
  if (maxIds > kMaxIds) {
    h->err = "icb_trace_pixel_static: maxIds " + std::to_string(maxIds) + " exceeds ICB_TRACE_STATIC_MAX_IDS (" +
             std::to_string(kMaxIds) + "), the size of the library's built-in staging array; use icb_trace_pixel";
    return ICB_ERR_INVALID;
  }
  if (!h->haveGrid) { h->err = "icb_trace_pixel_static: icb_set_grid has not been called"; return ICB_ERR_NO_GRID; }
  if (!h->dLut) { h->err = "icb_trace_pixel_static: icb_set_transfunc has not been called"; return ICB_ERR_NO_TRANSFUNC; }
  if (!h->haveCamera) { h->err = "icb_trace_pixel_static: icb_set_camera has not been called"; return ICB_ERR_INVALID; }
  if (width < 1 || height < 1 || (int64_t)width * height > (int64_t)1 << 30 || x < 0 || y < 0 || x >= width || y >= height) {
    h->err = "icb_trace_pixel_static: pixel outside the frame"; return ICB_ERR_INVALID;
  }
  if (h->device < 0 || h->device >= kMaxDevices) { h->err = "icb_trace_pixel_static: device ordinal out of range"; return ICB_ERR_INVALID; }
  CKS(cudaSetDevice(h->device));
  unsigned char *scratch = nullptr;
  CKS(scratchFor(h->device, &scratch));
  float4 *dRgba = reinterpret_cast<float4 *>(scratch);
  int *dN = reinterpret_cast<int *>(scratch + sizeof(float4));
  int *dIds = reinterpret_cast<int *>(scratch + kHeaderBytes);

  RenderParams p = h->rp;            // exactly what icb_trace_pixel hands the kernel
  p.width = width; p.height = height;
  int n = 0;
  CKS(cudaMemsetAsync(scratch, 0, kHeaderBytes, h->ownStream));
  CKS(launchTracePixel(h->grid, p, h->dLut, x, y, dIds, maxIds, dRgba, dN, h->ownStream));
  CKS(cudaMemcpyAsync(&n, dN, sizeof(int), cudaMemcpyDeviceToHost, h->ownStream));
  CKS(cudaStreamSynchronize(h->ownStream));
  if (maxIds > 0 && n > 0) CKS(cudaMemcpyAsync(ids_host, dIds, sizeof(int) * (size_t)std::min(n, maxIds), cudaMemcpyDeviceToHost, h->ownStream));
  if (rgba_host) CKS(cudaMemcpyAsync(rgba_host, dRgba, sizeof(float4), cudaMemcpyDeviceToHost, h->ownStream));
  CKS(cudaStreamSynchronize(h->ownStream));
  return n;
}
