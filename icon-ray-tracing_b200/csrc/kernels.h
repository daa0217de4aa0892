// Launch wrappers of the sm_100a kernels (kernels.cu).
#pragma once
#include <cuda_runtime.h>

#include "icb_types.h"

namespace icb {

#ifndef ICB_RENDER_THREADS
#define ICB_RENDER_THREADS 256
#endif
constexpr int kRenderThreads = ICB_RENDER_THREADS;
constexpr int kTile = 16;                 // ICB_TILE
constexpr int kTilePixels = kTile * kTile;
constexpr int kUnitPixels = 32;           // work unit: one 8x4 block of a tile, a ray per lane
constexpr int kUnitsPerTile = kTilePixels / kUnitPixels;
constexpr int kMaxLutEntries = 4096;      // transfer-function entries staged in shared memory (64 KB)

struct RenderTargets {
  unsigned *fb;        // RGBA8 (full frame, or compact tiles when compact != 0)
  float4   *accum;     // fp32 RGBA running mean
  float    *depth;     // cleared at accumID==0 like clearFramebuffer (pipeline.cu:171-199); never written otherwise
  int      *perPixel;  // TRACE only: {steps, hits, hashLo, hashHi} per pixel
  int      *tileCost;  // steps per screen tile (next frame's load-balancing estimate)
  Counters *counters;
  int       compact;
};

cudaError_t launchPrepareFrame(int *tileCounter, Counters *counters, int *tileCost, int nTiles, int *cancelFrom, int nCancel, cudaStream_t stream);
cudaError_t configureKernels();
int renderBlocksPerSM(int tfSize, bool prism = false);
cudaError_t launchPrismMask(const DeviceGrid &g, const int *nz, float tfLo, float tfHi, int tfSize, uint32_t *mask, cudaStream_t stream);
cudaError_t launchVertexScalars(const DeviceGrid &g, const int *colLayers, const int *vtxOff, const int *vtxCols, size_t nVert,
                                float *out, cudaStream_t stream);
cudaError_t launchRender(const DeviceGrid &g, const RenderParams &p, const float4 *lut, const RenderTargets &out,
                         const int *tileList, const int *order, int *tileCounter, int refillMin, int numBlocks, bool trace,
                         const SplitArgs &sp, cudaStream_t stream);
cudaError_t launchWoodcock(const DeviceGrid &g, const RenderParams &p, const float4 *lut, const float *logTable,
                           const RenderTargets &out, cudaStream_t stream);
cudaError_t launchWoodcockGrid(const DeviceGrid &g, const RenderParams &p, const float4 *lut, const float *logTable,
                               const MajorantGrid &mg, const RenderTargets &out, cudaStream_t stream);
cudaError_t launchWoodcockSphere(const DeviceGrid &g, const RenderParams &p, const float4 *lut, const float *logTable,
                                 const MajorantGrid &sg, const RenderTargets &out, cudaStream_t stream);
cudaError_t launchShellRanges(const DeviceGrid &g, const float *colTrig, const MajorantGrid &sg, float *ranges, cudaStream_t stream);
cudaError_t launchLogfCheck(const float *table, unsigned long long *mismatches, cudaStream_t stream);
cudaError_t launchGridRanges(const DeviceGrid &g, const float *colTrig, const MajorantGrid &mg, float *ranges, cudaStream_t stream);
cudaError_t launchMaxOpacities(const float *ranges, const MajorantGrid &mg, const float4 *lut, int size, float tfLo, float tfHi,
                               float *maxOp, uint32_t *occupied, cudaStream_t stream);
cudaError_t launchOrderTiles(const int *tileList, int numTiles, const int *tileCost, int *order, int minCostLog2, int maxHeavy, int *nHeavy,
                             cudaStream_t stream);
cudaError_t launchTracePixel(const DeviceGrid &g, const RenderParams &p, const float4 *lut, int px, int py, int *ids,
                             int maxIds, float4 *rgbaOut, int *nStepsOut, cudaStream_t stream);
cudaError_t launchLocate(const DeviceGrid &g, const float *pos, size_t n, int *ids, float *values, unsigned flags,
                         Counters *counters, cudaStream_t stream);
cudaError_t launchScatterScalars(const float *src, const int *origToInt, float *dst, size_t nRec, cudaStream_t stream);
cudaError_t launchScatterTiles(const unsigned *rgba8, const float *accumT, const int *tileIds, int nTiles, unsigned *fb,
                               float *accum, int width, int height, cudaStream_t stream);

}  // namespace icb
