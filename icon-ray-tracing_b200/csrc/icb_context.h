// Internal state behind an icb_handle (include/icon_b200.h), shared by the single-GPU entry points (icb_api.cu)
// and the multi-GPU group (icb_group.cu).  Not part of the ABI.
#pragma once
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "grid_build.h"
#include "icon_b200.h"
#include "kernels.h"

using namespace icb;

// Per-frame device state.  Two slots let a caller keep two frames in flight on two streams (icb_set_frame_slot): the
// tail of one frame -- a handful of long limb rays, a serial chain of dependent fetches -- is then filled by the head
// of the next, instead of idling the GPU.
struct FrameSlot {
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  int32_t *dTileList = nullptr;
  int nTileList = 0, tileListCap = 0;
  int *dTileCounter = nullptr;
  int *dTileCost = nullptr;
  int *dOrder = nullptr;
  int tileCostCap = 0;
  Counters *dCounters = nullptr;
  // split rays: number of heavy tiles of the frame (device), chunk results and cancel marks of their rays
  int32_t *dNHeavy = nullptr, *dCancel = nullptr;
  ChunkResult *dChunks = nullptr;
  int splitCap = 0;       // heavy tiles the two arrays above are sized for
  bool used = false;      // ev0/ev1 have been recorded
  // time series with two copies of the records' lines (icb_group_swap_scalars): behind every frame, the event of the copy
  // it read -- the swap that overwrites a copy waits for exactly the frames that still read that copy
  cudaEvent_t evBuf[2] = {nullptr, nullptr};
  bool bufUsed[2] = {false, false};
  // frame buffers owned by the handle for the host-buffer entry points (grow-only)
  uint32_t *dFrameFb = nullptr;
  float *dFrameAccum = nullptr;
  size_t frameCap = 0;
  cudaStream_t stream = nullptr;   // this slot's own stream (icb_render_host_async)
};
constexpr int kFrameSlots = ICB_FRAME_SLOTS;

struct icb_context {
  FrameSlot slots[kFrameSlots];
  int cur = 0;
  int device = 0;
  std::string err;
  cudaStream_t ownStream = nullptr;
  int numSMs = 148, blocksPerSM = 1, blocksForLut = 300;

  // grid
  bool haveGrid = false;
  size_t nRec = 0, nCol = 0;
  RecBlock *dRec = nullptr;
  float *dLines = nullptr;
  float *dLinesAlt = nullptr;   // second copy of the lines (time series: the next timestep is scattered into the copy no frame reads)
  int linesBuf = 0;             // which copy grid.lines points to: 0 dLines, 1 dLinesAlt
  int32_t *dOrigToInt = nullptr;
  ColBlock *dCol = nullptr;
  int32_t *dBinOff = nullptr, *dBinList = nullptr, *dGlobal = nullptr, *dSeed = nullptr;
  float *dStage = nullptr;      // staging for icb_set_scalars (n*32 floats), allocated on first use
  // prism sampler: vertex -> (column, corner) lists, layers per column, and (allocated by icb_set_sampler) the table
  int32_t *dVtxOff = nullptr, *dVtxCols = nullptr, *dColLayers = nullptr;
  float *dVtxScalars = nullptr;
  size_t nVert = 0;
  int vtxStride = 0;
  int sampler = ICB_SAMPLER_CELL;
  // majorant grid of ICB_INTEGRATOR_WOODCOCK_GRID (built lazily on the render stream)
  RawVec<float> colTrigHost;   // kept until first uploaded
  float *dColTrig = nullptr, *dRanges = nullptr, *dMaxOp = nullptr;
  int accelDims[3] = {256, 256, 256};
  // shell accel of ICB_INTEGRATOR_WOODCOCK_SPHERE: (rad, lat, lon) cells over the spherical bounds (hostCode.cu:654: 1 x 1024 x 1024)
  float *dShellRanges = nullptr, *dShellMaxOp = nullptr;
  int shellDims[3] = {1, 1024, 1024};
  int shellDirty = 2;
  cudaEvent_t shellEvent = nullptr;
  bool l2Persist = false;           // persisting-L2 carve-out available for the logf table
  size_t l2WindowMax = 0;
  int framesInFlight = 1;
  // prism sampler: per-record transparency masks (rebuilt lazily when the transfer function or the scalars changed)
  uint32_t *dPrismMask = nullptr;
  int *dNz = nullptr;
  int nzCap = 0;
  bool maskDirty = true;
  cudaEvent_t maskEvent = nullptr;
  cudaEvent_t majorEvent = nullptr; // recorded behind the last majorant build: every other render stream waits on it
  int majorDirty = 2;               // 2: value ranges + opacities stale, 1: opacities stale, 0: current
  int blocksPerSMPrism = 1, blocksForLutPrism = -1;
  DeviceGrid grid{};
  icb_bounds bounds{};
  icb_stats_t stats{};
  float marginMeters = 16.f;
  float marginUsed = 16.f;     // the margin the current grid was built with
  bool skipAhead = true;
  int maxWalk = 8;
  int refillMin = 8;
  int skipBack = 0;
  int chunkLog2 = 10;           // marcher: samples per compositing chunk (icb_set_ray_split)
  int splitMaxHeavy = 384;      // tiles per frame whose rays go through the chunk pass (0: no ray is split)
  int splitMode = 0;            // 0 never, 1 always, 2 auto: when the frame is a sort-first share (a tile list of at most half the tiles)
  int splitMinCostLog2 = 16;    // ... among those whose cost (lane iterations of their rays, an earlier frame) is >= 2^this
  int integrator = ICB_INTEGRATOR_MARCH;
  float *dLogTable = nullptr;   // logf(1 - k/2^24) from the host libm (Woodcock integrator), built on first use
  int logfForm = -1;            // -1 not checked yet; 0 the computed (FMA-build) form reproduces the host libm; 2 the kernels read dLogTable
  int blocksOverride = 0;
  int minBlocks = 0;                // ICB_MIN_BLOCKS: smallest grid a frame's persistent kernel is given (default: half the SM count)

  // transfer function
  float4 *dLut = nullptr;
  int lutSize = 0;

  RenderParams rp{};
  bool haveCamera = false;

  // tiles
  // frame buffers owned by the handle for the host-buffer entry point (grow-only)
};


// ---- shared between icb_api.cu and icb_group.cu ------------------------------------------------------------------
// everything icb_set_grid derives from the host builder besides the arrays themselves: what a rank that RECEIVES the
// finished device arrays from a peer (icb_group_set_grid) needs to adopt them
struct GridMeta {
  uint64_t nRec, nCol, nVert;
  uint64_t nBinOff, nBinList, nGlobal, nSeed, nVtxOff, nVtxCols, nColLayers, nColTrig;
  int32_t binN, recOfOrig0, maxColLayers, pad0;
  float cullLo, cullHi, marginUsed, pad1;
  int64_t numGeneric;
  icb_bounds bounds;
};
// points the handle's DeviceGrid at its (already filled) device arrays and fills bounds / params / stats from `m`
int icbAdoptGrid(icb_context *h, const GridMeta &m, double setupSeconds);
void icbFreeGrid(icb_context *h);
// one frame of whatever integrator is selected, into caller-provided device buffers (icb_render & co.)
int icbRenderCommon(icb_context *h, uint32_t *fb, float *accum, float *depth, int32_t *perPixel, int width, int height,
                    bool compact, bool trace, cudaStream_t stream);
