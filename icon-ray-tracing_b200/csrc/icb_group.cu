// Multi-GPU behind the C ABI (include/icon_b200.h "icb_group_*"): sort-first frames over a grid
// replicated per GPU.  The reference has no multi-GPU path -- nothing is mirrored here except the
// frame itself, which must be bit-identical to the single-GPU frame (no cross-tile blending).
//
// A group = this process's share of `world` ranks: one rank per process (icb_group_create, the torchrun layout) or
// all ranks in one process (icb_group_create_local, what the C++ command-line host uses).  Per frame and rank:
//   partition (C++, identical on every rank, from the all-reduced per-tile costs of the frame that left the device
//   `framesInFlight` frames ago) -> tile list up (pinned, stream-ordered) -> renderKernel into the rank's slice of the
//   compact tile buffer -> [comm stream] ncclSend/ncclRecv of the slices to rank 0 + ncclAllReduce of the tile costs
//   -> costs back to pinned host memory behind an event -> rank 0: scatterTilesKernel into the frame (+ D2H).
// Nothing in icb_group_render waits for the device except for the cost event recorded `framesInFlight` frames ago.
//
// NCCL is resolved at run time (dlopen "libnccl.so.2": the copy a host process has already loaded, e.g. torch's, else
// the system's), so that libicon_b200.so itself has no NCCL dependency for single-GPU callers.
#include <dlfcn.h>

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <initializer_list>
#include <string>
#include <vector>

#include "icb_context.h"

namespace {

// ---- the slice of the NCCL 2.x C API used here (ABI-stable across 2.x) ------------------------------------------
typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { ncclUint8 = 1, ncclInt32 = 2, ncclFloat32 = 7 };
enum { ncclSum = 0 };

struct Nccl {
  void *lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*CommSplit)(ncclComm_t, int, int, ncclComm_t *, void *) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*Send)(const void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void *, void *, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
  int (*GetVersion)(int *) = nullptr;
};

Nccl g_nccl;
std::string g_groupError;

bool loadNccl(std::string &err)
{
  if (g_nccl.lib) return true;
  void *lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);        // the copy this process already uses (e.g. torch's)
  if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
  if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
  if (!lib) { err = std::string("NCCL not found (dlopen libnccl.so.2): ") + dlerror(); return false; }
  Nccl n;
  n.lib = lib;
  bool ok = true;
  auto sym = [&](const char *name, bool required = true) -> void * {
    void *p = dlsym(lib, name);
    if (!p && required) { err = std::string("NCCL symbol missing: ") + name; ok = false; }
    return p;
  };
  *(void **)&n.GetUniqueId = sym("ncclGetUniqueId");
  *(void **)&n.CommInitRank = sym("ncclCommInitRank");
  *(void **)&n.CommInitAll = sym("ncclCommInitAll");
  *(void **)&n.CommDestroy = sym("ncclCommDestroy");
  *(void **)&n.CommSplit = sym("ncclCommSplit", false);
  *(void **)&n.GroupStart = sym("ncclGroupStart");
  *(void **)&n.GroupEnd = sym("ncclGroupEnd");
  *(void **)&n.Send = sym("ncclSend");
  *(void **)&n.Recv = sym("ncclRecv");
  *(void **)&n.AllReduce = sym("ncclAllReduce");
  *(void **)&n.AllGather = sym("ncclAllGather");
  *(void **)&n.Broadcast = sym("ncclBroadcast");
  *(void **)&n.GetErrorString = sym("ncclGetErrorString");
  *(void **)&n.GetVersion = sym("ncclGetVersion", false);
  if (!ok) return false;
  g_nccl = n;
  return true;
}

// ---- the partition (deterministic; every rank computes the same one from the same all-reduced costs) -------------
inline uint32_t spread16(uint32_t v)
{
  v &= 0xffffu;
  v = (v | (v << 8)) & 0x00ff00ffu;
  v = (v | (v << 4)) & 0x0f0f0f0fu;
  v = (v | (v << 2)) & 0x33333333u;
  v = (v | (v << 1)) & 0x55555555u;
  return v;
}

void mortonOrder(int tilesX, int tilesY, std::vector<int32_t> &order)
{
  const int n = tilesX * tilesY;
  std::vector<std::pair<uint32_t, int32_t>> key((size_t)n);
  for (int t = 0; t < n; ++t) key[(size_t)t] = {spread16((uint32_t)(t % tilesX)) | (spread16((uint32_t)(t / tilesX)) << 1), t};
  std::sort(key.begin(), key.end());
  order.resize((size_t)n);
  for (int i = 0; i < n; ++i) order[(size_t)i] = key[(size_t)i].second;
}

// tiles per dealt block: 16 Z-consecutive 16x16 tiles = a 64x64-pixel square (ICB_DEAL_BLOCK: experiments with larger blocks)
int dealBlock()
{
  static const int v = [] { const char *e = getenv("ICB_DEAL_BLOCK"); const int n = e ? atoi(e) : 16; return n >= 1 && n <= 65536 ? n : 16; }();
  return v;
}

// Blocks of dealBlock() Z-consecutive tiles, sorted by cost (heaviest first, ties by position), dealt to the ranks in
// snake order (0..n-1, n-1..0, ...): every rank receives the same mix of heavy (limb), medium and empty screen regions,
// and the frame's longest rays -- serial chains of dependent fetches that bound a rank's frame time from below -- are
// spread over all ranks.  Each rank's tiles stay in Z order.
void dealTiles(const std::vector<int32_t> &order, const int32_t *costs, int nRanks, std::vector<int32_t> &tiles, std::vector<int32_t> &offsets)
{
  const int n = (int)order.size();
  const int kDealBlock = dealBlock();
  const int nb = (n + kDealBlock - 1) / kDealBlock;
  std::vector<int64_t> bw((size_t)nb, 0);
  for (int i = 0; i < n; ++i) bw[(size_t)(i / kDealBlock)] += (costs ? (int64_t)costs[order[(size_t)i]] : 0) + 1;   // +1: a tile is never free
  std::vector<int32_t> byCost((size_t)nb);
  for (int b = 0; b < nb; ++b) byCost[(size_t)b] = b;
  std::stable_sort(byCost.begin(), byCost.end(), [&](int32_t a, int32_t b) { return bw[(size_t)a] > bw[(size_t)b]; });
  std::vector<int32_t> rankOfBlock((size_t)nb);
  for (int j = 0; j < nb; ++j) {
    const int m = j % (2 * nRanks);
    rankOfBlock[(size_t)byCost[(size_t)j]] = m < nRanks ? m : 2 * nRanks - 1 - m;
  }
  offsets.assign((size_t)nRanks + 1, 0);
  for (int b = 0; b < nb; ++b) offsets[(size_t)rankOfBlock[(size_t)b] + 1] += std::min(kDealBlock, n - b * kDealBlock);
  for (int r = 0; r < nRanks; ++r) offsets[(size_t)r + 1] += offsets[(size_t)r];
  tiles.resize((size_t)n);
  std::vector<int32_t> fill(offsets.begin(), offsets.end() - 1);
  for (int b = 0; b < nb; ++b) {
    int32_t &f = fill[(size_t)rankOfBlock[(size_t)b]];
    for (int i = b * kDealBlock; i < std::min(n, (b + 1) * kDealBlock); ++i) tiles[(size_t)f++] = order[(size_t)i];
  }
}

struct SlotRes {                 // what one frame in flight writes, per local rank
  uint32_t *compact = nullptr;   // compact tiles of ALL ranks, rank-major (rank 0 receives straight into place)
  int32_t *costsDev = nullptr;
  int32_t *tilesPinned = nullptr, *costsHost = nullptr;
  uint32_t *fb = nullptr;        // rank 0: the assembled frame
  int32_t *idsDev = nullptr, *idsPinned = nullptr;
  cudaEvent_t evRendered = nullptr, evComm = nullptr, evCosts = nullptr, evFrame = nullptr;
  bool costsPending = false;
};

struct RankState {
  icb_context *h = nullptr;
  int rank = 0;
  ncclComm_t comm = nullptr, comm2 = nullptr;      // comm2: time-series slab exchange (beside the tile gather)
  cudaStream_t commStream = nullptr, copyStream = nullptr, agStream = nullptr;   // copy: H2D of a slab; ag: its ncclAllGather (the next slab's H2D overlaps it)
  cudaEvent_t h2dDone[2] = {nullptr, nullptr};
  std::vector<SlotRes> slots;
  // time series
  float *stage[2] = {nullptr, nullptr};
  cudaEvent_t ready[2] = {nullptr, nullptr}, freeEv[2] = {nullptr, nullptr}, applied = nullptr;
  bool freeSet[2] = {false, false};
};

}  // namespace

struct icb_group_s {
  std::vector<RankState> local;
  int world = 1;
  bool ownsComms = true;
  std::string err;
  int width = 0, height = 0, tilesX = 0, tilesY = 0, ntiles = 0, inFlight = 1;
  bool begun = false;
  std::vector<int32_t> order, tiles, offsets;
  long frameNo = 0;
  double tHost = 0, tWait = 0;
  long statFrames = 0;
  // time series
  size_t slab = 0, nScalars = 0;
  long uploads = 0, swaps = 0;
  bool noLinesCopy = std::getenv("ICB_NO_LINES_COPY") != nullptr;     // experiment knob: time series swapped in place, behind every frame
};

#define GCK(call)                                                                                  \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      g->err = std::string(#call) + ": " + cudaGetErrorString(e_);                                 \
      return ICB_ERR_CUDA;                                                                         \
    }                                                                                              \
  } while (0)
#define GNC(call)                                                                                  \
  do {                                                                                             \
    ncclResult_t r_ = (call);                                                                      \
    if (r_ != 0) {                                                                                 \
      g->err = std::string(#call) + ": " + g_nccl.GetErrorString(r_);                              \
      return ICB_ERR_NCCL;                                                                         \
    }                                                                                              \
  } while (0)
// for calls made between ncclGroupStart and ncclGroupEnd: the group is closed before the error is returned, so that a failed
// call does not leave the thread inside an open NCCL group (every later NCCL call would silently be queued into it)
#define GNG(call)                                                                                  \
  do {                                                                                             \
    ncclResult_t r_ = (call);                                                                      \
    if (r_ != 0) {                                                                                 \
      g->err = std::string(#call) + ": " + g_nccl.GetErrorString(r_);                              \
      g_nccl.GroupEnd();                                                                           \
      return ICB_ERR_NCCL;                                                                         \
    }                                                                                              \
  } while (0)
#define GRC(call)                                                                                  \
  do {                                                                                             \
    int rc_ = (call);                                                                              \
    if (rc_ < 0) { g->err = std::string(#call) + ": " + icb_last_error(R.h); return rc_; }       \
  } while (0)

static int initRankStreams(icb_group_s *g, RankState &R)
{
  GCK(cudaSetDevice(R.h->device));
  // the comm stream's few NCCL CTAs go ahead of the frames' persistent kernels whenever CTA slots free up
  {
    int prLo = 0, prHi = 0;
    GCK(cudaDeviceGetStreamPriorityRange(&prLo, &prHi));
    GCK(cudaStreamCreateWithPriority(&R.commStream, cudaStreamNonBlocking, getenv("ICB_COMM_NO_PRIORITY") ? prLo : prHi));
  }
  GCK(cudaStreamCreateWithFlags(&R.copyStream, cudaStreamNonBlocking));
  GCK(cudaStreamCreateWithFlags(&R.agStream, cudaStreamNonBlocking));
  return ICB_OK;
}

static void freeSlots(RankState &R)
{
  cudaSetDevice(R.h->device);
  for (SlotRes &s : R.slots) {
    cudaFree(s.compact); cudaFree(s.costsDev); cudaFree(s.fb); cudaFree(s.idsDev);
    cudaFreeHost(s.tilesPinned); cudaFreeHost(s.costsHost); cudaFreeHost(s.idsPinned);
    for (cudaEvent_t e : {s.evRendered, s.evComm, s.evCosts, s.evFrame})
      if (e) cudaEventDestroy(e);
  }
  R.slots.clear();
}

extern "C" {

int icb_partition_tiles(int tilesX, int tilesY, const int32_t *costs, int nRanks, int32_t *tileIds, int32_t *offsets)
{
  if (tilesX < 1 || tilesY < 1 || nRanks < 1 || !tileIds || !offsets || (int64_t)tilesX * tilesY > (1 << 26)) return ICB_ERR_INVALID;
  std::vector<int32_t> order, tiles, off;
  mortonOrder(tilesX, tilesY, order);
  dealTiles(order, costs, nRanks, tiles, off);
  std::memcpy(tileIds, tiles.data(), tiles.size() * sizeof(int32_t));
  std::memcpy(offsets, off.data(), off.size() * sizeof(int32_t));
  return ICB_OK;
}

const char *icb_group_last_error(icb_group g) { return g ? g->err.c_str() : g_groupError.c_str(); }

int icb_group_unique_id(void *id)
{
  if (!id) return ICB_ERR_INVALID;
  if (!loadNccl(g_groupError)) return ICB_ERR_NCCL;
  ncclUniqueId u;
  const ncclResult_t r = g_nccl.GetUniqueId(&u);
  if (r != 0) { g_groupError = std::string("ncclGetUniqueId: ") + g_nccl.GetErrorString(r); return ICB_ERR_NCCL; }
  std::memcpy(id, &u, sizeof(u));
  return ICB_OK;
}

int icb_group_create(icb_handle h, int rank, int world, const void *id, icb_group *out)
{
  if (!h || !out || (!id && world > 1) || world < 1 || rank < 0 || rank >= world) return ICB_ERR_INVALID;
  *out = nullptr;
  if (world > 1 && !loadNccl(g_groupError)) return ICB_ERR_NCCL;        // (a group of one needs no NCCL)
  icb_group_s *g = new icb_group_s();
  g->world = world;
  g->local.resize(1);
  RankState &R = g->local[0];
  R.h = h; R.rank = rank;
  auto fail = [&](int rc) { g_groupError = g->err; delete g; return rc; };
  if (cudaSetDevice(h->device) != cudaSuccess) { g->err = "cudaSetDevice failed"; return fail(ICB_ERR_CUDA); }
  if (world > 1) {
    ncclUniqueId u;
    std::memcpy(&u, id, sizeof(u));
    const ncclResult_t r = g_nccl.CommInitRank(&R.comm, world, u, rank);
    if (r != 0) { g->err = std::string("ncclCommInitRank: ") + g_nccl.GetErrorString(r); return fail(ICB_ERR_NCCL); }
  }
  int rc = initRankStreams(g, R);
  if (rc != ICB_OK) return fail(rc);
  *out = g;
  return ICB_OK;
}

int icb_group_create_local(const icb_handle *handles, int nDev, icb_group *out)
{
  if (!handles || !out || nDev < 1 || nDev > 64) return ICB_ERR_INVALID;
  *out = nullptr;
  if (nDev > 1 && !loadNccl(g_groupError)) return ICB_ERR_NCCL;
  icb_group_s *g = new icb_group_s();
  g->world = nDev;
  g->local.resize((size_t)nDev);
  auto fail = [&](int rc) { g_groupError = g->err; delete g; return rc; };
  std::vector<int> devs((size_t)nDev);
  for (int i = 0; i < nDev; ++i) {
    if (!handles[i]) { g->err = "icb_group_create_local: null handle"; return fail(ICB_ERR_INVALID); }
    g->local[(size_t)i].h = handles[i];
    g->local[(size_t)i].rank = i;
    devs[(size_t)i] = handles[i]->device;
  }
  std::vector<ncclComm_t> comms((size_t)nDev, nullptr);
  if (nDev > 1) {
    const ncclResult_t r = g_nccl.CommInitAll(comms.data(), nDev, devs.data());
    if (r != 0) { g->err = std::string("ncclCommInitAll: ") + g_nccl.GetErrorString(r); return fail(ICB_ERR_NCCL); }
  }
  for (int i = 0; i < nDev; ++i) {
    g->local[(size_t)i].comm = comms[(size_t)i];
    int rc = initRankStreams(g, g->local[(size_t)i]);
    if (rc != ICB_OK) return fail(rc);
  }
  *out = g;
  return ICB_OK;
}

int icb_group_destroy(icb_group g)
{
  if (!g) return ICB_ERR_INVALID;
  for (RankState &R : g->local) {
    cudaSetDevice(R.h->device);
    cudaDeviceSynchronize();
    freeSlots(R);
    for (int s = 0; s < 2; ++s) {
      cudaFree(R.stage[s]);
      if (R.ready[s]) cudaEventDestroy(R.ready[s]);
      if (R.freeEv[s]) cudaEventDestroy(R.freeEv[s]);
    }
    if (R.applied) cudaEventDestroy(R.applied);
    if (R.comm2) g_nccl.CommDestroy(R.comm2);
    if (R.comm) g_nccl.CommDestroy(R.comm);
    if (R.commStream) cudaStreamDestroy(R.commStream);
    if (R.copyStream) cudaStreamDestroy(R.copyStream);
    if (R.agStream) cudaStreamDestroy(R.agStream);
    for (int s = 0; s < 2; ++s) if (R.h2dDone[s]) cudaEventDestroy(R.h2dDone[s]);
  }
  delete g;
  return ICB_OK;
}

int icb_group_size(icb_group g) { return g ? g->world : ICB_ERR_INVALID; }

// ---- scene: forwarded to every local handle ---------------------------------------------------------------------
int icb_group_set_transfunc(icb_group g, const float *rgba, int size, float lo, float hi, float opacityScale)
{
  if (!g) return ICB_ERR_INVALID;
  for (RankState &R : g->local) GRC(icb_set_transfunc(R.h, rgba, size, lo, hi, opacityScale));
  return ICB_OK;
}

int icb_group_set_camera(icb_group g, const float org[3], const float dir_00[3], const float dir_du[3], const float dir_dv[3])
{
  if (!g) return ICB_ERR_INVALID;
  for (RankState &R : g->local) GRC(icb_set_camera(R.h, org, dir_00, dir_du, dir_dv));
  return ICB_OK;
}

int icb_group_set_params(icb_group g, float unitDistance, float dt, const float ambientColor[3], float ambientRadiance, int accumID,
                         unsigned flags, float termAlpha)
{
  if (!g) return ICB_ERR_INVALID;
  for (RankState &R : g->local) GRC(icb_set_params(R.h, unitDistance, dt, ambientColor, ambientRadiance, accumID, flags, termAlpha));
  return ICB_OK;
}

// The grid is built ONCE, by the group's rank 0 (host preprocessing + upload: icb_set_grid), and the finished device
// arrays go to the other ranks by ncclBroadcast over NVLink -- 105 GB at R2B10 x 90 is a second or two, against
// minutes of host work per rank.  Ranks other than 0 pass records == NULL (their n is ignored).
int icb_group_set_grid(icb_group g, const void *records, size_t n)
{
  if (!g) return ICB_ERR_INVALID;
  const auto t0 = std::chrono::steady_clock::now();
  RankState *root = nullptr;
  for (RankState &R : g->local)
    if (R.rank == 0) root = &R;
  GridMeta meta{};
  if (root) {
    RankState &R = *root;
    if (!records) { g->err = "icb_group_set_grid: rank 0 needs the records"; return ICB_ERR_INVALID; }
    GRC(icb_set_grid(R.h, records, n));
    icb_context *h = R.h;
    GCK(cudaSetDevice(h->device));
    // the column table of the majorant builder travels with the grid (on a single GPU it waits on the host for first use)
    if (!h->dColTrig && !h->colTrigHost.empty()) {
      GCK(cudaMalloc((void **)&h->dColTrig, h->colTrigHost.size() * sizeof(float)));
      GCK(cudaMemcpy(h->dColTrig, h->colTrigHost.data(), h->colTrigHost.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
    meta.nRec = h->nRec; meta.nCol = h->nCol; meta.nVert = h->nVert;
    meta.nBinOff = (uint64_t)6 * h->grid.binN * h->grid.binN + 1;
    meta.nBinList = (uint64_t)h->stats.numBinEntries;
    meta.nGlobal = (uint64_t)h->grid.nGlobal;
    meta.nSeed = (uint64_t)12 * h->grid.binN * h->grid.binN;
    meta.nVtxOff = h->nVert + 1; meta.nVtxCols = h->nCol * 3; meta.nColLayers = h->nCol; meta.nColTrig = h->nCol * kColTrigStride;
    meta.binN = h->grid.binN; meta.recOfOrig0 = h->grid.recOfOrig0; meta.maxColLayers = h->vtxStride - 2;
    meta.cullLo = h->grid.cullLo; meta.cullHi = h->grid.cullHi; meta.marginUsed = h->marginUsed;
    meta.numGeneric = h->stats.numGenericColumns; meta.bounds = h->bounds;
    h->colTrigHost.clear(); h->colTrigHost.shrink_to_fit();
  }
  if (g->world == 1) return ICB_OK;
  // 1. the meta block
  std::vector<GridMeta *> dMeta(g->local.size(), nullptr);
  for (size_t i = 0; i < g->local.size(); ++i) {
    RankState &R = g->local[i];
    GCK(cudaSetDevice(R.h->device));
    GCK(cudaMalloc((void **)&dMeta[i], sizeof(GridMeta)));
    if (R.rank == 0) GCK(cudaMemcpy(dMeta[i], &meta, sizeof(GridMeta), cudaMemcpyHostToDevice));
  }
  GNC(g_nccl.GroupStart());
  for (size_t i = 0; i < g->local.size(); ++i) {
    RankState &R = g->local[i];
    GNG(g_nccl.Broadcast(dMeta[i], dMeta[i], sizeof(GridMeta), ncclUint8, 0, R.comm, R.commStream));
  }
  GNC(g_nccl.GroupEnd());
  for (size_t i = 0; i < g->local.size(); ++i) {
    RankState &R = g->local[i];
    GCK(cudaSetDevice(R.h->device));
    GCK(cudaStreamSynchronize(R.commStream));
    if (!root && i == 0) GCK(cudaMemcpy(&meta, dMeta[i], sizeof(GridMeta), cudaMemcpyDeviceToHost));
    GCK(cudaFree(dMeta[i]));
  }
  // 2. the arrays
  struct Arr { void **p; size_t bytes; };
  auto arrays = [&](icb_context *h) {
    return std::vector<Arr>{
        {(void **)&h->dRec, meta.nRec * sizeof(RecBlock)}, {(void **)&h->dLines, meta.nRec * kLineStride * sizeof(float)},
        {(void **)&h->dOrigToInt, meta.nRec * 4}, {(void **)&h->dCol, meta.nCol * sizeof(ColBlock)},
        {(void **)&h->dBinOff, meta.nBinOff * 4}, {(void **)&h->dBinList, meta.nBinList * 4},
        {(void **)&h->dGlobal, meta.nGlobal * 4}, {(void **)&h->dSeed, meta.nSeed * 4},
        {(void **)&h->dVtxOff, meta.nVtxOff * 4}, {(void **)&h->dVtxCols, meta.nVtxCols * 4},
        {(void **)&h->dColLayers, meta.nColLayers * 4}, {(void **)&h->dColTrig, meta.nColTrig * 4}};
  };
  for (RankState &R : g->local) {
    if (R.rank == 0) continue;
    icb_context *h = R.h;
    GCK(cudaSetDevice(h->device));
    GCK(cudaDeviceSynchronize());
    icbFreeGrid(h);
    for (const Arr &a : arrays(h)) GCK(cudaMalloc(a.p, std::max<size_t>(a.bytes, 4)));
  }
  const size_t nArr = arrays(g->local[0].h).size();
  for (size_t k = 0; k < nArr; ++k) {
    GNC(g_nccl.GroupStart());
    for (RankState &R : g->local) {
      const Arr a = arrays(R.h)[k];
      if (a.bytes > 0) GNG(g_nccl.Broadcast(*a.p, *a.p, a.bytes, ncclUint8, 0, R.comm, R.commStream));
    }
    GNC(g_nccl.GroupEnd());
  }
  const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  for (RankState &R : g->local) {
    GCK(cudaSetDevice(R.h->device));
    GCK(cudaStreamSynchronize(R.commStream));
    if (R.rank != 0) {
      int rc = icbAdoptGrid(R.h, meta, secs);
      if (rc != ICB_OK) { g->err = std::string("icb_group_set_grid: ") + icb_last_error(R.h); return rc; }
    }
  }
  return ICB_OK;
}

// ---- frames ------------------------------------------------------------------------------------------------------
int icb_group_begin(icb_group g, int width, int height, int framesInFlight)
{
  if (!g || width < 1 || height < 1 || framesInFlight < 1 || framesInFlight > ICB_FRAME_SLOTS) return ICB_ERR_INVALID;
  // one compositing arithmetic per group (icb_set_ray_split): chunked and plain frames differ in fp32 rounding
  for (RankState &R : g->local) {
    const icb_context *a = g->local[0].h, *b = R.h;
    const bool chunkedA = a->splitMode != ICB_SPLIT_NEVER, chunkedB = b->splitMode != ICB_SPLIT_NEVER;
    if (chunkedA != chunkedB || (chunkedA && a->chunkLog2 != b->chunkLog2)) {
      g->err = "icb_group_begin: the group's handles disagree on icb_set_ray_split (mode NEVER vs ALWAYS/AUTO, or chunk size): "
               "their tiles would not assemble into the single-GPU frame bit for bit";
      return ICB_ERR_INVALID;
    }
  }
  g->width = width; g->height = height;
  g->tilesX = (width + kTile - 1) / kTile; g->tilesY = (height + kTile - 1) / kTile;
  g->ntiles = g->tilesX * g->tilesY;
  g->inFlight = framesInFlight;
  mortonOrder(g->tilesX, g->tilesY, g->order);
  dealTiles(g->order, nullptr, g->world, g->tiles, g->offsets);
  g->frameNo = 0; g->tHost = g->tWait = 0; g->statFrames = 0;
  const size_t nt = (size_t)g->ntiles;
  for (RankState &R : g->local) {
    GCK(cudaSetDevice(R.h->device));
    GCK(cudaDeviceSynchronize());
    freeSlots(R);
    R.slots.resize((size_t)framesInFlight);
    for (SlotRes &s : R.slots) {
      GCK(cudaMalloc((void **)&s.compact, nt * kTilePixels * sizeof(uint32_t)));
      GCK(cudaMemset(s.compact, 0, nt * kTilePixels * sizeof(uint32_t)));
      GCK(cudaMalloc((void **)&s.costsDev, nt * sizeof(int32_t)));
      GCK(cudaMemset(s.costsDev, 0, nt * sizeof(int32_t)));
      GCK(cudaHostAlloc((void **)&s.tilesPinned, nt * sizeof(int32_t), cudaHostAllocDefault));
      GCK(cudaHostAlloc((void **)&s.costsHost, nt * sizeof(int32_t), cudaHostAllocDefault));
      if (R.rank == 0) {
        GCK(cudaMalloc((void **)&s.fb, (size_t)width * height * sizeof(uint32_t)));
        GCK(cudaMemset(s.fb, 0, (size_t)width * height * sizeof(uint32_t)));
        GCK(cudaMalloc((void **)&s.idsDev, nt * sizeof(int32_t)));
        GCK(cudaHostAlloc((void **)&s.idsPinned, nt * sizeof(int32_t), cudaHostAllocDefault));
      }
      for (cudaEvent_t *e : {&s.evRendered, &s.evComm, &s.evCosts, &s.evFrame}) GCK(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
      s.costsPending = false;
    }
  }
  g->begun = true;
  return ICB_OK;
}

// One sort-first frame with the handles' current camera / parameters, enqueued without waiting for the device.
// fbPointer_pinned (rank 0 only, may be NULL): page-locked host memory that receives the RGBA8 frame.
// Returns the slot (0 .. framesInFlight-1) the frame occupies, or a negative icb_error.
int icb_group_render(icb_group g, uint32_t *fbPointer_pinned)
{
  if (!g) return ICB_ERR_INVALID;
  if (!g->begun) { g->err = "icb_group_render: icb_group_begin has not been called"; return ICB_ERR_INVALID; }
  const auto tIn = std::chrono::steady_clock::now();
  const long f = g->frameNo;
  const int b = (int)(f % g->inFlight);
  const int W = g->width, H = g->height, nt = g->ntiles;
  // costs of the frame that occupied this slot last: recorded inFlight frames ago, they have (all but certainly) arrived
  if (f >= g->inFlight) {
    for (RankState &R : g->local) {
      GCK(cudaSetDevice(R.h->device));
      GCK(cudaEventSynchronize(R.slots[(size_t)b].evCosts));
    }
    g->tWait += std::chrono::duration<double>(std::chrono::steady_clock::now() - tIn).count();
    dealTiles(g->order, g->local[0].slots[(size_t)b].costsHost, g->world, g->tiles, g->offsets);
  }
  const std::vector<int32_t> &off = g->offsets;
  // 1. render this rank's tiles into its slice of the compact buffer
  for (RankState &R : g->local) {
    icb_context *h = R.h;
    SlotRes &s = R.slots[(size_t)b];
    GCK(cudaSetDevice(h->device));
    GRC(icb_set_frame_slot(h, b));
    cudaStream_t st = h->slots[b].stream;
    const int n = off[(size_t)R.rank + 1] - off[(size_t)R.rank];
    std::memcpy(s.tilesPinned, g->tiles.data() + off[(size_t)R.rank], (size_t)n * sizeof(int32_t));
    GRC(icb_set_tiles_async(h, s.tilesPinned, n, st));
    if (n > 0) {
      GRC(icb_render_tiles(h, s.compact + (size_t)off[(size_t)R.rank] * kTilePixels, nullptr, W, H, st));
      GRC(icb_tile_costs(h, s.costsDev, nt, st));
    } else GCK(cudaMemsetAsync(s.costsDev, 0, (size_t)nt * sizeof(int32_t), st));      // (more ranks than blocks of tiles)
    GCK(cudaEventRecord(s.evRendered, st));
    GCK(cudaStreamWaitEvent(R.commStream, s.evRendered, 0));
  }
  // 2. tiles to rank 0, costs to everybody (the comm stream: NCCL work of consecutive frames runs in frame order)
  if (g->world > 1) {
    GNC(g_nccl.GroupStart());
    for (RankState &R : g->local) {
      SlotRes &s = R.slots[(size_t)b];
      if (R.rank == 0) {
        for (int r = 1; r < g->world; ++r) {
          const size_t cnt = (size_t)(off[(size_t)r + 1] - off[(size_t)r]) * kTilePixels;
          if (cnt > 0) GNG(g_nccl.Recv(s.compact + (size_t)off[(size_t)r] * kTilePixels, cnt, ncclInt32, r, R.comm, R.commStream));
        }
      } else {
        const size_t cnt = (size_t)(off[(size_t)R.rank + 1] - off[(size_t)R.rank]) * kTilePixels;
        if (cnt > 0) GNG(g_nccl.Send(s.compact + (size_t)off[(size_t)R.rank] * kTilePixels, cnt, ncclInt32, 0, R.comm, R.commStream));
      }
      GNG(g_nccl.AllReduce(s.costsDev, s.costsDev, (size_t)nt, ncclInt32, ncclSum, R.comm, R.commStream));
    }
    GNC(g_nccl.GroupEnd());
  }
  // 3. costs home (next partition; the kernel's heaviest-first order), frame assembled on rank 0
  for (RankState &R : g->local) {
    icb_context *h = R.h;
    SlotRes &s = R.slots[(size_t)b];
    GCK(cudaSetDevice(h->device));
    cudaStream_t st = h->slots[b].stream;
    GCK(cudaEventRecord(s.evComm, R.commStream));
    GCK(cudaStreamWaitEvent(st, s.evComm, 0));
    if (g->world > 1) GRC(icb_set_tile_costs(h, s.costsDev, nt, st));
    if (R.rank == 0) {
      std::memcpy(s.idsPinned, g->tiles.data(), (size_t)nt * sizeof(int32_t));       // tile id of every compact tile, rank-major
      GCK(cudaMemcpyAsync(s.idsDev, s.idsPinned, (size_t)nt * sizeof(int32_t), cudaMemcpyHostToDevice, st));
      GRC(icb_scatter_tiles(h, s.compact, nullptr, s.idsDev, nt, s.fb, nullptr, W, H, st));
    }
    // (behind the copies FROM this slot's pinned lists: the host rewrites them once this event has passed)
    GCK(cudaMemcpyAsync(s.costsHost, s.costsDev, (size_t)nt * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    GCK(cudaEventRecord(s.evCosts, st));
    if (R.rank == 0 && fbPointer_pinned)
      GCK(cudaMemcpyAsync(fbPointer_pinned, s.fb, (size_t)W * H * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    GCK(cudaEventRecord(s.evFrame, st));
  }
  g->frameNo = f + 1;
  g->statFrames += 1;
  g->tHost += std::chrono::duration<double>(std::chrono::steady_clock::now() - tIn).count();
  return b;
}

// returns once the frame of `slot` (slot < 0: every slot) has left the device: rank 0's frame is assembled (and in host
// memory if a pinned buffer was given)
int icb_group_wait(icb_group g, int slot)
{
  if (!g || !g->begun || slot >= g->inFlight) return ICB_ERR_INVALID;
  for (RankState &R : g->local) {
    GCK(cudaSetDevice(R.h->device));
    for (int b = 0; b < g->inFlight; ++b)
      if (slot < 0 || slot == b) GCK(cudaEventSynchronize(R.slots[(size_t)b].evFrame));
  }
  return ICB_OK;
}

// rank 0's assembled frame of `slot` (device memory owned by the group; NULL on groups without rank 0)
int icb_group_frame(icb_group g, int slot, uint32_t **fbPointer_dev)
{
  if (!g || !g->begun || !fbPointer_dev || slot < 0 || slot >= g->inFlight) return ICB_ERR_INVALID;
  *fbPointer_dev = nullptr;
  for (RankState &R : g->local)
    if (R.rank == 0) *fbPointer_dev = R.slots[(size_t)slot].fb;
  return ICB_OK;
}

int icb_group_stats(icb_group g, icb_group_stats_t *out)
{
  if (!g || !out) return ICB_ERR_INVALID;
  std::memset(out, 0, sizeof(*out));
  out->world = g->world; out->localRanks = (int)g->local.size(); out->framesInFlight = g->inFlight;
  out->frames = g->statFrames;
  out->hostMsPerFrame = g->statFrames ? 1e3 * g->tHost / g->statFrames : 0.0;
  out->waitMsPerFrame = g->statFrames ? 1e3 * g->tWait / g->statFrames : 0.0;
  return ICB_OK;
}

int icb_group_reset_stats(icb_group g)
{
  if (!g) return ICB_ERR_INVALID;
  g->tHost = g->tWait = 0; g->statFrames = 0;
  return ICB_OK;
}

// ---- time series (BASELINE configs[3]): scalars streamed from pinned host memory --------------------------------
// Each rank uploads 1/world of the timestep over its own PCIe link, the equal slabs (the last one padded) are exchanged
// by ONE ncclAllGather on a communicator of its own (beside the tile gather), double-buffered against rendering.
// values_pinned: n*32 floats laid out like icb_set_scalars, page-locked; must stay untouched until the next call
// but one.  Asynchronous.
int icb_group_scalar_slab(icb_group g, size_t n, size_t *firstFloat, size_t *numFloats)
{
  if (!g || !firstFloat || !numFloats || g->local.empty()) return ICB_ERR_INVALID;
  const size_t total = n * (size_t)kHeightStride;
  const size_t slab = ((total + (size_t)g->world - 1) / (size_t)g->world + 3) & ~(size_t)3;
  size_t lo = total, hi = 0;
  for (RankState &R : g->local) {
    const size_t a = std::min(total, slab * (size_t)R.rank), e = std::min(total, slab * ((size_t)R.rank + 1));
    lo = std::min(lo, a); hi = std::max(hi, e);
  }
  *firstFloat = std::min(lo, hi);
  *numFloats = hi > lo ? hi - lo : 0;
  return ICB_OK;
}

int icb_group_stream_scalars(icb_group g, const float *values_pinned, size_t n)
{
  if (!g || !values_pinned) return ICB_ERR_INVALID;
  // two stage buffers: a third upload would overwrite the oldest timestep that is staged but not yet swapped in
  if (g->uploads - g->swaps >= 2) {
    g->err = "icb_group_stream_scalars: two streamed timesteps are already pending; call icb_group_swap_scalars first";
    return ICB_ERR_INVALID;
  }
  const size_t total = n * (size_t)kHeightStride;
  const size_t slab = ((total + (size_t)g->world - 1) / (size_t)g->world + 3) & ~(size_t)3;
  const int sb = (int)(g->uploads & 1);
  for (RankState &R : g->local) {
    icb_context *h = R.h;
    if (n != h->nRec) { g->err = "icb_group_stream_scalars: record count differs from the grid"; return ICB_ERR_INVALID; }
    GCK(cudaSetDevice(h->device));
    if (g->slab != slab || !R.stage[0]) {
      GCK(cudaDeviceSynchronize());
      for (int s = 0; s < 2; ++s) {
        cudaFree(R.stage[s]); R.stage[s] = nullptr;
        GCK(cudaMalloc((void **)&R.stage[s], slab * (size_t)g->world * sizeof(float)));
        if (!R.ready[s]) GCK(cudaEventCreateWithFlags(&R.ready[s], cudaEventDisableTiming));
        if (!R.freeEv[s]) GCK(cudaEventCreateWithFlags(&R.freeEv[s], cudaEventDisableTiming));
        if (!R.h2dDone[s]) GCK(cudaEventCreateWithFlags(&R.h2dDone[s], cudaEventDisableTiming));
        R.freeSet[s] = false;
      }
      if (!R.applied) GCK(cudaEventCreateWithFlags(&R.applied, cudaEventDisableTiming));
    }
  }
  g->slab = slab; g->nScalars = n;
  // a communicator of its own for the slab exchange (NCCL >= 2.18: ncclCommSplit), else the frame communicator
  if (g->world > 1 && g_nccl.CommSplit && !g->local[0].comm2) {
    GNC(g_nccl.GroupStart());
    for (RankState &R : g->local) GNG(g_nccl.CommSplit(R.comm, 0, R.rank, &R.comm2, nullptr));
    GNC(g_nccl.GroupEnd());
  }
  for (RankState &R : g->local) {
    GCK(cudaSetDevice(R.h->device));
    if (R.freeSet[sb]) GCK(cudaStreamWaitEvent(R.copyStream, R.freeEv[sb], 0));      // the scatter that last read this buffer has run
    const size_t a = std::min(total, slab * (size_t)R.rank), e = std::min(total, slab * ((size_t)R.rank + 1));
    if (e > a) GCK(cudaMemcpyAsync(R.stage[sb] + a, values_pinned + a, (e - a) * sizeof(float), cudaMemcpyHostToDevice, R.copyStream));
    if (g->world > 1) {
      GCK(cudaEventRecord(R.h2dDone[sb], R.copyStream));
      GCK(cudaStreamWaitEvent(R.agStream, R.h2dDone[sb], 0));
    }
  }
  if (g->world > 1) {
    GNC(g_nccl.GroupStart());
    for (RankState &R : g->local)
      GNG(g_nccl.AllGather(R.stage[sb] + slab * (size_t)R.rank, R.stage[sb], slab, ncclFloat32, R.comm2 ? R.comm2 : R.comm, R.agStream));
    GNC(g_nccl.GroupEnd());
  }
  for (RankState &R : g->local) {
    GCK(cudaSetDevice(R.h->device));
    GCK(cudaEventRecord(R.ready[sb], g->world > 1 ? R.agStream : R.copyStream));
  }
  g->uploads += 1;
  return ICB_OK;
}

// Swap the grid's scalars for the oldest streamed timestep not yet applied: after every frame in flight (they still read
// the old values), before every frame enqueued afterwards.  Stream-ordered; no host wait.
int icb_group_swap_scalars(icb_group g)
{
  if (!g) return ICB_ERR_INVALID;
  if (g->swaps >= g->uploads) { g->err = "icb_group_swap_scalars: no streamed timestep pending"; return ICB_ERR_INVALID; }
  const int sb = (int)(g->swaps & 1);
  for (RankState &R : g->local) {
    icb_context *h = R.h;
    GCK(cudaSetDevice(h->device));
    cudaStream_t st = h->ownStream;
    // Nothing but the lines depends on the scalars with the cell sampler and the marcher (no vertex table, no majorants,
    // no transparency masks): the timestep is scattered into a SECOND copy of the lines while the frames in flight go on
    // reading the current one, and the copies change roles -- the swap waits only for frames that read the copy it
    // overwrites (two swaps old), so rendering and the scalar exchange overlap.  Otherwise: in place, behind every frame.
    const bool twoCopies = h->sampler == ICB_SAMPLER_CELL && h->integrator == ICB_INTEGRATOR_MARCH && !g->noLinesCopy;
    if (twoCopies) {
      const size_t bytes = h->nRec * (size_t)kLineStride * sizeof(float);
      if (!h->dLinesAlt) {
        if (cudaMalloc((void **)&h->dLinesAlt, bytes) != cudaSuccess) { cudaGetLastError(); h->dLinesAlt = nullptr; }
        else GCK(cudaMemcpyAsync(h->dLinesAlt, h->dLines, bytes, cudaMemcpyDeviceToDevice, st));     // (the heights)
      }
    }
    if (twoCopies && h->dLinesAlt) {
      const int target = 1 - h->linesBuf;
      float *tl = target ? h->dLinesAlt : h->dLines;
      for (FrameSlot &sl : h->slots)
        if (sl.bufUsed[target]) GCK(cudaStreamWaitEvent(st, sl.evBuf[target], 0));
      GCK(cudaStreamWaitEvent(st, R.ready[sb], 0));
      GCK(launchScatterScalars(reinterpret_cast<const float *>(R.stage[sb]), h->dOrigToInt, tl, g->nScalars, st));
      h->linesBuf = target;
      h->grid.lines = tl;
      h->majorDirty = h->shellDirty = 2;
    } else {
      for (FrameSlot &sl : h->slots)
        if (sl.used) GCK(cudaStreamWaitEvent(st, sl.ev1, 0));
      GCK(cudaStreamWaitEvent(st, R.ready[sb], 0));
      GRC(icb_set_scalars_dev(h, R.stage[sb], g->nScalars, st));
    }
    GCK(cudaEventRecord(R.freeEv[sb], st));
    R.freeSet[sb] = true;
    GCK(cudaEventRecord(R.applied, st));
    for (FrameSlot &sl : h->slots)
      if (sl.stream) GCK(cudaStreamWaitEvent(sl.stream, R.applied, 0));
  }
  g->swaps += 1;
  return ICB_OK;
}

}  // extern "C"
