// tests/emu/locator_emu.cpp -- TEST INFRASTRUCTURE.
// Compiles the device locator/marcher text (icon-ray-tracing_b200/csrc/locator.cuh) for the
// host, with shims for the CUDA intrinsics, so the walk/exact-query logic and the host grid
// builder can be unit-tested against the oracle on machines without a GPU.  This harness is
// never linked into libicon_b200.so and is not a fallback of the product.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

struct float4 { float x, y, z, w; };
struct int4 { int x, y, z, w; };
struct int2 { int x, y; };
struct float2 { float x, y; };
static inline float2 make_float2(float x, float y) { return {x, y}; }
static inline int2 make_int2(int x, int y) { return {x, y}; }
static inline float4 make_float4(float x, float y, float z, float w) { return {x, y, z, w}; }
#define __device__
#define __forceinline__ inline
#define __noinline__
#define __restrict__
static inline float __fmul_rn(float a, float b) { return a * b; }
static inline float __fadd_rn(float a, float b) { return a + b; }
static inline float __fsub_rn(float a, float b) { return a - b; }
static inline float __fdiv_rn(float a, float b) { return a / b; }
static inline float __fsqrt_rn(float a) { return sqrtf(a); }
static inline float __fdividef(float a, float b) { return a / b; }
static inline float __fmaf_rn(float a, float b, float c) { return fmaf(a, b, c); }
static inline double __fma_rn(double a, double b, double c) { return fma(a, b, c); }
template <typename T> static inline T __ldg(const T *p) { return *p; }
static inline int __float_as_int(float f) { int i; std::memcpy(&i, &f, 4); return i; }
static inline float __int_as_float(int i) { float f; std::memcpy(&f, &i, 4); return f; }
using std::max;
using std::min;

#include "grid_build.h"
#include "locator.cuh"

using namespace icb;

static HostGrid g_host;
static DeviceGrid g_dev;
static std::string g_err;
static int g_lastSkipped = 0;
static int g_logTable = 0;      // 1: the ray-gens read logf(1-u) from the host-built table (round 1's form), 0: they compute it (the default)

extern "C" {

const char *emu_error(void) { return g_err.c_str(); }

int emu_build(const void *records, size_t n, float marginMeters, int maxWalk)
{
  // the way icb_set_grid builds: without lines, which are then streamed chunk by chunk (fillLines); here the chunks
  // (deliberately ragged) are assembled on the host, and checked against the builder's own whole-array form
  int rc = buildGrid(records, n, marginMeters, g_host, g_err, false);
  if (rc != 0) return rc;
  {
    HostGrid whole;
    rc = buildGrid(records, n, marginMeters, whole, g_err);
    if (rc != 0) return rc;
    std::vector<int32_t> intToOrig(n);
    for (size_t i = 0; i < n; ++i) intToOrig[(size_t)g_host.origToInt[i]] = (int32_t)i;
    g_host.lines.assign(n * (size_t)kLineStride, -1.f);
    const size_t chunk = n / 3 + 1;
    for (size_t a = 0; a < n; a += chunk) fillLines(records, intToOrig.data(), a, std::min(n, a + chunk), g_host.lines.data() + a * kLineStride);
    if (std::memcmp(g_host.lines.data(), whole.lines.data(), whole.lines.size() * sizeof(float)) != 0 || whole.lines.size() != g_host.lines.size()) {
      g_err = "fillLines differs from buildGrid's lines";
      return -1;
    }
  }
  g_dev.rec = g_host.rec.data(); g_dev.lines = g_host.lines.data();
  g_dev.col = g_host.col.data();
  g_dev.binOff = g_host.binOff.data(); g_dev.binList = g_host.binList.data(); g_dev.binN = g_host.binN; g_dev.seed = g_host.seed.data();
  g_dev.globalRecs = g_host.globalRecs.data(); g_dev.nGlobal = (int)g_host.globalRecs.size(); g_dev.recOfOrig0 = g_host.origToInt[0];
  g_dev.nRec = (int)g_host.rec.size(); g_dev.nCol = (int)g_host.col.size(); g_dev.maxWalk = maxWalk;
  g_dev.sampler = 0; g_dev.vtxScalars = nullptr; g_dev.vtxStride = 0; g_dev.prismMask = nullptr;
  g_dev.cullLo = g_host.cullLo; g_dev.cullHi = g_host.cullHi;
  return 0;
}

// sampler 1: prism (vertex scalars computed by the device function vertexScalarAt, on the host)
static std::vector<float> g_vtx;
void emu_set_sampler(int sampler)
{
  g_dev.sampler = sampler;
  if (!sampler) return;
  const int stride = g_host.maxColLayers + 2;
  g_vtx.assign((size_t)g_host.numVertices * stride, 0.f);
  for (int64_t v = 0; v < g_host.numVertices; ++v)
    for (int j = 0; j < stride; ++j)
      g_vtx[(size_t)v * stride + j] = vertexScalarAt(g_dev.rec, g_dev.lines, g_dev.col, g_host.colLayers.data(), g_host.vtxOff.data(),
                                                     g_host.vtxCols.data(), (size_t)v, j);
  g_dev.vtxScalars = g_vtx.data();
  g_dev.vtxStride = stride;
}

void emu_info(int64_t *out /*8*/)
{
  out[0] = (int64_t)g_host.rec.size(); out[1] = (int64_t)g_host.col.size(); out[2] = g_host.numVertices;
  out[3] = g_host.binN; out[4] = (int64_t)g_host.binList.size(); out[5] = g_host.numGeneric;
  int64_t nb = 0; for (auto &c : g_host.col) for (int k = 0; k < 3; ++k) nb += c.nbr[k] < 0;
  out[6] = nb; out[7] = (int64_t)g_host.globalRecs.size();
}

void emu_bounds(icb_bounds *b) { *b = g_host.bounds; }

// flags: 1 = walk (state carried across consecutive positions, restarted every `chunk`), 2 = exact only
void emu_locate(const float *pos, size_t n, int *ids, float *values, unsigned flags, int chunk, uint64_t *counters /*4*/)
{
  WalkState s;
  resetState(s);
  float L[kLineStride] = {0};
  unsigned nSlow = 0, nExact = 0;
  uint64_t hits = 0;
  for (size_t i = 0; i < n; ++i) {
    if (!(flags & 1u) || (chunk > 0 && i % (size_t)chunk == 0)) resetState(s);
    const float x = pos[3 * i], y = pos[3 * i + 1], z = pos[3 * i + 2];
    const float r = __fsqrt_rn(dot3(x, y, z, x, y, z));
    bool hit;
    if (flags & 2u) { ++nExact; hit = exactLocate(g_dev, x, y, z, r, s, L); }
    else hit = locateStep(g_dev, x, y, z, r, s, L, nSlow, nExact);
    if (hit) {
      ids[i] = s.orig; ++hits;
      float v = L[kHeightStride + s.layer];
      if (g_dev.sampler != 0 && !prismSample(g_dev, s, L, x, y, z, v)) v = NAN;
      values[i] = v;
    }
    else { ids[i] = -1; values[i] = 0.f; }
  }
  if (counters) { counters[0] = n; counters[1] = hits; counters[2] = nSlow; counters[3] = nExact; }
}

int emu_march_pixel(const RenderParams *p, const float *lut, int px, int py, int *ids, int maxIds, float *rgba, int useSkip)
{
  float4 C;
  int n = 0;
  int skipped = 0;
  float L[kLineStride] = {0};
  marchPixel(g_dev, *p, (const float4 *)lut, px, py, ids, maxIds, &C, &n, L, useSkip != 0, &skipped);
  g_lastSkipped = skipped;
  rgba[0] = C.x; rgba[1] = C.y; rgba[2] = C.z; rgba[3] = C.w;
  return n;
}

// one frame of the Woodcock integrator (woodcockPixel, the text the CUDA kernel runs), pixels in row order
void emu_woodcock_frame(const RenderParams *p, const float *lut, unsigned *fb, float *accum, uint64_t *counters /*3*/)
{
  static std::vector<float> table;
  if (g_logTable && table.empty()) buildLogTable(table);
  uint64_t steps = 0, hits = 0, rays = 0;
  for (int py = 0; py < p->height; ++py)
    for (int px = 0; px < p->width; ++px) {
      float L[kLineStride] = {0};
      float4 C;
      unsigned nSteps = 0, nHits = 0;
      if (woodcockPixel(g_dev, *p, (const float4 *)lut, g_logTable ? table.data() : nullptr, px, py, L, C, nSteps, nHits)) {
        writePixel(*p, C, (size_t)px + (size_t)p->width * py, fb, (float4 *)accum);
        ++rays;
      }
      steps += nSteps; hits += nHits;
    }
  if (counters) { counters[0] = steps; counters[1] = hits; counters[2] = rays; }
}

// one frame of the reference's default ray-gen (woodcockGridPixel: dda3 + per-macrocell Woodcock tracking)
void emu_woodcock_grid_frame(const RenderParams *p, const float *lut, const int *dims, const float *maxOp, unsigned *fb, float *accum,
                             uint64_t *counters /*3*/)
{
  static std::vector<float> table;
  if (g_logTable && table.empty()) buildLogTable(table);
  MajorantGrid mg;
  mg.maxOpacities = maxOp;
  mg.occupied = nullptr;
  for (int k = 0; k < 3; ++k) { mg.dims[k] = dims[k]; mg.lo[k] = p->volLo[k]; mg.hi[k] = p->volHi[k]; }
  uint64_t steps = 0, hits = 0, rays = 0;
  for (int py = 0; py < p->height; ++py)
    for (int px = 0; px < p->width; ++px) {
      float L[kLineStride + 4] = {0};     // the slot with its pad words (kSlotMask, kSlotLog)
      float4 C;
      unsigned nSteps = 0, nHits = 0;
      if (woodcockGridPixel(g_dev, *p, (const float4 *)lut, g_logTable ? table.data() : nullptr, mg, px, py, L, C, nSteps, nHits)) {
        writePixel(*p, C, (size_t)px + (size_t)p->width * py, fb, (float4 *)accum);
        ++rays;
      }
      steps += nSteps; hits += nHits;
    }
  if (counters) { counters[0] = steps; counters[1] = hits; counters[2] = rays; }
}

// one frame of the same ray-gen in SPHERE_ACCEL_MODE (woodcockSpherePixel: sdda over the shell accel)
void emu_woodcock_sphere_frame(const RenderParams *p, const float *lut, const int *dims, const float *sphLo, const float *sphHi, const float *maxOp,
                               unsigned *fb, float *accum, uint64_t *counters /*3*/)
{
  static std::vector<float> table;
  if (g_logTable && table.empty()) buildLogTable(table);
  MajorantGrid sg;
  sg.maxOpacities = maxOp;
  sg.occupied = nullptr;
  for (int k = 0; k < 3; ++k) { sg.dims[k] = dims[k]; sg.lo[k] = sphLo[k]; sg.hi[k] = sphHi[k]; }
  uint64_t steps = 0, hits = 0, rays = 0;
  for (int py = 0; py < p->height; ++py)
    for (int px = 0; px < p->width; ++px) {
      float L[kLineStride + 4] = {0};
      float4 C;
      unsigned nSteps = 0, nHits = 0;
      if (woodcockSpherePixel(g_dev, *p, (const float4 *)lut, g_logTable ? table.data() : nullptr, sg, px, py, L, C, nSteps, nHits)) {
        writePixel(*p, C, (size_t)px + (size_t)p->width * py, fb, (float4 *)accum);
        ++rays;
      }
      steps += nSteps; hits += nHits;
    }
  if (counters) { counters[0] = steps; counters[1] = hits; counters[2] = rays; }
}

int emu_last_skipped(void) { return g_lastSkipped; }

void emu_set_log_table(int on) { g_logTable = on; }

int emu_sizeof_params(void) { return (int)sizeof(RenderParams); }

}  // extern "C"
