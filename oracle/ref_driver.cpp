// oracle/ref_driver.cpp -- TEST INFRASTRUCTURE, not product code.
//
// C entry points around the *unmodified* reference sources, which are compiled
// where they lie under /root/reference (see oracle/Makefile; output goes to
// oracle/_ref/libicon_ref.so only).  This translation unit #includes the
// reference's icon_rt/deviceCode.cu so that its inline hot-path functions
// (generateRay, sampleVolume -> sample, postClassify, woodcockTracking and the
// two ray-gen programs) are callable from here; common/{fb,pipeline,transfunc}.cu
// are compiled as separate objects with `g++ -xc++`.
//
// Everything in this file is glue except march_pixel(), the CPU statement of the
// new front-to-back integrator -- the reference has no
// ray marcher (deviceCode.cu:149-186 is Woodcock tracking), so the marcher is
// composed from the reference's own functions: generateRay (deviceCode.cu:36-49),
// intersectSphere (ShellAccel.h:34-53), the segment case analysis of sdda
// (ShellAccel.h:85-111), the brute-force locate loop (deviceCode.cu:116-123 ->
// ICONGrid.h:181-208) and postClassify (deviceCode.cu:127-135).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference leg may load the resulting library.

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include <dvr_course-common.h>   // reference: common/ (pulls camera.h, pipeline.h, fb.h)
#include <thread_pool.h>         // reference: common/thread_pool.h
#include <for_each.h>            // reference: common/for_each.h (parallel::for_each, 64x64 tiles)
#include "deviceCode.cu"         // reference: icon_rt/deviceCode.cu, included in place
#include "ref_kernel_shims.h"    // ours: host stand-ins for blockIdx / atomicCAS / ...
#include "hostcode_kernels.gen.h" // reference: icon_rt/hostCode.cu:36-56,205-434 cut out at build time (oracle/Makefile)

// The global launch-parameter block the reference device code reads
// (deviceCode.cu:31; pipeline.h:28-30 DECL_LAUNCH_PARAMS in the CPU build).
namespace icon_rt {
extern "C" {
LaunchParams optixLaunchParams;
}
}

using namespace icon_rt;

namespace {

struct MarchCfg {
  int   jitter;        // 1: reference generateRay with LCG jitter; 0: rnd() := 0
  float dt;            // step length
  float termAlpha;     // early-ray-termination threshold (0.99f in A.3)
  float rInner, rOuter;
  int   chunk;         // > 0: chunked compositing (see march_pixel); 0: the plain statement of A.3
};

inline uint64_t fnv1a_step(uint64_t h, int32_t id)
{
  uint32_t u = (uint32_t)id;
  for (int b = 0; b < 4; ++b) {
    h ^= (u >> (8 * b)) & 0xffu;
    h *= 1099511628211ull;
  }
  return h;
}

// generateRay with the two rnd() draws replaced by 0 (SURVEY.md A.3 "jitter = 0").
// Same expression shape as deviceCode.cu:36-49.
inline Ray generateRayNoJitter(const vec2f screen)
{
  auto &lp = optixLaunchParams;
  vec3f org = lp.camera.org;
  vec3f dir = lp.camera.dir_00 + (screen.u + 0.f) * lp.camera.dir_du + (screen.v + 0.f) * lp.camera.dir_dv;
  dir = normalize(dir);
  if (fabsf(dir.x) < 1e-5f) dir.x = 1e-5f;
  if (fabsf(dir.y) < 1e-5f) dir.y = 1e-5f;
  if (fabsf(dir.z) < 1e-5f) dir.z = 1e-5f;
  return Ray(org, dir, 0.f, 1e10f);
}

// Segment selection of sdda (ShellAccel.h:85-111) + A.3 shrink/clamp.
// Returns number of segments (0..2) written to seg[][2].
inline int shellSegments(const Ray &ray, float rInner, float rOuter, float seg[2][2])
{
  const float sceneEPS = rInner * 1e-6f;
  float t1 = 0.f, t2 = 0.f, t3 = 0.f, t4 = 0.f;
  bool s1 = intersectSphere(ray, rOuter, t1, t4);
  bool s2 = intersectSphere(ray, rInner, t2, t3);
  if (!s1) return 0;               // reference reads uninitialised t1/t4 if s2&&!s1; geometrically impossible
  if (t4 < ray.tmin) return 0;
  box1f ranges[2] = {{INFINITY, -INFINITY}, {INFINITY, -INFINITY}};
  if (s1 && !s2)            ranges[0] = {t1, t4};
  else if (ray.tmin < t2) { ranges[0] = {t1, t2}; ranges[1] = {t3, t4}; }
  else                      ranges[0] = {t3, t4};
  int n = 0;
  const float tlo = fmaxf(ray.tmin, 0.f);
  for (int i = 0; i < 2; ++i) {
    if (ranges[i].empty()) break;   // ShellAccel.h:114
    float lo = fmaxf(ranges[i].lower + sceneEPS, tlo);
    float hi = fminf(ranges[i].upper - sceneEPS, ray.tmax);
    seg[n][0] = lo; seg[n][1] = hi; ++n;
  }
  return n;
}

struct PixelOut {
  vec4f    C;
  uint64_t hash;
  int      nSteps, nHits;
};

// maxTrace>0: write the first maxTrace located ids (-1 = miss) to trace[]
inline PixelOut march_pixel(int x, int y, int W, int H, const MarchCfg &cfg,
                            int32_t *trace, int maxTrace)
{
  auto &lp = optixLaunchParams;
  Random rnd(lp.accumID * W * H + (unsigned)x, (unsigned)y);
  Ray ray = cfg.jitter ? generateRay(vec2f(vec2i(x, y)) + vec2f(.5f), rnd)
                       : generateRayNoJitter(vec2f(vec2i(x, y)) + vec2f(.5f));
  PixelOut out;
  out.C = vec4f(0.f);
  out.hash = 1469598103934665603ull;
  out.nSteps = out.nHits = 0;

  float seg[2][2];
  int nseg = shellSegments(ray, cfg.rInner, cfg.rOuter, seg);
  const vec3f amb = lp.ambientColor;
  bool done = false;
  // Chunked compositing (cfg.chunk > 0; what the CUDA marcher computes): samples are grouped into chunks of cfg.chunk
  // consecutive sample indices per shell segment (cut positions a function of the ray alone), each chunk is composited
  // from zero into its own accumulator, chunks are merged in order with `over` (C += (1-A) c_loc, A += (1-A) a_loc), and
  // early termination is tested on the merged opacity after every contributing sample.  Associativity of `over` makes it
  // the plain recurrence up to fp32 rounding; the device may then march the chunks of a long ray on different lanes.
  const int chunk = cfg.chunk;
  vec4f loc(0.f);
  for (int s = 0; s < nseg && !done; ++s) {
    for (int i = 0;; ++i) {
      float t = seg[s][0] + (float(i) + 0.5f) * cfg.dt;
      if (t >= seg[s][1]) break;
      if (chunk > 0 && i % chunk == 0 && i > 0) {
        const float T = 1.f - out.C.w;
        out.C.x += T * loc.x; out.C.y += T * loc.y; out.C.z += T * loc.z; out.C.w += T * loc.w;
        loc = vec4f(0.f);
      }
      vec3f P = ray.org + ray.dir * t;      // Ray::eval, vecmath.h:1917-1918
      float value = 0.f;
      int id = -1;
      for (unsigned c = 0; c < (unsigned)lp.volume.numCells; ++c) {   // deviceCode.cu:119-122
        if (sample(lp.volume.cells[c], P, value)) { id = (int)c; break; }
      }
      if (trace && out.nSteps < maxTrace) trace[out.nSteps] = id;
      out.hash = fnv1a_step(out.hash, id);
      out.nSteps++;
      if (id < 0) continue;
      out.nHits++;
      vec4f sm = postClassify(lp.transfunc, value);
      float a = 1.f - expf(-sm.w * cfg.dt / lp.unitDistance);
      vec3f col = vec3f(sm.x, sm.y, sm.z) * amb * lp.ambientRadiance;
      if (chunk > 0) {
        float w = (1.f - loc.w) * a;
        loc.x += w * col.x;
        loc.y += w * col.y;
        loc.z += w * col.z;
        loc.w += w;
        if (out.C.w + (1.f - out.C.w) * loc.w >= cfg.termAlpha) { done = true; break; }
      } else {
        float w = (1.f - out.C.w) * a;
        out.C.x += w * col.x;
        out.C.y += w * col.y;
        out.C.z += w * col.z;
        out.C.w += w;
        if (out.C.w >= cfg.termAlpha) { done = true; break; }
      }
    }
    if (chunk > 0) {
      const float T = 1.f - out.C.w;
      out.C.x += T * loc.x; out.C.y += T * loc.y; out.C.z += T * loc.z; out.C.w += T * loc.w;
      loc = vec4f(0.f);
    }
  }
  return out;
}

inline void write_pixel(int pixelID, const vec4f &C)
{
  auto &lp = optixLaunchParams;
  // deviceCode.cu:333-340
  float accum = 1.f / (lp.accumID + 1);
  lp.accumBuffer[pixelID] = lerp(C, lp.accumBuffer[pixelID], accum);
  vec4f accumColor = lp.accumBuffer[pixelID];
  accumColor.r = linear_to_srgb(accumColor.r);
  accumColor.g = linear_to_srgb(accumColor.g);
  accumColor.b = linear_to_srgb(accumColor.b);
  lp.fbPointer[pixelID] = make_rgba(accumColor);
}

// 64x64 image tiles pulled from an atomic counter by N threads -- the dispatch scheme of the
// reference's CPU branch (parallel::for_each, for_each.h:70-85 -> thread_pool::run,
// thread_pool.h:88-107,146-161).  The reference's thread_pool itself is not used for the
// marcher: its start/reset handshake is racy when run() is called back to back (work-item
// counters are reset while late threads still spin on start_threads; observed as a hang in
// this container), so the tiles are dispatched on plain std::threads here.
template <typename F>
void forEachTile(int x0, int x1, int y0, int y1, int nthreads, F func)
{
  if (nthreads <= 0) nthreads = (int)std::thread::hardware_concurrency();   // pipeline.cu:767
  const int T = 64, tx = (x1 - x0 + T - 1) / T, ty = (y1 - y0 + T - 1) / T;
  std::atomic<int> next{0};
  auto work = [&]() {
    for (;;) {
      const int t = next.fetch_add(1);
      if (t >= tx * ty) break;
      const int bx = x0 + (t % tx) * T, by = y0 + (t / tx) * T;
      for (int y = by; y < std::min(by + T, y1); ++y)
        for (int x = bx; x < std::min(bx + T, x1); ++x) func(x, y);
    }
  };
  std::vector<std::thread> th;
  for (int i = 1; i < nthreads; ++i) th.emplace_back(work);
  work();
  for (auto &t : th) t.join();
}

} // namespace

static int g_chunk = 0;

extern "C" {

// chunked compositing for the marcher entry points below (0 = plain A.3 statement)
void ref_set_chunk(int chunk) { g_chunk = chunk; }

int ref_sizeof_cell(void) { return (int)sizeof(ICONCell); }
int ref_hardware_concurrency(void) { return (int)std::thread::hardware_concurrency(); }

// ---- building blocks -------------------------------------------------------

void ref_to_cartesian(const float *sph, int n, float *out)
{
  for (int i = 0; i < n; ++i) {
    vec3f c = toCartesian(vec3f(sph[3 * i], sph[3 * i + 1], sph[3 * i + 2]));
    out[3 * i] = c.x; out[3 * i + 1] = c.y; out[3 * i + 2] = c.z;
  }
}

// the three side planes sample() builds for a record (ICONGrid.h:187-203)
void ref_cell_planes(const void *cells, int n, float *planes /* n*12 */)
{
  const ICONCell *c = (const ICONCell *)cells;
  for (int i = 0; i < n; ++i) {
    const ICONCell &cell = c[i];
    vec3f bv1 = toCartesian({cell.height[0], cell.lat.x, cell.lon.x});
    vec3f bv2 = toCartesian({cell.height[0], cell.lat.y, cell.lon.y});
    vec3f bv3 = toCartesian({cell.height[0], cell.lat.z, cell.lon.z});
    vec3f tv1 = toCartesian({cell.height[cell.numLayers], cell.lat.x, cell.lon.x});
    vec3f tv2 = toCartesian({cell.height[cell.numLayers], cell.lat.y, cell.lon.y});
    vec3f tv3 = toCartesian({cell.height[cell.numLayers], cell.lat.z, cell.lon.z});
    Plane p[3] = {makePlane(bv1, bv2, tv2), makePlane(bv2, bv3, tv3), makePlane(bv3, bv1, tv1)};
    for (int k = 0; k < 3; ++k)
      for (int j = 0; j < 4; ++j) planes[i * 12 + k * 4 + j] = p[k][j];
  }
}

// brute-force locate: first record index accepting pos (deviceCode.cu:116-123)
void ref_locate(const void *cells, int numCells, const float *pos, int npos,
                int32_t *ids, float *values, int nthreads)
{
  const ICONCell *c = (const ICONCell *)cells;
  if (nthreads <= 0) nthreads = (int)std::thread::hardware_concurrency();
  std::atomic<int> next{0};
  auto work = [&]() {
    for (;;) {
      int i = next.fetch_add(1);
      if (i >= npos) break;
      vec3f P(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]);
      float v = 0.f;
      int id = -1;
      for (int k = 0; k < numCells; ++k)
        if (sample(c[k], P, v)) { id = k; break; }
      ids[i] = id;
      values[i] = id >= 0 ? v : 0.f;
    }
  };
  std::vector<std::thread> th;
  for (int t = 1; t < nthreads; ++t) th.emplace_back(work);
  work();
  for (auto &t : th) t.join();
}

void ref_find_height(const void *cell, const float *r, int n, int32_t *layer)
{
  const ICONCell &c = *(const ICONCell *)cell;
  for (int i = 0; i < n; ++i) layer[i] = c.findHeight(r[i]);
}

void ref_post_classify(const float *lut, int size, float lo, float hi, float opacityScale,
                       const float *v, int n, float *rgba)
{
  icon_rt::Transfunc tf;
  tf.valueRange = box1f(lo, hi);
  tf.opacityScale = opacityScale;
  tf.values = (vec4f *)lut;
  tf.size = size;
  for (int i = 0; i < n; ++i) {
    vec4f s = postClassify(tf, v[i]);
    rgba[4 * i] = s.x; rgba[4 * i + 1] = s.y; rgba[4 * i + 2] = s.z; rgba[4 * i + 3] = s.w;
  }
}

void ref_intersect_sphere(const float *org, const float *dir, float radius, int n,
                          int32_t *hit, float *tnear, float *tfar)
{
  for (int i = 0; i < n; ++i) {
    Ray ray(vec3f(org[3 * i], org[3 * i + 1], org[3 * i + 2]),
            vec3f(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]), 0.f, 1e10f);
    float a = 0.f, b = 0.f;
    hit[i] = intersectSphere(ray, radius, a, b) ? 1 : 0;
    tnear[i] = a; tfar[i] = b;
  }
}

void ref_set_camera(const float *org, const float *d00, const float *du, const float *dv)
{
  auto &lp = optixLaunchParams;
  lp.camera.org = vec3f(org[0], org[1], org[2]);
  lp.camera.dir_00 = vec3f(d00[0], d00[1], d00[2]);
  lp.camera.dir_du = vec3f(du[0], du[1], du[2]);
  lp.camera.dir_dv = vec3f(dv[0], dv[1], dv[2]);
}

// generateRay for a list of pixels; seeds as in the ray-gen programs (deviceCode.cu:246-249)
void ref_generate_rays(const int32_t *xy, int n, int W, int H, int accumID, int jitter,
                       float *org, float *dir)
{
  for (int i = 0; i < n; ++i) {
    int x = xy[2 * i], y = xy[2 * i + 1];
    Random rnd(accumID * W * H + (unsigned)x, (unsigned)y);
    Ray ray = jitter ? generateRay(vec2f(vec2i(x, y)) + vec2f(.5f), rnd)
                     : generateRayNoJitter(vec2f(vec2i(x, y)) + vec2f(.5f));
    org[3 * i] = ray.org.x; org[3 * i + 1] = ray.org.y; org[3 * i + 2] = ray.org.z;
    dir[3 * i] = ray.dir.x; dir[3 * i + 1] = ray.dir.y; dir[3 * i + 2] = ray.dir.z;
  }
}

void ref_lcg(unsigned v0, unsigned v1, int n, float *out)
{
  Random rnd(v0, v1);
  for (int i = 0; i < n; ++i) out[i] = rnd();
}

// A.3 segments, and the segments the reference's own sdda() hands to its callback
// (first callback of each radial range), for cross-checking the case analysis.
int ref_shell_segments(const float *org, const float *dir, float tmin, float tmax,
                       float rInner, float rOuter, float *seg4)
{
  Ray ray(vec3f(org[0], org[1], org[2]), vec3f(dir[0], dir[1], dir[2]), tmin, tmax);
  float seg[2][2];
  int n = shellSegments(ray, rInner, rOuter, seg);
  for (int i = 0; i < n; ++i) { seg4[2 * i] = seg[i][0]; seg4[2 * i + 1] = seg[i][1]; }
  return n;
}

int ref_sdda_callbacks(const float *org, const float *dir, float tmin, float tmax,
                       const float *sphLo, const float *sphHi, const int32_t *dims,
                       int maxCb, float *cbT /* maxCb*2 */, int32_t *cbLeaf)
{
  Ray ray(vec3f(org[0], org[1], org[2]), vec3f(dir[0], dir[1], dir[2]), tmin, tmax);
  ShellAccel accel;
  accel.dims = vec3i(dims[0], dims[1], dims[2]);
  accel.sphericalBounds = box3f(vec3f(sphLo[0], sphLo[1], sphLo[2]), vec3f(sphHi[0], sphHi[1], sphHi[2]));
  accel.valueRanges = nullptr;
  accel.maxOpacities = nullptr;
  int n = 0;
  auto cb = [&](const int leafID, float t0, float t1) {
    if (n < maxCb) { cbT[2 * n] = t0; cbT[2 * n + 1] = t1; cbLeaf[n] = leafID; }
    ++n;
    return n < 4096;   // traverse on (bounded)
  };
  sdda(ray, accel, cb, false);
  return n;
}

void ref_srgb_pack(const float *rgba, int n, uint32_t *out)
{
  for (int i = 0; i < n; ++i) {
    vec4f c(rgba[4 * i], rgba[4 * i + 1], rgba[4 * i + 2], rgba[4 * i + 3]);
    c.r = linear_to_srgb(c.r); c.g = linear_to_srgb(c.g); c.b = linear_to_srgb(c.b);
    out[i] = make_rgba(c);
  }
}

// record bounds / data range exactly as hostCode.cu:792-808 (ICONCell::getBounds ICONGrid.h:78-115)
void ref_grid_bounds(const void *cells, int n, float *volLo, float *volHi,
                     float *sphLo, float *sphHi, float *dataRange)
{
  const ICONCell *c = (const ICONCell *)cells;
  box3f vb({INFINITY, INFINITY, INFINITY}, {-INFINITY, -INFINITY, -INFINITY});
  box3f sb{vec3f(INFINITY), vec3f(-INFINITY)};
  box1f dr(INFINITY, -INFINITY);
  for (int i = 0; i < n; ++i) {
    const ICONCell &cell = c[i];
    float minLat = fminf(cell.lat.x, fminf(cell.lat.y, cell.lat.z));
    float maxLat = fmaxf(cell.lat.x, fmaxf(cell.lat.y, cell.lat.z));
    float minLon = fminf(cell.lon.x, fminf(cell.lon.y, cell.lon.z));
    float maxLon = fmaxf(cell.lon.x, fmaxf(cell.lon.y, cell.lon.z));
    sb.lower.x = fminf(sb.lower.x, cell.height[0]);
    sb.upper.x = fmaxf(sb.upper.x, cell.height[cell.numLayers]);
    sb.lower.y = fminf(sb.lower.y, minLat);
    sb.upper.y = fmaxf(sb.upper.y, maxLat);
    sb.lower.z = fminf(sb.lower.z, minLon);
    sb.upper.z = fmaxf(sb.upper.z, maxLon);
    vb.extend(cell.getBounds());
    for (int j = 0; j < cell.numLayers; ++j) dr.extend(cell.value[j]);
  }
  for (int k = 0; k < 3; ++k) {
    volLo[k] = vb.lower[k]; volHi[k] = vb.upper[k];
    sphLo[k] = sb.lower[k]; sphHi[k] = sb.upper[k];
  }
  dataRange[0] = dr.lower; dataRange[1] = dr.upper;
}

// Camera::viewAll + getScreen (camera.h:86-104) -> launch-param camera as hostCode.cu:936-945
void ref_camera_view_all(const float *volLo, const float *volHi, float fovy, float aspect,
                         int W, int H, float *org, float *d00, float *du, float *dv)
{
  dvr_course::Camera cam;
  cam.fovy = fovy;
  cam.setAspect(aspect);
  cam.viewAll(box3f(vec3f(volLo[0], volLo[1], volLo[2]), vec3f(volHi[0], volHi[1], volHi[2])));
  vec3f ll, hz, vt;
  cam.getScreen(ll, hz, vt);
  vec3f p = cam.getPosition(), u = hz / W, v = vt / H;
  for (int k = 0; k < 3; ++k) { org[k] = p[k]; d00[k] = ll[k]; du[k] = u[k]; dv[k] = v[k]; }
}

void ref_camera_orient(const float *eye, const float *poi, const float *up, float fovy,
                       float aspect, int W, int H, float *org, float *d00, float *du, float *dv)
{
  dvr_course::Camera cam;
  cam.setAspect(aspect);
  cam.setOrientation(vec3f(eye[0], eye[1], eye[2]), vec3f(poi[0], poi[1], poi[2]),
                     vec3f(up[0], up[1], up[2]), fovy);
  vec3f ll, hz, vt;
  cam.getScreen(ll, hz, vt);
  vec3f p = cam.getPosition(), u = hz / W, v = vt / H;
  for (int k = 0; k < 3; ++k) { org[k] = p[k]; d00[k] = ll[k]; du[k] = u[k]; dv[k] = v[k]; }
}

void ref_resample_lut(const float *src, int nsrc, float *dst, int ndst)
{
  std::vector<vec4f> s(nsrc), d(ndst);
  std::memcpy(s.data(), src, sizeof(vec4f) * nsrc);
  dvr_course::resampleLUT(d, s);     // dvr_course-common.h:44-70
  std::memcpy(dst, d.data(), sizeof(vec4f) * ndst);
}

// ---- frames ----------------------------------------------------------------

// Fills the launch-param block the way hostCode.cu:912-958 does.
void ref_set_scene(const void *cells, int numCells, const float *volLo, const float *volHi,
                   const float *lut, int lutSize, float lo, float hi, float opacityScale,
                   float unitDistance, const float *ambientColor, float ambientRadiance)
{
  auto &lp = optixLaunchParams;
  lp.volume.cells = (ICONCell *)cells;
  lp.volume.numCells = numCells;
  lp.volume.bounds = box3f(vec3f(volLo[0], volLo[1], volLo[2]), vec3f(volHi[0], volHi[1], volHi[2]));
  lp.volume.accelMode = SPHERE_ACCEL_MODE;
  lp.transfunc.valueRange = box1f(lo, hi);
  lp.transfunc.opacityScale = opacityScale;
  lp.transfunc.values = (vec4f *)lut;
  lp.transfunc.size = lutSize;
  lp.unitDistance = unitDistance;
  lp.ambientColor = vec3f(ambientColor[0], ambientColor[1], ambientColor[2]);
  lp.ambientRadiance = ambientRadiance;
}

// Ray-march frame (A.3) over the pixel rectangle [x0,x1)x[y0,y1) of a WxH image,
// dispatched like Pipeline::launch's CPU branch (pipeline.cu:1066-1071): 64x64 tiles pulled
// by hardware_concurrency() threads (see forEachTile).  deadline_s<=0: no limit.
// perPixel (optional, W*H*4 int32): nSteps, nHits, hash lo, hash hi.
// Returns the number of pixels completed; totals[0]=steps, totals[1]=hits, totals[2]=seconds.
long ref_march_frame(int W, int H, int x0, int y0, int x1, int y1, int accumID,
                     uint32_t *fbPointer, float *accumBuffer,
                     int jitter, float dt, float termAlpha, float rInner, float rOuter,
                     int32_t *perPixel, int nthreads, double deadline_s, double *totals)
{
  auto &lp = optixLaunchParams;
  lp.fbPointer = fbPointer;
  lp.fbDepth = nullptr;
  lp.accumBuffer = (vec4f *)accumBuffer;
  lp.accumID = accumID;
  MarchCfg cfg{jitter, dt, termAlpha, rInner, rOuter, g_chunk};
  std::atomic<long> steps{0}, hits{0}, pixels{0};
  auto t0 = std::chrono::steady_clock::now();
  auto expired = [&]() {
    if (deadline_s <= 0) return false;
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > deadline_s;
  };
  if (accumID == 0)    // pipeline.cu:1058-1059 clearFramebuffer
    for (int y = y0; y < y1; ++y)
      for (int x = x0; x < x1; ++x) {
        int p = x + W * y;
        fbPointer[p] = 0;
        for (int k = 0; k < 4; ++k) accumBuffer[4 * p + k] = 0.f;
      }
  forEachTile(x0, x1, y0, y1, nthreads, [&](int x, int y) {
    if (expired()) return;
    PixelOut o = march_pixel(x, y, W, H, cfg, nullptr, 0);
    int p = x + W * y;
    write_pixel(p, o.C);
    if (perPixel) {
      perPixel[4 * p + 0] = o.nSteps;
      perPixel[4 * p + 1] = o.nHits;
      perPixel[4 * p + 2] = (int32_t)(uint32_t)(o.hash & 0xffffffffu);
      perPixel[4 * p + 3] = (int32_t)(uint32_t)(o.hash >> 32);
    }
    steps += o.nSteps; hits += o.nHits; pixels += 1;
  });
  if (totals) {
    totals[0] = (double)steps.load();
    totals[1] = (double)hits.load();
    totals[2] = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  }
  return pixels.load();
}

// Full located-id sequence (-1 = miss) of one pixel; returns nSteps.
int ref_march_trace(int W, int H, int x, int y, int accumID, int jitter, float dt, float termAlpha,
                    float rInner, float rOuter, int32_t *ids, int maxIds, float *rgba)
{
  auto &lp = optixLaunchParams;
  lp.accumID = accumID;
  MarchCfg cfg{jitter, dt, termAlpha, rInner, rOuter, g_chunk};
  PixelOut o = march_pixel(x, y, W, H, cfg, ids, maxIds);
  rgba[0] = o.C.x; rgba[1] = o.C.y; rgba[2] = o.C.z; rgba[3] = o.C.w;
  return o.nSteps;
}

// ---- the reference's majorant builders, as the reference wrote them (hostCode.cu kernels on host shims) --------
// buildICONGrid (hostCode.cu:668-682): initGrid + buildGrid_ICON over a dims grid on worldBounds
void ref_grid_ranges(const void *cells, int numCells, const int *dims, const float *wLo, const float *wHi, float *ranges /* 2 per macrocell */)
{
  Grid grid;
  grid.dims = vec3i(dims[0], dims[1], dims[2]);
  grid.worldBounds = box3f(vec3f(wLo[0], wLo[1], wLo[2]), vec3f(wHi[0], wHi[1], wHi[2]));
  grid.valueRanges = (box1f *)ranges;
  grid.maxOpacities = nullptr;
  const size_t numMC = grid.dims.x * size_t(grid.dims.y) * grid.dims.z;
  refLaunch1024(numMC, (void (*)(Grid))initGrid, grid);
  refLaunch1024((size_t)numCells, buildGrid_ICON, grid, (const ICONCell *)cells, (size_t)numCells);
}

// buildShellAccel (hostCode.cu:652-666): initGrid + buildShell_ICON over the (rad, lat, lon) grid on sphericalBounds
void ref_shell_ranges(const void *cells, int numCells, const int *dims, const float *sLo, const float *sHi, float *ranges)
{
  ShellAccel shell;
  shell.dims = vec3i(dims[0], dims[1], dims[2]);
  shell.sphericalBounds = box3f(vec3f(sLo[0], sLo[1], sLo[2]), vec3f(sHi[0], sHi[1], sHi[2]));
  shell.valueRanges = (box1f *)ranges;
  shell.maxOpacities = nullptr;
  const size_t numMC = shell.dims.x * size_t(shell.dims.y) * shell.dims.z;
  refLaunch1024(numMC, (void (*)(ShellAccel))initGrid, shell);
  refLaunch1024((size_t)numCells, buildShell_ICON, shell, (const ICONCell *)cells, (size_t)numCells);
}

// computeMaxOpacities (hostCode.cu:362-434; shell != 0: the ShellAccel overload, else the Grid overload)
void ref_max_opacities(const float *ranges, long numMC, const float *lut, int size, float tfLo, float tfHi, int shell, float *maxOp)
{
  const box1f vr(tfLo, tfHi), rel(0.f, 1.f);
  if (shell) {
    ShellAccel a;
    a.dims = vec3i((int)numMC, 1, 1);
    a.valueRanges = (box1f *)ranges;
    a.maxOpacities = maxOp;
    refLaunch1024((size_t)numMC, (void (*)(ShellAccel, const vec4f *, int, box1f, box1f, float))computeMaxOpacities, a, (const vec4f *)lut, size, vr, rel, 1.f);
  } else {
    Grid g;
    g.dims = vec3i((int)numMC, 1, 1);
    g.valueRanges = (box1f *)ranges;
    g.maxOpacities = maxOp;
    refLaunch1024((size_t)numMC, (void (*)(Grid, const vec4f *, int, box1f, box1f, float))computeMaxOpacities, g, (const vec4f *)lut, size, vr, rel, 1.f);
  }
}

// The reference's wedge sampler, unmodified: intersectWedgeEXT (UElems.h:214-311) -- Newton inverse map + shape
// functions -- on n (point, wedge) pairs; V24 holds six float4 vertices (xyz + scalar) per wedge in
// buildCuBQLAccel's order (hostCode.cu:562-597), exactly what the CUBQL-mode lambda passes (deviceCode.cu:98-106).
void ref_wedge(const float *pos, const float *V24, int n, int32_t *accept, float *value)
{
  for (int i = 0; i < n; ++i) {
    const float *v = V24 + 24 * (size_t)i;
    float4 q[6];
    for (int k = 0; k < 6; ++k) { q[k].x = v[4 * k]; q[k].y = v[4 * k + 1]; q[k].z = v[4 * k + 2]; q[k].w = v[4 * k + 3]; }
    float val = 0.f;
    const bool ok = intersectWedgeEXT(val, vec3f(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]), q[0], q[1], q[2], q[3], q[4], q[5]);
    accept[i] = ok ? 1 : 0;
    value[i] = ok ? val : 0.f;
  }
}

// The reference's own integrator, unmodified: woodcockTrackingAE (deviceCode.cu:239-275)
// through Pipeline::launch's CPU branch (pipeline.cu:1038-1075), one frame per call with a
// fresh Pipeline -- the non-interactive pipeline renders exactly one frame (sampleLimit 1,
// pipeline.cu:762-765) and only clears the buffers when frameID==0 (:1058-1059), so progressive
// averaging over frames is done by the caller: with cleared buffers the frame written for
// accumID=f is new/(f+1) (deviceCode.cu:267-268).  accumID only seeds the RNG here.
// Returns seconds spent in launch().
double ref_woodcock_frame(int W, int H, int accumID, uint32_t *fbOut, float *accumOut)
{
  static char a0[] = "icon_rt";
  char *argv[] = {a0, nullptr};
  dvr_course::Pipeline pl(1, argv, "icon_rt_ref");
  dvr_course::Frame fb(W, H);
  dvr_course::Camera cam;   // launch params carry the camera; Pipeline only needs it non-null
  pl.setFrame(&fb);
  pl.setCamera(&cam);
  pl.setRayGen(woodcockTrackingAE);
  auto &lp = optixLaunchParams;
  lp.fbPointer = fb.fbPointer;
  lp.fbDepth = fb.fbDepth;
  lp.accumBuffer = fb.accumBuffer;
  lp.accumID = accumID;
  auto t0 = std::chrono::steady_clock::now();
  pl.launch();
  double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  std::memcpy(fbOut, fb.fbPointer, sizeof(uint32_t) * W * H);
  std::memcpy(accumOut, fb.accumBuffer, sizeof(float) * 4 * W * H);
  return sec;
}

// The reference's DEFAULT ray-gen, unmodified (hostCode.cu:860): woodcockTrackingWithAccel (deviceCode.cu:281-341) in
// GRID_ACCEL_MODE -- dda3 (DDA.h:35-136) over volume.gridAccel with per-macrocell majorants maxOpacities[leafID].
// The majorant array is an INPUT here (the builders are CUDA kernels inside hostCode.cu, which does not compile
// without nvcc + OWL; oracle/icon_oracle.c restates them).
double ref_woodcock_grid_frame(int W, int H, int accumID, const int *dims, const float *gLo, const float *gHi,
                               const float *maxOpacities, uint32_t *fbOut, float *accumOut)
{
  static char a0[] = "icon_rt";
  char *argv[] = {a0, nullptr};
  dvr_course::Pipeline pl(1, argv, "icon_rt_ref");
  dvr_course::Frame fb(W, H);
  dvr_course::Camera cam;
  pl.setFrame(&fb);
  pl.setCamera(&cam);
  pl.setRayGen(woodcockTrackingWithAccel);
  auto &lp = optixLaunchParams;
  lp.volume.gridAccel.valueRanges = nullptr;
  lp.volume.gridAccel.dims = vec3i(dims[0], dims[1], dims[2]);
  lp.volume.gridAccel.worldBounds = box3f(vec3f(gLo[0], gLo[1], gLo[2]), vec3f(gHi[0], gHi[1], gHi[2]));
  lp.volume.gridAccel.maxOpacities = (float *)maxOpacities;
  lp.volume.accelMode = GRID_ACCEL_MODE;
  lp.fbPointer = fb.fbPointer;
  lp.fbDepth = fb.fbDepth;
  lp.accumBuffer = fb.accumBuffer;
  lp.accumID = accumID;
  auto t0 = std::chrono::steady_clock::now();
  pl.launch();
  double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  lp.volume.accelMode = SPHERE_ACCEL_MODE;
  std::memcpy(fbOut, fb.fbPointer, sizeof(uint32_t) * W * H);
  std::memcpy(accumOut, fb.accumBuffer, sizeof(float) * 4 * W * H);
  return sec;
}

// The same ray-gen in SPHERE_ACCEL_MODE, the mode the reference starts in (hostCode.cu:75, :915): sdda (ShellAccel.h:83-199)
// over volume.accel with the shell's majorants -- the reference's own code, unmodified.
double ref_woodcock_sphere_frame(int W, int H, int accumID, const int *dims, const float *sLo, const float *sHi,
                                 const float *maxOpacities, uint32_t *fbOut, float *accumOut)
{
  static char a0[] = "icon_rt";
  char *argv[] = {a0, nullptr};
  dvr_course::Pipeline pl(1, argv, "icon_rt_ref");
  dvr_course::Frame fb(W, H);
  dvr_course::Camera cam;
  pl.setFrame(&fb);
  pl.setCamera(&cam);
  pl.setRayGen(woodcockTrackingWithAccel);
  auto &lp = optixLaunchParams;
  lp.volume.accel.dims = vec3i(dims[0], dims[1], dims[2]);
  lp.volume.accel.sphericalBounds = box3f(vec3f(sLo[0], sLo[1], sLo[2]), vec3f(sHi[0], sHi[1], sHi[2]));
  lp.volume.accel.valueRanges = nullptr;
  lp.volume.accel.maxOpacities = (float *)maxOpacities;
  lp.volume.accelMode = SPHERE_ACCEL_MODE;
  lp.fbPointer = fb.fbPointer;
  lp.fbDepth = fb.fbDepth;
  lp.accumBuffer = fb.accumBuffer;
  lp.accumID = accumID;
  auto t0 = std::chrono::steady_clock::now();
  pl.launch();
  double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  std::memcpy(fbOut, fb.fbPointer, sizeof(uint32_t) * W * H);
  std::memcpy(accumOut, fb.accumBuffer, sizeof(float) * 4 * W * H);
  return sec;
}

} // extern "C"
