// Device-side building blocks of the icon_rt hot path (included by kernels.cu; tests/emu compiles
// the same text for the host with intrinsic shims to unit-test the walk logic without a GPU --
// that harness is test infrastructure and is never linked into libicon_b200.so).
//
// Parity rules: every expression that feeds a located record id or a sample
// position uses __fmul_rn/__fadd_rn/__fsub_rn (never contracted into FMA) in the reference's
// operand order; kernels.cu is additionally compiled with -fmad=false, IEEE sqrt and division
// (nvcc defaults), and never with -use_fast_math.
//
// Reference functions restated here (file:line under /root/reference):
//   evalPlane            icon_rt/ICONGrid.h:176-179
//   sample() predicate   icon_rt/ICONGrid.h:181-208 (planes precomputed on the host, grid_build.cpp)
//   findHeight           icon_rt/ICONGrid.h:117-144
//   postClassify         icon_rt/deviceCode.cu:127-135
//   generateRay          icon_rt/deviceCode.cu:36-49
//   LCG<4>               common/dvr_course-common-both.h:42-88
//   intersectSphere      icon_rt/ShellAccel.h:34-53
//   sdda segment choice  icon_rt/ShellAccel.h:85-111
//   accumulate + pack    icon_rt/deviceCode.cu:333-340, common/dvr_course-common-both.h:30-35,89-110
#pragma once
#include "icb_types.h"

namespace icb {
namespace {

constexpr unsigned FULL = 0xffffffffu;

// ---------------------------------------------------------------------------------------------
// exact fp32 helpers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float dot3(float ax, float ay, float az, float bx, float by, float bz)
{
  return __fadd_rn(__fadd_rn(__fmul_rn(ax, bx), __fmul_rn(ay, by)), __fmul_rn(az, bz));
}

__device__ __forceinline__ float evalPlane(const float4 &p, float x, float y, float z)
{
  return __fsub_rn(dot3(x, y, z, p.x, p.y, p.z), p.w);
}

struct WalkState {
  int   col, rec, layer, nL;  // rec < 0: no planes loaded; layer < 0: (lo,hi) brackets a gap (miss)
  float lo, hi;               // open radius interval in which (rec,layer) stays the answer
  float hbot, htop, thr;
  float4 p0, p1, p2;
};

__device__ __forceinline__ void resetState(WalkState &s)
{
  s.col = -1; s.rec = -1; s.layer = -1; s.nL = 0;
  s.lo = 0.f; s.hi = 0.f; s.hbot = 0.f; s.htop = 0.f; s.thr = INFINITY;
  s.p0 = s.p1 = s.p2 = make_float4(0.f, 0.f, 0.f, 0.f);
}

__device__ __forceinline__ void loadRec(const DeviceGrid &g, int rec, WalkState &s)
{
  const float4 *q = reinterpret_cast<const float4 *>(g.rec + rec);
  s.p0 = __ldg(q + 0);
  s.p1 = __ldg(q + 1);
  s.p2 = __ldg(q + 2);
  const float4 m = __ldg(q + 3);
  s.hbot = m.x; s.htop = m.y; s.thr = m.z; s.nL = __float_as_int(m.w);
  s.rec = rec;
}

__device__ __forceinline__ ColBlock loadCol(const DeviceGrid &g, int c)
{
  const int4 *q = reinterpret_cast<const int4 *>(g.col + c);
  const int4 a = __ldg(q), b = __ldg(q + 1);
  ColBlock cb;
  cb.nbr[0] = a.x; cb.nbr[1] = a.y; cb.nbr[2] = a.z; cb.recBegin = a.w;
  cb.nRec = b.x; cb.flags = b.y; cb.rLo = __int_as_float(b.z); cb.rHi = __int_as_float(b.w);
  return cb;
}

// first i in [0,n) with r <= H[i+1]; H sorted, H[0] <= r <= H[n], n >= 1
__device__ __forceinline__ int layerSearchSorted(const float *__restrict__ H, int n, float r, int hint)
{
  int i = min(max(hint, 0), n - 1);
#pragma unroll 1
  for (int step = 0; step < 4; ++step) {
    if (i + 1 < n && r > __ldg(H + i + 1)) ++i;
    else if (i > 0 && r <= __ldg(H + i)) --i;
    else return i;
  }
  int lo = 0, hi = n - 1;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (r <= __ldg(H + mid + 1)) hi = mid; else lo = mid + 1;
  }
  return lo;
}

// the reference's search, probe for probe (ICONGrid.h:117-144) -- exact even for unsorted heights
__device__ __forceinline__ int findHeightLiteral(const float *__restrict__ H, int n, float r)
{
  int first = 0, count = n;
  while (count > 0) {
    const int step = count / 2;
    const int it = first + step;
    if (!(r <= __ldg(H + it + 1))) { first = it + 1; count -= step + 1; }
    else count = step;
  }
  return first;
}

__device__ __forceinline__ int cubeBin(const DeviceGrid &g, float x, float y, float z)
{
  const float ax = fabsf(x), ay = fabsf(y), az = fabsf(z);
  int face; float u, w;
  if (ax >= ay && ax >= az) { face = x > 0.f ? 0 : 1; u = y / ax; w = z / ax; }
  else if (ay >= az)        { face = y > 0.f ? 2 : 3; u = x / ay; w = z / ay; }
  else                      { face = z > 0.f ? 4 : 5; u = x / az; w = y / az; }
  const int N = g.binN;
  const int iu = min(N - 1, max(0, (int)floorf((u * 0.5f + 0.5f) * (float)N)));
  const int iw = min(N - 1, max(0, (int)floorf((w * 0.5f + 0.5f) * (float)N)));
  return (face * N + iw) * N + iu;
}

// Exact answer: lowest original record index among all records accepting P, looked for in the
// conservative candidate list of P's cube-map bin.  No topology, ordering or margin assumptions.
// Out of line and fed by value (no references), so that calling it costs the hot loop no stack traffic.
// Returns (internal record, column) or (-1,-1).
__device__ __noinline__ int2 exactQueryIds(const RecBlock *__restrict__ recs, const ColBlock *__restrict__ cols,
                                           const int *__restrict__ binOff, const int *__restrict__ binList,
                                           const int *__restrict__ recOrig, int binN, float x, float y, float z, float r)
{
  DeviceGrid g;
  g.rec = recs; g.col = cols; g.binN = binN;
  const int bin = cubeBin(g, x, y, z);
  const int j0 = __ldg(binOff + bin), j1 = __ldg(binOff + bin + 1);
  int best = 0x7fffffff, bestRec = -1, bestCol = -1;
  for (int j = j0; j < j1; ++j) {
    const int c = __ldg(binList + j);
    const ColBlock cb = loadCol(g, c);
    if (r < cb.rLo || r > cb.rHi) continue;
    for (int k = 0; k < cb.nRec; ++k) {
      const int rec = cb.recBegin + k;
      const float4 *q = reinterpret_cast<const float4 *>(recs + rec);
      const float4 m = __ldg(q + 3);
      if (r < m.x || r > m.y) continue;                   // ICONGrid.h:184
      if (evalPlane(__ldg(q + 0), x, y, z) > 0.f) continue;   // ICONGrid.h:201-203
      if (evalPlane(__ldg(q + 1), x, y, z) > 0.f) continue;
      if (evalPlane(__ldg(q + 2), x, y, z) > 0.f) continue;
      const int orig = __ldg(recOrig + rec);
      if (orig < best) { best = orig; bestRec = rec; bestCol = c; }
    }
  }
  return make_int2(bestRec, bestCol);
}

__device__ __forceinline__ bool exactQuery(const DeviceGrid &g, float x, float y, float z, float r, WalkState &s)
{
  const int2 id = exactQueryIds(g.rec, g.col, g.binOff, g.binList, g.recOrig, g.binN, x, y, z, r);
  if (id.x < 0) return false;
  loadRec(g, id.x, s);
  s.col = id.y;
  const float *H = g.heights + (size_t)id.x * kHeightStride;
  s.layer = findHeightLiteral(H, s.nL, r);
  s.lo = __ldg(H + s.layer);
  s.hi = __ldg(H + min(s.layer + 1, kHeightStride - 1));
  return true;
}

// Vertical relocation inside a well-formed column whose horizontal containment is already safe.
// Returns 1 hit, 0 miss (gap bracket set), -1 tie with a record boundary (caller asks exactQuery).
__device__ __forceinline__ int verticalRelocate(const DeviceGrid &g, const ColBlock &cb, float r, WalkState &s)
{
  int found = -1;
  bool tie = false;
  float gapLo = -INFINITY, gapHi = INFINITY;
  for (int k = 0; k < cb.nRec; ++k) {
    const float4 m = __ldg(reinterpret_cast<const float4 *>(g.rec + cb.recBegin + k) + 3);
    tie |= (r == m.x) | (r == m.y);
    if (r > m.x && r < m.y) found = k;
    if (m.y < r) gapLo = fmaxf(gapLo, m.y);
    if (m.x > r) gapHi = fminf(gapHi, m.x);
  }
  if (tie || !(r == r)) return -1;
  if (found < 0) { s.layer = -1; s.lo = gapLo; s.hi = gapHi; return 0; }
  const int rec = cb.recBegin + found;
  int hint = s.layer;
  if (rec != s.rec) { loadRec(g, rec, s); hint = -1; }
  const float *H = g.heights + (size_t)rec * kHeightStride;
  s.layer = layerSearchSorted(H, s.nL, r, hint);
  s.lo = __ldg(H + s.layer);
  s.hi = __ldg(H + s.layer + 1);
  return 1;
}

struct Evals {
  float e0, e1, e2;
  float m;      // max, NaN if any is NaN (never "safe")
  int kmax;
};

__device__ __forceinline__ Evals evalPlanes(const float4 &p0, const float4 &p1, const float4 &p2, float x, float y, float z)
{
  Evals e;
  e.e0 = evalPlane(p0, x, y, z); e.e1 = evalPlane(p1, x, y, z); e.e2 = evalPlane(p2, x, y, z);
  e.m = e.e0; e.kmax = 0;
  if (e.e1 > e.m) { e.m = e.e1; e.kmax = 1; }
  if (e.e2 > e.m) { e.m = e.e2; e.kmax = 2; }
  if (!(e.e0 == e.e0) || !(e.e1 == e.e1) || !(e.e2 == e.e2)) e.m = INFINITY * 0.f;
  return e;
}

__device__ __forceinline__ float maxEval(const WalkState &s, float x, float y, float z, int &kmax)
{
  const Evals e = evalPlanes(s.p0, s.p1, s.p2, x, y, z);
  kmax = e.kmax;
  return e.m;
}

__device__ __forceinline__ void setLayerSorted(const DeviceGrid &g, float r, int hint, WalkState &s)
{
  const float *H = g.heights + (size_t)s.rec * kHeightStride;
  s.layer = layerSearchSorted(H, s.nL, r, hint);
  s.lo = __ldg(H + s.layer);
  s.hi = __ldg(H + s.layer + 1);
}

__device__ __forceinline__ int pick3(int k, int a0, int a1, int a2) { return k == 0 ? a0 : (k == 1 ? a1 : a2); }

__device__ __forceinline__ void setRec(const float4 &b0, const float4 &b1, const float4 &b2, const float4 &bm, int rec, WalkState &s)
{
  s.p0 = b0; s.p1 = b1; s.p2 = b2;
  s.hbot = bm.x; s.htop = bm.y; s.thr = bm.z; s.nL = __float_as_int(bm.w);
  s.rec = rec;
}

// P lies within the margin band of exactly one side plane (kmax) of the current record `s` (record index k of
// column c), the other two planes are deeply satisfied and r is strictly inside the record: the only records
// that can accept P are this one and the one at the same radius across that edge.  Evaluate both with their own
// planes and let the lower caller index win, exactly as the reference's first-match loop does
// (deviceCode.cu:119-122).  Returns 1 hit (s updated), -1 when the situation is not that simple (caller asks
// the exact query).
__device__ __forceinline__ int resolveAcrossEdge(const DeviceGrid &g, const ColBlock &cb, int c, int k, const Evals &ea,
                                                 int hintLayer, float x, float y, float z, float r, WalkState &s)
{
  const float other = ea.kmax == 0 ? fmaxf(ea.e1, ea.e2) : (ea.kmax == 1 ? fmaxf(ea.e0, ea.e2) : fmaxf(ea.e0, ea.e1));
  if (!(other < -s.thr) || !(r > s.hbot && r < s.htop)) return -1;
  const int nb = pick3(ea.kmax, cb.nbr[0], cb.nbr[1], cb.nbr[2]);
  if (nb < 0) return -1;
  const ColBlock cn = loadCol(g, nb);
  if ((cn.flags & COL_GENERIC) || k >= cn.nRec) return -1;
  const int recB = cn.recBegin + k;
  const float4 *q = reinterpret_cast<const float4 *>(g.rec + recB);
  const float4 b0 = __ldg(q + 0), b1 = __ldg(q + 1), b2 = __ldg(q + 2), bm = __ldg(q + 3);
  if (!(r > bm.x && r < bm.y)) return -1;               // stair-stepped seam: not the same record index over there
  const Evals eb = evalPlanes(b0, b1, b2, x, y, z);
  if (!(eb.m == eb.m)) return -1;
  // the shared edge seen from the neighbour, and its other two planes deep as well (P is far from both end corners)
  const int kb = cn.nbr[0] == c ? 0 : (cn.nbr[1] == c ? 1 : (cn.nbr[2] == c ? 2 : -1));
  if (kb < 0) return -1;
  const float otherB = kb == 0 ? fmaxf(eb.e1, eb.e2) : (kb == 1 ? fmaxf(eb.e0, eb.e2) : fmaxf(eb.e0, eb.e1));
  if (!(otherB < -bm.z)) return -1;
  const bool accA = ea.m <= 0.f, accB = eb.m <= 0.f;
  if (!accA && !accB) return -1;
  const bool takeB = accB && (!accA || __ldg(g.recOrig + recB) < __ldg(g.recOrig + s.rec));
  if (takeB) { setRec(b0, b1, b2, bm, recB, s); s.col = nb; }
  setLayerSorted(g, r, hintLayer, s);
  return 1;
}

// Neighbour-triangle walk from the current column.  Returns 1 hit, 0 miss, -1 undecided (record-boundary ties,
// generic columns, corners, off the mesh, more than maxWalk hops): the caller then asks the exact query.
// The record index within the column and the layer index carry over from hop to hop as search hints (adjacent
// columns are layered alike), so a crossing costs one column block, one record block and one or two heights.
__device__ __forceinline__ int walkLocate(const DeviceGrid &g, float x, float y, float z, float r, WalkState &s)
{
  int c = s.col;
  if (c < 0) return -1;
  ColBlock cb = loadCol(g, c);
  if (cb.flags & COL_GENERIC) return -1;
  int k, hl = s.layer;
  if (s.rec >= 0) k = s.rec - cb.recBegin;             // planes already in registers
  else { k = cb.nRec - 1; hl = kHeightStride; loadRec(g, cb.recBegin + k, s); }     // a new ray: enters from the top
  s.layer = -1; s.lo = 0.f; s.hi = 0.f;
#pragma unroll 1
  for (int step = 0; step < g.maxWalk; ++step) {
    // the record of this column at radius r (stacked records, sorted by hbot)
#pragma unroll 1
    while (r > s.htop && k + 1 < cb.nRec) { ++k; loadRec(g, cb.recBegin + k, s); hl = 0; }
#pragma unroll 1
    while (r < s.hbot && k > 0) { --k; loadRec(g, cb.recBegin + k, s); hl = kHeightStride; }
    const Evals e = evalPlanes(s.p0, s.p1, s.p2, x, y, z);
    if (e.m < -s.thr) {
      if (r > s.hbot && r < s.htop) { setLayerSorted(g, r, hl, s); return 1; }
      const int rec0 = s.rec;
      const int v = verticalRelocate(g, cb, r, s);      // gaps and record-boundary ties
      if (v == 0) return 0;
      if (v == 1 && s.rec == rec0) return 1;
      return -1;
    }
    if (!(e.m == e.m)) return -1;
    if (e.m > s.thr) {                                  // clearly outside across plane kmax: hop
      const int nb = pick3(e.kmax, cb.nbr[0], cb.nbr[1], cb.nbr[2]);
      if (nb < 0) return -1;
      cb = loadCol(g, nb);
      if (cb.flags & COL_GENERIC) return -1;
      c = nb;
      k = min(k, cb.nRec - 1);
      loadRec(g, cb.recBegin + k, s);
      s.col = c;
      continue;
    }
    return resolveAcrossEdge(g, cb, c, k, e, hl, x, y, z, r, s);
  }
  return -1;
}

// Records whose acceptance region is not confined to their column's cone (grid_build.cpp "unbounded":
// zero/NaN plane normals, numLayers == 0, clockwise corners) are checked for every sample; the list is
// sorted by caller index, so the scan ends at the first entry that cannot beat the current answer.
__device__ __noinline__ int globalsQuery(const RecBlock *__restrict__ recs, const int *__restrict__ recOrig,
                                         const int *__restrict__ globalRecs, int nGlobal, float x, float y, float z, float r, int cur)
{
  for (int j = 0; j < nGlobal; ++j) {
    const int rec = __ldg(globalRecs + j);
    if (__ldg(recOrig + rec) >= cur) break;
    const float4 *q = reinterpret_cast<const float4 *>(recs + rec);
    const float4 m = __ldg(q + 3);
    if (r < m.x || r > m.y) continue;
    if (evalPlane(__ldg(q + 0), x, y, z) > 0.f) continue;
    if (evalPlane(__ldg(q + 1), x, y, z) > 0.f) continue;
    if (evalPlane(__ldg(q + 2), x, y, z) > 0.f) continue;
    return rec;
  }
  return -1;
}

__device__ __forceinline__ bool applyGlobals(const DeviceGrid &g, float x, float y, float z, float r, WalkState &s, bool hit)
{
  const int cur = hit ? __ldg(g.recOrig + s.rec) : 0x7fffffff;
  const int rec = globalsQuery(g.rec, g.recOrig, g.globalRecs, g.nGlobal, x, y, z, r, cur);
  if (rec < 0) return hit;
  loadRec(g, rec, s);                      // thr == +inf: the next sample re-locates from scratch
  s.col = -1;
  const float *H = g.heights + (size_t)rec * kHeightStride;
  s.layer = findHeightLiteral(H, s.nL, r);
  s.lo = s.hi = 0.f;
  return true;
}

// One located sample.  Fast path: still safely inside the current column and inside the current
// layer's open radius bracket -> no memory traffic at all.
__device__ __forceinline__ bool locateStepLocal(const DeviceGrid &g, float x, float y, float z, float r, WalkState &s,
                                                unsigned &nSlow, unsigned &nExact)
{
  if (s.rec >= 0) {
    int kmax;
    const float m = maxEval(s, x, y, z, kmax);
    if (m < -s.thr) {
      if (r > s.lo && r < s.hi) return s.layer >= 0;
      ++nSlow;
      if (r > s.hbot && r < s.htop) {              // another layer of the same record: one or two cached loads
        setLayerSorted(g, r, s.layer, s);
        return true;
      }
      const ColBlock cb = loadCol(g, s.col);
      const int oldRec = s.rec;
      const int v = verticalRelocate(g, cb, r, s);
      if (v == 0) return false;
      if (v == 1) {
        if (s.rec == oldRec) return true;
        if (maxEval(s, x, y, z, kmax) < -s.thr) return true;     // new record: its own planes decide
      }
      --nSlow;
    }
  }
  ++nSlow;
  const int w = walkLocate(g, x, y, z, r, s);
  if (w >= 0) return w != 0;
  ++nExact;
  return exactQuery(g, x, y, z, r, s);
}

// A position with a NaN coordinate has r == NaN, which fails every rejection of sample() (ICONGrid.h:184,
// 201-203 are all `<`/`>` tests): the brute-force loop returns record 0, layer findHeight(NaN) == numLayers.
__device__ __forceinline__ bool nanPosition(const DeviceGrid &g, WalkState &s)
{
  loadRec(g, g.recOfOrig0, s);
  s.col = -1; s.thr = INFINITY; s.lo = s.hi = 0.f;
  s.layer = s.nL;
  return true;
}

__device__ __forceinline__ bool locateStep(const DeviceGrid &g, float x, float y, float z, float r, WalkState &s,
                                           unsigned &nSlow, unsigned &nExact)
{
  if (!(r == r)) return nanPosition(g, s);
  const bool hit = locateStepLocal(g, x, y, z, r, s, nSlow, nExact);
  if (g.nGlobal == 0) return hit;
  return applyGlobals(g, x, y, z, r, s, hit);
}

__device__ __forceinline__ bool exactLocate(const DeviceGrid &g, float x, float y, float z, float r, WalkState &s)
{
  if (!(r == r)) return nanPosition(g, s);
  const bool hit = exactQuery(g, x, y, z, r, s);
  if (g.nGlobal == 0) return hit;
  return applyGlobals(g, x, y, z, r, s, hit);
}

// ---------------------------------------------------------------------------------------------
// classification / ray set-up
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 postClassify(const float4 *__restrict__ lut, const RenderParams &p, float v)
{
  v = __fdiv_rn(__fsub_rn(v, p.tfLo), __fsub_rn(p.tfHi, p.tfLo));
  const float vs = __fmul_rn(v, (float)p.tfSize);
  const int idx = (int)vs;
  const float frac = __fsub_rn(vs, (float)idx);
  const float4 a = lut[min(max(idx, 0), p.tfSize - 1)];
  const float4 b = lut[min(max(idx + 1, 0), p.tfSize - 1)];
  const float om = __fsub_rn(1.f, frac);
  float4 o;
  o.x = __fadd_rn(__fmul_rn(a.x, frac), __fmul_rn(b.x, om));
  o.y = __fadd_rn(__fmul_rn(a.y, frac), __fmul_rn(b.y, om));
  o.z = __fadd_rn(__fmul_rn(a.z, frac), __fmul_rn(b.z, om));
  o.w = __fadd_rn(__fmul_rn(a.w, frac), __fmul_rn(__fmul_rn(b.w, om), p.opacityScale));
  return o;
}

struct Lcg {
  unsigned state;
  __device__ __forceinline__ void init(unsigned v0, unsigned v1)
  {
    unsigned s0 = 0;
#pragma unroll
    for (int n = 0; n < 4; ++n) {
      s0 += 0x9e3779b9u;
      v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
      v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    state = v0;
  }
  __device__ __forceinline__ float next()
  {
    state = 1664525u * state + 1013904223u;
    return __fdiv_rn((float)(state & 0x00FFFFFFu), (float)0x01000000);
  }
};

__device__ __forceinline__ bool intersectSphere(float ox, float oy, float oz, float dx, float dy, float dz,
                                                float radius, float &tnear, float &tfar)
{
  const float A = dot3(dx, dy, dz, dx, dy, dz);
  const float B = __fmul_rn(dot3(dx, dy, dz, ox, oy, oz), 2.f);
  const float C = __fsub_rn(dot3(ox, oy, oz, ox, oy, oz), __fmul_rn(radius, radius));
  float d = __fsub_rn(__fmul_rn(B, B), __fmul_rn(__fmul_rn(4.f, A), C));
  if (d < 0.f) return false;
  d = __fsqrt_rn(d);
  const float q = B < 0.f ? __fmul_rn(-0.5f, __fsub_rn(B, d)) : __fmul_rn(-0.5f, __fadd_rn(B, d));
  const float t1 = __fdiv_rn(q, A);
  const float t2 = __fdiv_rn(C, q);
  tnear = fminf(t1, t2);
  tfar = fmaxf(t1, t2);
  return true;
}

struct RaySetup {
  float dx, dy, dz;
  float seg[2][2];
  int nseg;
};

__device__ __forceinline__ void setupRay(const RenderParams &p, int px, int py, RaySetup &r)
{
  Lcg rnd;
  rnd.init((unsigned)(p.accumID * p.width * p.height) + (unsigned)px, (unsigned)py);
  const bool jit = (p.flags & 1u) != 0;
  // deviceCode.cu:42-43 leaves the order of its two rnd() calls unsequenced; the reference as built by
  // g++ (oracle/_ref) draws the dir_dv sample first, then the dir_du sample.
  const float jv = jit ? rnd.next() : 0.f;
  const float b = __fadd_rn(__fadd_rn((float)py, .5f), jv);
  const float ju = jit ? rnd.next() : 0.f;
  const float a = __fadd_rn(__fadd_rn((float)px, .5f), ju);
  float dx = __fadd_rn(__fadd_rn(p.d00[0], __fmul_rn(a, p.du[0])), __fmul_rn(b, p.dv[0]));
  float dy = __fadd_rn(__fadd_rn(p.d00[1], __fmul_rn(a, p.du[1])), __fmul_rn(b, p.dv[1]));
  float dz = __fadd_rn(__fadd_rn(p.d00[2], __fmul_rn(a, p.du[2])), __fmul_rn(b, p.dv[2]));
  const float len = __fsqrt_rn(dot3(dx, dy, dz, dx, dy, dz));
  dx = __fdiv_rn(dx, len); dy = __fdiv_rn(dy, len); dz = __fdiv_rn(dz, len);
  if (fabsf(dx) < 1e-5f) dx = 1e-5f;
  if (fabsf(dy) < 1e-5f) dy = 1e-5f;
  if (fabsf(dz) < 1e-5f) dz = 1e-5f;
  r.dx = dx; r.dy = dy; r.dz = dz;
  r.nseg = 0;
  const float tmin = 0.f, tmax = 1e10f;                  // Ray(org,dir,0.f,1e10f)
  const float sceneEPS = __fmul_rn(p.rInner, 1e-6f);
  float t1 = 0.f, t2 = 0.f, t3 = 0.f, t4 = 0.f;
  const bool s1 = intersectSphere(p.org[0], p.org[1], p.org[2], dx, dy, dz, p.rOuter, t1, t4);
  const bool s2 = intersectSphere(p.org[0], p.org[1], p.org[2], dx, dy, dz, p.rInner, t2, t3);
  if (!s1) return;
  if (t4 < tmin) return;
  float r0lo, r0hi, r1lo = INFINITY, r1hi = -INFINITY;
  if (!s2)            { r0lo = t1; r0hi = t4; }
  else if (tmin < t2) { r0lo = t1; r0hi = t2; r1lo = t3; r1hi = t4; }
  else                { r0lo = t3; r0hi = t4; }
  const float tlo = fmaxf(tmin, 0.f);
  if (r0hi <= r0lo) return;
  r.seg[0][0] = fmaxf(__fadd_rn(r0lo, sceneEPS), tlo);
  r.seg[0][1] = fminf(__fsub_rn(r0hi, sceneEPS), tmax);
  r.nseg = 1;
  if (r1hi <= r1lo) return;
  r.seg[1][0] = fmaxf(__fadd_rn(r1lo, sceneEPS), tlo);
  r.seg[1][1] = fminf(__fsub_rn(r1hi, sceneEPS), tmax);
  r.nseg = 2;
}

__device__ __forceinline__ float linearToSrgb(float x)
{
  if (x <= 0.0031308f) return __fmul_rn(12.92f, x);
  return __fsub_rn(__fmul_rn(1.055f, powf(x, 1.f / 2.4f)), 0.055f);
}
__device__ __forceinline__ unsigned make8(float f)
{
  return (unsigned)fminf(255.f, fmaxf(0.f, (float)(int)__fmul_rn(f, 256.f)));
}

__device__ __forceinline__ void writePixel(const RenderParams &p, float4 C, size_t idx, unsigned *fb, float4 *accum)
{
  const float aw = __fdiv_rn(1.f, (float)(p.accumID + 1));
  const float om = __fsub_rn(1.f, aw);
  float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
  if (p.accumID != 0 && accum) o = accum[idx];
  float4 n;
  n.x = __fadd_rn(__fmul_rn(aw, C.x), __fmul_rn(om, o.x));
  n.y = __fadd_rn(__fmul_rn(aw, C.y), __fmul_rn(om, o.y));
  n.z = __fadd_rn(__fmul_rn(aw, C.z), __fmul_rn(om, o.z));
  n.w = __fadd_rn(__fmul_rn(aw, C.w), __fmul_rn(om, o.w));
  if (accum) accum[idx] = n;
  if (fb)
    fb[idx] = (make8(linearToSrgb(n.x)) << 0) + (make8(linearToSrgb(n.y)) << 8) +
              (make8(linearToSrgb(n.z)) << 16) + (make8(n.w) << 24);
}

__device__ __forceinline__ unsigned long long fnvStep(unsigned long long h, int id)
{
  const unsigned u = (unsigned)id;
#pragma unroll
  for (int b = 0; b < 4; ++b) {
    h ^= (u >> (8 * b)) & 0xffu;
    h *= 1099511628211ull;
  }
  return h;
}

__device__ __forceinline__ unsigned compact1by1(unsigned x)
{
  x &= 0x55555555u;
  x = (x ^ (x >> 1)) & 0x33333333u;
  x = (x ^ (x >> 2)) & 0x0f0f0f0fu;
  x = (x ^ (x >> 4)) & 0x00ff00ffu;
  return x;
}


// Skip-ahead.  The sample at (x,y,z) = P_i has just been located in (s.rec, s.layer) by the exact predicate.
// Returns n >= 0 such that the NEXT n samples of the ray are certain to take the fast path of locateStep with
// the same answer, so that neither their positions nor their plane tests need to be evaluated:
//   e_k(P_j) = e_k(P_i) + (t_j - t_i) * dot(dir, N_k) + rounding,   |P_j|^2 = |P_i + (t_j - t_i) dir|^2 + rounding,
// where every rounding term (t_j, the position, the dot products, the square root) is a few ulps of the largest
// coordinate magnitude M = max(|org|, rOuter).  RenderParams carries the slack that covers them with a factor
// of two to spare (skipRadSlack = 16 ulp(M) in metres, skipPlaneScale = 1 + slack/margin applied to the
// record's safe-interior threshold); DESIGN.md "skip-ahead" has the error budget.  The caller still clamps n to
// the end of the shell segment with the exact t_j.
__device__ __forceinline__ float divApprox(float a, float b) { return __fdividef(a, b); }
__device__ __forceinline__ float sqrtApprox(float a)
{
#ifdef __CUDA_ARCH__
  float r;
  asm("sqrt.approx.f32 %0, %1;" : "=f"(r) : "f"(a));
  return r;
#else
  return sqrtf(a);
#endif
}

// (the quotients and roots below only size the skip; their relative error, ~1e-6 of at most maxSkip*dt, is a
// fraction of a metre and far inside the slack, so the approximate MUFU forms are used)
__device__ __forceinline__ int skipCount(const WalkState &s, const RenderParams &p, float dx, float dy, float dz,
                                         float x, float y, float z, float r, float dt)
{
  const float thr2 = __fmul_rn(s.thr, p.skipPlaneScale);
  const Evals e = evalPlanes(s.p0, s.p1, s.p2, x, y, z);
  if (!(e.m < -thr2)) return 0;
  const float loS = s.lo + p.skipRadSlack, hiS = s.hi - p.skipRadSlack;
  if (!(r > loS && r < hiS)) return 0;
  float D = (float)p.maxSkip * dt;
  const float B0 = dot3(dx, dy, dz, s.p0.x, s.p0.y, s.p0.z);
  const float B1 = dot3(dx, dy, dz, s.p1.x, s.p1.y, s.p1.z);
  const float B2 = dot3(dx, dy, dz, s.p2.x, s.p2.y, s.p2.z);
  if (B0 > 0.f) D = fminf(D, divApprox(-thr2 - e.e0, B0));
  if (B1 > 0.f) D = fminf(D, divApprox(-thr2 - e.e1, B1));
  if (B2 > 0.f) D = fminf(D, divApprox(-thr2 - e.e2, B2));
  const float b = dot3(x, y, z, dx, dy, dz), dd = dot3(dx, dy, dz, dx, dy, dz);
  const float cHi = (r - hiS) * (r + hiS);                      // < 0
  const float sqHi = sqrtApprox(fmaxf(b * b - dd * cHi, 0.f));
  D = fminf(D, b < 0.f ? divApprox(sqHi - b, dd) : divApprox(-cHi, b + sqHi));  // leaves the layer through its top
  if (b < 0.f) {                                                // heading inwards: may reach the layer's bottom
    const float cLo = (r - loS) * (r + loS);                    // > 0
    const float disc = b * b - dd * cLo;
    if (disc > 0.f) D = fminf(D, divApprox(cLo, sqrtApprox(disc) - b));
  }
  const int n = (int)divApprox(D, dt) - 1;
  return n > 0 ? min(n, p.maxSkip) : 0;
}

// exact sample parameter of step i of a segment starting at lo: one product, one sum, no FMA
__device__ __forceinline__ float sampleT(float lo, int i, float dt)
{
  return __fadd_rn(lo, __fmul_rn(__fadd_rn((float)i, 0.5f), dt));
}

// clamp a skip of n samples starting at index i so that every skipped sample lies before the segment's end
__device__ __forceinline__ int clampSkip(int n, int i, float lo, float hi, float t, float dt)
{
  const float room = divApprox(hi - t, dt);
  n = min(n, room > 2.f ? (int)room - 2 : 0);
  while (n > 0 && !(sampleT(lo, i + n - 1, dt) < hi)) --n;
  return n;
}

// One pixel marched by one thread: the statement the render kernel's lanes
// follow step for step.  ids (optional) receives the located record id of every step (-1 = miss).
__device__ inline void marchPixel(const DeviceGrid &g, const RenderParams &p, const float4 *lut, int px, int py,
                                  int *ids, int maxIds, float4 *rgbaOut, int *nStepsOut, bool useSkip = false,
                                  int *nSkippedOut = nullptr)
{
  RaySetup rs;
  setupRay(p, px, py, rs);
  float4 C = make_float4(0.f, 0.f, 0.f, 0.f);
  WalkState s;
  resetState(s);
  unsigned nSlow = 0, nExact = 0;
  int n = 0, nSkipped = 0;
  bool done = false;
  for (int sg = 0; sg < rs.nseg && !done; ++sg) {
    for (int i = 0;; ++i) {
      const float t = sampleT(rs.seg[sg][0], i, p.dt);
      if (t >= rs.seg[sg][1]) break;
      const float x = __fadd_rn(p.org[0], __fmul_rn(rs.dx, t));
      const float y = __fadd_rn(p.org[1], __fmul_rn(rs.dy, t));
      const float z = __fadd_rn(p.org[2], __fmul_rn(rs.dz, t));
      const float r = __fsqrt_rn(dot3(x, y, z, x, y, z));
      const bool hit = locateStep(g, x, y, z, r, s, nSlow, nExact);
      if (ids && n < maxIds) ids[n] = hit ? g.recOrig[s.rec] : -1;
      ++n;
      if (!hit) continue;
      const float v = g.values[s.rec * kHeightStride + s.layer];
      const float4 sm = postClassify(lut, p, v);
      const float a = __fsub_rn(1.f, expf(__fdiv_rn(__fmul_rn(-sm.w, p.dt), p.unitDistance)));
      const float cr = __fmul_rn(__fmul_rn(sm.x, p.ambient[0]), p.ambientRadiance);
      const float cg = __fmul_rn(__fmul_rn(sm.y, p.ambient[1]), p.ambientRadiance);
      const float cb = __fmul_rn(__fmul_rn(sm.z, p.ambient[2]), p.ambientRadiance);
      int reps = 1;
      if (useSkip && p.maxSkip > 0 && g.nGlobal == 0)
        reps += clampSkip(skipCount(s, p, rs.dx, rs.dy, rs.dz, x, y, z, r, p.dt), i + 1, rs.seg[sg][0], rs.seg[sg][1], t, p.dt);
      for (int q = 0; q < reps; ++q) {
        if (q > 0) {                       // a skipped sample: same record, same layer, by construction
          if (ids && n < maxIds) ids[n] = g.recOrig[s.rec];
          ++n; ++i; ++nSkipped;
        }
        const float w = __fmul_rn(__fsub_rn(1.f, C.w), a);
        C.x = __fadd_rn(C.x, __fmul_rn(w, cr));
        C.y = __fadd_rn(C.y, __fmul_rn(w, cg));
        C.z = __fadd_rn(C.z, __fmul_rn(w, cb));
        C.w = __fadd_rn(C.w, w);
        if (C.w >= p.termAlpha) { done = true; break; }
      }
      if (done) break;
    }
  }
  *rgbaOut = C;
  *nStepsOut = n;
  if (nSkippedOut) *nSkippedOut = nSkipped;
}

}  // namespace
}  // namespace icb
