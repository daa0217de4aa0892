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
#include "fdlibm_trig.cuh"
#include "glibc_logf.cuh"

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

// The record a ray is in: its node (planes, radius range, links to the adjacent records) lives in
// registers; its height and value lines (2 x 128 bytes) live in a per-lane slot `L` -- shared memory on
// the device, filled by cp.async at the same time as the node is read, so that a crossing into an
// adjacent record costs ONE memory round trip and every layer change inside a record costs none.
struct WalkState {
  int   rec;                  // internal record id, -1: none
  int   layer;                // >= 0: located layer; -1: (lo,hi) brackets radii no record of the column covers
  int   info, orig;
  int   nbr0, nbr1, nbr2;
  int   hops;                 // hops spent on the current sample
  float lo, hi;               // open radius interval in which (rec,layer) stays the answer
  float hbot, htop, thr;
  float4 p0, p1, p2;
};

struct NodeRegs {
  float4 p0, p1, p2;
  float hbot, htop, thr;
  int orig, nbr0, nbr1, nbr2, info;
};

__device__ __forceinline__ int recK(int info) { return info & REC_K_MASK; }
__device__ __forceinline__ int recCount(int info) { return (info >> REC_NREC_SHIFT) & 0xff; }
__device__ __forceinline__ int recLayers(int info) { return (info >> REC_NL_SHIFT) & 31; }

__device__ __forceinline__ void resetState(WalkState &s)
{
  s.rec = -1; s.layer = -1; s.info = 0; s.orig = -1; s.nbr0 = s.nbr1 = s.nbr2 = -1; s.hops = 0;
  s.lo = 0.f; s.hi = 0.f; s.hbot = 0.f; s.htop = 0.f; s.thr = INFINITY;
  s.p0 = s.p1 = s.p2 = make_float4(0.f, 0.f, 0.f, 0.f);
}

__device__ __forceinline__ NodeRegs loadNode(const DeviceGrid &g, int rec)
{
  const float4 *q = reinterpret_cast<const float4 *>(g.rec + rec);
  NodeRegs n;
  n.p0 = __ldg(q + 0);
  n.p1 = __ldg(q + 1);
  n.p2 = __ldg(q + 2);
  const float4 m = __ldg(q + 3);
  const int4 l = __ldg(reinterpret_cast<const int4 *>(q) + 4);
  n.hbot = m.x; n.htop = m.y; n.thr = m.z; n.orig = __float_as_int(m.w);
  n.nbr0 = l.x; n.nbr1 = l.y; n.nbr2 = l.z; n.info = l.w;
  return n;
}

__device__ __forceinline__ void commitNode(const NodeRegs &n, int rec, WalkState &s)
{
  s.p0 = n.p0; s.p1 = n.p1; s.p2 = n.p2;
  s.hbot = n.hbot; s.htop = n.htop; s.thr = n.thr; s.orig = n.orig;
  s.nbr0 = n.nbr0; s.nbr1 = n.nbr1; s.nbr2 = n.nbr2; s.info = n.info;
  s.rec = rec;
}

// the node of `rec` straight into the state's registers
__device__ __forceinline__ void loadNodeInto(const DeviceGrid &g, int rec, WalkState &s)
{
  const float4 *q = reinterpret_cast<const float4 *>(g.rec + rec);
  s.p0 = __ldg(q + 0);
  s.p1 = __ldg(q + 1);
  s.p2 = __ldg(q + 2);
  const float4 m = __ldg(q + 3);
  const int4 l = __ldg(reinterpret_cast<const int4 *>(q) + 4);
  s.hbot = m.x; s.htop = m.y; s.thr = m.z; s.orig = __float_as_int(m.w);
  s.nbr0 = l.x; s.nbr1 = l.y; s.nbr2 = l.z; s.info = l.w;
  s.rec = rec;
}

// the record's two lines -> this lane's slot.  Device: 16 asynchronous 16-byte copies (LDGSTS), completed
// by linesWait(); the emulation harness defines ICB_EMU and gets a memcpy.
__device__ __forceinline__ void linesFetch(float *L, const float *__restrict__ src)
{
#if defined(__CUDA_ARCH__)
  const unsigned dst = (unsigned)__cvta_generic_to_shared(L);
#pragma unroll
  for (int i = 0; i < kLineStride / 4; ++i)
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst + 16u * i), "l"(src + 4 * i) : "memory");
#else
  for (int i = 0; i < kLineStride; ++i) L[i] = src[i];
#endif
}
__device__ __forceinline__ void linesWait()
{
#if defined(__CUDA_ARCH__)
  asm volatile("cp.async.wait_all;" ::: "memory");
#endif
}

// the record's lines, and with the prism sampler its transparency mask into the slot's pad word (same round trip)
__device__ __forceinline__ void recLinesFetch(const DeviceGrid &g, int rec, float *L)
{
  linesFetch(L, g.lines + (size_t)rec * kLineStride);
  if (g.prismMask) {
#if defined(__CUDA_ARCH__)
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(L + kSlotMask)), "l"(g.prismMask + rec) : "memory");
#else
    L[kSlotMask] = __int_as_float((int)g.prismMask[rec]);
#endif
  }
}

__device__ __forceinline__ void loadRecFull(const DeviceGrid &g, int rec, WalkState &s, float *L)
{
  recLinesFetch(g, rec, L);
  commitNode(loadNode(g, rec), rec, s);
  linesWait();
}

// first i in [0,n) with r <= H[i+1]; H sorted, H[0] <= r <= H[n], 1 <= n <= 31.  Branch-free binary lifting:
// the answer is the largest i <= n-1 with r > H[i] (0 if none); five probes of the lane's line slot, the same
// instruction stream for every lane whatever its layer.
__device__ __forceinline__ int layerSearchSorted(const float *H, int n, float r)
{
  int i = 0;
#pragma unroll
  for (int step = 16; step >= 1; step >>= 1) {
    const int mid = i + step;
    const float h = H[min(mid, kHeightStride - 1)];
    if (mid < n && r > h) i = mid;
  }
  return i;
}

// the reference's search, probe for probe (ICONGrid.h:117-144) -- exact even for unsorted heights
__device__ __forceinline__ int findHeightLiteral(const float *H, int n, float r)
{
  int first = 0, count = n;
  while (count > 0) {
    const int step = count / 2;
    const int it = first + step;
    if (!(r <= H[it + 1])) { first = it + 1; count -= step + 1; }
    else count = step;
  }
  return first;
}

__device__ __forceinline__ int cubeBin(int N, float x, float y, float z)
{
  const float ax = fabsf(x), ay = fabsf(y), az = fabsf(z);
  int face; float u, w;
  if (ax >= ay && ax >= az) { face = x > 0.f ? 0 : 1; u = y / ax; w = z / ax; }
  else if (ay >= az)        { face = y > 0.f ? 2 : 3; u = x / ay; w = z / ay; }
  else                      { face = z > 0.f ? 4 : 5; u = x / az; w = y / az; }
  const int iu = min(N - 1, max(0, (int)floorf((u * 0.5f + 0.5f) * (float)N)));
  const int iw = min(N - 1, max(0, (int)floorf((w * 0.5f + 0.5f) * (float)N)));
  return (face * N + iw) * N + iu;
}

// 64 bytes (three planes + {hbot, htop, thr, orig}) of a record -> staging slot, asynchronously
__device__ __forceinline__ void stageFetch(float *dst, const RecBlock *__restrict__ src)
{
#if defined(__CUDA_ARCH__)
  const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
  const float *q = reinterpret_cast<const float *>(src);
#pragma unroll
  for (int i = 0; i < 4; ++i)
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(d + 16u * i), "l"(q + 4 * i) : "memory");
#else
  const float *q = reinterpret_cast<const float *>(src);
  for (int i = 0; i < 16; ++i) dst[i] = q[i];
#endif
}

// Exact answer: lowest original record index among all records accepting P, looked for in the
// conservative candidate list of P's cube-map bin.  No topology or ordering assumptions; the one
// shortcut is the walk's own margin rule -- a well-formed column whose bottom record's planes put P
// more than the margin outside rejects P with every record (all its records share the corners).
// Candidates are staged four at a time through the lane's line slot (dead at this point: the caller
// reloads it), so that a query costs a handful of memory round trips instead of one or two per candidate.
// Out of line and fed by value (no references), so that calling it costs the hot loop no stack traffic.
// Returns the internal record; or -2 - r0 when nothing accepts P but P lies deeper than the margin inside the
// column whose bottom record is r0; or -1.
__device__ __noinline__ int exactQueryId(const RecBlock *__restrict__ recs, const ColBlock *__restrict__ cols,
                                         const int *__restrict__ binOff, const int *__restrict__ binList, int binN,
                                         float x, float y, float z, float r, float *stage)
{
  const int bin = cubeBin(binN, x, y, z);
  const int j0 = __ldg(binOff + bin), j1 = __ldg(binOff + bin + 1);
  int best = 0x7fffffff, bestRec = -1, safeRec = -1;
  const float4 *st = reinterpret_cast<const float4 *>(stage);
#pragma unroll 1
  for (int jb = j0; jb < j1; jb += 4) {
    int c[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) c[q] = jb + q < j1 ? __ldg(binList + jb + q) : -1;
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (c[q] >= 0) stageFetch(stage + 16 * q, recs + c[q]);
    linesWait();
    const int nb = min(4, j1 - jb);
#pragma unroll 1
    for (int q = 0; q < nb; ++q) {
      const int r0 = __ldg(binList + jb + q);          // (again, from L1: keeps c[] in registers)
      const float4 p0 = st[4 * q], p1 = st[4 * q + 1], p2 = st[4 * q + 2], m0 = st[4 * q + 3];
      const float e0 = evalPlane(p0, x, y, z), e1 = evalPlane(p1, x, y, z), e2 = evalPlane(p2, x, y, z);
      if (e0 > m0.z || e1 > m0.z || e2 > m0.z) continue;          // clearly outside the column (never for thr = +inf)
      if (e0 < -m0.z && e1 < -m0.z && e2 < -m0.z && safeRec < 0) safeRec = r0;
      const int4 lk = __ldg(reinterpret_cast<const int4 *>(recs + r0) + 4);      // {nbr0, nbr1, nbr2, info}
      int nrec = recCount(lk.w);
      if (lk.w & REC_GENERIC) nrec = __ldg(&cols[__ldg(&recs[r0].col)].nRec);   // (the 8-bit count saturates)
      // bottom record from the staged copy, the others from memory
      if (!(r < m0.x || r > m0.y) && !(e0 > 0.f) && !(e1 > 0.f) && !(e2 > 0.f)) {
        const int orig = __float_as_int(m0.w);
        if (orig < best) { best = orig; bestRec = r0; }
      }
      for (int k = 1; k < nrec; ++k) {
        const int rec = r0 + k;
        const float4 *qv = reinterpret_cast<const float4 *>(recs + rec);
        const float4 m = __ldg(qv + 3);
        if (r < m.x || r > m.y) continue;                        // ICONGrid.h:184
        if (evalPlane(__ldg(qv + 0), x, y, z) > 0.f) continue;   // ICONGrid.h:201-203
        if (evalPlane(__ldg(qv + 1), x, y, z) > 0.f) continue;
        if (evalPlane(__ldg(qv + 2), x, y, z) > 0.f) continue;
        const int orig = __float_as_int(m.w);
        if (orig < best) { best = orig; bestRec = rec; }
      }
    }
  }
  if (bestRec >= 0) return bestRec;
  return safeRec >= 0 ? -2 - safeRec : -1;
}

// radii around r that no record of the (well-formed) column of `rec` covers
__device__ __noinline__ float2 columnGap(const RecBlock *__restrict__ recs, int rec, int info, float r)
{
  const int rec0 = rec - recK(info), n = recCount(info);
  float gapLo = -INFINITY, gapHi = INFINITY;
  for (int k = 0; k < n; ++k) {
    const float4 m = __ldg(reinterpret_cast<const float4 *>(recs + rec0 + k) + 3);
    if (m.y < r) gapLo = fmaxf(gapLo, m.y);
    if (m.x > r) gapHi = fminf(gapHi, m.x);
  }
  return make_float2(gapLo, gapHi);
}

// On a miss the state is either left as it is (whatever it brackets stays true: a bracket is a property of
// positions) or moved to the column P is inside, bracketing the radius gap.
__device__ __forceinline__ bool exactQuery(const DeviceGrid &g, float x, float y, float z, float r, WalkState &s, float *L)
{
  const int rec = exactQueryId(g.rec, g.col, g.binOff, g.binList, g.binN, x, y, z, r, L);
  if (rec < 0) {
    if (rec <= -2) {
      loadRecFull(g, -2 - rec, s, L);
      const float2 gap = columnGap(g.rec, s.rec, s.info, r);
      s.layer = -1; s.lo = gap.x; s.hi = gap.y;
    } else if (s.rec >= 0) {               // the query staged its candidates through the slot: put the lines back
      recLinesFetch(g, s.rec, L);
      linesWait();
    }
    return false;
  }
  loadRecFull(g, rec, s, L);
  s.layer = findHeightLiteral(L, recLayers(s.info), r);
  s.lo = L[s.layer];
  s.hi = L[min(s.layer + 1, kHeightStride - 1)];
  return true;
}

// Start of a walk with no usable history (a new ray, the far side of the inner sphere): the column nearest to the
// centre of P's cube-map bin, bottom or top record whichever is nearer in radius.  Only a starting point -- the
// state machine takes it from there (a hop or two, or the exact query when it cannot).
__device__ __forceinline__ void seedState(const DeviceGrid &g, float rMid, float x, float y, float z, float r, WalkState &s, float *L)
{
  const int bin = cubeBin(g.binN, x, y, z);
  const int2 sd = __ldg(reinterpret_cast<const int2 *>(g.seed) + bin);
  const int rec = r > rMid ? sd.y : sd.x;
  resetState(s);
  if (rec >= 0) loadRecFull(g, rec, s, L);
}

__device__ __forceinline__ int opaque(int v)
{
#if defined(__CUDA_ARCH__)
  asm volatile("" : "+r"(v));      // keeps the compiler from threading the phases of locateAction back apart
#endif
  return v;
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

__device__ __forceinline__ Evals noEvals()
{
  Evals e;
  e.e0 = e.e1 = e.e2 = e.m = INFINITY * 0.f; e.kmax = 0;
  return e;
}

__device__ __forceinline__ void setLayerSorted(const float *L, float r, WalkState &s)
{
  s.layer = layerSearchSorted(L, recLayers(s.info), r);
  s.lo = L[s.layer];
  s.hi = L[s.layer + 1];
}

__device__ __forceinline__ int pick3(int k, int a0, int a1, int a2) { return k == 0 ? a0 : (k == 1 ? a1 : a2); }
__device__ __forceinline__ float others3(int k, float a0, float a1, float a2)
{
  return k == 0 ? fmaxf(a1, a2) : (k == 1 ? fmaxf(a0, a2) : fmaxf(a0, a1));
}

// One action of the locate state machine for the sample at P (radius r, not NaN).  Returns true when the sample
// is decided (`hit`, and for a hit s.rec/s.layer); false when the state moved one record closer and the caller
// must call again with the same P.  Each call performs at most one node fetch, at ONE place in the code, so the
// lanes of a warp that cross into another record this iteration all wait for memory together -- whatever the
// reason for the crossing:
//   same prism        all three planes deeper than the record's margin, r inside the layer bracket: no memory
//   layer change      same record, another layer: searched in the lane's line slot, no global memory
//   hop               clearly beyond side plane k (> margin): fetch the record across that edge (node + lines)
//   vertical          above/below the record's radius range: fetch the next record of the column
//   band              within the margin band of exactly one side plane, the other two deep, r strictly inside:
//                     only this record and the one across the edge can accept P; fetch that node, evaluate both
//                     with their own planes, the lower caller index wins (first match, deviceCode.cu:119-122)
//   anything else     (record-boundary ties, corners, generic columns, off the mesh, too many hops): the exact
//                     candidate query, which assumes nothing
// `e` returns the plane evaluations of P against the final record (NaN when not available); skipCount reuses it.
__device__ __forceinline__ bool locateAction(const DeviceGrid &g, float x, float y, float z, float r, WalkState &s, float *L,
                                             Evals &e, bool &hit, unsigned &nSlow, unsigned &nExact, unsigned &nFetch)
{
  // ---- 1. decide from registers -------------------------------------------------------------------
  int target = -1, kind = 0;          // kind: 1 hop, 2 band, 3 vertical
  bool decided = false, exact = false, search = false;
  hit = false;
  e = noEvals();
  if (s.rec >= 0) {
    e = evalPlanes(s.p0, s.p1, s.p2, x, y, z);
    const int back = s.hops >> 8;
    if (back != 0) {                    // the band test of the last call went to the record we came from: return to it (below)
      ++nSlow;
      target = pick3(back - 1, s.nbr0, s.nbr1, s.nbr2);
      kind = 4;
      s.hops &= 0xff;
      if (target < 0) exact = true;     // (cannot happen: the link we came through has a counterpart)
    } else if (e.m < -s.thr) {
      if (r > s.lo && r < s.hi) { decided = true; hit = s.layer >= 0; }
      else {
        ++nSlow;
        const int k = recK(s.info), nrec = recCount(s.info);
        if (r > s.hbot && r < s.htop) search = true;
        else if (r > s.htop) {
          if (k + 1 < nrec) { target = s.rec + 1; kind = 3; }
          else { s.layer = -1; s.lo = s.htop; s.hi = INFINITY; decided = true; }
        } else if (r < s.hbot) {
          if (k > 0) { target = s.rec - 1; kind = 3; }
          else { s.layer = -1; s.lo = -INFINITY; s.hi = s.hbot; decided = true; }
        } else exact = true;            // r == hbot or r == htop: a tie between stacked records
      }
    } else {
      ++nSlow;
      target = pick3(e.kmax, s.nbr0, s.nbr1, s.nbr2);
      if (e.m > s.thr) {
        kind = 1;
        if (target < 0 || ++s.hops > g.maxWalk) { exact = true; target = -1; }
      } else if (e.m == e.m) {
        const float other = others3(e.kmax, e.e0, e.e1, e.e2);
        const int kb = (s.info >> (REC_KB_SHIFT + 2 * e.kmax)) & 3;
        const bool sameK = ((s.info >> (REC_SAMEK_SHIFT + e.kmax)) & 1) != 0;
        kind = 2;
        if (!(other < -s.thr) || !(r > s.hbot && r < s.htop) || kb == 3 || !sameK || target < 0) { exact = true; target = -1; }
      } else { exact = true; target = -1; }
    }
  } else exact = true;

  // ---- 2. the one fetch site ------------------------------------------------------------------------
  // Every lane that fetches -- hop, vertical move, band test, return from a band test -- loads the target's node STRAIGHT
  // INTO ITS STATE (no staging registers to copy from afterwards) and the target's lines into its slot, and evaluates P
  // against the new planes in the same instructions as the other lanes.  What the decisions below need of the record
  // being left is a handful of scalars saved in front of the load.
  target = opaque(target);
  if (target >= 0) {
    ++nFetch;
    const float oldBot = s.hbot, oldTop = s.htop, oldM = e.m;
    const int oldOrig = s.orig;
    const int kbOld = (s.info >> (REC_KB_SHIFT + 2 * e.kmax)) & 3;      // (band: the shared edge seen from the neighbour)
    const bool up = target > s.rec;
    recLinesFetch(g, target, L);
    loadNodeInto(g, target, s);
    linesWait();
    s.layer = -1; s.lo = 0.f; s.hi = 0.f;
    e = evalPlanes(s.p0, s.p1, s.p2, x, y, z);
    if (s.info & REC_GENERIC) { exact = true; s.rec = -1; }
    else if (kind == 4) search = true;             // back in the record that won the band test (r inside its range: tested there)
    else if (kind != 2) {
      if (e.m < -s.thr) {                          // (else: not safely inside the new record either -- decide again from there)
        if (r > s.hbot && r < s.htop) search = true;
        else if (kind == 3) {                      // a gap between stacked records
          if (up && r < s.hbot) { s.lo = oldTop; s.hi = s.hbot; decided = true; }
          if (!up && r > s.htop) { s.lo = s.htop; s.hi = oldBot; decided = true; }
        }                                          // (else above/below this record: the next call moves vertically)
      }
    } else {
      // band: only the record left (A) and this one (B) can accept P; both with their own planes, the lower caller index
      // wins (first match, deviceCode.cu:119-122).  The shared edge seen from B, its other two planes deep as well (P is
      // far from both end corners).
      const float otherB = others3(kbOld, e.e0, e.e1, e.e2);
      const bool accA = oldM <= 0.f, accB = e.m <= 0.f;
      if (!(e.m == e.m) || !(r > s.hbot && r < s.htop) || !(otherB < -s.thr) || (!accA && !accB)) exact = true;
      else if (accB && (!accA || s.orig < oldOrig)) search = true;         // B: its node and lines are already in place
      else s.hops = (s.hops & 0xff) | ((kbOld + 1) << 8);                   // A: one more fetch, across B's side of the edge
    }
  }

  // ---- 3. the one layer search site (line slot, no global memory) -----------------------------------
  if (opaque(search ? 1 : 0)) { setLayerSorted(L, r, s); decided = true; hit = true; }

  // ---- 4. the exact candidate query ---------------------------------------------------------------------
  if (opaque(exact ? 1 : 0)) {
    ++nExact;
    hit = exactQuery(g, x, y, z, r, s, L);
    e = hit ? evalPlanes(s.p0, s.p1, s.p2, x, y, z) : noEvals();
    decided = true;
  }
  if (decided) s.hops = 0;
  return decided;
}

// Records whose acceptance region is not confined to their column's cone (grid_build.cpp "unbounded":
// zero/NaN plane normals, numLayers == 0, clockwise corners) are checked for every sample; the list is
// sorted by caller index, so the scan ends at the first entry that cannot beat the current answer.
__device__ __noinline__ int globalsQuery(const RecBlock *__restrict__ recs, const int *__restrict__ globalRecs, int nGlobal,
                                         float x, float y, float z, float r, int cur)
{
  for (int j = 0; j < nGlobal; ++j) {
    const int rec = __ldg(globalRecs + j);
    const float4 *q = reinterpret_cast<const float4 *>(recs + rec);
    const float4 m = __ldg(q + 3);
    if (__float_as_int(m.w) >= cur) break;
    if (r < m.x || r > m.y) continue;
    if (evalPlane(__ldg(q + 0), x, y, z) > 0.f) continue;
    if (evalPlane(__ldg(q + 1), x, y, z) > 0.f) continue;
    if (evalPlane(__ldg(q + 2), x, y, z) > 0.f) continue;
    return rec;
  }
  return -1;
}

__device__ __forceinline__ bool applyGlobals(const DeviceGrid &g, float x, float y, float z, float r, WalkState &s, float *L, bool hit)
{
  const int cur = hit ? s.orig : 0x7fffffff;
  const int rec = globalsQuery(g.rec, g.globalRecs, g.nGlobal, x, y, z, r, cur);
  if (rec < 0) return hit;
  loadRecFull(g, rec, s, L);               // thr == +inf: the next sample re-locates from scratch
  s.layer = findHeightLiteral(L, recLayers(s.info), r);
  s.lo = s.hi = 0.f;
  return true;
}

// A position with a NaN coordinate has r == NaN, which fails every rejection of sample() (ICONGrid.h:184,
// 201-203 are all `<`/`>` tests): the brute-force loop returns record 0, layer findHeight(NaN) == numLayers.
__device__ __forceinline__ bool nanPosition(const DeviceGrid &g, WalkState &s, float *L)
{
  loadRecFull(g, g.recOfOrig0, s, L);
  s.thr = INFINITY; s.lo = s.hi = 0.f;
  s.layer = recLayers(s.info);
  return true;
}

// One located sample (scalar form: the parity hooks and the emulation harness; the render kernel calls
// locateAction itself so that a warp's lanes retry together).
__device__ __forceinline__ bool locateStep(const DeviceGrid &g, float x, float y, float z, float r, WalkState &s, float *L,
                                           unsigned &nSlow, unsigned &nExact)
{
  if (!(r == r)) return nanPosition(g, s, L);
  bool hit = false;
  Evals e;
  unsigned nFetch = 0;
#pragma unroll 1
  while (!locateAction(g, x, y, z, r, s, L, e, hit, nSlow, nExact, nFetch)) {}
  if (g.nGlobal == 0) return hit;
  return applyGlobals(g, x, y, z, r, s, L, hit);
}

__device__ __forceinline__ bool exactLocate(const DeviceGrid &g, float x, float y, float z, float r, WalkState &s, float *L)
{
  if (!(r == r)) return nanPosition(g, s, L);
  const bool hit = exactQuery(g, x, y, z, r, s, L);
  if (g.nGlobal == 0) return hit;
  return applyGlobals(g, x, y, z, r, s, L, hit);
}

// ---------------------------------------------------------------------------------------------
// prism sampler: the wedge shape functions of UElems.h:176-187 evaluated at the closed-form
// parametric coordinates of P in the located prism -- what the reference's intersectWedgeEXT (UElems.h:214-311)
// finds by Newton iteration for the wedge {h_i U_k, h_{i+1} U_k}: central projection onto the chord triangle
// gives (r,s), the scale factor gives t (clamped: the predicate's layers are spherical, the wedge's flat).
// Everything the value needs beyond the planes in registers is staged in the lane's slot (kSlot* offsets).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void slotCopy(float *dst, const float *__restrict__ src, int bytes16)
{
#if defined(__CUDA_ARCH__)
  const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
  for (int i = 0; i < bytes16; ++i)
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(d + 16u * i), "l"(src + 4 * i) : "memory");
#else
  for (int i = 0; i < 4 * bytes16; ++i) dst[i] = src[i];
#endif
}
__device__ __forceinline__ void slotCopy4(float *dst, const float *__restrict__ src)
{
#if defined(__CUDA_ARCH__)
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
#else
  *dst = *src;
#endif
}

// scalar of welded vertex v at layer interface j (see vertexScalarsKernel)
__device__ __forceinline__ float vertexScalarAt(const RecBlock *__restrict__ recs, const float *__restrict__ lines,
                                                const ColBlock *__restrict__ cols, const int *__restrict__ colLayers,
                                                const int *__restrict__ vtxOff, const int *__restrict__ vtxCols, size_t v, int j)
{
  float acc = 0.f;
  int cnt = 0;
  for (int q = __ldg(vtxOff + v); q < __ldg(vtxOff + v + 1); ++q) {
    const int c = __ldg(vtxCols + q);
    const int Lc = __ldg(colLayers + c);
    if (Lc < 1 || j > Lc) continue;
    // values of global layers g0 = max(j-1,0) and g1 = min(j, Lc-1) of the column
    const int g0 = max(j - 1, 0), g1 = min(j, Lc - 1);
    float v0 = 0.f, v1 = 0.f;
    int rec = __ldg(&cols[c].recBegin), b = 0;
    for (int k = __ldg(&cols[c].nRec); k > 0; --k, ++rec) {
      const int n = __ldg(&recs[rec].nLayers);
      if (g0 >= b && g0 < b + n) v0 = __ldg(lines + (size_t)rec * kLineStride + kHeightStride + (g0 - b));
      if (g1 >= b && g1 < b + n) v1 = __ldg(lines + (size_t)rec * kLineStride + kHeightStride + (g1 - b));
      b += n;
      if (b > g1) break;
    }
    const float I = (j == 0) ? v0 : (j == Lc ? v1 : __fmul_rn(__fadd_rn(v0, v1), 0.5f));
    acc = __fadd_rn(acc, I);
    ++cnt;
  }
  return cnt > 0 ? __fdiv_rn(acc, (float)cnt) : 0.f;
}

// the node's prism tail (48 bytes: the rest of the node's cache line) -> slot
__device__ __forceinline__ void prismFetchTail(const DeviceGrid &g, int rec, float *L)
{
  slotCopy(L + kSlotTail, reinterpret_cast<const float *>(g.rec + rec) + 20, 3);
  linesWait();
}

// the six vertex scalars of (record whose tail is in the slot, layer) -> slot; also the layer's bottom height and thickness
__device__ __forceinline__ void prismFetchScalars(const DeviceGrid &g, int layer, float *L)
{
  const int li = min(max(layer, 0), 30);
  const size_t j = (size_t)(__float_as_int(L[kSlotBase]) + li);
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const float *src = g.vtxScalars + (size_t)__float_as_int(L[kSlotVert + k]) * (size_t)g.vtxStride + j;
    slotCopy4(L + kSlotScal + k, src);
    slotCopy4(L + kSlotScal + 3 + k, src + 1);
  }
  const float h0 = L[li], h1 = L[li + 1];
  L[kSlotH0] = h0;
  L[kSlotDh] = __fsub_rn(h1, h0);
  linesWait();
}

// value of the prism at P; false when it is not a number (such a sample adds nothing)
__device__ __forceinline__ bool prismValue(const float4 &p0, const float4 &p1, const float4 &p2, const float *L,
                                           float x, float y, float z, float &val)
{
  const float e0 = evalPlane(p0, x, y, z), e1 = evalPlane(p1, x, y, z), e2 = evalPlane(p2, x, y, z);
  const float a0 = __fmul_rn(e1, L[kSlotKappa + 1]), a1 = __fmul_rn(e2, L[kSlotKappa + 2]), a2 = __fmul_rn(e0, L[kSlotKappa]);
  const float sum = __fadd_rn(__fadd_rn(a0, a1), a2);
  const float r = __fdiv_rn(a1, sum), s = __fdiv_rn(a2, sum);
  const float rho = dot3(x, y, z, L[kSlotM], L[kSlotM + 1], L[kSlotM + 2]);
  const float h0 = L[kSlotH0], d = L[kSlotDh];
  float t = d > 0.f ? __fdiv_rn(__fsub_rn(rho, h0), d) : 0.f;
  t = fminf(fmaxf(t, 0.f), 1.f);
  const float w0 = __fsub_rn(__fsub_rn(1.f, r), s), om = __fsub_rn(1.f, t);
  float v = 0.f;
  v = __fadd_rn(v, __fmul_rn(__fmul_rn(w0, om), L[kSlotScal + 0]));
  v = __fadd_rn(v, __fmul_rn(__fmul_rn(r, om), L[kSlotScal + 1]));
  v = __fadd_rn(v, __fmul_rn(__fmul_rn(s, om), L[kSlotScal + 2]));
  v = __fadd_rn(v, __fmul_rn(__fmul_rn(w0, t), L[kSlotScal + 3]));
  v = __fadd_rn(v, __fmul_rn(__fmul_rn(r, t), L[kSlotScal + 4]));
  v = __fadd_rn(v, __fmul_rn(__fmul_rn(s, t), L[kSlotScal + 5]));
  val = v;
  return v == v;
}

// one-off form (parity hooks, Woodcock): stage everything for (s.rec, s.layer), then evaluate
__device__ __forceinline__ bool prismSample(const DeviceGrid &g, const WalkState &s, float *L, float x, float y, float z, float &val)
{
  prismFetchTail(g, s.rec, L);
  prismFetchScalars(g, s.layer, L);
  return prismValue(s.p0, s.p1, s.p2, L, x, y, z, val);
}

// ---------------------------------------------------------------------------------------------
// classification / ray set-up
// ---------------------------------------------------------------------------------------------
// a / b, IEEE round-to-nearest -- __fdiv_rn, except that a zero numerator over a finite non-zero denominator is answered
// directly (the correctly signed zero): the hardware divide sends every zero numerator through its ~65-instruction slow
// path, and the optical depth of a transparent sample, -0 * dt / unitDistance, is exactly that case on nearly every prism
__device__ __forceinline__ float divExact(float a, float b)
{
  if (a == 0.f && b != 0.f && fabsf(b) < INFINITY) return __int_as_float((__float_as_int(a) ^ __float_as_int(b)) & (int)0x80000000);
  return __fdiv_rn(a, b);
}

// 1 - exp(-sigma * dt / unitDistance), the sample opacity of SURVEY.md A.3, operation for operation; a zero optical depth
// (expf(+-0) = 1, 1 - 1 = +0) skips the exponential
__device__ __forceinline__ float sampleOpacity(float sigma, float dt, float unitDistance)
{
  const float q = divExact(__fmul_rn(-sigma, dt), unitDistance);
  if (q == 0.f) return 0.f;
  return __fsub_rn(1.f, expf(q));
}

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
  __device__ __forceinline__ unsigned nextBits()       // the 24 bits next() would have returned as k / 2^24
  {
    state = 1664525u * state + 1013904223u;
    return state & 0x00FFFFFFu;
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

// generateRay (deviceCode.cu:36-49); rnd keeps its state for the integrators that go on drawing from it
__device__ __forceinline__ void generateRayDir(const RenderParams &p, int px, int py, Lcg &rnd, bool jit, float &dx, float &dy, float &dz)
{
  rnd.init((unsigned)(p.accumID * p.width * p.height) + (unsigned)px, (unsigned)py);
  // deviceCode.cu:42-43 leaves the order of its two rnd() calls unsequenced; the reference as built by
  // g++ (oracle/_ref) draws the dir_dv sample first, then the dir_du sample.
  const float jv = jit ? rnd.next() : 0.f;
  const float b = __fadd_rn(__fadd_rn((float)py, .5f), jv);
  const float ju = jit ? rnd.next() : 0.f;
  const float a = __fadd_rn(__fadd_rn((float)px, .5f), ju);
  dx = __fadd_rn(__fadd_rn(p.d00[0], __fmul_rn(a, p.du[0])), __fmul_rn(b, p.dv[0]));
  dy = __fadd_rn(__fadd_rn(p.d00[1], __fmul_rn(a, p.du[1])), __fmul_rn(b, p.dv[1]));
  dz = __fadd_rn(__fadd_rn(p.d00[2], __fmul_rn(a, p.du[2])), __fmul_rn(b, p.dv[2]));
  const float len = __fsqrt_rn(dot3(dx, dy, dz, dx, dy, dz));
  dx = __fdiv_rn(dx, len); dy = __fdiv_rn(dy, len); dz = __fdiv_rn(dz, len);
  if (fabsf(dx) < 1e-5f) dx = 1e-5f;
  if (fabsf(dy) < 1e-5f) dy = 1e-5f;
  if (fabsf(dz) < 1e-5f) dz = 1e-5f;
}

__device__ __forceinline__ void setupRay(const RenderParams &p, int px, int py, RaySetup &r)
{
  Lcg rnd;
  float dx, dy, dz;
  generateRayDir(p, px, py, rnd, (p.flags & 1u) != 0, dx, dy, dz);
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
__device__ __forceinline__ float divApprox(float a, float b)
{
#ifdef __CUDA_ARCH__
  float r;                                   // MUFU.RCP + FMUL: two instructions (__fdividef adds a range test and a rescale)
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(b));
  return __fmul_rn(a, r);
#else
  return a / b;
#endif
}
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
__device__ __forceinline__ int skipCount(const WalkState &s, const Evals &e, const RenderParams &p, float dx, float dy, float dz,
                                         float x, float y, float z, float r, float dt)
{
  const float thr2 = __fmul_rn(s.thr, p.skipPlaneScale);
  if (!(e.m < -thr2)) return 0;
  const float loS = s.lo + p.skipRadSlack, hiS = s.hi - p.skipRadSlack;
  if (!(r > loS && r < hiS)) return 0;
  float D = (float)p.maxSkip * dt;
  const float B0 = __fmaf_rn(dz, s.p0.z, __fmaf_rn(dy, s.p0.y, __fmul_rn(dx, s.p0.x)));
  const float B1 = __fmaf_rn(dz, s.p1.z, __fmaf_rn(dy, s.p1.y, __fmul_rn(dx, s.p1.x)));
  const float B2 = __fmaf_rn(dz, s.p2.z, __fmaf_rn(dy, s.p2.y, __fmul_rn(dx, s.p2.x)));
  if (B0 > 0.f) D = fminf(D, divApprox(-thr2 - e.e0, B0));
  if (B1 > 0.f) D = fminf(D, divApprox(-thr2 - e.e1, B1));
  if (B2 > 0.f) D = fminf(D, divApprox(-thr2 - e.e2, B2));
  // (|dir|^2 = 1 to a few ulps -- generateRay normalises and then only lifts components below 1e-5 -- which is far
  // inside the slack: the quadratic's leading coefficient is taken as 1)
  const float b = __fmaf_rn(z, dz, __fmaf_rn(y, dy, __fmul_rn(x, dx)));
  if (hiS < INFINITY) {                                           // leaves the bracket through its top
    const float cHi = (r - hiS) * (r + hiS);                      // < 0
    const float sqHi = sqrtApprox(fmaxf(__fmaf_rn(b, b, -cHi), 0.f));
    D = fminf(D, b < 0.f ? sqHi - b : divApprox(-cHi, b + sqHi));
  }
  if (b < 0.f && loS > 0.f) {                                     // heading inwards: may reach the bracket's bottom
    const float cLo = (r - loS) * (r + loS);                      // > 0
    const float disc = __fmaf_rn(b, b, -cLo);
    if (disc > 0.f) D = fminf(D, divApprox(cLo, sqrtApprox(disc) - b));
  }
  const int n = (int)divApprox(D, dt) - p.skipBack;
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

// Chunked compositing (DESIGN.md "split rays"; CPU statements: oracle/ref_driver.cpp and oracle/icon_oracle.c march_pixel
// with chunk > 0).  `loc` composites the samples of the current chunk front to back from zero (the A.3 recurrence);
// `tot` holds the chunks merged so far.  A chunk = 2^chunkLog2 consecutive sample indices of one shell segment; `key`
// names the chunk of the sample being added (merging is lazy: a chunk nothing was added to merges to exactly nothing).
__device__ __forceinline__ void mergeChunk(float4 &tot, float4 &loc)
{
  const float T = __fsub_rn(1.f, tot.w);
  tot.x = __fadd_rn(tot.x, __fmul_rn(T, loc.x));
  tot.y = __fadd_rn(tot.y, __fmul_rn(T, loc.y));
  tot.z = __fadd_rn(tot.z, __fmul_rn(T, loc.z));
  tot.w = __fadd_rn(tot.w, __fmul_rn(T, loc.w));
  loc = make_float4(0.f, 0.f, 0.f, 0.f);
}
__device__ __forceinline__ int chunkKey(int segLeft, int sampleIdx, int chunkLog2) { return (segLeft << 24) | (sampleIdx >> chunkLog2); }

// adds one sample (opacity a, colour c) to its chunk; true when the ray's merged opacity has reached termAlpha
__device__ __forceinline__ bool compositeSample(float4 &tot, float4 &loc, int &curKey, int key, float a, float cr, float cg, float cb,
                                                float termAlpha)
{
  if (key != curKey) { mergeChunk(tot, loc); curKey = key; }
  const float w = __fmul_rn(__fsub_rn(1.f, loc.w), a);
  loc.x = __fadd_rn(loc.x, __fmul_rn(w, cr));
  loc.y = __fadd_rn(loc.y, __fmul_rn(w, cg));
  loc.z = __fadd_rn(loc.z, __fmul_rn(w, cb));
  loc.w = __fadd_rn(loc.w, w);
  return __fadd_rn(tot.w, __fmul_rn(__fsub_rn(1.f, tot.w), loc.w)) >= termAlpha;
}

// The marcher's compositing step.  CHUNKED = false: the plain front-to-back recurrence of SURVEY.md A.3 (the default);
// CHUNKED = true: the chunked form above, which split rays need (icb_set_ray_split mode != NEVER; RenderParams.chunkLog2 > 0).
template <bool CHUNKED>
__device__ __forceinline__ bool compositeStep(float4 &tot, float4 &loc, int &curKey, int key, float a, float cr, float cg, float cb,
                                              float termAlpha)
{
  if (CHUNKED) return compositeSample(tot, loc, curKey, key, a, cr, cg, cb, termAlpha);
  const float w = __fmul_rn(__fsub_rn(1.f, tot.w), a);
  tot.x = __fadd_rn(tot.x, __fmul_rn(w, cr));
  tot.y = __fadd_rn(tot.y, __fmul_rn(w, cg));
  tot.z = __fadd_rn(tot.z, __fmul_rn(w, cb));
  tot.w = __fadd_rn(tot.w, w);
  return tot.w >= termAlpha;
}
template <bool CHUNKED>
__device__ __forceinline__ void mergeLast(float4 &tot, float4 &loc)
{
  if (CHUNKED) mergeChunk(tot, loc);
}

// chunks of a shell segment [lo, hi): an upper bound on its sample count, rounded up to whole chunks (the same fp32
// expression wherever it is evaluated); -1 when the segment is too long to split (or not a number)
__device__ __forceinline__ int segmentChunks(float lo, float hi, float dt, int chunkLog2)
{
  const float v = __fdiv_rn(__fsub_rn(hi, lo), dt);
  if (!(v < (float)(kMaxChunks << chunkLog2))) return -1;
  const int n = v > 0.f ? (int)v + 1 : 0;
  return (n + (1 << chunkLog2) - 1) >> chunkLog2;
}

// One pixel marched by one thread: the statement the render kernel's lanes
// follow step for step.  ids (optional) receives the located record id of every step (-1 = miss).
// L: this thread's 64-float line slot.
__device__ inline void marchPixel(const DeviceGrid &g, const RenderParams &p, const float4 *lut, int px, int py,
                                  int *ids, int maxIds, float4 *rgbaOut, int *nStepsOut, float *L, bool useSkip = false,
                                  int *nSkippedOut = nullptr)
{
  RaySetup rs;
  setupRay(p, px, py, rs);
  float4 C = make_float4(0.f, 0.f, 0.f, 0.f), loc = make_float4(0.f, 0.f, 0.f, 0.f);
  int curKey = -1;
  WalkState s;
  resetState(s);
  unsigned nSlow = 0, nExact = 0;
  int n = 0, nSkipped = 0;
  bool done = false;
  const bool skipOn = useSkip && p.maxSkip > 0 && g.nGlobal == 0;
  const float rMid = 0.5f * (p.rInner + p.rOuter);
  for (int sg = 0; sg < rs.nseg && !done; ++sg) {
    s.rec = -2;                          // no walk history at the start of a segment: seed (as the render kernel does)
    for (int i = 0;; ++i) {
      const float t = sampleT(rs.seg[sg][0], i, p.dt);
      if (t >= rs.seg[sg][1]) break;
      const float x = __fadd_rn(p.org[0], __fmul_rn(rs.dx, t));
      const float y = __fadd_rn(p.org[1], __fmul_rn(rs.dy, t));
      const float z = __fadd_rn(p.org[2], __fmul_rn(rs.dz, t));
      const float r = __fsqrt_rn(dot3(x, y, z, x, y, z));
      bool hit = false;
      Evals e = noEvals();
      if (s.rec == -2) {
        if (r == r) seedState(g, rMid, x, y, z, r, s, L);
        else s.rec = -1;
      }
      if (!(r == r)) hit = nanPosition(g, s, L);
      else {
        unsigned nFetch = 0;
        while (!locateAction(g, x, y, z, r, s, L, e, hit, nSlow, nExact, nFetch)) {}
        if (g.nGlobal != 0) hit = applyGlobals(g, x, y, z, r, s, L, hit);
      }
      const int id = hit ? s.orig : -1;
      if (ids && n < maxIds) ids[n] = id;
      ++n;
      int reps = 1;
      if (skipOn && r == r && (hit || s.layer < 0))
        reps += clampSkip(skipCount(s, e, p, rs.dx, rs.dy, rs.dz, x, y, z, r, p.dt), i + 1, rs.seg[sg][0], rs.seg[sg][1], t, p.dt);
      if (!hit) {                            // skipped misses: same bracket, by construction
        for (int q = 1; q < reps; ++q) { if (ids && n < maxIds) ids[n] = -1; ++n; ++i; ++nSkipped; }
        continue;
      }
      float v = L[kHeightStride + s.layer];
      if (g.sampler != 0) {                // prism sampler: stage the prism once, evaluate it per sample
        prismFetchTail(g, s.rec, L);
        prismFetchScalars(g, s.layer, L);
      }
      float4 sm = make_float4(0.f, 0.f, 0.f, 0.f);
      float a = 0.f, cr = 0.f, cg = 0.f, cb = 0.f;
      for (int q = 0; q < reps; ++q) {
        float xq = x, yq = y, zq = z;
        if (q > 0) {                       // a skipped sample: same record, same layer, by construction
          if (ids && n < maxIds) ids[n] = id;
          ++n; ++i; ++nSkipped;
          if (g.sampler != 0) {
            const float tq = sampleT(rs.seg[sg][0], i, p.dt);
            xq = __fadd_rn(p.org[0], __fmul_rn(rs.dx, tq));
            yq = __fadd_rn(p.org[1], __fmul_rn(rs.dy, tq));
            zq = __fadd_rn(p.org[2], __fmul_rn(rs.dz, tq));
          }
        }
        if (q == 0 || g.sampler != 0) {
          if (g.sampler != 0 && !prismValue(s.p0, s.p1, s.p2, L, xq, yq, zq, v)) continue;
          sm = postClassify(lut, p, v);
          a = sampleOpacity(sm.w, p.dt, p.unitDistance);
          cr = __fmul_rn(__fmul_rn(sm.x, p.ambient[0]), p.ambientRadiance);
          cg = __fmul_rn(__fmul_rn(sm.y, p.ambient[1]), p.ambientRadiance);
          cb = __fmul_rn(__fmul_rn(sm.z, p.ambient[2]), p.ambientRadiance);
        }
        const bool sat = p.chunkLog2 > 0 ? compositeStep<true>(C, loc, curKey, chunkKey(rs.nseg - 1 - sg, i, p.chunkLog2), a, cr, cg, cb, p.termAlpha)
                                         : compositeStep<false>(C, loc, curKey, 0, a, cr, cg, cb, p.termAlpha);
        if (sat) { done = true; break; }
      }
      if (done) break;
    }
  }
  if (p.chunkLog2 > 0) mergeChunk(C, loc);
  *rgbaOut = C;
  *nStepsOut = n;
  if (nSkippedOut) *nSkippedOut = nSkipped;
}

// The reference's own integrator for one pixel: woodcockTrackingAE (deviceCode.cu:239-275) = generateRay, boxTest
// against the volume bounds (vecmath.h:1927-1938), woodcockTracking with majorant 1 (deviceCode.cu:149-186).
// logf(1 - k/2^24) is glibc's, bit for bit: computed (glibc_logf.cuh; logTable == nullptr) or, as in round 1, read from
// logTable[k] built by the host's libm (log_table.cpp; ICB_LOG_TABLE=1), so every free-flight distance, sampled
// position, located record and accepted collision is the one the reference's CPU path computes.
// Returns false when the ray misses the box (the reference returns without touching the pixel).
__device__ inline bool woodcockPixel(const DeviceGrid &g, const RenderParams &p, const float4 *lut, const float *__restrict__ logTable,
                                     int px, int py, float *L, float4 &Cout, unsigned &nSteps, unsigned &nHits)
{
  Lcg rnd;
  float dx, dy, dz;
  generateRayDir(p, px, py, rnd, true, dx, dy, dz);
  float t0, t1;
  {
    const float lx = __fdiv_rn(__fsub_rn(p.volLo[0], p.org[0]), dx), hx = __fdiv_rn(__fsub_rn(p.volHi[0], p.org[0]), dx);
    const float ly = __fdiv_rn(__fsub_rn(p.volLo[1], p.org[1]), dy), hy = __fdiv_rn(__fsub_rn(p.volHi[1], p.org[1]), dy);
    const float lz = __fdiv_rn(__fsub_rn(p.volLo[2], p.org[2]), dz), hz = __fdiv_rn(__fsub_rn(p.volHi[2], p.org[2]), dz);
    t0 = fmaxf(0.f, fmaxf(fmaxf(fminf(lx, hx), fminf(ly, hy)), fminf(lz, hz)));
    t1 = fminf(1e10f, fminf(fminf(fmaxf(lx, hx), fmaxf(ly, hy)), fmaxf(lz, hz)));
  }
  if (!(t0 < t1)) return false;
  const float majorant = 1.f;
  const float rate = __fdiv_rn(majorant, p.unitDistance);
  const float rMid = 0.5f * (p.rInner + p.rOuter);
  float ar = 0.f, ag = 0.f, ab = 0.f, extinction = 0.f;
  float t = t0;
  WalkState s;
  resetState(s);
  s.rec = -2;
  unsigned nSlow = 0, nExact = 0, nFetch = 0;
  // one loop, two steps per iteration: a free flight for the lanes that have no sample to locate, then ONE locateAction
  // for every lane (an undecided lane comes back with the same sample) -- the reference's nesting
  // (woodcockTracking -> sampleVolume) would leave the warp waiting on its slowest locate at every tracking step
  bool have = false;
  float x = 0.f, y = 0.f, z = 0.f, r = 0.f;
#pragma unroll 1
  for (;;) {
    if (!have) {
      const unsigned bits = rnd.nextBits();
      t = __fsub_rn(t, __fdiv_rn(logTable ? logTable[bits] : glibc::logOneMinus<true>(bits), rate));
      if (t > t1) break;
      x = __fadd_rn(p.org[0], __fmul_rn(dx, t));
      y = __fadd_rn(p.org[1], __fmul_rn(dy, t));
      z = __fadd_rn(p.org[2], __fmul_rn(dz, t));
      r = __fsqrt_rn(dot3(x, y, z, x, y, z));
      ++nSteps;
      // outside every record's radius range (ICONGrid.h:184): a miss, whatever column lies below -- no walking along
      // the ray while it is outside the shell; the walk is seeded afresh where the ray comes back in
      if (r < g.cullLo || r > g.cullHi) { s.rec = -2; continue; }
      if (s.rec == -2 && r == r) seedState(g, rMid, x, y, z, r, s, L);
      have = true;
    }
    bool hit = false;
    if (!(r == r)) hit = nanPosition(g, s, L);
    else {
      Evals e;
      if (!locateAction(g, x, y, z, r, s, L, e, hit, nSlow, nExact, nFetch)) continue;
      if (g.nGlobal != 0) hit = applyGlobals(g, x, y, z, r, s, L, hit);
    }
    have = false;
    if (!hit) continue;
    ++nHits;
    float v = L[kHeightStride + s.layer];
    if (g.sampler != 0) {
      if (!prismSample(g, s, L, x, y, z, v)) continue;      // (overwrites the slot's value half only; the walk reads heights)
    }
    const float4 sm = postClassify(lut, p, v);
    const float u = rnd.next();
    if (sm.w >= __fmul_rn(u, majorant)) { ar = sm.x; ag = sm.y; ab = sm.z; extinction = sm.w; break; }
  }
  Cout.x = __fmul_rn(__fmul_rn(ar, p.ambient[0]), p.ambientRadiance);
  Cout.y = __fmul_rn(__fmul_rn(ag, p.ambient[1]), p.ambientRadiance);
  Cout.z = __fmul_rn(__fmul_rn(ab, p.ambient[2]), p.ambientRadiance);
  Cout.w = extinction > 0.f ? 1.f : 0.f;
  return true;
}

// ---------------------------------------------------------------------------------------------
// The reference's DEFAULT ray-gen: woodcockTrackingWithAccel (deviceCode.cu:281-341) in GRID_ACCEL_MODE.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int gridCoord(float v, float lo, float hi, int dim)     // projectOnGrid, DDA.h:24-32
{
  const float s = __fmul_rn(__fdiv_rn(__fsub_rn(v, lo), __fsub_rn(hi, lo)), (float)dim);
  return max(0, min((int)s, dim - 1));
}

// One ray of the default ray-gen between two iterations of its loop.  The nested loops of the reference (dda3 ->
// per-macrocell woodcockTracking -> sampleVolume) run as ONE loop over a small state machine, so that the lanes of a warp
// meet at the expensive step -- locating a sample -- every iteration, whatever macrocell or tracking step each is in:
//   phase 0  advance the DDA to the next macrocell with a positive majorant (empty ones cost no random number,
//            deviceCode.cu:160-161), or leave the grid
//   phase 1  one free flight; beyond the macrocell's t1 -> step the DDA (phase 0), else the sample position (phase 2)
//   phase 2  one locateAction; undecided lanes come back; a located sample is classified and accepted/rejected
//            (accept strictly inside (t0,t1) ends the ray, deviceCode.cu:303-311; otherwise the DDA steps on)
struct TrackRay {              // what the per-cell woodcockTracking + sampleVolume keep (phases 1 and 2)
  Lcg rnd;
  float d[3];
  float c0, c1, majorant, rate, t;
  float x, y, z, r;
  int phase;
  bool drew;                   // a random number was drawn since the last free flight (which staged entry that flight takes)
  WalkState s;
  float4 C;                    // the pixel's result once the step function returns true
};

struct WgRay : TrackRay {
  float rt0, dtmax;
  float tn[3], dist[3];
  int cell[3];
  float t0;
  bool first;                  // the first macrocell is visited before any DDA step
};

// generateRay, boxTest, dda3 set-up (DDA.h:35-109: shifted ray copy, projectOnGrid).  False when the ray misses the
// volume's box (the reference returns without touching the pixel).
__device__ __forceinline__ void wgStageLog(const float *__restrict__ logTable, unsigned state, float *L)
{
  // The table entry of a free flight is a dependent random read (64 MB table) right in front of the locate.  The ray's
  // random sequence is known in advance: between two flights it draws nothing (a miss, or the DDA steps on) or exactly
  // one number (the accept/reject test of a located sample), so the next flight's entry is one of two -- both are
  // staged into the lane's slot while the current sample is being located.
  if (!logTable) return;             // (the flights compute logf(1-u) instead, glibc_logf.cuh: nothing to stage)
  const unsigned s1 = 1664525u * state + 1013904223u, s2 = 1664525u * s1 + 1013904223u;
  slotCopy4(L + kSlotLog, logTable + (s1 & 0x00FFFFFFu));
  slotCopy4(L + kSlotLog + 1, logTable + (s2 & 0x00FFFFFFu));
}

__device__ __forceinline__ bool wgInit(const RenderParams &p, const float *__restrict__ logTable, const MajorantGrid &mg, int px, int py, WgRay &w, float *L)
{
  generateRayDir(p, px, py, w.rnd, true, w.d[0], w.d[1], w.d[2]);
  float rt1;
  {
    float lo3[3], hi3[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const float l = __fdiv_rn(__fsub_rn(p.volLo[k], p.org[k]), w.d[k]), h = __fdiv_rn(__fsub_rn(p.volHi[k], p.org[k]), w.d[k]);
      lo3[k] = fminf(l, h); hi3[k] = fmaxf(l, h);
    }
    w.rt0 = fmaxf(0.f, fmaxf(fmaxf(lo3[0], lo3[1]), lo3[2]));
    rt1 = fminf(1e10f, fminf(fminf(hi3[0], hi3[1]), hi3[2]));
  }
  if (!(w.rt0 < rt1)) return false;
  w.C = make_float4(0.f, 0.f, 0.f, 0.f);
  resetState(w.s);
  w.s.rec = -2;
  w.dtmax = __fsub_rn(rt1, w.rt0);
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const float o = __fadd_rn(p.org[k], __fmul_rn(w.rt0, w.d[k]));
    const float rcp = __fdiv_rn(1.f, w.d[k]);
    const float l = __fmul_rn(__fsub_rn(mg.lo[k], o), rcp), h = __fmul_rn(__fsub_rn(mg.hi[k], o), rcp);
    float tnear = fminf(l, h);
    const float tfar = fmaxf(l, h);
    if (w.d[k] == 0.f) tnear = 3.402823466e+38f;
    w.cell[k] = gridCoord(o, mg.lo[k], mg.hi[k], mg.dims[k]);
    w.dist[k] = fmaxf(0.f, __fdiv_rn(__fsub_rn(tfar, tnear), (float)mg.dims[k]));
    w.tn[k] = w.d[k] > 0.f ? __fadd_rn(tnear, __fmul_rn((float)(w.cell[k] + 1), w.dist[k]))
                           : __fadd_rn(tnear, __fmul_rn((float)(mg.dims[k] - w.cell[k]), w.dist[k]));
  }
  w.t0 = 0.f; w.phase = 0; w.first = true; w.drew = false;
  wgStageLog(logTable, w.rnd.state, L);
  w.c0 = w.c1 = w.majorant = w.t = 0.f; w.rate = 1.f;
  w.x = w.y = w.z = w.r = 0.f;
  return true;
}

// phases 1 and 2 of a ray (one free flight | one locateAction): shared by the grid and the shell traversal
__device__ __forceinline__ bool trackStep(const DeviceGrid &g, const RenderParams &p, const float4 *lut, const float *__restrict__ logTable,
                                          TrackRay &w, float *L, float rMid, unsigned &nSteps, unsigned &nHits,
                                          unsigned &nSlow, unsigned &nExact, unsigned &nFetch)
{
  if (w.phase == 1) {
    const unsigned bits = w.rnd.nextBits();
    float lg;
    if (logTable) { linesWait(); lg = L[kSlotLog + (w.drew ? 1 : 0)]; }
    else lg = glibc::logOneMinus<true>(bits);
    w.drew = false;
    wgStageLog(logTable, w.rnd.state, L);
    w.t = __fsub_rn(w.t, __fdiv_rn(lg, w.rate));
    if (w.t > w.c1) w.phase = 0;      // woodcockTracking returns tmax: not inside (t0,t1) -> traverse on
    else {
      w.x = __fadd_rn(p.org[0], __fmul_rn(w.d[0], w.t));
      w.y = __fadd_rn(p.org[1], __fmul_rn(w.d[1], w.t));
      w.z = __fadd_rn(p.org[2], __fmul_rn(w.d[2], w.t));
      w.r = __fsqrt_rn(dot3(w.x, w.y, w.z, w.x, w.y, w.z));
      ++nSteps;
      if (w.r < g.cullLo || w.r > g.cullHi) w.s.rec = -2;      // (radial cull as in woodcockPixel: a miss, the next free flight)
      else {
        if (w.s.rec == -2 && w.r == w.r) seedState(g, rMid, w.x, w.y, w.z, w.r, w.s, L);
        w.phase = 2;
      }
    }
  }
  if (w.phase == 2) {
    bool hit = false, decided = true;
    if (!(w.r == w.r)) hit = nanPosition(g, w.s, L);
    else {
      Evals e;
      decided = locateAction(g, w.x, w.y, w.z, w.r, w.s, L, e, hit, nSlow, nExact, nFetch);
      if (decided && g.nGlobal != 0) hit = applyGlobals(g, w.x, w.y, w.z, w.r, w.s, L, hit);
    }
    if (decided) {
      w.phase = 1;                // (a miss or a rejected collision: the next free flight)
      if (hit) {
        ++nHits;
        float v = L[kHeightStride + w.s.layer];
        const bool ok = g.sampler == 0 || prismSample(g, w.s, L, w.x, w.y, w.z, v);
        if (ok) {
          const float4 sm = postClassify(lut, p, v);
          const float u = w.rnd.next();
          w.drew = true;
          if (sm.w >= __fmul_rn(u, w.majorant)) {
            if (w.t > w.c0 && w.t < w.c1) {
              w.C.x = __fmul_rn(__fmul_rn(sm.x, p.ambient[0]), p.ambientRadiance);
              w.C.y = __fmul_rn(__fmul_rn(sm.y, p.ambient[1]), p.ambientRadiance);
              w.C.z = __fmul_rn(__fmul_rn(sm.z, p.ambient[2]), p.ambientRadiance);
              w.C.w = sm.w > 0.f ? 1.f : 0.f;
              return true;
            }
            w.phase = 0;          // a collision exactly on the macrocell's boundary is dropped by the callback: traverse on
          }
        }
      }
    }
  }
  return false;
}

// one iteration of the ray's loop; true when the ray has ended (w.C = colour, alpha)
__device__ __forceinline__ bool wgStep(const DeviceGrid &g, const RenderParams &p, const float4 *lut, const float *__restrict__ logTable,
                                       const MajorantGrid &mg, WgRay &w, float *L, float rMid, unsigned &nSteps, unsigned &nHits,
                                       unsigned &nSlow, unsigned &nExact, unsigned &nFetch)
{
  if (w.phase == 0) {
    bool out = false;
#pragma unroll 1
    for (;;) {
      if (!w.first) {             // the DDA step that follows a macrocell's callback (DDA.h:111-132)
        const float tmin3 = fminf(fminf(w.tn[0], w.tn[1]), w.tn[2]);
        w.t0 = fminf(tmin3, w.dtmax);
#pragma unroll
        for (int k = 0; k < 3; ++k)
          if (!out && w.tn[k] == tmin3) {
            w.tn[k] = __fadd_rn(w.tn[k], w.dist[k]);
            w.cell[k] += w.d[k] > 0.f ? 1 : -1;
            if (w.cell[k] == (w.d[k] > 0.f ? mg.dims[k] : -1)) out = true;
          }
        if (out) break;
      }
      w.first = false;
      const float t1 = fminf(fminf(fminf(w.tn[0], w.tn[1]), w.tn[2]), w.dtmax);
      // (cell index in 32-bit arithmetic: icb_set_accel_grid admits at most 2^30 cells)
      const int leaf = (w.cell[2] * mg.dims[1] + w.cell[1]) * mg.dims[0] + w.cell[0];
      // (the bit mask keeps the DDA's chain of dependent reads in L1/L2: 2 MB for 256^3 against 64 MB of majorants)
      const bool occ = mg.occupied == nullptr || ((__ldg(mg.occupied + (leaf >> 5)) >> (unsigned)(leaf & 31)) & 1u) != 0u;
      w.majorant = occ ? __ldg(mg.maxOpacities + leaf) : 0.f;
      if (w.majorant > 0.f) {
        w.c0 = __fadd_rn(w.rt0, w.t0); w.c1 = __fadd_rn(w.rt0, t1);
        w.rate = __fdiv_rn(w.majorant, p.unitDistance);
        w.t = w.c0;
        w.phase = 1;
        break;
      }
    }
    if (out) return true;
  }
  return trackStep(g, p, lut, logTable, w, L, rMid, nSteps, nHits, nSlow, nExact, nFetch);
}

// one pixel by one thread
__device__ inline bool woodcockGridPixel(const DeviceGrid &g, const RenderParams &p, const float4 *lut, const float *__restrict__ logTable,
                                         const MajorantGrid &mg, int px, int py, float *L, float4 &Cout, unsigned &nSteps, unsigned &nHits)
{
  WgRay w;
  if (!wgInit(p, logTable, mg, px, py, w, L)) return false;
  const float rMid = 0.5f * (p.rInner + p.rOuter);
  unsigned nSlow = 0, nExact = 0, nFetch = 0;
#pragma unroll 1
  while (!wgStep(g, p, lut, logTable, mg, w, L, rMid, nSteps, nHits, nSlow, nExact, nFetch)) {}
  Cout = w.C;
  return true;
}

// ---------------------------------------------------------------------------------------------
// woodcockTrackingWithAccel in SPHERE_ACCEL_MODE -- the mode the reference starts in (hostCode.cu:75, :915): sdda
// (ShellAccel.h:83-199) over the (rad, lat, lon) shell accel.  As written there, the lat/lon cell planes go through three
// points at radius 0 (makePlane(0, toCartesian(0,..), toCartesian(0,..))): the plane is all-zero and evalPlane returns 0
// for every position, so tnext.y = tnext.z = 0.  A radial segment [lower, upper] therefore gets ONE callback over the
// whole segment with the majorant of its entry point's (lat, lon) cell, followed by a diagonal walk through (lat, lon)
// cells with zero-length callbacks [0, 0] -- each draws one random number when its majorant is positive -- until the
// lat or lon index reaches the exit point's.  Every comparison and update of that loop is restated; cells come from
// toSpherical (ICONGrid.h:36-42) through glibc's asinf / atan2f operation sequences (fdlibm_trig.cuh).
// ---------------------------------------------------------------------------------------------
struct WsRay : TrackRay {
  float rl[2], ru[2];          // the radial segments (box1f ranges[2])
  float tn[3], tcur;           // tnext, t of sdda's loop
  int cell[3], stop[3];
  int stepBits;                // bit k: step[k] == +1
  int seg;
  bool started;                // inside a segment's loop (else: set the next segment up)
  bool restage;                // random numbers were drawn outside a free flight since the log entries were staged
};

__device__ __forceinline__ int sphCoord(float v, float lo, float hi, int dim)     // projectToSphericalGrid, ShellAccel.h:57-69 (unbounded)
{
  return (int)__fmul_rn(__fdiv_rn(__fsub_rn(v, lo), __fsub_rn(hi, lo)), (float)(dim - 1));
}

__device__ __forceinline__ int wrapCoord(int c, int d)                            // normalizeGridCoord, ShellAccel.h:72-81
{
  c %= d;
  return c < 0 ? c + d : c;
}

__device__ __forceinline__ bool wsInit(const RenderParams &p, const float *__restrict__ logTable, const MajorantGrid &sg, int px, int py, WsRay &w, float *L)
{
  generateRayDir(p, px, py, w.rnd, true, w.d[0], w.d[1], w.d[2]);
  float rt0, rt1;
  {
    float lo3[3], hi3[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const float l = __fdiv_rn(__fsub_rn(p.volLo[k], p.org[k]), w.d[k]), h = __fdiv_rn(__fsub_rn(p.volHi[k], p.org[k]), w.d[k]);
      lo3[k] = fminf(l, h); hi3[k] = fmaxf(l, h);
    }
    rt0 = fmaxf(0.f, fmaxf(fmaxf(lo3[0], lo3[1]), lo3[2]));
    rt1 = fminf(1e10f, fminf(fminf(hi3[0], hi3[1]), hi3[2]));
  }
  if (!(rt0 < rt1)) return false;
  w.C = make_float4(0.f, 0.f, 0.f, 0.f);
  resetState(w.s);
  w.s.rec = -2;
  w.rl[0] = w.rl[1] = INFINITY; w.ru[0] = w.ru[1] = -INFINITY;
  float t1 = 0.f, t2 = 0.f, t3 = 0.f, t4 = 0.f;
  const bool s1 = intersectSphere(p.org[0], p.org[1], p.org[2], w.d[0], w.d[1], w.d[2], sg.hi[0], t1, t4);
  const bool s2 = intersectSphere(p.org[0], p.org[1], p.org[2], w.d[0], w.d[1], w.d[2], sg.lo[0], t2, t3);
  if ((s1 || s2) && !(t4 < rt0)) {
    if (s1 && !s2) { w.rl[0] = t1; w.ru[0] = t4; }
    else if (rt0 < t2) { w.rl[0] = t1; w.ru[0] = t2; w.rl[1] = t3; w.ru[1] = t4; }
    else { w.rl[0] = t3; w.ru[0] = t4; }
  }
  w.seg = 0; w.started = false; w.restage = false; w.stepBits = 0;
  w.phase = 0; w.drew = false;
  wgStageLog(logTable, w.rnd.state, L);
  w.c0 = w.c1 = w.majorant = w.t = w.tcur = 0.f; w.rate = 1.f;
  w.x = w.y = w.z = w.r = 0.f;
#pragma unroll
  for (int k = 0; k < 3; ++k) { w.tn[k] = 0.f; w.cell[k] = 0; w.stop[k] = 0; }
  return true;
}

__device__ __forceinline__ bool wsStep(const DeviceGrid &g, const RenderParams &p, const float4 *lut, const float *__restrict__ logTable,
                                       const MajorantGrid &sg, WsRay &w, float *L, float rMid, unsigned &nSteps, unsigned &nHits,
                                       unsigned &nSlow, unsigned &nExact, unsigned &nFetch)
{
  if (w.phase == 0) {
#pragma unroll 1
    for (;;) {
      if (!w.started) {
        const float lower = w.seg == 0 ? w.rl[0] : w.rl[1], upper = w.seg == 0 ? w.ru[0] : w.ru[1];
        if (w.seg >= 2 || upper < lower) return true;            // ranges[i].empty(): the traversal is over
        const float eps = __fmul_rn(sg.lo[0], 1e-6f);
        const float ta = __fadd_rn(lower, eps), tb = __fsub_rn(upper, eps);
        float sp[2][3];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const float te = e == 0 ? ta : tb;
          const float x = __fadd_rn(p.org[0], __fmul_rn(w.d[0], te)), y = __fadd_rn(p.org[1], __fmul_rn(w.d[1], te)),
                      z = __fadd_rn(p.org[2], __fmul_rn(w.d[2], te));
          const float r = __fsqrt_rn(dot3(x, y, z, x, y, z));
          sp[e][0] = r; sp[e][1] = fdlibm::asinF(__fdiv_rn(z, r)); sp[e][2] = fdlibm::atan2F(y, x);
        }
        w.stepBits = 0;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          w.cell[k] = sphCoord(sp[0][k], sg.lo[k], sg.hi[k], sg.dims[k]);
          const bool up = sp[0][k] < sp[1][k];
          if (up) w.stepBits |= 1 << k;
          w.stop[k] = sphCoord(sp[1][k], sg.lo[k], sg.hi[k], sg.dims[k]) + (up ? 1 : -1);
        }
        w.tn[0] = upper; w.tn[1] = 0.f; w.tn[2] = 0.f;
        w.tcur = lower;
        w.started = true;
      } else {
        const float tc = fminf(fminf(w.tn[0], w.tn[1]), w.tn[2]);
        bool brk = false;
#pragma unroll
        for (int k = 0; k < 3; ++k)
          if (!brk && w.tn[k] == tc) {
            w.cell[k] += ((w.stepBits >> k) & 1) ? 1 : -1;
            if (w.cell[k] == w.stop[k]) brk = true;
            else if (k > 0) w.tn[k] = 0.f;                                             // evalPlane of the all-zero plane
          }
        if (brk) { w.started = false; ++w.seg; continue; }
        w.tcur = tc;
      }
      float t1 = 3.402823466e+38f;
#pragma unroll
      for (int k = 0; k < 3; ++k)
        if (w.tn[k] < t1 && w.tn[k] >= w.tcur) t1 = w.tn[k];
      const int leaf = (wrapCoord(w.cell[2], sg.dims[2]) * sg.dims[1] + wrapCoord(w.cell[1], sg.dims[1])) * sg.dims[0] + wrapCoord(w.cell[0], sg.dims[0]);
      w.majorant = __ldg(sg.maxOpacities + leaf);
      if (w.majorant > 0.f) {
        w.rate = __fdiv_rn(w.majorant, p.unitDistance);
        // a zero-length callback [0,0]: one free flight 0 - logf(1-u)/rate, which leaves the interval unless u's 24 bits
        // are all zero (logf(1) = 0) -- the draw is taken here and the table read, flight and compare are skipped
        if (w.tcur == 0.f && t1 == 0.f && __fdiv_rn(-5.9604645e-8f, w.rate) != 0.f) {
          const unsigned s1 = 1664525u * w.rnd.state + 1013904223u;
          if ((s1 & 0x00FFFFFFu) != 0u) { w.rnd.state = s1; w.restage = true; continue; }
        }
        w.c0 = w.tcur; w.c1 = t1;
        w.t = w.c0;
        w.phase = 1;
        if (w.restage) { wgStageLog(logTable, w.rnd.state, L); w.drew = false; w.restage = false; }
        break;
      }
    }
  }
  return trackStep(g, p, lut, logTable, w, L, rMid, nSteps, nHits, nSlow, nExact, nFetch);
}

__device__ inline bool woodcockSpherePixel(const DeviceGrid &g, const RenderParams &p, const float4 *lut, const float *__restrict__ logTable,
                                           const MajorantGrid &sg, int px, int py, float *L, float4 &Cout, unsigned &nSteps, unsigned &nHits)
{
  WsRay w;
  if (!wsInit(p, logTable, sg, px, py, w, L)) return false;
  const float rMid = 0.5f * (p.rInner + p.rOuter);
  unsigned nSlow = 0, nExact = 0, nFetch = 0;
#pragma unroll 1
  while (!wsStep(g, p, lut, logTable, sg, w, L, rMid, nSteps, nHits, nSlow, nExact, nFetch)) {}
  Cout = w.C;
  return true;
}

}  // namespace
}  // namespace icb
