// sm_100a kernels of the icon_rt hot path: cell location (walk + exact candidate query),
// prism sampling, transfer-function lookup and front-to-back compositing.
//
// Parity rules: every expression that feeds a located record id or a sample
// position uses __fmul_rn/__fadd_rn/__fsub_rn (never contracted into FMA) in the reference's
// operand order; the file is additionally compiled with -fmad=false, IEEE sqrt and division
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
#include "kernels.h"
#include "locator.cuh"

#include <algorithm>
#include <cfloat>
#include <cuda_runtime.h>

namespace icb {

namespace {

constexpr int kSlotStride = kLineStride + 4;   // floats per thread line slot (272 bytes)
__host__ __device__ inline size_t nzInts(int tfSize) { return (size_t)((tfSize + 1 + 3) & ~3); }     // prism sampler: opacity prefix counts, 16-byte padded
inline size_t renderSmemBytes(int tfSize, bool prism = false)
{
  return (size_t)tfSize * sizeof(float4) + (prism ? nzInts(tfSize) * sizeof(int) : 0) + (size_t)kRenderThreads * kSlotStride * sizeof(float);
}

// Prism sampler, empty-space skipping: can any point of the staged prism classify to a non-zero opacity?  The
// value is a convex combination of the six vertex scalars (weights >= 0 for a well-formed record: the located
// point passes all three plane tests and every kappa is negative), postClassify is monotone in the value, and
// nz[] counts the transfer-function entries with non-zero opacity (or non-finite colour) below each index --
// one index of slack either side covers the rounding.  The per-macrocell value ranges and max-opacity tables of
// the reference (hostCode.cu:205-336, 362-434) reduced to the located prism itself.
// no value in [lo, hi] classifies to a non-zero opacity: postClassify's table index is monotone in the value, nz[] counts
// the non-clear entries below each index, two indices of slack either side cover the rounding
__device__ __forceinline__ bool rangeTransparent(float lo, float hi, const int *nz, float tfLo, float tfHi, int tfSize)
{
  const float den = __fsub_rn(tfHi, tfLo), size = (float)tfSize;
  const float va = __fmul_rn(__fdiv_rn(__fsub_rn(lo, tfLo), den), size);
  const float vb = __fmul_rn(__fdiv_rn(__fsub_rn(hi, tfLo), den), size);
  if (!(va == va) || !(vb == vb)) return false;
  const float a = fminf(fmaxf(fminf(va, vb), -2.f), size + 2.f), b = fminf(fmaxf(fmaxf(va, vb), -2.f), size + 2.f);
  const int i0 = min(max((int)a - 2, 0), tfSize - 1), i1 = min((int)b + 2, tfSize - 1);   // (values past either end classify to the end entries)
  return nz[i1 + 1] - nz[i0] == 0;
}

__device__ __forceinline__ bool prismTransparent(const float *L, const int *nz, const RenderParams &p, float thr)
{
  if (!(thr < INFINITY)) return false;
  if (!(L[kSlotKappa] < 0.f && L[kSlotKappa + 1] < 0.f && L[kSlotKappa + 2] < 0.f)) return false;
  float lo = L[kSlotScal], hi = L[kSlotScal];
  bool fin = fabsf(lo) < INFINITY;
#pragma unroll
  for (int k = 1; k < 6; ++k) {
    const float v = L[kSlotScal + k];
    fin = fin && fabsf(v) < INFINITY;
    lo = fminf(lo, v); hi = fmaxf(hi, v);
  }
  if (!fin) return false;
  return rangeTransparent(lo, hi, nz, p.tfLo, p.tfHi, p.tfSize);
}

#ifndef ICB_PRISM_BURST
#define ICB_PRISM_BURST 4
#endif
constexpr int kPrismBurst = ICB_PRISM_BURST;   // prism sampler: samples of one prism evaluated per loop iteration

// ---------------------------------------------------------------------------------------------
// render: persistent warps, rays pulled lane by lane from 16x16 screen tiles
// ---------------------------------------------------------------------------------------------
#ifndef ICB_RENDER_MIN_BLOCKS
#define ICB_RENDER_MIN_BLOCKS 2
#endif
// PIECES = false: the pixel pass (every pixel of the frame's tiles; a lane = a ray).
// PIECES = true:  the chunk pass that runs in front of it for the frame's heaviest tiles (sp.nHeavy leading entries of
//                 the tile order): a lane = ONE CHUNK of a ray, composited from zero into a ChunkResult.  The pixel pass
//                 then assembles such a ray from its chunk results instead of marching it (and re-marches, with the true
//                 prefix, only the chunk in which the ray terminates) -- the frame's critical path is one chunk long
//                 instead of one limb ray long, and the result is bit-identical to a lane marching the whole ray because
//                 the compositing arithmetic is chunked in both passes (CHUNKED; locator.cuh compositeSample).
// CHUNKED = false: the plain A.3 recurrence, no chunk state -- the default (split rays off).
template <bool TRACE, bool PRISM, bool PIECES, bool CHUNKED>
#ifdef ICB_RENDER_MAXNREG
__global__ void __maxnreg__(ICB_RENDER_MAXNREG)      // (occupancy experiments: an explicit register cap)
#else
__global__ void __launch_bounds__(kRenderThreads, ICB_RENDER_MIN_BLOCKS)
#endif
renderKernel(const DeviceGrid g, const RenderParams p, const float4 *__restrict__ lutG, RenderTargets out,
             const int *__restrict__ tileList, const int *__restrict__ order, int *tileCounter, int refillMin, const SplitArgs sp)
{
  static_assert(CHUNKED || !PIECES, "split rays need the chunked compositing arithmetic");
  const int nHeavy = (!TRACE && CHUNKED && sp.nHeavy) ? __ldg(sp.nHeavy) : 0;
  if (PIECES && nHeavy == 0) return;
  // shared memory: the transfer function, then one 272-byte line slot per thread (heights | values of the
  // record the thread's ray is in; 68-float stride keeps the 16-byte cp.async alignment and spreads banks)
  extern __shared__ float4 smem[];
  float4 *lut = smem;
  for (int i = threadIdx.x; i < p.tfSize; i += blockDim.x) lut[i] = lutG[i];
  int *nz = reinterpret_cast<int *>(smem + p.tfSize);
  float *L = reinterpret_cast<float *>(smem + p.tfSize) + (PRISM ? (int)nzInts(p.tfSize) : 0) + threadIdx.x * kSlotStride;
  __syncthreads();
  if (PRISM) {
    if (threadIdx.x == 0) {
      int run = 0;
      const bool scaleOk = fabsf(p.opacityScale) < INFINITY;
      for (int i = 0; i < p.tfSize; ++i) {
        nz[i] = run;
        const float4 c = lut[i];
        const bool clear = scaleOk && c.w == 0.f && fabsf(c.x) < INFINITY && fabsf(c.y) < INFINITY && fabsf(c.z) < INFINITY;
        run += clear ? 0 : 1;
      }
      nz[p.tfSize] = run;
    }
    __syncthreads();
  }

  const int lane = threadIdx.x & 31;
  const float dt = p.dt;
  const float rMid = 0.5f * (p.rInner + p.rOuter);
  const float ambR = p.ambient[0], ambG = p.ambient[1], ambB = p.ambient[2];

  // per-lane ray
  bool active = false;
  bool pending = false;        // a finished ray whose pixel is written when the warp next refills (lanes together)
  size_t outIdx = 0;
  int tileOfRay = 0, pixOfRay = 0;
  float dx = 0.f, dy = 0.f, dz = 0.f;
  float segLo = 0.f, segHi = 0.f, seg2Lo = 0.f, seg2Hi = 0.f;
  int segLeft = 0, si = 0;
  float4 C = make_float4(0.f, 0.f, 0.f, 0.f);      // chunks merged so far (the pixel's colour once the ray has ended)
  float4 Lc = make_float4(0.f, 0.f, 0.f, 0.f);     // the current chunk's own composite
  int curKey = -1;
  int endIdx = 0x7fffffff;                          // PIECES: first sample index behind this lane's chunk
  int resIdx = 0;                                   // PIECES: where the chunk's result goes
  WalkState s;
  resetState(s);
  int pend = 0, curId = -1;           // prism sampler: skipped samples of the current prism still to be evaluated
  int cKey = -1, cRec = -1;          // prism sampler: the record whose tail / the prism whose scalars the slot holds
  bool transparent = false;
  const bool skipOn = p.maxSkip > 0 && g.nGlobal == 0;
  float colR = 0.f, colG = 0.f, colB = 0.f, colA = 0.f;
  unsigned raySteps = 0, rayHits = 0;
  unsigned rayIters = 0;              // loop iterations the ray has occupied this lane: the tile's cost (a time, not a sample count)
  unsigned long long hash = 0;
  // per-thread totals
  unsigned long long totSteps = 0, totHits = 0;
  unsigned nSlow = 0, nExact = 0, nRays = 0, nFetch = 0;
  unsigned dWarp = 0, dLane = 0, dFast = 0, dLayer = 0, dRetry = 0, dNoSkip = 0;     // TRACE only
  // warp-uniform work cursor: a unit is one 8x4 block (32 Morton-consecutive pixels) of a 16x16 tile; units are
  // handed out heaviest tile first (order[] comes from the previous frame's per-tile sample counts), so that
  // the frame's tail is one ray long rather than one tile long
  int chunkTile = -1, chunkListIdx = 0, chunkNext = 0, chunkEnd = 0;
  int unitOrderIdx = 0, pieceK = 0;
  bool noMore = false;
  const int numUnits = PIECES ? nHeavy * kMaxChunks * kUnitsPerTile : p.numTiles * kUnitsPerTile;

  for (;;) {
    const unsigned need = __ballot_sync(FULL, !active);
    if (need != 0u && !noMore && (__popc(need) >= refillMin || need == FULL)) {
      if (pending) { writePixel(p, C, outIdx, out.fb, out.accum); pending = false; }
      for (;;) {
        if (chunkNext >= chunkEnd) {
          int u = 0;
          if (lane == 0) u = atomicAdd(tileCounter, 1);
          u = __shfl_sync(FULL, u, 0);
          if (u >= numUnits) { noMore = true; break; }
          // chunk pass: chunk 0 of every heavy tile's rays first, then chunk 1, ... -- a chunk that saturates cancels the
          // later chunks of its ray, which by then have mostly not been started
          unitOrderIdx = PIECES ? (u / kUnitsPerTile) % nHeavy : u / kUnitsPerTile;
          pieceK = PIECES ? u / (kUnitsPerTile * nHeavy) : 0;
          const int ti = __ldg(order + unitOrderIdx);
          chunkListIdx = ti;
          chunkTile = tileList ? __ldg(tileList + ti) : ti;
          chunkNext = (u % kUnitsPerTile) * kUnitPixels;
          chunkEnd = chunkNext + kUnitPixels;
        }
        const unsigned needNow = __ballot_sync(FULL, !active);
        if (needNow == 0u) break;
        const int rank = __popc(needNow & ((1u << lane) - 1u));
        const int take = min(__popc(needNow), chunkEnd - chunkNext);
        if (!active && rank < take) {
          const int slot = chunkNext + rank;
          const int lx = (int)compact1by1((unsigned)slot), ly = (int)compact1by1((unsigned)slot >> 1);
          const int px = (chunkTile % p.tilesX) * kTile + lx, py = (chunkTile / p.tilesX) * kTile + ly;
          if (px < p.width && py < p.height) {
            RaySetup rs;
            setupRay(p, px, py, rs);
            pixOfRay = px + p.width * py;
            outIdx = out.compact ? (size_t)chunkListIdx * kTilePixels + (size_t)(ly * kTile + lx) : (size_t)pixOfRay;
            tileOfRay = chunkTile;
            C = make_float4(0.f, 0.f, 0.f, 0.f);
            Lc = make_float4(0.f, 0.f, 0.f, 0.f);
            curKey = -1;
            raySteps = rayHits = 0;
            rayIters = 1;
            if (TRACE) hash = 1469598103934665603ull;
            // split rays: heavy tile, two or more chunks (and no more than kMaxChunks)
            int nC0 = 0, nC1 = 0;
            bool split = false;
            if (!TRACE && unitOrderIdx < nHeavy && rs.nseg > 0) {
              nC0 = segmentChunks(rs.seg[0][0], rs.seg[0][1], dt, p.chunkLog2);
              nC1 = rs.nseg > 1 ? segmentChunks(rs.seg[1][0], rs.seg[1][1], dt, p.chunkLog2) : 0;
              split = nC0 >= 0 && nC1 >= 0 && nC0 + nC1 >= 2 && nC0 + nC1 <= kMaxChunks;
            }
            const int rayIdx = unitOrderIdx * kTilePixels + ly * kTile + lx;
            if (PIECES) {
              // this lane's chunk of the ray, if the ray has one and no earlier chunk has saturated already
              if (split && pieceK < nC0 + nC1 && __ldcg(sp.cancelFrom + rayIdx) > pieceK) {
                const int sg = pieceK < nC0 ? 0 : 1, c = pieceK - (sg ? nC0 : 0);
                dx = rs.dx; dy = rs.dy; dz = rs.dz;
                segLo = rs.seg[sg][0]; segHi = rs.seg[sg][1];
                segLeft = 0;
                si = c << p.chunkLog2;
                endIdx = (c + 1) << p.chunkLog2;
                resIdx = rayIdx * kMaxChunks + pieceK;
                cKey = -1;
                pend = 0;
                s.rec = -2;
                active = true;
              }
            } else if (rs.nseg == 0) {
              writePixel(p, C, outIdx, out.fb, out.accum);
              if (out.depth && p.accumID == 0 && !out.compact) out.depth[outIdx] = 0.f;
              if (TRACE && out.perPixel) {
                int4 pp = make_int4(0, 0, (int)(unsigned)(hash & 0xffffffffu), (int)(unsigned)(hash >> 32));
                reinterpret_cast<int4 *>(out.perPixel)[pixOfRay] = pp;
              }
            } else {
              dx = rs.dx; dy = rs.dy; dz = rs.dz;
              segLo = rs.seg[0][0]; segHi = rs.seg[0][1];
              seg2Lo = rs.seg[1][0]; seg2Hi = rs.seg[1][1];
              segLeft = rs.nseg - 1;
              si = 0;
              cKey = -1;
              pend = 0;
              s.rec = -2;             // seeded in the march loop, where all lanes without a walk history do it together
              ++nRays;
              bool whole = false;     // every chunk merged, nothing left to march
              if (split) {
                // assemble the ray from its chunk results (the chunk pass ran in front of this kernel): merge in order until
                // the chunk in which the ray terminates; that one is marched again below with the true prefix, so that the
                // ray stops at exactly the sample a single lane would have stopped at
                const ChunkResult *R = sp.results + (size_t)rayIdx * kMaxChunks;
                int kstar = -1;
                for (int k = 0; k < nC0 + nC1; ++k) {
                  const float4 c4 = __ldcg(reinterpret_cast<const float4 *>(R + k));
                  const int4 m4 = __ldcg(reinterpret_cast<const int4 *>(R + k) + 1);
                  const float G = __fadd_rn(C.w, __fmul_rn(__fsub_rn(1.f, C.w), c4.w));
                  if ((m4.z & 1) || G >= p.termAlpha) { kstar = k; break; }
                  Lc = c4;
                  mergeChunk(C, Lc);
                  raySteps += (unsigned)m4.x; rayHits += (unsigned)m4.y;
                }
                if (kstar < 0) whole = true;
                else {
                  const int sg = kstar < nC0 ? 0 : 1, c = kstar - (sg ? nC0 : 0);
                  if (sg == 1) { segLo = seg2Lo; segHi = seg2Hi; segLeft = 0; }
                  si = c << p.chunkLog2;
                }
              }
              if (whole) {
                writePixel(p, C, outIdx, out.fb, out.accum);
                if (out.depth && p.accumID == 0 && !out.compact) out.depth[outIdx] = 0.f;
                if (out.tileCost) atomicAdd(out.tileCost + tileOfRay, 1);
                totSteps += raySteps; totHits += rayHits;
              } else active = true;
            }
          }
        }
        chunkNext += take;
      }
    }
    if (__ballot_sync(FULL, active) == 0u) {
      if (noMore) break;
      continue;
    }

    if (TRACE) { if (lane == 0) ++dWarp; if (active) ++dLane; }
    // chunk pass: a lane whose chunk can no longer contribute (an earlier chunk of the ray has saturated) drops it
    if (PIECES && active && (rayIters & 15u) == 0u && __ldcg(sp.cancelFrom + resIdx / kMaxChunks) <= resIdx % kMaxChunks) {
      if (out.tileCost) atomicAdd(out.tileCost + tileOfRay, (int)rayIters);
      active = false;
    }
    if (active) {
      ++rayIters;
      bool finished = false, truncated = false;
      // prism sampler: a located sample and the samples skipped behind it are evaluated ONE per loop iteration at a
      // single site below, so that the lanes of a warp evaluate (value, classification, expf, compositing) together
      // whatever their skip lengths; a lane with `pend` skipped samples left does no locating this iteration
      bool doValue = false;
      float vx = 0.f, vy = 0.f, vz = 0.f;
      if (PRISM && pend > 0) {
        const float tq = sampleT(segLo, si, dt);
        vx = __fadd_rn(p.org[0], __fmul_rn(dx, tq));
        vy = __fadd_rn(p.org[1], __fmul_rn(dy, tq));
        vz = __fadd_rn(p.org[2], __fmul_rn(dz, tq));
        ++si; ++raySteps; ++rayHits; --pend;
        if (TRACE) hash = fnvStep(hash, curId);
        doValue = true;
      } else {
      const float t = sampleT(segLo, si, dt);
      if (t >= segHi || (PIECES && si >= endIdx)) {
        if (segLeft > 0) { segLo = seg2Lo; segHi = seg2Hi; segLeft = 0; si = 0; s.rec = -2; }   // far side of the inner sphere: no walking there
        else finished = true;
      } else {
        const float x = __fadd_rn(p.org[0], __fmul_rn(dx, t));
        const float y = __fadd_rn(p.org[1], __fmul_rn(dy, t));
        const float z = __fadd_rn(p.org[2], __fmul_rn(dz, t));
        const float r = __fsqrt_rn(dot3(x, y, z, x, y, z));
        // one action of the locate state machine; an undecided lane comes back with the same sample next
        // iteration, together with whatever the other lanes of the warp have to do then
        bool hit = false, decided;
        Evals e = noEvals();
        if (s.rec == -2) {
          if (r == r) seedState(g, rMid, x, y, z, r, s, L);
          else s.rec = -1;
          if (PRISM) { cRec = -1; cKey = -1; }
        }
        if (r == r) {
          const unsigned slow0 = nSlow, fetch0 = nFetch, exact0 = nExact;
          decided = locateAction(g, x, y, z, r, s, L, e, hit, nSlow, nExact, nFetch);
          if (TRACE) {
            if (!decided) ++dRetry;
            else if (nSlow == slow0 && nExact == exact0) ++dFast;
            else if (nFetch == fetch0 && nExact == exact0) ++dLayer;
          }
          if (decided && g.nGlobal != 0) hit = applyGlobals(g, x, y, z, r, s, L, hit);
          // (any fetch rewrites the slot's value half, where the prism sampler keeps its staged prism)
          if (PRISM && (nFetch != fetch0 || nExact != exact0 || g.nGlobal != 0)) { cRec = -1; cKey = -1; }
        } else { hit = nanPosition(g, s, L); decided = true; if (PRISM) { cRec = -1; cKey = -1; } }
        if (decided) {
          ++si; ++raySteps;
          int id = -1;
          if (TRACE) { id = hit ? s.orig : -1; hash = fnvStep(hash, id); }
          // this sample, then as many following ones as are provably the same prism (skip-ahead)
          int reps = 1;
          if (skipOn && (hit || s.layer < 0))
            reps += clampSkip(skipCount(s, e, p, dx, dy, dz, x, y, z, r, dt), si, segLo, segHi, t, dt);
          if (PIECES) reps = min(reps, endIdx - (si - 1));          // (a chunk lane stays inside its chunk)
          if (TRACE && reps == 1) ++dNoSkip;
          if (hit) {
            ++rayHits;
            const int key = s.rec * kHeightStride + s.layer;
            if (PRISM) {
              if (key != cKey) {
                cKey = key;
                // the record's transparency mask came with its lines: a clear layer needs no vertex scalars at all
                if (g.prismMask && (p.flags & 4u) == 0u && ((__float_as_uint(L[kSlotMask]) >> s.layer) & 1u)) transparent = true;
                else {
                  if (s.rec != cRec) { prismFetchTail(g, s.rec, L); cRec = s.rec; }
                  prismFetchScalars(g, s.layer, L);
                  transparent = (p.flags & 4u) == 0u && prismTransparent(L, nz, p, s.thr);
                }
              }
              if (transparent) {
                if (TRACE) for (int q = 1; q < reps; ++q) hash = fnvStep(hash, id);
                si += reps - 1; raySteps += reps - 1; rayHits += reps - 1;
              } else {
                vx = x; vy = y; vz = z;
                doValue = true;
                pend = reps - 1;
                if (TRACE) curId = id;
              }
            } else {
            if (key != cKey) {
              cKey = key;
              const float v = L[kHeightStride + s.layer];
              const float4 sm = postClassify(lut, p, v);
              colA = sampleOpacity(sm.w, dt, p.unitDistance);
              colR = __fmul_rn(__fmul_rn(sm.x, ambR), p.ambientRadiance);
              colG = __fmul_rn(__fmul_rn(sm.y, ambG), p.ambientRadiance);
              colB = __fmul_rn(__fmul_rn(sm.z, ambB), p.ambientRadiance);
              // a sample that adds exactly nothing: w = (1-C.w)*0 = 0, C += 0*col  (col finite)
              transparent = colA == 0.f && fabsf(colR) < INFINITY && fabsf(colG) < INFINITY && fabsf(colB) < INFINITY;
            }
            if (transparent) {
              if (TRACE) for (int q = 1; q < reps; ++q) hash = fnvStep(hash, id);
              si += reps - 1; raySteps += reps - 1; rayHits += reps - 1;
            } else {
#pragma unroll 1
              for (int q = 0; q < reps; ++q) {
                if (q > 0) { ++si; ++raySteps; ++rayHits; if (TRACE) hash = fnvStep(hash, id); }
                if (compositeStep<CHUNKED>(C, Lc, curKey, chunkKey(segLeft, si - 1, p.chunkLog2), colA, colR, colG, colB, p.termAlpha)) { finished = true; truncated = true; break; }
              }
            }
            }
          } else {
            if (TRACE) for (int q = 1; q < reps; ++q) hash = fnvStep(hash, -1);
            si += reps - 1; raySteps += reps - 1;
          }
        }
      }
      }   // (pend == 0)
      if (PRISM && doValue) {
        // this sample, then up to kPrismBurst-1 of the samples skipped behind it (a short burst keeps the number of
        // loop iterations a long opaque run costs down; lanes without pending samples idle through it)
#pragma unroll 1
        for (int u = 0;; ++u) {
          float v;
          if (prismValue(s.p0, s.p1, s.p2, L, vx, vy, vz, v)) {
            const float4 sm = postClassify(lut, p, v);
            const float a = sampleOpacity(sm.w, dt, p.unitDistance);
            if (compositeStep<CHUNKED>(C, Lc, curKey, chunkKey(segLeft, si - 1, p.chunkLog2), a, __fmul_rn(__fmul_rn(sm.x, ambR), p.ambientRadiance),
                                __fmul_rn(__fmul_rn(sm.y, ambG), p.ambientRadiance), __fmul_rn(__fmul_rn(sm.z, ambB), p.ambientRadiance), p.termAlpha)) {
              finished = true; truncated = true; pend = 0;
            }
          }
          if (pend <= 0 || u + 1 >= kPrismBurst) break;
          const float tq = sampleT(segLo, si, dt);
          vx = __fadd_rn(p.org[0], __fmul_rn(dx, tq));
          vy = __fadd_rn(p.org[1], __fmul_rn(dy, tq));
          vz = __fadd_rn(p.org[2], __fmul_rn(dz, tq));
          ++si; ++raySteps; ++rayHits; --pend;
          if (TRACE) hash = fnvStep(hash, curId);
        }
      }
      if (finished) {
        if (PIECES) {
          // the chunk's own composite (nothing merged in front of it: C stays zero), for the lane that assembles the pixel
          mergeChunk(C, Lc);
          ChunkResult *R = sp.results + resIdx;
          __stcg(reinterpret_cast<float4 *>(R), C);
          __stcg(reinterpret_cast<int4 *>(R) + 1, make_int4((int)raySteps, (int)rayHits, truncated ? 1 : 0, 0));
          if (truncated) atomicMin(sp.cancelFrom + resIdx / kMaxChunks, resIdx % kMaxChunks + 1);
          if (out.tileCost) atomicAdd(out.tileCost + tileOfRay, (int)rayIters);
        } else {
          mergeLast<CHUNKED>(C, Lc);
          pending = true;
          if (out.depth && p.accumID == 0 && !out.compact) out.depth[outIdx] = 0.f;
          if (out.tileCost) atomicAdd(out.tileCost + tileOfRay, (int)rayIters);
          if (TRACE && out.perPixel) {
            int4 pp = make_int4((int)raySteps, (int)rayHits, (int)(unsigned)(hash & 0xffffffffu), (int)(unsigned)(hash >> 32));
            reinterpret_cast<int4 *>(out.perPixel)[pixOfRay] = pp;
          }
          totSteps += raySteps; totHits += rayHits;
        }
        active = false;
      }
    }
  }

  if (pending) writePixel(p, C, outIdx, out.fb, out.accum);

  // counters: warp reduce, one atomic per warp
  unsigned long long v[5] = {totSteps, totHits, (unsigned long long)nSlow, (unsigned long long)nExact, (unsigned long long)nRays};
#pragma unroll
  for (int k = 0; k < 5; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_down_sync(FULL, v[k], o);
  }
  if (lane == 0 && out.counters) {
    if (!PIECES) {         // (a chunk's samples are counted by the pixel pass, when it merges them)
      atomicAdd(&out.counters->steps, v[0]);
      atomicAdd(&out.counters->hits, v[1]);
      atomicAdd(&out.counters->totalSteps, v[0]);
      atomicAdd(&out.counters->totalHits, v[1]);
      atomicAdd(&out.counters->rays, v[4]);
    }
    atomicAdd(&out.counters->slow, v[2]);
    atomicAdd(&out.counters->exact, v[3]);
  }
  if (TRACE && out.counters) {
    unsigned long long d[7] = {dWarp, dLane, dFast, dLayer, nFetch, dRetry, dNoSkip};
#pragma unroll
    for (int k = 0; k < 7; ++k) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) d[k] += __shfl_down_sync(FULL, d[k], o);
      if (lane == 0) atomicAdd(&out.counters->dbg[k], d[k]);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// the reference's own integrator (ray-gen woodcockTrackingAE, deviceCode.cu:239-275): a thread = a pixel,
// a block = a 16x16 screen tile in Morton order
// ---------------------------------------------------------------------------------------------
template <int THREADS, int MIN_BLOCKS>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS)
woodcockKernel(const DeviceGrid g, const RenderParams p, const float4 *__restrict__ lutG, const float *__restrict__ logTable,
               RenderTargets out)
{
  constexpr int kBlocksPerTile = kTilePixels / THREADS;     // (block sizes as for woodcockGridKernel below)
  extern __shared__ float4 smem[];
  float4 *lut = smem;
  for (int i = threadIdx.x; i < p.tfSize; i += blockDim.x) lut[i] = lutG[i];
  float *L = reinterpret_cast<float *>(smem + p.tfSize) + threadIdx.x * kSlotStride;
  __syncthreads();
  const int tile = blockIdx.x / kBlocksPerTile;
  const unsigned slot = (blockIdx.x % kBlocksPerTile) * THREADS + threadIdx.x;
  const int lx = (int)compact1by1(slot), ly = (int)compact1by1(slot >> 1);
  const int px = (tile % p.tilesX) * kTile + lx, py = (tile / p.tilesX) * kTile + ly;
  unsigned nSteps = 0, nHits = 0, nRays = 0;
  if (px < p.width && py < p.height) {
    float4 C;
    if (woodcockPixel(g, p, lut, logTable, px, py, L, C, nSteps, nHits)) {
      writePixel(p, C, (size_t)px + (size_t)p.width * py, out.fb, out.accum);
      ++nRays;
    }
  }
  unsigned long long v[3] = {nSteps, nHits, nRays};
#pragma unroll
  for (int k = 0; k < 3; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_down_sync(FULL, v[k], o);
  }
  if ((threadIdx.x & 31) == 0 && out.counters) {
    atomicAdd(&out.counters->steps, v[0]);
    atomicAdd(&out.counters->hits, v[1]);
    atomicAdd(&out.counters->totalSteps, v[0]);
    atomicAdd(&out.counters->totalHits, v[1]);
    atomicAdd(&out.counters->rays, v[2]);
  }
}

// the reference's default ray-gen (woodcockTrackingWithAccel, GRID_ACCEL_MODE): same thread/tile mapping.  A thread keeps
// its pixel to the end: refilling finished lanes from other pixels (renderKernel's scheme) was measured and is SLOWER here
// (33.9 ms -> 36.8 / 46.0 / 65.8 ms per C3 frame at refill thresholds 16 / 4 / 1) -- a ray ends at its first accepted
// collision, and what the warp gains in busy lanes it loses in coherence (neighbouring pixels walk the same macrocells
// and records).  Two CTAs per SM: three (80 registers, 100 B of spills) measured 37.1 ms.
// Block size: a CTA's slot frees only when its longest ray ends, and the kernel is latency-bound at 16 warps per SM
// (108 registers), so small CTAs pay twice -- 64 threads x 10 CTAs/SM at 96 registers measured 27.0 ms per C3 frame,
// 128 x 5: 27.8, 128 x 4 (108 registers): 31.2, 256 x 2: 33.7.  Every CTA stages the transfer function, so larger tables
// take the larger blocks (launchWoodcockGrid).
template <int THREADS, int MIN_BLOCKS>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS)
woodcockGridKernel(const DeviceGrid g, const RenderParams p, const float4 *__restrict__ lutG, const float *__restrict__ logTable,
                   const MajorantGrid mg, RenderTargets out)
{
  constexpr int kWgBlocksPerTile = kTilePixels / THREADS;   // a block = THREADS Morton-consecutive pixels of a 16x16 tile
  extern __shared__ float4 smem[];
  float4 *lut = smem;
  for (int i = threadIdx.x; i < p.tfSize; i += blockDim.x) lut[i] = lutG[i];
  float *L = reinterpret_cast<float *>(smem + p.tfSize) + threadIdx.x * kSlotStride;
  __syncthreads();
  const int tile = blockIdx.x / kWgBlocksPerTile;
  const unsigned slot = (blockIdx.x % kWgBlocksPerTile) * THREADS + threadIdx.x;
  const int lx = (int)compact1by1(slot), ly = (int)compact1by1(slot >> 1);
  const int px = (tile % p.tilesX) * kTile + lx, py = (tile / p.tilesX) * kTile + ly;
  unsigned nSteps = 0, nHits = 0, nRays = 0;
  if (px < p.width && py < p.height) {
    float4 C;
    if (woodcockGridPixel(g, p, lut, logTable, mg, px, py, L, C, nSteps, nHits)) {
      writePixel(p, C, (size_t)px + (size_t)p.width * py, out.fb, out.accum);
      ++nRays;
    }
  }
  unsigned long long v[3] = {nSteps, nHits, nRays};
#pragma unroll
  for (int k = 0; k < 3; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_down_sync(FULL, v[k], o);
  }
  if ((threadIdx.x & 31) == 0 && out.counters) {
    atomicAdd(&out.counters->steps, v[0]);
    atomicAdd(&out.counters->hits, v[1]);
    atomicAdd(&out.counters->totalSteps, v[0]);
    atomicAdd(&out.counters->totalHits, v[1]);
    atomicAdd(&out.counters->rays, v[2]);
  }
}

// woodcockTrackingWithAccel in SPHERE_ACCEL_MODE (the mode the reference starts in): same thread/tile mapping and block
// sizes as the grid ray-gen; the traversal is sdda over the shell accel (locator.cuh wsStep).
template <int THREADS, int MIN_BLOCKS>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS)
woodcockSphereKernel(const DeviceGrid g, const RenderParams p, const float4 *__restrict__ lutG, const float *__restrict__ logTable,
                     const MajorantGrid sg, RenderTargets out)
{
  constexpr int kBlocksPerTile = kTilePixels / THREADS;
  extern __shared__ float4 smem[];
  float4 *lut = smem;
  for (int i = threadIdx.x; i < p.tfSize; i += blockDim.x) lut[i] = lutG[i];
  float *L = reinterpret_cast<float *>(smem + p.tfSize) + threadIdx.x * kSlotStride;
  __syncthreads();
  const int tile = blockIdx.x / kBlocksPerTile;
  const unsigned slot = (blockIdx.x % kBlocksPerTile) * THREADS + threadIdx.x;
  const int lx = (int)compact1by1(slot), ly = (int)compact1by1(slot >> 1);
  const int px = (tile % p.tilesX) * kTile + lx, py = (tile / p.tilesX) * kTile + ly;
  unsigned nSteps = 0, nHits = 0, nRays = 0;
  if (px < p.width && py < p.height) {
    float4 C;
    if (woodcockSpherePixel(g, p, lut, logTable, sg, px, py, L, C, nSteps, nHits)) {
      writePixel(p, C, (size_t)px + (size_t)p.width * py, out.fb, out.accum);
      ++nRays;
    }
  }
  unsigned long long v[3] = {nSteps, nHits, nRays};
#pragma unroll
  for (int k = 0; k < 3; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_down_sync(FULL, v[k], o);
  }
  if ((threadIdx.x & 31) == 0 && out.counters) {
    atomicAdd(&out.counters->steps, v[0]);
    atomicAdd(&out.counters->hits, v[1]);
    atomicAdd(&out.counters->totalSteps, v[0]);
    atomicAdd(&out.counters->totalHits, v[1]);
    atomicAdd(&out.counters->rays, v[2]);
  }
}

// ---------------------------------------------------------------------------------------------
// majorant grid builders (hostCode.cu:62-71 initGrid, :36-56 float atomicMin/Max, :205-255 buildGrid_ICON with
// rasterizeBox :73-91, :400-434 computeMaxOpacities).  A thread = one layer of one record; the corner
// directions come as host-computed cosf/sinf (glibc's, like the side planes) so that toCartesian is the
// reference's bit for bit: x = (r*cos(lat))*cos(lon), y = (r*cos(lat))*sin(lon), z = r*sin(lat).
// ---------------------------------------------------------------------------------------------
__global__ void initRangesKernel(float2 *ranges, size_t n)
{
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) ranges[i] = make_float2(FLT_MAX, -FLT_MAX);
}

__device__ __forceinline__ void atomicMinF(float *a, float val)
{
  int ret = __float_as_int(*a);
  while (val < __int_as_float(ret)) {
    const int old = ret;
    if ((ret = atomicCAS(reinterpret_cast<int *>(a), old, __float_as_int(val))) == old) break;
  }
}
__device__ __forceinline__ void atomicMaxF(float *a, float val)
{
  int ret = __float_as_int(*a);
  while (val > __int_as_float(ret)) {
    const int old = ret;
    if ((ret = atomicCAS(reinterpret_cast<int *>(a), old, __float_as_int(val))) == old) break;
  }
}

__global__ void __launch_bounds__(256)
gridRangesKernel(const RecBlock *__restrict__ recs, const float *__restrict__ lines, const float *__restrict__ colTrig, size_t nRec,
                 MajorantGrid mg, float2 *ranges, const float *__restrict__ vtx, int vtxStride)
{
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t rec = tid >> 5;
  const int i = (int)(tid & 31);
  if (rec >= nRec) return;
  const int n = __ldg(&recs[rec].nLayers);
  if (i >= n) return;
  const float *H = lines + rec * kLineStride;
  const float *T = colTrig + (size_t)__ldg(&recs[rec].col) * kColTrigStride;
  const float h0 = __ldg(H + i), h1 = __ldg(H + i + 1);
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  float tv[3][3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const float cl = __ldg(T + 4 * k), sl = __ldg(T + 4 * k + 1), co = __ldg(T + 4 * k + 2), so = __ldg(T + 4 * k + 3);
    const float b[3] = {__fmul_rn(__fmul_rn(h0, cl), co), __fmul_rn(__fmul_rn(h0, cl), so), __fmul_rn(h0, sl)};
    tv[k][0] = __fmul_rn(__fmul_rn(h1, cl), co); tv[k][1] = __fmul_rn(__fmul_rn(h1, cl), so); tv[k][2] = __fmul_rn(h1, sl);
#pragma unroll
    for (int q = 0; q < 3; ++q) { lo[q] = fminf(lo[q], b[q]); hi[q] = fmaxf(hi[q], b[q]); }
  }
  float bary[3];
#pragma unroll
  for (int q = 0; q < 3; ++q) bary[q] = __fdiv_rn(__fadd_rn(__fadd_rn(tv[0][q], tv[1][q]), tv[2][q]), 3.f);
  const float D = __fsub_rn(h1, __fsqrt_rn(dot3(bary[0], bary[1], bary[2], bary[0], bary[1], bary[2])));
  const float off = __fdiv_rn(D, h1);
#pragma unroll
  for (int k = 0; k < 3; ++k)
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      const float t = __fadd_rn(tv[k][q], __fmul_rn(tv[k][q], off));
      lo[q] = fminf(lo[q], t); hi[q] = fmaxf(hi[q], t);
    }
  // getValue(height[i]), getValue(height[i+1]) (ICONGrid.h:146-153 through findHeight's own probe sequence)
  float vlo, vhi;
  if (vtx) {
    // prism sampler: a sample's value is a convex combination of the prism's six vertex scalars (prismValue), which are
    // means over the neighbouring columns -- the cell values above do not bound it.  The value range of the prism is the
    // range of its six vertex scalars, the shape of the reference's wedge builder (buildGrid_UMesh, hostCode.cu:338-360:
    // min/max over the element's vertex scalars into the macrocells its box overlaps).
    vlo = INFINITY; vhi = -INFINITY;
    const int base = __ldg(&recs[rec].layerBase) + i;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const float *row = vtx + (size_t)__ldg(&recs[rec].vert[k]) * (size_t)vtxStride + base;
      const float s0 = __ldg(row), s1 = __ldg(row + 1);
      vlo = fminf(vlo, fminf(s0, s1)); vhi = fmaxf(vhi, fmaxf(s0, s1));
    }
  } else {
    const float a = __ldg(H + kHeightStride + findHeightLiteral(H, n, h0)), b = __ldg(H + kHeightStride + findHeightLiteral(H, n, h1));
    vlo = fminf(fminf(INFINITY, a), b); vhi = fmaxf(fmaxf(-INFINITY, a), b);
  }
  int l[3], u[3];
#pragma unroll
  for (int q = 0; q < 3; ++q) { l[q] = gridCoord(lo[q], mg.lo[q], mg.hi[q], mg.dims[q]); u[q] = gridCoord(hi[q], mg.lo[q], mg.hi[q], mg.dims[q]); }
  for (int z = l[2]; z <= u[2]; ++z)
    for (int y = l[1]; y <= u[1]; ++y)
      for (int x = l[0]; x <= u[0]; ++x) {
        float2 *r = ranges + ((size_t)z * mg.dims[0] * mg.dims[1] + (size_t)y * mg.dims[0] + x);
        atomicMinF(&r->x, vlo);
        atomicMaxF(&r->y, vhi);
      }
}

// buildShell_ICON (hostCode.cu:298-334): per layer, loMC = component-wise min of the three corners' cells at height[i],
// upMC = max at height[i+1] (projectToSphericalGrid, unbounded truncation -- the reference indexes with them as they are);
// every shell cell in between takes atomicMin(lower, getValue(height[i])), atomicMax(upper, getValue(height[i+1])).
// A thread = a record.  The (lat, lon) indices do not depend on the layer, and with one radial cell (the reference's
// 1 x 1024 x 1024) neither does the radial one: the layers' values are then folded first and the cells written once --
// min/max are associative, the table is the same.  (A column across the +-pi seam covers a whole row of lon cells.)
__global__ void __launch_bounds__(256)
shellRangesKernel(const RecBlock *__restrict__ recs, const float *__restrict__ lines, const float *__restrict__ colTrig, size_t nRec,
                  MajorantGrid sg, float2 *ranges, const float *__restrict__ vtx, int vtxStride)
{
  const size_t rec = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (rec >= nRec) return;
  const int n = __ldg(&recs[rec].nLayers);
  const float *H = lines + rec * kLineStride;
  const float *T = colTrig + (size_t)__ldg(&recs[rec].col) * kColTrigStride;
  int l[3] = {INT_MAX, INT_MAX, INT_MAX}, u[3] = {INT_MIN, INT_MIN, INT_MIN};
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int cy = sphCoord(__ldg(T + 12 + k), sg.lo[1], sg.hi[1], sg.dims[1]), cz = sphCoord(__ldg(T + 15 + k), sg.lo[2], sg.hi[2], sg.dims[2]);
    l[1] = min(l[1], cy); u[1] = max(u[1], cy);
    l[2] = min(l[2], cz); u[2] = max(u[2], cz);
  }
  const bool fold = sg.dims[0] == 1;
  float flo = INFINITY, fhi = -INFINITY;
  for (int i = 0; i < n; ++i) {
    const float h0 = __ldg(H + i), h1 = __ldg(H + i + 1);
    float vlo, vhi;
    if (vtx) {       // prism sampler: the range of the prism's six vertex scalars (as in gridRangesKernel)
      vlo = INFINITY; vhi = -INFINITY;
      const int base = __ldg(&recs[rec].layerBase) + i;
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const float *row = vtx + (size_t)__ldg(&recs[rec].vert[k]) * (size_t)vtxStride + base;
        const float s0 = __ldg(row), s1 = __ldg(row + 1);
        vlo = fminf(vlo, fminf(s0, s1)); vhi = fmaxf(vhi, fmaxf(s0, s1));
      }
    } else {
      vlo = __ldg(H + kHeightStride + findHeightLiteral(H, n, h0));
      vhi = __ldg(H + kHeightStride + findHeightLiteral(H, n, h1));
    }
    if (fold) {
      // (atomicMin/Max as written take `val < current` / `val > current`: NaN never enters, like fminf/fmaxf on a non-NaN start)
      flo = vlo < flo ? vlo : flo; fhi = vhi > fhi ? vhi : fhi;
      if (i + 1 < n) continue;
      vlo = flo; vhi = fhi;
      l[0] = u[0] = sphCoord(h0, sg.lo[0], sg.hi[0], 1);
    } else {
      l[0] = sphCoord(h0, sg.lo[0], sg.hi[0], sg.dims[0]); u[0] = sphCoord(h1, sg.lo[0], sg.hi[0], sg.dims[0]);
    }
    for (int z = l[2]; z <= u[2]; ++z)
      for (int y = l[1]; y <= u[1]; ++y)
        for (int x = l[0]; x <= u[0]; ++x) {
          float2 *r = ranges + ((size_t)z * sg.dims[0] * sg.dims[1] + (size_t)y * sg.dims[0] + x);
          atomicMinF(&r->x, vlo);
          atomicMaxF(&r->y, vhi);
        }
  }
}

__global__ void maxOpacitiesKernel(const float2 *__restrict__ ranges, size_t n, const float4 *__restrict__ lut, int size, float tfLo, float tfHi,
                                   float *maxOp, uint32_t *occupied)
{
  const size_t m = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  float mo = 0.f;
  if (m < n) {
    float2 r = ranges[m];
    if (!(r.y < r.x)) {
      const float den = __fsub_rn(tfHi, tfLo);
      r.x = __fdiv_rn(__fsub_rn(r.x, tfLo), den);
      r.y = __fdiv_rn(__fsub_rn(r.y, tfLo), den);
      const int lo = max(0, min((int)__fmul_rn(r.x, (float)(size - 1)), size - 1));
      const int hi = max(0, min((int)__fmul_rn(r.y, (float)(size - 1)) + 1, size - 1));
      for (int i = lo; i <= hi; ++i) mo = fmaxf(mo, __ldg(&lut[i].w));
    }
    maxOp[m] = mo;
  }
  // the DDA's "anything here?" test as one bit per macrocell (a warp = 32 consecutive macrocells = one word)
  const unsigned bits = __ballot_sync(FULL, mo > 0.f);
  if ((threadIdx.x & 31) == 0 && m < n) occupied[m >> 5] = bits;
}

// parity hook: logf(1 - k/2^24) as the ray-gens compute it against the host libm's table, all 2^24 arguments
__global__ void logfCheckKernel(const float *__restrict__ table, unsigned long long *mismatches)
{
  const unsigned k = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned bad = 0;
  if (k < (1u << 24)) {
    const unsigned want = __float_as_uint(table[k]);
    bad = (__float_as_uint(glibc::logOneMinus<true>(k)) != want ? 1u : 0u) | (__float_as_uint(glibc::logOneMinus<false>(k)) != want ? 0x10000u : 0u);
  }
  const unsigned a = __popc(__ballot_sync(FULL, bad & 1u)), b = __popc(__ballot_sync(FULL, bad >> 16));
  if ((threadIdx.x & 31) == 0 && (a | b)) { atomicAdd(mismatches, (unsigned long long)a); atomicAdd(mismatches + 1, (unsigned long long)b); }
}

// one pixel, full id sequence (parity hook); a single thread marches exactly like a render lane
__global__ void tracePixelKernel(const DeviceGrid g, const RenderParams p, const float4 *__restrict__ lut,
                                 int px, int py, int *ids, int maxIds, float4 *rgbaOut, int *nStepsOut)
{
  __shared__ __align__(16) float L[kLineStride + 4];
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  marchPixel(g, p, lut, px, py, ids, maxIds, rgbaOut, nStepsOut, L);
}

// ---------------------------------------------------------------------------------------------
// locate (parity hook for sampleVolume, deviceCode.cu:58-125)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
locateKernel(const DeviceGrid g, const float *__restrict__ pos, size_t n, int *ids, float *values, unsigned flags,
             Counters *counters)
{
  extern __shared__ float4 smem[];
  float *L = reinterpret_cast<float *>(smem) + threadIdx.x * kSlotStride;
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool walk = (flags & 1u) != 0, exactOnly = (flags & 2u) != 0;
  const size_t per = walk ? 64 : 1;
  const size_t i0 = tid * per;
  WalkState s;
  resetState(s);
  unsigned nSlow = 0, nExact = 0, nHits = 0, nSteps = 0;
  for (size_t i = i0; i < i0 + per && i < n; ++i) {
    const float x = pos[3 * i], y = pos[3 * i + 1], z = pos[3 * i + 2];
    const float r = __fsqrt_rn(dot3(x, y, z, x, y, z));
    bool hit;
    if (exactOnly) { ++nExact; hit = exactLocate(g, x, y, z, r, s, L); }
    else hit = locateStep(g, x, y, z, r, s, L, nSlow, nExact);
    ++nSteps;
    if (hit) {
      ++nHits;
      ids[i] = s.orig;
      float v = L[kHeightStride + s.layer];
      if (g.sampler != 0 && !prismSample(g, s, L, x, y, z, v)) v = INFINITY * 0.f;
      values[i] = v;
    } else {
      ids[i] = -1;
      values[i] = 0.f;
    }
  }
  if (counters) {
    unsigned v[4] = {nSteps, nHits, nSlow, nExact};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_down_sync(FULL, v[k], o);
    }
    if ((threadIdx.x & 31) == 0) {
      atomicAdd(&counters->steps, (unsigned long long)v[0]);
      atomicAdd(&counters->hits, (unsigned long long)v[1]);
      atomicAdd(&counters->slow, (unsigned long long)v[2]);
      atomicAdd(&counters->exact, (unsigned long long)v[3]);
    }
  }
}

// frame prologue: zero the tile cursor, the counters and the per-tile costs (one launch instead of three memsets)
__global__ void prepareFrameKernel(int *tileCounter, Counters *counters, int *tileCost, int nTiles, int *cancelFrom, int nCancel)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  for (int j = i; j < nCancel; j += gridDim.x * blockDim.x) cancelFrom[j] = kMaxChunks;
  if (i == 0) {
    tileCounter[0] = 0; tileCounter[1] = 0;      // work cursors of the pixel pass and of the chunk pass
    counters->steps = counters->hits = counters->slow = counters->exact = counters->rays = 0ull;
    for (int k = 0; k < 8; ++k) counters->dbg[k] = 0ull;
  }
  if (i < nTiles) tileCost[i] = 0;
}

// Work order of the next frame: positions of the tile list sorted by the previous frame's per-tile sample count,
// heaviest first (bucketed by log2; the order inside a bucket does not matter).  One block.
__global__ void __launch_bounds__(1024)
orderTilesKernel(const int *__restrict__ tileList, int numTiles, const int *__restrict__ tileCost, int *order, int minCostLog2, int maxHeavy,
                 int *nHeavy)
{
  __shared__ int count[32], cursor[32];
  if (threadIdx.x < 32) count[threadIdx.x] = 0;
  __syncthreads();
  for (int i = threadIdx.x; i < numTiles; i += blockDim.x) {
    const int tile = tileList ? tileList[i] : i;
    atomicAdd(&count[31 - __clz(max(tileCost[tile], 0) + 1)], 1);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int run = 0, heavy = 0;
    for (int b = 31; b >= 0; --b) {
      cursor[b] = run; run += count[b];
      if (b >= minCostLog2) heavy = run;
    }
    // split rays: the leading `heavy` entries of the order are the tiles whose rays go through the chunk pass (which tiles
    // of a bucket make the cut is arbitrary -- and irrelevant: splitting a ray does not change its result)
    if (nHeavy) *nHeavy = min(heavy, maxHeavy);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < numTiles; i += blockDim.x) {
    const int tile = tileList ? tileList[i] : i;
    order[atomicAdd(&cursor[31 - __clz(max(tileCost[tile], 0) + 1)], 1)] = i;
  }
}

// scalars arrive in the caller's record order (n*32); scatter to the internal order
__global__ void scatterScalarsKernel(const float4 *__restrict__ src, const int *__restrict__ origToInt, float4 *dst, size_t nRec)
{
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;   // one float4 (of 8 per record) per thread
  if (i >= nRec * 8) return;
  const size_t rec = i >> 3, q = i & 7;
  dst[(size_t)__ldg(origToInt + rec) * (kLineStride / 4) + (kHeightStride / 4) + q] = src[i];
}

// Prism sampler set-up (and again after every scalar swap): scalar of welded vertex v at layer interface j =
// mean over the (column, corner) pairs welded to v of the column's interface value I[0] = v[0],
// I[j] = (v[j-1] + v[j]) / 2, I[L] = v[L-1] (mid-level values, cf. hostCode.cu:574-575), summed in list order.
// A thread = one (vertex, interface); j runs fastest so a warp writes one row segment.
__global__ void __launch_bounds__(256)
vertexScalarsKernel(const RecBlock *__restrict__ recs, const float *__restrict__ lines, const ColBlock *__restrict__ cols,
                    const int *__restrict__ colLayers, const int *__restrict__ vtxOff, const int *__restrict__ vtxCols,
                    size_t nVert, int stride, float *__restrict__ out)
{
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= nVert * (size_t)stride) return;
  const size_t v = tid / (size_t)stride;
  out[tid] = vertexScalarAt(recs, lines, cols, colLayers, vtxOff, vtxCols, v, (int)(tid - v * (size_t)stride));
}

// Prism sampler, empty-space table (the per-macrocell value range -> max opacity tables of hostCode.cu:205-336, 362-434 at
// the granularity of a prism): bit i of mask[rec] = layer i of the record is certainly clear under the current transfer
// function -- same test as prismTransparent (well-formed record, every kappa negative, no opaque table entry in the
// index range of the six vertex scalars).  A thread = a record.  Rerun after the transfer function or the scalars change.
__global__ void __launch_bounds__(256)
prismMaskKernel(const RecBlock *__restrict__ recs, size_t nRec, const float *__restrict__ vtx, int stride, const int *__restrict__ nz,
                float tfLo, float tfHi, int tfSize, uint32_t *__restrict__ mask)
{
  const size_t rec = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (rec >= nRec) return;
  const RecBlock &r = recs[rec];
  uint32_t m = 0;
  const int nl = __ldg(&r.nLayers);
  const bool ok = __ldg(&r.thr) < INFINITY && __ldg(&r.kappa[0]) < 0.f && __ldg(&r.kappa[1]) < 0.f && __ldg(&r.kappa[2]) < 0.f;
  if (ok) {
    const float *row[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) row[k] = vtx + (size_t)__ldg(&r.vert[k]) * (size_t)stride + __ldg(&r.layerBase);
    float b0 = __ldg(row[0]), b1 = __ldg(row[1]), b2 = __ldg(row[2]);
    for (int i = 0; i < nl && i < 31; ++i) {
      const float t0 = __ldg(row[0] + i + 1), t1 = __ldg(row[1] + i + 1), t2 = __ldg(row[2] + i + 1);
      const float lo = fminf(fminf(fminf(b0, b1), fminf(b2, t0)), fminf(t1, t2));
      const float hi = fmaxf(fmaxf(fmaxf(b0, b1), fmaxf(b2, t0)), fmaxf(t1, t2));
      const bool fin = fabsf(b0) < INFINITY && fabsf(b1) < INFINITY && fabsf(b2) < INFINITY && fabsf(t0) < INFINITY && fabsf(t1) < INFINITY && fabsf(t2) < INFINITY;
      if (fin && rangeTransparent(lo, hi, nz, tfLo, tfHi, tfSize)) m |= 1u << i;
      b0 = t0; b1 = t1; b2 = t2;
    }
  }
  mask[rec] = m;
}

__global__ void scatterTilesKernel(const unsigned *__restrict__ rgba8, const float4 *__restrict__ accumT,
                                   const int *__restrict__ tileIds, int nTiles, unsigned *fb, float4 *accum,
                                   int width, int height, int tilesX)
{
  const int ti = blockIdx.x;
  if (ti >= nTiles) return;
  const int tile = tileIds[ti];
  const int lx = threadIdx.x % kTile, ly = threadIdx.x / kTile;
  const int px = (tile % tilesX) * kTile + lx, py = (tile / tilesX) * kTile + ly;
  if (px >= width || py >= height) return;
  const size_t src = (size_t)ti * kTilePixels + threadIdx.x, dst = (size_t)py * width + px;
  if (fb && rgba8) fb[dst] = rgba8[src];
  if (accum && accumT) accum[dst] = accumT[src];
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------------------------
cudaError_t launchRender(const DeviceGrid &g, const RenderParams &p, const float4 *lut, const RenderTargets &out,
                         const int *tileList, const int *order, int *tileCounter, int refillMin, int numBlocks, bool trace,
                         const SplitArgs &sp, cudaStream_t stream)
{
  const bool prism = g.sampler != 0, chunked = p.chunkLog2 > 0;
  const size_t smem = renderSmemBytes(p.tfSize, prism);
  SplitArgs none{};
#define ICB_LAUNCH_RENDER(T, P, PC, CH, CNT, SP) \
  renderKernel<T, P, PC, CH><<<numBlocks, kRenderThreads, smem, stream>>>(g, p, lut, out, tileList, order, CNT, refillMin, SP)
  if (trace) {
    if (prism) { if (chunked) ICB_LAUNCH_RENDER(true, true, false, true, tileCounter, none); else ICB_LAUNCH_RENDER(true, true, false, false, tileCounter, none); }
    else       { if (chunked) ICB_LAUNCH_RENDER(true, false, false, true, tileCounter, none); else ICB_LAUNCH_RENDER(true, false, false, false, tileCounter, none); }
  } else if (!chunked) {
    if (prism) ICB_LAUNCH_RENDER(false, true, false, false, tileCounter, none); else ICB_LAUNCH_RENDER(false, false, false, false, tileCounter, none);
  } else {
    // the chunk pass of the frame's heaviest tiles (exits at once when there are none), then the pixel pass
    if (prism) {
      if (sp.nHeavy) ICB_LAUNCH_RENDER(false, true, true, true, tileCounter + 1, sp);
      ICB_LAUNCH_RENDER(false, true, false, true, tileCounter, sp);
    } else {
      if (sp.nHeavy) ICB_LAUNCH_RENDER(false, false, true, true, tileCounter + 1, sp);
      ICB_LAUNCH_RENDER(false, false, false, true, tileCounter, sp);
    }
  }
#undef ICB_LAUNCH_RENDER
  return cudaGetLastError();
}

constexpr int kWgLutSmall = 320, kWgLutMedium = 1408;

template <int THREADS, int MIN_BLOCKS>
static cudaError_t launchWoodcockT(const DeviceGrid &g, const RenderParams &p, const float4 *lut, const float *logTable,
                                   const RenderTargets &out, cudaStream_t stream)
{
  const size_t smem = (size_t)p.tfSize * sizeof(float4) + (size_t)THREADS * kSlotStride * sizeof(float);
  woodcockKernel<THREADS, MIN_BLOCKS><<<p.tilesX * p.tilesY * (kTilePixels / THREADS), THREADS, smem, stream>>>(g, p, lut, logTable, out);
  return cudaGetLastError();
}

cudaError_t launchWoodcock(const DeviceGrid &g, const RenderParams &p, const float4 *lut, const float *logTable,
                           const RenderTargets &out, cudaStream_t stream)
{
  if (p.tfSize <= kWgLutSmall) return launchWoodcockT<64, 10>(g, p, lut, logTable, out, stream);
  if (p.tfSize <= kWgLutMedium) return launchWoodcockT<128, 5>(g, p, lut, logTable, out, stream);
  return launchWoodcockT<256, 2>(g, p, lut, logTable, out, stream);
}

template <int THREADS, int MIN_BLOCKS>
static cudaError_t launchWoodcockGridT(const DeviceGrid &g, const RenderParams &p, const float4 *lut, const float *logTable,
                                       const MajorantGrid &mg, const RenderTargets &out, cudaStream_t stream)
{
  const size_t smem = (size_t)p.tfSize * sizeof(float4) + (size_t)THREADS * kSlotStride * sizeof(float);
  woodcockGridKernel<THREADS, MIN_BLOCKS><<<p.tilesX * p.tilesY * (kTilePixels / THREADS), THREADS, smem, stream>>>(g, p, lut, logTable, mg, out);
  return cudaGetLastError();
}

cudaError_t launchWoodcockGrid(const DeviceGrid &g, const RenderParams &p, const float4 *lut, const float *logTable,
                               const MajorantGrid &mg, const RenderTargets &out, cudaStream_t stream)
{
  // the block size that keeps most threads resident next to the CTA's copy of the table (228 KB of shared memory per SM)
  if (p.tfSize <= kWgLutSmall) return launchWoodcockGridT<64, 10>(g, p, lut, logTable, mg, out, stream);      // 10 x 64
  if (p.tfSize <= kWgLutMedium) return launchWoodcockGridT<128, 5>(g, p, lut, logTable, mg, out, stream);     // 5..4 x 128
  return launchWoodcockGridT<256, 2>(g, p, lut, logTable, mg, out, stream);
}

template <int THREADS, int MIN_BLOCKS>
static cudaError_t launchWoodcockSphereT(const DeviceGrid &g, const RenderParams &p, const float4 *lut, const float *logTable,
                                         const MajorantGrid &sg, const RenderTargets &out, cudaStream_t stream)
{
  const size_t smem = (size_t)p.tfSize * sizeof(float4) + (size_t)THREADS * kSlotStride * sizeof(float);
  woodcockSphereKernel<THREADS, MIN_BLOCKS><<<p.tilesX * p.tilesY * (kTilePixels / THREADS), THREADS, smem, stream>>>(g, p, lut, logTable, sg, out);
  return cudaGetLastError();
}

cudaError_t launchWoodcockSphere(const DeviceGrid &g, const RenderParams &p, const float4 *lut, const float *logTable,
                                 const MajorantGrid &sg, const RenderTargets &out, cudaStream_t stream)
{
  if (p.tfSize <= kWgLutSmall) return launchWoodcockSphereT<64, 10>(g, p, lut, logTable, sg, out, stream);
  if (p.tfSize <= kWgLutMedium) return launchWoodcockSphereT<128, 5>(g, p, lut, logTable, sg, out, stream);
  return launchWoodcockSphereT<256, 2>(g, p, lut, logTable, sg, out, stream);
}

cudaError_t launchShellRanges(const DeviceGrid &g, const float *colTrig, const MajorantGrid &sg, float *ranges, cudaStream_t stream)
{
  const size_t numMC = (size_t)sg.dims[0] * sg.dims[1] * sg.dims[2];
  initRangesKernel<<<(unsigned)((numMC + 255) / 256), 256, 0, stream>>>(reinterpret_cast<float2 *>(ranges), numMC);
  if (g.nRec > 0)
    shellRangesKernel<<<(unsigned)(((size_t)g.nRec + 255) / 256), 256, 0, stream>>>(g.rec, g.lines, colTrig, (size_t)g.nRec, sg,
                                                                                   reinterpret_cast<float2 *>(ranges),
                                                                                   g.sampler != 0 ? g.vtxScalars : nullptr, g.vtxStride);
  return cudaGetLastError();
}

cudaError_t launchLogfCheck(const float *table, unsigned long long *mismatches, cudaStream_t stream)
{
  logfCheckKernel<<<(1u << 24) / 256, 256, 0, stream>>>(table, mismatches);
  return cudaGetLastError();
}

cudaError_t launchGridRanges(const DeviceGrid &g, const float *colTrig, const MajorantGrid &mg, float *ranges, cudaStream_t stream)
{
  const size_t numMC = (size_t)mg.dims[0] * mg.dims[1] * mg.dims[2];
  initRangesKernel<<<(unsigned)((numMC + 255) / 256), 256, 0, stream>>>(reinterpret_cast<float2 *>(ranges), numMC);
  const size_t threads = (size_t)g.nRec * 32;
  if (threads > 0)
    gridRangesKernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(g.rec, g.lines, colTrig, (size_t)g.nRec, mg,
                                                                            reinterpret_cast<float2 *>(ranges),
                                                                            g.sampler != 0 ? g.vtxScalars : nullptr, g.vtxStride);
  return cudaGetLastError();
}

cudaError_t launchMaxOpacities(const float *ranges, const MajorantGrid &mg, const float4 *lut, int size, float tfLo, float tfHi,
                               float *maxOp, uint32_t *occupied, cudaStream_t stream)
{
  const size_t numMC = (size_t)mg.dims[0] * mg.dims[1] * mg.dims[2];
  maxOpacitiesKernel<<<(unsigned)((numMC + 255) / 256), 256, 0, stream>>>(reinterpret_cast<const float2 *>(ranges), numMC, lut, size,
                                                                          tfLo, tfHi, maxOp, occupied);
  return cudaGetLastError();
}

cudaError_t launchOrderTiles(const int *tileList, int numTiles, const int *tileCost, int *order, int minCostLog2, int maxHeavy, int *nHeavy,
                             cudaStream_t stream)
{
  orderTilesKernel<<<1, 1024, 0, stream>>>(tileList, numTiles, tileCost, order, minCostLog2, maxHeavy, nHeavy);
  return cudaGetLastError();
}

cudaError_t launchPrepareFrame(int *tileCounter, Counters *counters, int *tileCost, int nTiles, int *cancelFrom, int nCancel, cudaStream_t stream)
{
  const int blocks = (std::max(nTiles, 1) + 255) / 256;
  prepareFrameKernel<<<blocks, 256, 0, stream>>>(tileCounter, counters, tileCost, nTiles, cancelFrom, nCancel);
  return cudaGetLastError();
}

cudaError_t configureKernels()
{
  cudaError_t e = cudaSuccess;
  const int smemCell = (int)renderSmemBytes(kMaxLutEntries), smemPrism = (int)renderSmemBytes(kMaxLutEntries, true);
#define ICB_CFG_RENDER(T, P, PC, CH) \
  if ((e = cudaFuncSetAttribute(renderKernel<T, P, PC, CH>, cudaFuncAttributeMaxDynamicSharedMemorySize, P ? smemPrism : smemCell)) != cudaSuccess) return e
  ICB_CFG_RENDER(false, false, false, false); ICB_CFG_RENDER(false, false, false, true); ICB_CFG_RENDER(false, false, true, true);
  ICB_CFG_RENDER(true, false, false, false);  ICB_CFG_RENDER(true, false, false, true);
  ICB_CFG_RENDER(false, true, false, false);  ICB_CFG_RENDER(false, true, false, true);  ICB_CFG_RENDER(false, true, true, true);
  ICB_CFG_RENDER(true, true, false, false);   ICB_CFG_RENDER(true, true, false, true);
#undef ICB_CFG_RENDER
  e = cudaFuncSetAttribute(woodcockKernel<64, 10>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWgLutSmall * 16 + 64 * kSlotStride * 4);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(woodcockKernel<128, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWgLutMedium * 16 + 128 * kSlotStride * 4);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(woodcockKernel<256, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxLutEntries * 16 + 256 * kSlotStride * 4);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(woodcockGridKernel<64, 10>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWgLutSmall * 16 + 64 * kSlotStride * 4);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(woodcockGridKernel<128, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWgLutMedium * 16 + 128 * kSlotStride * 4);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(woodcockGridKernel<256, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxLutEntries * 16 + 256 * kSlotStride * 4);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(woodcockSphereKernel<64, 10>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWgLutSmall * 16 + 64 * kSlotStride * 4);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(woodcockSphereKernel<128, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWgLutMedium * 16 + 128 * kSlotStride * 4);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(woodcockSphereKernel<256, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxLutEntries * 16 + 256 * kSlotStride * 4);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(locateKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 256 * kSlotStride * 4);
}

int renderBlocksPerSM(int tfSize, bool prism)
{
  int n = 0;
  cudaError_t e = prism ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, renderKernel<false, true, false, true>, kRenderThreads, renderSmemBytes(tfSize, true))
                        : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, renderKernel<false, false, false, true>, kRenderThreads, renderSmemBytes(tfSize));
  if (e != cudaSuccess) return 1;
  return n > 0 ? n : 1;
}

cudaError_t launchPrismMask(const DeviceGrid &g, const int *nz, float tfLo, float tfHi, int tfSize, uint32_t *mask, cudaStream_t stream)
{
  const unsigned blocks = (unsigned)(((size_t)g.nRec + 255) / 256);
  if (blocks == 0) return cudaSuccess;
  prismMaskKernel<<<blocks, 256, 0, stream>>>(g.rec, (size_t)g.nRec, g.vtxScalars, g.vtxStride, nz, tfLo, tfHi, tfSize, mask);
  return cudaGetLastError();
}

cudaError_t launchVertexScalars(const DeviceGrid &g, const int *colLayers, const int *vtxOff, const int *vtxCols, size_t nVert,
                                float *out, cudaStream_t stream)
{
  const size_t total = nVert * (size_t)g.vtxStride;
  const unsigned blocks = (unsigned)((total + 255) / 256);
  if (blocks == 0) return cudaSuccess;
  vertexScalarsKernel<<<blocks, 256, 0, stream>>>(g.rec, g.lines, g.col, colLayers, vtxOff, vtxCols, nVert, g.vtxStride, out);
  return cudaGetLastError();
}

cudaError_t launchTracePixel(const DeviceGrid &g, const RenderParams &p, const float4 *lut, int px, int py, int *ids,
                             int maxIds, float4 *rgbaOut, int *nStepsOut, cudaStream_t stream)
{
  tracePixelKernel<<<1, 32, 0, stream>>>(g, p, lut, px, py, ids, maxIds, rgbaOut, nStepsOut);
  return cudaGetLastError();
}

cudaError_t launchLocate(const DeviceGrid &g, const float *pos, size_t n, int *ids, float *values, unsigned flags,
                         Counters *counters, cudaStream_t stream)
{
  const size_t per = (flags & 1u) ? 64 : 1;
  const size_t threads = (n + per - 1) / per;
  const unsigned blocks = (unsigned)((threads + 255) / 256);
  if (blocks == 0) return cudaSuccess;
  locateKernel<<<blocks, 256, 256 * kSlotStride * 4, stream>>>(g, pos, n, ids, values, flags, counters);
  return cudaGetLastError();
}

cudaError_t launchScatterScalars(const float *src, const int *origToInt, float *dst, size_t nRec, cudaStream_t stream)
{
  const size_t total = nRec * 8;
  const unsigned blocks = (unsigned)((total + 255) / 256);
  if (blocks == 0) return cudaSuccess;
  scatterScalarsKernel<<<blocks, 256, 0, stream>>>(reinterpret_cast<const float4 *>(src), origToInt,
                                                   reinterpret_cast<float4 *>(dst), nRec);
  return cudaGetLastError();
}

cudaError_t launchScatterTiles(const unsigned *rgba8, const float *accumT, const int *tileIds, int nTiles, unsigned *fb,
                               float *accum, int width, int height, cudaStream_t stream)
{
  if (nTiles <= 0) return cudaSuccess;
  const int tilesX = (width + kTile - 1) / kTile;
  scatterTilesKernel<<<nTiles, kTilePixels, 0, stream>>>(rgba8, reinterpret_cast<const float4 *>(accumT), tileIds, nTiles, fb,
                                                         reinterpret_cast<float4 *>(accum), width, height, tilesX);
  return cudaGetLastError();
}

}  // namespace icb
