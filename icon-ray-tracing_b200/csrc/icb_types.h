// Shared host/device data layout of the B200 icon_rt path.  See DESIGN.md "HBM layout".
#pragma once
#include <cstdint>

namespace icb {

// One 128-byte node per record (ICONCell, icon_rt/ICONGrid.h:59-76), read as 16-byte vectors:
// the three side planes exactly as sample() builds them (ICONGrid.h:187-203), computed on the
// host with the reference's operand order so that the device predicate is bit-identical, plus
// everything a ray needs to step to an adjacent record without touching any other array
// (DESIGN.md "one round trip per crossing").
struct alignas(16) RecBlock {
  float p0[4], p1[4], p2[4];  // makePlane(bv1,bv2,tv2), (bv2,bv3,tv3), (bv3,bv1,tv1): N.xyz, w
  float hbot, htop;           // height[0], height[numLayers]
  float thr;                  // safe-interior threshold = marginMeters * max_k |N_k|  (+inf: never trust)
  int32_t orig;               // index in the caller's array (the reference's record id)
  int32_t nbr[3];             // record across side plane k at the same stacking index (clamped to the
                              // neighbour column's top record), -1 = none
  int32_t info;               // REC_* bit fields below (includes numLayers)
  // -- the 80 bytes above are what a crossing reads (five 16-byte vectors, three 32-byte sectors) --
  int32_t nLayers;            // numLayers (also in info)
  int32_t col;                // internal column id
  // -- prism sampler (read at value set-up only: the rest of the node's 128-byte line) --
  float   kappa[3];           // evalPlane(p_k, P) * kappa[k] = unnormalised barycentric weight of the corner opposite plane k
  float   M[3];               // dot(P, M) = scale factor rho of P's central projection onto the column's chord triangle
  int32_t vert[3];            // welded vertex ids of the three corners (rows of DeviceGrid::vtxScalars)
  int32_t layerBase;          // layers of the column below this record
};
static_assert(sizeof(RecBlock) == 128, "RecBlock must be 128 bytes");

// RecBlock::info
enum : int32_t {
  REC_K_MASK     = 0xff,      // stacking index of the record inside its column (bottom = 0)
  REC_NREC_SHIFT = 8,         // bits 8..15: records in the column
  REC_GENERIC    = 1 << 16,   // column topology/ordering not trusted: exact candidate query only
  REC_KB_SHIFT   = 17,        // bits 17..22: for edge k, which side plane of the neighbour is the shared edge
                              // (2 bits each, 3 = unknown -> no two-candidate resolution across that edge)
  REC_SAMEK_SHIFT = 23,       // bits 23..25: neighbour across edge k has a record with the same stacking index
  REC_NL_SHIFT   = 26,        // bits 26..30: numLayers (0..31)
};

enum : int32_t { COL_GENERIC = 1 };

// One 32-byte block per column (= set of records sharing the same lat/lon triple)
struct alignas(16) ColBlock {
  int32_t nbr[3];    // column across side plane k (internal id), -1 = none
  int32_t recBegin;  // first internal record id; the column's records are contiguous, sorted by hbot
  int32_t nRec;
  int32_t flags;     // COL_GENERIC: topology/ordering not trusted -> always exact candidate query
  float   rLo, rHi;  // min hbot / max htop over the column's records
};
static_assert(sizeof(ColBlock) == 32, "ColBlock must be 32 bytes");

constexpr int kHeightStride = 32;  // floats per record of heights and of values (one 128-byte line each)
constexpr int kLineStride = 64;    // floats per record in lines[]: height[0..31] then value[0..31] (256 bytes)

struct DeviceGrid {
  const RecBlock *rec;
  const float    *lines;     // nRec * 64: height[0..numLayers] | value[0..numLayers) of each record
  const ColBlock *col;
  const int32_t  *seed;      // 6*binN*binN * 2: {bottom, top} record of the column nearest to the bin's centre (-1: none)
  const int32_t  *binOff;    // 6*binN*binN + 1
  const int32_t  *binList;   // column ids
  const int32_t  *globalRecs;  // records checked for every sample (degenerate geometry), ascending recOrig
  int32_t nGlobal;
  int32_t recOfOrig0;          // internal id of the caller's record 0 (answer for NaN positions)
  int32_t binN;
  int32_t nRec, nCol;
  int32_t maxWalk;
  // prism sampler (icb_set_sampler): per welded vertex, the scalar at every layer interface of its columns
  const float *vtxScalars;     // numVertices * vtxStride
  int32_t vtxStride;
  int32_t sampler;             // 0: value[layer] (ICONGrid.h:146-153), 1: wedge shape functions (UElems.h:176-187)
  // prism sampler, empty-space skipping: bit i of prismMask[rec] = every point of layer i of the record classifies to
  // exactly zero opacity under the current transfer function (prismMaskKernel); travels with the record's lines
  const uint32_t *prismMask;   // nRec, or null
  // radial cull of the Woodcock ray-gens: no record accepts a radius below min(height[0]) or above
  // max(height[numLayers]) (ICONGrid.h:184); -INF / +INF when a record carries a NaN there (its test passes everything)
  float cullLo, cullHi;
};

// per-lane line slot, prism sampler: floats 0..31 heights, 32..43 the node's prism tail, 44..49 the located
// prism's six vertex scalars (bottom 1,2,3, top 1,2,3), 50..51 its bottom height and thickness
constexpr int kSlotTail = 32, kSlotKappa = 34, kSlotM = 37, kSlotVert = 40, kSlotBase = 43, kSlotScal = 44, kSlotH0 = 50, kSlotDh = 51,
              kSlotMask = 64,   // (first of the slot's four pad words) the record's transparency mask
              kSlotLog = 65;    // two pad words: the log-table entries the default ray-gen's next free flight can need

// the reference's Grid accel (Params.h:44-49) as the default ray-gen reads it
constexpr int kColTrigStride = 18;   // per column: (cos lat, sin lat, cos lon, sin lon) x 3 corners (host libm), then lat[3], lon[3]

struct MajorantGrid {
  const float *maxOpacities;   // dims.x*dims.y*dims.z (hostCode.cu:400-434)
  const uint32_t *occupied;    // one bit per macrocell: majorant > 0 (null: test the majorant itself)
  int32_t dims[3];
  float lo[3], hi[3];          // worldBounds = the volume's Cartesian bounds (hostCode.cu:668-670)
};

struct RenderParams {
  float org[3], d00[3], du[3], dv[3];
  float ambient[3];
  float ambientRadiance;
  float unitDistance, dt, termAlpha;
  float rInner, rOuter;
  float tfLo, tfHi, opacityScale;
  int32_t tfSize;
  int32_t width, height;
  int32_t accumID;
  uint32_t flags;
  float skipPlaneScale, skipRadSlack;  // skip-ahead slack (locator.cuh skipCount); maxSkip == 0 disables it
  int32_t maxSkip;
  int32_t skipBack;                   // samples held back from the computed skip length (0 or 1)
  int32_t tilesX, tilesY, numTiles;   // numTiles = entries of tileList if non-null else tilesX*tilesY
  float volLo[3], volHi[3];           // Cartesian bounds of the volume (boxTest of the Woodcock ray-gen)
  // marcher: a ray's samples are composited in chunks of 2^chunkLog2 consecutive sample indices per shell segment, the
  // chunks merged in order with `over` (locator.cuh compositeSample / mergeChunk); the chunks of a long ray may then be
  // marched by different lanes without changing a bit of the result
  int32_t chunkLog2;
};

// what one chunk of a split ray leaves behind for the lane that assembles the pixel (32 bytes)
struct alignas(16) ChunkResult {
  float c[4];          // the chunk's own front-to-back composite (premultiplied colour, opacity)
  int32_t steps, hits; // samples taken / located inside the chunk
  int32_t flags;       // 1: the chunk stopped early (its own opacity reached the termination threshold)
  int32_t pad;
};
constexpr int kMaxChunks = 16;        // chunks per split ray (rays with more are marched by one lane)

// split-ray working set of one frame slot (render kernels: chunk pass writes, pixel pass reads)
struct SplitArgs {
  const int32_t *nHeavy;   // device: number of leading entries of the frame's tile order whose rays are split (null: none)
  ChunkResult *results;    // [heavy tile][pixel in tile][chunk]
  int32_t *cancelFrom;     // [heavy tile][pixel in tile]: chunks >= this index are known not to contribute
};

struct Counters {
  unsigned long long steps, hits, slow, exact, rays;
  unsigned long long totalSteps, totalHits;   // never reset by a frame (icb_reset_totals)
  // icb_render_trace only (icb_debug_counters): warp iterations of the march loop, lane iterations, lane
  // iterations that stayed in their prism / changed layer / fetched a node / came back undecided / skipped nothing
  unsigned long long dbg[8];
};

}  // namespace icb
