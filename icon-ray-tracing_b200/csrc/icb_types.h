// Shared host/device data layout of the B200 icon_rt path.  See DESIGN.md "HBM layout".
#pragma once
#include <cstdint>

namespace icb {

// One 64-byte block per record (ICONCell, icon_rt/ICONGrid.h:59-76), 16-byte vectors:
// the three side planes exactly as sample() builds them (ICONGrid.h:187-203), computed on the
// host with the reference's operand order so that the device predicate is bit-identical.
struct alignas(16) RecBlock {
  float p0[4], p1[4], p2[4];  // makePlane(bv1,bv2,tv2), (bv2,bv3,tv3), (bv3,bv1,tv1): N.xyz, w
  float hbot, htop;           // height[0], height[numLayers]
  float thr;                  // safe-interior threshold = marginMeters * max_k |N_k|  (+inf: never trust)
  int32_t nLayers;            // numLayers
};
static_assert(sizeof(RecBlock) == 64, "RecBlock must be 64 bytes");

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

constexpr int kHeightStride = 32;  // floats per record in heights[] and values[] (one 128-byte line each)

struct DeviceGrid {
  const RecBlock *rec;
  const float    *heights;   // nRec * 32: height[0..numLayers] of each record
  const float    *values;    // nRec * 32: value[0..numLayers)
  const int32_t  *recOrig;   // internal record -> index in the caller's array (the reference's record id)
  const ColBlock *col;
  const int32_t  *binOff;    // 6*binN*binN + 1
  const int32_t  *binList;   // column ids
  const int32_t  *globalRecs;  // records checked for every sample (degenerate geometry), ascending recOrig
  int32_t nGlobal;
  int32_t recOfOrig0;          // internal id of the caller's record 0 (answer for NaN positions)
  int32_t binN;
  int32_t nRec, nCol;
  int32_t maxWalk;
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
  int32_t tilesX, tilesY, numTiles;   // numTiles = entries of tileList if non-null else tilesX*tilesY
};

struct Counters {
  unsigned long long steps, hits, slow, exact, rays;
  unsigned long long totalSteps, totalHits;   // never reset by a frame (icb_reset_totals)
};

}  // namespace icb
