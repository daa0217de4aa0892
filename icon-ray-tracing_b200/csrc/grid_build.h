// Host-side preprocessing of an ICONCell[] array into the device layout (icb_types.h).
#pragma once
#include <algorithm>
#include <cstddef>
#include <cstdint>
#include <memory>
#include <string>
#include <utility>
#include <vector>

#include "icb_types.h"
#include <cmath>
#include "icon_b200.h"

namespace icb {

// std::vector without the serial zero fill of resize(): the builder writes every element of these arrays from its
// parallel loops (first touch by the thread that fills the page), 2 GB of nodes at R2B8 x 90, 32 GB at R2B10 x 90
template <typename T>
struct NoInit : std::allocator<T> {
  template <typename U> struct rebind { using other = NoInit<U>; };
  NoInit() = default;
  template <typename U> NoInit(const NoInit<U> &) {}
  template <typename U> void construct(U *p) { ::new ((void *)p) U; }
  template <typename U, typename... A> void construct(U *p, A &&...a) { ::new ((void *)p) U(std::forward<A>(a)...); }
};
template <typename T> using RawVec = std::vector<T, NoInit<T>>;

struct HostGrid {
  RawVec<RecBlock> rec;               // internal (space-filling-curve) order
  std::vector<float>    lines;        // nRec*64: height[32] | value[32]
  float cullLo = -INFINITY, cullHi = INFINITY;   // DeviceGrid::cullLo/cullHi
  RawVec<int32_t>  origToInt;         // caller index -> internal
  RawVec<ColBlock> col;
  std::vector<int32_t>  binOff, binList;
  std::vector<int32_t>  seed;         // 2 per bin: bottom and top record of the column nearest to the bin centre
  std::vector<int32_t>  globalRecs;   // internal ids of records with an unbounded acceptance region
  // prism sampler: welded vertex -> (column, corner) pairs in ascending caller order (internal column ids),
  // layers per column, and the longest column (+1 = interfaces per vertex row)
  RawVec<int32_t>  vtxOff, vtxCols;
  RawVec<int32_t>  colLayers;
  // majorant-grid builder: per column and corner {cosf(lat), sinf(lat), cosf(lon), sinf(lon)} from the host's libm
  RawVec<float>    colTrig;
  int32_t maxColLayers = 0;
  int32_t binN = 0;
  icb_bounds bounds{};
  int64_t numVertices = 0, numGeneric = 0;
};

// marginMeters: half-width of the band around side planes inside which the walk defers to the
// exact candidate query.  Returns ICB_OK or an icb_error with `err` filled.
// withLines = false leaves out.lines empty: the caller streams them with fillLines (icb_set_grid does, so that the
// 256 bytes per record never exist as a whole in host memory -- 64 GB for R2B10 x 90).
int buildGrid(const void *records, size_t n, float marginMeters, HostGrid &out, std::string &err, bool withLines = true);

// lines (height[32] | value[32]) of the internal records [begin, end) into dst[(end-begin)*64]; intToOrig = inverse of origToInt
void fillLines(const void *records, const int32_t *intToOrig, size_t begin, size_t end, float *dst);

// memcpy by all host threads
void parallelCopy(void *dst, const void *src, size_t bytes);

// inverse of a permutation (parallel)
void invertPermutation(const int32_t *perm, size_t n, RawVec<int32_t> &inv);

// logf(1 - k/2^24), k = 0 .. 2^24-1, by the host's libm (log_table.cpp)
void buildLogTable(std::vector<float> &table);

// cube-map bin of a direction (double precision, host side); mirrors the device function.
void cubeProject(const double v[3], int &face, double &u, double &w);

}  // namespace icb
