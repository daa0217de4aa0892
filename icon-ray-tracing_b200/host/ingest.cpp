// convert_icon's array -> `.ic` half (reference: tools/convert_icon/convert_icon.cpp:317-391), without NetCDF: the caller
// hands over what the converter has in memory after reading its files -- corner angles (clon_vertices / clat_vertices,
// double, 3 per cell), HSURF (double per cell), the HHL level fields and the data level fields (float per cell, each
// with its `height` level index) -- and gets the 284-byte ICONCell records icb_set_grid takes.
//
// ICB_INGEST_REFERENCE reproduces the converter statement for statement, defects included:
//   * levels are sorted by `height` DESCENDING (:296, :331) -- bottom-up for ICON's top-down level numbering;
//   * numLayers = number of data files, silently capped at maxLayers (:345-349; the converter's default is 5);
//   * records of LMAX-1 = 31 layers; the last one has `numLayers % LMAX-1` = (numLayers % 32) - 1 layers (:363-366):
//     90 levels become 31 + 31 + 25, 32 levels become 31 and a record with numLayers = -1;
//   * H[0] of a column's first record is R + hsurf, of a later record the last height of the record before; H[j] is
//     R + HHL_j - hsurf (:360-373), float R = 6.371229e6, each sum in the converter's operand types.
// The converter leaves H[numLayersLocal+1..31] and value[numLayersLocal..31] as uninitialised stack; they are zero here.
// ICB_INGEST_FIXED differs in one place: the last record gets the layers that are left (31 + 31 + 28).
// ICB_INGEST_SURFACE_LEVEL (a flag, or-ed to either mode) repairs what SURVEY's list does not name: the converter starts a column
// at R + hsurf and then continues with R + HHL - hsurf, whose first value -- the lowest half level IS the surface -- is R: over
// land the bottom layer is upside down (and every data level sits one layer too low, the lowest one in that inverted layer).
// With the flag the column starts at the lowest half level, H[0] = R + HHL_bottom - hsurf, and data level j lies between its own
// two half levels; heights are then non-decreasing wherever the half levels are.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <thread>
#include <vector>

#include "icon_b200.h"

namespace {

#pragma pack(push, 1)
struct Record {
  float   lat[3], lon[3];
  int32_t numLayers;
  float   height[32], value[32];
};
#pragma pack(pop)
static_assert(sizeof(Record) == ICB_RECORD_BYTES, "ICONCell record must be 284 bytes");

constexpr int LMAX = 32;
inline int divUp(int a, int b) { return (a + b - 1) / b; }

inline int cappedLayers(int nData, int maxLayers) { return maxLayers >= 0 && nData > maxLayers ? maxLayers : nData; }

// layers of record i of a column with numLayers layers (convert_icon.cpp:362-366)
inline int localLayers(int numLayers, int i, int mode)
{
  int nl = LMAX - 1;
  if ((i + 1) * nl > numLayers) nl = mode == ICB_INGEST_FIXED ? numLayers - i * (LMAX - 1) : numLayers % LMAX - 1;
  return nl;
}

// order of the level fields: descending `height` (std::sort at :296/:331; ties keep file order here)
std::vector<int> descending(const int32_t *heights, int n)
{
  std::vector<int> o(n);
  std::iota(o.begin(), o.end(), 0);
  std::stable_sort(o.begin(), o.end(), [&](int a, int b) { return heights[a] > heights[b]; });
  return o;
}

} // namespace

extern "C" {

int64_t icb_ingest_num_records(int64_t cells, int numDataLevels, int maxLayers)
{
  if (cells < 0 || numDataLevels < 0) return ICB_ERR_INVALID;
  return cells * divUp(cappedLayers(numDataLevels, maxLayers), LMAX - 1);
}

// convert_icon.cpp:317-333: a data level is shifted and scaled to [0,1] by its own extrema (fmin/fmax: NaNs are passed
// over), in double, then narrowed to float
int icb_ingest_normalize(const double *var, int64_t n, float *out)
{
  if (!var || !out || n < 0) return ICB_ERR_INVALID;
  double lo = DBL_MAX, hi = -DBL_MAX;
  for (int64_t j = 0; j < n; ++j) { lo = fmin(lo, var[j]); hi = fmax(hi, var[j]); }
  for (int64_t j = 0; j < n; ++j) {
    double v = var[j];
    v -= lo;
    v /= hi - lo;
    out[j] = (float)v;
  }
  return ICB_OK;
}

int64_t icb_ingest_columns(int64_t cells, const double *clon_vertices, const double *clat_vertices, const double *hsurf,
                           int numHhl, const int32_t *hhlHeights, const float *const *hhlValues,
                           int numData, const int32_t *dataHeights, const float *const *dataValues,
                           int maxLayers, int mode, void *records, int64_t capacity)
{
  if (cells < 0 || !clon_vertices || !clat_vertices || !hsurf || numHhl < 0 || numData < 0 || (numHhl && (!hhlHeights || !hhlValues)) ||
      (numData && (!dataHeights || !dataValues)) || (mode & ~(ICB_INGEST_FIXED | ICB_INGEST_SURFACE_LEVEL)) != 0)
    return ICB_ERR_INVALID;
  const bool surfaceLevel = (mode & ICB_INGEST_SURFACE_LEVEL) != 0;
  mode &= ICB_INGEST_FIXED;
  const int numLayers = cappedLayers(numData, maxLayers);
  const int perColumn = divUp(numLayers, LMAX - 1);
  const int64_t total = cells * perColumn;
  if (!records) return total;
  if (capacity < total) return ICB_ERR_INVALID;
  // the converter indexes hhl[] and values[] with running counters and no bound: say so instead of reading past the end
  int needHhl = 0, needVal = 0;
  for (int i = 0; i < perColumn; ++i) { const int nl = std::max(0, localLayers(numLayers, i, mode)); needHhl += nl; needVal += nl; }
  if (surfaceLevel && perColumn > 0) needHhl += 1;
  if (needHhl > numHhl || needVal > numData) return ICB_ERR_INVALID;
  const std::vector<int> hhlOrder = descending(hhlHeights, numHhl), valOrder = descending(dataHeights, numData);
  Record *out = static_cast<Record *>(records);

  auto work = [&](int64_t c0, int64_t c1) {
    for (int64_t cellID = c0; cellID < c1; ++cellID) {
      float lat[3] = {(float)clat_vertices[cellID * 3], (float)clat_vertices[cellID * 3 + 1], (float)clat_vertices[cellID * 3 + 2]};
      float lon[3] = {(float)clon_vertices[cellID * 3], (float)clon_vertices[cellID * 3 + 1], (float)clon_vertices[cellID * 3 + 2]};
      constexpr float R = 6.371229E6f;
      int valueIt = 0, hhlIt = 0;
      float prevH = (float)(R + hsurf[cellID]);                  // float + double, narrowed on assignment (:360)
      if (surfaceLevel && perColumn > 0) prevH = (float)((R + hhlValues[hhlOrder[hhlIt++]][cellID]) - hsurf[cellID]);
      for (int i = 0; i < perColumn; ++i) {
        const int nl = localLayers(numLayers, i, mode);
        Record &o = out[cellID * perColumn + i];
        std::memset(&o, 0, sizeof o);
        std::memcpy(o.lat, lat, sizeof lat);
        std::memcpy(o.lon, lon, sizeof lon);
        o.numLayers = nl;
        o.height[0] = prevH;
        for (int j = 1; j <= nl; ++j) {
          o.height[j] = (float)((R + hhlValues[hhlOrder[hhlIt++]][cellID]) - hsurf[cellID]);   // (float + float) - double (:371)
          prevH = o.height[j];
        }
        for (int j = 0; j < nl; ++j) o.value[j] = dataValues[valOrder[valueIt++]][cellID];
      }
    }
  };
  unsigned nt = std::max(1u, std::min(64u, std::thread::hardware_concurrency()));
  if (cells < 4096) nt = 1;
  std::vector<std::thread> th;
  for (unsigned t = 0; t < nt; ++t) th.emplace_back(work, cells * t / nt, cells * (t + 1) / nt);
  for (auto &t : th) t.join();
  return total;
}

} // extern "C"
