// Host-side grid preprocessing.  Compiled with -ffp-contract=off: the per-record side planes
// must come out bit-identical to what the reference's sample() recomputes per call
// (icon_rt/ICONGrid.h:187-203 with toCartesian :44-54 and makePlane :168-174), because glibc's
// cosf/sinf cannot be reproduced on the device.  Everything the device evaluates per sample
// (dot products, length, comparisons) is plain IEEE fp32 and is done there without FMA.
#include "grid_build.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <thread>

namespace icb {
namespace {

#pragma pack(push, 1)
struct Record {  // ICONCell, icon_rt/ICONGrid.h:59-76; tools/convert_icon/convert_icon.cpp:383-387
  float   lat[3], lon[3];
  int32_t numLayers;
  float   height[32], value[32];
};
#pragma pack(pop)
static_assert(sizeof(Record) == ICB_RECORD_BYTES, "record size");

struct F3 { float x, y, z; };

// fminf(acc, v) / fmaxf(acc, v) for an accumulator that is never NaN (it starts at +-inf): a NaN v leaves it alone and a
// tie keeps it, as the libm functions do -- inline, where the library calls cost more than the loop around them
inline float accMin(float acc, float v) { return v < acc ? v : acc; }
inline float accMax(float acc, float v) { return v > acc ? v : acc; }

// (r,lat,lon) -> Cartesian with the reference's operand order: (r*cos(lat))*cos(lon), ...
inline F3 sphToCart(float r, float lat, float lon)
{
  F3 p;
  p.x = r * cosf(lat) * cosf(lon);
  p.y = r * cosf(lat) * sinf(lon);
  p.z = r * sinf(lat);
  return p;
}

// plane through a,b,c: N = (b-a) x (c-a), w = a.N, each product/sum rounded separately
inline void planeThrough(const F3 &a, const F3 &b, const F3 &c, float out[4])
{
  const float ux = b.x - a.x, uy = b.y - a.y, uz = b.z - a.z;
  const float vx = c.x - a.x, vy = c.y - a.y, vz = c.z - a.z;
  const float nx = uy * vz - uz * vy;
  const float ny = uz * vx - ux * vz;
  const float nz = ux * vy - uy * vx;
  out[0] = nx; out[1] = ny; out[2] = nz;
  out[3] = a.x * nx + a.y * ny + a.z * nz;
}

template <typename F>
void parallelFor(size_t n, F fn)
{
  unsigned nt = std::max(1u, std::min(64u, std::thread::hardware_concurrency()));
  if (n < 8192) nt = 1;
  if (nt == 1) { fn(0, n, 0u); return; }
  std::vector<std::thread> th;
  for (unsigned t = 0; t < nt; ++t) th.emplace_back(fn, n * t / nt, n * (t + 1) / nt, t);
  for (auto &t : th) t.join();
}
inline unsigned numWorkers(size_t n)
{
  unsigned nt = std::max(1u, std::min(64u, std::thread::hardware_concurrency()));
  return n < 8192 ? 1u : nt;
}

inline uint64_t mix64(uint64_t x)
{
  x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33;
  return x;
}

inline uint32_t fbits(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }

// canonical (lat,lon) key of a corner: poles lose their longitude, -pi folds onto +pi, -0 onto +0
inline uint64_t vertexKey(float lat, float lon)
{
  const float halfPi = 1.57079637f, pi = 3.14159274f;
  if (lat == 0.f) lat = 0.f;
  if (lon == 0.f) lon = 0.f;
  if (std::fabs(lat) >= halfPi) lon = 0.f;
  if (lon <= -pi) lon = pi;
  return ((uint64_t)fbits(lat) << 32) | fbits(lon);
}

inline void unitFromLatLon(float lat, float lon, double v[3])
{
  const double cl = std::cos((double)lat);
  v[0] = cl * std::cos((double)lon);
  v[1] = cl * std::sin((double)lon);
  v[2] = std::sin((double)lat);
}

// libm values of a column's six angles: cosf/sinf for the corner positions (sphToCart's operand order), the double
// unit vectors and the prism sampler's M row, which depend on the angles alone
struct CornerTrig {
  float  angles[6];                 // lat[3] | lon[3], the bytes the values below belong to
  float  cla[3], sla[3], clo[3], slo[3];
  double U[3][3];
  float  M[3];
  void set(const float *lat, const float *lon)
  {
    std::memcpy(angles, lat, 12); std::memcpy(angles + 3, lon, 12);
    for (int k = 0; k < 3; ++k) {
      cla[k] = cosf(lat[k]); sla[k] = sinf(lat[k]); clo[k] = cosf(lon[k]); slo[k] = sinf(lon[k]);
      unitFromLatLon(lat[k], lon[k], U[k]);
    }
    const double ax = U[1][0] - U[0][0], ay = U[1][1] - U[0][1], az = U[1][2] - U[0][2];
    const double bx = U[2][0] - U[0][0], by = U[2][1] - U[0][1], bz = U[2][2] - U[0][2];
    const double nx = ay * bz - az * by, ny = az * bx - ax * bz, nz = ax * by - ay * bx;
    const double w = U[0][0] * nx + U[0][1] * ny + U[0][2] * nz;
    const float m[3] = {(float)(nx / w), (float)(ny / w), (float)(nz / w)};
    for (int k = 0; k < 3; ++k) M[k] = std::isfinite(m[k]) ? m[k] : 0.f;
  }
  F3 toCart(float r, int k) const     // sphToCart(r, lat[k], lon[k])
  {
    F3 p;
    p.x = r * cla[k] * clo[k];
    p.y = r * cla[k] * slo[k];
    p.z = r * sla[k];
    return p;
  }
};

inline uint32_t part1by1(uint32_t x)
{
  x &= 0x0000ffff;
  x = (x | (x << 8)) & 0x00ff00ff;
  x = (x | (x << 4)) & 0x0f0f0f0f;
  x = (x | (x << 2)) & 0x33333333;
  x = (x | (x << 1)) & 0x55555555;
  return x;
}

struct Sortable { uint64_t key; uint32_t idx; };

// cube-map bins per face edge, in units of sqrt(columns per face): 1.5 leaves two to three candidate
// columns in a bin (one staged batch of the device query)
constexpr double kBinsPerColumnEdge = 1.5;

inline bool sortableLess(const Sortable &a, const Sortable &b) { return a.key < b.key || (a.key == b.key && a.idx < b.idx); }

// Parallel sample sort on the total order (key, idx) -- the result is the same array whatever the thread count.
// Splitters from a fixed-stride sample, one counting pass and one scatter pass per thread slice, then the buckets
// sorted independently (handed out by an atomic cursor); no serial merge at the end.
void sortByKey(RawVec<Sortable> &v)
{
  const size_t n = v.size();
  const unsigned nt = numWorkers(n);
  if (nt == 1) { std::sort(v.begin(), v.end(), sortableLess); return; }
  const unsigned nb = nt * 16;                       // buckets
  const size_t ns = (size_t)nb * 32;                 // sample size
  std::vector<Sortable> sample(ns);
  for (size_t i = 0; i < ns; ++i) sample[i] = v[(size_t)(((unsigned __int128)i * n) / ns)];
  std::sort(sample.begin(), sample.end(), sortableLess);
  std::vector<Sortable> split(nb - 1);
  for (unsigned b = 1; b < nb; ++b) split[b - 1] = sample[(size_t)b * 32];
  auto bucketOf = [&](const Sortable &x) {           // first bucket whose splitter is > x
    return (unsigned)(std::upper_bound(split.begin(), split.end(), x, sortableLess) - split.begin());
  };
  std::vector<size_t> cnt((size_t)nt * nb, 0);
  RawVec<uint16_t> bkt(n);
  parallelFor(n, [&](size_t a, size_t b, unsigned t) {
    size_t *c = &cnt[(size_t)t * nb];
    for (size_t i = a; i < b; ++i) { const unsigned k = bucketOf(v[i]); bkt[i] = (uint16_t)k; ++c[k]; }
  });
  std::vector<size_t> start(nb + 1, 0);
  {
    size_t run = 0;
    for (unsigned k = 0; k < nb; ++k) {
      start[k] = run;
      for (unsigned t = 0; t < nt; ++t) { const size_t c = cnt[(size_t)t * nb + k]; cnt[(size_t)t * nb + k] = run; run += c; }
    }
    start[nb] = run;
  }
  RawVec<Sortable> tmp(n);
  parallelFor(n, [&](size_t a, size_t b, unsigned t) {
    size_t *c = &cnt[(size_t)t * nb];
    for (size_t i = a; i < b; ++i) tmp[c[bkt[i]]++] = v[i];
  });
  RawVec<uint16_t>().swap(bkt);
  std::atomic<unsigned> next{0};
  {
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nt; ++t)
      th.emplace_back([&] {
        for (;;) {
          const unsigned k = next.fetch_add(1);
          if (k >= nb) break;
          std::sort(tmp.begin() + start[k], tmp.begin() + start[k + 1], sortableLess);
        }
      });
    for (auto &t : th) t.join();
  }
  v.swap(tmp);
}

}  // namespace

void cubeProject(const double v[3], int &face, double &u, double &w)
{
  const double ax = std::fabs(v[0]), ay = std::fabs(v[1]), az = std::fabs(v[2]);
  if (ax >= ay && ax >= az) { face = v[0] > 0 ? 0 : 1; u = v[1] / ax; w = v[2] / ax; }
  else if (ay >= az)        { face = v[1] > 0 ? 2 : 3; u = v[0] / ay; w = v[2] / ay; }
  else                      { face = v[2] > 0 ? 4 : 5; u = v[0] / az; w = v[1] / az; }
}

void parallelCopy(void *dst, const void *src, size_t bytes)
{
  parallelFor(bytes, [&](size_t a, size_t b, unsigned) { std::memcpy((char *)dst + a, (const char *)src + a, b - a); });
}

void invertPermutation(const int32_t *perm, size_t n, RawVec<int32_t> &inv)
{
  inv.resize(n);
  parallelFor(n, [&](size_t a, size_t b, unsigned) {
    for (size_t i = a; i < b; ++i) inv[(size_t)perm[i]] = (int32_t)i;
  });
}

void fillLines(const void *records, const int32_t *intToOrig, size_t begin, size_t end, float *dst)
{
  const Record *R = static_cast<const Record *>(records);
  parallelFor(end - begin, [&](size_t a, size_t b, unsigned) {
    for (size_t i = a; i < b; ++i) {
      const Record &r = R[intToOrig[begin + i]];
      std::memcpy(dst + i * kLineStride, r.height, sizeof(float) * 32);
      std::memcpy(dst + i * kLineStride + kHeightStride, r.value, sizeof(float) * 32);
    }
  });
}

int buildGrid(const void *records, size_t n, float marginMeters, HostGrid &out, std::string &err, bool withLines)
{
  const Record *R = (const Record *)records;
  if (!R || n == 0) { err = "icb_set_grid: empty grid"; return ICB_ERR_INVALID; }
  if (n >= (size_t)1 << 28) { err = "icb_set_grid: more than 2^28 records"; return ICB_ERR_INVALID; }
  out = HostGrid();
  // ICB_BUILD_TIMING=1: seconds per phase on stderr
  const bool timing = std::getenv("ICB_BUILD_TIMING") != nullptr;
  auto tPrev = std::chrono::steady_clock::now();
  auto lap = [&](const char *what) {
    if (!timing) return;
    const auto t = std::chrono::steady_clock::now();
    std::fprintf(stderr, "buildGrid %-34s %.3f s\n", what, std::chrono::duration<double>(t - tPrev).count());
    tPrev = t;
  };

  // ---- 1. validate -----------------------------------------------------------------
  {
    std::atomic<long> bad{-1};
    parallelFor(n, [&](size_t a, size_t b, unsigned) {
      for (size_t i = a; i < b; ++i)
        if (R[i].numLayers < 0 || R[i].numLayers > 31) { long e = -1; bad.compare_exchange_strong(e, (long)i); }
    });
    if (bad.load() >= 0) {
      err = "icb_set_grid: record " + std::to_string(bad.load()) + " has numLayers outside [0,31]";
      return ICB_ERR_BAD_RECORD;
    }
  }

  lap("1 validate");
  // ---- 3. columns = records with bit-identical lat/lon triples ---------------------------
  std::vector<int32_t> colOfRec(n);
  RawVec<int32_t> colFirst;        // representative record of each column (lowest index)
  {
    // run heads: records whose six angles differ from the record in front of them.  A column's stacked records usually
    // follow each other, so only the heads (one per column then) are hashed, sorted and compared
    const unsigned nt = numWorkers(n);
    std::vector<size_t> headCnt(nt + 1, 0);
    parallelFor(n, [&](size_t a, size_t b, unsigned t) {
      size_t c = 0;
      for (size_t i = a; i < b; ++i) c += (i == 0 || std::memcmp(R[i].lat, R[i - 1].lat, 24) != 0);
      headCnt[t + 1] = c;
    });
    for (unsigned t = 0; t < nt; ++t) headCnt[t + 1] += headCnt[t];
    const size_t nHead = headCnt[nt];
    RawVec<Sortable> s(nHead);
    RawVec<int32_t> headOf(n);          // record -> index of its run head in record order
    parallelFor(n, [&](size_t a, size_t b, unsigned t) {
      size_t hI = headCnt[t];
      int32_t cur = -1;
      if (a > 0 && a < b && std::memcmp(R[a].lat, R[a - 1].lat, 24) == 0) {     // the slice starts inside a run
        size_t q = a - 1;
        while (q > 0 && std::memcmp(R[q].lat, R[q - 1].lat, 24) == 0) --q;
        cur = (int32_t)q;
      }
      for (size_t i = a; i < b; ++i) {
        if (i == 0 || std::memcmp(R[i].lat, R[i - 1].lat, 24) != 0) {
          uint64_t w[3];
          std::memcpy(w, R[i].lat, 24);
          s[hI].key = mix64(w[0] ^ mix64(w[1] ^ mix64(w[2] + 0x9e3779b97f4a7c15ull)));
          s[hI].idx = (uint32_t)i;
          ++hI;
          cur = (int32_t)i;
        }
        headOf[i] = cur;
      }
    });
      sortByKey(s);
      // runs of equal hash; split on real inequality (hash collisions).  rep: head record -> representative record
    // (a slice handles the runs that START inside it; only head records are written, each by one thread)
    RawVec<int32_t> rep(n);
    parallelFor(nHead, [&](size_t a, size_t b, unsigned) {
      size_t i = a;
      while (i < b && i > 0 && s[i].key == s[i - 1].key) ++i;
      while (i < b) {
        size_t j = i + 1;
        while (j < nHead && s[j].key == s[i].key) ++j;
        // within [i,j): idx ascending.  Group by actual bytes.
        for (size_t q = i; q < j; ++q) {
          int32_t r = -1;
          for (size_t w = i; w < q; ++w)
            if (std::memcmp(R[s[q].idx].lat, R[s[w].idx].lat, 24) == 0) { r = rep[s[w].idx]; break; }
          rep[s[q].idx] = r >= 0 ? r : (int32_t)s[q].idx;
        }
        i = j;
      }
    });
    // number columns by their representative (= lowest record index) in ascending order: count, prefix, assign
    RawVec<int32_t> colOfRep(n);
    std::vector<size_t> repCnt(nt + 1, 0);
    parallelFor(n, [&](size_t a, size_t b, unsigned t) {
      size_t c = 0;
      for (size_t r = a; r < b; ++r) c += (headOf[r] == (int32_t)r && rep[r] == (int32_t)r);
      repCnt[t + 1] = c;
    });
    for (unsigned t = 0; t < nt; ++t) repCnt[t + 1] += repCnt[t];
    colFirst.resize(repCnt[nt]);
    parallelFor(n, [&](size_t a, size_t b, unsigned t) {
      size_t c = repCnt[t];
      for (size_t r = a; r < b; ++r)
        if (headOf[r] == (int32_t)r && rep[r] == (int32_t)r) { colOfRep[r] = (int32_t)c; colFirst[c] = (int32_t)r; ++c; }
    });
    parallelFor(n, [&](size_t a, size_t b, unsigned) {
      for (size_t r = a; r < b; ++r) colOfRec[r] = colOfRep[rep[headOf[r]]];
    });
  }
  const size_t nCol = colFirst.size();

  // records of each column, sorted by (hbot, index): histogram, prefix, fill by per-column cursors, then each column's list sorted
  RawVec<int32_t> colRecOff(nCol + 1), colRecs(n);
  {
    RawVec<int32_t> cnt(nCol);
    parallelFor(nCol, [&](size_t a, size_t b, unsigned) { for (size_t c = a; c < b; ++c) cnt[c] = 0; });
    parallelFor(n, [&](size_t a, size_t b, unsigned) {
      for (size_t r = a; r < b; ++r) __atomic_fetch_add(&cnt[colOfRec[r]], 1, __ATOMIC_RELAXED);
    });
    {
      // prefix over the columns: per-slice sums, then offsets
      const unsigned ntc = numWorkers(nCol);
      std::vector<int64_t> part(ntc + 1, 0);
      parallelFor(nCol, [&](size_t a, size_t b, unsigned t) {
        int64_t sum = 0;
        for (size_t c = a; c < b; ++c) sum += cnt[c];
        part[t + 1] = sum;
      });
      for (unsigned t = 0; t < ntc; ++t) part[t + 1] += part[t];
      parallelFor(nCol, [&](size_t a, size_t b, unsigned t) {
        int64_t run = part[t];
        for (size_t c = a; c < b; ++c) { colRecOff[c] = (int32_t)run; run += cnt[c]; cnt[c] = 0; }
      });
      colRecOff[nCol] = (int32_t)n;
    }
    parallelFor(n, [&](size_t a, size_t b, unsigned) {
      for (size_t r = a; r < b; ++r) {
        const int32_t c = colOfRec[r];
        colRecs[(size_t)colRecOff[c] + __atomic_fetch_add(&cnt[c], 1, __ATOMIC_RELAXED)] = (int32_t)r;
      }
    });
    parallelFor(nCol, [&](size_t a, size_t b, unsigned) {
      for (size_t c = a; c < b; ++c) {
        const int32_t lo = colRecOff[c], hi = colRecOff[c + 1];
        if (hi - lo < 2) continue;
        // the cursors above hand out a column's slots in whatever order the workers arrive: put the slice back into
        // caller order first.  The comparator below is not a strict weak order when a height[0] is NaN (degenerate
        // records), so the result of the second sort depends on the order it starts from -- it must not depend on timing.
        std::sort(colRecs.begin() + lo, colRecs.begin() + hi);
        std::sort(colRecs.begin() + lo, colRecs.begin() + hi, [&](int32_t x, int32_t y) {
          return R[x].height[0] < R[y].height[0] || (R[x].height[0] == R[y].height[0] && x < y);
        });
      }
    });
  }

  lap("3 columns");
  // ---- 4. vertex welding + edge adjacency ---------------------------------------------------
  RawVec<int32_t> colVert(nCol * 3);
  {
    RawVec<Sortable> s(nCol * 3);
    parallelFor(nCol, [&](size_t a, size_t b, unsigned) {
      for (size_t c = a; c < b; ++c) {
        const Record &r = R[colFirst[c]];
        for (int k = 0; k < 3; ++k) { s[c * 3 + k].key = vertexKey(r.lat[k], r.lon[k]); s[c * 3 + k].idx = (uint32_t)(c * 3 + k); }
      }
    });
      sortByKey(s);
      // vertex id = number of distinct keys in front of the entry; per slice: count the run starts, prefix, assign
    const size_t m = s.size();
    const unsigned nt = numWorkers(m);
    std::vector<int64_t> starts(nt + 1, 0);
    parallelFor(m, [&](size_t a, size_t b, unsigned t) {
      int64_t c = 0;
      for (size_t i = a; i < b; ++i) c += (i == 0 || s[i].key != s[i - 1].key);
      starts[t + 1] = c;
    });
    for (unsigned t = 0; t < nt; ++t) starts[t + 1] += starts[t];
    out.numVertices = starts[nt];
    // vertex -> (column, corner) pairs, ascending caller order (the sort's tie-break); caller-order column ids
    // for now, translated to internal ids once the space-filling-curve order exists
    out.vtxOff.resize((size_t)out.numVertices + 1);
    out.vtxCols.resize(m);
    parallelFor(m, [&](size_t a, size_t b, unsigned t) {
      int64_t vid = starts[t] - 1;
      for (size_t i = a; i < b; ++i) {
        if (i == 0 || s[i].key != s[i - 1].key) { ++vid; out.vtxOff[(size_t)vid] = (int32_t)i; }
        colVert[s[i].idx] = (int32_t)vid;
        out.vtxCols[i] = (int32_t)(s[i].idx / 3);
      }
    });
    out.vtxOff[(size_t)out.numVertices] = (int32_t)m;
  }
  RawVec<int32_t> nbr(nCol * 3);
  std::vector<uint8_t> generic(nCol, 0);
  {
    // side plane k of a column spans corners (k, k+1 mod 3): p0 <-> (v1,v2), p1 <-> (v2,v3), p2 <-> (v3,v1)
    RawVec<Sortable> s(nCol * 3);
    parallelFor(nCol, [&](size_t a, size_t b, unsigned) {
      for (size_t c = a; c < b; ++c)
        for (int k = 0; k < 3; ++k) {
          uint32_t va = (uint32_t)colVert[c * 3 + k], vb = (uint32_t)colVert[c * 3 + (k + 1) % 3];
          s[c * 3 + k].key = ((uint64_t)std::min(va, vb) << 32) | std::max(va, vb);
          s[c * 3 + k].idx = (uint32_t)(c * 3 + k);
          nbr[c * 3 + k] = -1;
        }
    });
      sortByKey(s);
      // runs of equal keys = the columns on one edge; a slice handles the runs that START inside it
    const size_t m = s.size();
    auto flagGeneric = [&](size_t c) { __atomic_store_n(&generic[c], (uint8_t)1, __ATOMIC_RELAXED); };
    parallelFor(m, [&](size_t a, size_t b, unsigned) {
      size_t i = a;
      while (i < b && i > 0 && s[i].key == s[i - 1].key) ++i;
      while (i < b) {
        size_t j = i + 1;
        while (j < m && s[j].key == s[i].key) ++j;
        if (j - i == 2) {
          const uint32_t ea = s[i].idx, eb = s[i + 1].idx;
          const size_t ca = ea / 3, cb = eb / 3;
          // consistent orientation: the shared edge runs in opposite directions in the two triangles
          const bool opposite = colVert[ca * 3 + ea % 3] == colVert[cb * 3 + (eb % 3 + 1) % 3] &&
                                colVert[ca * 3 + (ea % 3 + 1) % 3] == colVert[cb * 3 + eb % 3];
          if (ca != cb && opposite) { nbr[ea] = (int32_t)cb; nbr[eb] = (int32_t)ca; }
          else { flagGeneric(ca); flagGeneric(cb); }
        } else if (j - i > 2) {
          for (size_t q = i; q < j; ++q) flagGeneric(s[q].idx / 3);   // non-manifold edge
        }
        i = j;
      }
    });
  }

  lap("4 vertex welding, adjacency");
  // ---- 5. column geometry (what the walk assumes), double unit vectors of the corners ----------------------------
  // `unbounded`: the set of points the reference predicate accepts for these records is not confined to
  // the cone over the column's triangle -- zero or non-finite plane normals (numLayers == 0 makes all three
  // planes exactly zero: ICONGrid.h:187-203 accepts the whole sphere r == height[0]; a repeated corner gives a
  // great-circle sliver; NaN passes every `> 0` rejection) or clockwise corners (the predicate then accepts
  // the ANTIPODAL triangle).  Such records go to a global list every located sample is checked against.
  // (the per-record half of these checks needs the side planes and runs in step 7, where they are computed)
  std::vector<uint8_t> unbounded(nCol, 0);
  RawVec<double> colCentroid(nCol * 3), colU(nCol * 9);
  parallelFor(nCol, [&](size_t a, size_t b, unsigned) {
    for (size_t c = a; c < b; ++c) {
      const Record &r0 = R[colFirst[c]];
      double (*v)[3] = reinterpret_cast<double (*)[3]>(&colU[c * 9]);
      for (int k = 0; k < 3; ++k) unitFromLatLon(r0.lat[k], r0.lon[k], v[k]);
      double cen[3] = {v[0][0] + v[1][0] + v[2][0], v[0][1] + v[1][1] + v[2][1], v[0][2] + v[1][2] + v[2][2]};
      const double l = std::sqrt(cen[0] * cen[0] + cen[1] * cen[1] + cen[2] * cen[2]);
      for (int k = 0; k < 3; ++k) colCentroid[c * 3 + k] = l > 0 ? cen[k] / l : (k == 2 ? 1.0 : 0.0);
      // ccw seen from outside and not degenerate
      const double ux = v[1][0] - v[0][0], uy = v[1][1] - v[0][1], uz = v[1][2] - v[0][2];
      const double wx = v[2][0] - v[0][0], wy = v[2][1] - v[0][1], wz = v[2][2] - v[0][2];
      const double nx = uy * wz - uz * wy, ny = uz * wx - ux * wz, nz = ux * wy - uy * wx;
      const double orient = nx * cen[0] + ny * cen[1] + nz * cen[2];
      const double area2 = std::sqrt(nx * nx + ny * ny + nz * nz);
      if (!(orient > 0) || !(area2 > 1e-14)) generic[c] = unbounded[c] = 1;
      if (colVert[c * 3] == colVert[c * 3 + 1] || colVert[c * 3 + 1] == colVert[c * 3 + 2] || colVert[c * 3] == colVert[c * 3 + 2]) generic[c] = unbounded[c] = 1;
      for (int k = 0; k < 3; ++k)
        if (!std::isfinite(r0.lat[k]) || !std::isfinite(r0.lon[k])) generic[c] = unbounded[c] = 1;
    }
  });

  lap("5 column sanity");
  // ---- 6. space-filling-curve order of the columns ---------------------------------------------------
  std::vector<int32_t> order(nCol), newOf(nCol);
  {
    RawVec<Sortable> s(nCol);
    parallelFor(nCol, [&](size_t a, size_t b, unsigned) {
      for (size_t c = a; c < b; ++c) {
        int face; double u, w;
        cubeProject(&colCentroid[c * 3], face, u, w);
        const uint32_t iu = (uint32_t)std::min(65535.0, std::max(0.0, (u * 0.5 + 0.5) * 65536.0));
        const uint32_t iw = (uint32_t)std::min(65535.0, std::max(0.0, (w * 0.5 + 0.5) * 65536.0));
        s[c].key = ((uint64_t)face << 32) | (part1by1(iu) | (part1by1(iw) << 1));
        s[c].idx = (uint32_t)c;
      }
    });
    sortByKey(s);
    for (size_t i = 0; i < nCol; ++i) { order[i] = (int32_t)s[i].idx; newOf[s[i].idx] = (int32_t)i; }
  }

  lap("6 curve order");
  // ---- 7. emit internal arrays; per-record side planes and bounds (icon_rt/hostCode.cu:792-808, ICONGrid.h:78-115) ------
  // The planes are computed here, straight into the record's node in internal order: a column's libm values (cosf/sinf of
  // its six angles) are taken once and serve all of its stacked records and the majorant builders' column table.
  out.rec.resize(n);
  if (withLines) out.lines.assign(n * (size_t)kLineStride, 0.f);
  else out.lines.clear();
  out.origToInt.resize(n);
  out.col.resize(nCol);
  out.colLayers.resize(nCol);
  out.colTrig.resize(nCol * kColTrigStride);
  {
    std::vector<int32_t> recBegin(nCol + 1, 0);
    for (size_t i = 0; i < nCol; ++i) recBegin[i + 1] = recBegin[i] + (colRecOff[order[i] + 1] - colRecOff[order[i]]);
    const unsigned nt = numWorkers(nCol);
    struct Acc { float vlo[3], vhi[3], slo[3], shi[3], dlo, dhi; bool nanH; int32_t maxLayers; int64_t nGeneric; };
    std::vector<Acc> acc(nt);
    for (auto &a : acc) {
      for (int k = 0; k < 3; ++k) { a.vlo[k] = a.slo[k] = INFINITY; a.vhi[k] = a.shi[k] = -INFINITY; }
      a.dlo = INFINITY; a.dhi = -INFINITY; a.nanH = false; a.maxLayers = 0; a.nGeneric = 0;
    }
    parallelFor(nCol, [&](size_t a, size_t b, unsigned tid) {
      Acc &A = acc[tid];
      for (size_t i = a; i < b; ++i) {
        const size_t c = order[i];
        ColBlock &cb = out.col[i];
        const Record &r0 = R[colFirst[c]];
        const double (*U)[3] = reinterpret_cast<const double (*)[3]>(&colU[c * 9]);
        float *T = &out.colTrig[i * kColTrigStride];
        for (int k = 0; k < 3; ++k) { T[4 * k] = cosf(r0.lat[k]); T[4 * k + 1] = sinf(r0.lat[k]); T[4 * k + 2] = cosf(r0.lon[k]); T[4 * k + 3] = sinf(r0.lon[k]); }
        for (int k = 0; k < 3; ++k) { T[12 + k] = r0.lat[k]; T[15 + k] = r0.lon[k]; }          // (the shell accel projects the raw angles)
        for (int k = 18; k < kColTrigStride; ++k) T[k] = 0.f;
        // (r,lat,lon) -> Cartesian with the reference's operand order: (r*cos(lat))*cos(lon), (r*cos(lat))*sin(lon), r*sin(lat)
        auto toCart = [&](float r, int k) { F3 p; p.x = r * T[4 * k] * T[4 * k + 2]; p.y = r * T[4 * k] * T[4 * k + 3]; p.z = r * T[4 * k + 1]; return p; };
        // prism sampler: the M row depends on the corner directions alone
        float Mrow[3];
        {
          const double ax = U[1][0] - U[0][0], ay = U[1][1] - U[0][1], az = U[1][2] - U[0][2];
          const double bx = U[2][0] - U[0][0], by = U[2][1] - U[0][1], bz = U[2][2] - U[0][2];
          const double nx = ay * bz - az * by, ny = az * bx - ax * bz, nz = ax * by - ay * bx;
          const double w = U[0][0] * nx + U[0][1] * ny + U[0][2] * nz;
          const float m[3] = {(float)(nx / w), (float)(ny / w), (float)(nz / w)};
          for (int k = 0; k < 3; ++k) Mrow[k] = std::isfinite(m[k]) ? m[k] : 0.f;
        }
        A.slo[1] = fminf(A.slo[1], fminf(r0.lat[0], fminf(r0.lat[1], r0.lat[2])));
        A.shi[1] = fmaxf(A.shi[1], fmaxf(r0.lat[0], fmaxf(r0.lat[1], r0.lat[2])));
        A.slo[2] = fminf(A.slo[2], fminf(r0.lon[0], fminf(r0.lon[1], r0.lon[2])));
        A.shi[2] = fmaxf(A.shi[2], fmaxf(r0.lon[0], fmaxf(r0.lon[1], r0.lon[2])));

        cb.recBegin = recBegin[i];
        cb.nRec = colRecOff[c + 1] - colRecOff[c];
        cb.rLo = INFINITY; cb.rHi = -INFINITY;
        // per edge: the neighbour column (internal id), its record count, and which of ITS edges faces this column
        int32_t nbCol[3], nbN[3], nbKb[3];
        for (int k = 0; k < 3; ++k) {
          const int32_t oc = nbr[c * 3 + k];      // old (caller-order) column id
          nbCol[k] = oc >= 0 ? newOf[oc] : -1;
          cb.nbr[k] = nbCol[k];
          nbN[k] = oc >= 0 ? colRecOff[oc + 1] - colRecOff[oc] : 0;
          nbKb[k] = 3;
          if (oc >= 0)
            for (int q = 0; q < 3; ++q)
              if (nbr[(size_t)oc * 3 + q] == (int32_t)c) { nbKb[k] = q; break; }
        }
        // pass 1 over the column's records: planes, prism factors, margins, bounds, and what the walk assumes of them
        // (>= 1 layer, finite non-zero planes, heights non-decreasing and finite, stacked without overlap)
        bool gen = generic[c] != 0, unb = unbounded[c] != 0;
        float prevTop = -INFINITY;
        int32_t layerBase = 0;
        for (int32_t q = 0; q < cb.nRec; ++q) {
          const int32_t o = colRecs[colRecOff[c] + q];
          const Record &rc = R[o];
          const size_t ir = (size_t)cb.recBegin + q;
          RecBlock &rb = out.rec[ir];
          std::memset(&rb, 0, sizeof(rb));
          const float h0 = rc.height[0], h1 = rc.height[rc.numLayers];
          F3 bv[3], tv[3];
          for (int k = 0; k < 3; ++k) { bv[k] = toCart(h0, k); tv[k] = toCart(h1, k); }
          planeThrough(bv[0], bv[1], tv[1], rb.p0);
          planeThrough(bv[1], bv[2], tv[2], rb.p1);
          planeThrough(bv[2], bv[0], tv[0], rb.p2);
          rb.hbot = h0; rb.htop = h1; rb.nLayers = rc.numLayers;
          {
            // prism sampler factors (closed-form inverse of the wedge map of UElems.h:214-311): double arithmetic
            // on the fp32 planes above and the corner directions, rounded once
            const float *pl[3] = {rb.p0, rb.p1, rb.p2};
            for (int k = 0; k < 3; ++k) {
              const int opp = (k + 2) % 3;      // p0 spans corners (0,1), p1 (1,2), p2 (2,0)
              const double d = U[opp][0] * (double)pl[k][0] + U[opp][1] * (double)pl[k][1] + U[opp][2] * (double)pl[k][2];
              const float kap = (float)(1.0 / d);
              rb.kappa[k] = std::isfinite(kap) ? kap : 0.f;
            }
            for (int k = 0; k < 3; ++k) rb.M[k] = Mrow[k];
          }
          double nmax = 0;
          for (const float *pl : {rb.p0, rb.p1, rb.p2}) {
            const double n2 = (double)pl[0] * pl[0] + (double)pl[1] * pl[1] + (double)pl[2] * pl[2];
            nmax = std::max(nmax, std::sqrt(n2));
            if (!(n2 > 0) || !std::isfinite(n2) || !std::isfinite(pl[3])) gen = unb = true;
          }
          rb.thr = (float)(marginMeters * nmax);
          if (!(rb.thr > 0.f) || !std::isfinite(rb.thr)) rb.thr = INFINITY;
          if (rc.numLayers < 1) gen = unb = true;
          for (int j = 0; j < rc.numLayers; ++j)
            if (!(rc.height[j] <= rc.height[j + 1])) gen = true;
          if (!std::isfinite(h0) || !std::isfinite(h1) || !(h0 > 0.f)) gen = true;
          if (!(prevTop <= h0)) gen = true;
          prevTop = h1;
          if (!std::isfinite(rb.thr)) gen = true;

          // spherical bounds / data range
          A.slo[0] = accMin(A.slo[0], h0); A.shi[0] = accMax(A.shi[0], h1);
          if (h0 != h0 || h1 != h1) A.nanH = true;
          for (int j = 0; j < rc.numLayers; ++j) { A.dlo = accMin(A.dlo, rc.value[j]); A.dhi = accMax(A.dhi, rc.value[j]); }
          // Cartesian bounds of the column: bottom corners, and top corners pushed out so the
          // flat top triangle reaches radius h1 at its barycentre
          F3 bary = {(tv[0].x + tv[1].x + tv[2].x) / 3.f, (tv[0].y + tv[1].y + tv[2].y) / 3.f, (tv[0].z + tv[1].z + tv[2].z) / 3.f};
          const float D = h1 - sqrtf(bary.x * bary.x + bary.y * bary.y + bary.z * bary.z);
          const float off = D / h1;
          for (int k = 0; k < 3; ++k) {
            F3 t = tv[k];
            t.x = t.x + t.x * off; t.y = t.y + t.y * off; t.z = t.z + t.z * off;
            const float lo[3] = {fminf(bv[k].x, t.x), fminf(bv[k].y, t.y), fminf(bv[k].z, t.z)};
            const float hi[3] = {fmaxf(bv[k].x, t.x), fmaxf(bv[k].y, t.y), fmaxf(bv[k].z, t.z)};
            for (int w = 0; w < 3; ++w) { A.vlo[w] = accMin(A.vlo[w], lo[w]); A.vhi[w] = accMax(A.vhi[w], hi[w]); }
          }

          rb.orig = o;
          rb.col = (int32_t)i;
          for (int k = 0; k < 3; ++k) rb.vert[k] = colVert[c * 3 + k];
          rb.layerBase = layerBase;
          layerBase += rc.numLayers;
          out.origToInt[o] = (int32_t)ir;
          if (withLines) {
            std::memcpy(&out.lines[ir * kLineStride], rc.height, sizeof(float) * 32);
            std::memcpy(&out.lines[ir * kLineStride + kHeightStride], rc.value, sizeof(float) * 32);
          }
          cb.rLo = accMin(cb.rLo, h0);
          cb.rHi = accMax(cb.rHi, h1);
        }
        if (cb.nRec > 64) gen = true;
        generic[c] = gen; unbounded[c] = unb;
        cb.flags = gen ? COL_GENERIC : 0;
        A.nGeneric += gen;
        // pass 2: links and flags (they depend on the column's verdict)
        for (int32_t q = 0; q < cb.nRec; ++q) {
          RecBlock &rb = out.rec[(size_t)cb.recBegin + q];
          if (gen) rb.thr = INFINITY;
          int32_t info = (q > 255 ? 255 : q) | ((cb.nRec > 255 ? 255 : cb.nRec) << REC_NREC_SHIFT);
          if (gen || cb.nRec > 255) info |= REC_GENERIC;
          for (int k = 0; k < 3; ++k) {
            rb.nbr[k] = nbCol[k] >= 0 ? recBegin[nbCol[k]] + std::min(q, nbN[k] - 1) : -1;
            info |= nbKb[k] << (REC_KB_SHIFT + 2 * k);
            if (nbCol[k] >= 0 && q < nbN[k]) info |= 1 << (REC_SAMEK_SHIFT + k);
          }
          info |= (rb.nLayers & 31) << REC_NL_SHIFT;
          rb.info = info;
        }
        out.colLayers[i] = layerBase;
        A.maxLayers = std::max(A.maxLayers, layerBase);
      }
    });
    icb_bounds &B = out.bounds;
    for (int k = 0; k < 3; ++k) { B.volLo[k] = B.sphLo[k] = INFINITY; B.volHi[k] = B.sphHi[k] = -INFINITY; }
    B.dataRange[0] = INFINITY; B.dataRange[1] = -INFINITY;
    bool nanH = false;
    for (auto &a : acc) {
      for (int k = 0; k < 3; ++k) {
        B.volLo[k] = fminf(B.volLo[k], a.vlo[k]); B.volHi[k] = fmaxf(B.volHi[k], a.vhi[k]);
        B.sphLo[k] = fminf(B.sphLo[k], a.slo[k]); B.sphHi[k] = fmaxf(B.sphHi[k], a.shi[k]);
      }
      B.dataRange[0] = fminf(B.dataRange[0], a.dlo); B.dataRange[1] = fmaxf(B.dataRange[1], a.dhi);
      nanH = nanH || a.nanH;
      out.numGeneric += a.nGeneric;
      out.maxColLayers = std::max(out.maxColLayers, a.maxLayers);
    }
    out.cullLo = nanH ? -INFINITY : B.sphLo[0];
    out.cullHi = nanH ? INFINITY : B.sphHi[0];
    // icon_rt/hostCode.cu:838-840
    const float magnitude = floorf(log10f(B.sphLo[0]));
    B.unitDistance = 1.0f * powf(10.f, magnitude - 3);

    parallelFor(out.vtxCols.size(), [&](size_t a, size_t b, unsigned) {
      for (size_t i = a; i < b; ++i) out.vtxCols[i] = newOf[out.vtxCols[i]];
    });
    // global records, ascending caller index (the scan stops at the first index above the current answer)
    std::vector<int32_t> globOrig;
    for (size_t c = 0; c < nCol; ++c)
      if (unbounded[c])
        for (int32_t q = colRecOff[c]; q < colRecOff[c + 1]; ++q) globOrig.push_back(colRecs[q]);
    std::sort(globOrig.begin(), globOrig.end());
    out.globalRecs.resize(globOrig.size());
    for (size_t i = 0; i < globOrig.size(); ++i) out.globalRecs[i] = out.origToInt[globOrig[i]];
  }

  lap("7 emit arrays");
  // ---- 8. cube-map candidate bins (conservative) --------------------------------------------------------------
  {
    int32_t N = (int32_t)std::lround(std::sqrt((double)nCol / 6.0) * kBinsPerColumnEdge);
    N = std::max(1, std::min(4096, N));
    out.binN = N;
    const size_t nBins = (size_t)6 * N * N;
    const double pad = 2e-5;
    // pass 1 counts, pass 2 fills; per column the covered (face, iu0..iu1, iw0..iw1) boxes are recomputed
    auto boxes = [&](size_t i, int faceBox[6][4]) -> void {
      const size_t c = order[i];
      if (unbounded[c]) {   // reached through the global list instead
        for (int f = 0; f < 6; ++f) { faceBox[f][0] = 0; faceBox[f][1] = -1; faceBox[f][2] = 0; faceBox[f][3] = -1; }
        return;
      }
      const double (*v)[3] = reinterpret_cast<const double (*)[3]>(&colU[c * 9]);
      double extent = 0;
      for (int a = 0; a < 3; ++a)
        for (int b = a + 1; b < 3; ++b) {
          double d = v[a][0] * v[b][0] + v[a][1] * v[b][1] + v[a][2] * v[b][2];
          extent = std::max(extent, std::acos(std::max(-1.0, std::min(1.0, d))));
        }
      for (int f = 0; f < 6; ++f) {
        const int ax = f / 2, sgn = (f % 2) ? -1 : 1;
        const int ua = ax == 0 ? 1 : 0, wa = ax == 2 ? 1 : 2;
        double m[3];
        bool allPos = true, anyFar = false;
        for (int k = 0; k < 3; ++k) { m[k] = sgn * v[k][ax]; if (m[k] <= 0.15) allPos = false; if (m[k] > 0.5) anyFar = true; }
        faceBox[f][0] = 0; faceBox[f][1] = -1; faceBox[f][2] = 0; faceBox[f][3] = -1;   // empty
        if (!allPos) {
          if (extent < 0.3 && !anyFar) continue;                 // cannot reach |u|,|w| <= 1 on this face
          if (extent < 0.3) {
            // small triangle with a corner beyond the face's useful hemisphere cannot exist (m<=0.15 and m>0.5
            // are > 0.3 rad apart), so this is unreachable; keep the conservative answer anyway
          }
          faceBox[f][0] = 0; faceBox[f][1] = N - 1; faceBox[f][2] = 0; faceBox[f][3] = N - 1;   // whole face
          continue;
        }
        double ulo = 1e30, uhi = -1e30, wlo = 1e30, whi = -1e30;
        for (int k = 0; k < 3; ++k) {
          const double u = v[k][ua] / m[k], w = v[k][wa] / m[k];
          ulo = std::min(ulo, u); uhi = std::max(uhi, u); wlo = std::min(wlo, w); whi = std::max(whi, w);
        }
        ulo -= pad; uhi += pad; wlo -= pad; whi += pad;
        if (uhi < -1.0 - pad || ulo > 1.0 + pad || whi < -1.0 - pad || wlo > 1.0 + pad) continue;
        auto toBin = [&](double t) { return (int)std::min((double)(N - 1), std::max(0.0, std::floor((t * 0.5 + 0.5) * N))); };
        // `pad` (2e-5 in u) exceeds both the predicate's noise band (~1e-6 rad) and the device's fp32
        // rounding of (u*0.5+0.5)*N (<= N*2e-7 bins), so the covered bin range is conservative
        faceBox[f][0] = toBin(ulo); faceBox[f][1] = toBin(uhi);
        faceBox[f][2] = toBin(wlo); faceBox[f][3] = toBin(whi);
      }
    };
    std::vector<std::atomic<int32_t>> cnt(nBins);
    for (auto &c : cnt) c.store(0, std::memory_order_relaxed);
    parallelFor(nCol, [&](size_t a, size_t b, unsigned) {
      int fb[6][4];
      for (size_t i = a; i < b; ++i) {
        boxes(i, fb);
        for (int f = 0; f < 6; ++f)
          for (int w = fb[f][2]; w <= fb[f][3]; ++w)
            for (int u = fb[f][0]; u <= fb[f][1]; ++u)
              cnt[((size_t)f * N + w) * N + u].fetch_add(1, std::memory_order_relaxed);
      }
    });
    out.binOff.resize(nBins + 1);
    int64_t total = 0;
    for (size_t bI = 0; bI < nBins; ++bI) { out.binOff[bI] = (int32_t)total; total += cnt[bI].load(std::memory_order_relaxed); }
    if (total > std::numeric_limits<int32_t>::max()) { err = "icb_set_grid: bin lists overflow int32"; return ICB_ERR_INVALID; }
    out.binOff[nBins] = (int32_t)total;
    out.binList.resize((size_t)total);
    // fill in parallel (a bin's entries land in arrival order), then each bin's short list is sorted: ascending bottom
    // record = ascending internal column id, the order a serial fill over the columns would have produced
    for (auto &c : cnt) c.store(0, std::memory_order_relaxed);
    parallelFor(nCol, [&](size_t a, size_t b, unsigned) {
      int fb[6][4];
      for (size_t i = a; i < b; ++i) {
        boxes(i, fb);
        for (int f = 0; f < 6; ++f)
          for (int w = fb[f][2]; w <= fb[f][3]; ++w)
            for (int u = fb[f][0]; u <= fb[f][1]; ++u) {
              const size_t bI = ((size_t)f * N + w) * N + u;
              out.binList[(size_t)out.binOff[bI] + cnt[bI].fetch_add(1, std::memory_order_relaxed)] = out.col[i].recBegin;   // the column's bottom record
            }
      }
    });
    parallelFor(nBins, [&](size_t a, size_t b, unsigned) {
      for (size_t bI = a; bI < b; ++bI)
        if (out.binOff[bI + 1] - out.binOff[bI] > 1) std::sort(out.binList.begin() + out.binOff[bI], out.binList.begin() + out.binOff[bI + 1]);
    });
    lap("8 candidate bins");
    // ---- 9. walk seeds: per bin, the well-formed column whose centroid is nearest to the bin's centre -----------
    out.seed.assign(nBins * 2, -1);
    parallelFor(nBins, [&](size_t a, size_t b, unsigned) {
      for (size_t bI = a; bI < b; ++bI) {
        const int f = (int)(bI / ((size_t)N * N)), w = (int)((bI / N) % N), u = (int)(bI % N);
        const double cu = ((u + 0.5) / N) * 2.0 - 1.0, cw = ((w + 0.5) / N) * 2.0 - 1.0, sg = (f & 1) ? -1.0 : 1.0;
        double d[3];
        if (f < 2)      { d[0] = sg; d[1] = cu; d[2] = cw; }
        else if (f < 4) { d[0] = cu; d[1] = sg; d[2] = cw; }
        else            { d[0] = cu; d[1] = cw; d[2] = sg; }
        double bestDot = -2.0;
        const double dl = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        for (int32_t j = out.binOff[bI]; j < out.binOff[bI + 1]; ++j) {
          const int32_t r0 = out.binList[j];
          const int32_t ci = out.rec[r0].col;
          if (out.col[ci].flags & COL_GENERIC) continue;
          const double *cen = &colCentroid[(size_t)order[ci] * 3];
          const double dt = (cen[0] * d[0] + cen[1] * d[1] + cen[2] * d[2]) / dl;
          if (dt > bestDot) { bestDot = dt; out.seed[bI * 2] = r0; out.seed[bI * 2 + 1] = r0 + out.col[ci].nRec - 1; }
        }
      }
    });
  }
  lap("9 walk seeds");
  return ICB_OK;
}

}  // namespace icb
