// Synthetic ICON-like grid generator (replaces tools/convert_icon for this repo: no
// netCDF, no DWD data, no network -- SURVEY.md §2.1 #18, §8d "Synthetic grid").
//
// Output format is the reference's `.ic` record stream exactly as
// tools/convert_icon/convert_icon.cpp:383-387 writes it and icon_rt/hostCode.cu:717-734
// reads it: packed 284-byte records {float lat[3]; float lon[3]; int32 numLayers;
// float height[32]; float value[32]} (icon_rt/ICONGrid.h:59-76), corners ccw seen
// from outside (fixture icon_rt/hostCode.cu:771-777), <=31 layers per record
// (convert_icon.cpp:351-388), stacked bottom-up per column.
//
// RnBk: icosahedron, every edge split into n parts, then k bisections, vertices
// pushed to the unit sphere -> 20*n^2*4^k columns.  lat/lon are computed in double
// and cast to float (convert_icon.cpp:357-358); R = 6.371229e6 (convert_icon.cpp:359).
#include "icon_b200.h"

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

namespace {

struct D3 { double x, y, z; };
inline D3 add(D3 a, D3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline D3 sub(D3 a, D3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline D3 mul(D3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
inline double dot(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline D3 cross(D3 a, D3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline D3 unit(D3 a) { double l = std::sqrt(dot(a, a)); return {a.x / l, a.y / l, a.z / l}; }

struct Tri { D3 a, b, c; };

// midpoint on the sphere; a+b is commutative in IEEE arithmetic, so both triangles
// sharing the edge get bit-identical vertices whatever the traversal order
inline D3 mid(D3 a, D3 b) { return unit(add(a, b)); }

void icosahedron(std::vector<Tri> &out)
{
  const double t = (1.0 + std::sqrt(5.0)) / 2.0;
  D3 v[12] = {{-1, t, 0}, {1, t, 0}, {-1, -t, 0}, {1, -t, 0}, {0, -1, t}, {0, 1, t},
              {0, -1, -t}, {0, 1, -t}, {t, 0, -1}, {t, 0, 1}, {-t, 0, -1}, {-t, 0, 1}};
  for (auto &p : v) p = unit(p);
  static const int f[20][3] = {{0, 11, 5}, {0, 5, 1}, {0, 1, 7}, {0, 7, 10}, {0, 10, 11},
                               {1, 5, 9}, {5, 11, 4}, {11, 10, 2}, {10, 7, 6}, {7, 1, 8},
                               {3, 9, 4}, {3, 4, 2}, {3, 2, 6}, {3, 6, 8}, {3, 8, 9},
                               {4, 9, 5}, {2, 4, 11}, {6, 2, 10}, {8, 6, 7}, {9, 8, 1}};
  for (auto &i : f) {
    Tri tr{v[i[0]], v[i[1]], v[i[2]]};
    // ccw as seen from outside: normal of (b-a)x(c-a) points away from the origin
    if (dot(cross(sub(tr.b, tr.a), sub(tr.c, tr.a)), add(add(tr.a, tr.b), tr.c)) < 0) std::swap(tr.b, tr.c);
    out.push_back(tr);
  }
}

// split every edge of a root face into n parts (barycentric lattice, then normalise).
// lattice point (i,j) = (n-i-j)*a + i*b + j*c; integer weights keep shared-edge points
// bit-identical between neighbouring faces (the zero-weight corner contributes exact 0).
void rootDivide(const Tri &t, int n, std::vector<Tri> &out)
{
  auto P = [&](int i, int j) {
    int k = n - i - j;
    // sum the three terms in an order independent of which face we are in: sort by weight owner
    D3 ta = mul(t.a, (double)k), tb = mul(t.b, (double)i), tc = mul(t.c, (double)j);
    // terms with weight 0 are exact zeros; for edge points exactly two terms are non-zero and
    // their sum is commutative.  Interior points belong to one face only.
    D3 s;
    if (k == 0) s = add(tb, tc);
    else if (i == 0) s = add(ta, tc);
    else if (j == 0) s = add(ta, tb);
    else s = add(add(ta, tb), tc);
    return unit(s);
  };
  for (int j = 0; j < n; ++j)
    for (int i = 0; i < n - j; ++i) {
      out.push_back({P(i, j), P(i + 1, j), P(i, j + 1)});
      if (i + j < n - 1) out.push_back({P(i + 1, j), P(i + 1, j + 1), P(i, j + 1)});
    }
}

void bisect(const std::vector<Tri> &in, std::vector<Tri> &out)
{
  out.resize(in.size() * 4);
  auto work = [&](size_t a, size_t b) {
    for (size_t q = a; q < b; ++q) {
      const Tri &t = in[q];
      D3 ab = mid(t.a, t.b), bc = mid(t.b, t.c), ca = mid(t.c, t.a);
      out[4 * q + 0] = {t.a, ab, ca};
      out[4 * q + 1] = {ab, t.b, bc};
      out[4 * q + 2] = {ca, bc, t.c};
      out[4 * q + 3] = {ab, bc, ca};
    }
  };
  unsigned nt = std::max(1u, std::min(64u, std::thread::hardware_concurrency()));
  if (in.size() < 4096) nt = 1;
  if (nt == 1) { work(0, in.size()); return; }
  std::vector<std::thread> th;
  for (unsigned i = 0; i < nt; ++i) th.emplace_back(work, in.size() * i / nt, in.size() * (i + 1) / nt);
  for (auto &t : th) t.join();
}

inline void latlon(D3 p, double &lat, double &lon)
{
  double z = std::min(1.0, std::max(-1.0, p.z));
  lat = std::asin(z);
  double y = p.y == 0.0 ? 0.0 : p.y;   // -0 -> +0 so the +-pi seam has one representation
  lon = (std::fabs(p.x) < 1e-300 && y == 0.0) ? 0.0 : std::atan2(y, p.x);
}

#pragma pack(push, 1)
struct Record {
  float   lat[3], lon[3];
  int32_t numLayers;
  float   height[32], value[32];
};
#pragma pack(pop)
static_assert(sizeof(Record) == 284, "ICONCell record must be 284 bytes");

inline int recordsPerColumn(int levels) { return (levels + 30) / 31; }

} // namespace

extern "C" {

int64_t icb_gen_num_columns(int n, int k)
{
  if (n < 1 || k < 0 || k > 12) return -1;
  int64_t c = 20ll * n * n;
  for (int i = 0; i < k; ++i) c *= 4;
  return c;
}

int64_t icb_gen_num_records(int n, int k, int levels)
{
  int64_t c = icb_gen_num_columns(n, k);
  if (c < 0 || levels < 1) return -1;
  return c * recordsPerColumn(levels);
}

// h_j = R + E*(hsurf + (Htop-hsurf)*(j/L)^2), hsurf = 2000*max(0,sin 3lat*cos 2lon)
// at the cell centre; T = 0.5 + 0.35*cos(2lat)*(1-z/Htop) + 0.15*sin(3(lon+2pi*s/24)+4lat)*exp(-z/2e4),
// clamped to [0,1], z = un-exaggerated mid-height of the layer.
} // extern "C"

namespace {

void buildMesh(int n, int k, std::vector<Tri> &cur)
{
  std::vector<Tri> base, nxt;
  icosahedron(base);
  for (const Tri &t : base) rootDivide(t, n, cur);
  for (int i = 0; i < k; ++i) { bisect(cur, nxt); cur.swap(nxt); }
}

constexpr double kR = 6.371229e6, kHtop = 75000.0;

// a column's centre, surface height and (optionally) its corner angles
inline void columnGeometry(const Tri &t, double &clat, double &clon, double &hsurf, double *la, double *lo)
{
  if (la) { latlon(t.a, la[0], lo[0]); latlon(t.b, la[1], lo[1]); latlon(t.c, la[2], lo[2]); }
  latlon(unit(add(add(t.a, t.b), t.c)), clat, clon);
  hsurf = 2000.0 * std::max(0.0, std::sin(3 * clat) * std::cos(2 * clon));
}

inline void columnValues(double clat, double clon, double hsurf, int L, double lonShift, double *val)
{
  for (int j = 0; j < L; ++j) {
    double f = (j + 0.5) / L;
    double z = hsurf + (kHtop - hsurf) * f * f;
    double T = 0.5 + 0.35 * std::cos(2 * clat) * (1.0 - z / kHtop)
             + 0.15 * std::sin(3 * (clon + lonShift) + 4 * clat) * std::exp(-z / 2.0e4);
    val[j] = std::min(1.0, std::max(0.0, T));
  }
}

template <typename F>
void forColumns(int64_t c0, int64_t c1, F work)
{
  unsigned nt = std::max(1u, std::min(64u, std::thread::hardware_concurrency()));
  if (c1 - c0 < 4096) nt = 1;
  std::vector<std::thread> th;
  for (unsigned i = 0; i < nt; ++i)
    th.emplace_back(work, c0 + (c1 - c0) * i / nt, c0 + (c1 - c0) * (i + 1) / nt);
  for (auto &t : th) t.join();
}

} // namespace

extern "C" {

int icb_gen_grid(int n, int k, int levels, float exaggeration, int timestep, void *records, int64_t capacity)
{
  const int64_t ncol = icb_gen_num_columns(n, k);
  const int64_t nrec = icb_gen_num_records(n, k, levels);
  if (ncol < 0 || nrec < 0 || !records || capacity < nrec) return ICB_ERR_INVALID;

  std::vector<Tri> cur;
  buildMesh(n, k, cur);

  const double E = exaggeration;
  const int L = levels, rpc = recordsPerColumn(levels);
  const double lonShift = 2.0 * M_PI * (double)timestep / 24.0;
  Record *rec = (Record *)records;

  forColumns(0, ncol, [&](int64_t c0, int64_t c1) {
    std::vector<double> h(L + 1), val(L);
    for (int64_t c = c0; c < c1; ++c) {
      double la[3], lo[3], clat, clon, hsurf;
      columnGeometry(cur[c], clat, clon, hsurf, la, lo);
      for (int j = 0; j <= L; ++j) {
        double f = (double)j / L;
        h[j] = kR + E * (hsurf + (kHtop - hsurf) * f * f);
      }
      columnValues(clat, clon, hsurf, L, lonShift, val.data());
      for (int r = 0; r < rpc; ++r) {
        Record &o = rec[c * rpc + r];
        std::memset(&o, 0, sizeof o);
        for (int q = 0; q < 3; ++q) { o.lat[q] = (float)la[q]; o.lon[q] = (float)lo[q]; }
        const int j0 = r * 31, nl = std::min(31, L - j0);
        o.numLayers = nl;
        for (int j = 0; j <= nl; ++j) o.height[j] = (float)h[j0 + j];
        for (int j = 0; j < nl; ++j) o.value[j] = (float)val[j0 + j];
      }
    }
  });
  return ICB_OK;
}

// The scalars alone (ICONCell::value, 32 floats per record, icb_set_scalars layout) of records [firstRecord,
// firstRecord + numRecords) for ntimesteps consecutive timesteps: values[t][numRecords * 32].  What a rank of the time-series
// workload holds in pinned memory: its slab of every timestep, bit-identical to the value arrays of icb_gen_grid.
int icb_gen_values(int n, int k, int levels, int timestep0, int ntimesteps, int64_t firstRecord, int64_t numRecords, float *values)
{
  const int64_t ncol = icb_gen_num_columns(n, k);
  const int64_t nrec = icb_gen_num_records(n, k, levels);
  if (ncol < 0 || nrec < 0 || !values || ntimesteps < 1 || firstRecord < 0 || numRecords < 0 || firstRecord + numRecords > nrec) return ICB_ERR_INVALID;
  if (numRecords == 0) return ICB_OK;
  std::vector<Tri> cur;
  buildMesh(n, k, cur);
  const int L = levels, rpc = recordsPerColumn(levels);
  const int64_t c0 = firstRecord / rpc, c1 = (firstRecord + numRecords + rpc - 1) / rpc;
  forColumns(c0, c1, [&](int64_t a, int64_t b) {
    std::vector<double> val(L);
    for (int64_t c = a; c < b; ++c) {
      double clat, clon, hsurf;
      columnGeometry(cur[c], clat, clon, hsurf, nullptr, nullptr);
      for (int t = 0; t < ntimesteps; ++t) {
        columnValues(clat, clon, hsurf, L, 2.0 * M_PI * (double)(timestep0 + t) / 24.0, val.data());
        for (int r = 0; r < rpc; ++r) {
          const int64_t ir = c * rpc + r;
          if (ir < firstRecord || ir >= firstRecord + numRecords) continue;
          float *o = values + ((size_t)t * (size_t)numRecords + (size_t)(ir - firstRecord)) * 32;
          const int j0 = r * 31, nl = std::min(31, L - j0);
          for (int j = 0; j < 32; ++j) o[j] = j < nl ? (float)val[j0 + j] : 0.f;
        }
      }
    }
  });
  return ICB_OK;
}

} // extern "C"
