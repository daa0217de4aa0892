// Host grid builder: timing and array checksums (no GPU).  Generates an RnBk x L grid with the synthetic generator, optionally
// shuffles the records (mode 1) and plants degenerate ones (mode 2: zero layers, NaN / repeated / clockwise corners, unsorted and
// overlapping heights, duplicates), runs icb::buildGrid and prints an FNV-1a checksum of every array it produces.
// tests/test_builder_pins.py compares that output with tests/golden/builder_checksums.txt (written by round 1's builder).
//   g++ -std=c++17 -O2 -ffp-contract=off -Iinclude -Iicon-ray-tracing_b200/csrc tools/time_build.cpp \
//       icon-ray-tracing_b200/build/{grid_build,gen_grid}.o -o time_build -lpthread;  ICB_BUILD_TIMING=1 ./time_build 2 8 90 [mode] [L]
#include "grid_build.h"
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
extern "C" int64_t icb_gen_num_records(int,int,int);
extern "C" int icb_gen_grid(int n, int k, int levels, float exaggeration, int timestep, void *records, int64_t capacity);
#pragma pack(push, 1)
struct Rec { float lat[3], lon[3]; int32_t nl; float h[32], v[32]; };
#pragma pack(pop)
int main(int argc, char **argv) {
  int n = atoi(argv[1]), k = atoi(argv[2]), L = atoi(argv[3]);
  int mode = argc > 4 ? atoi(argv[4]) : 0;   // 1: shuffle records, 2: shuffle + degenerate records
  int64_t cnt = icb_gen_num_records(n, k, L);
  fprintf(stderr, "count %ld\n", (long)cnt);
  std::vector<Rec> rec((size_t)cnt);
  auto t0 = std::chrono::steady_clock::now();
  icb_gen_grid(n, k, L, 10.f, 0, rec.data(), cnt);
  if (mode >= 1) { std::mt19937_64 g(7); std::shuffle(rec.begin(), rec.end(), g); }
  if (mode >= 2) {
    std::mt19937_64 g(11);
    for (int q = 0; q < 40; ++q) {
      Rec &r = rec[g() % cnt];
      switch (q % 8) {
        case 0: r.nl = 0; break;
        case 1: r.lat[1] = NAN; break;
        case 2: std::swap(r.lat[0], r.lat[1]); std::swap(r.lon[0], r.lon[1]); break;   // clockwise
        case 3: r.lat[1] = r.lat[0]; r.lon[1] = r.lon[0]; break;                        // repeated corner
        case 4: r.h[3] = r.h[1]; break;                                                 // unsorted heights
        case 5: r.h[0] = NAN; break;
        case 6: r.h[r.nl] = r.h[r.nl] + 5000.f; break;                                  // overlap with the record above
        case 7: rec[g() % cnt] = r; break;                                              // exact duplicate record
      }
    }
  }
  auto t1 = std::chrono::steady_clock::now();
  fprintf(stderr, "gen %.2f s\n", std::chrono::duration<double>(t1 - t0).count());
  icb::HostGrid g; std::string err;
  int rc = icb::buildGrid(rec.data(), cnt, 16.f, g, err, argc > 5);
  auto t2 = std::chrono::steady_clock::now();
  fprintf(stderr, "build rc=%d %.2f s\n", rc, std::chrono::duration<double>(t2 - t1).count());
  auto h = [](const void *p, size_t n) { uint64_t x = 1469598103934665603ull; const unsigned char *c = (const unsigned char *)p; for (size_t i = 0; i < n; ++i) { x ^= c[i]; x *= 1099511628211ull; } return x; };
  printf("rec %016lx col %016lx binOff %016lx binList %016lx seed %016lx o2i %016lx vtxOff %016lx vtxCols %016lx colLayers %016lx colTrig %016lx glob %016lx lines %016lx bounds %016lx\n",
    h(g.rec.data(), g.rec.size()*sizeof(g.rec[0])), h(g.col.data(), g.col.size()*sizeof(g.col[0])), h(g.binOff.data(), g.binOff.size()*4), h(g.binList.data(), g.binList.size()*4), h(g.seed.data(), g.seed.size()*4), h(g.origToInt.data(), g.origToInt.size()*4), h(g.vtxOff.data(), g.vtxOff.size()*4), h(g.vtxCols.data(), g.vtxCols.size()*4), h(g.colLayers.data(), g.colLayers.size()*4), h(g.colTrig.data(), g.colTrig.size()*4), h(g.globalRecs.data(), g.globalRecs.size()*4), h(g.lines.data(), g.lines.size()*4), h(&g.bounds, sizeof g.bounds));
  printf("nv %ld ng %ld nglob %zu binN %d maxCL %d cull %a %a\n", (long)g.numVertices, (long)g.numGeneric, g.globalRecs.size(), g.binN, g.maxColLayers, g.cullLo, g.cullHi);
}
