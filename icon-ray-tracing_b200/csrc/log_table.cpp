// logf(1 - u) for every value u = k / 2^24 the reference's LCG can return (common/dvr_course-common-both.h:80-87),
// tabulated on the HOST with the libm the reference's CPU path itself calls (woodcockTracking, deviceCode.cu:163).
// CUDA's logf and glibc's differ in the last place for a share of the arguments; a table makes the free-flight
// distances of the Woodcock integrator -- and with them every sampled position -- bit-identical to the reference.
#include "grid_build.h"

#include <cmath>
#include <thread>

namespace icb {

void buildLogTable(std::vector<float> &table)
{
  const size_t n = (size_t)1 << 24;
  table.resize(n);
  unsigned nt = std::max(1u, std::min(32u, std::thread::hardware_concurrency()));
  std::vector<std::thread> th;
  for (unsigned t = 0; t < nt; ++t)
    th.emplace_back([&, t] {
      for (size_t k = n * t / nt; k < n * (t + 1) / nt; ++k) {
        const float u = (float)k / (float)0x01000000;
        table[k] = logf(1.f - u);
      }
    });
  for (auto &t : th) t.join();
}

}  // namespace icb
