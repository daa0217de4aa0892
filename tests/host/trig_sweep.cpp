// Sweep of csrc/fdlibm_trig.cuh (compiled for the host, -ffp-contract=off) against the host's libm: the operation
// sequence the device runs must give glibc's asinf/atanf/atan2f bit for bit.  Usage: trig_sweep <stride> <pairs> [threads]
//   stride  every stride-th float bit pattern for the one-argument functions (1 = all 2^32)
//   pairs   random (y,x) pairs for atan2f, plus a fixed grid of special values
// Prints "asinf <tested> <mismatches>" etc.; exit code 1 on any mismatch.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>

#include "fdlibm_trig.cuh"

static inline bool sameBits(float a, float b)
{
  if (std::isnan(a) && std::isnan(b)) return true;
  return icbF2U(a) == icbF2U(b);
}

int main(int argc, char **argv)
{
  const uint64_t stride = argc > 1 ? strtoull(argv[1], nullptr, 10) : 1;
  const uint64_t pairs = argc > 2 ? strtoull(argv[2], nullptr, 10) : 100000000ull;
  const int nt = argc > 3 ? atoi(argv[3]) : (int)std::thread::hardware_concurrency();
  std::vector<uint64_t> bad((size_t)nt * 3, 0), cnt((size_t)nt * 3, 0);
  std::vector<std::thread> th;
  for (int t = 0; t < nt; ++t)
    th.emplace_back([&, t] {
      for (uint64_t u = (uint64_t)t * stride; u < (1ull << 32); u += stride * (uint64_t)nt) {
        const float x = icbU2F((uint32_t)u);
        volatile float vx = x;                      // (keeps the compiler from folding libm calls)
        if (!sameBits(icb::fdlibm::asinF(x), asinf(vx))) { if (bad[t * 3]++ < 3) fprintf(stderr, "asinf %08x\n", (unsigned)u); }
        if (!sameBits(icb::fdlibm::atanF(x), atanf(vx))) { if (bad[t * 3 + 1]++ < 3) fprintf(stderr, "atanf %08x\n", (unsigned)u); }
        cnt[t * 3] += 1; cnt[t * 3 + 1] += 1;
      }
      uint64_t s = 0x9E3779B97F4A7C15ull * (uint64_t)(t + 1);
      auto next = [&]() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; };
      for (uint64_t i = (uint64_t)t; i < pairs; i += (uint64_t)nt) {
        const uint64_t r = next();
        float y, x;
        switch (i & 3) {
          case 0: y = icbU2F((uint32_t)r); x = icbU2F((uint32_t)(r >> 32)); break;                      // any two bit patterns
          case 1: {                                                                                        // points of a planet-sized shell
            y = (float)((double)(int32_t)(uint32_t)r * (6.5e6 / 2147483648.0));
            x = (float)((double)(int32_t)(uint32_t)(r >> 32) * (6.5e6 / 2147483648.0));
            break;
          }
          case 2: y = icbU2F((uint32_t)r); x = icbU2F((uint32_t)r + (uint32_t)((r >> 32) % 2000001u) - 1000000u); break;   // |y/x| near 1
          default: y = (float)((double)(int32_t)(uint32_t)r / 2147483648.0); x = (float)((double)(int32_t)(uint32_t)(r >> 32) / 2147483648.0); break;
        }
        volatile float vy = y, vx = x;
        if (!sameBits(icb::fdlibm::atan2F(y, x), atan2f(vy, vx))) { if (bad[t * 3 + 2]++ < 3) fprintf(stderr, "atan2f %08x %08x\n", icbF2U(y), icbF2U(x)); }
        cnt[t * 3 + 2] += 1;
      }
    });
  for (auto &t : th) t.join();
  uint64_t b[3] = {0, 0, 0}, c[3] = {0, 0, 0};
  for (int t = 0; t < nt; ++t) for (int k = 0; k < 3; ++k) { b[k] += bad[t * 3 + k]; c[k] += cnt[t * 3 + k]; }
  // special values for atan2f: every pair of a fixed list
  const float sp[] = {0.f, -0.f, 1.f, -1.f, INFINITY, -INFINITY, NAN, 1e-45f, -1e-45f, 1.17549435e-38f, 3.4028235e38f, -3.4028235e38f,
                      0.5f, 2.f, 1e-30f, 1e30f, -1e-30f, -1e30f, 6.371e6f, -6.371e6f, 1e-20f, 1e20f};
  for (float y : sp) for (float x : sp) {
    volatile float vy = y, vx = x;
    if (!sameBits(icb::fdlibm::atan2F(y, x), atan2f(vy, vx))) { ++b[2]; fprintf(stderr, "atan2f special %g %g\n", y, x); }
    ++c[2];
  }
  printf("asinf %llu %llu\natanf %llu %llu\natan2f %llu %llu\n", (unsigned long long)c[0], (unsigned long long)b[0], (unsigned long long)c[1],
         (unsigned long long)b[1], (unsigned long long)c[2], (unsigned long long)b[2]);
  return (b[0] | b[1] | b[2]) ? 1 : 0;
}
