// Sweep of csrc/glibc_logf.cuh (compiled for the host, -ffp-contract=off) against the host's libm logf: (a) the 2^24
// arguments 1 - k/2^24 the Woodcock ray-gens evaluate, (b) every stride-th positive normal float.
// Prints "<set> <tested> <mismatches fma form> <mismatches sse2 form>"; exit code 1 when NEITHER form reproduces libm
// on all of (a).   Usage: logf_sweep <stride> [threads]
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>

#include "glibc_logf.cuh"

int main(int argc, char **argv)
{
  const uint64_t stride = argc > 1 ? strtoull(argv[1], nullptr, 10) : 1;
  const int nt = argc > 2 ? atoi(argv[2]) : (int)std::thread::hardware_concurrency();
  std::vector<uint64_t> bad((size_t)nt * 4, 0), cnt((size_t)nt * 2, 0);
  std::vector<std::thread> th;
  for (int t = 0; t < nt; ++t)
    th.emplace_back([&, t] {
      for (uint32_t k = (uint32_t)t; k < (1u << 24); k += (uint32_t)nt) {
        volatile float x = 1.f - (float)k / (float)0x01000000;
        const float want = logf(x);
        if (icbLogfF2U(icb::glibc::logOneMinus<true>(k)) != icbLogfF2U(want)) ++bad[t * 4];
        if (icbLogfF2U(icb::glibc::logOneMinus<false>(k)) != icbLogfF2U(want)) ++bad[t * 4 + 1];
        ++cnt[t * 2];
      }
      for (uint64_t u = 0x00800000ull + (uint64_t)t * stride; u < 0x7f800000ull; u += stride * (uint64_t)nt) {
        volatile float x = icbLogfU2F((uint32_t)u);
        const float want = logf(x);
        if (icbLogfF2U(icb::glibc::logfNormal<true>(x)) != icbLogfF2U(want)) ++bad[t * 4 + 2];
        if (icbLogfF2U(icb::glibc::logfNormal<false>(x)) != icbLogfF2U(want)) ++bad[t * 4 + 3];
        ++cnt[t * 2 + 1];
      }
    });
  for (auto &t : th) t.join();
  uint64_t b[4] = {0, 0, 0, 0}, c[2] = {0, 0};
  for (int t = 0; t < nt; ++t) { for (int q = 0; q < 4; ++q) b[q] += bad[t * 4 + q]; c[0] += cnt[t * 2]; c[1] += cnt[t * 2 + 1]; }
  printf("draws %llu %llu %llu\nnormals %llu %llu %llu\n", (unsigned long long)c[0], (unsigned long long)b[0], (unsigned long long)b[1],
         (unsigned long long)c[1], (unsigned long long)b[2], (unsigned long long)b[3]);
  return (b[0] != 0 && b[1] != 0) ? 1 : 0;
}
