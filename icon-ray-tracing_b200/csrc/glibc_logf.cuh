// logf as the reference's CPU path computes it: glibc 2.39 x86-64 (sysdeps/ieee754/flt-32/e_logf.c -- table-driven,
// evaluated in double precision, rounded once to float).  libm.so.6 carries two builds behind an ifunc: __logf_fma
// (vfmadd throughout; what a CPU with FMA+AVX2 runs) and __logf_sse2 (separate mulsd/addsd); both operation sequences
// are restated here from the disassembly, constants from .rodata (they are the published __logf_data).  The two agree
// except where the double result lands within ~1e-16 of a float rounding boundary.
//
// Why: a free flight of the Woodcock ray-gens is t -= logf(1 - u) / rate (deviceCode.cu:163).  Round 1 read the
// value from a 64 MB table built by the host's libm -- a dependent random read per tracking step, 88 % of the AE
// ray-gen's stall samples (profiles/r02_e_wae_full_c3.txt).  Computing it costs a 16-entry table lookup, five
// double-precision FMAs and two conversions.  Which build the host runs is not assumed: the first icb_set_integrator
// that selects a Woodcock ray-gen builds the host table once and checks the FMA form -- the one the kernels evaluate,
// logOneMinus<true> -- against all 2^24 arguments on the device (launchLogfCheck).  If it reproduces the table entirely the
// table is dropped and the kernels compute; if not (SSE2 ifunc path, another glibc) they go on reading the table, as
// round 1 did.  icb_selftest_logf reports the mismatch counts of both forms.
//
// Compiles as device code (nvcc) and as host C++ (g++ -ffp-contract=off) for tests/host/logf_sweep.cpp.
#pragma once
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define ICB_LOGF_FN __device__ __forceinline__
#define ICB_DFMA(a, b, c) __fma_rn((a), (b), (c))
#define ICB_DMUL(a, b) __dmul_rn((a), (b))
#define ICB_DADD(a, b) __dadd_rn((a), (b))
#define ICB_LOGF_F2U(a) __float_as_uint((a))
#define ICB_LOGF_U2F(a) __uint_as_float((a))
#define ICB_LOGF_U2D(a) __longlong_as_double((long long)(a))
#else
#include <cmath>
#define ICB_LOGF_FN static inline
#define ICB_DFMA(a, b, c) fma((a), (b), (c))
#define ICB_DMUL(a, b) ((a) * (b))
#define ICB_DADD(a, b) ((a) + (b))
static inline uint32_t icbLogfF2U(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
static inline float icbLogfU2F(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
static inline double icbLogfU2D(uint64_t u) { double d; std::memcpy(&d, &u, 8); return d; }
#define ICB_LOGF_F2U(a) icbLogfF2U((a))
#define ICB_LOGF_U2F(a) icbLogfU2F((a))
#define ICB_LOGF_U2D(a) icbLogfU2D((a))
#endif

namespace icb {
namespace glibc {

// __logf_data.tab: {invc, logc} for the 16 subintervals of [0x1.66p-1, 0x1.66p0) (bit patterns of the doubles)
#if defined(__CUDACC__)
__device__
#endif
static const uint64_t kLogfTab[32] = {
    0x3ff661ec79f8f3beull, 0xbfd57bf7808caadeull, 0x3ff571ed4aaf883dull, 0xbfd2bef0a7c06ddbull, 0x3ff49539f0f010b0ull, 0xbfd01eae7f513a67ull,
    0x3ff3c995b0b80385ull, 0xbfcb31d8a68224e9ull, 0x3ff30d190c8864a5ull, 0xbfc6574f0ac07758ull, 0x3ff25e227b0b8ea0ull, 0xbfc1aa2bc79c8100ull,
    0x3ff1bb4a4a1a343full, 0xbfba4e76ce8c0e5eull, 0x3ff12358f08ae5baull, 0xbfb1973c5a611cccull, 0x3ff0953f419900a7ull, 0xbfa252f438e10c1eull,
    0x3ff0000000000000ull, 0x0000000000000000ull, 0x3fee608cfd9a47acull, 0x3faaa5aa5df25984ull, 0x3feca4b31f026aa0ull, 0x3fbc5e53aa362eb4ull,
    0x3feb2036576afce6ull, 0x3fc526e57720db08ull, 0x3fe9c2d163a1aa2dull, 0x3fcbc2860d224770ull, 0x3fe886e6037841edull, 0x3fd1058bc8a07ee1ull,
    0x3fe767dcf5534862ull, 0x3fd4043057b6ee09ull};

// x positive, finite, normal (the callers' arguments are 1 - k/2^24, k < 2^24).  useFma: the __logf_fma sequence
template <bool USE_FMA>
ICB_LOGF_FN float logfNormal(float x)
{
  const uint32_t ix = ICB_LOGF_F2U(x);
  if (ix == 0x3f800000u) return 0.f;
  const uint32_t tmp = ix - 0x3f330000u;
  const int i = (int)((tmp >> 19) & 15u);
  const int k = (int)tmp >> 23;
  const uint32_t iz = ix - (tmp & 0xff800000u);
  const double invc = ICB_LOGF_U2D(kLogfTab[2 * i]), logc = ICB_LOGF_U2D(kLogfTab[2 * i + 1]);
  const double ln2 = ICB_LOGF_U2D(0x3fe62e42fefa39efull), A0 = ICB_LOGF_U2D(0xbfd00ea348b88334ull), A1 = ICB_LOGF_U2D(0x3fd5575b0be00b6aull),
               A2 = ICB_LOGF_U2D(0xbfdffffef20a4123ull);
  const double z = (double)ICB_LOGF_U2F(iz), kd = (double)k;
  if (USE_FMA) {
    const double y0 = ICB_DFMA(kd, ln2, logc);
    const double r = ICB_DFMA(z, invc, -1.0);
    double y = ICB_DFMA(r, A1, A2);
    const double r2 = ICB_DMUL(r, r);
    const double s = ICB_DADD(r, y0);
    y = ICB_DFMA(r2, A0, y);
    return (float)ICB_DFMA(r2, y, s);
  }
  const double r = ICB_DADD(ICB_DMUL(z, invc), -1.0);
  const double r2 = ICB_DMUL(r, r);
  double y = ICB_DADD(ICB_DMUL(A1, r), A2);
  y = ICB_DADD(ICB_DMUL(A0, r2), y);
  const double y0 = ICB_DADD(ICB_DMUL(kd, ln2), logc);
  return (float)ICB_DADD(ICB_DADD(y0, r), ICB_DMUL(y, r2));
}

// logf(1 - k/2^24) for the 24 random bits k of the reference's LCG draw: the argument is exact in fp32
template <bool USE_FMA>
ICB_LOGF_FN float logOneMinus(uint32_t k)
{
  return logfNormal<USE_FMA>((float)(0x01000000u - k) * 5.9604644775390625e-8f);
}

}  // namespace glibc
}  // namespace icb
