// asinf / atanf / atan2f as the reference's CPU path computes them: glibc 2.39 x86-64 (sysdeps/ieee754/flt-32
// e_asinf.c, s_atanf.c, e_atan2f.c -- the fdlibm single-precision kernels, compiled to plain SSE mulss/addss/divss/sqrtss,
// no FMA; libm.so.6 __asinf_finite / atanf / __atan2f_finite).  The reference's SPHERE_ACCEL_MODE projects ray points to
// (lat, lon) cells through toSpherical (ICONGrid.h:36-42: asinf(z/r), atan2f(y,x)); a last-bit difference moves a point
// across a cell boundary and with it the ray's whole random sequence, so the device evaluates the same operation
// sequence with round-to-nearest, non-contracted arithmetic (__f*_rn) and is bit-identical on finite input.
// tests/test_fdlibm_trig.py sweeps this header (compiled for the host) against the host's libm.
//
// Compiles as device code (nvcc) and as host C++ (g++ -ffp-contract=off) for that sweep.
#pragma once
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define ICB_TRIG_FN __device__ __forceinline__
#define ICB_FMUL(a, b) __fmul_rn((a), (b))
#define ICB_FADD(a, b) __fadd_rn((a), (b))
#define ICB_FSUB(a, b) __fsub_rn((a), (b))
#define ICB_FDIV(a, b) __fdiv_rn((a), (b))
#define ICB_FSQRT(a) __fsqrt_rn((a))
#define ICB_F2U(a) __float_as_uint((a))
#define ICB_U2F(a) __uint_as_float((a))
#else
#include <cmath>
#define ICB_TRIG_FN static inline
#define ICB_FMUL(a, b) ((a) * (b))
#define ICB_FADD(a, b) ((a) + (b))
#define ICB_FSUB(a, b) ((a) - (b))
#define ICB_FDIV(a, b) ((a) / (b))
#define ICB_FSQRT(a) sqrtf((a))
static inline uint32_t icbF2U(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
static inline float icbU2F(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
#define ICB_F2U(a) icbF2U((a))
#define ICB_U2F(a) icbU2F((a))
#endif

namespace icb {
namespace fdlibm {

// e_asinf.c: |x| < 0.5 -> x + x*P(x^2); 0.5 <= |x| < 1 -> pi/2 - 2*asin(sqrt((1-|x|)/2)), in two precisions
ICB_TRIG_FN float asinF(float x)
{
  const float pio2_hi = ICB_U2F(0x3fc90fdbu), pio2_lo = ICB_U2F(0xb33bbd2eu), pio4_hi = ICB_U2F(0x3f490fdbu);
  const float p0 = ICB_U2F(0x3e2aaae4u), p1 = ICB_U2F(0x3d9980f2u), p2 = ICB_U2F(0x3d3a3f25u), p3 = ICB_U2F(0x3cc6141eu),
              p4 = ICB_U2F(0x3d2cb694u);
  const uint32_t hx = ICB_F2U(x), ix = hx & 0x7fffffffu;
  if (ix == 0x3f800000u) return ICB_FADD(ICB_FMUL(x, pio2_hi), ICB_FMUL(x, pio2_lo));     // asin(+-1) = +-pi/2, inexact
  if (ix > 0x3f800000u) { const float d = ICB_FSUB(x, x); return ICB_FDIV(d, d); }       // |x| > 1 (or NaN): NaN
  if (ix < 0x3f000000u) {                                                                 // |x| < 0.5
    if (ix < 0x32000000u) return x;                                                       // |x| < 2^-27
    const float t = ICB_FMUL(x, x);
    float w = ICB_FMUL(p4, t);
    w = ICB_FMUL(ICB_FADD(w, p3), t);
    w = ICB_FMUL(ICB_FADD(w, p2), t);
    w = ICB_FMUL(ICB_FADD(w, p1), t);
    w = ICB_FMUL(ICB_FADD(w, p0), t);
    return ICB_FADD(ICB_FMUL(w, x), x);
  }
  const float ax = ICB_U2F(ix);
  const float t = ICB_FMUL(ICB_FSUB(1.f, ax), 0.5f);
  float p = ICB_FMUL(p4, t);
  p = ICB_FMUL(ICB_FADD(p, p3), t);
  p = ICB_FMUL(ICB_FADD(p, p2), t);
  p = ICB_FMUL(ICB_FADD(p, p1), t);
  p = ICB_FMUL(ICB_FADD(p, p0), t);
  const float s = ICB_FSQRT(t);
  float r;
  if (ix >= 0x3f79999au) {                                                                // |x| > 0.975
    const float a = ICB_FADD(ICB_FMUL(p, s), s);                                          // s + s*p
    r = ICB_FSUB(pio2_hi, ICB_FADD(ICB_FADD(a, a), ICB_U2F(0x333bbd2eu)));                // pio2_hi - (2(s+s*p) - pio2_lo)
  } else {
    const float w = ICB_U2F(ICB_F2U(s) & 0xfffff000u);
    const float c = ICB_FDIV(ICB_FSUB(t, ICB_FMUL(w, w)), ICB_FADD(s, w));
    const float pp = ICB_FSUB(ICB_FMUL(ICB_FADD(s, s), p), ICB_FSUB(pio2_lo, ICB_FADD(c, c)));
    const float q = ICB_FSUB(pio4_hi, ICB_FADD(w, w));
    r = ICB_FSUB(pio4_hi, ICB_FSUB(pp, q));
  }
  return (hx >> 31) ? -r : r;
}

// s_atanf.c: argument reduction to |x| < 7/16 around atan(0.5), atan(1), atan(1.5), atan(inf), odd/even split polynomial
ICB_TRIG_FN float atanF(float x)
{
  const uint32_t hx = ICB_F2U(x), ix = hx & 0x7fffffffu;
  const float hi3 = ICB_U2F(0x3fc90fdau), lo3 = ICB_U2F(0x33a22168u);
  if (ix >= 0x4c000000u) {                                                                // |x| >= 2^25
    if (ix > 0x7f800000u) return ICB_FADD(x, x);
    return (hx >> 31) ? ICB_FSUB(-hi3, lo3) : ICB_FADD(lo3, hi3);
  }
  int id;
  float hi = 0.f, lo = 0.f;
  if (ix < 0x3ee00000u) {                                                                 // |x| < 0.4375
    if (ix < 0x31000000u) return x;                                                       // |x| < 2^-29
    id = -1;
  } else {
    const float ax = ICB_U2F(ix);
    if (ix < 0x3f980000u) {                                                               // |x| < 1.1875
      if (ix < 0x3f300000u) {                                                             // 7/16 <= |x| < 11/16
        id = 0; hi = ICB_U2F(0x3eed6338u); lo = ICB_U2F(0x31ac3769u);
        x = ICB_FDIV(ICB_FSUB(ICB_FADD(ax, ax), 1.f), ICB_FADD(ax, 2.f));
      } else {
        id = 1; hi = ICB_U2F(0x3f490fdau); lo = ICB_U2F(0x33222168u);
        x = ICB_FDIV(ICB_FSUB(ax, 1.f), ICB_FADD(ax, 1.f));
      }
    } else if (ix < 0x401c0000u) {                                                        // |x| < 2.4375
      id = 2; hi = ICB_U2F(0x3f7b985eu); lo = ICB_U2F(0x33140fb4u);
      x = ICB_FDIV(ICB_FSUB(ax, 1.5f), ICB_FADD(ICB_FMUL(ax, 1.5f), 1.f));
    } else {
      id = 3; hi = hi3; lo = lo3;
      x = ICB_FDIV(-1.f, ax);
    }
  }
  const float z = ICB_FMUL(x, x), w = ICB_FMUL(z, z);
  float s1 = ICB_FMUL(ICB_U2F(0x3c8569d7u), w);
  s1 = ICB_FMUL(ICB_FADD(s1, ICB_U2F(0x3d4bda59u)), w);
  s1 = ICB_FMUL(ICB_FADD(s1, ICB_U2F(0x3d886b35u)), w);
  s1 = ICB_FMUL(ICB_FADD(s1, ICB_U2F(0x3dba2e6eu)), w);
  s1 = ICB_FMUL(ICB_FADD(s1, ICB_U2F(0x3e124925u)), w);
  s1 = ICB_FMUL(ICB_FADD(s1, ICB_U2F(0x3eaaaaabu)), z);
  float s2 = ICB_FMUL(ICB_U2F(0xbd15a221u), w);
  s2 = ICB_FMUL(ICB_FSUB(s2, ICB_U2F(0x3d6ef16bu)), w);
  s2 = ICB_FMUL(ICB_FSUB(s2, ICB_U2F(0x3d9d8795u)), w);
  s2 = ICB_FMUL(ICB_FSUB(s2, ICB_U2F(0x3de38e38u)), w);
  s2 = ICB_FMUL(ICB_FSUB(s2, ICB_U2F(0x3e4ccccdu)), w);
  const float xs = ICB_FMUL(ICB_FADD(s1, s2), x);
  if (id < 0) return ICB_FSUB(x, xs);
  const float r = ICB_FSUB(hi, ICB_FSUB(ICB_FSUB(xs, lo), x));
  return (hx >> 31) ? -r : r;
}

// e_atan2f.c: quadrant bookkeeping around atanf(|y/x|)
ICB_TRIG_FN float atan2F(float y, float x)
{
  const float pi = ICB_U2F(0x40490fdbu), pio2 = ICB_U2F(0x3fc90fdbu), pio4 = ICB_U2F(0x3f490fdbu), tiny = ICB_U2F(0x0da24260u);
  const float negPiLo = ICB_U2F(0x33bbbd2eu);                                             // -pi_lo = 8.7422777e-08
  const uint32_t hx = ICB_F2U(x), hy = ICB_F2U(y), ix = hx & 0x7fffffffu, iy = hy & 0x7fffffffu;
  if (ix > 0x7f800000u || iy > 0x7f800000u) return ICB_FADD(x, y);
  if (hx == 0x3f800000u) return atanF(y);
  const int m = (int)(hy >> 31) | (int)((hx >> 30) & 2u);
  if (iy == 0u) {
    if (m < 2) return y;
    return m == 2 ? ICB_FADD(tiny, pi) : ICB_FSUB(-pi, tiny);
  }
  if (ix == 0u) return (hy >> 31) ? ICB_FSUB(-pio2, tiny) : ICB_FADD(tiny, pio2);
  if (ix == 0x7f800000u) {
    if (iy == 0x7f800000u) {
      switch (m) {
        case 0: return ICB_FADD(tiny, pio4);
        case 1: return ICB_FSUB(-pio4, tiny);
        case 2: return ICB_FADD(ICB_FMUL(3.f, pio4), tiny);
        default: return ICB_FSUB(ICB_FMUL(-3.f, pio4), tiny);
      }
    }
    switch (m) {
      case 0: return 0.f;
      case 1: return -0.f;
      case 2: return ICB_FADD(tiny, pi);
      default: return ICB_FSUB(-pi, tiny);
    }
  }
  if (iy == 0x7f800000u) return (hy >> 31) ? ICB_FSUB(-pio2, tiny) : ICB_FADD(tiny, pio2);
  const int d = (int)iy - (int)ix, k = d >> 23;
  float z;
  if (d > 0x1e7fffff) z = ICB_FSUB(pio2, ICB_U2F(0x333bbd2eu));                           // |y/x| > 2^60
  else if ((hx >> 31) && k < -60) z = 0.f;                                                // |y/x| < 2^-60, x < 0
  else z = atanF(ICB_U2F(ICB_F2U(ICB_FDIV(y, x)) & 0x7fffffffu));
  switch (m) {
    case 0: return z;
    case 1: return ICB_U2F(ICB_F2U(z) + 0x80000000u);
    case 2: return ICB_FSUB(pi, ICB_FADD(z, negPiLo));
    default: return ICB_FSUB(ICB_FADD(z, negPiLo), pi);
  }
}

}  // namespace fdlibm
}  // namespace icb
