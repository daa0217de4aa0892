// oracle/ref_kernel_shims.h -- TEST INFRASTRUCTURE.
// Host stand-ins for the few CUDA built-ins the reference's majorant-builder kernels use (icon_rt/hostCode.cu:36-56,
// 205-434: initGrid, rasterizeBox, buildGrid_ICON, buildShell_ICON, buildGrid_UMesh, computeMaxOpacities and the float
// atomicMin/atomicMax CAS loops), so that those kernels -- cut out of hostCode.cu by line range into
// oracle/_ref/hostcode_kernels.gen.h at build time (oracle/Makefile; the file never enters the repo) -- run on the CPU
// exactly as the reference wrote them.  hostCode.cu as a whole needs nvcc + OWL and does not compile here.
#pragma once
#include <cstring>

#ifndef __global__
#define __global__
#endif

struct RefDim3 { unsigned x, y, z; };
static thread_local RefDim3 blockIdx = {0, 0, 0}, blockDim = {1, 1, 1}, threadIdx = {0, 0, 0};

static inline int __float_as_int(float f) { int i; std::memcpy(&i, &f, 4); return i; }
static inline float __int_as_float(int i) { float f; std::memcpy(&f, &i, 4); return f; }
static inline int atomicCAS(int *address, int compare, int val) { return __sync_val_compare_and_swap(address, compare, val); }

// launches `kernel<<<ceil(n/1024), 1024>>>(args...)` the way hostCode.cu:661-699 does, one "thread" after the other
template <typename K, typename... A>
static inline void refLaunch1024(size_t n, K kernel, A... args)
{
  blockDim = {1024, 1, 1};
  for (size_t b = 0; b < (n + 1023) / 1024; ++b)
    for (unsigned t = 0; t < 1024; ++t) {
      blockIdx = {(unsigned)b, 0, 0};
      threadIdx = {t, 0, 0};
      kernel(args...);
    }
}
