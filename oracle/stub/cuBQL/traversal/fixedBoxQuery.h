// Stand-in for the cuBQL submodule header that icon_rt/Params.h:20,42 includes
// unconditionally. cuBQL (pinned eb7763e5) is absent from /root/reference; the
// CPU (non-RTCORE) branch of the reference never calls into it, it only needs
// the type name to exist.
#pragma once
namespace cuBQL { template <typename T, int D> struct BinaryBVH { int unused; }; }
