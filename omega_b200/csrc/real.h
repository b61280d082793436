// real.h -- Real of the reference (src/base/DataTypes.h:30-34) inside the library: OmegaReal of the public
// header and its two-level vector.  fp64 unless built with -DOMEGA_SINGLE_PRECISION (libomega_b200_sp.so).
#pragma once
#include <cuda_runtime.h>

#include "omega_b200.h"

namespace omega {
using Real = OmegaReal;
#ifdef OMEGA_SINGLE_PRECISION
using Real2 = float2;
__host__ __device__ __forceinline__ Real2 make_real2(float a, float b) { return make_float2(a, b); }
constexpr bool kSinglePrecision = true;
#else
using Real2 = double2;
__host__ __device__ __forceinline__ Real2 make_real2(double a, double b) { return make_double2(a, b); }
constexpr bool kSinglePrecision = false;
#endif
} // namespace omega
