// kernel_util.cuh -- thread mapping and vector load/store shared by all kernels.
//
// Fields are [element][level] with level innermost.  One thread owns V
// consecutive levels (V=2 -> one 16-byte Real2 access) of one element; the
// work index runs level-fastest so a warp covers 32*V consecutive levels of
// as few elements as possible: with 64 levels a warp is exactly one 512-byte
// element row, every neighbour gather is one fully coalesced row read and the
// element's connectivity/geometry loads are warp-uniform broadcasts.
#pragma once
#include <cuda_runtime.h>

#include "real.h"

namespace omega {

constexpr int kBlock = 256;

template <int V> __device__ __forceinline__ void ldv(Real (&r)[V], const Real *__restrict__ p);
template <> __device__ __forceinline__ void ldv<1>(Real (&r)[1], const Real *__restrict__ p) {
   r[0] = __ldg(p);
}
template <> __device__ __forceinline__ void ldv<2>(Real (&r)[2], const Real *__restrict__ p) {
   const Real2 t = __ldg(reinterpret_cast<const Real2 *>(p));
   r[0] = t.x;
   r[1] = t.y;
}
// plain (coherent) load for read-modify-write targets
template <int V> __device__ __forceinline__ void ldp(Real (&r)[V], const Real *p);
template <> __device__ __forceinline__ void ldp<1>(Real (&r)[1], const Real *p) { r[0] = *p; }
template <> __device__ __forceinline__ void ldp<2>(Real (&r)[2], const Real *p) {
   const Real2 t = *reinterpret_cast<const Real2 *>(p);
   r[0] = t.x;
   r[1] = t.y;
}
template <int V> __device__ __forceinline__ void stv(Real *p, const Real (&r)[V]);
template <> __device__ __forceinline__ void stv<1>(Real *p, const Real (&r)[1]) { *p = r[0]; }
template <> __device__ __forceinline__ void stv<2>(Real *p, const Real (&r)[2]) {
   *reinterpret_cast<Real2 *>(p) = make_real2(r[0], r[1]);
}

// fire-and-forget prefetch of this lane's part of a row: raises the bytes in flight
// per SM without landing registers (mode 1: into L1, mode 2: into L2 only)
__device__ __forceinline__ void prefetch_row(const Real *p, int mode) {
   if (mode == 1) asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
   else if (mode == 2) asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}

// work index -> (element i, first level k); returns false past the end
template <int V> __device__ __forceinline__ bool elem_k(int N, int K, int &i, int &k) {
   const int KV      = K / V;
   const long long w = blockIdx.x * (long long)blockDim.x + threadIdx.x;
   if (w >= (long long)N * KV) return false;
   i = (int)(w / KV);
   k = ((int)(w - (long long)i * KV)) * V;
   return true;
}

inline unsigned grid_for(long long NWork) { return (unsigned)((NWork + kBlock - 1) / kBlock); }

// dispatch on vector width: V=2 when K is even, else 1
#define OMEGA_LAUNCH_V(KERNEL, N, K, STREAM, ...)                                              \
   do {                                                                                        \
      if ((N) > 0) {                                                                           \
         if (((K) & 1) == 0)                                                                   \
            KERNEL<2><<<::omega::grid_for((long long)(N) * ((K) / 2)), ::omega::kBlock, 0, STREAM>>>(__VA_ARGS__); \
         else                                                                                  \
            KERNEL<1><<<::omega::grid_for((long long)(N) * (K)), ::omega::kBlock, 0, STREAM>>>(__VA_ARGS__);       \
      }                                                                                        \
   } while (0)

} // namespace omega
