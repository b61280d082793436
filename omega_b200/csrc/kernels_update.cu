// kernels_update.cu -- the stage-update kernels of TimeStepper
// (components/omega/src/timeStepping/TimeStepper.cpp:378-524) and the row
// permutation used by field transfers and the halo pack/unpack.
#include <algorithm>

#include "kernel_util.cuh"
#include "omega_dev.h"

namespace omega {

// updateThicknessByTend / updateVelocityByTend (:378-430): Y1 = Y2 + c*T
template <int V>
__global__ void __launch_bounds__(kBlock)
k_axpy(int N, int K, Real *Y1, const Real *Y2, Real C, const Real *__restrict__ T) {
   int i, k;
   if (!elem_k<V>(N, K, i, k)) return;
   const size_t o = (size_t)i * K + k;
   Real y[V], t[V];
   ldp<V>(y, Y2 + o);
   ldv<V>(t, T + o);
#pragma unroll
   for (int q = 0; q < V; ++q) y[q] = y[q] + C * t[q];
   stv<V>(Y1 + o, y);
}

// Mode 0 updateTracersByTend (:446-469)  Next = (Cur*H2 + c*T)/H1
// Mode 1 weightTracers (:473-487)        Next = Cur*H1
// Mode 2 accumulateTracersUpdate (:492-507) Next += c*T
// Mode 3 finalizeTracersUpdate (:511-524)   Next /= H1
template <int V, int Mode>
__global__ void __launch_bounds__(kBlock)
k_tracer_update(int NAll, int NSize, int K, int NTracers, Real *Next, const Real *Cur,
                const Real *H1, const Real *H2, Real C, const Real *__restrict__ T) {
   int i, k;
   if (!elem_k<V>(NAll, K, i, k)) return;
   const size_t o = (size_t)i * K + k, CS = (size_t)NSize * K;
   Real h1[V], h2[V];
   if (Mode != 2) ldp<V>(h1, H1 + o);
   if (Mode == 0) ldp<V>(h2, H2 + o);
   for (int l = 0; l < NTracers; ++l) {
      Real r[V], a[V], t[V];
      if (Mode == 0 || Mode == 1) ldp<V>(a, Cur + l * CS + o);
      if (Mode == 2 || Mode == 3) ldp<V>(a, Next + l * CS + o);
      if (Mode == 0 || Mode == 2) ldv<V>(t, T + l * CS + o);
#pragma unroll
      for (int q = 0; q < V; ++q) {
         if (Mode == 0) r[q] = (a[q] * h2[q] + C * t[q]) / h1[q];
         if (Mode == 1) r[q] = a[q] * h1[q];
         if (Mode == 2) r[q] = a[q] + C * t[q];
         if (Mode == 3) r[q] = a[q] / h1[q];
      }
      stv<V>(Next + l * CS + o, r);
   }
}

// Dst[o][r][:] = Src[o][SrcRowOfDst[r]][:]
__global__ void __launch_bounds__(kBlock)
k_permute_rows(int NOuter, int NRows, int NInner, const int *__restrict__ SrcRowOfDst,
               const Real *__restrict__ Src, Real *__restrict__ Dst) {
   const long long w     = blockIdx.x * (long long)blockDim.x + threadIdx.x;
   const long long total = (long long)NOuter * NRows * NInner;
   if (w >= total) return;
   const int k       = (int)(w % NInner);
   const long long t = w / NInner;
   const int r       = (int)(t % NRows);
   const int o       = (int)(t / NRows);
   Dst[w] = Src[((size_t)o * NRows + SrcRowOfDst[r]) * NInner + k];
}

void launch_axpy(const Launch &L, int NAll, int K, Real *Y1, const Real *Y2, Real C, const Real *T) {
   OMEGA_LAUNCH_V(k_axpy, NAll, K, L.Stream, NAll, K, Y1, Y2, C, T);
   ++*L.Count;
}

#define OMEGA_LAUNCH_TR(MODE, ...)                                                                \
   do {                                                                                           \
      if (NAll > 0 && NTracers > 0) {                                                             \
         if ((K & 1) == 0)                                                                        \
            k_tracer_update<2, MODE><<<grid_for((long long)NAll * (K / 2)), kBlock, 0, L.Stream>>>(__VA_ARGS__); \
         else                                                                                     \
            k_tracer_update<1, MODE><<<grid_for((long long)NAll * K), kBlock, 0, L.Stream>>>(__VA_ARGS__);       \
         ++*L.Count;                                                                              \
      }                                                                                           \
   } while (0)

void launch_update_tracers(const Launch &L, int NAll, int NSize, int K, int NTracers, Real *Next,
                           const Real *Cur, const Real *H1, const Real *H2, Real C, const Real *T) {
   OMEGA_LAUNCH_TR(0, NAll, NSize, K, NTracers, Next, Cur, H1, H2, C, T);
}
void launch_weight_tracers(const Launch &L, int NAll, int NSize, int K, int NTracers, Real *Next,
                           const Real *Cur, const Real *H) {
   OMEGA_LAUNCH_TR(1, NAll, NSize, K, NTracers, Next, Cur, H, nullptr, 0.0, nullptr);
}
void launch_accumulate_tracers(const Launch &L, int NAll, int NSize, int K, int NTracers, Real *Accum,
                               Real C, const Real *T) {
   OMEGA_LAUNCH_TR(2, NAll, NSize, K, NTracers, Accum, nullptr, nullptr, nullptr, C, T);
}
void launch_finalize_tracers(const Launch &L, int NAll, int NSize, int K, int NTracers, Real *Next,
                             const Real *H) {
   OMEGA_LAUNCH_TR(3, NAll, NSize, K, NTracers, Next, nullptr, H, nullptr, 0.0, nullptr);
}

// One slab of HOST rows [Row0, Row0 + NSlab) of one [NRows][NInner] field, staged contiguously in Stage:
// ToDevice: Dev[DevRowOfHost[Row0 + i]] = Stage[i];  else: Stage[i] = Dev[DevRowOfHost[Row0 + i]]
template <bool ToDevice>
__global__ void __launch_bounds__(kBlock)
k_slab_rows(int Row0, int NSlab, int NInner, const int *__restrict__ DevRowOfHost, Real *__restrict__ Dev,
            Real *__restrict__ Stage) {
   const long long w = blockIdx.x * (long long)blockDim.x + threadIdx.x;
   if (w >= (long long)NSlab * NInner) return;
   const int i = (int)(w / NInner), k = (int)(w - (long long)i * NInner);
   const size_t d = (size_t)DevRowOfHost[Row0 + i] * NInner + k;
   if (ToDevice) Dev[d] = Stage[w];
   else Stage[w] = Dev[d];
}

void launch_slab_rows(cudaStream_t S, bool ToDevice, int Row0, int NSlab, int NInner, const int *DevRowOfHost, Real *Dev,
                      Real *Stage) {
   const long long total = (long long)NSlab * NInner;
   if (total <= 0) return;
   if (ToDevice) k_slab_rows<true><<<grid_for(total), kBlock, 0, S>>>(Row0, NSlab, NInner, DevRowOfHost, Dev, Stage);
   else k_slab_rows<false><<<grid_for(total), kBlock, 0, S>>>(Row0, NSlab, NInner, DevRowOfHost, Dev, Stage);
}

// The whole field between PINNED host memory (read / written by the kernel itself over PCIe, no staging) and
// the device-order array: Dev[o][DevRowOfHost[r]][:] <-> Host[o][r][:].  A warp moves one row at a time with
// 16-byte accesses (a 64-level row = 512 contiguous bytes on both sides).  Few small CTAs: the kernel lives
// beside the stage kernels of another stream and is bound by the link, not by the SMs.
template <bool ToDevice, int V>
__global__ void __launch_bounds__(128)
k_rows_hostmem(int NOuter, int NRows, int NInner, const int *__restrict__ DevRowOfHost, Real *__restrict__ Dev,
               Real *__restrict__ Host) {
   const int KV = NInner / V;
   const long long total = (long long)NOuter * NRows * KV;
   for (long long w = blockIdx.x * (long long)blockDim.x + threadIdx.x; w < total; w += (long long)gridDim.x * blockDim.x) {
      const int kv      = (int)(w % KV);
      const long long t = w / KV;
      const int r = (int)(t % NRows), o = (int)(t / NRows);
      const size_t h = (size_t)t * NInner + (size_t)kv * V;
      const size_t d = ((size_t)o * NRows + DevRowOfHost[r]) * NInner + (size_t)kv * V;
      if (V == 2) {
         if (ToDevice) *reinterpret_cast<Real2 *>(Dev + d) = __ldcs(reinterpret_cast<const Real2 *>(Host + h));
         else __stcs(reinterpret_cast<Real2 *>(Host + h), *reinterpret_cast<const Real2 *>(Dev + d));
      } else {
         if (ToDevice) Dev[d] = __ldcs(Host + h);
         else __stcs(Host + h, Dev[d]);
      }
   }
}

void launch_rows_hostmem(cudaStream_t S, bool ToDevice, int NOuter, int NRows, int NInner, const int *DevRowOfHost, Real *Dev,
                         Real *Host, int NCtas) {
   const long long total = (long long)NOuter * NRows * NInner;
   if (total <= 0) return;
   const unsigned g = (unsigned)std::max(1, NCtas);
   if ((NInner & 1) == 0) {
      if (ToDevice) k_rows_hostmem<true, 2><<<g, 128, 0, S>>>(NOuter, NRows, NInner, DevRowOfHost, Dev, Host);
      else k_rows_hostmem<false, 2><<<g, 128, 0, S>>>(NOuter, NRows, NInner, DevRowOfHost, Dev, Host);
   } else {
      if (ToDevice) k_rows_hostmem<true, 1><<<g, 128, 0, S>>>(NOuter, NRows, NInner, DevRowOfHost, Dev, Host);
      else k_rows_hostmem<false, 1><<<g, 128, 0, S>>>(NOuter, NRows, NInner, DevRowOfHost, Dev, Host);
   }
}

void launch_permute_rows(cudaStream_t S, int NOuter, int NRows, int NInner, const int *SrcRowOfDst,
                         const Real *Src, Real *Dst) {
   const long long total = (long long)NOuter * NRows * NInner;
   if (total > 0) k_permute_rows<<<grid_for(total), kBlock, 0, S>>>(NOuter, NRows, NInner, SrcRowOfDst, Src, Dst);
}

} // namespace omega
