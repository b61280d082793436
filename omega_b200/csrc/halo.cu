// halo.cu -- device halves of Halo::exchangeFullArrayHalo: pack the rows a
// neighbour needs into one contiguous buffer per neighbour, unpack received
// buffers into the halo rows.  Buffer order follows the reference
// (components/omega/src/base/Halo.h:304-502 pack, :545-760 unpack):
// [outer][exchange index][inner].  Lists arrive in host order
// (Halo.cpp:543-597) and are translated once to device order.
#include <cstring>

#include "kernel_util.cuh"
#include "omega_dev.h"

struct OmegaHalo {
   const OmegaMesh *M;
   int NNeighbors;
   std::vector<int> NeighborTasks;
   std::vector<int> SendOff[3], RecvOff[3]; // [NNeighbors+1]
   int *DSendIdx[3], *DRecvIdx[3];          // device order
   std::vector<void *> Allocs;
};

namespace omega {

// Buf[n-block][o][i][k] <-> Array[o][Idx[i]][k]; one launch covers every neighbour:
// the per-neighbour blocks are contiguous in list order, block n starts at
// NOuter*NInner*Off[n].
template <bool Pack>
__global__ void __launch_bounds__(kBlock)
k_halo_rows(int NNbr, const int *__restrict__ Off, const int *__restrict__ Idx, int NOuter, int NRows,
            int NInner, double *__restrict__ Array, double *__restrict__ Buf) {
   const long long w     = blockIdx.x * (long long)blockDim.x + threadIdx.x;
   const long long total = (long long)Off[NNbr] * NOuter * NInner;
   if (w >= total) return;
   // find neighbour block (few neighbours: linear scan)
   int n = 0;
   while (n + 1 < NNbr && w >= (long long)Off[n + 1] * NOuter * NInner) ++n;
   const long long base = (long long)Off[n] * NOuter * NInner;
   const int cnt        = Off[n + 1] - Off[n];
   const long long r    = w - base;
   const int k          = (int)(r % NInner);
   const long long t    = r / NInner;
   const int i          = (int)(t % cnt);
   const int o          = (int)(t / cnt);
   const size_t a       = ((size_t)o * NRows + Idx[Off[n] + i]) * NInner + k;
   if (Pack) Buf[w] = Array[a];
   else Array[a] = Buf[w];
}

__global__ void k_copy_off(int n, const int *src, int *dst) {
   int i = blockIdx.x * blockDim.x + threadIdx.x;
   if (i < n) dst[i] = src[i];
}

} // namespace omega

using namespace omega;

extern "C" {

int omega_halo_create(const OmegaMesh *M, const OmegaHaloDesc *D, OmegaHalo **Out) {
   OMEGA_REQUIRE(M && D && Out, "omega_halo_create: null argument");
   OmegaHalo *H = new OmegaHalo();
   H->M = M;
   H->NNeighbors = D->NNeighbors;
   H->NeighborTasks.assign(D->NeighborTasks, D->NeighborTasks + D->NNeighbors);
   for (int e = 0; e < 3; ++e) {
      H->SendOff[e].assign(D->SendOffsets[e], D->SendOffsets[e] + D->NNeighbors + 1);
      H->RecvOff[e].assign(D->RecvOffsets[e], D->RecvOffsets[e] + D->NNeighbors + 1);
      for (int dir = 0; dir < 2; ++dir) {
         const std::vector<int> &off = dir ? H->RecvOff[e] : H->SendOff[e];
         const int *idx = dir ? D->RecvIdx[e] : D->SendIdx[e];
         const int n = off[D->NNeighbors];
         // [offsets (NNeighbors+1) | indices (n)] in one allocation
         std::vector<int> h(D->NNeighbors + 1 + n);
         std::memcpy(h.data(), off.data(), sizeof(int) * (D->NNeighbors + 1));
         for (int i = 0; i < n; ++i) h[D->NNeighbors + 1 + i] = M->IPerm[e][idx[i]];
         int *d = nullptr;
         OMEGA_CUDA_CHECK(cudaMalloc((void **)&d, sizeof(int) * h.size()));
         H->Allocs.push_back(d);
         OMEGA_CUDA_CHECK(cudaMemcpy(d, h.data(), sizeof(int) * h.size(), cudaMemcpyHostToDevice));
         (dir ? H->DRecvIdx[e] : H->DSendIdx[e]) = d;
      }
   }
   *Out = H;
   return 0;
}

int omega_halo_destroy(OmegaHalo *H) {
   if (!H) return 0;
   for (void *p : H->Allocs) cudaFree(p);
   delete H;
   return 0;
}

int omega_halo_buffer_counts(const OmegaHalo *H, OmegaMeshElement Elem, int NOuter, int NInner,
                             long long *SendCounts, long long *RecvCounts) {
   OMEGA_REQUIRE(H && Elem >= 0 && Elem < 3, "omega_halo_buffer_counts: bad argument");
   for (int n = 0; n < H->NNeighbors; ++n) {
      if (SendCounts) SendCounts[n] = (long long)(H->SendOff[Elem][n + 1] - H->SendOff[Elem][n]) * NOuter * NInner;
      if (RecvCounts) RecvCounts[n] = (long long)(H->RecvOff[Elem][n + 1] - H->RecvOff[Elem][n]) * NOuter * NInner;
   }
   return 0;
}

int omega_halo_pack(const OmegaHalo *H, OmegaMeshElement Elem, int NOuter, int NInner, const double *Array,
                    double *SendBuf, void *Stream) {
   OMEGA_REQUIRE(H && Array && SendBuf && Elem >= 0 && Elem < 3, "omega_halo_pack: bad argument");
   const int NN = H->NNeighbors;
   const long long total = (long long)H->SendOff[Elem][NN] * NOuter * NInner;
   if (total == 0) return 0;
   const int *d = H->DSendIdx[Elem];
   k_halo_rows<true><<<grid_for(total), kBlock, 0, (cudaStream_t)Stream>>>(
       NN, d, d + NN + 1, NOuter, n_size(H->M, Elem), NInner, const_cast<double *>(Array), SendBuf);
   OMEGA_CUDA_CHECK(cudaGetLastError());
   return 0;
}

int omega_halo_unpack(const OmegaHalo *H, OmegaMeshElement Elem, int NOuter, int NInner, const double *RecvBuf,
                      double *Array, void *Stream) {
   OMEGA_REQUIRE(H && Array && RecvBuf && Elem >= 0 && Elem < 3, "omega_halo_unpack: bad argument");
   const int NN = H->NNeighbors;
   const long long total = (long long)H->RecvOff[Elem][NN] * NOuter * NInner;
   if (total == 0) return 0;
   const int *d = H->DRecvIdx[Elem];
   k_halo_rows<false><<<grid_for(total), kBlock, 0, (cudaStream_t)Stream>>>(
       NN, d, d + NN + 1, NOuter, n_size(H->M, Elem), NInner, Array, const_cast<double *>(RecvBuf));
   OMEGA_CUDA_CHECK(cudaGetLastError());
   return 0;
}

} // extern "C"
