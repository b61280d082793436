// halo.cu -- Halo::exchangeFullArrayHalo behind the C ABI.
//
// Reference: components/omega/src/base/Halo.h:768-915 (post receives, pack,
// send, unpack as messages land), Halo.cpp:455-757 (lists, neighbours,
// buffers).  Here an exchange ROUND moves every array of one exchange point of
// the stepper (h, u and all tracers; RungeKutta4Stepper.cpp:108-114,130-131)
// with ONE pack launch, ONE grouped NCCL send/recv per neighbour over NVLink
// and ONE unpack launch:
//   * k_halo_round<Pack>: every (neighbour, array) pair is a blockIdx.y slice;
//     a thread moves a Real2 (two levels) of one row, so a 64-level row is
//     one coalesced 512-byte access on both sides;
//   * the message of neighbour n is the concatenation of its per-array
//     buffers, each in the reference's order [outer][exchange index][level]
//     (Halo.h:304-502 pack, :545-760 unpack);
//   * NCCL is loaded at run time (dlopen libnccl.so.2: the host process -- the
//     reference's MPI ranks, or a torch process -- usually has it loaded
//     already); the communicator is created from a unique id the host
//     broadcasts (MPI_Bcast / torch.distributed), one rank per GPU.
// The two device halves stay exported (omega_halo_pack_round / unpack_round and
// the per-array omega_halo_pack / unpack) for callers that own the transport
// and for the one-GPU N-rank emulation of the tests.
//
// omega_halo_create also classifies the mesh tiles (omega_dev.h) for the
// overlap of the exchange with interior compute: boundary tiles hold send-list
// rows and are launched first by the stepper (step.cu), halo tiles are skipped
// in the stages whose result the exchange overwrites.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstring>

#include "halo.h"
#include "kernel_util.cuh"
#include "omega_dev.h"

namespace omega {

// ---- per-array kernels (kept: one array, caller-owned buffers) ---------------
// Buf[n-block][o][i][k] <-> Array[o][Idx[i]][k]; one launch covers every neighbour:
// the per-neighbour blocks are contiguous in list order, block n starts at
// NOuter*NInner*Off[n].
template <bool Pack, class T = Real>
__global__ void __launch_bounds__(kBlock)
k_halo_rows(int NNbr, const int *__restrict__ Off, const int *__restrict__ Idx, int NOuter, int NRows,
            int NInner, T *__restrict__ Array, T *__restrict__ Buf) {
   const long long w     = blockIdx.x * (long long)blockDim.x + threadIdx.x;
   const long long total = (long long)Off[NNbr] * NOuter * NInner;
   if (w >= total) return;
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

// ---- one exchange round: all arrays, all neighbours, one launch ----------------
constexpr int kRoundThreads = 128; // small CTAs: they must find room beside a resident tile CTA

template <bool Pack, int V>
__global__ void __launch_bounds__(kRoundThreads)
k_halo_round(HaloRoundArgs A, Real *__restrict__ Buf) {
   const int n = blockIdx.y / A.NArr, a = blockIdx.y - n * A.NArr;
   const int *Off = A.Off[a];
   const int o0 = Off[n], cnt = Off[n + 1] - o0;
   const int KV = A.K / V;
   const long long items = (long long)cnt * A.NOuter[a] * KV;
   Real *Arr = A.Arr[a];
   Real *B   = Buf + A.Start[a][n];
   const int *Idx = A.Idx[a] + o0;
   for (long long w = blockIdx.x * (long long)kRoundThreads + threadIdx.x; w < items; w += (long long)gridDim.x * kRoundThreads) {
      const int kv      = (int)(w % KV);
      const long long t = w / KV;
      const int i = (int)(t % cnt), o = (int)(t / cnt);
      const size_t src = ((size_t)o * A.NRows[a] + Idx[i]) * A.K + (size_t)kv * V;
      const size_t dst = (size_t)t * A.K + (size_t)kv * V;
      if (V == 2) {
         if (Pack) *reinterpret_cast<Real2 *>(B + dst) = *reinterpret_cast<const Real2 *>(Arr + src);
         else *reinterpret_cast<Real2 *>(Arr + src) = *reinterpret_cast<const Real2 *>(B + dst);
      } else {
         if (Pack) B[dst] = Arr[src];
         else Arr[src] = B[dst];
      }
   }
}

// ---- NCCL, resolved at run time -------------------------------------------------
struct NcclApi {
   void *Lib = nullptr;
   ncclResult_t (*GetUniqueId)(ncclUniqueId *)                                                   = nullptr;
   ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int)                            = nullptr;
   ncclResult_t (*CommDestroy)(ncclComm_t)                                                       = nullptr;
   ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t)     = nullptr;
   ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t)           = nullptr;
   ncclResult_t (*GroupStart)()                                                                  = nullptr;
   ncclResult_t (*GroupEnd)()                                                                    = nullptr;
   const char *(*GetErrorString)(ncclResult_t)                                                   = nullptr;
};

static NcclApi g_nccl;

static int load_nccl() {
   if (g_nccl.Lib) return 0;
   const char *names[] = {getenv("OMEGA_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
   void *lib = nullptr;
   for (const char *n : names)
      if (n && (lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break;
   OMEGA_REQUIRE(lib, std::string("NCCL not found (dlopen libnccl.so.2): ") + (dlerror() ? dlerror() : ""));
#define OMEGA_NCCL_SYM(FIELD, NAME)                                                          \
   g_nccl.FIELD = reinterpret_cast<decltype(g_nccl.FIELD)>(dlsym(lib, NAME));                \
   OMEGA_REQUIRE(g_nccl.FIELD, std::string("NCCL symbol missing: ") + NAME)
   OMEGA_NCCL_SYM(GetUniqueId, "ncclGetUniqueId");
   OMEGA_NCCL_SYM(CommInitRank, "ncclCommInitRank");
   OMEGA_NCCL_SYM(CommDestroy, "ncclCommDestroy");
   OMEGA_NCCL_SYM(Send, "ncclSend");
   OMEGA_NCCL_SYM(Recv, "ncclRecv");
   OMEGA_NCCL_SYM(GroupStart, "ncclGroupStart");
   OMEGA_NCCL_SYM(GroupEnd, "ncclGroupEnd");
   OMEGA_NCCL_SYM(GetErrorString, "ncclGetErrorString");
#undef OMEGA_NCCL_SYM
   g_nccl.Lib = lib;
   return 0;
}

#define OMEGA_NCCL_CHECK(expr)                                                                \
   do {                                                                                       \
      ncclResult_t r_ = (expr);                                                               \
      if (r_ != ncclSuccess) {                                                                \
         ::omega::set_error(std::string(#expr) + ": " + g_nccl.GetErrorString(r_));           \
         return 1;                                                                            \
      }                                                                                       \
   } while (0)

// (re)size the library-owned round buffers and fill the launch arguments
static int round_args(OmegaHalo *H, Real *const *Ptrs, const int *Elems, const int *NOuter, int N, bool Send,
                      HaloRoundArgs &A, long long &Total) {
   const int NN = H->NNeighbors, K = H->M->D.K;
   OMEGA_REQUIRE(N >= 1 && N <= kHaloMaxArrays, "halo round: 1..3 arrays");
   OMEGA_REQUIRE(NN <= kHaloMaxNeighbors, "halo round: more neighbours than kHaloMaxNeighbors; use the per-array entry points");
   std::memset(&A, 0, sizeof(A));
   A.NArr = N; A.K = K;
   long long pos = 0;
   for (int n = 0; n < NN; ++n)
      for (int a = 0; a < N; ++a) {
         const int e = Elems[a];
         const std::vector<int> &off = Send ? H->SendOff[e] : H->RecvOff[e];
         A.Start[a][n] = pos;
         pos += (long long)(off[n + 1] - off[n]) * NOuter[a] * K;
      }
   for (int a = 0; a < N; ++a) {
      const int e = Elems[a];
      OMEGA_REQUIRE(e >= 0 && e < 3 && Ptrs[a], "halo round: bad array");
      A.Arr[a] = Ptrs[a]; A.NOuter[a] = NOuter[a]; A.NRows[a] = n_size(H->M, e);
      const int *d = Send ? H->DSendIdx[e] : H->DRecvIdx[e];
      A.Off[a] = d; A.Idx[a] = d + NN + 1;
   }
   Total = pos;
   return 0;
}

static int ensure_buffer(OmegaHalo *H, Real **Buf, size_t *Cap, long long Need) {
   if ((size_t)Need <= *Cap) return 0;
   // growing frees the old block: wait for exchanges in flight (rare: first round, or a wider round later)
   OMEGA_CUDA_CHECK(cudaDeviceSynchronize());
   if (*Buf) cudaFree(*Buf);
   OMEGA_CUDA_CHECK(cudaMalloc((void **)Buf, (size_t)Need * sizeof(Real)));
   *Cap = (size_t)Need;
   return 0;
}

template <bool Pack> static int launch_round(const OmegaHalo *H, const HaloRoundArgs &A, long long MaxItems, Real *Buf, cudaStream_t St) {
   if (H->NNeighbors == 0 || MaxItems == 0) return 0;
   const int V = (A.K & 1) ? 1 : 2;
   const unsigned gx = (unsigned)std::min<long long>((MaxItems / V + kRoundThreads - 1) / kRoundThreads, 1024);
   dim3 grid(std::max(gx, 1u), (unsigned)(H->NNeighbors * A.NArr));
   if (V == 2) k_halo_round<Pack, 2><<<grid, kRoundThreads, 0, St>>>(A, Buf);
   else k_halo_round<Pack, 1><<<grid, kRoundThreads, 0, St>>>(A, Buf);
   OMEGA_CUDA_CHECK(cudaGetLastError());
   return 0;
}

static long long max_items(const OmegaHalo *H, const int *Elems, const int *NOuter, int N, bool Send) {
   long long m = 0;
   for (int n = 0; n < H->NNeighbors; ++n)
      for (int a = 0; a < N; ++a) {
         const std::vector<int> &off = Send ? H->SendOff[Elems[a]] : H->RecvOff[Elems[a]];
         m = std::max(m, (long long)(off[n + 1] - off[n]) * NOuter[a] * H->M->D.K);
      }
   return m;
}

int halo_pack_round(OmegaHalo *H, Real *const *Ptrs, const int *Elems, const int *NOuter, int N, cudaStream_t St) {
   HaloRoundArgs A;
   long long total = 0;
   if (round_args(H, Ptrs, Elems, NOuter, N, true, A, total)) return 2;
   if (ensure_buffer(H, &H->SendBuf, &H->SendCap, total)) return 1;
   H->LastSendDoubles = total;
   return launch_round<true>(H, A, max_items(H, Elems, NOuter, N, true), H->SendBuf, St);
}

int halo_unpack_round(OmegaHalo *H, Real *const *Ptrs, const int *Elems, const int *NOuter, int N, cudaStream_t St) {
   HaloRoundArgs A;
   long long total = 0;
   if (round_args(H, Ptrs, Elems, NOuter, N, false, A, total)) return 2;
   if (ensure_buffer(H, &H->RecvBuf, &H->RecvCap, total)) return 1;
   return launch_round<false>(H, A, max_items(H, Elems, NOuter, N, false), H->RecvBuf, St);
}

// message sizes (doubles) of neighbour n for this round
static void round_counts(const OmegaHalo *H, const int *Elems, const int *NOuter, int N, int n, long long &S, long long &R) {
   S = R = 0;
   for (int a = 0; a < N; ++a) {
      const int e = Elems[a];
      S += (long long)(H->SendOff[e][n + 1] - H->SendOff[e][n]) * NOuter[a] * H->M->D.K;
      R += (long long)(H->RecvOff[e][n + 1] - H->RecvOff[e][n]) * NOuter[a] * H->M->D.K;
   }
}

// pack -> grouped NCCL send/recv with every neighbour -> unpack, all on St
int halo_exchange_round(OmegaHalo *H, Real *const *Ptrs, const int *Elems, const int *NOuter, int N, cudaStream_t St) {
   OMEGA_REQUIRE(H->NNeighbors == 0 || H->Comm, "omega_halo_exchange_round: no communicator (omega_halo_comm_init)");
   if (H->NNeighbors == 0) return 0;
   // both buffers sized before anything is enqueued (ensure_buffer may synchronise)
   {
      long long ts = 0, tr = 0;
      for (int n = 0; n < H->NNeighbors; ++n) {
         long long s, r;
         round_counts(H, Elems, NOuter, N, n, s, r);
         ts += s; tr += r;
      }
      if (ensure_buffer(H, &H->SendBuf, &H->SendCap, ts) || ensure_buffer(H, &H->RecvBuf, &H->RecvCap, tr)) return 1;
   }
   HaloTimed tm;
   const bool timed = H->Profiling != 0;
   if (timed) {
      OMEGA_CUDA_CHECK(cudaEventCreate(&tm.Start));
      OMEGA_CUDA_CHECK(cudaEventCreate(&tm.Stop));
      OMEGA_CUDA_CHECK(cudaEventRecord(tm.Start, St));
   }
   if (halo_pack_round(H, Ptrs, Elems, NOuter, N, St)) return 1;
   ncclComm_t comm = (ncclComm_t)H->Comm;
   OMEGA_NCCL_CHECK(g_nccl.GroupStart());
   long long so = 0, ro = 0;
   for (int n = 0; n < H->NNeighbors; ++n) {
      long long s, r;
      round_counts(H, Elems, NOuter, N, n, s, r);
      const int peer = H->NeighborTasks[n];
      if (s) OMEGA_NCCL_CHECK(g_nccl.Send(H->SendBuf + so, (size_t)s, (kSinglePrecision ? ncclFloat : ncclDouble), peer, comm, St));
      if (r) OMEGA_NCCL_CHECK(g_nccl.Recv(H->RecvBuf + ro, (size_t)r, (kSinglePrecision ? ncclFloat : ncclDouble), peer, comm, St));
      so += s; ro += r;
   }
   OMEGA_NCCL_CHECK(g_nccl.GroupEnd());
   if (halo_unpack_round(H, Ptrs, Elems, NOuter, N, St)) return 1;
   if (timed) {
      OMEGA_CUDA_CHECK(cudaEventRecord(tm.Stop, St));
      tm.Bytes = (so + ro) * (long long)sizeof(Real);
      H->Events.push_back(tm);
   }
   H->Rounds += 1;
   return 0;
}

} // namespace omega

using namespace omega;

extern "C" {

int omega_halo_create(const OmegaMesh *M, const OmegaHaloDesc *D, OmegaHalo **Out) {
   OMEGA_REQUIRE(M && D && Out, "omega_halo_create: null argument");
   OmegaHalo *H = new OmegaHalo();
   H->M = M;
   H->MyTask = D->MyTask;
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
   // ---- tile classes for the overlapped exchange (step.cu): [boundary | interior owned | halo]
   const int nt = (int)M->TileC0.size();
   if (nt > 0) {
      std::vector<char> bnd(nt, 0);
      const int ns_c = H->SendOff[0][D->NNeighbors], ns_e = H->SendOff[1][D->NNeighbors];
      for (int i = 0; i < ns_c; ++i) {
         const int t = M->TileOfCell[M->IPerm[0][D->SendIdx[0][i]]];
         if (t >= 0) bnd[t] = 1;
      }
      for (int i = 0; i < ns_e; ++i) {
         const int t = M->TileOfEdge[M->IPerm[1][D->SendIdx[1][i]]];
         if (t >= 0) bnd[t] = 1;
      }
      std::vector<int> order;
      int nOwned = 0;
      for (int t = 0; t < nt; ++t) nOwned += M->TileC0[t] < M->D.NCellsOwned;
      for (int t = 0; t < nOwned; ++t)
         if (bnd[t]) order.push_back(t);
      H->NTilesBoundary = (int)order.size();
      for (int t = 0; t < nOwned; ++t)
         if (!bnd[t]) order.push_back(t);
      H->NTilesOwned = nOwned;
      for (int t = nOwned; t < nt; ++t) order.push_back(t);
      int *d = nullptr;
      OMEGA_CUDA_CHECK(cudaMalloc((void **)&d, sizeof(int) * order.size()));
      H->Allocs.push_back(d);
      OMEGA_CUDA_CHECK(cudaMemcpy(d, order.data(), sizeof(int) * order.size(), cudaMemcpyHostToDevice));
      H->DTileOrder = d;
      std::vector<int> order2(order.begin(), order.begin() + H->NTilesBoundary);
      order2.insert(order2.end(), order.begin() + nOwned, order.end());
      order2.insert(order2.end(), order.begin() + H->NTilesBoundary, order.begin() + nOwned);
      OMEGA_CUDA_CHECK(cudaMalloc((void **)&d, sizeof(int) * order2.size()));
      H->Allocs.push_back(d);
      OMEGA_CUDA_CHECK(cudaMemcpy(d, order2.data(), sizeof(int) * order2.size(), cudaMemcpyHostToDevice));
      H->DTileOrderBH = d;
      // every local non-owned cell / edge must be covered by the receive lists, or skipping the
      // halo tiles in the stages before an exchange would leave stale rows
      H->HaloCovered = H->RecvOff[0][D->NNeighbors] == M->D.NCellsAll - M->D.NCellsOwned &&
                       H->RecvOff[1][D->NNeighbors] == M->D.NEdgesAll - M->D.NEdgesOwned;
   }
   {
      int dev = 0;
      OMEGA_CUDA_CHECK(cudaGetDevice(&dev));
      OMEGA_CUDA_CHECK(cudaDeviceGetAttribute(&H->NumSMs, cudaDevAttrMultiProcessorCount, dev));
   }
   int lo = 0, hi = 0;
   OMEGA_CUDA_CHECK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
   OMEGA_CUDA_CHECK(cudaStreamCreateWithPriority(&H->CommStream, cudaStreamNonBlocking, hi));
   OMEGA_CUDA_CHECK(cudaEventCreateWithFlags(&H->EvReady, cudaEventDisableTiming));
   OMEGA_CUDA_CHECK(cudaEventCreateWithFlags(&H->EvDone, cudaEventDisableTiming));
   *Out = H;
   return 0;
}

int omega_halo_destroy(OmegaHalo *H) {
   if (!H) return 0;
   cudaDeviceSynchronize();
   if (H->Comm && H->OwnsComm && g_nccl.CommDestroy) g_nccl.CommDestroy((ncclComm_t)H->Comm);
   for (void *p : H->Allocs) cudaFree(p);
   if (H->SendBuf) cudaFree(H->SendBuf);
   if (H->RecvBuf) cudaFree(H->RecvBuf);
   for (auto &t : H->Events) { cudaEventDestroy(t.Start); cudaEventDestroy(t.Stop); }
   if (H->CommStream) cudaStreamDestroy(H->CommStream);
   if (H->EvReady) cudaEventDestroy(H->EvReady);
   if (H->EvDone) cudaEventDestroy(H->EvDone);
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

int omega_halo_pack(const OmegaHalo *H, OmegaMeshElement Elem, int NOuter, int NInner, const Real *Array,
                    Real *SendBuf, void *Stream) {
   OMEGA_REQUIRE(H && Array && SendBuf && Elem >= 0 && Elem < 3, "omega_halo_pack: bad argument");
   const int NN = H->NNeighbors;
   const long long total = (long long)H->SendOff[Elem][NN] * NOuter * NInner;
   if (total == 0) return 0;
   const int *d = H->DSendIdx[Elem];
   k_halo_rows<true><<<grid_for(total), kBlock, 0, (cudaStream_t)Stream>>>(
       NN, d, d + NN + 1, NOuter, n_size(H->M, Elem), NInner, const_cast<Real *>(Array), SendBuf);
   OMEGA_CUDA_CHECK(cudaGetLastError());
   return 0;
}

int omega_halo_unpack(const OmegaHalo *H, OmegaMeshElement Elem, int NOuter, int NInner, const Real *RecvBuf,
                      Real *Array, void *Stream) {
   OMEGA_REQUIRE(H && Array && RecvBuf && Elem >= 0 && Elem < 3, "omega_halo_unpack: bad argument");
   const int NN = H->NNeighbors;
   const long long total = (long long)H->RecvOff[Elem][NN] * NOuter * NInner;
   if (total == 0) return 0;
   const int *d = H->DRecvIdx[Elem];
   k_halo_rows<false><<<grid_for(total), kBlock, 0, (cudaStream_t)Stream>>>(
       NN, d, d + NN + 1, NOuter, n_size(H->M, Elem), NInner, Array, const_cast<Real *>(RecvBuf));
   OMEGA_CUDA_CHECK(cudaGetLastError());
   return 0;
}

// ---- arrays of any element type (Halo.h:62-917 exchanges I4, I8, R4, R8 arrays of 1 to 5 dimensions) ----------
// An array [d0]..[element][last] is NOuter = product of the dimensions in front of the element dimension and
// NInner = the one behind it (1 for a 1-D array); elements are moved as 4- or 8-byte words.
extern "C++" {
template <bool Pack>
static int rows_any(const OmegaHalo *H, int Elem, int NOuter, int NInner, int ElemBytes, void *Array, void *Buf, cudaStream_t St) {
   const int NN = H->NNeighbors;
   const std::vector<int> &off = Pack ? H->SendOff[Elem] : H->RecvOff[Elem];
   const long long total = (long long)off[NN] * NOuter * NInner;
   if (total == 0) return 0;
   const int *d = Pack ? H->DSendIdx[Elem] : H->DRecvIdx[Elem];
   if (ElemBytes == 8)
      k_halo_rows<Pack, unsigned long long><<<grid_for(total), kBlock, 0, St>>>(NN, d, d + NN + 1, NOuter, n_size(H->M, Elem), NInner,
                                                                                 (unsigned long long *)Array, (unsigned long long *)Buf);
   else
      k_halo_rows<Pack, unsigned><<<grid_for(total), kBlock, 0, St>>>(NN, d, d + NN + 1, NOuter, n_size(H->M, Elem), NInner,
                                                                      (unsigned *)Array, (unsigned *)Buf);
   OMEGA_CUDA_CHECK(cudaGetLastError());
   return 0;
}
} // extern "C++"

int omega_halo_pack_any(const OmegaHalo *H, OmegaMeshElement Elem, int NOuter, int NInner, int ElemBytes, const void *Array,
                        void *SendBuf, void *Stream) {
   OMEGA_REQUIRE(H && Array && SendBuf && Elem >= 0 && Elem < 3 && NOuter >= 1 && NInner >= 1 && (ElemBytes == 4 || ElemBytes == 8),
                 "omega_halo_pack_any: bad argument (elements of 4 or 8 bytes)");
   return rows_any<true>(H, Elem, NOuter, NInner, ElemBytes, const_cast<void *>(Array), SendBuf, (cudaStream_t)Stream);
}

int omega_halo_unpack_any(const OmegaHalo *H, OmegaMeshElement Elem, int NOuter, int NInner, int ElemBytes, const void *RecvBuf,
                          void *Array, void *Stream) {
   OMEGA_REQUIRE(H && Array && RecvBuf && Elem >= 0 && Elem < 3 && NOuter >= 1 && NInner >= 1 && (ElemBytes == 4 || ElemBytes == 8),
                 "omega_halo_unpack_any: bad argument (elements of 4 or 8 bytes)");
   return rows_any<false>(H, Elem, NOuter, NInner, ElemBytes, Array, const_cast<void *>(RecvBuf), (cudaStream_t)Stream);
}

int omega_halo_exchange_any(OmegaHalo *H, void *Array, OmegaMeshElement Elem, int NOuter, int NInner, int ElemBytes, void *Stream) {
   OMEGA_REQUIRE(H && Array && Elem >= 0 && Elem < 3 && NOuter >= 1 && NInner >= 1 && (ElemBytes == 4 || ElemBytes == 8),
                 "omega_halo_exchange_any: bad argument (elements of 4 or 8 bytes)");
   OMEGA_REQUIRE(H->NNeighbors == 0 || H->Comm, "omega_halo_exchange_any: no communicator (omega_halo_comm_init)");
   const int NN = H->NNeighbors;
   if (NN == 0) return 0;
   cudaStream_t St = (cudaStream_t)Stream;
   const size_t per = (size_t)NOuter * NInner * ElemBytes; // bytes per list entry
   const size_t sb = (size_t)H->SendOff[Elem][NN] * per, rb = (size_t)H->RecvOff[Elem][NN] * per;
   // the library's round buffers, sized in Reals
   if (ensure_buffer(H, &H->SendBuf, &H->SendCap, (long long)((sb + sizeof(Real) - 1) / sizeof(Real))) ||
       ensure_buffer(H, &H->RecvBuf, &H->RecvCap, (long long)((rb + sizeof(Real) - 1) / sizeof(Real))))
      return 1;
   if (rows_any<true>(H, Elem, NOuter, NInner, ElemBytes, Array, H->SendBuf, St)) return 1;
   ncclComm_t comm = (ncclComm_t)H->Comm;
   OMEGA_NCCL_CHECK(g_nccl.GroupStart());
   for (int n = 0; n < NN; ++n) {
      const size_t s = (size_t)(H->SendOff[Elem][n + 1] - H->SendOff[Elem][n]) * per, so = (size_t)H->SendOff[Elem][n] * per;
      const size_t r = (size_t)(H->RecvOff[Elem][n + 1] - H->RecvOff[Elem][n]) * per, ro = (size_t)H->RecvOff[Elem][n] * per;
      const int peer = H->NeighborTasks[n];
      if (s) OMEGA_NCCL_CHECK(g_nccl.Send((const char *)H->SendBuf + so, s, ncclChar, peer, comm, St));
      if (r) OMEGA_NCCL_CHECK(g_nccl.Recv((char *)H->RecvBuf + ro, r, ncclChar, peer, comm, St));
   }
   OMEGA_NCCL_CHECK(g_nccl.GroupEnd());
   return rows_any<false>(H, Elem, NOuter, NInner, ElemBytes, Array, H->RecvBuf, St);
}

// ---- exchange rounds ---------------------------------------------------------------
int omega_halo_round_counts(const OmegaHalo *H, const int *Elems, const int *NOuter, int N, long long *SendCounts,
                            long long *RecvCounts) {
   OMEGA_REQUIRE(H && Elems && NOuter && N >= 1 && N <= kHaloMaxArrays, "omega_halo_round_counts: bad argument");
   for (int n = 0; n < H->NNeighbors; ++n) {
      long long s, r;
      round_counts(H, Elems, NOuter, N, n, s, r);
      if (SendCounts) SendCounts[n] = s;
      if (RecvCounts) RecvCounts[n] = r;
   }
   return 0;
}

int omega_halo_pack_round(OmegaHalo *H, Real *const *Ptrs, const int *Elems, const int *NOuter, int N,
                          Real **SendBuf, void *Stream) {
   OMEGA_REQUIRE(H && Ptrs && Elems && NOuter, "omega_halo_pack_round: null argument");
   int rc = halo_pack_round(H, Ptrs, Elems, NOuter, N, (cudaStream_t)Stream);
   if (SendBuf) *SendBuf = H->SendBuf;
   return rc;
}

int omega_halo_recv_buffer(OmegaHalo *H, const int *Elems, const int *NOuter, int N, Real **RecvBuf) {
   OMEGA_REQUIRE(H && Elems && NOuter && RecvBuf, "omega_halo_recv_buffer: null argument");
   long long tr = 0;
   for (int n = 0; n < H->NNeighbors; ++n) {
      long long s, r;
      round_counts(H, Elems, NOuter, N, n, s, r);
      tr += r;
   }
   if (ensure_buffer(H, &H->RecvBuf, &H->RecvCap, std::max(tr, 1LL))) return 1;
   *RecvBuf = H->RecvBuf;
   return 0;
}

int omega_halo_unpack_round(OmegaHalo *H, Real *const *Ptrs, const int *Elems, const int *NOuter, int N, void *Stream) {
   OMEGA_REQUIRE(H && Ptrs && Elems && NOuter, "omega_halo_unpack_round: null argument");
   return halo_unpack_round(H, Ptrs, Elems, NOuter, N, (cudaStream_t)Stream);
}

int omega_halo_nccl_unique_id(char *Id128) {
   OMEGA_REQUIRE(Id128, "omega_halo_nccl_unique_id: null argument");
   if (load_nccl()) return 1;
   static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
   ncclUniqueId id;
   OMEGA_NCCL_CHECK(g_nccl.GetUniqueId(&id));
   std::memcpy(Id128, &id, sizeof(id));
   return 0;
}

int omega_halo_comm_init(OmegaHalo *H, const char *Id128, int NRanks, int Rank) {
   OMEGA_REQUIRE(H && Id128 && NRanks >= 1 && Rank >= 0 && Rank < NRanks, "omega_halo_comm_init: bad argument");
   OMEGA_REQUIRE(Rank == H->MyTask, "omega_halo_comm_init: Rank must equal the MyTask of the halo lists");
   if (load_nccl()) return 1;
   ncclUniqueId id;
   std::memcpy(&id, Id128, sizeof(id));
   ncclComm_t comm = nullptr;
   OMEGA_NCCL_CHECK(g_nccl.CommInitRank(&comm, NRanks, id, Rank));
   H->Comm = comm;
   H->OwnsComm = 1;
   H->NRanks = NRanks;
   return 0;
}

int omega_halo_set_comm(OmegaHalo *H, void *NcclComm, int NRanks) {
   OMEGA_REQUIRE(H && NcclComm, "omega_halo_set_comm: null argument");
   if (load_nccl()) return 1;
   H->Comm = NcclComm;
   H->OwnsComm = 0;
   H->NRanks = NRanks;
   return 0;
}

int omega_halo_exchange_round(OmegaHalo *H, Real *const *Ptrs, const int *Elems, const int *NOuter, int N, void *Stream) {
   OMEGA_REQUIRE(H && Ptrs && Elems && NOuter, "omega_halo_exchange_round: null argument");
   return halo_exchange_round(H, Ptrs, Elems, NOuter, N, (cudaStream_t)Stream);
}

int omega_halo_exchange(OmegaHalo *H, Real *Array, OmegaMeshElement Elem, int NOuter, void *Stream) {
   OMEGA_REQUIRE(H && Array, "omega_halo_exchange: null argument");
   Real *p[1] = {Array};
   int e[1] = {(int)Elem}, o[1] = {NOuter};
   return halo_exchange_round(H, p, e, o, 1, (cudaStream_t)Stream);
}

int omega_halo_tile_classes(const OmegaHalo *H, int *NBoundary, int *NOwned, int *NTiles) {
   OMEGA_REQUIRE(H, "omega_halo_tile_classes: null argument");
   if (NBoundary) *NBoundary = H->NTilesBoundary;
   if (NOwned) *NOwned = H->NTilesOwned;
   if (NTiles) *NTiles = (int)H->M->TileC0.size();
   return 0;
}

int omega_halo_set_profiling(OmegaHalo *H, int Enable) {
   OMEGA_REQUIRE(H, "omega_halo_set_profiling: null argument");
   H->Profiling = Enable;
   return 0;
}

int omega_halo_get_profile(OmegaHalo *H, double *Ms, long long *Rounds, long long *Bytes) {
   OMEGA_REQUIRE(H && Ms && Rounds && Bytes, "omega_halo_get_profile: null argument");
   OMEGA_CUDA_CHECK(cudaDeviceSynchronize());
   *Ms = 0; *Rounds = 0; *Bytes = 0;
   for (auto &t : H->Events) {
      float ms = 0;
      OMEGA_CUDA_CHECK(cudaEventElapsedTime(&ms, t.Start, t.Stop));
      *Ms += ms; *Rounds += 1; *Bytes += t.Bytes;
      cudaEventDestroy(t.Start);
      cudaEventDestroy(t.Stop);
   }
   H->Events.clear();
   return 0;
}

} // extern "C"
