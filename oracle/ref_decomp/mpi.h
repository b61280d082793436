// mpi.h -- TEST INFRASTRUCTURE: the handful of MPI calls src/base/Halo.cpp and Halo.h make, over the THREADS of one
// process (one thread = one task; the task id is thread-local, set by ref_decomp_driver.cpp).  Sends are eager copies
// into a mailbox; receives match (source, tag or MPI_ANY_TAG) in arrival order, as MPI's non-overtaking rule requires.
#pragma once
#include <condition_variable>
#include <cstring>
#include <list>
#include <mutex>
#include <vector>

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef struct { int MPI_SOURCE, MPI_TAG; } MPI_Status;
#define MPI_COMM_WORLD 0
#define MPI_SUCCESS 0
#define MPI_INT 4
#define MPI_FLOAT 5
#define MPI_DOUBLE 8
#define MPI_LONG_LONG 9
#define MPI_ANY_TAG (-1)
#define MPI_STATUS_IGNORE ((MPI_Status *)0)
#define MPI_STATUSES_IGNORE ((MPI_Status *)0)

struct RefMpiReq {
   bool Recv = false, Done = true;
   void *Buf = nullptr;
   size_t Bytes = 0;
   int Src = 0, Tag = 0;
};
typedef RefMpiReq *MPI_Request;

struct RefMpiWorld {
   struct Msg { int Src, Dst, Tag; std::vector<char> Data; };
   int NumTasks = 1;
   std::mutex Mu;
   std::condition_variable Cv;
   std::list<Msg> Box;
   std::vector<const void *> A2A;
   int Arrived = 0;
   long Generation = 0;
   static RefMpiWorld &get() {
      static RefMpiWorld W;
      return W;
   }
   static int &rank() {
      static thread_local int R = 0;
      return R;
   }
   void barrier() {
      std::unique_lock<std::mutex> L(Mu);
      const long g = Generation;
      if (++Arrived == NumTasks) {
         Arrived = 0;
         ++Generation;
         Cv.notify_all();
      } else
         Cv.wait(L, [&] { return Generation != g; });
   }
};
inline size_t ref_mpi_size(MPI_Datatype T) { return T == MPI_INT || T == MPI_FLOAT ? 4 : 8; }

inline int MPI_Isend(const void *Buf, int Count, MPI_Datatype T, int Dest, int Tag, MPI_Comm, MPI_Request *Req) {
   RefMpiWorld &W = RefMpiWorld::get();
   RefMpiWorld::Msg M{RefMpiWorld::rank(), Dest, Tag, {}};
   M.Data.assign((const char *)Buf, (const char *)Buf + (size_t)Count * ref_mpi_size(T));
   {
      std::lock_guard<std::mutex> L(W.Mu);
      W.Box.push_back(std::move(M));
   }
   W.Cv.notify_all();
   *Req = new RefMpiReq();
   return MPI_SUCCESS;
}
inline int MPI_Irecv(void *Buf, int Count, MPI_Datatype T, int Source, int Tag, MPI_Comm, MPI_Request *Req) {
   RefMpiReq *R = new RefMpiReq();
   R->Recv = true; R->Done = false; R->Buf = Buf; R->Bytes = (size_t)Count * ref_mpi_size(T); R->Src = Source; R->Tag = Tag;
   *Req = R;
   return MPI_SUCCESS;
}
// (caller holds W.Mu)
inline bool ref_mpi_try(RefMpiWorld &W, RefMpiReq *R) {
   if (R->Done) return true;
   const int me = RefMpiWorld::rank();
   for (auto it = W.Box.begin(); it != W.Box.end(); ++it)
      if (it->Dst == me && it->Src == R->Src && (R->Tag == MPI_ANY_TAG || R->Tag == it->Tag)) {
         std::memcpy(R->Buf, it->Data.data(), std::min(R->Bytes, it->Data.size()));
         W.Box.erase(it);
         R->Done = true;
         return true;
      }
   return false;
}
inline int MPI_Test(MPI_Request *Req, int *Flag, MPI_Status *) {
   RefMpiWorld &W = RefMpiWorld::get();
   std::lock_guard<std::mutex> L(W.Mu);
   *Flag = ref_mpi_try(W, *Req) ? 1 : 0;
   return MPI_SUCCESS;
}
inline int MPI_Waitall(int N, MPI_Request *Reqs, MPI_Status *) {
   RefMpiWorld &W = RefMpiWorld::get();
   std::unique_lock<std::mutex> L(W.Mu);
   for (int i = 0; i < N; ++i)
      while (!ref_mpi_try(W, Reqs[i])) W.Cv.wait(L);
   return MPI_SUCCESS;
}
inline int MPI_Wait(MPI_Request *Req, MPI_Status *S) { return MPI_Waitall(1, Req, S); }
inline int MPI_Alltoall(const void *Send, int SCount, MPI_Datatype ST, void *Recv, int RCount, MPI_Datatype RT, MPI_Comm) {
   RefMpiWorld &W = RefMpiWorld::get();
   const int me = RefMpiWorld::rank(), n = W.NumTasks;
   {
      std::lock_guard<std::mutex> L(W.Mu);
      if ((int)W.A2A.size() != n) W.A2A.assign(n, nullptr);
      W.A2A[me] = Send;
   }
   W.barrier();
   const size_t sb = (size_t)SCount * ref_mpi_size(ST), rb = (size_t)RCount * ref_mpi_size(RT);
   for (int t = 0; t < n; ++t) std::memcpy((char *)Recv + t * rb, (const char *)W.A2A[t] + me * sb, std::min(sb, rb));
   W.barrier();
   return MPI_SUCCESS;
}
