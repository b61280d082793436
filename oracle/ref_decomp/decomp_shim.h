// decomp_shim.h -- TEST INFRASTRUCTURE.  Force-included (-include) in front of the reference's own
// src/base/Decomp.cpp so that it compiles UNMODIFIED where it lies under /root/reference without MPI, SCORPIO,
// METIS, yaml-cpp and spdlog, and runs as N tasks = N THREADS of one process:
//   MachEnv      one object per thread (task id, task count);
//   Broadcast    rendezvous of the N threads on a shared buffer (the only communication Decomp.cpp uses);
//   IO::         the "mesh file" is the global connectivity handed in by the driver (file values, 1-based, 0 = none);
//                readArray gathers through the offsets of the linear decomposition Decomp.cpp itself builds;
//   METIS_PartGraphKway   returns the partition handed in by the driver (METIS is third party and absent; the
//                result on owned elements does not depend on the partition).
//   mpi.h        (this directory) the few MPI calls of src/base/Halo.cpp over the same threads.
// Everything that numbers cells, edges and vertices, builds the halos, the local connectivity and the exchange lists
// is the reference's code.  Used by tests/test_decomp_vs_reference.py to pin omega_b200/csrc/decomp.cpp and the halo
// lists on several tasks.
#pragma once

#define OMEGA_MACH_ENV_H
#define OMEGA_BROADCAST_H
#define OMEGA_IO_H
#define OMEGA_CONFIG_H
#define OMEGA_ERROR_H
#define OMEGA_LOG_H
#define OMEGA_LOG_FORMATTERS_H
#define OMEGA_DIMENSION_H
#define OMEGA_VERTCOORD_H // HorzMesh.h includes it; HorzMesh::init (compiled, never called) reads NVertLayers from it

#include <Kokkos_Core.hpp>

#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "DataTypes.h"   // reference's own
#include "OmegaKokkos.h" // reference's own
#include "mpi.h"         // stand-in: MPI_Comm only

#define LOG_TRACE(...) ((void)0)
#define LOG_DEBUG(...) ((void)0)
#define LOG_INFO(...) ((void)0)
#define LOG_WARN(...) ((void)0)
#define LOG_ERROR(...) ((void)0)
#define LOG_CRITICAL(...) (std::fprintf(stderr, "omega_ref_decomp: critical error at %s:%d\n", __FILE__, __LINE__))

namespace OMEGA {

enum class ErrorCode { Success = 0, Warn, Fail, Critical };
class Error {
 public:
   ErrorCode Code = ErrorCode::Success;
   Error() = default;
   explicit Error(ErrorCode C) : Code(C) {}
   bool isSuccess() const { return Code == ErrorCode::Success; }
   bool isFail() const { return Code == ErrorCode::Fail || Code == ErrorCode::Critical; }
   void reset() { Code = ErrorCode::Success; }
};
} // namespace OMEGA
#define ABORT_ERROR(_Msg, ...)                                                                      \
   {                                                                                                \
      std::fprintf(stderr, "omega_ref_decomp: ABORT_ERROR %s (%s:%d)\n", _Msg, __FILE__, __LINE__);  \
      std::abort();                                                                                 \
   }
#define CHECK_ERROR_ABORT(_ErrChk, _NewMsg, ...) \
   if (_ErrChk.isFail()) ABORT_ERROR(_NewMsg)

namespace OMEGA {

// ---- MachEnv.h: one per thread ---------------------------------------------------------------------
class MachEnv {
 public:
   int MyTask = 0, NumTasks = 1;
   MachEnv() = default;
   MachEnv(int T, int N) : MyTask(T), NumTasks(N) {}
   static MachEnv *getDefault() {
      static MachEnv E;
      return &E;
   }
   MPI_Comm getComm() const { return 0; }
   int getMyTask() const { return MyTask; }
   int getNumTasks() const { return NumTasks; }
   int getMasterTask() const { return 0; }
   bool isMasterTask() const { return MyTask == 0; }
};

// ---- Broadcast.h: the N threads meet on the root's vector ---------------------------------------------
struct RefDecompWorld { // set by the driver before the threads start
   int NumTasks = 1;
   std::mutex Mu;
   std::condition_variable Cv;
   int Arrived = 0;
   long Generation = 0;
   const std::vector<I4> *Root = nullptr;
   static RefDecompWorld &get() {
      static RefDecompWorld W;
      return W;
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
inline void Broadcast(std::vector<I4> &Value, const MachEnv *InEnv, const int RankBcast) {
   RefDecompWorld &W = RefDecompWorld::get();
   if (W.NumTasks == 1) return;
   if (InEnv->getMyTask() == RankBcast) W.Root = &Value;
   W.barrier(); // the root's vector is published
   if (InEnv->getMyTask() != RankBcast) Value = *W.Root;
   W.barrier(); // everybody has copied: the root may go on and change it
}

// ---- Config.h: only Decomp::init (compiled, never called here) touches it -------------------------------
class Config {
 public:
   explicit Config(const std::string & = "") {}
   static Config *getOmegaConfig() {
      static Config C;
      return &C;
   }
   Error get(Config &) { return Error(); }
   template <class T> Error get(const std::string &, T &) { return Error(ErrorCode::Fail); }
};

// ---- IO.h: the mesh "file" is memory ----------------------------------------------------------------------
namespace IO {
enum Mode { ModeUnknown, ModeRead, ModeWrite };
enum FileFmt { FmtDefault };
enum class IfExists { Fail, Replace, Append };
enum Rearranger { RearrBox, RearrSubset };
enum IODataType { IOTypeI4, IOTypeI8, IOTypeR4, IOTypeR8, IOTypeChar, IOTypeLogical };

struct RefMeshFile { // global arrays as a mesh file holds them: 1-based IDs, 0 = none
   std::map<std::string, int> Dims;
   std::map<std::string, const int *> Arrays;
   std::map<std::string, const double *> ArraysR8; // geometry of the global mesh (HorzMesh.cpp:420-520)
   std::mutex Mu;
   std::map<int, std::vector<int>> Decomps;
   std::map<int, int> DecompType;
   int NextDecomp = 1;
   static RefMeshFile &get() {
      static RefMeshFile F;
      return F;
   }
};
inline void openFile(int &FileID, const std::string &, Mode, FileFmt = FmtDefault, IfExists = IfExists::Fail) { FileID = 1; }
inline void closeFile(int &) {}
inline Error getDimFromFile(int, const std::string &DimName, int &DimID, int &DimLength) {
   RefMeshFile &F = RefMeshFile::get();
   auto it = F.Dims.find(DimName);
   if (it == F.Dims.end()) {
      DimLength = -1;
      return Error(ErrorCode::Fail);
   }
   DimID = 0;
   DimLength = it->second;
   return Error();
}
inline int createDecomp(IODataType T, int, const std::vector<int> &, int, const std::vector<int> &GlobalIndx, Rearranger) {
   RefMeshFile &F = RefMeshFile::get();
   std::lock_guard<std::mutex> L(F.Mu);
   const int id = F.NextDecomp++;
   F.Decomps[id]    = GlobalIndx;
   F.DecompType[id] = (int)T;
   return id;
}
inline void destroyDecomp(int &DecompID) {
   RefMeshFile &F = RefMeshFile::get();
   std::lock_guard<std::mutex> L(F.Mu);
   F.Decomps.erase(DecompID);
}
inline Error readArray(void *Array, int Size, const std::string &VarName, int, int DecompID, int &VarID, int = -1) {
   RefMeshFile &F = RefMeshFile::get();
   std::vector<int> Idx;
   int Type;
   {
      std::lock_guard<std::mutex> L(F.Mu);
      Idx  = F.Decomps.at(DecompID);
      Type = F.DecompType.at(DecompID);
   }
   if (Type == (int)IOTypeR8) {
      auto r = F.ArraysR8.find(VarName);
      if (r == F.ArraysR8.end()) return Error(ErrorCode::Fail);
      double *Out = static_cast<double *>(Array);
      for (int i = 0; i < Size && i < (int)Idx.size(); ++i)
         if (Idx[i] >= 0) Out[i] = r->second ? r->second[Idx[i]] : 0.0; // (a null entry: a variable the test does not supply)
      VarID = 0;
      return Error();
   }
   auto it = F.Arrays.find(VarName);
   if (it == F.Arrays.end()) return Error(ErrorCode::Fail);
   int *Out = static_cast<int *>(Array);
   for (int i = 0; i < Size; ++i) // entries past a task's share of the linear distribution keep offset -1
      if (Idx[i] >= 0) Out[i] = it->second[Idx[i]];
   VarID = 0;
   return Error();
}
} // namespace IO

// ---- Dimension.h, VertCoord.h: names only ---------------------------------------------------------------------
class Dimension {
 public:
   template <class... A> static std::shared_ptr<Dimension> create(A &&...) { return std::make_shared<Dimension>(); }
   static bool exists(const std::string &) { return false; }
   static void destroy(const std::string &) {}
};
class VertCoord {
 public:
   I4 NVertLayers = 1;
   static VertCoord *getDefault() {
      static VertCoord V;
      return &V;
   }
};
} // namespace OMEGA
