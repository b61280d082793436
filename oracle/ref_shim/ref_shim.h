// ref_shim.h -- TEST INFRASTRUCTURE.  Force-included (-include) in front of the
// reference's own sources so that its hot-path files compile UNMODIFIED where
// they lie under /root/reference although Kokkos, MPI, spdlog, cpptrace,
// yaml-cpp, SCORPIO and METIS are not available here.
//
// It (1) pre-defines the include guards of the reference headers that drag in
// those dependencies and (2) supplies small stand-ins with the same names and
// the members the compiled sources touch.  Everything numerical -- the
// functors (TendencyTerms.h, auxiliaryVars/*.h, HorzOperators.h), their
// constructors, AuxiliaryState.cpp, Tendencies.cpp, CustomTendencyTerms.cpp,
// OceanState.cpp, TimeStepper.cpp, the three steppers and TimeMgr.cpp -- is the
// reference's code.  Stand-ins that carry behaviour (HorzMesh derived fields,
// Tracers time-level ring) restate the cited reference lines.
#pragma once

// ---- guards of reference headers replaced by the stand-ins below -------------
#define OMEGA_CONFIG_H
#define OMEGA_HALO_H
#define OMEGA_HORZMESH_H
#define OMEGA_TRACERS_H
#ifndef REF_SHIM_REAL_VERTCOORD // (ref_vertcoord.cpp compiles the reference's own VertCoord.cpp)
#define OMEGA_VERTCOORD_H
#endif
#define OMEGA_DECOMP_H
#define OMEGA_MACH_ENV_H
#define OMEGA_LOG_H
#define OMEGA_LOG_FORMATTERS_H
#define OMEGA_ERROR_H
#define OMEGA_FIELD_H
#define OMEGA_IO_H
#define OMEGA_IOSTREAM_H
#define OMEGA_DIMENSION_H

#include <Kokkos_Core.hpp>

#include <cstdio>
#include <cstdlib>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include "DataTypes.h"   // reference's own (src/base/DataTypes.h)
#include "OmegaKokkos.h" // reference's own (src/infra/OmegaKokkos.h)
#include "mpi.h"         // stand-in: MPI_Comm only

// ---- Logging.h / Error.h (spdlog, cpptrace) ------------------------------------
#define LOG_TRACE(...) ((void)0)
#define LOG_DEBUG(...) ((void)0)
#define LOG_INFO(...) ((void)0)
#define LOG_WARN(...) ((void)0)
#define LOG_ERROR(...) ((void)0)
#define LOG_CRITICAL(...) (std::fprintf(stderr, "omega_ref: critical error at %s:%d\n", __FILE__, __LINE__))

namespace OMEGA {

enum class ErrorCode { Success = 0, Warn, Fail, Critical };

class Error {
 public:
   ErrorCode Code = ErrorCode::Success;
   std::string Msg;
   Error() = default;
   template <class... A> Error(ErrorCode C, const std::string &M, A &&...) : Code(C), Msg(M) {}
   template <class... A> Error(ErrorCode C, int, const char *, const std::string &M, A &&...) : Code(C), Msg(M) {}
   bool isSuccess() const { return Code == ErrorCode::Success; }
   bool isFail() const { return Code == ErrorCode::Fail || Code == ErrorCode::Critical; }
   void reset() { Code = ErrorCode::Success; Msg.clear(); }
   Error &operator+=(const Error &O) {
      if (static_cast<int>(O.Code) > static_cast<int>(Code)) Code = O.Code;
      if (!O.Msg.empty()) Msg += (Msg.empty() ? "" : "; ") + O.Msg;
      return *this;
   }
   [[noreturn]] static void abort() { std::abort(); }
};

} // namespace OMEGA

#define ABORT_ERROR(_Msg, ...)                                                              \
   {                                                                                        \
      std::fprintf(stderr, "omega_ref: ABORT_ERROR %s (%s:%d)\n", _Msg, __FILE__, __LINE__); \
      OMEGA::Error::abort();                                                                \
   }
#define OMEGA_ASSERT(_Condition, _ErrMsg, ...)
#define OMEGA_REQUIRE(_Condition, _ErrMsg, ...) \
   if (!(_Condition)) ABORT_ERROR(_ErrMsg)
#define RETURN_ERROR(_ReturnErr, _ErrCode, _ErrMsg, ...)            \
   {                                                                \
      OMEGA::Error _NewErr(_ErrCode, __LINE__, __FILE__, _ErrMsg);  \
      return (_ReturnErr += _NewErr);                               \
   }
#define CHECK_ERROR(_ErrChk, _NewMsg, ...) \
   if (!_ErrChk.isSuccess()) { _ErrChk.reset(); }
#define CHECK_ERROR_WARN(_ErrChk, _NewMsg, ...) \
   if (!_ErrChk.isSuccess()) { _ErrChk.reset(); }
#define CHECK_ERROR_ABORT(_ErrChk, _NewMsg, ...) \
   if (_ErrChk.isFail()) ABORT_ERROR(_NewMsg)

namespace OMEGA {

// ---- MachEnv.h --------------------------------------------------------------------
#ifndef OMEGA_VECTOR_LENGTH
#define OMEGA_VECTOR_LENGTH 1
#endif
constexpr int VecLength = OMEGA_VECTOR_LENGTH; // src/base/MachEnv.h:32

class MachEnv {
 public:
   static MachEnv *getDefault() {
      static MachEnv E;
      return &E;
   }
   MPI_Comm getComm() const { return 0; }
   int getMyTask() const { return 0; }
   int getNumTasks() const { return 1; }
};

// ---- Config.h (yaml-cpp): flat key -> value store filled by the driver -------------
class Config {
 public:
   std::string Name;
   explicit Config(const std::string &N = "") : Name(N) {}
   static std::map<std::string, std::string> &store() {
      static std::map<std::string, std::string> S;
      return S;
   }
   static Config *getOmegaConfig() {
      static Config C("Omega");
      return &C;
   }
   Error get(Config &) { return Error(); } // every group exists
   template <class T> Error get(const std::string &Key, T &Val) {
      auto It = store().find(Key);
      if (It == store().end()) return Error(ErrorCode::Fail, "Config: key not found: " + Key);
      if constexpr (std::is_same_v<T, std::string>) {
         Val = It->second;
      } else if constexpr (std::is_same_v<T, bool>) {
         Val = (It->second == "true" || It->second == "1");
      } else {
         std::istringstream S(It->second);
         S.precision(17);
         S >> Val;
      }
      return Error();
   }
   template <class T> static void set(const std::string &Key, const T &Val) {
      std::ostringstream S;
      S.precision(17);
      S << Val;
      store()[Key] = S.str();
   }
};

// ---- Field.h / Dimension.h (metadata registry only) -----------------------------------
class Field {
 public:
   template <class... A> static std::shared_ptr<Field> create(A &&...) { return std::make_shared<Field>(); }
   static void destroy(const std::string &) {}
   static std::shared_ptr<Field> get(const std::string &) { return std::make_shared<Field>(); }
   template <class T> void attachData(const T &) {}
   template <class T> static void attachFieldData(const std::string &, const T &) {}
};
class FieldGroup {
 public:
   static std::shared_ptr<FieldGroup> create(const std::string &) { return std::make_shared<FieldGroup>(); }
   static void destroy(const std::string &) {}
   static bool exists(const std::string &) { return true; }
   static void addFieldToGroup(const std::string &, const std::string &) {}
   void addField(const std::string &) {}
};

// ---- Decomp.h ------------------------------------------------------------------------
class Decomp {
 public:
   // sizes the reference's VertCoord constructor reads (VertCoord.cpp:50-128); filled by ref_vertcoord.cpp
   I4 NCellsOwned = 0, NCellsAll = 0, NCellsSize = 0, NEdgesOwned = 0, NEdgesAll = 0, NEdgesSize = 0, NVerticesOwned = 0,
      NVerticesAll = 0, NVerticesSize = 0, VertexDegree = 3, MaxEdges = 6, HaloWidth = 1;
   std::string MeshFileName = "memory";
   Array2DI4 CellsOnEdge, CellsOnVertex;
   static Decomp *getDefault() { return nullptr; }
};

// ---- Halo.h: one task, closed mesh -> the exchange is the identity.  A multi-task test may install a callback:
// the reference's steppers then call it at their own exchange points (OceanState.cpp:333-344,
// RungeKutta4Stepper.cpp:108-114,130-131, Tracers.cpp:457-476) with the array to exchange, and the test moves the
// halo values between its task processes (tests/test_multirank_cpu.py).  Arrays are [NOuter][N*Size][NInner] Real.
enum MeshElement { OnCell, OnEdge, OnVertex }; // src/base/Halo.h:45
class Halo {
 public:
   using Callback = void (*)(void *Data, int Elem, int NOuter, int NInner);
   static Callback &callback() {
      static Callback C = nullptr;
      return C;
   }
   static Halo *getDefault() {
      static Halo H;
      return &H;
   }
   MPI_Comm getComm() const { return 0; }
   template <class T> int exchangeFullArrayHalo(T &A, MeshElement E) {
      if (!callback()) return 0;
      const int R = A.rank();
      const int NOuter = R == 3 ? A.extent_int(0) : 1, NInner = R == 1 ? 1 : A.extent_int(R - 1);
      callback()(static_cast<void *>(A.data()), static_cast<int>(E), NOuter, NInner);
      return 0;
   }
};

#ifdef REF_SHIM_REAL_VERTCOORD
// ---- what the reference's VertCoord.cpp touches besides the stand-ins above (ref_vertcoord.cpp) ----
inline int &refNVertLayers() { // the "nVertLevels" dimension of the mesh "file"
   static int N = 1;
   return N;
}
namespace IO {
enum Mode { ModeUnknown, ModeRead, ModeWrite };
inline void openFile(int &FileID, const std::string &, Mode) { FileID = 1; }
inline void closeFile(int &) {}
inline Error getDimFromFile(int, const std::string &, int &DimID, int &DimLength) {
   DimID     = 0;
   DimLength = refNVertLayers();
   return Error();
}
} // namespace IO
class Dimension {
 public:
   static std::shared_ptr<Dimension> create(const std::string &, int) { return std::make_shared<Dimension>(); }
   static void destroy(const std::string &) {}
};
class IOStream {
 public:
   static std::shared_ptr<IOStream> get(const std::string &) { return std::make_shared<IOStream>(); }
   bool validate() const { return false; } // no InitialVertCoord stream: the driver fills the members itself
   static Error read(const std::string &) { return Error(ErrorCode::Fail, "no stream"); }
   template <class... A> static Error read(const std::string &, A &&...) { return Error(ErrorCode::Fail, "no stream"); }
};
#endif

// ---- VertCoord.h: the hot path only reads NVertLayers -------------------------------------
#ifndef REF_SHIM_REAL_VERTCOORD
class VertCoord {
 public:
   I4 NVertLayers = 1;
   static VertCoord *&defaultPtr() {
      static VertCoord *P = nullptr;
      return P;
   }
   static VertCoord *getDefault() { return defaultPtr(); }
};
#endif

// ---- HorzMesh.h: public members the compiled sources read (src/ocn/HorzMesh.h:100-265) ------
class HorzMesh {
 public:
   std::string MeshName = "Default";
   I4 NCellsOwned, NCellsAll, NCellsSize;
   I4 NEdgesOwned, NEdgesAll, NEdgesSize;
   I4 NVerticesOwned, NVerticesAll, NVerticesSize;
   I4 MaxEdges, MaxEdges2, VertexDegree;
   Array1DI4 NEdgesOnCell, NEdgesOnEdge;
   Array2DI4 EdgesOnCell, CellsOnEdge, VerticesOnEdge, EdgesOnEdge, CellsOnVertex, EdgesOnVertex;
   Array1DReal AreaCell, BottomDepth, XCell, YCell, DcEdge, DvEdge, AngleEdge, FEdge, XEdge, YEdge,
       AreaTriangle, FVertex, MeshScalingDel2, MeshScalingDel4;
   HostArray1DReal BottomDepthH;
   Array2DReal WeightsOnEdge, KiteAreasOnVertex, EdgeSignOnCell, EdgeSignOnVertex, EdgeMask;
   static HorzMesh *&defaultPtr() {
      static HorzMesh *P = nullptr;
      return P;
   }
   static HorzMesh *getDefault() { return defaultPtr(); }
};

// ---- Tracers.h: static arrays + time-level ring (src/ocn/Tracers.cpp:269-277,457-525) --------
class Tracers {
 public:
   static std::vector<Array3DReal> &arrays() {
      static std::vector<Array3DReal> A;
      return A;
   }
   static I4 &numTracers() {
      static I4 N = 0;
      return N;
   }
   static I4 &curTimeIndex() {
      static I4 I = 0;
      return I;
   }
   static I4 &numTimeLevels() {
      static I4 N = 2;
      return N;
   }
   static I4 getNumTracers() { return numTracers(); }
   static I4 getTimeIndex(I4 &TimeIndex, const I4 TimeLevel) {
      const I4 NTL = numTimeLevels();
      if (NTL > 1 && (TimeLevel > 1 || (TimeLevel + NTL) <= 1)) return -1;
      TimeIndex = (TimeLevel + curTimeIndex() + NTL) % NTL;
      return 0;
   }
   static I4 getAll(Array3DReal &TracerArray, const I4 TimeLevel) {
      I4 TimeIndex;
      I4 Err      = getTimeIndex(TimeIndex, TimeLevel);
      TracerArray = arrays()[TimeIndex];
      return Err;
   }
   static I4 exchangeHalo(const I4 TimeLevel) { // Tracers.cpp:457-470
      I4 TimeIndex;
      getTimeIndex(TimeIndex, TimeLevel);
      if (numTracers() > 0) Halo::getDefault()->exchangeFullArrayHalo(arrays()[TimeIndex], OnCell);
      return 0;
   }
   static void updateTimeLevels() {
      exchangeHalo(1);
      curTimeIndex() = (curTimeIndex() + 1) % numTimeLevels();
   }
};

} // namespace OMEGA
