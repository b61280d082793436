// OmegaB200.h -- C++ host mirror of the reference's hot-path classes over the
// C ABI of include/omega_b200.h.  Same class names, method names, argument
// meaning and error behaviour (int 0 = ok / abort on unusable state) as
// components/omega: HorzMesh (src/ocn/HorzMesh.h), OceanState
// (src/ocn/OceanState.h:113-129), Tracers (src/ocn/Tracers.h),
// AuxiliaryState (src/ocn/AuxiliaryState.h:74-82), Tendencies
// (src/ocn/Tendencies.h:73-102), TimeStepper (src/timeStepping/TimeStepper.h:82-84),
// Halo (src/base/Halo.h:768-771).  Header-only; link libomega_b200.so.
// Arrays are device allocations in DEVICE order behind a small View-like
// handle (pointer + extents + label), the role Kokkos::View plays upstream.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <map>
#include <memory>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

#include "omega_b200.h"

namespace OMEGA {

using I4   = int;
using R8   = double;
using Real = OmegaReal;
static_assert(sizeof(Real) == sizeof(double), "this mirror header is written for the fp64 build (libomega_b200.so); "
                                              "the single-precision library is driven through the same C ABI with float arrays");

[[noreturn]] inline void abortError(const std::string &Msg) { // ABORT_ERROR, src/infra/Error.h:207-210
   std::fprintf(stderr, "[omega_b200] %s: %s\n", Msg.c_str(), omega_last_error());
   std::abort();
}
#define OMEGA_B200_CHECK(call, what) \
   do { if ((call) != 0) ::OMEGA::abortError(what); } while (0)

enum MeshElement { OnCell = OMEGA_ON_CELL, OnEdge = OMEGA_ON_EDGE, OnVertex = OMEGA_ON_VERTEX };

/// The ownership convention of the reference's classes (e.g. src/ocn/OceanState.cpp:30-62,369-407,
/// src/ocn/Tendencies.cpp:75-121, src/ocn/HorzMesh.cpp:29-47,651-690): objects are made by a static create(Name, ...)
/// that refuses a name in use, live in a static map, are found again with get(Name) / getDefault() (the one
/// called "Default") and end with erase(Name) / clear().  T::create(Name, args...) builds T(Name, args...) where T
/// has such a constructor, T(args...) otherwise.  Plain construction stays possible (the tests use both).
template <class T> class Registry {
 public:
   template <class... A> static T *create(const std::string &Name, A &&...Args) {
      if (all().count(Name)) {
         std::fprintf(stderr, "[omega_b200] Attempted to create an object with name %s but it already exists\n", Name.c_str());
         return nullptr;
      }
      T *Obj;
      if constexpr (std::is_constructible_v<T, const std::string &, A...>) Obj = new T(Name, std::forward<A>(Args)...);
      else Obj = new T(std::forward<A>(Args)...);
      all()[Name].reset(Obj);
      if (Name == "Default") defaultPtr() = Obj;
      return Obj;
   }
   static T *getDefault() { return defaultPtr(); }
   static T *get(const std::string &Name) {
      auto It = all().find(Name);
      if (It == all().end()) {
         std::fprintf(stderr, "[omega_b200] Attempt to retrieve non-existent object: %s\n", Name.c_str());
         return nullptr;
      }
      return It->second.get();
   }
   static void erase(const std::string &Name) {
      auto It = all().find(Name);
      if (It == all().end()) return;
      if (It->second.get() == defaultPtr()) defaultPtr() = nullptr;
      all().erase(It);
   }
   /// objects go in reverse order of creation is NOT promised: erase dependants first, as upstream's finalize does
   static void clear() {
      all().clear();
      defaultPtr() = nullptr;
   }

 private:
   static std::map<std::string, std::unique_ptr<T>> &all() {
      static std::map<std::string, std::unique_ptr<T>> M;
      return M;
   }
   static T *&defaultPtr() {
      static T *D = nullptr;
      return D;
   }
};

/// Device array handle: ref-counted, zero-initialised (like Kokkos::View)
template <int Rank> class DeviceArray {
 public:
   DeviceArray() = default;
   DeviceArray(const std::string &Label, std::initializer_list<long> Ext) : Label_(Label) {
      int i = 0;
      Size_ = 1;
      for (long e : Ext) { Ext_[i++] = e; Size_ *= e; }
      void *p = nullptr;
      OMEGA_B200_CHECK(omega_dev_alloc(&p, (unsigned long long)Size_ * sizeof(double)), "DeviceArray alloc " + Label);
      Data_ = std::shared_ptr<double>(static_cast<double *>(p), [](double *q) { omega_dev_free(q); });
   }
   /// non-owning alias of library-owned memory (tendency / aux arrays)
   DeviceArray(const std::string &Label, double *Ptr, std::initializer_list<long> Ext) : Label_(Label) {
      int i = 0;
      Size_ = 1;
      for (long e : Ext) { Ext_[i++] = e; Size_ *= e; }
      Data_ = std::shared_ptr<double>(Ptr, [](double *) {});
   }
   /// [i][...] of this array as an array of one rank less that SHARES the allocation: the role of
   /// Kokkos::subview(A, i, ALL, ALL) in Tracers::getByIndex (src/ocn/Tracers.cpp:292-293)
   DeviceArray<Rank - 1> slice(long i, const std::string &Label = "") const {
      static_assert(Rank >= 2, "slice needs a leading dimension");
      DeviceArray<Rank - 1> Out;
      long Inner = 1;
      for (int d = 1; d < Rank; ++d) { Out.Ext_[d - 1] = Ext_[d]; Inner *= Ext_[d]; }
      Out.Size_  = Inner;
      Out.Label_ = Label.empty() ? Label_ : Label;
      Out.Data_  = std::shared_ptr<double>(Data_, Data_.get() + i * Inner); // aliasing constructor: same owner
      return Out;
   }
   double *data() const { return Data_.get(); }
   long extent(int i) const { return Ext_[i]; }
   int extent_int(int i) const { return (int)Ext_[i]; }
   long size() const { return Size_; }
   const std::string &label() const { return Label_; }

 private:
   template <int R2> friend class DeviceArray;
   std::string Label_;
   std::shared_ptr<double> Data_;
   long Ext_[Rank > 0 ? Rank : 1] = {0};
   long Size_ = 0;
};
using Array1DReal = DeviceArray<1>;
using Array2DReal = DeviceArray<2>;
using Array3DReal = DeviceArray<3>;

/// HorzMesh: device-resident mesh built from host-order arrays
class HorzMesh : public Registry<HorzMesh> {
 public:
   I4 NCellsOwned, NCellsAll, NCellsSize, NEdgesOwned, NEdgesAll, NEdgesSize;
   I4 NVerticesOwned, NVerticesAll, NVerticesSize, MaxEdges, MaxEdges2, VertexDegree, NVertLayers;
   OmegaMesh *Handle = nullptr;

   HorzMesh(const OmegaMeshDesc &D, bool Reorder = true) {
      NCellsOwned = D.NCellsOwned; NCellsAll = D.NCellsAll; NCellsSize = D.NCellsSize;
      NEdgesOwned = D.NEdgesOwned; NEdgesAll = D.NEdgesAll; NEdgesSize = D.NEdgesSize;
      NVerticesOwned = D.NVerticesOwned; NVerticesAll = D.NVerticesAll; NVerticesSize = D.NVerticesSize;
      MaxEdges = D.MaxEdges; MaxEdges2 = D.MaxEdges2; VertexDegree = D.VertexDegree; NVertLayers = D.NVertLayers;
      OMEGA_B200_CHECK(omega_mesh_create(&D, Reorder ? 1 : 0, &Handle), "HorzMesh: omega_mesh_create");
   }
   ~HorzMesh() { omega_mesh_destroy(Handle); }
   HorzMesh(const HorzMesh &) = delete;
   I4 size(MeshElement E) const { return E == OnCell ? NCellsSize : E == OnEdge ? NEdgesSize : NVerticesSize; }
   /// host-order -> device-order copy (OceanState::copyToDevice) and back
   void copyToDevice(MeshElement E, int NOuter, int NInner, const double *Host, double *Dev, void *Stream = nullptr) const {
      OMEGA_B200_CHECK(omega_field_to_device(Handle, (OmegaMeshElement)E, NOuter, NInner, Host, Dev, Stream), "copyToDevice");
   }
   void copyToHost(MeshElement E, int NOuter, int NInner, const double *Dev, double *Host, void *Stream = nullptr) const {
      OMEGA_B200_CHECK(omega_field_to_host(Handle, (OmegaMeshElement)E, NOuter, NInner, Dev, Host, Stream), "copyToHost");
   }
};

/// Halo: exchange lists on device and the exchange itself (src/base/Halo.h:62-917): pack kernel,
/// NCCL send/recv with every neighbour, unpack kernel, all inside the library.  The host only
/// bootstraps the communicator: one rank calls uniqueId(), the 128 bytes travel by the host's own
/// means (MPI_Bcast over MachEnv's communicator in Omega) and every rank calls initComm().
class Halo : public Registry<Halo> {
 public:
   OmegaHalo *Handle = nullptr;
   const HorzMesh *Mesh;
   std::vector<int> Neighbors;

   Halo(const HorzMesh *M, const OmegaHaloDesc &D) : Mesh(M), Neighbors(D.NeighborTasks, D.NeighborTasks + D.NNeighbors) {
      OMEGA_B200_CHECK(omega_halo_create(M->Handle, &D, &Handle), "Halo: omega_halo_create");
   }
   ~Halo() { omega_halo_destroy(Handle); }
   Halo(const Halo &) = delete;

   static void uniqueId(char (&Id)[128]) { OMEGA_B200_CHECK(omega_halo_nccl_unique_id(Id), "Halo: omega_halo_nccl_unique_id"); }
   void initComm(const char (&Id)[128], int NRanks, int Rank) {
      OMEGA_B200_CHECK(omega_halo_comm_init(Handle, Id, NRanks, Rank), "Halo: omega_halo_comm_init");
   }

   /// Halo::exchangeFullArrayHalo(Array, Elem), src/base/Halo.h:768-915.  The
   /// element dimension is second to last, as upstream.
   template <int Rank> int exchangeFullArrayHalo(DeviceArray<Rank> &Array, MeshElement Elem, void *Stream = nullptr) {
      static_assert(Rank == 2 || Rank == 3, "state arrays are [elem][level] or [tracer][elem][level]");
      const int NOuter = Rank == 3 ? Array.extent_int(0) : 1;
      if (Neighbors.empty()) return 0;
      return omega_halo_exchange(Handle, Array.data(), (OmegaMeshElement)Elem, NOuter, Stream) ? -1 : 0;
   }
   /// The other instantiations of the reference's template (I4, I8, R4, R8 arrays of one to five dimensions,
   /// test/base/HaloTest.cpp:41-99): Ptr is the device array [NOuter][N*Size][NInner], NOuter the product of the
   /// dimensions in front of the element dimension, NInner the one behind it (1 for a 1-D array).
   template <class T> int exchangeFullArrayHalo(T *Ptr, MeshElement Elem, int NOuter, int NInner, void *Stream = nullptr) {
      static_assert(sizeof(T) == 4 || sizeof(T) == 8, "elements of 4 or 8 bytes");
      if (Neighbors.empty()) return 0;
      return omega_halo_exchange_any(Handle, Ptr, (OmegaMeshElement)Elem, NOuter, NInner, (int)sizeof(T), Stream) ? -1 : 0;
   }
};

/// OceanState: h and u time-level rings (src/ocn/OceanState.cpp:65-117,333-407)
class OceanState : public Registry<OceanState> {
 public:
   std::vector<Array2DReal> LayerThickness, NormalVelocity;
   I4 NTimeLevels, NVertLayers, CurTimeIndex = 0;
   const HorzMesh *Mesh;
   Halo *MeshHalo;

   OceanState(const std::string &Name, const HorzMesh *M, Halo *H, int NVertLayers_, int NTimeLevels_)
       : NTimeLevels(NTimeLevels_), NVertLayers(NVertLayers_), Mesh(M), MeshHalo(H) {
      for (int i = 0; i < NTimeLevels; ++i) {
         LayerThickness.emplace_back("LayerThickness" + Name + std::to_string(i), std::initializer_list<long>{M->NCellsSize, NVertLayers});
         NormalVelocity.emplace_back("NormalVelocity" + Name + std::to_string(i), std::initializer_list<long>{M->NEdgesSize, NVertLayers});
      }
   }
   /// one time level over arrays somebody else owns: the state a custom tendency is handed for a stage
   /// (the reference passes its one-level ProvisState there, RungeKutta4Stepper.cpp:57-58)
   OceanState(const HorzMesh *M, double *H, double *U, int NVertLayers_) : NTimeLevels(1), NVertLayers(NVertLayers_), Mesh(M), MeshHalo(nullptr) {
      LayerThickness.emplace_back("LayerThicknessStage", H, std::initializer_list<long>{M->NCellsSize, NVertLayers});
      NormalVelocity.emplace_back("NormalVelocityStage", U, std::initializer_list<long>{M->NEdgesSize, NVertLayers});
   }
   I4 getTimeIndex(I4 &TimeIndex, I4 TimeLevel) const { // OceanState.cpp:394-407
      if (NTimeLevels > 1 && (TimeLevel > 1 || (TimeLevel + NTimeLevels) <= 1)) return -1;
      TimeIndex = (TimeLevel + CurTimeIndex + NTimeLevels) % NTimeLevels;
      return 0;
   }
   I4 getLayerThickness(Array2DReal &Out, I4 TimeLevel) const {
      I4 i; I4 Err = getTimeIndex(i, TimeLevel);
      if (!Err) Out = LayerThickness[i];
      return Err;
   }
   I4 getNormalVelocity(Array2DReal &Out, I4 TimeLevel) const {
      I4 i; I4 Err = getTimeIndex(i, TimeLevel);
      if (!Err) Out = NormalVelocity[i];
      return Err;
   }
   I4 copyToDevice(I4 TimeLevel, const double *H, const double *U) {
      I4 i; if (getTimeIndex(i, TimeLevel)) return -1;
      Mesh->copyToDevice(OnCell, 1, NVertLayers, H, LayerThickness[i].data());
      Mesh->copyToDevice(OnEdge, 1, NVertLayers, U, NormalVelocity[i].data());
      return 0;
   }
   I4 copyToHost(I4 TimeLevel, double *H, double *U) const {
      I4 i; if (getTimeIndex(i, TimeLevel)) return -1;
      Mesh->copyToHost(OnCell, 1, NVertLayers, LayerThickness[i].data(), H);
      Mesh->copyToHost(OnEdge, 1, NVertLayers, NormalVelocity[i].data(), U);
      return 0;
   }
   I4 exchangeHalo(I4 TimeLevel) { // OceanState.cpp:333-344
      I4 i; if (getTimeIndex(i, TimeLevel)) return -1;
      if (!MeshHalo) return 0;
      I4 Err = MeshHalo->exchangeFullArrayHalo(LayerThickness[i], OnCell);
      Err |= MeshHalo->exchangeFullArrayHalo(NormalVelocity[i], OnEdge);
      return Err;
   }
   void updateTimeLevels() { // OceanState.cpp:348-365
      if (NTimeLevels == 1) abortError("OceanState: can't update time levels for NTimeLevels == 1");
      exchangeHalo(1);
      CurTimeIndex = (CurTimeIndex + 1) % NTimeLevels;
   }
};

/// Tracer names and indices (src/ocn/Tracers.cpp:135-190: define; :244-263: getIndex / getName).  The reference
/// fills them from the Tracers: group of the configuration (Base: [Temperature, Salinity], ...); here the host hands
/// the names over, in index order.
class TracerNames {
 public:
   static constexpr I4 IndxInvalid = -1; // Tracers.h:43
   void define(const std::vector<std::string> &InNames) {
      Names = InNames;
      Indexes.clear();
      for (I4 i = 0; i < (I4)Names.size(); ++i) Indexes[Names[i]] = i;
   }
   I4 getIndex(I4 &TracerIndex, const std::string &TracerName) const {
      auto It = Indexes.find(TracerName);
      if (It == Indexes.end()) {
         std::fprintf(stderr, "[omega_b200] Tracers: Tracer index for '%s' is not found.\n", TracerName.c_str());
         return -1;
      }
      TracerIndex = It->second;
      return 0;
   }
   I4 getName(std::string &TracerName, I4 TracerIndex) const {
      if (TracerIndex < 0 || TracerIndex >= (I4)Names.size()) {
         std::fprintf(stderr, "[omega_b200] Tracers: Tracer name for index '%d' is not found.\n", TracerIndex);
         return -1;
      }
      TracerName = Names[TracerIndex];
      return 0;
   }
   I4 size() const { return (I4)Names.size(); }

 private:
   std::vector<std::string> Names;
   std::map<std::string, I4> Indexes;
};

/// Tracers: [NTracers][NCellsSize][NVertLayers] per time level (src/ocn/Tracers.cpp:269-277,457-525).
/// An instance instead of the reference's static class; NTracers is a plain integer.
class Tracers {
 public:
   std::vector<Array3DReal> TracerArrays;
   I4 NTracers, NTimeLevels, CurTimeIndex = 0;
   Halo *MeshHalo;
   TracerNames Names;
   /// InNames: one name per tracer in index order; left out, the tracers are called Tracer0, Tracer1, ...
   Tracers(const HorzMesh *M, Halo *H, int NTracers_, int NVertLayers, int NTimeLevels_, std::vector<std::string> InNames = {})
       : NTracers(NTracers_), NTimeLevels(NTimeLevels_), MeshHalo(H) {
      if (InNames.empty())
         for (int i = 0; i < NTracers; ++i) InNames.push_back("Tracer" + std::to_string(i));
      if ((int)InNames.size() != NTracers) abortError("Tracers: one name per tracer");
      Names.define(InNames);
      for (int i = 0; i < NTimeLevels; ++i)
         TracerArrays.emplace_back("Tracers" + std::to_string(i), std::initializer_list<long>{NTracers, M->NCellsSize, NVertLayers});
   }
   I4 getNumTracers() const { return NTracers; }
   I4 getIndex(I4 &TracerIndex, const std::string &TracerName) const { return Names.getIndex(TracerIndex, TracerName); }
   I4 getName(std::string &TracerName, I4 TracerIndex) const { return Names.getName(TracerName, TracerIndex); }
   /// Tracers.cpp:279-295: one tracer of a time level, sharing the allocation
   I4 getByIndex(Array2DReal &Out, I4 TimeLevel, I4 TracerIndex) const {
      if (TracerIndex < 0 || TracerIndex >= NTracers) return -2;
      I4 i; I4 Err = getTimeIndex(i, TimeLevel);
      if (!Err) Out = TracerArrays[i].slice(TracerIndex);
      return Err;
   }
   /// Tracers.cpp:297-310
   I4 getByName(Array2DReal &Out, I4 TimeLevel, const std::string &TracerName) const {
      I4 Idx;
      if (Names.getIndex(Idx, TracerName)) return -1;
      return getByIndex(Out, TimeLevel, Idx) ? -2 : 0;
   }
   I4 getTimeIndex(I4 &TimeIndex, I4 TimeLevel) const {
      if (NTimeLevels > 1 && (TimeLevel > 1 || (TimeLevel + NTimeLevels) <= 1)) return -1;
      TimeIndex = (TimeLevel + CurTimeIndex + NTimeLevels) % NTimeLevels;
      return 0;
   }
   I4 getAll(Array3DReal &Out, I4 TimeLevel) const {
      I4 i; I4 Err = getTimeIndex(i, TimeLevel);
      if (!Err) Out = TracerArrays[i];
      return Err;
   }
   I4 exchangeHalo(I4 TimeLevel) {
      I4 i; if (getTimeIndex(i, TimeLevel)) return -1;
      return MeshHalo ? MeshHalo->exchangeFullArrayHalo(TracerArrays[i], OnCell) : 0;
   }
   void updateTimeLevels() {
      exchangeHalo(1);
      CurTimeIndex = (CurTimeIndex + 1) % NTimeLevels;
   }
};

} // namespace OMEGA

#include "OmegaB200Stepper.h"
