// OmegaB200Vertical.h -- C++ host mirror of the Omega-1 column pieces over the C ABI of
// include/omega_b200.h: same class and method names and argument meaning as
//   src/ocn/VertCoord.h:28-209   (computePressure / computeZHeight / computeGeopotential / computeTargetThickness),
//   src/ocn/Eos.h:272-328        (computeSpecVol / computeSpecVolDisp, EosType),
//   src/base/TriDiagSolvers.h    (the array forms of PCRSolver::solve and PCRDiffusionSolver::solve),
// each method one C entry point.  Arrays are DeviceArray handles in DEVICE order (HorzMesh::copyToDevice).
#pragma once
#include <vector>

#include "OmegaB200.h"

namespace OMEGA {

/// VertCoord: the active layer range of every column and the diagnostic column integrals
class VertCoord {
 public:
   const HorzMesh *Mesh;
   I4 NVertLayers, NVertLayersP1;
   R8 Rho0 = 1026.0; ///< constants of VertCoord.cpp:656-659 (Rho0 from the configuration there)
   // public members of the reference class (VertCoord.h:71-118), device order
   Array2DReal PressureInterface, PressureMid, ZInterface, ZMid, GeopotentialMid, LayerThicknessTarget;
   Array1DReal VertCoordMovementWeights;
   Array2DReal RefLayerThickness;
   int *MinLayerCell = nullptr, *MaxLayerCell = nullptr; ///< [NCellsSize] on the device, device order

   /// MinLayerCellH / MaxLayerCellH: host order, as read from the mesh file (VertCoord.cpp completeSetup)
   VertCoord(const HorzMesh *M, int NVertLayers_, const int *MinLayerCellH, const int *MaxLayerCellH)
       : Mesh(M), NVertLayers(NVertLayers_), NVertLayersP1(NVertLayers_ + 1),
         PressureInterface("PressureInterface", {M->NCellsSize, NVertLayers_ + 1}), PressureMid("PressureMid", {M->NCellsSize, NVertLayers_}),
         ZInterface("ZInterface", {M->NCellsSize, NVertLayers_ + 1}), ZMid("ZMid", {M->NCellsSize, NVertLayers_}),
         GeopotentialMid("GeopotentialMid", {M->NCellsSize, NVertLayers_}),
         LayerThicknessTarget("LayerThicknessTarget", {M->NCellsSize, NVertLayers_}),
         VertCoordMovementWeights("VertCoordMovementWeights", {NVertLayers_}),
         RefLayerThickness("RefLayerThickness", {M->NCellsSize, NVertLayers_}) {
      const int n = M->NCellsSize;
      std::vector<int> perm(n), iperm(n), lo(n), hi(n);
      OMEGA_B200_CHECK(omega_mesh_perm(M->Handle, OMEGA_ON_CELL, perm.data(), iperm.data()), "VertCoord: omega_mesh_perm");
      for (int i = 0; i < n; ++i) { lo[i] = MinLayerCellH[perm[i]]; hi[i] = MaxLayerCellH[perm[i]]; }
      void *p = nullptr;
      OMEGA_B200_CHECK(omega_dev_alloc(&p, 2ull * n * sizeof(int)), "VertCoord: alloc layer ranges");
      MinLayerCell = static_cast<int *>(p);
      MaxLayerCell = MinLayerCell + n;
      OMEGA_B200_CHECK(omega_dev_upload(MinLayerCell, lo.data(), (unsigned long long)n * sizeof(int)), "VertCoord: upload");
      OMEGA_B200_CHECK(omega_dev_upload(MaxLayerCell, hi.data(), (unsigned long long)n * sizeof(int)), "VertCoord: upload");
   }
   ~VertCoord() { omega_dev_free(MinLayerCell); }
   VertCoord(const VertCoord &) = delete;

   /// VertCoord.cpp:648-690
   void computePressure(const Array2DReal &LayerThickness, const Array1DReal &SurfacePressure, void *Stream = nullptr) {
      OMEGA_B200_CHECK(omega_vert_compute_pressure(Mesh->Handle, MinLayerCell, MaxLayerCell, Rho0, LayerThickness.data(),
                                                   SurfacePressure.data(), PressureInterface.data(), PressureMid.data(), Stream),
                       "VertCoord::computePressure");
   }
   /// VertCoord.cpp:697-735
   void computeZHeight(const Array2DReal &LayerThickness, const Array2DReal &SpecVol, void *Stream = nullptr) {
      OMEGA_B200_CHECK(omega_vert_compute_z_height(Mesh->Handle, MinLayerCell, MaxLayerCell, Rho0, LayerThickness.data(), SpecVol.data(),
                                                   ZInterface.data(), ZMid.data(), Stream),
                       "VertCoord::computeZHeight");
   }
   /// VertCoord.cpp:742-778
   void computeGeopotential(const Array1DReal &TidalPotential, const Array1DReal &SelfAttractionLoading, void *Stream = nullptr) {
      OMEGA_B200_CHECK(omega_vert_compute_geopotential(Mesh->Handle, MinLayerCell, MaxLayerCell, ZMid.data(), TidalPotential.data(),
                                                       SelfAttractionLoading.data(), GeopotentialMid.data(), Stream),
                       "VertCoord::computeGeopotential");
   }
   /// VertCoord.cpp:786-851
   void computeTargetThickness(void *Stream = nullptr) {
      OMEGA_B200_CHECK(omega_vert_compute_target_thickness(Mesh->Handle, MinLayerCell, MaxLayerCell, Rho0, PressureInterface.data(),
                                                           RefLayerThickness.data(), VertCoordMovementWeights.data(),
                                                           LayerThicknessTarget.data(), Stream),
                       "VertCoord::computeTargetThickness");
   }
};

/// Eos.h:24-27
enum class EosType { LinearEos, Teos10Eos };

/// Eos: specific volume from conservative temperature, absolute salinity and pressure
class Eos {
 public:
   const HorzMesh *Mesh;
   EosType EosChoice = EosType::Teos10Eos;
   R8 DRhodT = -0.2, DRhodS = 0.8, RhoT0S0 = 1000.0; ///< linear coefficients (Default.yml Eos: Linear:)
   Array2DReal SpecVol, SpecVolDisplaced;            ///< Eos.h:281-282

   Eos(const HorzMesh *M, int NVertLayers)
       : Mesh(M), SpecVol("SpecVol", {M->NCellsSize, NVertLayers}), SpecVolDisplaced("SpecVolDisplaced", {M->NCellsSize, NVertLayers}) {}

   /// Eos.cpp:113-141
   void computeSpecVol(const Array2DReal &ConservTemp, const Array2DReal &AbsSalinity, const Array2DReal &Pressure, void *Stream = nullptr) {
      run(ConservTemp, AbsSalinity, Pressure, 0, SpecVol, Stream);
   }
   /// Eos.cpp:144-180: the parcel displaced adiabatically by KDisp layers
   void computeSpecVolDisp(const Array2DReal &ConservTemp, const Array2DReal &AbsSalinity, const Array2DReal &Pressure, I4 KDisp,
                           void *Stream = nullptr) {
      run(ConservTemp, AbsSalinity, Pressure, KDisp, SpecVolDisplaced, Stream);
   }

 private:
   void run(const Array2DReal &Ct, const Array2DReal &Sa, const Array2DReal &P, I4 KDisp, Array2DReal &Out, void *Stream) {
      OMEGA_B200_CHECK(omega_eos_compute_spec_vol(Mesh->Handle, EosChoice == EosType::Teos10Eos ? 1 : 0, Ct.data(), Sa.data(), P.data(),
                                                  KDisp, DRhodT, DRhodS, RhoT0S0, Out.data(), Stream),
                       "Eos::computeSpecVol");
   }
};

/// TriDiagSolvers.h:217-245: NBatch general systems, arrays [NBatch][NRow]; X holds the right-hand side, then the solution
struct PCRSolver {
   static void solve(const Array2DReal &DL, const Array2DReal &D, const Array2DReal &DU, Array2DReal &X, void *Stream = nullptr) {
      OMEGA_B200_CHECK(omega_tridiag_solve(X.extent_int(0), X.extent_int(1), DL.data(), D.data(), DU.data(), X.data(), Stream),
                       "PCRSolver::solve");
   }
};
/// TriDiagSolvers.h:458-484: diffusion-type systems D = H + G(k-1) + G(k), off-diagonals -G
struct PCRDiffusionSolver {
   static void solve(const Array2DReal &G, const Array2DReal &H, Array2DReal &X, void *Stream = nullptr) {
      OMEGA_B200_CHECK(omega_tridiag_diffusion_solve(X.extent_int(0), X.extent_int(1), G.data(), H.data(), X.data(), Stream),
                       "PCRDiffusionSolver::solve");
   }
};
using TriDiagSolver     = PCRSolver;          // the reference's device choice (TriDiagSolvers.h:27-37)
using TriDiagDiffSolver = PCRDiffusionSolver;

} // namespace OMEGA
