// ref_vertcoord.cpp -- TEST INFRASTRUCTURE.  The reference's own src/ocn/VertCoord.cpp, compiled unmodified over
// ref_shim.h (with REF_SHIM_REAL_VERTCOORD: no stand-in VertCoord; IO / Dimension / IOStream reduced to what the
// constructor touches; Kokkos team loops = one-thread teams in league order, Kokkos_Core.hpp), and its four column
// integrals -- computePressure, computeZHeight, computeGeopotential, computeTargetThickness (VertCoord.cpp:648-860) --
// called on plain arrays.  Pins oracle/vertical_oracle.c (tests/test_vertical.py).
#define private public // the constructor is private (create() stores the object in a static map)
#include "VertCoord.h"
#undef private

using namespace OMEGA;

extern "C" int ref_vertcoord_run(int NCells, int K, const int *KMin, const int *KMax, double Rho0, const double *BottomDepth,
                                 const double *H, const double *SurfP, const double *SpecVol, const double *Tidal, const double *Sal,
                                 const double *RefH, const double *Wgt, double *PInt, double *PMid, double *ZInt, double *ZMid,
                                 double *Geop, double *Target) {
   static bool kokkos = false;
   if (!kokkos) {
      Kokkos::initialize();
      kokkos = true;
   }
   refNVertLayers() = K;
   Decomp D;
   D.NCellsOwned = D.NCellsAll = NCells;
   D.NCellsSize  = NCells + 1;
   D.NEdgesOwned = D.NEdgesAll = 0; D.NEdgesSize = 1;
   D.NVerticesOwned = D.NVerticesAll = 0; D.NVerticesSize = 1;
   D.CellsOnEdge   = Array2DI4("CellsOnEdge", 1, 2);
   D.CellsOnVertex = Array2DI4("CellsOnVertex", 1, 3);
   VertCoord V("RefVertCoord", &D);
   V.Rho0 = Rho0;
   V.VertCoordMovementWeights = Array1DReal("VertCoordMovementWeights", K);
   for (int k = 0; k < K; ++k) V.VertCoordMovementWeights(k) = Wgt[k];
   Array2DReal Hh("h", NCells + 1, K), Sv("specvol", NCells + 1, K);
   Array1DReal Sp("surfp", NCells + 1), Td("tidal", NCells + 1), Sl("sal", NCells + 1);
   for (int i = 0; i < NCells; ++i) {
      V.MinLayerCell(i) = KMin[i];
      V.MaxLayerCell(i) = KMax[i];
      V.BottomDepth(i)  = BottomDepth[i];
      Sp(i) = SurfP[i]; Td(i) = Tidal[i]; Sl(i) = Sal[i];
      for (int k = 0; k < K; ++k) {
         Hh(i, k) = H[(size_t)i * K + k];
         Sv(i, k) = SpecVol[(size_t)i * K + k];
         V.RefLayerThickness(i, k) = RefH[(size_t)i * K + k];
      }
   }
   V.computePressure(Hh, Sp);
   V.computeZHeight(Hh, Sv);
   V.computeGeopotential(Td, Sl);
   V.computeTargetThickness();
   for (int i = 0; i < NCells; ++i) {
      for (int k = 0; k <= K; ++k) {
         PInt[(size_t)i * (K + 1) + k] = V.PressureInterface(i, k);
         ZInt[(size_t)i * (K + 1) + k] = V.ZInterface(i, k);
      }
      for (int k = 0; k < K; ++k) {
         PMid[(size_t)i * K + k]   = V.PressureMid(i, k);
         ZMid[(size_t)i * K + k]   = V.ZMid(i, k);
         Geop[(size_t)i * K + k]   = V.GeopotentialMid(i, k);
         Target[(size_t)i * K + k] = V.LayerThicknessTarget(i, k);
      }
   }
   return 0;
}
