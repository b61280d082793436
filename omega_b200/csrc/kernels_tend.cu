// kernels_tend.cu -- Tendencies kernels that fill the public tendency arrays
// (behind Tendencies::compute{Thickness,Velocity,Tracer}TendenciesOnly,
// components/omega/src/ocn/Tendencies.cpp:257-486).  The reference launches
// one kernel per enabled term, each re-reading and re-writing the tendency
// array; here every term of one equation is applied in ONE pass, in the
// reference's term order, so the array is written once.
#include "kernel_util.cuh"
#include "omega_dev.h"

namespace omega {

// deepCopy(0) + ThicknessFluxDivOnCell (src/ocn/TendencyTerms.h:35-58) +
// ManufacturedThicknessTendency (src/ocn/CustomTendencyTerms.cpp:107-140)
template <int V>
__global__ void __launch_bounds__(kBlock)
k_thick_tend(DMesh M, int FluxDivOn, const Real *__restrict__ Flux, const Real *__restrict__ U,
             Real *__restrict__ HTend, ManufacturedParams MS) {
   int c, k;
   if (!elem_k<V>(M.NCellsAll, M.K, c, k)) return;
   Real tend[V] = {0};
   if (FluxDivOn) {
      const Real InvA = 1.0 / M.AreaCell[c];
      const int n       = M.NEdgesOnCell[c];
      Real d[V] = {0};
      for (int j = 0; j < n; ++j) {
         const int e     = M.EdgesOnCell[c * M.MaxEdges + j];
         const Real cf = M.DvEdge[e] * M.EdgeSignOnCell[c * M.MaxEdges + j];
         Real f[V], u[V];
         ldv<V>(f, Flux + (size_t)e * M.K + k);
         ldv<V>(u, U + (size_t)e * M.K + k);
#pragma unroll
         for (int q = 0; q < V; ++q) d[q] -= cf * f[q] * u[q] * InvA;
      }
#pragma unroll
      for (int q = 0; q < V; ++q) tend[q] -= d[q];
   }
   if (MS.Enabled) {
      const Real Phase = MS.Kx * M.XCell[c] + MS.Ky * M.YCell[c] - MS.AngFreq * MS.Time;
      const Real src   = MS.Eta0 * (-MS.H0 * (MS.Kx + MS.Ky) * sin(Phase) - MS.AngFreq * cos(Phase) +
                                    MS.Eta0 * (MS.Kx + MS.Ky) * cos(2.0 * Phase));
#pragma unroll
      for (int q = 0; q < V; ++q) tend[q] += src;
   }
   stv<V>(HTend + (size_t)c * M.K + k, tend);
}

// deepCopy(0) + PotentialVortHAdvOnEdge, KEGradOnEdge, SSHGradOnEdge,
// VelocityDiffusionOnEdge, VelocityHyperDiffOnEdge, WindForcingOnEdge,
// BottomDragOnEdge (src/ocn/TendencyTerms.h:81-334) +
// ManufacturedVelocityTendency (src/ocn/CustomTendencyTerms.cpp:145-210),
// applied in the order of Tendencies.cpp:301-425.
template <int V>
__global__ void __launch_bounds__(kBlock)
k_vel_tend(DMesh M, OmegaTendConfig G, AuxPtrs A, const Real *__restrict__ U,
           const Real *__restrict__ URaw, Real *__restrict__ UTend, ManufacturedParams MS) {
   int e, k;
   if (!elem_k<V>(M.NEdgesAll, M.K, e, k)) return;
   const size_t o    = (size_t)e * M.K + k;
   const Real mask = M.EdgeMask[e];
   const int2 cc = M.CellsOnEdge[e], vv = M.VerticesOnEdge[e];
   Real tend[V] = {0};
   if (G.PVTendencyEnable) {
      Real qe[V], fe[V], acc[V] = {0};
      ldv<V>(qe, A.NormRelVortEdge + o);
      ldv<V>(fe, A.NormPlanetVortEdge + o);
      const int n = M.NEdgesOnEdge[e];
      for (int j = 0; j < n; ++j) {
         const int je   = M.EdgesOnEdge[e * M.MaxEdges2 + j];
         const Real w = M.WeightsOnEdge[e * M.MaxEdges2 + j];
         const size_t oj = (size_t)je * M.K + k;
         Real qj[V], fj[V], fl[V], uj[V];
         ldv<V>(qj, A.NormRelVortEdge + oj);
         ldv<V>(fj, A.NormPlanetVortEdge + oj);
         ldv<V>(fl, A.Flux + oj);
         ldv<V>(uj, U + oj);
#pragma unroll
         for (int q = 0; q < V; ++q) {
            const Real NormVort = (qe[q] + fe[q] + qj[q] + fj[q]) * 0.5;
            acc[q] += w * fl[q] * uj[q] * NormVort;
         }
      }
#pragma unroll
      for (int q = 0; q < V; ++q) tend[q] += mask * acc[q];
   }
   const Real InvDc = 1.0 / M.DcEdge[e], InvDv = 1.0 / M.DvEdge[e];
   if (G.KETendencyEnable) {
      Real a0[V], a1[V];
      ldv<V>(a0, A.KE + (size_t)cc.x * M.K + k);
      ldv<V>(a1, A.KE + (size_t)cc.y * M.K + k);
#pragma unroll
      for (int q = 0; q < V; ++q) tend[q] -= mask * (a1[q] - a0[q]) * InvDc;
   }
   if (G.SSHTendencyEnable) {
      const Real Grav = 9.80665;
      Real a0[V], a1[V];
      ldv<V>(a0, A.Ssh + (size_t)cc.x * M.K + k);
      ldv<V>(a1, A.Ssh + (size_t)cc.y * M.K + k);
#pragma unroll
      for (int q = 0; q < V; ++q) tend[q] -= mask * Grav * (a1[q] - a0[q]) * InvDc;
   }
   if (G.VelDiffTendencyEnable) {
      Real d0[V], d1[V], r0[V], r1[V];
      ldv<V>(d0, A.Div + (size_t)cc.x * M.K + k);
      ldv<V>(d1, A.Div + (size_t)cc.y * M.K + k);
      ldv<V>(r0, A.RelVort + (size_t)vv.x * M.K + k);
      ldv<V>(r1, A.RelVort + (size_t)vv.y * M.K + k);
      const Real cf = mask * G.ViscDel2 * M.MeshScalingDel2[e];
#pragma unroll
      for (int q = 0; q < V; ++q) {
         const Real Del2U = ((d1[q] - d0[q]) * InvDc - (r1[q] - r0[q]) * InvDv);
         tend[q] += cf * Del2U;
      }
   }
   if (G.VelHyperDiffTendencyEnable) {
      Real d0[V], d1[V], r0[V], r1[V];
      ldv<V>(d0, A.Del2Div + (size_t)cc.x * M.K + k);
      ldv<V>(d1, A.Del2Div + (size_t)cc.y * M.K + k);
      ldv<V>(r0, A.Del2RelVort + (size_t)vv.x * M.K + k);
      ldv<V>(r1, A.Del2RelVort + (size_t)vv.y * M.K + k);
      const Real cf = mask * G.ViscDel4 * M.MeshScalingDel4[e];
#pragma unroll
      for (int q = 0; q < V; ++q) {
         const Real Del2U = (G.DivFactor * (d1[q] - d0[q]) * InvDc - (r1[q] - r0[q]) * InvDv);
         tend[q] -= cf * Del2U;
      }
   }
   if (G.WindForcingTendencyEnable && k == 0) { // acts on layer 0 only (:294-300)
      const Real InvThick = 1.0 / A.Mean[(size_t)e * M.K];
      tend[0] += mask * InvThick * A.NormalStress[e] / G.Density0;
   }
   if (G.BottomDragTendencyEnable) { // bottom layer only (:323-333)
      const int KBot = M.K - 1;
      if (KBot >= k && KBot < k + V) {
         const Real VelNorm  = sqrt(A.KE[(size_t)cc.x * M.K + KBot] + A.KE[(size_t)cc.y * M.K + KBot]);
         const Real InvThick = 1.0 / A.Mean[(size_t)e * M.K + KBot];
         tend[KBot - k] -= mask * G.BottomDragCoeff * VelNorm * InvThick * URaw[(size_t)e * M.K + KBot];
      }
   }
   if (MS.Enabled) {
      const Real Grav  = 9.80665;
      const Real Phase = MS.Kx * M.XEdge[e] + MS.Ky * M.YEdge[e] - MS.AngFreq * MS.Time;
      const Real Src0  = MS.AngFreq * sin(Phase) - 0.5 * MS.Eta0 * (MS.Kx + MS.Ky) * sin(2.0 * Phase);
      const Real f     = M.FEdge[e];
      Real Uu = MS.Eta0 * ((-f + Grav * MS.Kx) * cos(Phase) + Src0);
      Real Vv = MS.Eta0 * ((f + Grav * MS.Ky) * cos(Phase) + Src0);
      const Real Kx2 = MS.Kx * MS.Kx, Ky2 = MS.Ky * MS.Ky;
      if (G.VelDiffTendencyEnable) {
         Uu += G.ViscDel2 * MS.Eta0 * (Kx2 + Ky2) * cos(Phase);
         Vv += G.ViscDel2 * MS.Eta0 * (Kx2 + Ky2) * cos(Phase);
      }
      if (G.VelHyperDiffTendencyEnable) {
         const Real Kx4 = Kx2 * Kx2, Ky4 = Ky2 * Ky2;
         Uu -= G.ViscDel4 * MS.Eta0 * ((Kx4 + Ky4 + Kx2 * Ky2) * cos(Phase));
         Vv -= G.ViscDel4 * MS.Eta0 * ((Kx4 + Ky4 + Kx2 * Ky2) * cos(Phase));
      }
      const Real a   = M.AngleEdge[e];
      const Real src = cos(a) * Uu + sin(a) * Vv;
#pragma unroll
      for (int q = 0; q < V; ++q) tend[q] += src;
   }
   stv<V>(UTend + o, tend);
}

// deepCopy(0) + TracerHorzAdvOnCell, TracerDiffOnCell, TracerHyperDiffOnCell
// (src/ocn/TendencyTerms.h:349-480) for every tracer in one pass over cells.
template <int V>
__global__ void __launch_bounds__(kBlock)
k_tracer_tend(DMesh M, OmegaTendConfig G, AuxPtrs A, const Real *__restrict__ URaw,
              const Real *__restrict__ Tr, int NTracers, Real *__restrict__ TrTend) {
   int c, k;
   if (!elem_k<V>(M.NCellsAll, M.K, c, k)) return;
   const Real InvA = 1.0 / M.AreaCell[c];
   const int n       = M.NEdgesOnCell[c];
   const size_t CS = (size_t)M.NCellsSize * M.K, ES = (size_t)M.NEdgesSize * M.K;
   for (int l = 0; l < NTracers; ++l) {
      const Real *T = Tr + l * CS, *HTr = A.HTr + l * ES, *D2 = A.Del2Tr + l * CS;
      Real tend[V] = {0};
      if (G.TracerHorzAdvTendencyEnable) {
         Real t[V] = {0};
         for (int j = 0; j < n; ++j) {
            const int e     = M.EdgesOnCell[c * M.MaxEdges + j];
            const Real cf = M.EdgeMask[e] * M.DvEdge[e] * M.EdgeSignOnCell[c * M.MaxEdges + j];
            Real ht[V], u[V];
            ldv<V>(ht, HTr + (size_t)e * M.K + k);
            ldv<V>(u, URaw + (size_t)e * M.K + k);
#pragma unroll
            for (int q = 0; q < V; ++q) t[q] -= cf * ht[q] * u[q] * InvA;
         }
#pragma unroll
         for (int q = 0; q < V; ++q) tend[q] -= t[q];
      }
      if (G.TracerDiffTendencyEnable) {
         Real t[V] = {0};
         for (int j = 0; j < n; ++j) {
            const int e   = M.EdgesOnCell[c * M.MaxEdges + j];
            const int2 cc = M.CellsOnEdge[e];
            const Real RTemp = M.MeshScalingDel2[e] * M.DvEdge[e] / M.DcEdge[e];
            const Real cf    = M.EdgeMask[e] * M.EdgeSignOnCell[c * M.MaxEdges + j] * RTemp;
            Real m[V], t0[V], t1[V];
            ldv<V>(m, A.Mean + (size_t)e * M.K + k);
            ldv<V>(t0, T + (size_t)cc.x * M.K + k);
            ldv<V>(t1, T + (size_t)cc.y * M.K + k);
#pragma unroll
            for (int q = 0; q < V; ++q) t[q] -= cf * m[q] * (t1[q] - t0[q]);
         }
#pragma unroll
         for (int q = 0; q < V; ++q) tend[q] += G.EddyDiff2 * t[q] * InvA;
      }
      if (G.TracerHyperDiffTendencyEnable) {
         Real t[V] = {0};
         for (int j = 0; j < n; ++j) {
            const int e   = M.EdgesOnCell[c * M.MaxEdges + j];
            const int2 cc = M.CellsOnEdge[e];
            const Real RTemp = M.MeshScalingDel4[e] * M.DvEdge[e] / M.DcEdge[e];
            const Real cf    = M.EdgeMask[e] * M.EdgeSignOnCell[c * M.MaxEdges + j] * RTemp;
            Real t0[V], t1[V];
            ldv<V>(t0, D2 + (size_t)cc.x * M.K + k);
            ldv<V>(t1, D2 + (size_t)cc.y * M.K + k);
#pragma unroll
            for (int q = 0; q < V; ++q) t[q] -= cf * (t1[q] - t0[q]);
         }
#pragma unroll
         for (int q = 0; q < V; ++q) tend[q] -= G.EddyDiff4 * t[q] * InvA;
      }
      stv<V>(TrTend + l * CS + (size_t)c * M.K + k, tend);
   }
}

void launch_thick_tend(const Launch &L, const DMesh &M, const OmegaTendConfig &G, const AuxPtrs &A,
                       const Real *U, Real *HTend, const ManufacturedParams &MS) {
   OMEGA_LAUNCH_V(k_thick_tend, M.NCellsAll, M.K, L.Stream, M, G.ThicknessFluxTendencyEnable, A.Flux, U, HTend, MS);
   ++*L.Count;
}
void launch_vel_tend(const Launch &L, const DMesh &M, const OmegaTendConfig &G, const AuxPtrs &A,
                     const Real *U, const Real *URaw, Real *UTend, const ManufacturedParams &MS) {
   OMEGA_LAUNCH_V(k_vel_tend, M.NEdgesAll, M.K, L.Stream, M, G, A, U, URaw, UTend, MS);
   ++*L.Count;
}
void launch_tracer_tend(const Launch &L, const DMesh &M, const OmegaTendConfig &G, const AuxPtrs &A,
                        const Real *URaw, const Real *Tr, int NTracers, Real *TrTend) {
   if (NTracers == 0) return;
   OMEGA_LAUNCH_V(k_tracer_tend, M.NCellsAll, M.K, L.Stream, M, G, A, URaw, Tr, NTracers, TrTend);
   ++*L.Count;
}

} // namespace omega
