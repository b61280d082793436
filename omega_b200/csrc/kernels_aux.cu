// kernels_aux.cu -- AuxiliaryState kernels that materialise every public aux
// array (the path behind AuxiliaryState::computeMomAux / computeAll,
// components/omega/src/ocn/AuxiliaryState.cpp:60-185).  One kernel per
// reference launch group; the per-element operation order of each functor is
// kept (J-loop order, association) so results agree with the reference to
// rounding of FMA contraction only.
#include "kernel_util.cuh"
#include "omega_dev.h"

namespace omega {

// vertexAuxState1: VorticityAuxVars::computeVarsOnVertex
// (src/ocn/auxiliaryVars/VorticityAuxVars.h:24-59)
template <int V>
__global__ void __launch_bounds__(kBlock)
k_vertex_aux1(DMesh M, const Real *__restrict__ H, const Real *__restrict__ U,
              Real *__restrict__ RelVort, Real *__restrict__ NormRelVort,
              Real *__restrict__ NormPlanetVort) {
   int v, k;
   if (!elem_k<V>(M.NVerticesAll, M.K, v, k)) return;
   const Real InvAt = 1.0 / M.AreaTriangle[v];
   Real hv[V] = {0}, rv[V] = {0};
#pragma unroll
   for (int j = 0; j < 3; ++j) {
      const int c       = M.CellsOnVertex[v * 3 + j];
      const int e       = M.EdgesOnVertex[v * 3 + j];
      const Real kite = M.KiteAreasOnVertex[v * 3 + j];
      const Real dcs  = InvAt * M.DcEdge[e] * M.EdgeSignOnVertex[v * 3 + j];
      Real h[V], u[V];
      ldv<V>(h, H + (size_t)c * M.K + k);
      ldv<V>(u, U + (size_t)e * M.K + k);
#pragma unroll
      for (int q = 0; q < V; ++q) {
         hv[q] += InvAt * kite * h[q];
         rv[q] += dcs * u[q];
      }
   }
   Real nr[V], np[V];
   const Real f = M.FVertex[v];
#pragma unroll
   for (int q = 0; q < V; ++q) {
      const Real inv = 1.0 / hv[q];
      nr[q]            = rv[q] * inv;
      np[q]            = f * inv;
   }
   const size_t o = (size_t)v * M.K + k;
   stv<V>(RelVort + o, rv);
   stv<V>(NormRelVort + o, nr);
   stv<V>(NormPlanetVort + o, np);
}

// cellAuxState1 + cellAuxState3: KineticAuxVars::computeVarsOnCell
// (KineticAuxVars.h:20-47) and LayerThicknessAuxVars::computeVarsOnCells
// (LayerThicknessAuxVars.h:63-82)
template <int V>
__global__ void __launch_bounds__(kBlock)
k_cell_aux1(DMesh M, const Real *__restrict__ H, const Real *__restrict__ U,
            Real *__restrict__ KE, Real *__restrict__ Div, Real *__restrict__ Ssh) {
   int c, k;
   if (!elem_k<V>(M.NCellsAll, M.K, c, k)) return;
   const Real InvA = 1.0 / M.AreaCell[c];
   const int n       = M.NEdgesOnCell[c];
   Real ke[V] = {0}, dv[V] = {0};
   for (int j = 0; j < n; ++j) {
      const int e        = M.EdgesOnCell[c * M.MaxEdges + j];
      const Real dve   = M.DvEdge[e];
      const Real area  = 0.5 * dve * M.DcEdge[e];
      const Real kcoef = area * 0.5 * InvA;
      const Real dcoef = dve * InvA * M.EdgeSignOnCell[c * M.MaxEdges + j];
      Real u[V];
      ldv<V>(u, U + (size_t)e * M.K + k);
#pragma unroll
      for (int q = 0; q < V; ++q) {
         ke[q] += kcoef * u[q] * u[q];
         dv[q] -= dcoef * u[q];
      }
   }
   const size_t o = (size_t)c * M.K + k;
   stv<V>(KE + o, ke);
   stv<V>(Div + o, dv);
   Real h[V], s[V];
   ldv<V>(h, H + o);
   const Real b = M.BottomDepth[c];
#pragma unroll
   for (int q = 0; q < V; ++q) s[q] = h[q] - b;
   stv<V>(Ssh + o, s);
}

// edgeAuxState1: WindForcingAuxVars::computeVarsOnEdge (WindForcingAuxVars.h:22-29)
// with InterpCellToEdge (src/ocn/HorzOperators.h:137-187)
__device__ __forceinline__ Real interp_c2e(const DMesh &M, int e, const Real *Cell, int Aniso) {
   if (Aniso) {
      const int2 c = M.CellsOnEdge[e];
      return 0.5 * (Cell[c.x] + Cell[c.y]);
   }
   Real acc = 0, area = 0;
   const int2 vv = M.VerticesOnEdge[e];
#pragma unroll
   for (int j = 0; j < 2; ++j) {
      const int v = j == 0 ? vv.x : vv.y;
#pragma unroll
      for (int l = 0; l < 3; ++l) {
         const Real kite = M.KiteAreasOnVertex[v * 3 + l];
         acc += Cell[M.CellsOnVertex[v * 3 + l]] * kite;
         area += kite;
      }
   }
   return acc * (1.0 / area);
}

__global__ void __launch_bounds__(kBlock)
k_wind_aux(DMesh M, const Real *__restrict__ Zonal, const Real *__restrict__ Merid,
           Real *__restrict__ NormalStress, int Aniso) {
   const int e = blockIdx.x * blockDim.x + threadIdx.x;
   if (e >= M.NEdgesAll) return;
   const Real zs = interp_c2e(M, e, Zonal, Aniso);
   const Real ms = interp_c2e(M, e, Merid, Aniso);
   const Real a  = M.AngleEdge[e];
   NormalStress[e] = cos(a) * zs + sin(a) * ms;
}

// edgeAuxState2: VorticityAuxVars::computeVarsOnEdge (VorticityAuxVars.h:61-76),
// LayerThicknessAuxVars::computeVarsOnEdge (LayerThicknessAuxVars.h:25-61),
// VelocityDel2AuxVars::computeVarsOnEdge (VelocityDel2AuxVars.h:21-45).
// ThickOnly = the computeLayerThickAux launch of Tendencies.cpp:507-512.
template <int V>
__global__ void __launch_bounds__(kBlock)
k_edge_aux2(DMesh M, const Real *__restrict__ H, const Real *__restrict__ U, AuxPtrs A,
            int FluxUpwind, int ThickOnly) {
   int e, k;
   if (!elem_k<V>(M.NEdgesAll, M.K, e, k)) return;
   const int2 cc  = M.CellsOnEdge[e];
   const size_t o = (size_t)e * M.K + k;
   Real h0[V], h1[V], mean[V], flux[V];
   ldv<V>(h0, H + (size_t)cc.x * M.K + k);
   ldv<V>(h1, H + (size_t)cc.y * M.K + k);
#pragma unroll
   for (int q = 0; q < V; ++q) mean[q] = 0.5 * (h0[q] + h1[q]);
   if (!FluxUpwind) {
#pragma unroll
      for (int q = 0; q < V; ++q) flux[q] = mean[q];
   } else {
      Real u[V];
      ldv<V>(u, U + o);
#pragma unroll
      for (int q = 0; q < V; ++q)
         flux[q] = u[q] > 0 ? h0[q] : (u[q] < 0 ? h1[q] : fmax(h0[q], h1[q]));
   }
   stv<V>(A.Mean + o, mean);
   stv<V>(A.Flux + o, flux);
   if (ThickOnly) return;
   const int2 vv = M.VerticesOnEdge[e];
   Real a0[V], a1[V], r[V];
   ldv<V>(a0, A.NormRelVort + (size_t)vv.x * M.K + k);
   ldv<V>(a1, A.NormRelVort + (size_t)vv.y * M.K + k);
#pragma unroll
   for (int q = 0; q < V; ++q) r[q] = 0.5 * (a0[q] + a1[q]);
   stv<V>(A.NormRelVortEdge + o, r);
   ldv<V>(a0, A.NormPlanetVort + (size_t)vv.x * M.K + k);
   ldv<V>(a1, A.NormPlanetVort + (size_t)vv.y * M.K + k);
#pragma unroll
   for (int q = 0; q < V; ++q) r[q] = 0.5 * (a0[q] + a1[q]);
   stv<V>(A.NormPlanetVortEdge + o, r);
   const Real dc = M.DcEdge[e], dv = M.DvEdge[e];
   const Real InvDc = 1.0 / dc, InvDv = 1.0 / fmax(dv, (Real)(0.25 * dc));
   const Real mask = M.EdgeMask[e];
   Real d0[V], d1[V];
   ldv<V>(d0, A.Div + (size_t)cc.x * M.K + k);
   ldv<V>(d1, A.Div + (size_t)cc.y * M.K + k);
   ldv<V>(a0, A.RelVort + (size_t)vv.x * M.K + k);
   ldv<V>(a1, A.RelVort + (size_t)vv.y * M.K + k);
#pragma unroll
   for (int q = 0; q < V; ++q) {
      const Real GradDiv  = (d1[q] - d0[q]) * InvDc;
      const Real CurlVort = -(a1[q] - a0[q]) * InvDv;
      r[q]                  = mask * GradDiv + CurlVort;
   }
   stv<V>(A.Del2Edge + o, r);
}

// vertexAuxState2: VelocityDel2AuxVars::computeVarsOnVertex (VelocityDel2AuxVars.h:69-89)
template <int V>
__global__ void __launch_bounds__(kBlock)
k_vertex_aux2(DMesh M, const Real *__restrict__ Del2Edge, Real *__restrict__ Del2RelVort) {
   int v, k;
   if (!elem_k<V>(M.NVerticesAll, M.K, v, k)) return;
   const Real InvAt = 1.0 / M.AreaTriangle[v];
   Real t[V] = {0};
#pragma unroll
   for (int j = 0; j < 3; ++j) {
      const int e    = M.EdgesOnVertex[v * 3 + j];
      const Real c = InvAt * M.DcEdge[e] * M.EdgeSignOnVertex[v * 3 + j];
      Real d[V];
      ldv<V>(d, Del2Edge + (size_t)e * M.K + k);
#pragma unroll
      for (int q = 0; q < V; ++q) t[q] += c * d[q];
   }
   stv<V>(Del2RelVort + (size_t)v * M.K + k, t);
}

// cellAuxState2: VelocityDel2AuxVars::computeVarsOnCell (VelocityDel2AuxVars.h:47-67)
template <int V>
__global__ void __launch_bounds__(kBlock)
k_cell_aux2(DMesh M, const Real *__restrict__ Del2Edge, Real *__restrict__ Del2Div) {
   int c, k;
   if (!elem_k<V>(M.NCellsAll, M.K, c, k)) return;
   const Real InvA = 1.0 / M.AreaCell[c];
   const int n       = M.NEdgesOnCell[c];
   Real t[V] = {0};
   for (int j = 0; j < n; ++j) {
      const int e     = M.EdgesOnCell[c * M.MaxEdges + j];
      const Real cf = M.DvEdge[e] * InvA * M.EdgeSignOnCell[c * M.MaxEdges + j];
      Real d[V];
      ldv<V>(d, Del2Edge + (size_t)e * M.K + k);
#pragma unroll
      for (int q = 0; q < V; ++q) t[q] -= cf * d[q];
   }
   stv<V>(Del2Div + (size_t)c * M.K + k, t);
}

// edgeAuxState4: TracerAuxVars::computeVarsOnEdge (TracerAuxVars.h:25-59), all tracers
template <int V>
__global__ void __launch_bounds__(kBlock)
k_tracer_aux_edge(DMesh M, const Real *__restrict__ H, const Real *__restrict__ U,
                  const Real *__restrict__ Tr, int NTracers, Real *__restrict__ HTr, int Upwind) {
   int e, k;
   if (!elem_k<V>(M.NEdgesAll, M.K, e, k)) return;
   const int2 cc = M.CellsOnEdge[e];
   Real h0[V], h1[V], u[V];
   ldv<V>(h0, H + (size_t)cc.x * M.K + k);
   ldv<V>(h1, H + (size_t)cc.y * M.K + k);
   if (Upwind) ldv<V>(u, U + (size_t)e * M.K + k);
   const size_t CS = (size_t)M.NCellsSize * M.K, ES = (size_t)M.NEdgesSize * M.K;
   for (int l = 0; l < NTracers; ++l) {
      Real t0[V], t1[V], r[V];
      ldv<V>(t0, Tr + l * CS + (size_t)cc.x * M.K + k);
      ldv<V>(t1, Tr + l * CS + (size_t)cc.y * M.K + k);
#pragma unroll
      for (int q = 0; q < V; ++q) {
         const Real a = h0[q] * t0[q], b = h1[q] * t1[q];
         if (!Upwind) r[q] = 0.5 * (a + b);
         else r[q] = u[q] > 0 ? a : (u[q] < 0 ? b : fmax(a, b));
      }
      stv<V>(HTr + l * ES + (size_t)e * M.K + k, r);
   }
}

// cellAuxState4: TracerAuxVars::computeVarsOnCells (TracerAuxVars.h:61-91), all tracers
template <int V>
__global__ void __launch_bounds__(kBlock)
k_tracer_aux_cell(DMesh M, const Real *__restrict__ Mean, const Real *__restrict__ Tr,
                  int NTracers, Real *__restrict__ Del2Tr) {
   int c, k;
   if (!elem_k<V>(M.NCellsAll, M.K, c, k)) return;
   const Real InvA = 1.0 / M.AreaCell[c];
   const int n       = M.NEdgesOnCell[c];
   const size_t CS   = (size_t)M.NCellsSize * M.K;
   for (int l = 0; l < NTracers; ++l) {
      Real t[V] = {0};
      const Real *T = Tr + l * CS;
      for (int j = 0; j < n; ++j) {
         const int e     = M.EdgesOnCell[c * M.MaxEdges + j];
         const int2 cc   = M.CellsOnEdge[e];
         const Real cf = M.EdgeMask[e] * M.EdgeSignOnCell[c * M.MaxEdges + j] * (M.DvEdge[e] / M.DcEdge[e]);
         Real m[V], t0[V], t1[V];
         ldv<V>(m, Mean + (size_t)e * M.K + k);
         ldv<V>(t0, T + (size_t)cc.x * M.K + k);
         ldv<V>(t1, T + (size_t)cc.y * M.K + k);
#pragma unroll
         for (int q = 0; q < V; ++q) t[q] -= cf * m[q] * (t1[q] - t0[q]);
      }
#pragma unroll
      for (int q = 0; q < V; ++q) t[q] *= InvA;
      stv<V>(Del2Tr + l * CS + (size_t)c * M.K + k, t);
   }
}

// ---------------------------------------------------------------------------
void launch_vertex_aux1(const Launch &L, const DMesh &M, const Real *H, const Real *U, const AuxPtrs &A) {
   OMEGA_LAUNCH_V(k_vertex_aux1, M.NVerticesAll, M.K, L.Stream, M, H, U, A.RelVort, A.NormRelVort, A.NormPlanetVort);
   ++*L.Count;
}
void launch_cell_aux1(const Launch &L, const DMesh &M, const Real *H, const Real *U, const AuxPtrs &A) {
   OMEGA_LAUNCH_V(k_cell_aux1, M.NCellsAll, M.K, L.Stream, M, H, U, A.KE, A.Div, A.Ssh);
   ++*L.Count;
}
void launch_wind_aux(const Launch &L, const DMesh &M, const AuxPtrs &A, int Aniso) {
   if (M.NEdgesAll > 0)
      k_wind_aux<<<grid_for(M.NEdgesAll), kBlock, 0, L.Stream>>>(M, A.ZonalStress, A.MeridStress, A.NormalStress, Aniso);
   ++*L.Count;
}
void launch_edge_aux2(const Launch &L, const DMesh &M, const Real *H, const Real *U, const AuxPtrs &A,
                      int FluxUpwind, int ThickOnly) {
   OMEGA_LAUNCH_V(k_edge_aux2, M.NEdgesAll, M.K, L.Stream, M, H, U, A, FluxUpwind, ThickOnly);
   ++*L.Count;
}
void launch_vertex_aux2(const Launch &L, const DMesh &M, const AuxPtrs &A) {
   OMEGA_LAUNCH_V(k_vertex_aux2, M.NVerticesAll, M.K, L.Stream, M, A.Del2Edge, A.Del2RelVort);
   ++*L.Count;
}
void launch_cell_aux2(const Launch &L, const DMesh &M, const AuxPtrs &A) {
   OMEGA_LAUNCH_V(k_cell_aux2, M.NCellsAll, M.K, L.Stream, M, A.Del2Edge, A.Del2Div);
   ++*L.Count;
}
void launch_tracer_aux_edge(const Launch &L, const DMesh &M, const Real *H, const Real *U,
                            const Real *Tr, int NTracers, const AuxPtrs &A, int Upwind) {
   if (NTracers == 0) return;
   OMEGA_LAUNCH_V(k_tracer_aux_edge, M.NEdgesAll, M.K, L.Stream, M, H, U, Tr, NTracers, A.HTr, Upwind);
   ++*L.Count;
}
void launch_tracer_aux_cell(const Launch &L, const DMesh &M, const Real *Tr, int NTracers,
                            const AuxPtrs &A) {
   if (NTracers == 0) return;
   OMEGA_LAUNCH_V(k_tracer_aux_cell, M.NCellsAll, M.K, L.Stream, M, A.Mean, Tr, NTracers, A.Del2Tr);
   ++*L.Count;
}

} // namespace omega
