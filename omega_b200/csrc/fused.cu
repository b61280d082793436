// fused.cu -- the fused per-stage RK4 pipeline (TimeStepper::doStep hot path).
//
// One RK4 stage of the reference is ~27 kernel launches that stream
// (177 + 24 T) doubles per cell-layer through HBM.  Here a stage
// is at most five launches and no tendency, edge-aux or tracer-flux array is
// ever materialised:
//
//   fA  cells+vertices : h,u            -> KE, Div | RelVort, PV(vertex)
//   fB  cells+vertices : Div, RelVort   -> Del2Div | Del2RelVort   (Del4 only;
//                        Del2Edge is recomputed on the fly, never stored)
//   fT2 cells          : tracers, h     -> Del2Tracers             (tracer Del4 only)
//   fC  cells          : thickness + every tracer tendency of every tracer,
//                        with the RK updates of h and tracers as epilogue
//   fE  edges          : every velocity tendency term + the RK update of u
//
// The stage update  q_next (+)= b_s dt R,  q_provis' = q_cur + a_{s+1} dt R  is
// applied in the epilogue of the kernel that produced R, into the OTHER of two
// provisional buffers (neighbours still read this stage's provisional state).
//
// Work assignment: a CTA owns a contiguous block of rows in space-filling-curve
// order and its warps sweep it together, so neighbour rows fetched by one warp
// are L1/L2 hits for the next; fA/fB process the vertex range that belongs to
// the same patch in the same CTA so u / Div rows are fetched once.
//
// Reference semantics restated: src/timeStepping/RungeKutta4Stepper.cpp:68-137,
// src/timeStepping/TimeStepper.cpp:378-524, src/ocn/Tendencies.cpp:257-486,
// src/ocn/AuxiliaryState.cpp:60-185 and the functors cited at each block.
#include <nvtx3/nvToolsExt.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "fused_args.h"
#include "halo.h"
#include "kernel_util.cuh"
#include "omega_dev.h"
#include "stepper.h"

namespace omega {

// rows [r0, r0+n) of this CTA for an index space of N rows split over gridDim.x CTAs
__device__ __forceinline__ void cta_rows(int N, int &r0, int &n) {
   const long long per = ((long long)N + gridDim.x - 1) / gridDim.x;
   const long long a   = per * blockIdx.x;
   r0                  = (int)(a < N ? a : N);
   const long long b   = a + per;
   n                   = (int)((b < N ? b : N) - r0);
}

// sweep (row, level-vector) items of a row block: consecutive threads take
// consecutive level vectors of a row, then the next row
#define OMEGA_FOR_ROWS(ROW, KLEV, R0, NROWS, KV, V)                                        \
   for (int w_ = threadIdx.x, ROW = (R0) + w_ / (KV), KLEV = (w_ % (KV)) * (V);           \
        w_ < (NROWS) * (KV);                                                               \
        w_ += blockDim.x, ROW = (R0) + w_ / (KV), KLEV = (w_ % (KV)) * (V))

// ---------------------------------------------------------------------------
// fA: KineticAuxVars::computeVarsOnCell (KineticAuxVars.h:20-47) on the CTA's
// cells, VorticityAuxVars::computeVarsOnVertex (VorticityAuxVars.h:24-59) on
// the matching vertex range.  PV(vertex) = NormRelVort + NormPlanetVort is
// stored as ONE array (the edge kernel only ever needs their sum).
// ---------------------------------------------------------------------------
template <int V>
__global__ void __launch_bounds__(kBlock)
k_fA(DMesh M, const Real *__restrict__ H, const Real *__restrict__ U, Real *__restrict__ KE,
     Real *__restrict__ Div, Real *__restrict__ RelVort, Real *__restrict__ PVv, int XMode, Real aKE,
     Real aSSH, Real *__restrict__ KEBot) {
   const int KV = M.K / V;
   int r0, n;
   cta_rows(M.NCellsAll, r0, n);
   OMEGA_FOR_ROWS(c, k, r0, n, KV, V) {
      Real ke[V] = {0}, dv[V] = {0};
      // six table entries per cell (omega_dev.h: ACell*; edge row = boundary slot and zero
      // coefficients when the cell has fewer edges): all u-row gathers in flight together
#pragma unroll
      for (int j = 0; j < 6; ++j) {
         const int e      = M.ACellEdge[c * 6 + j];
         const Real2 cf = M.ACellCf[c * 6 + j];
         Real u[V];
         ldv<V>(u, U + (size_t)e * M.K + k);
#pragma unroll
         for (int q = 0; q < V; ++q) {
            ke[q] += cf.x * u[q] * u[q];
            dv[q] -= cf.y * u[q];
         }
      }
      if (KEBot && k + V == M.K) KEBot[c] = ke[V - 1]; // plain KE of the bottom layer for BottomDragOnEdge (k_edge_layer_terms)
      if (XMode) {
         // tiled edge pass: KE and SSH gradients share (X(c1)-X(c0))/dc, X = aKE*KE + aSSH*g*(h - b)
         // (KEGradOnEdge, SSHGradOnEdge, TendencyTerms.h:127-173; LayerThicknessAuxVars.h:63-82)
         Real h[V];
         ldv<V>(h, H + (size_t)c * M.K + k);
         const Real b = M.BottomDepth[c];
#pragma unroll
         for (int q = 0; q < V; ++q) ke[q] = aKE * ke[q] + aSSH * (9.80665 * (h[q] - b));
      }
      stv<V>(KE + (size_t)c * M.K + k, ke);
      stv<V>(Div + (size_t)c * M.K + k, dv);
   }
   cta_rows(M.NVerticesAll, r0, n);
   OMEGA_FOR_ROWS(v, k, r0, n, KV, V) {
      Real hv[V] = {0}, rv[V] = {0};
#pragma unroll
      for (int j = 0; j < 3; ++j) {
         const int2 ix    = M.AVertIdx[v * 3 + j];
         const Real2 cf = M.AVertCf[v * 3 + j];
         Real h[V], u[V];
         ldv<V>(h, H + (size_t)ix.x * M.K + k);
         ldv<V>(u, U + (size_t)ix.y * M.K + k);
#pragma unroll
         for (int q = 0; q < V; ++q) {
            hv[q] += cf.x * h[q];
            rv[q] += cf.y * u[q];
         }
      }
      const Real f = M.FVertex[v];
      Real pv[V];
#pragma unroll
      for (int q = 0; q < V; ++q) {
         const Real inv = 1.0 / hv[q];
         pv[q]            = rv[q] * inv + f * inv;
      }
      stv<V>(RelVort + (size_t)v * M.K + k, rv);
      stv<V>(PVv + (size_t)v * M.K + k, pv);
   }
}

// ---------------------------------------------------------------------------
// fB: VelocityDel2AuxVars::computeVarsOnCell / computeVarsOnVertex
// (VelocityDel2AuxVars.h:47-89) with Del2Edge (VelocityDel2AuxVars.h:21-45) never
// stored: its two differences are folded, per (cell, edge) and per (vertex,
// edge), into the flattened stencils of omega_dev.h (B2*), so one element costs
// two table loads and three gathered rows per edge instead of nine loads and four rows.
// ---------------------------------------------------------------------------
template <int V>
__global__ void __launch_bounds__(kBlock)
k_fB(DMesh M, const Real *__restrict__ Div, const Real *__restrict__ RelVort,
     Real *__restrict__ Del2Div, Real *__restrict__ Del2RelVort, int YMode, Real Nu2, Real Nu4Div,
     Real Nu4) {
   const int KV = M.K / V;
   int r0, n;
   cta_rows(M.NCellsAll, r0, n);
   OMEGA_FOR_ROWS(c, k, r0, n, KV, V) {
      Real ds[V], t[V] = {0};
      ldv<V>(ds, Div + (size_t)c * M.K + k);
#pragma unroll
      for (int j = 0; j < 6; ++j) {
         const int4 ix    = M.B2CellIdx[c * 6 + j];
         const Real2 cf = M.B2CellCf[c * 6 + j];
         Real dn[V], z0[V], z1[V];
         ldv<V>(dn, Div + (size_t)ix.x * M.K + k);
         ldv<V>(z0, RelVort + (size_t)ix.y * M.K + k);
         ldv<V>(z1, RelVort + (size_t)ix.z * M.K + k);
#pragma unroll
         for (int q = 0; q < V; ++q) t[q] += cf.x * (dn[q] - ds[q]) + cf.y * (z1[q] - z0[q]);
      }
      if (YMode) {
         // tiled edge pass: Del2 and Del4 velocity terms share one difference per element kind,
         // Y(c) = nu2*Div - nu4*DivFactor*Del2Div, Z(v) = nu2*RelVort - nu4*Del2RelVort
         // (VelocityDiffusionOnEdge, VelocityHyperDiffOnEdge, TendencyTerms.h:195-269)
#pragma unroll
         for (int q = 0; q < V; ++q) t[q] = Nu2 * ds[q] - Nu4Div * t[q];
      }
      stv<V>(Del2Div + (size_t)c * M.K + k, t);
   }
   cta_rows(M.NVerticesAll, r0, n);
   OMEGA_FOR_ROWS(v, k, r0, n, KV, V) {
      Real zs[V], t[V] = {0};
      ldv<V>(zs, RelVort + (size_t)v * M.K + k);
#pragma unroll
      for (int j = 0; j < 3; ++j) {
         const int4 ix    = M.B2VertIdx[v * 3 + j];
         const Real2 cf = M.B2VertCf[v * 3 + j];
         Real d0[V], d1[V], zo[V];
         ldv<V>(d0, Div + (size_t)ix.x * M.K + k);
         ldv<V>(d1, Div + (size_t)ix.y * M.K + k);
         ldv<V>(zo, RelVort + (size_t)ix.z * M.K + k);
#pragma unroll
         for (int q = 0; q < V; ++q) t[q] += cf.x * (d1[q] - d0[q]) + cf.y * (zo[q] - zs[q]);
      }
      if (YMode) {
#pragma unroll
         for (int q = 0; q < V; ++q) t[q] = Nu2 * zs[q] - Nu4 * t[q];
      }
      stv<V>(Del2RelVort + (size_t)v * M.K + k, t);
   }
}

// k_fB with every gathered row read once (tables B3*, mesh.cu): 13 rows per cell and 7 per vertex
// instead of 19 and 10.  X != nullptr: the cell result is W = Y - X, the one cell field the tiled
// edge pass then differences for KE + SSH gradient, Del2 and Del4 divergence parts together.
template <int V>
__global__ void __launch_bounds__(kBlock)
k_fB3(DMesh M, const Real *__restrict__ Div, const Real *__restrict__ RelVort, Real *__restrict__ Del2Div,
      Real *__restrict__ Del2RelVort, int YMode, Real Nu2, Real Nu4Div, Real Nu4, const Real *__restrict__ X) {
   const int KV = M.K / V;
   int r0, n;
   cta_rows(M.NCellsAll, r0, n);
   OMEGA_FOR_ROWS(c, k, r0, n, KV, V) {
      int ix[12];
      Real cf[12];
#pragma unroll
      for (int q = 0; q < 3; ++q) {
         const int4 t4 = M.B3CellIdx[c * 3 + q];
         ix[4 * q] = t4.x; ix[4 * q + 1] = t4.y; ix[4 * q + 2] = t4.z; ix[4 * q + 3] = t4.w;
      }
#pragma unroll
      for (int q = 0; q < 6; ++q) {
         const Real2 t2 = M.B3CellCf[c * 6 + q];
         cf[2 * q] = t2.x; cf[2 * q + 1] = t2.y;
      }
      Real ds[V], xs[V] = {0}, t[V] = {0}, g[12][V];
      ldv<V>(ds, Div + (size_t)c * M.K + k);
      if (X) ldv<V>(xs, X + (size_t)c * M.K + k);
#pragma unroll
      for (int j = 0; j < 6; ++j) ldv<V>(g[j], Div + (size_t)ix[j] * M.K + k);
#pragma unroll
      for (int j = 6; j < 12; ++j) ldv<V>(g[j], RelVort + (size_t)ix[j] * M.K + k);
#pragma unroll
      for (int j = 0; j < 6; ++j)
#pragma unroll
         for (int q = 0; q < V; ++q) t[q] += cf[j] * (g[j][q] - ds[q]);
#pragma unroll
      for (int j = 6; j < 12; ++j)
#pragma unroll
         for (int q = 0; q < V; ++q) t[q] += cf[j] * g[j][q];
      if (YMode) {
#pragma unroll
         for (int q = 0; q < V; ++q) t[q] = Nu2 * ds[q] - Nu4Div * t[q] - xs[q];
      }
      stv<V>(Del2Div + (size_t)c * M.K + k, t);
   }
   cta_rows(M.NVerticesAll, r0, n);
   OMEGA_FOR_ROWS(v, k, r0, n, KV, V) {
      int ix[8];
      Real cf[8];
#pragma unroll
      for (int q = 0; q < 2; ++q) {
         const int4 t4 = M.B3VertIdx[v * 2 + q];
         ix[4 * q] = t4.x; ix[4 * q + 1] = t4.y; ix[4 * q + 2] = t4.z; ix[4 * q + 3] = t4.w;
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
         const Real2 t2 = M.B3VertCf[v * 4 + q];
         cf[2 * q] = t2.x; cf[2 * q + 1] = t2.y;
      }
      Real zs[V], t[V] = {0}, g[7][V];
      ldv<V>(zs, RelVort + (size_t)v * M.K + k);
#pragma unroll
      for (int j = 0; j < 4; ++j) ldv<V>(g[j], Div + (size_t)ix[j] * M.K + k);
#pragma unroll
      for (int j = 4; j < 7; ++j) ldv<V>(g[j], RelVort + (size_t)ix[j] * M.K + k);
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
         for (int q = 0; q < V; ++q) t[q] += cf[j] * g[j][q];
#pragma unroll
      for (int j = 4; j < 7; ++j)
#pragma unroll
         for (int q = 0; q < V; ++q) t[q] += cf[j] * (g[j][q] - zs[q]);
      if (YMode) {
#pragma unroll
         for (int q = 0; q < V; ++q) t[q] = Nu2 * zs[q] - Nu4 * t[q];
      }
      stv<V>(Del2RelVort + (size_t)v * M.K + k, t);
   }
}

// ---------------------------------------------------------------------------
// fT2: TracerAuxVars::computeVarsOnCells (TracerAuxVars.h:61-91) for every
// tracer; MeanLayerThickEdge is recomputed from h.  s*(phi(c1)-phi(c0)) ==
// phi(self)-phi(nbr) exactly, so only the neighbour row is fetched.
// ---------------------------------------------------------------------------
template <int V>
__global__ void __launch_bounds__(kBlock)
k_fT2(DMesh M, const Real *__restrict__ H, const Real *__restrict__ Tr, int NTracers,
      Real *__restrict__ Del2Tr) {
   const int KV = M.K / V;
   const size_t CS = (size_t)M.NCellsSize * M.K;
   int r0, n;
   cta_rows(M.NCellsAll, r0, n);
   OMEGA_FOR_ROWS(c, k, r0, n, KV, V) {
      const Real InvA = 1.0 / M.AreaCell[c];
      const int ne      = M.NEdgesOnCell[c];
      const size_t o    = (size_t)c * M.K + k;
      Real hs[V];
      ldv<V>(hs, H + o);
      // per-edge coefficient  mask*sign*(dv/dc)*mean  and neighbour offset
      Real cf[6][V];
      size_t on[6];
#pragma unroll
      for (int j = 0; j < 6; ++j) {
         if (j < ne) {
            const int e   = M.EdgesOnCell[c * M.MaxEdges + j];
            const int2 cc = M.CellsOnEdge[e];
            const int nb  = cc.x == c ? cc.y : cc.x;
            on[j]         = (size_t)nb * M.K + k;
            const Real g = M.EdgeMask[e] * M.EdgeSignOnCell[c * M.MaxEdges + j] * (M.DvEdge[e] / M.DcEdge[e]);
            Real h0[V], h1[V];
            ldv<V>(h0, H + (size_t)cc.x * M.K + k);
            ldv<V>(h1, H + (size_t)cc.y * M.K + k);
            const Real sgn = cc.x == c ? -1.0 : 1.0; // phi(c1)-phi(c0) = sgn*(phi(self)-phi(nbr))
#pragma unroll
            for (int q = 0; q < V; ++q) cf[j][q] = g * (0.5 * (h0[q] + h1[q])) * sgn;
         }
      }
      for (int l = 0; l < NTracers; ++l) {
         const Real *T = Tr + l * CS;
         Real ts[V], t[V] = {0};
         ldv<V>(ts, T + o);
#pragma unroll
         for (int j = 0; j < 6; ++j) {
            if (j < ne) {
               Real tn[V];
               ldv<V>(tn, T + on[j]);
#pragma unroll
               for (int q = 0; q < V; ++q) t[q] -= cf[j][q] * (ts[q] - tn[q]);
            }
         }
#pragma unroll
         for (int q = 0; q < V; ++q) t[q] *= InvA;
         stv<V>(Del2Tr + l * CS + o, t);
      }
   }
}


// ---------------------------------------------------------------------------
// fC: cell pass.  ThicknessFluxDivOnCell (TendencyTerms.h:35-58) with
// LayerThicknessAuxVars::computeVarsOnEdge (LayerThicknessAuxVars.h:25-61)
// recomputed per edge; TracerHorzAdvOnCell / TracerDiffOnCell /
// TracerHyperDiffOnCell (TendencyTerms.h:349-480) with
// TracerAuxVars::computeVarsOnEdge (TracerAuxVars.h:25-59) recomputed; then
// updateThicknessByTend, weightTracers, accumulateTracersUpdate,
// updateTracersByTend, finalizeTracersUpdate (TimeStepper.cpp:378-524) as the
// epilogue.  Per-edge coefficients live in registers and serve every tracer.
// ---------------------------------------------------------------------------

template <int V, bool UPW>
__global__ void __launch_bounds__(kBlock)
k_fC(DMesh M, CellArgs P) {
   const int KV = M.K / V;
   const size_t CS = (size_t)M.NCellsSize * M.K;
   int r0, n;
   cta_rows(M.NCellsAll, r0, n);
   OMEGA_FOR_ROWS(c, k, r0, n, KV, V) {
      const Real InvA = 1.0 / M.AreaCell[c];
      const int ne      = M.NEdgesOnCell[c];
      const size_t o    = (size_t)c * M.K + k;
      Real hs[V];
      ldv<V>(hs, P.Hs + o);
      Real hn[kME][V], w[kME][V]; // neighbour thickness; mask*dv*sign*u_tr/A
      Real cd[kME], ch[kME];      // -k2/A*mask*sign*scale2*dv/dc*sgn ; +k4/A*mask*sign*scale4*dv/dc*sgn
      int nb[kME];
      bool self0[kME];
      Real su[UPW ? kME : 1][V]; // sign of the stage velocity (upwind options only)
      Real hT[V] = {0};
#pragma unroll
      for (int j = 0; j < kME; ++j) {
         if (j < ne) {
            const int e       = M.EdgesOnCell[c * M.MaxEdges + j];
            const int2 cc     = M.CellsOnEdge[e];
            self0[j]          = cc.x == c;
            nb[j]             = self0[j] ? cc.y : cc.x;
            const Real sg   = M.EdgeSignOnCell[c * M.MaxEdges + j];
            const Real mask = M.EdgeMask[e], dve = M.DvEdge[e];
            const Real rdc  = dve / M.DcEdge[e];
            const Real sgn  = self0[j] ? -1.0 : 1.0;
            cd[j] = -(P.EddyDiff2 * InvA) * (mask * sg * (M.MeshScalingDel2[e] * rdc)) * sgn;
            ch[j] = (P.EddyDiff4 * InvA) * (mask * sg * (M.MeshScalingDel4[e] * rdc)) * sgn;
            Real us[V], ut[V];
            ldv<V>(hn[j], P.Hs + (size_t)nb[j] * M.K + k);
            ldv<V>(us, P.Us + (size_t)e * M.K + k);
            if (P.UTr == P.Us) {
#pragma unroll
               for (int q = 0; q < V; ++q) ut[q] = us[q];
            } else {
               ldv<V>(ut, P.UTr + (size_t)e * M.K + k);
            }
            const Real cf = dve * sg;
#pragma unroll
            for (int q = 0; q < V; ++q) {
               // FluxLayerThickEdge: centred, or upwind on the sign of u (h0 is cell 0 of the edge)
               Real flux;
               if (UPW) su[j][q] = us[q] > 0 ? 1.0 : (us[q] < 0 ? -1.0 : 0.0);
               if (!UPW || !P.FluxUpwind) flux = 0.5 * (hs[q] + hn[j][q]);
               else {
                  const Real h0 = self0[j] ? hs[q] : hn[j][q], h1 = self0[j] ? hn[j][q] : hs[q];
                  flux = us[q] > 0 ? h0 : (us[q] < 0 ? h1 : fmax(h0, h1));
               }
               hT[q] += cf * flux * us[q] * InvA; // Tend -= (-sum) (TendencyTerms.h:48-56)
               w[j][q] = mask * dve * sg * ut[q] * InvA;
            }
         }
      }
      if (!P.FluxDivOn) {
#pragma unroll
         for (int q = 0; q < V; ++q) hT[q] = 0.0;
      }
      if (P.MS.Enabled) {
         const Real Phase = P.MS.Kx * M.XCell[c] + P.MS.Ky * M.YCell[c] - P.MS.AngFreq * P.MS.Time;
         const Real src   = P.MS.Eta0 * (-P.MS.H0 * (P.MS.Kx + P.MS.Ky) * sin(Phase) - P.MS.AngFreq * cos(Phase) +
                                         P.MS.Eta0 * (P.MS.Kx + P.MS.Ky) * cos(2.0 * Phase));
#pragma unroll
         for (int q = 0; q < V; ++q) hT[q] += src;
      }
      // ---- thickness updates (updateThicknessByTend) ----
      Real hcur[V], hnext_old[V], hnext[V], hprov[V];
      if (P.First && P.Hcur == P.Hs) {
#pragma unroll
         for (int q = 0; q < V; ++q) hcur[q] = hs[q];
      } else {
         ldp<V>(hcur, P.Hcur + o);
      }
      if (!P.First || P.WSel) ldp<V>(hnext_old, P.Hnext + o);
#pragma unroll
      for (int q = 0; q < V; ++q) {
         hnext[q] = (P.First ? hcur[q] : hnext_old[q]) + P.B * hT[q];
         hprov[q] = hcur[q] + P.A * hT[q];
      }
      stv<V>(P.Hnext + o, hnext);
      if (!P.Last) stv<V>(P.Hprov + o, hprov);
      // ---- tracers ----
      Real hw[V], h2[V], hfin[V];
#pragma unroll
      for (int q = 0; q < V; ++q) {
         hw[q]   = P.WSel ? hnext_old[q] : hcur[q];
         h2[q]   = P.H2Sel == 2 ? hprov[q] : (P.H2Sel ? hnext[q] : hcur[q]);
         hfin[q] = P.FinSel ? hcur[q] : hnext[q];
      }
      for (int l = 0; l < P.NTracers; ++l) {
         const Real *T  = P.Trs + l * CS;
         const Real *D2 = P.Del2Tr + l * CS;
         Real ts[V], d2s[V], tend[V] = {0};
         ldv<V>(ts, T + o);
         if (P.HypOn) ldv<V>(d2s, D2 + o);
#pragma unroll
         for (int j = 0; j < kME; ++j) {
            if (j < ne) {
               const size_t on = (size_t)nb[j] * M.K + k;
               Real tn[V];
               ldv<V>(tn, T + on);
#pragma unroll
               for (int q = 0; q < V; ++q) {
                  if (P.AdvOn) {
                     const Real a = hs[q] * ts[q], b = hn[j][q] * tn[q];
                     Real htr;
                     if (!UPW || !P.TrUpwind) htr = 0.5 * (a + b);
                     else {
                        // TracerAuxVars upwinds on the velocity of the aux call (the stage velocity)
                        const Real h0t = self0[j] ? a : b, h1t = self0[j] ? b : a;
                        const Real sgnu = su[UPW ? j : 0][q];
                        htr = sgnu > 0 ? h0t : (sgnu < 0 ? h1t : fmax(h0t, h1t));
                     }
                     tend[q] += htr * w[j][q];
                  }
                  if (P.DiffOn) tend[q] += cd[j] * (0.5 * (hs[q] + hn[j][q])) * (ts[q] - tn[q]);
               }
               if (P.HypOn) {
                  Real d2n[V];
                  ldv<V>(d2n, D2 + on);
#pragma unroll
                  for (int q = 0; q < V; ++q) tend[q] += ch[j] * (d2s[q] - d2n[q]);
               }
            }
         }
         // epilogue: weight/accumulate/provisional/finalise (TimeStepper.cpp:446-524)
         Real tcur[V], acc[V], out[V];
         if (P.First && P.TrCur == P.Trs) {
#pragma unroll
            for (int q = 0; q < V; ++q) tcur[q] = ts[q];
         } else if (P.First || !P.Last) {
            ldp<V>(tcur, P.TrCur + l * CS + o);
         }
         if (!P.First) ldp<V>(acc, P.TrNext + l * CS + o);
#pragma unroll
         for (int q = 0; q < V; ++q) {
            const Real base = P.First ? tcur[q] * hw[q] : acc[q];
            out[q]            = base + P.B * tend[q];
            if (P.Last) out[q] = out[q] / hfin[q];
         }
         stv<V>(P.TrNext + l * CS + o, out);
         if (!P.Last) {
#pragma unroll
            for (int q = 0; q < V; ++q) out[q] = (tcur[q] * h2[q] + P.A * tend[q]) / (P.DivSel ? hcur[q] : hprov[q]);
            stv<V>(P.TrProv + l * CS + o, out);
         }
      }
   }
}


// ---------------------------------------------------------------------------
// fC, centred-flux variant (the Default.yml configuration and every BASELINE
// workload): same arithmetic as k_fC above with the per-edge work of one cell
// folded into ONE coefficient per (edge, level):
//    tend_l = phi_l(self)*S0 + sum_j E_j*phi_l(nbr_j)
//           + d2_l(self)*SH  - sum_j ch_j*d2_l(nbr_j)
//    E_j  = 0.5*h(nbr_j)*w_j - D_j,   S0 = sum_j (0.5*h(self)*w_j + D_j)
//    w_j  = mask*dv*sign*u/A,         D_j = -(k2/A)*mask*sign*(dv/dc)*mean_h_j*sgn_j
// so a thread keeps 6*(V+1) doubles for all tracers instead of 6*(2V+2); that
// is what lets 16 cells (512 threads) share one SM's L1 in lockstep.  The
// reassociation moves results by a few ulp of the largest term (parity tests
// hold the 1e-12 bound on every case).
// ---------------------------------------------------------------------------
constexpr int kCellThreads = 512;

template <int V, int NT, int MINB>
__global__ void __launch_bounds__(NT, MINB)
k_fC_center(DMesh M, CellArgs P) {
   const int KV = M.K / V;
   const size_t CS = (size_t)M.NCellsSize * M.K;
   int r0, n;
   cta_rows(M.NCellsAll, r0, n);
   OMEGA_FOR_ROWS(c, k, r0, n, KV, V) {
      const Real InvA = 1.0 / M.AreaCell[c];
      const int ne      = M.NEdgesOnCell[c];
      const size_t o    = (size_t)c * M.K + k;
      Real hs[V];
      ldv<V>(hs, P.Hs + o);
      Real E[kME][V], ch[kME], S0[V] = {0}, SH = 0.0, hT[V] = {0};
      int nb[kME];
      const Real k2 = P.DiffOn ? P.EddyDiff2 * InvA : 0.0, k4 = P.HypOn ? P.EddyDiff4 * InvA : 0.0;
#pragma unroll
      for (int j = 0; j < kME; ++j) {
         if (j < ne) {
            const int e       = M.EdgesOnCell[c * M.MaxEdges + j];
            const int2 cc     = M.CellsOnEdge[e];
            const bool self0  = cc.x == c;
            nb[j]             = self0 ? cc.y : cc.x;
            const Real sg   = M.EdgeSignOnCell[c * M.MaxEdges + j];
            const Real mask = M.EdgeMask[e], dve = M.DvEdge[e];
            const Real rdc  = dve / M.DcEdge[e];
            const Real sgn  = self0 ? -1.0 : 1.0;
            const Real cd   = -k2 * (mask * sg * (M.MeshScalingDel2[e] * rdc)) * sgn;
            ch[j]             = k4 * (mask * sg * (M.MeshScalingDel4[e] * rdc)) * sgn;
            SH += ch[j];
            Real hn[V], us[V], ut[V];
            ldv<V>(hn, P.Hs + (size_t)nb[j] * M.K + k);
            ldv<V>(us, P.Us + (size_t)e * M.K + k);
            if (P.UTr == P.Us) {
#pragma unroll
               for (int q = 0; q < V; ++q) ut[q] = us[q];
            } else {
               ldv<V>(ut, P.UTr + (size_t)e * M.K + k);
            }
            const Real cf = dve * sg;
            const Real wa = P.AdvOn ? mask * dve * sg : 0.0;
#pragma unroll
            for (int q = 0; q < V; ++q) {
               const Real mean = 0.5 * (hs[q] + hn[q]);
               hT[q] += cf * mean * us[q] * InvA;
               const Real w = wa * ut[q] * InvA;
               const Real D = cd * mean;
               E[j][q]        = 0.5 * hn[q] * w - D;
               S0[q] += 0.5 * hs[q] * w + D;
            }
         }
      }
      if (!P.FluxDivOn) {
#pragma unroll
         for (int q = 0; q < V; ++q) hT[q] = 0.0;
      }
      if (P.MS.Enabled) {
         const Real Phase = P.MS.Kx * M.XCell[c] + P.MS.Ky * M.YCell[c] - P.MS.AngFreq * P.MS.Time;
         const Real src   = P.MS.Eta0 * (-P.MS.H0 * (P.MS.Kx + P.MS.Ky) * sin(Phase) - P.MS.AngFreq * cos(Phase) +
                                         P.MS.Eta0 * (P.MS.Kx + P.MS.Ky) * cos(2.0 * Phase));
#pragma unroll
         for (int q = 0; q < V; ++q) hT[q] += src;
      }
      Real hcur[V], hnext_old[V], hnext[V], hprov[V];
      if (P.First && P.Hcur == P.Hs) {
#pragma unroll
         for (int q = 0; q < V; ++q) hcur[q] = hs[q];
      } else {
         ldp<V>(hcur, P.Hcur + o);
      }
      if (!P.First || P.WSel) ldp<V>(hnext_old, P.Hnext + o);
#pragma unroll
      for (int q = 0; q < V; ++q) {
         hnext[q] = (P.First ? hcur[q] : hnext_old[q]) + P.B * hT[q];
         hprov[q] = hcur[q] + P.A * hT[q];
      }
      stv<V>(P.Hnext + o, hnext);
      if (!P.Last) stv<V>(P.Hprov + o, hprov);
      // hx: stage 0 -> weight thickness; last stage -> final thickness; else unused
      Real hx[V], h2[V];
#pragma unroll
      for (int q = 0; q < V; ++q) {
         hx[q] = P.First ? (P.WSel ? hnext_old[q] : hcur[q]) : (P.FinSel ? hcur[q] : hnext[q]);
         h2[q] = P.H2Sel == 2 ? hprov[q] : (P.H2Sel ? hnext[q] : hcur[q]);
      }
      for (int l = 0; l < P.NTracers; ++l) {
         const Real *T  = P.Trs + l * CS;
         const Real *D2 = P.Del2Tr + l * CS;
         if (P.Prefetch && l + 1 < P.NTracers) {
            const Real *Tn = T + CS, *Dn = D2 + CS;
            prefetch_row(Tn + o, P.Prefetch);
            if (P.HypOn) prefetch_row(Dn + o, P.Prefetch);
#pragma unroll
            for (int j = 0; j < kME; ++j) {
               if (j < ne) {
                  prefetch_row(Tn + (size_t)nb[j] * M.K + k, P.Prefetch);
                  if (P.HypOn) prefetch_row(Dn + (size_t)nb[j] * M.K + k, P.Prefetch);
               }
            }
            if (!(P.First && P.TrCur == P.Trs) && (P.First || !P.Last)) prefetch_row(P.TrCur + (l + 1) * CS + o, P.Prefetch);
            if (!P.First) prefetch_row(P.TrNext + (l + 1) * CS + o, P.Prefetch);
         }
         Real ts[V], d2s[V], tend[V];
         ldv<V>(ts, T + o);
         if (P.HypOn) ldv<V>(d2s, D2 + o);
#pragma unroll
         for (int q = 0; q < V; ++q) tend[q] = ts[q] * S0[q] + (P.HypOn ? d2s[q] * SH : 0.0);
#pragma unroll
         for (int j = 0; j < kME; ++j) {
            if (j < ne) {
               const size_t on = (size_t)nb[j] * M.K + k;
               Real tn[V];
               ldv<V>(tn, T + on);
#pragma unroll
               for (int q = 0; q < V; ++q) tend[q] += E[j][q] * tn[q];
               if (P.HypOn) {
                  Real d2n[V];
                  ldv<V>(d2n, D2 + on);
#pragma unroll
                  for (int q = 0; q < V; ++q) tend[q] -= ch[j] * d2n[q];
               }
            }
         }
         Real tcur[V], acc[V], out[V];
         if (P.First && P.TrCur == P.Trs) {
#pragma unroll
            for (int q = 0; q < V; ++q) tcur[q] = ts[q];
         } else if (P.First || !P.Last) {
            ldp<V>(tcur, P.TrCur + l * CS + o);
         }
         if (!P.First) ldp<V>(acc, P.TrNext + l * CS + o);
#pragma unroll
         for (int q = 0; q < V; ++q) {
            const Real base = P.First ? tcur[q] * hx[q] : acc[q];
            out[q]            = base + P.B * tend[q];
            if (P.Last) out[q] = out[q] / hx[q];
         }
         stv<V>(P.TrNext + l * CS + o, out);
         if (!P.Last) {
#pragma unroll
            for (int q = 0; q < V; ++q) out[q] = (tcur[q] * h2[q] + P.A * tend[q]) / (P.DivSel ? hcur[q] : hprov[q]);
            stv<V>(P.TrProv + l * CS + o, out);
         }
      }
   }
}

// ---------------------------------------------------------------------------
// fE: edge pass.  PotentialVortHAdvOnEdge, KEGradOnEdge, SSHGradOnEdge,
// VelocityDiffusionOnEdge, VelocityHyperDiffOnEdge, WindForcingOnEdge,
// BottomDragOnEdge (TendencyTerms.h:81-334) + the manufactured source, with
// the edge aux vars they consume recomputed from cell/vertex fields
// (VorticityAuxVars.h:61-76, LayerThicknessAuxVars.h:25-82), then
// updateVelocityByTend (TimeStepper.cpp:407-430) as the epilogue.
// ---------------------------------------------------------------------------

template <int V> __device__ __forceinline__ void flux_edge(const DMesh &M, int e, int k, const Real *H,
                                                           const Real (&u)[V], int Upwind, Real (&flux)[V]) {
   const int2 cc = M.CellsOnEdge[e];
   Real h0[V], h1[V];
   ldv<V>(h0, H + (size_t)cc.x * M.K + k);
   ldv<V>(h1, H + (size_t)cc.y * M.K + k);
#pragma unroll
   for (int q = 0; q < V; ++q) {
      if (!Upwind) flux[q] = 0.5 * (h0[q] + h1[q]);
      else flux[q] = u[q] > 0 ? h0[q] : (u[q] < 0 ? h1[q] : fmax(h0[q], h1[q]));
   }
}

template <int V> __device__ __forceinline__ void pv_edge(const DMesh &M, int e, int k, const Real *PVv,
                                                         Real (&pv)[V]) {
   const int2 vv = M.VerticesOnEdge[e];
   Real a[V], b[V];
   ldv<V>(a, PVv + (size_t)vv.x * M.K + k);
   ldv<V>(b, PVv + (size_t)vv.y * M.K + k);
#pragma unroll
   for (int q = 0; q < V; ++q) pv[q] = 0.5 * (a[q] + b[q]);
}

template <int V, int MINB>
__global__ void __launch_bounds__(kBlock, MINB)
k_fE(DMesh M, EdgeArgs P) {
   const int KV = M.K / V;
   const OmegaTendConfig &G = P.G;
   int r0, n;
   cta_rows(M.NEdgesAll, r0, n);
   OMEGA_FOR_ROWS(e, k, r0, n, KV, V) {
      const size_t o    = (size_t)e * M.K + k;
      const Real mask = M.EdgeMask[e];
      const int2 cc = M.CellsOnEdge[e], vv = M.VerticesOnEdge[e];
      const Real InvDc = 1.0 / M.DcEdge[e], InvDv = 1.0 / M.DvEdge[e];
      Real tend[V] = {0};
      Real us[V], h0[V], h1[V];
      ldv<V>(us, P.Us + o);
      ldv<V>(h0, P.Hs + (size_t)cc.x * M.K + k);
      ldv<V>(h1, P.Hs + (size_t)cc.y * M.K + k);
      if (G.PVTendencyEnable) {
         Real qe[V], acc[V] = {0};
         pv_edge<V>(M, e, k, P.PVv, qe);
         const int ne = M.NEdgesOnEdge[e];
         for (int j = 0; j < ne; ++j) {
            const int je   = M.EdgesOnEdge[e * M.MaxEdges2 + j];
            const Real w = M.WeightsOnEdge[e * M.MaxEdges2 + j];
            Real uj[V], fl[V], qj[V];
            ldv<V>(uj, P.Us + (size_t)je * M.K + k);
            flux_edge<V>(M, je, k, P.Hs, uj, G.FluxThicknessType, fl);
            pv_edge<V>(M, je, k, P.PVv, qj);
#pragma unroll
            for (int q = 0; q < V; ++q) acc[q] += w * fl[q] * uj[q] * ((qe[q] + qj[q]) * 0.5);
         }
#pragma unroll
         for (int q = 0; q < V; ++q) tend[q] += mask * acc[q];
      }
      if (G.KETendencyEnable || G.BottomDragTendencyEnable) {
         Real a0[V], a1[V];
         ldv<V>(a0, P.KE + (size_t)cc.x * M.K + k);
         ldv<V>(a1, P.KE + (size_t)cc.y * M.K + k);
         if (G.KETendencyEnable) {
#pragma unroll
            for (int q = 0; q < V; ++q) tend[q] -= mask * (a1[q] - a0[q]) * InvDc;
         }
         if (G.BottomDragTendencyEnable) { // applied after the diffusion terms in the reference; the
            // terms are independent so only the summation order differs (within tolerance)
            const int KBot = M.K - 1;
            if (KBot >= k && KBot < k + V) {
               const int q           = KBot - k;
               const Real VelNorm  = sqrt(a0[q] + a1[q]);
               const Real InvThick = 1.0 / (0.5 * (h0[q] + h1[q]));
               tend[q] -= mask * G.BottomDragCoeff * VelNorm * InvThick * __ldg(P.URaw + (size_t)e * M.K + KBot);
            }
         }
      }
      if (G.SSHTendencyEnable) {
         const Real Grav = 9.80665, b0 = M.BottomDepth[cc.x], b1 = M.BottomDepth[cc.y];
#pragma unroll
         for (int q = 0; q < V; ++q) tend[q] -= mask * Grav * ((h1[q] - b1) - (h0[q] - b0)) * InvDc;
      }
      if (G.VelDiffTendencyEnable) {
         Real d0[V], d1[V], q0[V], q1[V];
         ldv<V>(d0, P.Div + (size_t)cc.x * M.K + k);
         ldv<V>(d1, P.Div + (size_t)cc.y * M.K + k);
         ldv<V>(q0, P.RelVort + (size_t)vv.x * M.K + k);
         ldv<V>(q1, P.RelVort + (size_t)vv.y * M.K + k);
         const Real cf = mask * G.ViscDel2 * M.MeshScalingDel2[e];
#pragma unroll
         for (int q = 0; q < V; ++q) tend[q] += cf * ((d1[q] - d0[q]) * InvDc - (q1[q] - q0[q]) * InvDv);
      }
      if (G.VelHyperDiffTendencyEnable) {
         Real d0[V], d1[V], q0[V], q1[V];
         ldv<V>(d0, P.Del2Div + (size_t)cc.x * M.K + k);
         ldv<V>(d1, P.Del2Div + (size_t)cc.y * M.K + k);
         ldv<V>(q0, P.Del2RelVort + (size_t)vv.x * M.K + k);
         ldv<V>(q1, P.Del2RelVort + (size_t)vv.y * M.K + k);
         const Real cf = mask * G.ViscDel4 * M.MeshScalingDel4[e];
#pragma unroll
         for (int q = 0; q < V; ++q)
            tend[q] -= cf * (G.DivFactor * (d1[q] - d0[q]) * InvDc - (q1[q] - q0[q]) * InvDv);
      }
      if (G.WindForcingTendencyEnable && k == 0) {
         const Real InvThick = 1.0 / (0.5 * (h0[0] + h1[0]));
         tend[0] += mask * InvThick * P.NormalStress[e] / G.Density0;
      }
      if (P.MS.Enabled) {
         const Real Grav  = 9.80665;
         const Real Phase = P.MS.Kx * M.XEdge[e] + P.MS.Ky * M.YEdge[e] - P.MS.AngFreq * P.MS.Time;
         const Real Src0  = P.MS.AngFreq * sin(Phase) - 0.5 * P.MS.Eta0 * (P.MS.Kx + P.MS.Ky) * sin(2.0 * Phase);
         const Real f     = M.FEdge[e];
         Real Uu = P.MS.Eta0 * ((-f + Grav * P.MS.Kx) * cos(Phase) + Src0);
         Real Vv = P.MS.Eta0 * ((f + Grav * P.MS.Ky) * cos(Phase) + Src0);
         const Real Kx2 = P.MS.Kx * P.MS.Kx, Ky2 = P.MS.Ky * P.MS.Ky;
         if (G.VelDiffTendencyEnable) {
            Uu += G.ViscDel2 * P.MS.Eta0 * (Kx2 + Ky2) * cos(Phase);
            Vv += G.ViscDel2 * P.MS.Eta0 * (Kx2 + Ky2) * cos(Phase);
         }
         if (G.VelHyperDiffTendencyEnable) {
            const Real Kx4 = Kx2 * Kx2, Ky4 = Ky2 * Ky2;
            Uu -= G.ViscDel4 * P.MS.Eta0 * ((Kx4 + Ky4 + Kx2 * Ky2) * cos(Phase));
            Vv -= G.ViscDel4 * P.MS.Eta0 * ((Kx4 + Ky4 + Kx2 * Ky2) * cos(Phase));
         }
         const Real a   = M.AngleEdge[e];
         const Real src = cos(a) * Uu + sin(a) * Vv;
#pragma unroll
         for (int q = 0; q < V; ++q) tend[q] += src;
      }
      // epilogue: updateVelocityByTend for the next-level and provisional states
      Real ucur[V], base[V], out[V];
      if (P.First && P.Ucur == P.Us) {
#pragma unroll
         for (int q = 0; q < V; ++q) ucur[q] = us[q];
      } else if (P.First || !P.Last) {
         ldp<V>(ucur, P.Ucur + o);
      }
      if (!P.First) ldp<V>(base, P.Unext + o);
#pragma unroll
      for (int q = 0; q < V; ++q) out[q] = (P.First ? ucur[q] : base[q]) + P.B * tend[q];
      stv<V>(P.Unext + o, out);
      if (!P.Last) {
#pragma unroll
         for (int q = 0; q < V; ++q) out[q] = ucur[q] + P.A * tend[q];
         stv<V>(P.Uprov + o, out);
      }
   }
}

// ---------------------------------------------------------------------------
// Terms the tiled kernels take as precomputed per-element values or as a single-layer correction.
//
// k_ms_sources: the manufactured-solution sources of src/ocn/CustomTendencyTerms.cpp:107-210 are the same on
// every level, so they are evaluated once per cell / edge and stage (not per level in the hot kernels):
// SrcH feeds the thickness tendency of the tiled cell pass, SrcU the velocity tendency of the tiled edge pass.
//
// k_edge_layer_terms: WindForcingOnEdge touches the top layer only and BottomDragOnEdge the bottom layer only
// (TendencyTerms.h:291-334).  The RK update is linear in the tendency, u_next = base + B*(R + r), so the tiled
// edge pass updates with R and this kernel adds B*r / A*r on the one layer concerned (the sum differs from the
// reference's term order by rounding only).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock)
k_ms_sources(DMesh M, ManufacturedParams MS, Real TimeEdge, int Del2On, Real Visc2, int Del4On, Real Visc4,
             Real *__restrict__ SrcH, Real *__restrict__ SrcU) {
   const int i = blockIdx.x * blockDim.x + threadIdx.x;
   if (i < M.NCellsAll) {
      const Real Phase = MS.Kx * M.XCell[i] + MS.Ky * M.YCell[i] - MS.AngFreq * MS.Time;
      SrcH[i] = MS.Eta0 * (-MS.H0 * (MS.Kx + MS.Ky) * sin(Phase) - MS.AngFreq * cos(Phase) +
                           MS.Eta0 * (MS.Kx + MS.Ky) * cos(2.0 * Phase));
   }
   if (i < M.NEdgesAll) {
      const int e = i;
      const Real Grav  = 9.80665;
      const Real Phase = MS.Kx * M.XEdge[e] + MS.Ky * M.YEdge[e] - MS.AngFreq * TimeEdge;
      const Real Src0  = MS.AngFreq * sin(Phase) - 0.5 * MS.Eta0 * (MS.Kx + MS.Ky) * sin(2.0 * Phase);
      const Real f     = M.FEdge[e];
      Real Uu = MS.Eta0 * ((-f + Grav * MS.Kx) * cos(Phase) + Src0);
      Real Vv = MS.Eta0 * ((f + Grav * MS.Ky) * cos(Phase) + Src0);
      const Real Kx2 = MS.Kx * MS.Kx, Ky2 = MS.Ky * MS.Ky;
      if (Del2On) {
         Uu += Visc2 * MS.Eta0 * (Kx2 + Ky2) * cos(Phase);
         Vv += Visc2 * MS.Eta0 * (Kx2 + Ky2) * cos(Phase);
      }
      if (Del4On) {
         const Real Kx4 = Kx2 * Kx2, Ky4 = Ky2 * Ky2;
         Uu -= Visc4 * MS.Eta0 * ((Kx4 + Ky4 + Kx2 * Ky2) * cos(Phase));
         Vv -= Visc4 * MS.Eta0 * ((Kx4 + Ky4 + Kx2 * Ky2) * cos(Phase));
      }
      const Real a = M.AngleEdge[e];
      SrcU[e] = cos(a) * Uu + sin(a) * Vv;
   }
}

__global__ void __launch_bounds__(kBlock)
k_edge_layer_terms(DMesh M, const Real *__restrict__ Hs, const Real *__restrict__ URaw,
                   const Real *__restrict__ KEBot, const Real *__restrict__ NormalStress, int WindOn, int DragOn,
                   Real InvRho0, Real Cd, Real *__restrict__ RTop, Real *__restrict__ RBot) {
   // evaluated BEFORE the edge pass: with raw time-level indexing URaw may be the next-level array the edge
   // pass is about to overwrite
   const int e = blockIdx.x * blockDim.x + threadIdx.x;
   if (e >= M.NEdgesAll) return;
   const int2 cc     = M.CellsOnEdge[e];
   const Real mask = M.EdgeMask[e];
   const int K = M.K, KBot = K - 1;
   Real r0 = 0.0, rb = 0.0; // tendency of the top / bottom layer
   if (WindOn) {
      const Real InvThick = 1.0 / (0.5 * (Hs[(size_t)cc.x * K] + Hs[(size_t)cc.y * K]));
      r0 = mask * InvThick * NormalStress[e] * InvRho0;
   }
   if (DragOn) {
      const Real VelNorm  = sqrt(KEBot[cc.x] + KEBot[cc.y]);
      const Real InvThick = 1.0 / (0.5 * (Hs[(size_t)cc.x * K + KBot] + Hs[(size_t)cc.y * K + KBot]));
      rb = -mask * Cd * VelNorm * InvThick * URaw[(size_t)e * K + KBot];
   }
   if (KBot == 0) {
      r0 += rb;
      rb = 0.0;
   }
   RTop[e] = r0;
   RBot[e] = rb;
}

__global__ void __launch_bounds__(kBlock)
k_edge_layer_apply(DMesh M, const Real *__restrict__ RTop, const Real *__restrict__ RBot, Real *__restrict__ Unext,
                   Real *__restrict__ Uprov, Real A, Real B, TileRange TR, const int *__restrict__ EdgeRows,
                   const TileHdr *__restrict__ Hdr) {
   // EdgeRows == nullptr: every edge once; else the OWN edges of the tiles of TR (halo split: the same edges,
   // in the same two groups, as the two launches of the tiled edge pass)
   int e = blockIdx.x * blockDim.x + threadIdx.x;
   if (EdgeRows) {
      const int t = e / kTileEdgesOwn, q = e - t * kTileEdgesOwn; // kTileEdgesOwn slots per tile
      if (t >= TR.Count) return;
      const TileHdr h = Hdr[TR.tile(t)];
      if (q >= h.NEOwn) return;
      e = EdgeRows[h.ERowOff + q];
   } else if (e >= M.NEdgesAll) return;
   const int K = M.K, KBot = K - 1;
   const Real r0 = RTop[e], rb = RBot[e];
   const size_t o = (size_t)e * K;
   if (r0 != 0.0) {
      Unext[o] += B * r0;
      if (Uprov) Uprov[o] += A * r0;
   }
   if (rb != 0.0) {
      Unext[o + KBot] += B * rb;
      if (Uprov) Uprov[o + KBot] += A * rb;
   }
}

// ---------------------------------------------------------------------------
// host side: one RK4 phase on the fused kernels
// ---------------------------------------------------------------------------
static int env_int(const char *name, int dflt) {
   const char *v = getenv(name);
   return v ? atoi(v) : dflt;
}

static unsigned fused_grid(int NRows, int K, int V, int Threads = kBlock, int Iters = 1) {
   // a CTA sweeps its row block in iterations of kBlock/KV rows; aim at 8
   // iterations per CTA but keep >= 4 CTAs per SM (148 SMs) when the mesh allows
   // one sweep iteration per CTA: consecutive CTAs (same wave, different SMs) then work on
   // SFC-adjacent row groups at the same time and meet in L2 (profiles/r01: a CTA that
   // sweeps 8 groups re-reads its neighbours' rows from DRAM, 1.85x read amplification)
   const int KV          = K / V;
   const int rowsPerIter = KV >= Threads ? 1 : Threads / KV;
   const long long per = (long long)rowsPerIter * std::max(Iters, 1);
   return (unsigned)std::max<long long>(((long long)NRows + per - 1) / per, 1);
}

// NVTX range per stage and kernel (the reference marks its timers the same way through Pacer:
// share/pacer/Pacer.cpp:157-158,191-192); a no-op without a profiler attached.
struct NvtxRange {
   explicit NvtxRange(const char *Name) { nvtxRangePushA(Name); }
   ~NvtxRange() { nvtxRangePop(); }
};

struct ScopedTimer {
   OmegaStepper *S;
   cudaStream_t St;
   OmegaStepper::Timed T;
   bool On;
   ScopedTimer(OmegaStepper *S_, cudaStream_t St_, int Kind) : S(S_), St(St_), On(S_->Profiling != 0) {
      if (On) {
         T.Kind = Kind;
         cudaEventCreate(&T.Start);
         cudaEventCreate(&T.Stop);
         cudaEventRecord(T.Start, St);
      }
   }
   ~ScopedTimer() {
      if (On) {
         cudaEventRecord(T.Stop, St);
         S->Events.push_back(T);
      }
   }
};


// One stage of a fused stepper as the kernels see it: R = RHS(stage state), then
//    next (+)= B R   (first stage: next = cur + B R; tracers weighted by thickness, TimeStepper.cpp:446-524)
//    prov  = cur + A R   (tracers: (cur*h_mul + A tend) / h_div)
// RK4 uses both outputs (accumulated next-level state + provisional state of the following stage); RK2 and
// forward-backward are written with the provisional output alone (their "next" output goes to a sink buffer
// with B = 0): a stage cannot update the next-level arrays in place, neighbours still read them.
struct StageDesc {
   const Real *Hs, *Us, *Trs;      // stage state the tendencies are evaluated on
   const Real *HsMom;              // thickness the momentum aux + edge pass see (forward-backward: the NEW thickness)
   const Real *UTr, *URaw;         // velocity seen by tracer advection / bottom drag (raw time-level indexing, P1)
   Real *Hnext, *Unext, *TrNext;   // accumulated output
   Real *Hprov, *Uprov, *TrProv;   // provisional output (nullptr: none)
   Real A, B, TimeCell, TimeEdge;
   int First, Last, WSel, H2Sel, FinSel, DivSel;
   int BeforeExchange;               // the stage's results are exchanged right after it (halo split applies)
   int SkipHaloOK;                   // ... and nothing else reads its halo rows (see fused_phase)
};

static int run_stage(OmegaStepper *S, const StageDesc &sd, const Real *Hcur, const Real *Ucur, const Real *TrCur,
                     cudaStream_t St) {
   const DMesh &D           = S->M->D;
   const OmegaTendConfig &G = S->Cfg;
   const int V = (D.K & 1) ? 1 : 2, T = S->NTracers;
   const bool vel4 = G.VelHyperDiffTendencyEnable, tr4 = G.TracerHyperDiffTendencyEnable && T > 0;
   const bool upw = G.FluxThicknessType || G.FluxTracerType;
   // sweep iterations per CTA (tuning knobs, defaults from profiles/r01 measurements)
   // A mesh with fewer (row, level-vector) items than two 256-thread CTAs per SM (config 1: 4 096 cells x 1
   // level) is bound by the latency of each thread's dependent gathers: 64-thread CTAs, one sweep iteration,
   // put the same threads on four times as many SMs.
   const bool small = (long long)D.NCellsAll * (D.K / V) < 2LL * 148 * kBlock && env_int("OMEGA_SMALL_MESH", 1);
   const int blk = small ? 64 : kBlock;
   const int itA = small ? 1 : env_int("OMEGA_ITERS_A", 1), itB = small ? 1 : env_int("OMEGA_ITERS_B", 2),
             itT = small ? 1 : env_int("OMEGA_ITERS_T2", 1), itC = small ? 1 : env_int("OMEGA_ITERS_C", 1),
             itE = small ? 1 : env_int("OMEGA_ITERS_E", 8);
   const unsigned gc = fused_grid(D.NCellsAll, D.K, V, blk, itC), ge = fused_grid(D.NEdgesAll, D.K, V, blk, itE);
   const unsigned gA = fused_grid(D.NCellsAll, D.K, V, blk, itA), gB = fused_grid(D.NCellsAll, D.K, V, blk, itB),
                  gT = fused_grid(D.NCellsAll, D.K, V, blk, itT);
#define LAUNCH_V(KIND, KERNEL, GRID, ...)                               \
   do {                                                                 \
      ScopedTimer tm_(S, St, KIND);                                     \
      NvtxRange nv_(#KERNEL);                                           \
      if (V == 2) KERNEL<2><<<GRID, blk, 0, St>>>(__VA_ARGS__);          \
      else KERNEL<1><<<GRID, blk, 0, St>>>(__VA_ARGS__);                 \
      ++S->Launches;                                                    \
   } while (0)
   // shared-memory staged passes (fused_tile.cu) replace fT2 + fC and fE when they apply; fA / fB then write
   // the combined fields the edge pass differences.  Upwind fluxes, the manufactured sources, wind forcing
   // and bottom drag all run on the tiled kernels (coefficient selection in the prologues, k_ms_sources,
   // k_edge_layer_terms).
   const bool tiledE = V == 2 && tile_edge_path_ok(D) && env_int("OMEGA_TILE_E", 1);
   const bool tiled  = V == 2 && (T >= 2 || (T == 1 && (!tr4 || tile_v2_enabled()))) && tile_path_ok(D) && env_int("OMEGA_TILE", 1) &&
                       (!tile_partial_chunk(D) || (tr4 && tile_v2_enabled()));
   const bool msOn = G.ManufacturedSolutionTendency != 0;
   const bool drag = G.BottomDragTendencyEnable != 0, wind = G.WindForcingTendencyEnable != 0;
   const Real nu2 = G.VelDiffTendencyEnable ? G.ViscDel2 : 0.0;
   if (msOn && (tiled || tiledE)) {
      const int n = std::max(D.NCellsAll, D.NEdgesAll);
      k_ms_sources<<<grid_for(n), kBlock, 0, St>>>(D, manufactured(S, sd.TimeCell), sd.TimeEdge, G.VelDiffTendencyEnable, G.ViscDel2,
                                                  G.VelHyperDiffTendencyEnable, G.ViscDel4, S->SrcH, S->SrcU);
      ++S->Launches;
   }
   // ---- cell pass first: it depends on the stage state only (and, at stage 0 with raw indexing, on the
   // stale next-level velocity the edge pass overwrites)
   CellArgs ca;
   ca.Hs = sd.Hs; ca.Us = sd.Us; ca.Trs = sd.Trs; ca.Del2Tr = S->A.Del2Tr;
   ca.UTr = sd.UTr;
   ca.Hcur = Hcur; ca.Hnext = sd.Hnext; ca.Hprov = sd.Hprov;
   ca.TrCur = TrCur; ca.TrNext = sd.TrNext; ca.TrProv = sd.TrProv;
   ca.A = sd.A; ca.B = sd.B;
   ca.First = sd.First; ca.Last = sd.Last; ca.NTracers = T;
   ca.WSel = sd.WSel; ca.H2Sel = sd.H2Sel; ca.FinSel = sd.FinSel; ca.DivSel = sd.DivSel;
   ca.FluxDivOn = G.ThicknessFluxTendencyEnable; ca.AdvOn = G.TracerHorzAdvTendencyEnable;
   ca.DiffOn = G.TracerDiffTendencyEnable; ca.HypOn = tr4;
   ca.FluxUpwind = G.FluxThicknessType; ca.TrUpwind = G.FluxTracerType;
   ca.EddyDiff2 = G.EddyDiff2; ca.EddyDiff4 = G.EddyDiff4;
   ca.Prefetch = env_int("OMEGA_FC_PREFETCH", 0);
   ca.MS = manufactured(S, sd.TimeCell);
   ca.SrcH = msOn ? S->SrcH : nullptr;
   S->LastPath = OMEGA_PATH_FUSED | (tiled ? OMEGA_PATH_TILED_CELL : 0) | (tiled && tile_v2_used(ca) ? OMEGA_PATH_CELL_V2 : 0) |
                 (tiledE ? OMEGA_PATH_TILED_EDGE : 0);
   // With a halo attached (multi-rank): the tiles that hold send-list rows run first in a stage right
   // before an exchange; once they are done the exchange starts on the halo's own stream (step.cu) while the
   // interior tiles still compute.  The halo tiles are skipped there -- the exchange overwrites every
   // non-owned row -- unless the stage's halo rows are read again before that (SkipHaloOK = 0: RK4 stage 1
   // with raw time-level indexing and CurTimeIndex == 1, SURVEY 3.5 P1: the stage-2 tracer update multiplies
   // by the ACCUMULATING next-level thickness, whose stage-1 contribution on the first halo layer is valid and
   // needed); the halo tiles then run with the boundary tiles, ahead of the exchange that writes the same rows.
   const OmegaHalo *Hl = S->Halo;
   const bool split = Hl && Hl->DTileOrder && Hl->HaloCovered && tiled && tiledE && sd.BeforeExchange &&
                      env_int("OMEGA_HALO_OVERLAP", 1);
   const TileRange trAll = all_tiles(D);
   TileRange trBnd = trAll, trInt = trAll;
   if (split) {
      const int nB = Hl->NTilesBoundary, nO = Hl->NTilesOwned, nH = D.Tiles.NTiles - nO;
      // the first launch is filled up with interior tiles to whole rounds of CTAs (one tile CTA per SM in the
      // cell pass, 2 x 4 chunk CTAs per SM = 74 tiles per round in the edge pass): a launch of the boundary
      // tiles alone is a fraction of a round and costs a whole one
      const int first = sd.SkipHaloOK ? nB : nB + nH, total = sd.SkipHaloOK ? nO : nO + nH;
      const int n1 = std::min(total, (first + Hl->NumSMs - 1) / Hl->NumSMs * Hl->NumSMs);
      const int *map = sd.SkipHaloOK ? Hl->DTileOrder : Hl->DTileOrderBH;
      trBnd = TileRange{map, 0, n1};
      trInt = TileRange{map, n1, total - n1};
      S->LastPath |= OMEGA_PATH_HALO_SPLIT;
   }
   if (tr4 && !tiled) LAUNCH_V(kFT2, k_fT2, gT, D, sd.Hs, sd.Trs, T, S->A.Del2Tr);
   {
      ScopedTimer tm_(S, St, kFC);
      NvtxRange nv_("cell pass");
      const int variant = env_int("OMEGA_FC_VARIANT", 3);
      if (tiled) {
         if (launch_fCT(D, ca, St, trBnd)) return 1;
      } else if (upw) {
         if (V == 2) k_fC<2, true><<<gc, blk, 0, St>>>(D, ca); else k_fC<1, true><<<gc, blk, 0, St>>>(D, ca);
      } else if (small) {
         if (V == 2) k_fC_center<2, 512, 1><<<gc, blk, 0, St>>>(D, ca); else k_fC_center<1, 512, 1><<<gc, blk, 0, St>>>(D, ca);
      } else if (V == 1) {
         k_fC_center<1, 512, 1><<<fused_grid(D.NCellsAll, D.K, 1, 512, itC), 512, 0, St>>>(D, ca);
      } else if (variant == 0) {
         k_fC_center<2, 512, 1><<<fused_grid(D.NCellsAll, D.K, 2, 512, itC), 512, 0, St>>>(D, ca);
      } else if (variant == 1) { // scalar levels, 2 CTAs of 512 (<= 64 regs)
         k_fC_center<1, 512, 2><<<fused_grid(D.NCellsAll, D.K, 1, 512, itC), 512, 0, St>>>(D, ca);
      } else if (variant == 2) { // 3 CTAs of 256 (<= 85 regs)
         k_fC_center<2, 256, 3><<<fused_grid(D.NCellsAll, D.K, 2, 256, itC), 256, 0, St>>>(D, ca);
      } else if (variant == 3) { // 1024 threads (<= 64 regs)
         k_fC_center<2, 1024, 1><<<fused_grid(D.NCellsAll, D.K, 2, 1024, itC), 1024, 0, St>>>(D, ca);
      } else if (variant == 4) { // scalar levels, 1 CTA of 1024
         k_fC_center<1, 1024, 1><<<fused_grid(D.NCellsAll, D.K, 1, 1024, itC), 1024, 0, St>>>(D, ca);
      } else { // 2 CTAs of 256, V=2 (<= 128 regs)
         k_fC_center<2, 256, 2><<<fused_grid(D.NCellsAll, D.K, 2, 256, itC), 256, 0, St>>>(D, ca);
      }
      ++S->Launches;
   }
   // ---- momentum: aux fields, then the edge pass
   const Real *Hm = sd.HsMom;
   // OMEGA_PDL=1 (measured, off by default): k_fA reads the stage state only and writes auxiliary fields the cell
   // pass never touches, so behind the tiled cell pass (which signals griddepcontrol.launch_dependents) it can be
   // launched with programmatic stream serialization and run on the SMs the cell pass's last round of CTAs
   // leaves idle; the same for the interior tiles' edge pass behind their cell pass.  Config 3 on one GPU:
   // 36.9 -> 37.0 ms/step -- the tail it fills is 0.3 of a round on a fifth of the SMs (< 0.01 ms of work) and the
   // early kernel competes with the cell pass's last CTAs for HBM.  Not while per-kernel events are recorded
   // between the launches (profiling), and not when the momentum thickness is one the cell pass writes
   // (forward-backward).
   const bool pdl = tiled && V == 2 && !S->Profiling && Hm != sd.Hnext && Hm != sd.Hprov && env_int("OMEGA_PDL", 0);
   if (pdl) {
      NvtxRange nv_("k_fA");
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(gA);
      cfg.blockDim = dim3(blk);
      cfg.stream = St;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      at[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = at;
      cfg.numAttrs = 1;
      OMEGA_CUDA_CHECK(cudaLaunchKernelEx(&cfg, k_fA<2>, D, Hm, sd.Us, S->A.KE, S->A.Div, S->A.RelVort, S->A.NormRelVort, tiledE ? 1 : 0,
                                          (Real)(G.KETendencyEnable ? 1.0 : 0.0), (Real)(G.SSHTendencyEnable ? 1.0 : 0.0),
                                          tiledE && drag ? S->A.ZonalStress2 : (Real *)nullptr));
      ++S->Launches;
   } else
   LAUNCH_V(kFA, k_fA, gA, D, Hm, sd.Us, S->A.KE, S->A.Div, S->A.RelVort, S->A.NormRelVort, tiledE ? 1 : 0,
            G.KETendencyEnable ? 1.0 : 0.0, G.SSHTendencyEnable ? 1.0 : 0.0, tiledE && drag ? S->A.ZonalStress2 : nullptr);
   // W = Y - X folded into fB's cell result when the tiled edge pass follows (one gathered cell field less)
   const bool xOn    = G.KETendencyEnable || G.SSHTendencyEnable;
   const bool fB3    = vel4 && D.B3OK && env_int("OMEGA_FB3", 1);
   const bool wMerge = fB3 && tiledE && xOn && env_int("OMEGA_FUSE_W", 1);
   if (fB3)
      LAUNCH_V(kFB, k_fB3, gB, D, S->A.Div, S->A.RelVort, S->A.Del2Div, S->A.Del2RelVort, tiledE ? 1 : 0, nu2,
               G.ViscDel4 * G.DivFactor, G.ViscDel4, wMerge ? S->A.KE : nullptr);
   else if (vel4)
      LAUNCH_V(kFB, k_fB, gB, D, S->A.Div, S->A.RelVort, S->A.Del2Div, S->A.Del2RelVort, tiledE ? 1 : 0, nu2,
               G.ViscDel4 * G.DivFactor, G.ViscDel4);
   EdgeArgs ea;
   ea.Hs = Hm; ea.Us = sd.Us; ea.URaw = sd.URaw;
   ea.KE = S->A.KE; ea.Div = S->A.Div; ea.RelVort = S->A.RelVort; ea.PVv = S->A.NormRelVort;
   ea.Del2Div = S->A.Del2Div; ea.Del2RelVort = S->A.Del2RelVort; ea.NormalStress = S->A.NormalStress;
   ea.Ucur = Ucur; ea.Unext = sd.Unext; ea.Uprov = sd.Uprov;
   ea.A = sd.A; ea.B = sd.B; ea.First = sd.First; ea.Last = sd.Last;
   ea.G = G; ea.MS = manufactured(S, sd.TimeEdge);
   EdgeTileArgs ta;
   if (tiledE && (wind || drag)) {
      k_edge_layer_terms<<<grid_for(D.NEdgesAll), kBlock, 0, St>>>(D, Hm, ea.URaw, S->A.ZonalStress2, S->A.NormalStress, wind, drag,
                                                                  1.0 / G.Density0, G.BottomDragCoeff, S->RTop, S->RBot);
      ++S->Launches;
   }
   if (tiledE) {
      ScopedTimer tm_(S, St, kFE);
      NvtxRange nv_("edge pass");
      ta.Us = sd.Us; ta.Hs = Hm; ta.PVv = S->A.NormRelVort; ta.X = S->A.KE;
      ta.Y = vel4 ? S->A.Del2Div : S->A.Div; ta.Z = vel4 ? S->A.Del2RelVort : S->A.RelVort;
      ta.SY = vel4 ? 1.0 : nu2; ta.SZ = ta.SY;
      ta.Ucur = Ucur; ta.Unext = sd.Unext; ta.Uprov = sd.Uprov;
      ta.A = sd.A; ta.B = sd.B; ta.First = sd.First; ta.Last = sd.Last;
      ta.PVOn = G.PVTendencyEnable; ta.XOn = xOn && !wMerge;
      ta.FluxUpwind = G.FluxThicknessType; ta.SrcU = msOn ? S->SrcU : nullptr;
      if (launch_fET(D, ta, St, trBnd)) return 1;
      ++S->Launches;
   } else {
      // occupancy variants measured on B200 (profiles/r01): 4 CTAs/SM (64 regs, some spills) is fastest
      const int fev = env_int("OMEGA_FE_VARIANT", 3);
      ScopedTimer tm_(S, St, kFE);
      NvtxRange nv_("edge pass");
      if (V == 1) k_fE<1, 2><<<ge, blk, 0, St>>>(D, ea);
      else if (fev == 2) k_fE<2, 3><<<ge, blk, 0, St>>>(D, ea);
      else if (fev == 3) k_fE<2, 4><<<ge, blk, 0, St>>>(D, ea);
      else k_fE<2, 2><<<ge, blk, 0, St>>>(D, ea);
      ++S->Launches;
   }
   if (split) {
      // boundary tiles of both passes are queued: the exchange may start behind them.  (The single-layer
      // wind / drag correction below touches send rows too, so it must precede the event.)
      if ((wind || drag) && trBnd.Count > 0) {
         k_edge_layer_apply<<<grid_for((long long)trBnd.Count * kTileEdgesOwn), kBlock, 0, St>>>(D, S->RTop, S->RBot, sd.Unext, sd.Uprov, sd.A,
                                                                                                sd.B, trBnd, D.Tiles.EdgeRows, D.Tiles.Hdr);
         ++S->Launches;
      }
      OMEGA_CUDA_CHECK(cudaEventRecord(Hl->EvReady, St));
      S->SplitRecorded = 1;
      {
         ScopedTimer tm_(S, St, kFC);
         if (launch_fCT(D, ca, St, trInt)) return 1;
         ++S->Launches;
      }
      {
         // the edge pass of the interior tiles does not depend on their cell pass queued right before it (unless
         // it writes the velocity level the cell pass still reads: stage 0 with raw time-level indexing)
         const bool pdlE = pdl && sd.Unext != sd.Us && sd.Unext != sd.UTr && sd.Uprov != sd.Us && sd.Uprov != sd.UTr;
         ScopedTimer tm_(S, St, kFE);
         if (launch_fET(D, ta, St, trInt, pdlE)) return 1;
         ++S->Launches;
      }
      if ((wind || drag) && trInt.Count > 0) {
         k_edge_layer_apply<<<grid_for((long long)trInt.Count * kTileEdgesOwn), kBlock, 0, St>>>(D, S->RTop, S->RBot, sd.Unext, sd.Uprov, sd.A,
                                                                                                sd.B, trInt, D.Tiles.EdgeRows, D.Tiles.Hdr);
         ++S->Launches;
      }
   } else if (tiledE && (wind || drag)) {
      k_edge_layer_apply<<<grid_for(D.NEdgesAll), kBlock, 0, St>>>(D, S->RTop, S->RBot, sd.Unext, sd.Uprov, sd.A, sd.B,
                                                                  TileRange{nullptr, 0, 0}, nullptr, nullptr);
      ++S->Launches;
   }
#undef LAUNCH_V
   OMEGA_CUDA_CHECK(cudaGetLastError());
   return 0;
}

static inline int tix(int Level, int Cur, int NTL) { return (Level + Cur + NTL) % NTL; }

// One phase (everything up to the next halo exchange point) of a stepper on the fused kernels:
// RungeKutta4Stepper.cpp:68-137 (two phases), RungeKutta2Stepper.cpp:27-73, ForwardBackwardStepper.cpp:49-68.
int fused_phase(OmegaStepper *S, OmegaStepperType Type, OmegaStateArrays *X, const OmegaStepCoeffs *C, int Phase, int Compat,
                cudaStream_t St) {
   const DMesh &D           = S->M->D;
   const OmegaTendConfig &G = S->Cfg;
   OMEGA_REQUIRE(S->M->MaxEdgesUsed <= kME, "fused path supports at most 6 edges per cell; use Fused=0");
   const int ci = X->CurTimeIndex, ni = tix(1, ci, 2);
   const int tci = X->TracerCurTimeIndex, tni = tix(1, tci, 2);
   Real *Hcur = X->LayerThickness[ci], *Hnext = X->LayerThickness[ni];
   Real *Ucur = X->NormalVelocity[ci], *Unext = X->NormalVelocity[ni];
   Real *TrCur = X->Tracers[tci], *TrNext = X->Tracers[tni];
   // raw time-level indexing of the reference: with CurTimeIndex == 1 the "current" raw
   // slot is the NEXT buffer and vice versa
   const bool swapped = Compat && ci == 1;
   S->ProvisExchangeSlot = 0;
   Launch L{St, &S->Launches};
   if (G.WindForcingTendencyEnable && Phase == 0) launch_wind_aux(L, D, S->A, G.WindStressInterpType);
   StageDesc sd;
   std::memset(&sd, 0, sizeof(sd));
   if (Type == OMEGA_STEPPER_RK4) {
      const int s0 = Phase == 0 ? 0 : 2, s1 = Phase == 0 ? 1 : 3;
      for (int s = s0; s <= s1; ++s) {
         NvtxRange nv_(s == 0 ? "RK4 stage 0" : s == 1 ? "RK4 stage 1" : s == 2 ? "RK4 stage 2" : "RK4 stage 3");
         // provisional state of stage s lives in P[s % 2] (stage 2 -> slot 0, the exchanged one)
         sd.Hs = s == 0 ? Hcur : S->P.H[s % 2];
         sd.Us = s == 0 ? Ucur : S->P.U[s % 2];
         sd.Trs = s == 0 ? TrCur : S->P.Tr[s % 2];
         sd.HsMom = sd.Hs;
         sd.Hprov = s == 3 ? nullptr : S->P.H[(s + 1) % 2];
         sd.Uprov = s == 3 ? nullptr : S->P.U[(s + 1) % 2];
         sd.TrProv = s == 3 ? nullptr : S->P.Tr[(s + 1) % 2];
         sd.Hnext = Hnext; sd.Unext = Unext; sd.TrNext = TrNext;
         sd.UTr = (s == 0 && swapped) ? Unext : sd.Us;
         sd.URaw = sd.UTr;
         sd.A = s < 3 ? C->A[s + 1] : 0.0; sd.B = C->B[s];
         sd.TimeCell = sd.TimeEdge = C->StageTime[s];
         sd.First = s == 0; sd.Last = s == 3;
         sd.WSel = swapped; sd.H2Sel = swapped; sd.FinSel = swapped; sd.DivSel = 0;
         sd.BeforeExchange = s == 1 || s == 3;
         sd.SkipHaloOK = s == 3 || !swapped;
         if (run_stage(S, sd, Hcur, Ucur, TrCur, St)) return 1;
      }
   } else if (Type == OMEGA_STEPPER_RK2) {
      // both stages are "prov = cur + c R(in)": stage 0 into the provisional slot, stage 1 into the next-level arrays
      // (the reference overwrites the next level in place, its tendencies sit in separate arrays).  Tracers:
      // (cur * hraw(Cur) + c tend) / hraw(Next) -- raw slots, so with CurTimeIndex == 1 the roles of the
      // current and the freshly updated thickness swap (RungeKutta2Stepper.cpp:44-47,64-67; P1).
      for (int s = 0; s < 2; ++s) {
         NvtxRange nv_(s == 0 ? "RK2 stage 0" : "RK2 stage 1");
         sd.Hs = s == 0 ? Hcur : S->P.H[0];
         sd.Us = s == 0 ? Ucur : S->P.U[0];
         sd.Trs = s == 0 ? TrCur : S->P.Tr[0];
         sd.HsMom = sd.Hs;
         sd.Hprov = s == 0 ? S->P.H[0] : Hnext;
         sd.Uprov = s == 0 ? S->P.U[0] : Unext;
         sd.TrProv = s == 0 ? S->P.Tr[0] : TrNext;
         sd.Hnext = S->P.H[1]; sd.Unext = S->P.U[1]; sd.TrNext = S->P.Tr[1]; // sink (B = 0)
         // velocity by raw index VelTimeLevel: stage 0 raw 0, stage 1 raw 1
         sd.UTr = !swapped ? sd.Us : (s == 0 ? Unext : Ucur);
         sd.URaw = sd.UTr;
         sd.A = s == 0 ? C->A[1] : C->B[1]; sd.B = 0.0;
         sd.TimeCell = sd.TimeEdge = C->StageTime[s];
         sd.First = s == 0; sd.Last = 0;
         sd.WSel = 0; sd.FinSel = 0;
         sd.H2Sel = swapped ? 2 : 0; sd.DivSel = swapped ? 1 : 0;
         sd.BeforeExchange = s == 1; sd.SkipHaloOK = 1;
         if (run_stage(S, sd, Hcur, Ucur, TrCur, St)) return 1;
      }
   } else {
      // forward-backward: thickness and tracers from the current state, then the velocity tendency on the NEW
      // thickness (ForwardBackwardStepper.cpp:49-68).  With raw indexing and CurTimeIndex == 1 the tracer
      // fluxes would see the new thickness while the thickness flux sees the old one; step.cu sends that case
      // to the kernel-per-term path.
      NvtxRange nv_("forward-backward");
      sd.Hs = Hcur; sd.Us = Ucur; sd.Trs = TrCur;
      sd.HsMom = Hnext;
      sd.Hprov = Hnext; sd.Uprov = Unext; sd.TrProv = TrNext;
      sd.Hnext = S->P.H[1]; sd.Unext = S->P.U[1]; sd.TrNext = S->P.Tr[1]; // sink (B = 0)
      sd.UTr = Ucur; sd.URaw = Ucur;
      sd.A = C->Dt; sd.B = 0.0;
      sd.TimeCell = C->StageTime[0]; sd.TimeEdge = C->StageTime[1];
      sd.First = 1; sd.Last = 0;
      sd.WSel = 0; sd.FinSel = 0; sd.H2Sel = 0; sd.DivSel = 0;
      sd.BeforeExchange = 1; sd.SkipHaloOK = 1;
      if (run_stage(S, sd, Hcur, Ucur, TrCur, St)) return 1;
   }
   return 0;
}

} // namespace omega
