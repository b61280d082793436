/* omega_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C, single-thread-or-OpenMP CPU restatement of the Omega per-step hot
 * path, used only as the parity checker by tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs.  Nothing under
 * omega_b200/ may call it.
 *
 * Parity pin: the reference's own tests hold no golden vector for a full
 * right-hand side or a full step, only analytic error
 * norms on three downloaded meshes.  This file is pinned two ways:
 *   (1) bit-for-bit against oracle/_ref/libomega_ref.so, which is the
 *       reference's UNMODIFIED functor headers and sequencing sources
 *       compiled over a host Kokkos stand-in (oracle/ref_shim/,
 *       tests/test_oracle_vs_ref.py);
 *   (2) against the plane known-answer norms of
 *       test/ocn/TendencyTermsTest.cpp:43-59, AuxiliaryVarsTest.cpp:34-68,
 *       HorzOperatorsTest.cpp:33-43 on a regenerated 48x48 periodic hex mesh
 *       (tests/test_oracle_kat.py).
 *
 * Every function cites the reference file:line (relative to
 * components/omega/) whose operation order it follows.  Operation order per
 * element is kept exactly (J-loop order, association), build with
 * -ffp-contract=off for the strict variant.
 */
#include "omega_b200.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define PFOR _Pragma("omp parallel for schedule(static)")

typedef struct OracleStepper {
   OmegaMeshDesc M;
   OmegaTendConfig Cfg;
   int NTracers, K;
   /* HorzMesh derived fields, src/ocn/HorzMesh.cpp:527-626 */
   double *EdgeSignOnCell, *EdgeSignOnVertex, *EdgeMask, *MeshScalingDel2, *MeshScalingDel4;
   OmegaAuxArrays Aux;
   OmegaTendArrays Tend;
   /* RungeKutta4Stepper::finalizeInit, src/timeStepping/RungeKutta4Stepper.cpp:43-64 */
   double *ProvisH, *ProvisU, *ProvisTr;
   /* Tendencies::CustomThicknessTend / CustomVelocityTend, src/ocn/Tendencies.h:51-53,182-183 */
   OmegaCustomTendency CustomThick, CustomVel;
   void *CustomUser;
} OracleStepper;

static double *zalloc(size_t n) {
   double *p = (double *)calloc(n ? n : 1, sizeof(double));
   if (!p) {
      fprintf(stderr, "oracle: out of memory\n");
      abort();
   }
   return p;
}

/* ------------------------------------------------------------------------ */
/* HorzMesh::computeEdgeSign / setMasks / setMeshScaling                     */
/* src/ocn/HorzMesh.cpp:527-575, 581-602, 607-626                            */
/* ------------------------------------------------------------------------ */
static void derive_mesh(OracleStepper *S) {
   const OmegaMeshDesc *M = &S->M;
   S->EdgeSignOnCell   = zalloc((size_t)M->NCellsSize * M->MaxEdges);
   S->EdgeSignOnVertex = zalloc((size_t)M->NVerticesSize * M->VertexDegree);
   S->EdgeMask         = zalloc((size_t)M->NEdgesSize);
   S->MeshScalingDel2  = zalloc((size_t)M->NEdgesSize);
   S->MeshScalingDel4  = zalloc((size_t)M->NEdgesSize);
   for (int c = 0; c < M->NCellsAll; ++c)
      for (int i = 0; i < M->NEdgesOnCell[c]; ++i) {
         int e = M->EdgesOnCell[c * M->MaxEdges + i];
         S->EdgeSignOnCell[c * M->MaxEdges + i] = (c == M->CellsOnEdge[2 * e]) ? -1.0 : 1.0;
      }
   for (int v = 0; v < M->NVerticesAll; ++v)
      for (int i = 0; i < M->VertexDegree; ++i) {
         int e = M->EdgesOnVertex[v * M->VertexDegree + i];
         S->EdgeSignOnVertex[v * M->VertexDegree + i] =
             (v == M->VerticesOnEdge[2 * e]) ? -1.0 : 1.0;
      }
   /* deepCopy(EdgeMask, 1.0) covers the padding row too */
   for (int e = 0; e < M->NEdgesSize; ++e)
      S->EdgeMask[e] = 1.0;
   for (int e = 0; e < M->NEdgesAll; ++e) {
      int c1 = M->CellsOnEdge[2 * e], c2 = M->CellsOnEdge[2 * e + 1];
      if (!(c1 >= 0 && c1 < M->NCellsAll) || !(c2 >= 0 && c2 < M->NCellsAll))
         S->EdgeMask[e] = 0.0;
   }
   for (int e = 0; e < M->NEdgesAll; ++e) {
      S->MeshScalingDel2[e] = 1.0;
      S->MeshScalingDel4[e] = 1.0;
   }
}

int oracle_stepper_create(const OmegaMeshDesc *Desc, const OmegaTendConfig *Cfg, int NTracers,
                          OracleStepper **Out) {
   OracleStepper *S = (OracleStepper *)calloc(1, sizeof(OracleStepper));
   S->M        = *Desc;
   S->Cfg      = *Cfg;
   S->NTracers = NTracers;
   S->K        = Desc->NVertLayers;
   derive_mesh(S);
   const size_t K = S->K, C = Desc->NCellsSize, E = Desc->NEdgesSize, V = Desc->NVerticesSize,
                T = NTracers;
   OmegaAuxArrays *A = &S->Aux;
   A->KineticEnergyCell = zalloc(C * K);   A->VelocityDivCell = zalloc(C * K);
   A->FluxLayerThickEdge = zalloc(E * K);  A->MeanLayerThickEdge = zalloc(E * K);
   A->SshCell = zalloc(C * K);
   A->RelVortVertex = zalloc(V * K);       A->NormRelVortVertex = zalloc(V * K);
   A->NormPlanetVortVertex = zalloc(V * K);
   A->NormRelVortEdge = zalloc(E * K);     A->NormPlanetVortEdge = zalloc(E * K);
   A->Del2Edge = zalloc(E * K);            A->Del2DivCell = zalloc(C * K);
   A->Del2RelVortVertex = zalloc(V * K);
   A->HTracersEdge = zalloc(T * E * K);    A->Del2TracersCell = zalloc(T * C * K);
   A->NormalStressEdge = zalloc(E);        A->ZonalStressCell = zalloc(C);
   A->MeridStressCell = zalloc(C);
   S->Tend.LayerThicknessTend = zalloc(C * K);
   S->Tend.NormalVelocityTend = zalloc(E * K);
   S->Tend.TracerTend         = zalloc(T * C * K);
   S->ProvisH = zalloc(C * K);  S->ProvisU = zalloc(E * K);  S->ProvisTr = zalloc(T * C * K);
   *Out = S;
   return 0;
}

int oracle_stepper_destroy(OracleStepper *S) {
   if (!S) return 0;
   double **p = (double **)&S->Aux;
   for (size_t i = 0; i < sizeof(OmegaAuxArrays) / sizeof(double *); ++i) free(p[i]);
   free(S->Tend.LayerThicknessTend); free(S->Tend.NormalVelocityTend); free(S->Tend.TracerTend);
   free(S->ProvisH); free(S->ProvisU); free(S->ProvisTr);
   free(S->EdgeSignOnCell); free(S->EdgeSignOnVertex); free(S->EdgeMask);
   free(S->MeshScalingDel2); free(S->MeshScalingDel4);
   free(S);
   return 0;
}

int oracle_stepper_set_config(OracleStepper *S, const OmegaTendConfig *Cfg) { S->Cfg = *Cfg; return 0; }
int oracle_stepper_aux_arrays(OracleStepper *S, OmegaAuxArrays *A) { *A = S->Aux; return 0; }
int oracle_stepper_tend_arrays(OracleStepper *S, OmegaTendArrays *T) { *T = S->Tend; return 0; }
int oracle_stepper_provis_arrays(OracleStepper *S, double **H, double **U, double **Tr) {
   *H = S->ProvisH; *U = S->ProvisU; *Tr = S->ProvisTr; return 0;
}
int oracle_mesh_derived(OracleStepper *S, double *ESC, double *ESV, double *Mask) {
   memcpy(ESC, S->EdgeSignOnCell, sizeof(double) * S->M.NCellsSize * S->M.MaxEdges);
   memcpy(ESV, S->EdgeSignOnVertex, sizeof(double) * S->M.NVerticesSize * S->M.VertexDegree);
   memcpy(Mask, S->EdgeMask, sizeof(double) * S->M.NEdgesSize);
   return 0;
}
int oracle_stepper_set_wind_stress(OracleStepper *S, const double *Zonal, const double *Merid) {
   memcpy(S->Aux.ZonalStressCell, Zonal, sizeof(double) * S->M.NCellsSize);
   memcpy(S->Aux.MeridStressCell, Merid, sizeof(double) * S->M.NCellsSize);
   return 0;
}

/* ------------------------------------------------------------------------ */
/* auxiliary variables                                                       */
/* ------------------------------------------------------------------------ */

/* InterpCellToEdge, src/ocn/HorzOperators.h:137-187 */
static double interp_cell_to_edge(const OracleStepper *S, int e, const double *Cell, int Aniso) {
   const OmegaMeshDesc *M = &S->M;
   if (Aniso) {
      int c0 = M->CellsOnEdge[2 * e], c1 = M->CellsOnEdge[2 * e + 1];
      return 0.5 * (Cell[c0] + Cell[c1]);
   }
   double Accum = 0, AreaAccum = 0;
   for (int j = 0; j < 2; ++j) {
      int v = M->VerticesOnEdge[2 * e + j];
      for (int l = 0; l < M->VertexDegree; ++l) {
         double Kite = M->KiteAreasOnVertex[v * M->VertexDegree + l];
         int c       = M->CellsOnVertex[v * M->VertexDegree + l];
         Accum += Cell[c] * Kite;
         AreaAccum += Kite;
      }
   }
   double Inv = 1.0 / AreaAccum;
   return Accum * Inv;
}

/* AuxiliaryState::computeMomAux, src/ocn/AuxiliaryState.cpp:60-143 */
int oracle_aux_compute_mom(OracleStepper *S, const double *H, const double *U) {
   const OmegaMeshDesc *M = &S->M;
   const OmegaAuxArrays *A = &S->Aux;
   const int K = S->K, VD = M->VertexDegree, ME = M->MaxEdges;

   /* vertexAuxState1: VorticityAuxVars::computeVarsOnVertex, VorticityAuxVars.h:24-59 */
   PFOR for (int v = 0; v < M->NVerticesAll; ++v) {
      const double InvAt = 1.0 / M->AreaTriangle[v];
      for (int k = 0; k < K; ++k) {
         double hv = 0, rv = 0;
         for (int j = 0; j < VD; ++j) {
            int c = M->CellsOnVertex[v * VD + j], e = M->EdgesOnVertex[v * VD + j];
            hv += InvAt * M->KiteAreasOnVertex[v * VD + j] * H[(size_t)c * K + k];
            rv += InvAt * M->DcEdge[e] * S->EdgeSignOnVertex[v * VD + j] * U[(size_t)e * K + k];
         }
         const double InvHv = 1.0 / hv;
         A->RelVortVertex[(size_t)v * K + k]        = rv;
         A->NormRelVortVertex[(size_t)v * K + k]    = rv * InvHv;
         A->NormPlanetVortVertex[(size_t)v * K + k] = M->FVertex[v] * InvHv;
      }
   }
   /* cellAuxState1: KineticAuxVars::computeVarsOnCell, KineticAuxVars.h:20-47 */
   PFOR for (int c = 0; c < M->NCellsAll; ++c) {
      const double InvA = 1.0 / M->AreaCell[c];
      for (int k = 0; k < K; ++k) {
         double ke = 0, dv = 0;
         for (int j = 0; j < M->NEdgesOnCell[c]; ++j) {
            int e             = M->EdgesOnCell[c * ME + j];
            double AreaEdge   = 0.5 * M->DvEdge[e] * M->DcEdge[e];
            double u          = U[(size_t)e * K + k];
            ke += AreaEdge * 0.5 * InvA * u * u;
            dv -= M->DvEdge[e] * InvA * S->EdgeSignOnCell[c * ME + j] * u;
         }
         A->KineticEnergyCell[(size_t)c * K + k] = ke;
         A->VelocityDivCell[(size_t)c * K + k]   = dv;
      }
   }
   /* edgeAuxState1: WindForcingAuxVars::computeVarsOnEdge, WindForcingAuxVars.h:22-29 */
   PFOR for (int e = 0; e < M->NEdgesAll; ++e) {
      double zs = interp_cell_to_edge(S, e, A->ZonalStressCell, S->Cfg.WindStressInterpType);
      double ms = interp_cell_to_edge(S, e, A->MeridStressCell, S->Cfg.WindStressInterpType);
      A->NormalStressEdge[e] = cos(M->AngleEdge[e]) * zs + sin(M->AngleEdge[e]) * ms;
   }
   /* edgeAuxState2: VorticityAuxVars.h:61-76, LayerThicknessAuxVars.h:25-61,
      VelocityDel2AuxVars.h:21-45 */
   PFOR for (int e = 0; e < M->NEdgesAll; ++e) {
      const int c0 = M->CellsOnEdge[2 * e], c1 = M->CellsOnEdge[2 * e + 1];
      const int v0 = M->VerticesOnEdge[2 * e], v1 = M->VerticesOnEdge[2 * e + 1];
      const double InvDc = 1.0 / M->DcEdge[e];
      const double InvDv = 1.0 / fmax(M->DvEdge[e], 0.25 * M->DcEdge[e]);
      for (int k = 0; k < K; ++k) {
         size_t i = (size_t)e * K + k;
         A->NormRelVortEdge[i] = 0.5 * (A->NormRelVortVertex[(size_t)v0 * K + k] +
                                        A->NormRelVortVertex[(size_t)v1 * K + k]);
         A->NormPlanetVortEdge[i] = 0.5 * (A->NormPlanetVortVertex[(size_t)v0 * K + k] +
                                           A->NormPlanetVortVertex[(size_t)v1 * K + k]);
         double h0 = H[(size_t)c0 * K + k], h1 = H[(size_t)c1 * K + k];
         A->MeanLayerThickEdge[i] = 0.5 * (h0 + h1);
         if (S->Cfg.FluxThicknessType == 0) {
            A->FluxLayerThickEdge[i] = 0.5 * (h0 + h1);
         } else {
            double u = U[i];
            if (u > 0) A->FluxLayerThickEdge[i] = h0;
            else if (u < 0) A->FluxLayerThickEdge[i] = h1;
            else A->FluxLayerThickEdge[i] = fmax(h0, h1);
         }
         double GradDiv = (A->VelocityDivCell[(size_t)c1 * K + k] -
                           A->VelocityDivCell[(size_t)c0 * K + k]) * InvDc;
         double CurlVort = -(A->RelVortVertex[(size_t)v1 * K + k] -
                             A->RelVortVertex[(size_t)v0 * K + k]) * InvDv;
         A->Del2Edge[i] = S->EdgeMask[e] * GradDiv + CurlVort;
      }
   }
   /* vertexAuxState2: VelocityDel2AuxVars::computeVarsOnVertex, VelocityDel2AuxVars.h:69-89 */
   PFOR for (int v = 0; v < M->NVerticesAll; ++v) {
      const double InvAt = 1.0 / M->AreaTriangle[v];
      for (int k = 0; k < K; ++k) {
         double t = 0;
         for (int j = 0; j < VD; ++j) {
            int e = M->EdgesOnVertex[v * VD + j];
            t += InvAt * M->DcEdge[e] * S->EdgeSignOnVertex[v * VD + j] * A->Del2Edge[(size_t)e * K + k];
         }
         A->Del2RelVortVertex[(size_t)v * K + k] = t;
      }
   }
   /* cellAuxState2: VelocityDel2AuxVars::computeVarsOnCell, VelocityDel2AuxVars.h:47-67 */
   PFOR for (int c = 0; c < M->NCellsAll; ++c) {
      const double InvA = 1.0 / M->AreaCell[c];
      for (int k = 0; k < K; ++k) {
         double t = 0;
         for (int j = 0; j < M->NEdgesOnCell[c]; ++j) {
            int e = M->EdgesOnCell[c * ME + j];
            t -= M->DvEdge[e] * InvA * S->EdgeSignOnCell[c * ME + j] * A->Del2Edge[(size_t)e * K + k];
         }
         A->Del2DivCell[(size_t)c * K + k] = t;
      }
   }
   /* cellAuxState3: LayerThicknessAuxVars::computeVarsOnCells, LayerThicknessAuxVars.h:63-82 */
   PFOR for (int c = 0; c < M->NCellsAll; ++c)
      for (int k = 0; k < K; ++k)
         A->SshCell[(size_t)c * K + k] = H[(size_t)c * K + k] - M->BottomDepth[c];
   return 0;
}

/* LayerThicknessAuxVars::computeVarsOnEdge alone (Tendencies.cpp:507-512) */
static void aux_thick_edge(OracleStepper *S, const double *H, const double *U) {
   const OmegaMeshDesc *M = &S->M;
   const OmegaAuxArrays *A = &S->Aux;
   const int K = S->K;
   PFOR for (int e = 0; e < M->NEdgesAll; ++e) {
      const int c0 = M->CellsOnEdge[2 * e], c1 = M->CellsOnEdge[2 * e + 1];
      for (int k = 0; k < K; ++k) {
         size_t i  = (size_t)e * K + k;
         double h0 = H[(size_t)c0 * K + k], h1 = H[(size_t)c1 * K + k];
         A->MeanLayerThickEdge[i] = 0.5 * (h0 + h1);
         if (S->Cfg.FluxThicknessType == 0) {
            A->FluxLayerThickEdge[i] = 0.5 * (h0 + h1);
         } else {
            double u = U[i];
            if (u > 0) A->FluxLayerThickEdge[i] = h0;
            else if (u < 0) A->FluxLayerThickEdge[i] = h1;
            else A->FluxLayerThickEdge[i] = fmax(h0, h1);
         }
      }
   }
}

/* edgeAuxState4 + cellAuxState4: TracerAuxVars.h:25-59, 61-91
   (AuxiliaryState.cpp:166-181, Tendencies.cpp:552-568) */
static void aux_tracer(OracleStepper *S, const double *H, const double *U, const double *Tr) {
   const OmegaMeshDesc *M = &S->M;
   const OmegaAuxArrays *A = &S->Aux;
   const int K = S->K, ME = M->MaxEdges;
   const size_t CS = (size_t)M->NCellsSize * K, ES = (size_t)M->NEdgesSize * K;
   for (int l = 0; l < S->NTracers; ++l) {
      const double *T = Tr + l * CS;
      PFOR for (int e = 0; e < M->NEdgesAll; ++e) {
         const int c0 = M->CellsOnEdge[2 * e], c1 = M->CellsOnEdge[2 * e + 1];
         for (int k = 0; k < K; ++k) {
            double a = H[(size_t)c0 * K + k] * T[(size_t)c0 * K + k];
            double b = H[(size_t)c1 * K + k] * T[(size_t)c1 * K + k];
            double r;
            if (S->Cfg.FluxTracerType == 0) r = 0.5 * (a + b);
            else {
               double u = U[(size_t)e * K + k];
               r = (u > 0) ? a : (u < 0) ? b : fmax(a, b);
            }
            A->HTracersEdge[l * ES + (size_t)e * K + k] = r;
         }
      }
   }
   for (int l = 0; l < S->NTracers; ++l) {
      const double *T = Tr + l * CS;
      PFOR for (int c = 0; c < M->NCellsAll; ++c) {
         const double InvA = 1.0 / M->AreaCell[c];
         for (int k = 0; k < K; ++k) {
            double t = 0;
            for (int j = 0; j < M->NEdgesOnCell[c]; ++j) {
               int e  = M->EdgesOnCell[c * ME + j];
               int c0 = M->CellsOnEdge[2 * e], c1 = M->CellsOnEdge[2 * e + 1];
               double DvDc = M->DvEdge[e] / M->DcEdge[e];
               double Grad = T[(size_t)c1 * K + k] - T[(size_t)c0 * K + k];
               t -= S->EdgeMask[e] * S->EdgeSignOnCell[c * ME + j] * DvDc *
                    A->MeanLayerThickEdge[(size_t)e * K + k] * Grad;
            }
            A->Del2TracersCell[l * CS + (size_t)c * K + k] = t * InvA;
         }
      }
   }
}

/* AuxiliaryState::computeAll, src/ocn/AuxiliaryState.cpp:146-185 */
int oracle_aux_compute_all(OracleStepper *S, const double *H, const double *U, const double *Tr) {
   oracle_aux_compute_mom(S, H, U);
   aux_tracer(S, H, U, Tr);
   return 0;
}

/* ------------------------------------------------------------------------ */
/* tendencies                                                                */
/* ------------------------------------------------------------------------ */

/* ManufacturedThicknessTendency / ManufacturedVelocityTendency,
   src/ocn/CustomTendencyTerms.cpp:107-140, 145-210 (constants :83-102) */
static void manufactured_consts(const OracleStepper *S, double *H0, double *Kx, double *Ky,
                                double *AngFreq) {
   const double Grav = 9.80665, Pii = 3.141592653589793;
   *H0      = S->M.BottomDepth[0];
   *Kx      = 2.0 * Pii / S->Cfg.ManufacturedWavelengthX;
   *Ky      = 2.0 * Pii / S->Cfg.ManufacturedWavelengthY;
   *AngFreq = sqrt(*H0 * Grav * (*Kx * *Kx + *Ky * *Ky));
}

static void manufactured_thick(OracleStepper *S, double Time) {
   const OmegaMeshDesc *M = &S->M;
   double H0, Kx, Ky, W;
   manufactured_consts(S, &H0, &Kx, &Ky, &W);
   const double Eta0 = S->Cfg.ManufacturedAmplitude;
   const int K = S->K;
   PFOR for (int c = 0; c < M->NCellsAll; ++c)
      for (int k = 0; k < K; ++k) {
         double Phase = Kx * M->XCell[c] + Ky * M->YCell[c] - W * Time;
         S->Tend.LayerThicknessTend[(size_t)c * K + k] +=
             Eta0 * (-H0 * (Kx + Ky) * sin(Phase) - W * cos(Phase) +
                     Eta0 * (Kx + Ky) * cos(2.0 * Phase));
      }
}

static void manufactured_vel(OracleStepper *S, double Time) {
   const OmegaMeshDesc *M = &S->M;
   double H0, Kx, Ky, W;
   manufactured_consts(S, &H0, &Kx, &Ky, &W);
   const double Eta0 = S->Cfg.ManufacturedAmplitude, Grav = 9.80665;
   const double Kx2 = Kx * Kx, Ky2 = Ky * Ky, Kx4 = Kx2 * Kx2, Ky4 = Ky2 * Ky2;
   const int K = S->K;
   PFOR for (int e = 0; e < M->NEdgesAll; ++e)
      for (int k = 0; k < K; ++k) {
         double Phase = Kx * M->XEdge[e] + Ky * M->YEdge[e] - W * Time;
         double Src0  = W * sin(Phase) - 0.5 * Eta0 * (Kx + Ky) * sin(2.0 * Phase);
         double Uu    = Eta0 * ((-M->FEdge[e] + Grav * Kx) * cos(Phase) + Src0);
         double Vv    = Eta0 * ((M->FEdge[e] + Grav * Ky) * cos(Phase) + Src0);
         if (S->Cfg.VelDiffTendencyEnable) {
            Uu += S->Cfg.ViscDel2 * Eta0 * (Kx2 + Ky2) * cos(Phase);
            Vv += S->Cfg.ViscDel2 * Eta0 * (Kx2 + Ky2) * cos(Phase);
         }
         if (S->Cfg.VelHyperDiffTendencyEnable) {
            Uu -= S->Cfg.ViscDel4 * Eta0 * ((Kx4 + Ky4 + Kx2 * Ky2) * cos(Phase));
            Vv -= S->Cfg.ViscDel4 * Eta0 * ((Kx4 + Ky4 + Kx2 * Ky2) * cos(Phase));
         }
         S->Tend.NormalVelocityTend[(size_t)e * K + k] +=
             cos(M->AngleEdge[e]) * Uu + sin(M->AngleEdge[e]) * Vv;
      }
}

/* Tendencies::computeThicknessTendenciesOnly, src/ocn/Tendencies.cpp:257-297
   functor ThicknessFluxDivOnCell, src/ocn/TendencyTerms.h:35-58 */
static int thick_tend(OracleStepper *S, const double *H, const double *U, double Time) {
   const OmegaMeshDesc *M = &S->M;
   const int K = S->K, ME = M->MaxEdges;
   double *Tend = S->Tend.LayerThicknessTend;
   memset(Tend, 0, sizeof(double) * M->NCellsSize * K);
   if (S->Cfg.ThicknessFluxTendencyEnable) {
      PFOR for (int c = 0; c < M->NCellsAll; ++c) {
         const double InvA = 1.0 / M->AreaCell[c];
         for (int k = 0; k < K; ++k) {
            double d = 0;
            for (int j = 0; j < M->NEdgesOnCell[c]; ++j) {
               int e = M->EdgesOnCell[c * ME + j];
               d -= M->DvEdge[e] * S->EdgeSignOnCell[c * ME + j] *
                    S->Aux.FluxLayerThickEdge[(size_t)e * K + k] * U[(size_t)e * K + k] * InvA;
            }
            Tend[(size_t)c * K + k] -= d;
         }
      }
   }
   if (S->Cfg.ManufacturedSolutionTendency) manufactured_thick(S, Time);
   /* a custom term installed by the host, Tendencies.cpp:288-293 */
   if (S->CustomThick && S->CustomThick(S->CustomUser, Tend, H, U, Time, NULL)) return 1;
   return 0;
}
int oracle_tend_thickness_only(OracleStepper *S, const double *U, double Time) { return thick_tend(S, NULL, U, Time); }

/* Tendencies::computeVelocityTendenciesOnly, src/ocn/Tendencies.cpp:301-425;
   functors src/ocn/TendencyTerms.h:81-334.  U = getNormalVelocity(VelTimeLevel)
   view, URaw = State->NormalVelocity[VelTimeLevel] (raw index, :322). */
static int vel_tend(OracleStepper *S, const double *H, const double *U, const double *URaw, double Time) {
   const OmegaMeshDesc *M = &S->M;
   const OmegaAuxArrays *A = &S->Aux;
   const OmegaTendConfig *G = &S->Cfg;
   const int K = S->K, ME2 = M->MaxEdges2;
   double *Tend = S->Tend.NormalVelocityTend;
   memset(Tend, 0, sizeof(double) * M->NEdgesSize * K);
   if (G->PVTendencyEnable) { /* PotentialVortHAdvOnEdge :81-108 */
      PFOR for (int e = 0; e < M->NEdgesAll; ++e)
         for (int k = 0; k < K; ++k) {
            double t = 0;
            for (int j = 0; j < M->NEdgesOnEdge[e]; ++j) {
               int je = M->EdgesOnEdge[e * ME2 + j];
               double NormVort = (A->NormRelVortEdge[(size_t)e * K + k] + A->NormPlanetVortEdge[(size_t)e * K + k] +
                                  A->NormRelVortEdge[(size_t)je * K + k] + A->NormPlanetVortEdge[(size_t)je * K + k]) * 0.5;
               t += M->WeightsOnEdge[e * ME2 + j] * A->FluxLayerThickEdge[(size_t)je * K + k] *
                    U[(size_t)je * K + k] * NormVort;
            }
            Tend[(size_t)e * K + k] += S->EdgeMask[e] * t;
         }
   }
   if (G->KETendencyEnable) { /* KEGradOnEdge :127-140 */
      PFOR for (int e = 0; e < M->NEdgesAll; ++e) {
         const int c0 = M->CellsOnEdge[2 * e], c1 = M->CellsOnEdge[2 * e + 1];
         const double InvDc = 1.0 / M->DcEdge[e];
         for (int k = 0; k < K; ++k)
            Tend[(size_t)e * K + k] -= S->EdgeMask[e] *
                (A->KineticEnergyCell[(size_t)c1 * K + k] - A->KineticEnergyCell[(size_t)c0 * K + k]) * InvDc;
      }
   }
   if (G->SSHTendencyEnable) { /* SSHGradOnEdge :159-173 */
      const double Grav = 9.80665;
      PFOR for (int e = 0; e < M->NEdgesAll; ++e) {
         const int c0 = M->CellsOnEdge[2 * e], c1 = M->CellsOnEdge[2 * e + 1];
         const double InvDc = 1.0 / M->DcEdge[e];
         for (int k = 0; k < K; ++k)
            Tend[(size_t)e * K + k] -= S->EdgeMask[e] * Grav *
                (A->SshCell[(size_t)c1 * K + k] - A->SshCell[(size_t)c0 * K + k]) * InvDc;
      }
   }
   if (G->VelDiffTendencyEnable) { /* VelocityDiffusionOnEdge :195-219 */
      PFOR for (int e = 0; e < M->NEdgesAll; ++e) {
         const int c0 = M->CellsOnEdge[2 * e], c1 = M->CellsOnEdge[2 * e + 1];
         const int v0 = M->VerticesOnEdge[2 * e], v1 = M->VerticesOnEdge[2 * e + 1];
         const double DcInv = 1.0 / M->DcEdge[e], DvInv = 1.0 / M->DvEdge[e];
         for (int k = 0; k < K; ++k) {
            double Del2U = ((A->VelocityDivCell[(size_t)c1 * K + k] - A->VelocityDivCell[(size_t)c0 * K + k]) * DcInv -
                            (A->RelVortVertex[(size_t)v1 * K + k] - A->RelVortVertex[(size_t)v0 * K + k]) * DvInv);
            Tend[(size_t)e * K + k] += S->EdgeMask[e] * G->ViscDel2 * S->MeshScalingDel2[e] * Del2U;
         }
      }
   }
   if (G->VelHyperDiffTendencyEnable) { /* VelocityHyperDiffOnEdge :244-269 */
      PFOR for (int e = 0; e < M->NEdgesAll; ++e) {
         const int c0 = M->CellsOnEdge[2 * e], c1 = M->CellsOnEdge[2 * e + 1];
         const int v0 = M->VerticesOnEdge[2 * e], v1 = M->VerticesOnEdge[2 * e + 1];
         const double DcInv = 1.0 / M->DcEdge[e], DvInv = 1.0 / M->DvEdge[e];
         for (int k = 0; k < K; ++k) {
            double Del2U = (G->DivFactor * (A->Del2DivCell[(size_t)c1 * K + k] - A->Del2DivCell[(size_t)c0 * K + k]) * DcInv -
                            (A->Del2RelVortVertex[(size_t)v1 * K + k] - A->Del2RelVortVertex[(size_t)v0 * K + k]) * DvInv);
            Tend[(size_t)e * K + k] -= S->EdgeMask[e] * G->ViscDel4 * S->MeshScalingDel4[e] * Del2U;
         }
      }
   }
   if (G->WindForcingTendencyEnable) { /* WindForcingOnEdge :291-301 */
      PFOR for (int e = 0; e < M->NEdgesAll; ++e) {
         double InvThick = 1.0 / A->MeanLayerThickEdge[(size_t)e * K];
         Tend[(size_t)e * K] += S->EdgeMask[e] * InvThick * A->NormalStressEdge[e] / G->Density0;
      }
   }
   if (G->BottomDragTendencyEnable) { /* BottomDragOnEdge :319-334 */
      const int KBot = K - 1;
      PFOR for (int e = 0; e < M->NEdgesAll; ++e) {
         const int c0 = M->CellsOnEdge[2 * e], c1 = M->CellsOnEdge[2 * e + 1];
         double VelNorm  = sqrt(A->KineticEnergyCell[(size_t)c0 * K + KBot] + A->KineticEnergyCell[(size_t)c1 * K + KBot]);
         double InvThick = 1.0 / A->MeanLayerThickEdge[(size_t)e * K + KBot];
         Tend[(size_t)e * K + KBot] -= S->EdgeMask[e] * G->BottomDragCoeff * VelNorm * InvThick *
                                       URaw[(size_t)e * K + KBot];
      }
   }
   if (G->ManufacturedSolutionTendency) manufactured_vel(S, Time);
   /* a custom term installed by the host, Tendencies.cpp:416-421 */
   if (S->CustomVel && S->CustomVel(S->CustomUser, Tend, H, U, Time, NULL)) return 1;
   return 0;
}
int oracle_tend_velocity_only(OracleStepper *S, const double *U, const double *URaw, double Time) {
   return vel_tend(S, NULL, U, URaw, Time);
}
int oracle_stepper_set_custom_tendencies(OracleStepper *S, OmegaCustomTendency Thickness, OmegaCustomTendency Velocity, void *User) {
   S->CustomThick = Thickness;
   S->CustomVel   = Velocity;
   S->CustomUser  = User;
   return 0;
}

/* Tendencies::computeTracerTendenciesOnly, src/ocn/Tendencies.cpp:427-486;
   functors src/ocn/TendencyTerms.h:349-480 */
int oracle_tend_tracer_only(OracleStepper *S, const double *URaw, const double *Tr) {
   const OmegaMeshDesc *M = &S->M;
   const OmegaAuxArrays *A = &S->Aux;
   const OmegaTendConfig *G = &S->Cfg;
   const int K = S->K, ME = M->MaxEdges;
   const size_t CS = (size_t)M->NCellsSize * K, ES = (size_t)M->NEdgesSize * K;
   memset(S->Tend.TracerTend, 0, sizeof(double) * S->NTracers * CS);
   for (int l = 0; l < S->NTracers; ++l) {
      double *Tend       = S->Tend.TracerTend + l * CS;
      const double *T    = Tr + l * CS;
      const double *HTr  = A->HTracersEdge + l * ES;
      const double *D2   = A->Del2TracersCell + l * CS;
      if (G->TracerHorzAdvTendencyEnable) { /* :349-373 */
         PFOR for (int c = 0; c < M->NCellsAll; ++c) {
            const double InvA = 1.0 / M->AreaCell[c];
            for (int k = 0; k < K; ++k) {
               double t = 0;
               for (int j = 0; j < M->NEdgesOnCell[c]; ++j) {
                  int e = M->EdgesOnCell[c * ME + j];
                  t -= S->EdgeMask[e] * M->DvEdge[e] * S->EdgeSignOnCell[c * ME + j] *
                       HTr[(size_t)e * K + k] * URaw[(size_t)e * K + k] * InvA;
               }
               Tend[(size_t)c * K + k] -= t;
            }
         }
      }
      if (G->TracerDiffTendencyEnable) { /* :394-426 */
         PFOR for (int c = 0; c < M->NCellsAll; ++c) {
            const double InvA = 1.0 / M->AreaCell[c];
            for (int k = 0; k < K; ++k) {
               double t = 0;
               for (int j = 0; j < M->NEdgesOnCell[c]; ++j) {
                  int e  = M->EdgesOnCell[c * ME + j];
                  int c0 = M->CellsOnEdge[2 * e], c1 = M->CellsOnEdge[2 * e + 1];
                  double RTemp = S->MeshScalingDel2[e] * M->DvEdge[e] / M->DcEdge[e];
                  double Grad  = T[(size_t)c1 * K + k] - T[(size_t)c0 * K + k];
                  t -= S->EdgeMask[e] * S->EdgeSignOnCell[c * ME + j] * RTemp *
                       A->MeanLayerThickEdge[(size_t)e * K + k] * Grad;
               }
               Tend[(size_t)c * K + k] += G->EddyDiff2 * t * InvA;
            }
         }
      }
      if (G->TracerHyperDiffTendencyEnable) { /* :449-480 */
         PFOR for (int c = 0; c < M->NCellsAll; ++c) {
            const double InvA = 1.0 / M->AreaCell[c];
            for (int k = 0; k < K; ++k) {
               double t = 0;
               for (int j = 0; j < M->NEdgesOnCell[c]; ++j) {
                  int e  = M->EdgesOnCell[c * ME + j];
                  int c0 = M->CellsOnEdge[2 * e], c1 = M->CellsOnEdge[2 * e + 1];
                  double RTemp = S->MeshScalingDel4[e] * M->DvEdge[e] / M->DcEdge[e];
                  double Grad  = D2[(size_t)c1 * K + k] - D2[(size_t)c0 * K + k];
                  t -= S->EdgeMask[e] * S->EdgeSignOnCell[c * ME + j] * RTemp * Grad;
               }
               Tend[(size_t)c * K + k] -= G->EddyDiff4 * t * InvA;
            }
         }
      }
   }
   return 0;
}

/* Tendencies::computeThicknessTendencies, src/ocn/Tendencies.cpp:488-519 */
int oracle_tend_thickness(OracleStepper *S, const double *H, const double *U, double Time) {
   aux_thick_edge(S, H, U);
   return thick_tend(S, H, U, Time);
}
/* Tendencies::computeVelocityTendencies, :521-535 */
int oracle_tend_velocity(OracleStepper *S, const double *H, const double *U, const double *URaw,
                         double Time) {
   oracle_aux_compute_mom(S, H, U);
   return vel_tend(S, H, U, URaw, Time);
}
/* Tendencies::computeTracerTendencies, :537-575 (raw-indexed H and U, :546-547) */
int oracle_tend_tracer(OracleStepper *S, const double *HRaw, const double *URaw, const double *Tr) {
   aux_tracer(S, HRaw, URaw, Tr);
   return oracle_tend_tracer_only(S, URaw, Tr);
}
/* Tendencies::computeAllTendencies, :579-600 */
int oracle_tend_all(OracleStepper *S, const double *H, const double *U, const double *URaw,
                    const double *Tr, double Time) {
   oracle_aux_compute_all(S, H, U, Tr);
   thick_tend(S, H, U, Time);
   vel_tend(S, H, U, URaw, Time);
   oracle_tend_tracer_only(S, URaw, Tr);
   return 0;
}

/* ------------------------------------------------------------------------ */
/* time stepping                                                             */
/* ------------------------------------------------------------------------ */

/* OceanState::getTimeIndex, src/ocn/OceanState.cpp:394-407 */
static int time_index(int Level, int Cur, int NTL) { return (Level + Cur + NTL) % NTL; }

/* A light view of one OceanState: ring of NTL buffers */
typedef struct StateView {
   double **H, **U;
   int NTL, Cur;
} StateView;

static double *get_h(const StateView *s, int Level) { return s->H[time_index(Level, s->Cur, s->NTL)]; }
static double *get_u(const StateView *s, int Level) { return s->U[time_index(Level, s->Cur, s->NTL)]; }
/* raw: State->LayerThickness[Level]; Compat=0 maps through the ring instead */
static double *raw_h(const StateView *s, int Level, int Compat) { return Compat ? s->H[Level] : get_h(s, Level); }
static double *raw_u(const StateView *s, int Level, int Compat) { return Compat ? s->U[Level] : get_u(s, Level); }

/* TimeStepper::updateThicknessByTend / updateVelocityByTend,
   src/timeStepping/TimeStepper.cpp:378-430 */
static void update_thick(OracleStepper *S, double *H1, const double *H2, double Coeff) {
   const int K = S->K;
   const double *T = S->Tend.LayerThicknessTend;
   PFOR for (int c = 0; c < S->M.NCellsAll; ++c)
      for (int k = 0; k < K; ++k)
         H1[(size_t)c * K + k] = H2[(size_t)c * K + k] + Coeff * T[(size_t)c * K + k];
}
static void update_vel(OracleStepper *S, double *U1, const double *U2, double Coeff) {
   const int K = S->K;
   const double *T = S->Tend.NormalVelocityTend;
   PFOR for (int e = 0; e < S->M.NEdgesAll; ++e)
      for (int k = 0; k < K; ++k)
         U1[(size_t)e * K + k] = U2[(size_t)e * K + k] + Coeff * T[(size_t)e * K + k];
}
/* TimeStepper::updateTracersByTend, :446-469 */
static void update_tracers(OracleStepper *S, double *Next, const double *Cur, const double *H1,
                           const double *H2, double Coeff) {
   const int K = S->K;
   const size_t CS = (size_t)S->M.NCellsSize * K;
   for (int l = 0; l < S->NTracers; ++l) {
      PFOR for (int c = 0; c < S->M.NCellsAll; ++c)
         for (int k = 0; k < K; ++k) {
            size_t i = (size_t)c * K + k;
            Next[l * CS + i] = (Cur[l * CS + i] * H2[i] + Coeff * S->Tend.TracerTend[l * CS + i]) / H1[i];
         }
   }
}
/* weightTracers :473-487, accumulateTracersUpdate :492-507, finalizeTracersUpdate :511-524 */
static void weight_tracers(OracleStepper *S, double *Next, const double *Cur, const double *H) {
   const int K = S->K;
   const size_t CS = (size_t)S->M.NCellsSize * K;
   for (int l = 0; l < S->NTracers; ++l) {
      PFOR for (int c = 0; c < S->M.NCellsAll; ++c)
         for (int k = 0; k < K; ++k)
            Next[l * CS + (size_t)c * K + k] = Cur[l * CS + (size_t)c * K + k] * H[(size_t)c * K + k];
   }
}
static void accumulate_tracers(OracleStepper *S, double *Accum, double Coeff) {
   const int K = S->K;
   const size_t CS = (size_t)S->M.NCellsSize * K;
   for (int l = 0; l < S->NTracers; ++l) {
      PFOR for (int c = 0; c < S->M.NCellsAll; ++c)
         for (int k = 0; k < K; ++k)
            Accum[l * CS + (size_t)c * K + k] += Coeff * S->Tend.TracerTend[l * CS + (size_t)c * K + k];
   }
}
static void finalize_tracers(OracleStepper *S, double *Next, const double *H) {
   const int K = S->K;
   const size_t CS = (size_t)S->M.NCellsSize * K;
   for (int l = 0; l < S->NTracers; ++l) {
      PFOR for (int c = 0; c < S->M.NCellsAll; ++c)
         for (int k = 0; k < K; ++k)
            Next[l * CS + (size_t)c * K + k] /= H[(size_t)c * K + k];
   }
}

/* computeAllTendencies(State, Aux, Tracers, ThickLevel, VelLevel, Time) with the
   reference's mix of ring-mapped and raw views */
static void all_tend(OracleStepper *S, const StateView *St, const double *Tr, int TL, int VL,
                     double Time, int Compat) {
   oracle_tend_all(S, get_h(St, TL), get_u(St, VL), raw_u(St, VL, Compat), Tr, Time);
}

/* Exchange points: after phase 0 the arrays in oracle_stepper_exchange_list
   must be halo-exchanged by the caller (no-op on one rank of a closed mesh). */
static void rk4_phase(OracleStepper *S, OmegaStateArrays *X, const OmegaStepCoeffs *C, int Phase,
                      int Compat) {
   /* RungeKutta4Stepper::doStep, src/timeStepping/RungeKutta4Stepper.cpp:68-137 */
   StateView St   = {X->LayerThickness, X->NormalVelocity, X->NTimeLevels, X->CurTimeIndex};
   double *PH[1]  = {S->ProvisH}, *PU[1] = {S->ProvisU};
   StateView Pv   = {PH, PU, 1, 0};
   const int CurLevel = 0, NextLevel = 1;
   double *CurTr  = X->Tracers[time_index(CurLevel, X->TracerCurTimeIndex, X->NTimeLevels)];
   double *NextTr = X->Tracers[time_index(NextLevel, X->TracerCurTimeIndex, X->NTimeLevels)];
   const int s0 = Phase == 0 ? 0 : 2, s1 = Phase == 0 ? 2 : 3;
   for (int Stage = s0; Stage <= s1; ++Stage) {
      if (Stage == 0) {
         weight_tracers(S, NextTr, CurTr, raw_h(&St, CurLevel, Compat));
         all_tend(S, &St, CurTr, CurLevel, CurLevel, C->StageTime[0], Compat);
         update_thick(S, get_h(&St, NextLevel), get_h(&St, CurLevel), C->B[0]);
         update_vel(S, get_u(&St, NextLevel), get_u(&St, CurLevel), C->B[0]);
         accumulate_tracers(S, NextTr, C->B[0]);
      } else {
         if (!(Phase == 1 && Stage == 2)) {
            update_thick(S, get_h(&Pv, CurLevel), get_h(&St, CurLevel), C->A[Stage]);
            update_vel(S, get_u(&Pv, CurLevel), get_u(&St, CurLevel), C->A[Stage]);
            update_tracers(S, S->ProvisTr, CurTr, raw_h(&Pv, CurLevel, Compat),
                           raw_h(&St, CurLevel, Compat), C->A[Stage]);
         }
         if (Phase == 0 && Stage == 2) return; /* exchange point 0: provis h,u,tracers */
         all_tend(S, &Pv, S->ProvisTr, CurLevel, CurLevel, C->StageTime[Stage], Compat);
         update_thick(S, get_h(&St, NextLevel), get_h(&St, NextLevel), C->B[Stage]);
         update_vel(S, get_u(&St, NextLevel), get_u(&St, NextLevel), C->B[Stage]);
         accumulate_tracers(S, NextTr, C->B[Stage]);
      }
   }
   finalize_tracers(S, NextTr, raw_h(&St, NextLevel, Compat));
   /* exchange point 1: State level 1 h,u and tracers level 1, then rotate */
}

static void rk2_phase(OracleStepper *S, OmegaStateArrays *X, const OmegaStepCoeffs *C, int Compat) {
   /* RungeKutta2Stepper::doStep, src/timeStepping/RungeKutta2Stepper.cpp:27-73 */
   StateView St = {X->LayerThickness, X->NormalVelocity, X->NTimeLevels, X->CurTimeIndex};
   const int CurLevel = 0, NextLevel = 1;
   double *CurTr  = X->Tracers[time_index(CurLevel, X->TracerCurTimeIndex, X->NTimeLevels)];
   double *NextTr = X->Tracers[time_index(NextLevel, X->TracerCurTimeIndex, X->NTimeLevels)];
   all_tend(S, &St, CurTr, CurLevel, CurLevel, C->StageTime[0], Compat);
   update_thick(S, get_h(&St, NextLevel), get_h(&St, CurLevel), C->A[1]);
   update_vel(S, get_u(&St, NextLevel), get_u(&St, CurLevel), C->A[1]);
   update_tracers(S, NextTr, CurTr, raw_h(&St, NextLevel, Compat), raw_h(&St, CurLevel, Compat), C->A[1]);
   all_tend(S, &St, NextTr, NextLevel, NextLevel, C->StageTime[1], Compat);
   update_thick(S, get_h(&St, NextLevel), get_h(&St, CurLevel), C->B[1]);
   update_vel(S, get_u(&St, NextLevel), get_u(&St, CurLevel), C->B[1]);
   update_tracers(S, NextTr, CurTr, raw_h(&St, NextLevel, Compat), raw_h(&St, CurLevel, Compat), C->B[1]);
}

static void fb_phase(OracleStepper *S, OmegaStateArrays *X, const OmegaStepCoeffs *C, int Compat) {
   /* ForwardBackwardStepper::doStep, src/timeStepping/ForwardBackwardStepper.cpp:27-82 */
   StateView St = {X->LayerThickness, X->NormalVelocity, X->NTimeLevels, X->CurTimeIndex};
   const int CurLevel = 0, NextLevel = 1;
   double *CurTr  = X->Tracers[time_index(CurLevel, X->TracerCurTimeIndex, X->NTimeLevels)];
   double *NextTr = X->Tracers[time_index(NextLevel, X->TracerCurTimeIndex, X->NTimeLevels)];
   oracle_tend_thickness(S, get_h(&St, CurLevel), get_u(&St, CurLevel), C->StageTime[0]);
   update_thick(S, get_h(&St, NextLevel), get_h(&St, CurLevel), C->Dt);
   oracle_tend_tracer(S, raw_h(&St, CurLevel, Compat), raw_u(&St, CurLevel, Compat), CurTr);
   update_tracers(S, NextTr, CurTr, raw_h(&St, NextLevel, Compat), raw_h(&St, CurLevel, Compat), C->Dt);
   oracle_tend_velocity(S, get_h(&St, NextLevel), get_u(&St, CurLevel), raw_u(&St, CurLevel, Compat),
                        C->StageTime[1]);
   update_vel(S, get_u(&St, NextLevel), get_u(&St, CurLevel), C->Dt);
}

/* number of phases (= exchange points) per step */
static int n_phases(OmegaStepperType Type) { return Type == OMEGA_STEPPER_RK4 ? 2 : 1; }

int oracle_stepper_step_phase(OracleStepper *S, OmegaStepperType Type, OmegaStateArrays *X,
                              const OmegaStepCoeffs *C, int Phase, int Fused, int Compat) {
   (void)Fused;
   if (Type == OMEGA_STEPPER_RK4) rk4_phase(S, X, C, Phase, Compat);
   else if (Type == OMEGA_STEPPER_RK2) rk2_phase(S, X, C, Compat);
   else fb_phase(S, X, C, Compat);
   return 0;
}

/* OceanState::updateTimeLevels + Tracers::updateTimeLevels minus the halo
   exchange: src/ocn/OceanState.cpp:348-365, src/ocn/Tracers.cpp:473-496 */
int oracle_state_rotate(OmegaStateArrays *X) {
   X->CurTimeIndex       = (X->CurTimeIndex + 1) % X->NTimeLevels;
   X->TracerCurTimeIndex = (X->TracerCurTimeIndex + 1) % X->NTimeLevels;
   return 0;
}

/* Arrays to exchange after Phase (RungeKutta4Stepper.cpp:108-114; updateTimeLevels) */
int oracle_stepper_exchange_list(OracleStepper *S, OmegaStepperType Type, const OmegaStateArrays *X,
                                 int Phase, double **Ptrs, int *Elems, int *NOuter, int *NArrays) {
   if (Type == OMEGA_STEPPER_RK4 && Phase == 0) {
      Ptrs[0] = S->ProvisH;  Elems[0] = OMEGA_ON_CELL; NOuter[0] = 1;
      Ptrs[1] = S->ProvisU;  Elems[1] = OMEGA_ON_EDGE; NOuter[1] = 1;
      Ptrs[2] = S->ProvisTr; Elems[2] = OMEGA_ON_CELL; NOuter[2] = S->NTracers;
   } else {
      Ptrs[0] = X->LayerThickness[time_index(1, X->CurTimeIndex, X->NTimeLevels)];
      Ptrs[1] = X->NormalVelocity[time_index(1, X->CurTimeIndex, X->NTimeLevels)];
      Ptrs[2] = X->Tracers[time_index(1, X->TracerCurTimeIndex, X->NTimeLevels)];
      Elems[0] = OMEGA_ON_CELL; Elems[1] = OMEGA_ON_EDGE; Elems[2] = OMEGA_ON_CELL;
      NOuter[0] = 1; NOuter[1] = 1; NOuter[2] = S->NTracers;
   }
   *NArrays = 3;
   return 0;
}

/* one rank, closed mesh: exchanges are identities */
int oracle_stepper_do_step(OracleStepper *S, OmegaStepperType Type, OmegaStateArrays *X,
                           const OmegaStepCoeffs *C, int Fused, int Compat) {
   for (int p = 0; p < n_phases(Type); ++p)
      oracle_stepper_step_phase(S, Type, X, C, p, Fused, Compat);
   return oracle_state_rotate(X);
}

/* ------------------------------------------------------------------------ */
/* generic operators used only by the known-answer tests                    */
/* src/ocn/HorzOperators.h:9-135                                             */
/* ------------------------------------------------------------------------ */
int oracle_op_divergence(OracleStepper *S, const double *VecEdge, double *DivCell) {
   const OmegaMeshDesc *M = &S->M;
   const int K = S->K, ME = M->MaxEdges;
   PFOR for (int c = 0; c < M->NCellsAll; ++c) {
      const double InvA = 1.0 / M->AreaCell[c];
      for (int k = 0; k < K; ++k) {
         double t = 0;
         for (int j = 0; j < M->NEdgesOnCell[c]; ++j) {
            int e = M->EdgesOnCell[c * ME + j];
            t -= M->DvEdge[e] * S->EdgeSignOnCell[c * ME + j] * VecEdge[(size_t)e * K + k] * InvA;
         }
         DivCell[(size_t)c * K + k] = t;
      }
   }
   return 0;
}
int oracle_op_gradient(OracleStepper *S, const double *ScalarCell, double *GradEdge) {
   const OmegaMeshDesc *M = &S->M;
   const int K = S->K;
   PFOR for (int e = 0; e < M->NEdgesAll; ++e) {
      const double InvDc = 1.0 / M->DcEdge[e];
      const int c0 = M->CellsOnEdge[2 * e], c1 = M->CellsOnEdge[2 * e + 1];
      for (int k = 0; k < K; ++k)
         GradEdge[(size_t)e * K + k] = InvDc * (ScalarCell[(size_t)c1 * K + k] - ScalarCell[(size_t)c0 * K + k]);
   }
   return 0;
}
int oracle_op_curl(OracleStepper *S, const double *VecEdge, double *CurlVertex) {
   const OmegaMeshDesc *M = &S->M;
   const int K = S->K, VD = M->VertexDegree;
   PFOR for (int v = 0; v < M->NVerticesAll; ++v) {
      const double InvAt = 1.0 / M->AreaTriangle[v];
      for (int k = 0; k < K; ++k) {
         double t = 0;
         for (int j = 0; j < VD; ++j) {
            int e = M->EdgesOnVertex[v * VD + j];
            t += M->DcEdge[e] * S->EdgeSignOnVertex[v * VD + j] * VecEdge[(size_t)e * K + k] * InvAt;
         }
         CurlVertex[(size_t)v * K + k] = t;
      }
   }
   return 0;
}
int oracle_op_tangential_recon(OracleStepper *S, const double *VecEdge, double *ReconEdge) {
   const OmegaMeshDesc *M = &S->M;
   const int K = S->K, ME2 = M->MaxEdges2;
   PFOR for (int e = 0; e < M->NEdgesAll; ++e)
      for (int k = 0; k < K; ++k) {
         double t = 0;
         for (int j = 0; j < M->NEdgesOnEdge[e]; ++j) {
            int je = M->EdgesOnEdge[e * ME2 + j];
            t += M->WeightsOnEdge[e * ME2 + j] * VecEdge[(size_t)je * K + k];
         }
         ReconEdge[(size_t)e * K + k] = t;
      }
   return 0;
}
int oracle_op_interp_cell_to_edge(OracleStepper *S, const double *Cell, int Aniso, double *Edge) {
   for (int e = 0; e < S->M.NEdgesAll; ++e) Edge[e] = interp_cell_to_edge(S, e, Cell, Aniso);
   return 0;
}
