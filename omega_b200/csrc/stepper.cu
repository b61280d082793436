// stepper.cu -- Tendencies + AuxiliaryState + TimeStepper behind the C ABI.
// Sequencing restated from components/omega/src/ocn/AuxiliaryState.cpp:60-185,
// src/ocn/Tendencies.cpp:257-600, src/timeStepping/TimeStepper.cpp:378-524,
// RungeKutta4Stepper.cpp:68-137, RungeKutta2Stepper.cpp:27-73,
// ForwardBackwardStepper.cpp:27-82.
#include <cmath>
#include <cstring>

#include "omega_dev.h"
#include "stepper.h"

using namespace omega;

namespace omega {

int stepper_alloc(OmegaStepper *S, Real **p, size_t n) {
   void *d = nullptr;
   const size_t bytes = (n ? n : 2) * sizeof(Real);
   OMEGA_CUDA_CHECK(cudaMalloc(&d, bytes));
   // the memset runs on the legacy default stream, which callers' non-blocking streams do not
   // synchronise with: wait for it here so no later kernel can race with it (allocations are rare)
   OMEGA_CUDA_CHECK(cudaMemset(d, 0, bytes));
   OMEGA_CUDA_CHECK(cudaStreamSynchronize(0));
   S->Allocs.push_back(d);
   *p = (Real *)d;
   return 0;
}

ManufacturedParams manufactured(const OmegaStepper *S, double Time) {
   // constants of ManufacturedSolution::init, src/ocn/CustomTendencyTerms.cpp:83-102
   ManufacturedParams P;
   std::memset(&P, 0, sizeof(P));
   P.Enabled = S->Cfg.ManufacturedSolutionTendency;
   if (!P.Enabled) return P;
   const Real Grav = 9.80665, Pii = 3.141592653589793;
   P.H0      = S->BottomDepth0;
   P.Eta0    = S->Cfg.ManufacturedAmplitude;
   P.Kx      = 2.0 * Pii / S->Cfg.ManufacturedWavelengthX;
   P.Ky      = 2.0 * Pii / S->Cfg.ManufacturedWavelengthY;
   P.AngFreq = std::sqrt(P.H0 * Grav * (P.Kx * P.Kx + P.Ky * P.Ky));
   P.Time    = Time;
   return P;
}

// HTracersEdge is the largest array of the model (3*T doubles per cell-layer)
// and only the kernel-per-term path needs it: allocate on first use.
int ensure_htr(OmegaStepper *S) {
   if (S->A.HTr || S->NTracers == 0) return 0;
   return stepper_alloc(S, &S->A.HTr, (size_t)S->NTracers * S->M->D.NEdgesSize * S->M->D.K);
}

} // namespace omega

extern "C" {

int omega_stepper_create(const OmegaMesh *Mesh, const OmegaTendConfig *Cfg, int NTracers,
                         OmegaStepper **Out) {
   OMEGA_REQUIRE(Mesh && Cfg && Out && NTracers >= 0, "omega_stepper_create: bad argument");
   OmegaStepper *S = new OmegaStepper();
   S->M = Mesh;
   S->Cfg = *Cfg;
   S->NTracers = NTracers;
   S->Launches = 0;
   S->ProvisExchangeSlot = 0;
   S->LastPath = 0;
   S->Halo = nullptr;
   S->SplitRecorded = 0;
   S->Profiling = 0;
   std::memset(&S->A, 0, sizeof(S->A));
   std::memset(&S->T, 0, sizeof(S->T));
   std::memset(&S->P, 0, sizeof(S->P));
   const DMesh &D = Mesh->D;
   const size_t K = D.K, C = (size_t)D.NCellsSize * K, E = (size_t)D.NEdgesSize * K,
                V = (size_t)D.NVerticesSize * K, T = NTracers;
   int rc = 0;
   AuxPtrs &A = S->A;
   rc |= stepper_alloc(S, &A.KE, C);       rc |= stepper_alloc(S, &A.Div, C);
   rc |= stepper_alloc(S, &A.Flux, E);     rc |= stepper_alloc(S, &A.Mean, E);
   rc |= stepper_alloc(S, &A.Ssh, C);      rc |= stepper_alloc(S, &A.RelVort, V);
   rc |= stepper_alloc(S, &A.NormRelVort, V);     rc |= stepper_alloc(S, &A.NormPlanetVort, V);
   rc |= stepper_alloc(S, &A.NormRelVortEdge, E); rc |= stepper_alloc(S, &A.NormPlanetVortEdge, E);
   rc |= stepper_alloc(S, &A.Del2Edge, E); rc |= stepper_alloc(S, &A.Del2Div, C);
   rc |= stepper_alloc(S, &A.Del2RelVort, V);
   rc |= stepper_alloc(S, &A.Del2Tr, T * C);
   rc |= stepper_alloc(S, &A.NormalStress, D.NEdgesSize);
   rc |= stepper_alloc(S, &A.ZonalStress, D.NCellsSize);
   rc |= stepper_alloc(S, &A.MeridStress, D.NCellsSize);
   rc |= stepper_alloc(S, &A.ZonalStress2, D.NCellsSize);
   rc |= stepper_alloc(S, &S->SrcH, D.NCellsSize);
   rc |= stepper_alloc(S, &S->SrcU, D.NEdgesSize);
   rc |= stepper_alloc(S, &S->RTop, D.NEdgesSize);
   rc |= stepper_alloc(S, &S->RBot, D.NEdgesSize);
   rc |= stepper_alloc(S, &S->T.LayerThicknessTend, C);
   rc |= stepper_alloc(S, &S->T.NormalVelocityTend, E);
   rc |= stepper_alloc(S, &S->T.TracerTend, T * C);
   // provisional state (RungeKutta4Stepper::finalizeInit, RungeKutta4Stepper.cpp:43-64);
   // two sets so the fused pipeline can ping-pong
   for (int i = 0; i < 2; ++i) {
      rc |= stepper_alloc(S, &S->P.H[i], C);
      rc |= stepper_alloc(S, &S->P.U[i], E);
      rc |= stepper_alloc(S, &S->P.Tr[i], T * C);
   }
   if (rc) {
      omega_stepper_destroy(S);
      return 1;
   }
   // BottomDepthH(0) of the HOST-order mesh (CustomTendencyTerms.cpp:80)
   Real b0 = 0;
   const int dev_row_of_host0 = Mesh->IPerm[0][0];
   OMEGA_CUDA_CHECK(cudaMemcpy(&b0, D.BottomDepth + dev_row_of_host0, sizeof(Real), cudaMemcpyDeviceToHost));
   S->BottomDepth0 = b0;
   *Out = S;
   return 0;
}

int omega_stepper_destroy(OmegaStepper *S) {
   if (!S) return 0;
   for (void *p : S->Allocs) cudaFree(p);
   delete S;
   return 0;
}

int omega_stepper_set_config(OmegaStepper *S, const OmegaTendConfig *Cfg) {
   OMEGA_REQUIRE(S && Cfg, "omega_stepper_set_config: null argument");
   S->Cfg = *Cfg;
   return 0;
}

int omega_stepper_aux_arrays(OmegaStepper *S, OmegaAuxArrays *Aux) {
   OMEGA_REQUIRE(S && Aux, "omega_stepper_aux_arrays: null argument");
   if (ensure_htr(S)) return 1;
   const AuxPtrs &A = S->A;
   Aux->KineticEnergyCell = A.KE;           Aux->VelocityDivCell = A.Div;
   Aux->FluxLayerThickEdge = A.Flux;        Aux->MeanLayerThickEdge = A.Mean;
   Aux->SshCell = A.Ssh;                    Aux->RelVortVertex = A.RelVort;
   Aux->NormRelVortVertex = A.NormRelVort;  Aux->NormPlanetVortVertex = A.NormPlanetVort;
   Aux->NormRelVortEdge = A.NormRelVortEdge; Aux->NormPlanetVortEdge = A.NormPlanetVortEdge;
   Aux->Del2Edge = A.Del2Edge;              Aux->Del2DivCell = A.Del2Div;
   Aux->Del2RelVortVertex = A.Del2RelVort;  Aux->HTracersEdge = A.HTr;
   Aux->Del2TracersCell = A.Del2Tr;         Aux->NormalStressEdge = A.NormalStress;
   Aux->ZonalStressCell = A.ZonalStress;    Aux->MeridStressCell = A.MeridStress;
   return 0;
}

int omega_stepper_tend_arrays(OmegaStepper *S, OmegaTendArrays *Tend) {
   OMEGA_REQUIRE(S && Tend, "omega_stepper_tend_arrays: null argument");
   *Tend = S->T;
   return 0;
}

int omega_stepper_provis_arrays(OmegaStepper *S, Real **H, Real **U, Real **Tr) {
   OMEGA_REQUIRE(S, "omega_stepper_provis_arrays: null argument");
   *H = S->P.H[0]; *U = S->P.U[0]; *Tr = S->P.Tr[0];
   return 0;
}

int omega_stepper_set_wind_stress(OmegaStepper *S, const Real *Zonal, const Real *Merid, void *Stream) {
   OMEGA_REQUIRE(S && Zonal && Merid, "omega_stepper_set_wind_stress: null argument");
   int rc = omega_field_to_device(S->M, OMEGA_ON_CELL, 1, 1, Zonal, S->A.ZonalStress, Stream);
   rc |= omega_field_to_device(S->M, OMEGA_ON_CELL, 1, 1, Merid, S->A.MeridStress, Stream);
   return rc;
}

long long omega_stepper_launch_count(const OmegaStepper *S) { return S ? S->Launches : 0; }
int omega_stepper_last_path(const OmegaStepper *S) { return S ? S->LastPath : 0; }

int omega_stepper_attach_halo(OmegaStepper *S, OmegaHalo *Halo) {
   OMEGA_REQUIRE(S, "omega_stepper_attach_halo: null stepper");
   S->Halo = Halo;
   return 0;
}

int omega_stepper_set_profiling(OmegaStepper *S, int Enable) {
   OMEGA_REQUIRE(S, "omega_stepper_set_profiling: null argument");
   S->Profiling = Enable;
   return 0;
}

// Sums (ms) and launch counts per fused kernel kind since the last call;
// synchronises the device. Kinds: 0 fA, 1 fB, 2 fT2, 3 fC, 4 fE.
int omega_stepper_get_profile(OmegaStepper *S, double *Ms, int *Counts, int NKinds) {
   OMEGA_REQUIRE(S && Ms && Counts && NKinds >= kNumFused, "omega_stepper_get_profile: bad argument");
   OMEGA_CUDA_CHECK(cudaDeviceSynchronize());
   for (int i = 0; i < NKinds; ++i) { Ms[i] = 0; Counts[i] = 0; }
   for (auto &t : S->Events) {
      float ms = 0;
      OMEGA_CUDA_CHECK(cudaEventElapsedTime(&ms, t.Start, t.Stop));
      Ms[t.Kind] += ms;
      Counts[t.Kind] += 1;
      cudaEventDestroy(t.Start);
      cudaEventDestroy(t.Stop);
   }
   S->Events.clear();
   return 0;
}

// ---- AuxiliaryState ---------------------------------------------------------
int omega_aux_compute_mom(OmegaStepper *S, const Real *H, const Real *U, void *Stream) {
   OMEGA_REQUIRE(S && H && U, "omega_aux_compute_mom: null argument");
   const DMesh &D = S->M->D;
   Launch L{(cudaStream_t)Stream, &S->Launches};
   // order of AuxiliaryState.cpp:79-139; cellAuxState3 (Ssh) rides in cellAuxState1
   launch_vertex_aux1(L, D, H, U, S->A);
   launch_cell_aux1(L, D, H, U, S->A);
   launch_wind_aux(L, D, S->A, S->Cfg.WindStressInterpType);
   launch_edge_aux2(L, D, H, U, S->A, S->Cfg.FluxThicknessType, 0);
   launch_vertex_aux2(L, D, S->A);
   launch_cell_aux2(L, D, S->A);
   OMEGA_CUDA_CHECK(cudaGetLastError());
   return 0;
}

static int tracer_aux(OmegaStepper *S, const Real *H, const Real *U, const Real *Tr, cudaStream_t St) {
   if (S->NTracers == 0) return 0;
   if (ensure_htr(S)) return 1;
   Launch L{St, &S->Launches};
   launch_tracer_aux_edge(L, S->M->D, H, U, Tr, S->NTracers, S->A, S->Cfg.FluxTracerType);
   launch_tracer_aux_cell(L, S->M->D, Tr, S->NTracers, S->A);
   OMEGA_CUDA_CHECK(cudaGetLastError());
   return 0;
}

int omega_aux_compute_all(OmegaStepper *S, const Real *H, const Real *U, const Real *Tr, void *Stream) {
   OMEGA_REQUIRE(S && H && U && (Tr || S->NTracers == 0), "omega_aux_compute_all: null argument");
   int rc = omega_aux_compute_mom(S, H, U, Stream);
   if (rc) return rc;
   return tracer_aux(S, H, U, Tr, (cudaStream_t)Stream);
}

// ---- Tendencies ---------------------------------------------------------------
// built-in terms, then the host's custom term if one is installed (Tendencies.cpp:288-293, 416-421); H is the
// thickness of the evaluation when the caller has one (the ...Only entry points of the C ABI do not)
static int thick_tend(OmegaStepper *S, const Real *H, const Real *U, double Time, void *Stream) {
   Launch L{(cudaStream_t)Stream, &S->Launches};
   launch_thick_tend(L, S->M->D, S->Cfg, S->A, U, S->T.LayerThicknessTend, manufactured(S, Time));
   OMEGA_CUDA_CHECK(cudaGetLastError());
   if (S->CustomThick) {
      const int rc = S->CustomThick(S->CustomUser, S->T.LayerThicknessTend, H, U, Time, Stream);
      OMEGA_REQUIRE(rc == 0, "custom thickness tendency returned an error");
   }
   return 0;
}
static int vel_tend(OmegaStepper *S, const Real *H, const Real *U, const Real *URaw, double Time, void *Stream) {
   Launch L{(cudaStream_t)Stream, &S->Launches};
   launch_vel_tend(L, S->M->D, S->Cfg, S->A, U, URaw, S->T.NormalVelocityTend, manufactured(S, Time));
   OMEGA_CUDA_CHECK(cudaGetLastError());
   if (S->CustomVel) {
      const int rc = S->CustomVel(S->CustomUser, S->T.NormalVelocityTend, H, U, Time, Stream);
      OMEGA_REQUIRE(rc == 0, "custom velocity tendency returned an error");
   }
   return 0;
}

int omega_stepper_set_custom_tendencies(OmegaStepper *S, OmegaCustomTendency Thickness, OmegaCustomTendency Velocity, void *User) {
   OMEGA_REQUIRE(S, "omega_stepper_set_custom_tendencies: null stepper");
   S->CustomThick = Thickness;
   S->CustomVel   = Velocity;
   S->CustomUser  = User;
   return 0;
}

int omega_tend_thickness_only(OmegaStepper *S, const Real *U, double Time, void *Stream) {
   OMEGA_REQUIRE(S && U, "omega_tend_thickness_only: null argument");
   return thick_tend(S, nullptr, U, Time, Stream);
}

int omega_tend_velocity_only(OmegaStepper *S, const Real *U, const Real *URaw, double Time, void *Stream) {
   OMEGA_REQUIRE(S && U && URaw, "omega_tend_velocity_only: null argument");
   return vel_tend(S, nullptr, U, URaw, Time, Stream);
}

int omega_tend_tracer_only(OmegaStepper *S, const Real *URaw, const Real *Tr, void *Stream) {
   OMEGA_REQUIRE(S && URaw && (Tr || S->NTracers == 0), "omega_tend_tracer_only: null argument");
   if (S->NTracers == 0) return 0;
   if (ensure_htr(S)) return 1;
   Launch L{(cudaStream_t)Stream, &S->Launches};
   launch_tracer_tend(L, S->M->D, S->Cfg, S->A, URaw, Tr, S->NTracers, S->T.TracerTend);
   OMEGA_CUDA_CHECK(cudaGetLastError());
   return 0;
}

int omega_tend_thickness(OmegaStepper *S, const Real *H, const Real *U, double Time, void *Stream) {
   OMEGA_REQUIRE(S && H && U, "omega_tend_thickness: null argument");
   Launch L{(cudaStream_t)Stream, &S->Launches};
   launch_edge_aux2(L, S->M->D, H, U, S->A, S->Cfg.FluxThicknessType, 1);
   return thick_tend(S, H, U, Time, Stream);
}

int omega_tend_velocity(OmegaStepper *S, const Real *H, const Real *U, const Real *URaw, double Time,
                        void *Stream) {
   OMEGA_REQUIRE(URaw, "omega_tend_velocity: null argument");
   int rc = omega_aux_compute_mom(S, H, U, Stream);
   if (rc) return rc;
   return vel_tend(S, H, U, URaw, Time, Stream);
}

int omega_tend_tracer(OmegaStepper *S, const Real *HRaw, const Real *URaw, const Real *Tr, void *Stream) {
   OMEGA_REQUIRE(S && HRaw && URaw, "omega_tend_tracer: null argument");
   int rc = tracer_aux(S, HRaw, URaw, Tr, (cudaStream_t)Stream);
   if (rc) return rc;
   return omega_tend_tracer_only(S, URaw, Tr, Stream);
}

int omega_tend_all(OmegaStepper *S, const Real *H, const Real *U, const Real *URaw, const Real *Tr,
                   double Time, void *Stream) {
   int rc = omega_aux_compute_all(S, H, U, Tr, Stream);
   if (rc) return rc;
   OMEGA_REQUIRE(URaw, "omega_tend_all: null argument");
   rc = thick_tend(S, H, U, Time, Stream);
   if (rc) return rc;
   rc = vel_tend(S, H, U, URaw, Time, Stream);
   if (rc) return rc;
   return omega_tend_tracer_only(S, URaw, Tr, Stream);
}

} // extern "C"
