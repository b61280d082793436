// step.cu -- TimeStepper::doStep for RK4 / RK2 / forward-backward.
// Kernel-per-term path: the reference's stage sequence on the kernels of
// kernels_{aux,tend,update}.cu, every public array filled.  Fused path: fused.cu.
// Reference: components/omega/src/timeStepping/RungeKutta4Stepper.cpp:68-137,
// RungeKutta2Stepper.cpp:27-73, ForwardBackwardStepper.cpp:27-82,
// TimeStepper.cpp:378-524; time-level ring: src/ocn/OceanState.cpp:394-407.
#include "halo.h"
#include "omega_dev.h"
#include "stepper.h"

using namespace omega;

namespace {

inline int time_index(int Level, int Cur, int NTL) { return (Level + Cur + NTL) % NTL; }

// One OceanState as the steppers see it: ring-mapped getters and the raw
// vector indexing several reference call sites use.
struct StateView {
   Real *const *H, *const *U;
   int NTL, Cur;
   Real *h(int Level) const { return H[time_index(Level, Cur, NTL)]; }
   Real *u(int Level) const { return U[time_index(Level, Cur, NTL)]; }
   Real *hraw(int Level, int Compat) const { return Compat ? H[Level] : h(Level); }
   Real *uraw(int Level, int Compat) const { return Compat ? U[Level] : u(Level); }
};

struct Ctx {
   OmegaStepper *S;
   cudaStream_t St;
   int Compat;
   const DMesh &D() const { return S->M->D; }
   Launch L() const { return Launch{St, &S->Launches}; }
   void update_thick(Real *H1, const Real *H2, Real C) const {
      launch_axpy(L(), D().NCellsAll, D().K, H1, H2, C, S->T.LayerThicknessTend);
   }
   void update_vel(Real *U1, const Real *U2, Real C) const {
      launch_axpy(L(), D().NEdgesAll, D().K, U1, U2, C, S->T.NormalVelocityTend);
   }
   void update_tracers(Real *Next, const Real *Cur, const Real *H1, const Real *H2, Real C) const {
      launch_update_tracers(L(), D().NCellsAll, D().NCellsSize, D().K, S->NTracers, Next, Cur, H1, H2, C, S->T.TracerTend);
   }
   int all_tend(const StateView &V, const Real *Tr, int TL, int VL, Real Time) const {
      return omega_tend_all(S, V.h(TL), V.u(VL), V.uraw(VL, Compat), Tr, Time, St);
   }
};

int rk4_phase(const Ctx &X, OmegaStateArrays *A, const OmegaStepCoeffs *C, int Phase) {
   OmegaStepper *S = X.S;
   const DMesh &D  = X.D();
   StateView St{A->LayerThickness, A->NormalVelocity, A->NTimeLevels, A->CurTimeIndex};
   Real *PH[1] = {S->P.H[0]}, *PU[1] = {S->P.U[0]};
   StateView Pv{PH, PU, 1, 0};
   const int Cur = 0, Next = 1, Compat = X.Compat;
   Real *CurTr  = A->Tracers[time_index(Cur, A->TracerCurTimeIndex, A->NTimeLevels)];
   Real *NextTr = A->Tracers[time_index(Next, A->TracerCurTimeIndex, A->NTimeLevels)];
   Real *PTr    = S->P.Tr[0];
   const int s0 = Phase == 0 ? 0 : 2, s1 = Phase == 0 ? 2 : 3;
   for (int Stage = s0; Stage <= s1; ++Stage) {
      if (Stage == 0) {
         launch_weight_tracers(X.L(), D.NCellsAll, D.NCellsSize, D.K, S->NTracers, NextTr, CurTr, St.hraw(Cur, Compat));
         if (X.all_tend(St, CurTr, Cur, Cur, C->StageTime[0])) return 1;
         X.update_thick(St.h(Next), St.h(Cur), C->B[0]);
         X.update_vel(St.u(Next), St.u(Cur), C->B[0]);
         launch_accumulate_tracers(X.L(), D.NCellsAll, D.NCellsSize, D.K, S->NTracers, NextTr, C->B[0], S->T.TracerTend);
      } else {
         if (!(Phase == 1 && Stage == 2)) {
            X.update_thick(Pv.h(Cur), St.h(Cur), C->A[Stage]);
            X.update_vel(Pv.u(Cur), St.u(Cur), C->A[Stage]);
            X.update_tracers(PTr, CurTr, Pv.hraw(Cur, Compat), St.hraw(Cur, Compat), C->A[Stage]);
         }
         if (Phase == 0 && Stage == 2) return 0; // exchange point 0 (RungeKutta4Stepper.cpp:108-114)
         if (X.all_tend(Pv, PTr, Cur, Cur, C->StageTime[Stage])) return 1;
         X.update_thick(St.h(Next), St.h(Next), C->B[Stage]);
         X.update_vel(St.u(Next), St.u(Next), C->B[Stage]);
         launch_accumulate_tracers(X.L(), D.NCellsAll, D.NCellsSize, D.K, S->NTracers, NextTr, C->B[Stage], S->T.TracerTend);
      }
   }
   launch_finalize_tracers(X.L(), D.NCellsAll, D.NCellsSize, D.K, S->NTracers, NextTr, St.hraw(Next, Compat));
   return 0;
}

int rk2_phase(const Ctx &X, OmegaStateArrays *A, const OmegaStepCoeffs *C) {
   StateView St{A->LayerThickness, A->NormalVelocity, A->NTimeLevels, A->CurTimeIndex};
   const int Cur = 0, Next = 1, Compat = X.Compat;
   Real *CurTr  = A->Tracers[time_index(Cur, A->TracerCurTimeIndex, A->NTimeLevels)];
   Real *NextTr = A->Tracers[time_index(Next, A->TracerCurTimeIndex, A->NTimeLevels)];
   if (X.all_tend(St, CurTr, Cur, Cur, C->StageTime[0])) return 1;
   X.update_thick(St.h(Next), St.h(Cur), C->A[1]);
   X.update_vel(St.u(Next), St.u(Cur), C->A[1]);
   X.update_tracers(NextTr, CurTr, St.hraw(Next, Compat), St.hraw(Cur, Compat), C->A[1]);
   if (X.all_tend(St, NextTr, Next, Next, C->StageTime[1])) return 1;
   X.update_thick(St.h(Next), St.h(Cur), C->B[1]);
   X.update_vel(St.u(Next), St.u(Cur), C->B[1]);
   X.update_tracers(NextTr, CurTr, St.hraw(Next, Compat), St.hraw(Cur, Compat), C->B[1]);
   return 0;
}

int fb_phase(const Ctx &X, OmegaStateArrays *A, const OmegaStepCoeffs *C) {
   OmegaStepper *S = X.S;
   StateView St{A->LayerThickness, A->NormalVelocity, A->NTimeLevels, A->CurTimeIndex};
   const int Cur = 0, Next = 1, Compat = X.Compat;
   Real *CurTr  = A->Tracers[time_index(Cur, A->TracerCurTimeIndex, A->NTimeLevels)];
   Real *NextTr = A->Tracers[time_index(Next, A->TracerCurTimeIndex, A->NTimeLevels)];
   if (omega_tend_thickness(S, St.h(Cur), St.u(Cur), C->StageTime[0], X.St)) return 1;
   X.update_thick(St.h(Next), St.h(Cur), C->Dt);
   if (S->NTracers) {
      if (omega_tend_tracer(S, St.hraw(Cur, Compat), St.uraw(Cur, Compat), CurTr, X.St)) return 1;
      X.update_tracers(NextTr, CurTr, St.hraw(Next, Compat), St.hraw(Cur, Compat), C->Dt);
   }
   if (omega_tend_velocity(S, St.h(Next), St.u(Cur), St.uraw(Cur, Compat), C->StageTime[1], X.St)) return 1;
   X.update_vel(St.u(Next), St.u(Cur), C->Dt);
   return 0;
}

int n_phases(OmegaStepperType Type) { return Type == OMEGA_STEPPER_RK4 ? 2 : 1; }

} // namespace

extern "C" {

int omega_stepper_step_phase(OmegaStepper *S, OmegaStepperType Type, OmegaStateArrays *State,
                             const OmegaStepCoeffs *Coeffs, int Phase, int Fused, int RawTimeLevelCompat,
                             void *Stream) {
   OMEGA_REQUIRE(S && State && Coeffs, "omega_stepper_step_phase: null argument");
   OMEGA_REQUIRE(State->NTimeLevels == 2, "omega_stepper_step_phase: NTimeLevels must be 2");
   OMEGA_REQUIRE(State->NTracers == S->NTracers, "omega_stepper_step_phase: NTracers mismatch");
   OMEGA_REQUIRE(Phase >= 0 && Phase < n_phases(Type), "omega_stepper_step_phase: bad phase");
   Ctx X{S, (cudaStream_t)Stream, RawTimeLevelCompat};
   int rc;
   S->LastPath = 0;
   S->SplitRecorded = 0;
   // forward-backward with raw time-level indexing and CurTimeIndex == 1: the tracer fluxes see the NEW thickness
   // while the thickness flux sees the old one (ForwardBackwardStepper.cpp:55-60, P1): not one fused cell pass
   const bool fbRaw = Type == OMEGA_STEPPER_FB && RawTimeLevelCompat && State->CurTimeIndex == 1;
   // the fused kernels unroll six edges per cell (fused.cu kME); a mesh with heptagons steps on the general kernels
   const bool wide = S->M->MaxEdgesUsed > 6;
   // a host custom tendency runs between the built-in terms and the stage update: not inside a fused stage
   const bool custom = S->CustomThick || S->CustomVel;
   if (Fused && !fbRaw && !wide && !custom)
      rc = fused_phase(S, Type, State, Coeffs, Phase, RawTimeLevelCompat, (cudaStream_t)Stream);
   else if (Type == OMEGA_STEPPER_RK4)
      rc = rk4_phase(X, State, Coeffs, Phase);
   else if (Type == OMEGA_STEPPER_RK2)
      rc = rk2_phase(X, State, Coeffs);
   else
      rc = fb_phase(X, State, Coeffs);
   if (rc) return rc;
   OMEGA_CUDA_CHECK(cudaGetLastError());
   return 0;
}

int omega_stepper_exchange_list(OmegaStepper *S, OmegaStepperType Type, const OmegaStateArrays *X, int Phase,
                                Real **Ptrs, int *Elems, int *NOuter, int *NArrays) {
   OMEGA_REQUIRE(S && X && Ptrs && Elems && NOuter && NArrays, "omega_stepper_exchange_list: null argument");
   if (Type == OMEGA_STEPPER_RK4 && Phase == 0) {
      const int b = S->ProvisExchangeSlot;
      Ptrs[0] = S->P.H[b];
      Ptrs[1] = S->P.U[b];
      Ptrs[2] = S->P.Tr[b];
   } else {
      Ptrs[0] = X->LayerThickness[time_index(1, X->CurTimeIndex, X->NTimeLevels)];
      Ptrs[1] = X->NormalVelocity[time_index(1, X->CurTimeIndex, X->NTimeLevels)];
      Ptrs[2] = X->Tracers[time_index(1, X->TracerCurTimeIndex, X->NTimeLevels)];
   }
   Elems[0] = OMEGA_ON_CELL; Elems[1] = OMEGA_ON_EDGE; Elems[2] = OMEGA_ON_CELL;
   NOuter[0] = 1; NOuter[1] = 1; NOuter[2] = S->NTracers;
   *NArrays = S->NTracers ? 3 : 2;
   return 0;
}

// OceanState::updateTimeLevels + Tracers::updateTimeLevels without the halo
// exchange (src/ocn/OceanState.cpp:348-365, src/ocn/Tracers.cpp:473-496)
int omega_state_rotate(OmegaStateArrays *X) {
   OMEGA_REQUIRE(X, "omega_state_rotate: null argument");
   X->CurTimeIndex       = (X->CurTimeIndex + 1) % X->NTimeLevels;
   X->TracerCurTimeIndex = (X->TracerCurTimeIndex + 1) % X->NTimeLevels;
   return 0;
}

// TimeStepper::doStep.  With a Halo the step includes its halo exchanges (RungeKutta4Stepper.cpp:108-114,
// 130-131; OceanState::updateTimeLevels / Tracers::updateTimeLevels, OceanState.cpp:348-365): after phase p
// the arrays of omega_stepper_exchange_list go through ONE exchange round (halo.cu).  When the fused tiled
// kernels ran the phase's last stage boundary-tiles-first, the round runs on the halo's own high-priority
// stream behind the boundary tiles, concurrently with the interior tiles, and Stream waits for it before
// the next phase; otherwise it runs in order on Stream.
int omega_stepper_do_step(OmegaStepper *S, OmegaStepperType Type, OmegaStateArrays *State,
                          const OmegaStepCoeffs *Coeffs, OmegaHalo *Halo, int Fused, int RawTimeLevelCompat,
                          void *Stream) {
   OMEGA_REQUIRE(S && State && Coeffs, "omega_stepper_do_step: null argument");
   cudaStream_t St = (cudaStream_t)Stream;
   if (Halo) {
      OMEGA_REQUIRE(Halo->M == S->M, "omega_stepper_do_step: the Halo belongs to another mesh");
      S->Halo = Halo;
   }
   for (int p = 0; p < n_phases(Type); ++p) {
      int rc = omega_stepper_step_phase(S, Type, State, Coeffs, p, Fused, RawTimeLevelCompat, Stream);
      if (rc) return rc;
      if (!Halo) continue;
      Real *ptrs[3];
      int elems[3], nouter[3], n = 0;
      rc = omega_stepper_exchange_list(S, Type, State, p, ptrs, elems, nouter, &n);
      if (rc) return rc;
      if (S->SplitRecorded) {
         OMEGA_CUDA_CHECK(cudaStreamWaitEvent(Halo->CommStream, Halo->EvReady, 0));
         rc = halo_exchange_round(Halo, ptrs, elems, nouter, n, Halo->CommStream);
         if (rc) return rc;
         OMEGA_CUDA_CHECK(cudaEventRecord(Halo->EvDone, Halo->CommStream));
         OMEGA_CUDA_CHECK(cudaStreamWaitEvent(St, Halo->EvDone, 0));
      } else {
         rc = halo_exchange_round(Halo, ptrs, elems, nouter, n, St);
         if (rc) return rc;
      }
   }
   return omega_state_rotate(State);
}

} // extern "C"
