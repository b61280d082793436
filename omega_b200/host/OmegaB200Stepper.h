// OmegaB200Stepper.h -- AuxiliaryState, Tendencies, TimeStepper mirrors (see
// OmegaB200.h).  One OmegaStepper handle of the C library backs the three
// objects; their public array members alias the library-owned device arrays
// (src/ocn/AuxiliaryState.h:37-42, src/ocn/Tendencies.h:55-57).
#pragma once
#include <functional>
#include <cmath>
#include "OmegaB200.h"

namespace OMEGA {

class AuxiliaryState : public Registry<AuxiliaryState> {
 public:
   // the public arrays of the six aux-var groups, flattened
   Array2DReal KineticEnergyCell, VelocityDivCell, FluxLayerThickEdge, MeanLayerThickEdge, SshCell,
       RelVortVertex, NormRelVortVertex, NormPlanetVortVertex, NormRelVortEdge, NormPlanetVortEdge,
       Del2Edge, Del2DivCell, Del2RelVortVertex;
   Array3DReal HTracersEdge, Del2TracersCell;
   Array1DReal ZonalStressCell, MeridStressCell, NormalStressEdge; ///< WindForcingAuxVars.h:16-20
   OmegaStepper *Handle = nullptr;
   const HorzMesh *Mesh;
   Halo *MeshHalo;
   I4 NTracers;

   AuxiliaryState(const std::string &Name, const HorzMesh *M, Halo *H, int NVertLayers, int NTracers_,
                  const OmegaTendConfig &Cfg)
       : Mesh(M), MeshHalo(H), NTracers(NTracers_) {
      OMEGA_B200_CHECK(omega_stepper_create(M->Handle, &Cfg, NTracers, &Handle), "AuxiliaryState: omega_stepper_create");
      OmegaAuxArrays A;
      OMEGA_B200_CHECK(omega_stepper_aux_arrays(Handle, &A), "omega_stepper_aux_arrays");
      const long K = NVertLayers, C = M->NCellsSize, E = M->NEdgesSize, V = M->NVerticesSize, T = NTracers;
#define OMEGA_ALIAS2(N, R) N = Array2DReal(#N + Name, A.N, {R, K})
      OMEGA_ALIAS2(KineticEnergyCell, C); OMEGA_ALIAS2(VelocityDivCell, C); OMEGA_ALIAS2(FluxLayerThickEdge, E);
      OMEGA_ALIAS2(MeanLayerThickEdge, E); OMEGA_ALIAS2(SshCell, C); OMEGA_ALIAS2(RelVortVertex, V);
      OMEGA_ALIAS2(NormRelVortVertex, V); OMEGA_ALIAS2(NormPlanetVortVertex, V); OMEGA_ALIAS2(NormRelVortEdge, E);
      OMEGA_ALIAS2(NormPlanetVortEdge, E); OMEGA_ALIAS2(Del2Edge, E); OMEGA_ALIAS2(Del2DivCell, C);
      OMEGA_ALIAS2(Del2RelVortVertex, V);
#undef OMEGA_ALIAS2
      HTracersEdge    = Array3DReal("HTracersEdge" + Name, A.HTracersEdge, {T, E, K});
      Del2TracersCell = Array3DReal("Del2TracersCell" + Name, A.Del2TracersCell, {T, C, K});
      ZonalStressCell  = Array1DReal("ZonalStressCell" + Name, A.ZonalStressCell, {C});
      MeridStressCell  = Array1DReal("MeridStressCell" + Name, A.MeridStressCell, {C});
      NormalStressEdge = Array1DReal("NormalStressEdge" + Name, A.NormalStressEdge, {E});
   }
   /// the two forcing arrays a host reads from its forcing file (host order, [NCellsSize])
   void setWindStress(const double *ZonalHost, const double *MeridHost, void *Stream = nullptr) {
      OMEGA_B200_CHECK(omega_stepper_set_wind_stress(Handle, ZonalHost, MeridHost, Stream), "AuxiliaryState::setWindStress");
   }
   /// src/ocn/AuxiliaryState.cpp:313-323: only the variables that are not computed need an exchange
   I4 exchangeHalo(void *Stream = nullptr) {
      if (!MeshHalo) return 0;
      I4 Err = MeshHalo->exchangeFullArrayHalo(ZonalStressCell.data(), OnCell, 1, 1, Stream);
      Err += MeshHalo->exchangeFullArrayHalo(MeridStressCell.data(), OnCell, 1, 1, Stream);
      return Err;
   }
   ~AuxiliaryState() { omega_stepper_destroy(Handle); }
   AuxiliaryState(const AuxiliaryState &) = delete;

   /// src/ocn/AuxiliaryState.cpp:60-143
   void computeMomAux(const OceanState *State, int ThickTimeLevel, int VelTimeLevel, void *Stream = nullptr) const {
      Array2DReal H, U;
      State->getLayerThickness(H, ThickTimeLevel);
      State->getNormalVelocity(U, VelTimeLevel);
      OMEGA_B200_CHECK(omega_aux_compute_mom(Handle, H.data(), U.data(), Stream), "computeMomAux");
   }
   /// src/ocn/AuxiliaryState.cpp:146-191
   void computeAll(const OceanState *State, const Array3DReal &TracerArray, int ThickTimeLevel, int VelTimeLevel,
                   void *Stream = nullptr) const {
      Array2DReal H, U;
      State->getLayerThickness(H, ThickTimeLevel);
      State->getNormalVelocity(U, VelTimeLevel);
      OMEGA_B200_CHECK(omega_aux_compute_all(Handle, H.data(), U.data(), TracerArray.data(), Stream), "computeAll");
   }
   void computeAll(const OceanState *State, const Array3DReal &TracerArray, int TimeLevel) const {
      computeAll(State, TracerArray, TimeLevel, TimeLevel);
   }
};

class Tendencies : public Registry<Tendencies> {
 public:
   /// src/ocn/Tendencies.h:51-53: a host term called after the built-in terms of a thickness / velocity tendency
   /// evaluation; it adds into the tendency array (first argument).  Here the arrays are device arrays in device
   /// order, the state is a one-level view of the thickness and velocity of the evaluation (levels 0, 0; the
   /// thickness is empty in the ...Only evaluations), the time is in seconds, and the work goes on customStream().
   using CustomTendencyType = std::function<void(Array2DReal, const OceanState *, const AuxiliaryState *, int, int, double)>;

   Array2DReal LayerThicknessTend, NormalVelocityTend;
   Array3DReal TracerTend;
   OmegaStepper *Handle; // shared with the AuxiliaryState it was built on

   /// Tendencies::create with custom terms (Tendencies.h:150-157); while one is installed TimeStepper::doStep runs
   /// the kernel-per-term path, the callbacks called in every stage (Tendencies.cpp:288-293, 416-421)
   Tendencies(const std::string &Name, const HorzMesh *M, const AuxiliaryState *Aux, int NVertLayers, CustomTendencyType CustomThick,
              CustomTendencyType CustomVel)
       : Tendencies(Name, M, Aux, NVertLayers) {
      setCustomTendencies(std::move(CustomThick), std::move(CustomVel));
   }
   void setCustomTendencies(CustomTendencyType CustomThick, CustomTendencyType CustomVel) {
      CustomThicknessTend = std::move(CustomThick);
      CustomVelocityTend  = std::move(CustomVel);
      OMEGA_B200_CHECK(omega_stepper_set_custom_tendencies(Handle, CustomThicknessTend ? &thickTrampoline : nullptr,
                                                           CustomVelocityTend ? &velTrampoline : nullptr, this),
                       "Tendencies: omega_stepper_set_custom_tendencies");
   }
   void *customStream() const { return CustomStream; }
   Tendencies(const Tendencies &) = delete; // the library holds `this` while a custom term is installed

   Tendencies(const std::string &Name, const HorzMesh *M, const AuxiliaryState *Aux, int NVertLayers)
       : Handle(Aux->Handle), Mesh(M), AuxState(Aux), NVertLayersT(NVertLayers) {
      OmegaTendArrays T;
      OMEGA_B200_CHECK(omega_stepper_tend_arrays(Handle, &T), "omega_stepper_tend_arrays");
      const long K = NVertLayers;
      LayerThicknessTend = Array2DReal("LayerThicknessTend" + Name, T.LayerThicknessTend, {M->NCellsSize, K});
      NormalVelocityTend = Array2DReal("NormalVelocityTend" + Name, T.NormalVelocityTend, {M->NEdgesSize, K});
      TracerTend         = Array3DReal("TracerTend" + Name, T.TracerTend, {Aux->NTracers, M->NCellsSize, K});
   }
   void readTendConfig(const OmegaTendConfig &Cfg) { OMEGA_B200_CHECK(omega_stepper_set_config(Handle, &Cfg), "readTendConfig"); }

   // raw vector indexing where the reference indexes raw (src/ocn/Tendencies.cpp:322,445,546-547)
   static double *rawU(const OceanState *S, int Level) { return S->NormalVelocity[Level].data(); }
   static double *rawH(const OceanState *S, int Level) { return S->LayerThickness[Level].data(); }
   static double *mapU(const OceanState *S, int Level) { Array2DReal A; S->getNormalVelocity(A, Level); return A.data(); }
   static double *mapH(const OceanState *S, int Level) { Array2DReal A; S->getLayerThickness(A, Level); return A.data(); }

   /// src/ocn/Tendencies.cpp:257-486
   void computeThicknessTendenciesOnly(const OceanState *S, const AuxiliaryState *, int, int VelLvl, double Time, void *St = nullptr) {
      OMEGA_B200_CHECK(omega_tend_thickness_only(Handle, mapU(S, VelLvl), Time, St), "computeThicknessTendenciesOnly");
   }
   void computeVelocityTendenciesOnly(const OceanState *S, const AuxiliaryState *, int, int VelLvl, double Time, void *St = nullptr) {
      OMEGA_B200_CHECK(omega_tend_velocity_only(Handle, mapU(S, VelLvl), rawU(S, VelLvl), Time, St), "computeVelocityTendenciesOnly");
   }
   void computeTracerTendenciesOnly(const OceanState *S, const AuxiliaryState *, const Array3DReal &Tr, int, int VelLvl, double, void *St = nullptr) {
      OMEGA_B200_CHECK(omega_tend_tracer_only(Handle, rawU(S, VelLvl), Tr.data(), St), "computeTracerTendenciesOnly");
   }
   /// src/ocn/Tendencies.cpp:488-600
   void computeThicknessTendencies(const OceanState *S, const AuxiliaryState *, int ThickLvl, int VelLvl, double Time, void *St = nullptr) {
      OMEGA_B200_CHECK(omega_tend_thickness(Handle, mapH(S, ThickLvl), mapU(S, VelLvl), Time, St), "computeThicknessTendencies");
   }
   void computeVelocityTendencies(const OceanState *S, const AuxiliaryState *, int ThickLvl, int VelLvl, double Time, void *St = nullptr) {
      OMEGA_B200_CHECK(omega_tend_velocity(Handle, mapH(S, ThickLvl), mapU(S, VelLvl), rawU(S, VelLvl), Time, St), "computeVelocityTendencies");
   }
   void computeTracerTendencies(const OceanState *S, const AuxiliaryState *, const Array3DReal &Tr, int ThickLvl, int VelLvl, double, void *St = nullptr) {
      OMEGA_B200_CHECK(omega_tend_tracer(Handle, rawH(S, ThickLvl), rawU(S, VelLvl), Tr.data(), St), "computeTracerTendencies");
   }
   void computeAllTendencies(const OceanState *S, const AuxiliaryState *, const Array3DReal &Tr, int ThickLvl, int VelLvl, double Time, void *St = nullptr) {
      OMEGA_B200_CHECK(omega_tend_all(Handle, mapH(S, ThickLvl), mapU(S, VelLvl), rawU(S, VelLvl), Tr.data(), Time, St), "computeAllTendencies");
   }

 private:
   const HorzMesh *Mesh;
   const AuxiliaryState *AuxState;
   int NVertLayersT;
   CustomTendencyType CustomThicknessTend, CustomVelocityTend;
   void *CustomStream = nullptr;
   static int trampoline(void *User, bool Vel, double *, const double *H, const double *U, double Time, void *Stream) {
      auto *Self = static_cast<Tendencies *>(User);
      OceanState Stage(Self->Mesh, const_cast<double *>(H), const_cast<double *>(U), Self->NVertLayersT);
      Self->CustomStream = Stream;
      if (Vel) Self->CustomVelocityTend(Self->NormalVelocityTend, &Stage, Self->AuxState, 0, 0, Time);
      else Self->CustomThicknessTend(Self->LayerThicknessTend, &Stage, Self->AuxState, 0, 0, Time);
      return 0;
   }
   static int thickTrampoline(void *User, double *T, const double *H, const double *U, double Time, void *Stream) {
      return trampoline(User, false, T, H, U, Time, Stream);
   }
   static int velTrampoline(void *User, double *T, const double *H, const double *U, double Time, void *Stream) {
      return trampoline(User, true, T, H, U, Time, Stream);
   }
};

enum class TimeStepperType { ForwardBackward, RungeKutta4, RungeKutta2, Invalid };

/// TimeStepper::doStep (src/timeStepping/TimeStepper.h:82-84).  SimTime is
/// elapsed seconds; Coeffs are the stage coefficients in seconds as
/// TimeInterval::get evaluates them (src/infra/TimeMgr.cpp:382-391).
class TimeStepper : public Registry<TimeStepper> {
 public:
   TimeStepperType Type;
   OmegaStepCoeffs Coeffs;
   Tendencies *Tend;
   AuxiliaryState *AuxState;
   Tracers *Tr;
   bool Fused = true, RawTimeLevelCompat = true;

   TimeStepper(TimeStepperType T, double Dt, Tendencies *Td, AuxiliaryState *Ax, Tracers *Trc) : Type(T), Tend(Td), AuxState(Ax), Tr(Trc) {
      setTimeStep(Dt);
   }
   /// Stage coefficients in seconds the way the reference forms them: RKB[s] * TimeStep in exact rational
   /// arithmetic, evaluated as Whole + Numer/Denom (src/infra/TimeMgr.cpp:382-391,747-766).
   /// Dt is taken as the rational with denominator <= 1e9 closest to it (continued fraction), which is
   /// exact for every time step written with a few decimals (0.5, 30.5, 0.1 ...), as capi.rk_coeffs does.
   static void toFraction(double X, long long &Num, long long &Den) {
      long long p0 = 0, q0 = 1, p1 = 1, q1 = 0;
      double r = X;
      for (int it = 0; it < 64; ++it) {
         const double fl = std::floor(r);
         const long long a = (long long)fl;
         const long long p2 = a * p1 + p0, q2 = a * q1 + q0;
         if (q2 > 1000000000LL) break;
         p0 = p1; q0 = q1; p1 = p2; q1 = q2;
         const double f = r - fl;
         if (f < 1e-15 || (double)p1 / (double)q1 == X) break;
         r = 1.0 / f;
      }
      Num = p1; Den = q1;
   }
   void setTimeStep(double Dt) {
      Coeffs = OmegaStepCoeffs();
      long long dn, dd;
      toFraction(Dt, dn, dd);
      auto frac = [&](long long num, long long den) { // (num/den) * (dn/dd) as Whole + Numer/Denom
         long long n = num * dn, d = den * dd;
         long long a = n < 0 ? -n : n, b = d; while (b) { long long t = a % b; a = b; b = t; }
         if (a > 1) { n /= a; d /= a; }
         const long long whole = n / d, rem = n % d;
         if (rem == 0) return (double)whole;
         return (double)whole + (double)rem / (double)d;
      };
      Coeffs.Dt = frac(1, 1);
      if (Type == TimeStepperType::RungeKutta4) {
         Coeffs.NStages = 4;
         const long long an[4] = {0, 1, 1, 1}, ad[4] = {1, 2, 2, 1}, bn[4] = {1, 1, 1, 1}, bd[4] = {6, 3, 3, 6};
         for (int s = 0; s < 4; ++s) { Coeffs.A[s] = frac(an[s], ad[s]); Coeffs.B[s] = frac(bn[s], bd[s]); }
      } else if (Type == TimeStepperType::RungeKutta2) {
         Coeffs.NStages = 2; Coeffs.A[1] = frac(1, 2); Coeffs.B[1] = frac(1, 1);
      } else {
         Coeffs.NStages = 1;
      }
   }
   void doStep(OceanState *State, double &SimTime, void *Stream = nullptr) {
      OmegaStateArrays X;
      X.NTimeLevels = State->NTimeLevels; X.CurTimeIndex = State->CurTimeIndex;
      X.TracerCurTimeIndex = Tr->CurTimeIndex; X.NTracers = Tr->NTracers;
      for (int i = 0; i < 2; ++i) {
         X.LayerThickness[i] = State->LayerThickness[i].data();
         X.NormalVelocity[i] = State->NormalVelocity[i].data();
         X.Tracers[i] = Tr->TracerArrays[i].data();
      }
      OmegaStepperType T = Type == TimeStepperType::RungeKutta4 ? OMEGA_STEPPER_RK4
                           : Type == TimeStepperType::RungeKutta2 ? OMEGA_STEPPER_RK2 : OMEGA_STEPPER_FB;
      for (int s = 0; s < 4; ++s) Coeffs.StageTime[s] = SimTime + (T == OMEGA_STEPPER_FB ? (s ? Coeffs.Dt : 0.0) : Coeffs.A[s]);
      // the whole step in the library: stage kernels, the halo exchange rounds (overlapped with the
      // interior tiles) and the rotation of both time-level rings (OceanState / Tracers::updateTimeLevels)
      OmegaHalo *H = State->MeshHalo && !State->MeshHalo->Neighbors.empty() ? State->MeshHalo->Handle : nullptr;
      OMEGA_B200_CHECK(omega_stepper_do_step(AuxState->Handle, T, &X, &Coeffs, H, Fused, RawTimeLevelCompat, Stream), "doStep");
      State->CurTimeIndex = X.CurTimeIndex;
      Tr->CurTimeIndex    = X.TracerCurTimeIndex;
      SimTime += Coeffs.Dt;
   }
};

} // namespace OMEGA
