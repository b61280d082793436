// ref_driver.cpp -- TEST INFRASTRUCTURE.  C entry points (the same oracle_* API
// as oracle/omega_oracle.c) that drive the REFERENCE's own classes:
// AuxiliaryState::computeAll/computeMomAux, Tendencies::compute*,
// TimeStepper::doStep (RK4 / RK2 / forward-backward), with the reference's own
// init sequence (TimeStepper::init1 -> OceanState::init -> AuxiliaryState::init
// -> Tendencies::init -> TimeStepper::init2, src/ocn/OceanInit.cpp:107-142) fed
// from a key/value Config stand-in.  Only marshalling lives here; no tendency
// or stepping arithmetic is restated in this file.
#include <cstring>
#include <sstream>

#include "AuxiliaryState.h"
#include "HorzOperators.h"
#include "OceanState.h"
#include "Tendencies.h"
#include "TimeStepper.h"
#include "omega_b200.h"

using namespace OMEGA;

namespace {

struct RefStepper {
   HorzMesh Mesh;
   VertCoord VC;
   OmegaTendConfig Cfg;
   int T, K;
   int StateIdx; // CurTimeIndex of the default OceanState (private there; flips once per doStep)
   int StepperKind;
   double Dt;
};

RefStepper *Live = nullptr;

template <class V, class S> void fill1(V &View, const S *Src, int N) {
   for (int i = 0; i < N; ++i) View(i) = Src[i];
}
template <class V, class S> void fill2(V &View, const S *Src, int N, int M) {
   for (int i = 0; i < N; ++i)
      for (int j = 0; j < M; ++j) View(i, j) = Src[(size_t)i * M + j];
}

void push_config(const OmegaTendConfig &G) {
   auto B = [](int v) { return std::string(v ? "true" : "false"); };
   auto &S = Config::store();
   S["ThicknessFluxTendencyEnable"] = B(G.ThicknessFluxTendencyEnable);
   S["PVTendencyEnable"] = B(G.PVTendencyEnable);
   S["KETendencyEnable"] = B(G.KETendencyEnable);
   S["SSHTendencyEnable"] = B(G.SSHTendencyEnable);
   S["VelDiffTendencyEnable"] = B(G.VelDiffTendencyEnable);
   S["VelHyperDiffTendencyEnable"] = B(G.VelHyperDiffTendencyEnable);
   S["WindForcingTendencyEnable"] = B(G.WindForcingTendencyEnable);
   S["BottomDragTendencyEnable"] = B(G.BottomDragTendencyEnable);
   S["TracerHorzAdvTendencyEnable"] = B(G.TracerHorzAdvTendencyEnable);
   S["TracerDiffTendencyEnable"] = B(G.TracerDiffTendencyEnable);
   S["TracerHyperDiffTendencyEnable"] = B(G.TracerHyperDiffTendencyEnable);
   S["UseCustomTendency"] = B(G.ManufacturedSolutionTendency);
   S["ManufacturedSolutionTendency"] = B(G.ManufacturedSolutionTendency);
   Config::set("ViscDel2", G.ViscDel2);
   Config::set("ViscDel4", G.ViscDel4);
   Config::set("DivFactor", G.DivFactor);
   Config::set("Density0", G.Density0);
   Config::set("BottomDragCoeff", G.BottomDragCoeff);
   Config::set("EddyDiff2", G.EddyDiff2);
   Config::set("EddyDiff4", G.EddyDiff4);
   Config::set("WavelengthX", G.ManufacturedWavelengthX);
   Config::set("WavelengthY", G.ManufacturedWavelengthY);
   Config::set("Amplitude", G.ManufacturedAmplitude);
   S["FluxThicknessType"] = G.FluxThicknessType ? "Upwind" : "Center";
   S["FluxTracerType"] = G.FluxTracerType ? "Upwind" : "Center";
   S["InterpType"] = G.WindStressInterpType ? "Anisotropic" : "Isotropic";
}

const char *stepper_name(int Kind) {
   return Kind == OMEGA_STEPPER_RK4 ? "RungeKutta4" : Kind == OMEGA_STEPPER_RK2 ? "RungeKutta2" : "Forward-Backward";
}

// Host custom tendencies (TimeStepperTest.cpp:49-72 installs one through the public Tendencies::create): a second
// Tendencies object "Custom" whose two std::function members forward to C function pointers
OmegaCustomTendency GThick = nullptr, GVel = nullptr;
void *GUser                = nullptr;
Tendencies *CustomTend     = nullptr;
Tendencies *active_tend() { return CustomTend ? CustomTend : Tendencies::getDefault(); }

Tendencies::CustomTendencyType forward_to(OmegaCustomTendency *Slot) {
   return [Slot](Array2DReal Tend, const OceanState *State, const AuxiliaryState *, int ThickLevel, int VelLevel, TimeInstant Time) {
      if (!*Slot) return;
      Array2DReal H, U;
      State->getLayerThickness(H, ThickLevel);
      State->getNormalVelocity(U, VelLevel);
      R8 Seconds;
      (Time - TimeStepper::getDefault()->getStartTime()).get(Seconds, TimeUnits::Seconds);
      (*Slot)(GUser, Tend.data(), H.data(), U.data(), Seconds, nullptr);
   };
}

// (re)create the default time stepper with the reference's two-phase init
void init_stepper(RefStepper *R, int Kind, double Dt, bool Attach) {
   TimeStepper::clear();
   OceanState::erase("ProvisDefault");
   Calendar::reset();
   auto &S = Config::store();
   S["CalendarType"] = "No Leap";
   S["TimeStepper"] = stepper_name(Kind);
   std::ostringstream TS;
   TS.precision(17);
   TS << "0000_00:00:" << Dt;
   S["TimeStep"] = TS.str();
   S["StartTime"] = "0001-01-01_00:00:00";
   S["StopTime"] = "0101-01-01_00:00:00";
   S["RunDuration"] = "none";
   TimeStepper::init1(); // src/timeStepping/TimeStepper.cpp:226-305
   if (Attach) { // TimeStepper::init2 (TimeStepper.cpp:302-312) with the tendencies in force
      TimeStepper::getDefault()->attachData(active_tend(), AuxiliaryState::getDefault(), HorzMesh::getDefault(), Halo::getDefault());
   }
   R->StepperKind = Kind;
   R->Dt = Dt;
}

TimeInstant time_at(double Seconds) {
   TimeInstant T = TimeStepper::getDefault()->getStartTime();
   TimeInterval D(Seconds, TimeUnits::Seconds);
   return T + D;
}

// put caller arrays into the default state's CURRENT slots
void load_state(RefStepper *R, const double *H, const double *U) {
   OceanState *St = OceanState::getDefault();
   const size_t K = R->K;
   if (H) std::memcpy(St->LayerThickness[R->StateIdx].data(), H, sizeof(double) * R->Mesh.NCellsSize * K);
   if (U) std::memcpy(St->NormalVelocity[R->StateIdx].data(), U, sizeof(double) * R->Mesh.NEdgesSize * K);
}

Array3DReal wrap_tracers(RefStepper *R, const double *Tr) {
   return Array3DReal(const_cast<double *>(Tr), R->T, R->Mesh.NCellsSize, R->K);
}

} // namespace

extern "C" {

// The reference's time arithmetic on a stage coefficient: Mult * TimeStep as the steppers form it
// (RungeKutta4Stepper.cpp:96-118: RKA[Stage] * TimeStep), evaluated in seconds (TimeMgr.cpp:382-391)
int ref_time_coeff(double Dt, double Mult, double *Out) {
   TimeInterval TimeStep(Dt, TimeUnits::Seconds);
   TimeInterval Coeff = Mult * TimeStep;
   R8 Seconds;
   Coeff.get(Seconds, TimeUnits::Seconds);
   *Out = Seconds;
   return 0;
}

// Multi-task tests: Fn is called at every halo exchange the reference's steppers perform (see ref_shim.h)
int ref_set_halo_callback(void (*Fn)(void *, int, int, int)) {
   Halo::callback() = Fn;
   return 0;
}

int oracle_stepper_destroy(RefStepper *R);

int oracle_stepper_create(const OmegaMeshDesc *D, const OmegaTendConfig *Cfg, int NTracers, RefStepper **Out) {
   if (Live) oracle_stepper_destroy(Live); // the reference keeps one "Default" instance of everything
   RefStepper *R = new RefStepper();
   R->Cfg = *Cfg;
   R->T = NTracers;
   R->K = D->NVertLayers;
   R->StateIdx = 0;
   HorzMesh &M = R->Mesh;
   M.NCellsOwned = D->NCellsOwned; M.NCellsAll = D->NCellsAll; M.NCellsSize = D->NCellsSize;
   M.NEdgesOwned = D->NEdgesOwned; M.NEdgesAll = D->NEdgesAll; M.NEdgesSize = D->NEdgesSize;
   M.NVerticesOwned = D->NVerticesOwned; M.NVerticesAll = D->NVerticesAll; M.NVerticesSize = D->NVerticesSize;
   M.MaxEdges = D->MaxEdges; M.MaxEdges2 = D->MaxEdges2; M.VertexDegree = D->VertexDegree;
   const int C = D->NCellsSize, E = D->NEdgesSize, V = D->NVerticesSize, ME = D->MaxEdges, ME2 = D->MaxEdges2,
             VD = D->VertexDegree, K = D->NVertLayers;
   M.NEdgesOnCell = Array1DI4("NEdgesOnCell", C);       fill1(M.NEdgesOnCell, D->NEdgesOnCell, C);
   M.EdgesOnCell = Array2DI4("EdgesOnCell", C, ME);      fill2(M.EdgesOnCell, D->EdgesOnCell, C, ME);
   M.CellsOnEdge = Array2DI4("CellsOnEdge", E, 2);       fill2(M.CellsOnEdge, D->CellsOnEdge, E, 2);
   M.VerticesOnEdge = Array2DI4("VerticesOnEdge", E, 2); fill2(M.VerticesOnEdge, D->VerticesOnEdge, E, 2);
   M.NEdgesOnEdge = Array1DI4("NEdgesOnEdge", E);        fill1(M.NEdgesOnEdge, D->NEdgesOnEdge, E);
   M.EdgesOnEdge = Array2DI4("EdgesOnEdge", E, ME2);     fill2(M.EdgesOnEdge, D->EdgesOnEdge, E, ME2);
   M.CellsOnVertex = Array2DI4("CellsOnVertex", V, VD);  fill2(M.CellsOnVertex, D->CellsOnVertex, V, VD);
   M.EdgesOnVertex = Array2DI4("EdgesOnVertex", V, VD);  fill2(M.EdgesOnVertex, D->EdgesOnVertex, V, VD);
#define R1(NAME, N) M.NAME = Array1DReal(#NAME, N); fill1(M.NAME, D->NAME, N)
   R1(AreaCell, C); R1(BottomDepth, C); R1(XCell, C); R1(YCell, C);
   R1(DcEdge, E); R1(DvEdge, E); R1(AngleEdge, E); R1(FEdge, E); R1(XEdge, E); R1(YEdge, E);
   R1(AreaTriangle, V); R1(FVertex, V);
#undef R1
   M.BottomDepthH = HostArray1DReal("BottomDepthH", C); fill1(M.BottomDepthH, D->BottomDepth, C);
   M.WeightsOnEdge = Array2DReal("WeightsOnEdge", E, ME2); fill2(M.WeightsOnEdge, D->WeightsOnEdge, E, ME2);
   M.KiteAreasOnVertex = Array2DReal("KiteAreasOnVertex", V, VD); fill2(M.KiteAreasOnVertex, D->KiteAreasOnVertex, V, VD);
   // derived fields: restated from HorzMesh::computeEdgeSign / setMasks / setMeshScaling
   // (src/ocn/HorzMesh.cpp:527-626), the one part of HorzMesh.cpp this build cannot compile
   M.EdgeSignOnCell = Array2DReal("EdgeSignOnCell", C, ME);
   for (int c = 0; c < M.NCellsAll; ++c)
      for (int i = 0; i < M.NEdgesOnCell(c); ++i)
         M.EdgeSignOnCell(c, i) = (c == M.CellsOnEdge(M.EdgesOnCell(c, i), 0)) ? -1.0 : 1.0;
   M.EdgeSignOnVertex = Array2DReal("EdgeSignOnVertex", V, VD);
   for (int v = 0; v < M.NVerticesAll; ++v)
      for (int i = 0; i < VD; ++i)
         M.EdgeSignOnVertex(v, i) = (v == M.VerticesOnEdge(M.EdgesOnVertex(v, i), 0)) ? -1.0 : 1.0;
   M.EdgeMask = Array2DReal("EdgeMask", E, K);
   deepCopy(M.EdgeMask, 1.0);
   for (int e = 0; e < M.NEdgesAll; ++e) {
      const I4 c1 = M.CellsOnEdge(e, 0), c2 = M.CellsOnEdge(e, 1);
      if (!(c1 >= 0 && c1 < M.NCellsAll) || !(c2 >= 0 && c2 < M.NCellsAll))
         for (int k = 0; k < K; ++k) M.EdgeMask(e, k) = 0.0;
   }
   M.MeshScalingDel2 = Array1DReal("MeshScalingDel2", E);
   M.MeshScalingDel4 = Array1DReal("MeshScalingDel4", E);
   for (int e = 0; e < M.NEdgesAll; ++e) M.MeshScalingDel2(e) = M.MeshScalingDel4(e) = 1.0;

   HorzMesh::defaultPtr() = &R->Mesh;
   R->VC.NVertLayers = K;
   VertCoord::defaultPtr() = &R->VC;
   Tracers::numTracers() = NTracers;
   Tracers::curTimeIndex() = 0;
   Tracers::arrays().clear();
   for (int i = 0; i < 2; ++i) Tracers::arrays().push_back(Array3DReal("Tracers", NTracers, C, K));

   push_config(*Cfg);
   Tendencies::clear();
   AuxiliaryState::clear();
   OceanState::clear();
   init_stepper(R, OMEGA_STEPPER_RK4, 60.0, false);
   OceanState::init();      // src/ocn/OceanState.cpp:30-60
   AuxiliaryState::init();  // src/ocn/AuxiliaryState.cpp:206-219
   Tendencies::init();      // src/ocn/Tendencies.cpp:26-72 (reads the config, builds the manufactured terms)
   CustomTend = nullptr;
   GThick = GVel = nullptr;
   TimeStepper::init2();
   Live = R;
   *Out = R;
   return 0;
}

int oracle_stepper_destroy(RefStepper *R) {
   if (!R) return 0;
   if (Live == R) {
      TimeStepper::clear();
      Tendencies::clear();
      CustomTend = nullptr;
      GThick = GVel = nullptr;
      AuxiliaryState::clear();
      OceanState::clear();
      Tracers::arrays().clear();
      Live = nullptr;
   }
   delete R;
   return 0;
}

int oracle_stepper_set_config(RefStepper *R, const OmegaTendConfig *Cfg) {
   if (Cfg->ManufacturedSolutionTendency != R->Cfg.ManufacturedSolutionTendency) return 2; // needs a new instance
   R->Cfg = *Cfg;
   push_config(*Cfg);
   Config TendConfig("Tendencies");
   Tendencies::getDefault()->readTendConfig(&TendConfig);
   if (CustomTend) CustomTend->readTendConfig(&TendConfig);
   AuxiliaryState::getDefault()->readConfigOptions(Config::getOmegaConfig());
   return 0;
}

int oracle_stepper_aux_arrays(RefStepper *, OmegaAuxArrays *A) {
   AuxiliaryState *X = AuxiliaryState::getDefault();
   A->KineticEnergyCell = X->KineticAux.KineticEnergyCell.data();
   A->VelocityDivCell = X->KineticAux.VelocityDivCell.data();
   A->FluxLayerThickEdge = X->LayerThicknessAux.FluxLayerThickEdge.data();
   A->MeanLayerThickEdge = X->LayerThicknessAux.MeanLayerThickEdge.data();
   A->SshCell = X->LayerThicknessAux.SshCell.data();
   A->RelVortVertex = X->VorticityAux.RelVortVertex.data();
   A->NormRelVortVertex = X->VorticityAux.NormRelVortVertex.data();
   A->NormPlanetVortVertex = X->VorticityAux.NormPlanetVortVertex.data();
   A->NormRelVortEdge = X->VorticityAux.NormRelVortEdge.data();
   A->NormPlanetVortEdge = X->VorticityAux.NormPlanetVortEdge.data();
   A->Del2Edge = X->VelocityDel2Aux.Del2Edge.data();
   A->Del2DivCell = X->VelocityDel2Aux.Del2DivCell.data();
   A->Del2RelVortVertex = X->VelocityDel2Aux.Del2RelVortVertex.data();
   A->HTracersEdge = X->TracerAux.HTracersEdge.data();
   A->Del2TracersCell = X->TracerAux.Del2TracersCell.data();
   A->NormalStressEdge = X->WindForcingAux.NormalStressEdge.data();
   A->ZonalStressCell = X->WindForcingAux.ZonalStressCell.data();
   A->MeridStressCell = X->WindForcingAux.MeridStressCell.data();
   return 0;
}

int oracle_stepper_set_custom_tendencies(RefStepper *R, OmegaCustomTendency Thickness, OmegaCustomTendency Velocity, void *User) {
   if (R->Cfg.ManufacturedSolutionTendency) return 2; // Tendencies::init gave those slots to the manufactured terms
   GThick = Thickness;
   GVel   = Velocity;
   GUser  = User;
   if ((Thickness || Velocity) && !CustomTend) {
      Config TendConfig("Tendencies");
      Config::getOmegaConfig()->get(TendConfig);
      CustomTend = Tendencies::create("Custom", HorzMesh::getDefault(), VertCoord::getDefault(), R->T, &TendConfig, forward_to(&GThick),
                                      forward_to(&GVel));
      if (!CustomTend) return 1;
      CustomTend->readTendConfig(&TendConfig);
   }
   init_stepper(R, R->StepperKind, R->Dt, true);
   return 0;
}

int oracle_stepper_tend_arrays(RefStepper *, OmegaTendArrays *T) {
   Tendencies *X = active_tend();
   T->LayerThicknessTend = X->LayerThicknessTend.data();
   T->NormalVelocityTend = X->NormalVelocityTend.data();
   T->TracerTend = X->TracerTend.data();
   return 0;
}

int oracle_mesh_derived(RefStepper *R, double *ESC, double *ESV, double *Mask) {
   HorzMesh &M = R->Mesh;
   std::memcpy(ESC, M.EdgeSignOnCell.data(), sizeof(double) * M.NCellsSize * M.MaxEdges);
   std::memcpy(ESV, M.EdgeSignOnVertex.data(), sizeof(double) * M.NVerticesSize * M.VertexDegree);
   for (int e = 0; e < M.NEdgesSize; ++e) Mask[e] = M.EdgeMask(e, 0);
   return 0;
}

int oracle_stepper_set_wind_stress(RefStepper *R, const double *Zonal, const double *Merid) {
   auto &W = AuxiliaryState::getDefault()->WindForcingAux;
   std::memcpy(W.ZonalStressCell.data(), Zonal, sizeof(double) * R->Mesh.NCellsSize);
   std::memcpy(W.MeridStressCell.data(), Merid, sizeof(double) * R->Mesh.NCellsSize);
   return 0;
}

int oracle_aux_compute_mom(RefStepper *R, const double *H, const double *U) {
   load_state(R, H, U);
   AuxiliaryState::getDefault()->computeMomAux(OceanState::getDefault(), 0, 0);
   return 0;
}

int oracle_aux_compute_all(RefStepper *R, const double *H, const double *U, const double *Tr) {
   load_state(R, H, U);
   AuxiliaryState::getDefault()->computeAll(OceanState::getDefault(), wrap_tracers(R, Tr), 0, 0);
   return 0;
}

// the *Only entry points read the velocity through the state; raw == mapped
// indexing requires StateIdx == 0 (always true before the first doStep)
int oracle_tend_thickness_only(RefStepper *R, const double *U, double Time) {
   load_state(R, nullptr, U);
   active_tend()->computeThicknessTendenciesOnly(OceanState::getDefault(), AuxiliaryState::getDefault(), 0, 0, time_at(Time));
   return 0;
}

int oracle_tend_velocity_only(RefStepper *R, const double *U, const double *URaw, double Time) {
   if (U != URaw || R->StateIdx != 0) return 2;
   load_state(R, nullptr, U);
   active_tend()->computeVelocityTendenciesOnly(OceanState::getDefault(), AuxiliaryState::getDefault(), 0, 0, time_at(Time));
   return 0;
}

int oracle_tend_tracer_only(RefStepper *R, const double *URaw, const double *Tr) {
   if (R->StateIdx != 0) return 2;
   load_state(R, nullptr, URaw);
   active_tend()->computeTracerTendenciesOnly(OceanState::getDefault(), AuxiliaryState::getDefault(), wrap_tracers(R, Tr), 0, 0, time_at(0.0));
   return 0;
}

int oracle_tend_all(RefStepper *R, const double *H, const double *U, const double *URaw, const double *Tr, double Time) {
   if (U != URaw || R->StateIdx != 0) return 2;
   load_state(R, H, U);
   active_tend()->computeAllTendencies(OceanState::getDefault(), AuxiliaryState::getDefault(), wrap_tracers(R, Tr), 0, 0, time_at(Time));
   return 0;
}

int oracle_stepper_do_step(RefStepper *R, OmegaStepperType Type, OmegaStateArrays *X, const OmegaStepCoeffs *C,
                           int /*Fused*/, int Compat) {
   if (!Compat) return 2; // the reference IS the raw-index behaviour
   if (Type != R->StepperKind || C->Dt != R->Dt) init_stepper(R, Type, C->Dt, true);
   OceanState *St = OceanState::getDefault();
   // bring the reference's ring indices in line with the caller's
   while (R->StateIdx != X->CurTimeIndex) {
      St->updateTimeLevels();
      R->StateIdx = (R->StateIdx + 1) % 2;
   }
   Tracers::curTimeIndex() = X->TracerCurTimeIndex;
   const size_t K = R->K, CS = (size_t)R->Mesh.NCellsSize * K, ES = (size_t)R->Mesh.NEdgesSize * K;
   for (int i = 0; i < 2; ++i) {
      std::memcpy(St->LayerThickness[i].data(), X->LayerThickness[i], sizeof(double) * CS);
      std::memcpy(St->NormalVelocity[i].data(), X->NormalVelocity[i], sizeof(double) * ES);
      if (R->T) std::memcpy(Tracers::arrays()[i].data(), X->Tracers[i], sizeof(double) * R->T * CS);
   }
   TimeInstant SimTime = time_at(C->StageTime[0]);
   TimeStepper::getDefault()->doStep(St, SimTime);
   R->StateIdx = (R->StateIdx + 1) % 2;
   for (int i = 0; i < 2; ++i) {
      std::memcpy(X->LayerThickness[i], St->LayerThickness[i].data(), sizeof(double) * CS);
      std::memcpy(X->NormalVelocity[i], St->NormalVelocity[i].data(), sizeof(double) * ES);
      if (R->T) std::memcpy(X->Tracers[i], Tracers::arrays()[i].data(), sizeof(double) * R->T * CS);
   }
   X->CurTimeIndex = R->StateIdx;
   X->TracerCurTimeIndex = Tracers::curTimeIndex();
   return 0;
}

// ---- src/ocn/HorzOperators.h functors for the known-answer tests --------------------
int oracle_op_divergence(RefStepper *R, const double *VecEdge, double *DivCell) {
   DivergenceOnCell Op(&R->Mesh);
   Array2DReal In(const_cast<double *>(VecEdge), R->Mesh.NEdgesSize, R->K), Out(DivCell, R->Mesh.NCellsSize, R->K);
   parallelFor({R->Mesh.NCellsAll, R->K / VecLength}, KOKKOS_LAMBDA(int I, int KC) { Op(Out, I, KC, In); });
   return 0;
}
int oracle_op_gradient(RefStepper *R, const double *ScalarCell, double *GradEdge) {
   GradientOnEdge Op(&R->Mesh);
   Array2DReal In(const_cast<double *>(ScalarCell), R->Mesh.NCellsSize, R->K), Out(GradEdge, R->Mesh.NEdgesSize, R->K);
   parallelFor({R->Mesh.NEdgesAll, R->K / VecLength}, KOKKOS_LAMBDA(int I, int KC) { Op(Out, I, KC, In); });
   return 0;
}
int oracle_op_curl(RefStepper *R, const double *VecEdge, double *CurlVertex) {
   CurlOnVertex Op(&R->Mesh);
   Array2DReal In(const_cast<double *>(VecEdge), R->Mesh.NEdgesSize, R->K), Out(CurlVertex, R->Mesh.NVerticesSize, R->K);
   parallelFor({R->Mesh.NVerticesAll, R->K / VecLength}, KOKKOS_LAMBDA(int I, int KC) { Op(Out, I, KC, In); });
   return 0;
}
int oracle_op_tangential_recon(RefStepper *R, const double *VecEdge, double *ReconEdge) {
   TangentialReconOnEdge Op(&R->Mesh);
   Array2DReal In(const_cast<double *>(VecEdge), R->Mesh.NEdgesSize, R->K), Out(ReconEdge, R->Mesh.NEdgesSize, R->K);
   parallelFor({R->Mesh.NEdgesAll, R->K / VecLength}, KOKKOS_LAMBDA(int I, int KC) { Op(Out, I, KC, In); });
   return 0;
}

} // extern "C"
