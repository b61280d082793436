// stepper.h -- the object behind OmegaStepper*: config, aux/tendency arrays
// (owned, as AuxiliaryState and Tendencies own theirs) and the provisional
// state of the Runge-Kutta steppers.
#pragma once
#include "omega_dev.h"

struct OmegaHalo;

struct ProvisArrays {
   omega::Real *H[2], *U[2], *Tr[2];
};

struct OmegaStepper {
   const OmegaMesh *M;
   OmegaTendConfig Cfg;
   int NTracers;
   omega::Real BottomDepth0;
   omega::AuxPtrs A;
   OmegaTendArrays T;
   ProvisArrays P;
   omega::Real *SrcH, *SrcU; // manufactured sources per cell / edge of the current stage (fused tiled path)
   omega::Real *RTop, *RBot; // wind / bottom-drag tendency of the top / bottom layer per edge (k_edge_layer_terms)
   long long Launches;
   int ProvisExchangeSlot; // which P.*[i] holds the stage-2 provisional state to exchange
   int LastPath;           // OMEGA_PATH_* bits of the kernels the last step phase ran
   struct OmegaHalo *Halo; // attached halo (omega_stepper_attach_halo): boundary tiles first, halo tiles skipped
   int SplitRecorded;      // the last phase recorded Halo->EvReady after its boundary tiles
   // host custom tendencies (Tendencies.h:51-53): called after the built-in terms of the kernel-per-term path
   OmegaCustomTendency CustomThick = nullptr, CustomVel = nullptr;
   void *CustomUser = nullptr;
   std::vector<void *> Allocs;
   // optional per-kernel CUDA-event timing of the fused pipeline (bench roofline)
   int Profiling;
   struct Timed { int Kind; cudaEvent_t Start, Stop; };
   std::vector<Timed> Events;
};

namespace omega {
enum FusedKernel { kFA = 0, kFB = 1, kFT2 = 2, kFC = 3, kFE = 4, kNumFused = 5 };
int stepper_alloc(OmegaStepper *S, omega::Real **p, size_t n);
ManufacturedParams manufactured(const OmegaStepper *S, double Time);
int ensure_htr(OmegaStepper *S);
// fused.cu: one phase of RK4 / RK2 / forward-backward on the fused per-stage pipeline
int fused_phase(OmegaStepper *S, OmegaStepperType Type, OmegaStateArrays *X, const OmegaStepCoeffs *C, int Phase,
                int Compat, cudaStream_t Stream);
} // namespace omega
