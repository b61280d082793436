// ref_vertical.cpp -- TEST INFRASTRUCTURE.  The reference's equation-of-state functors (src/ocn/Eos.h: Teos10Eos,
// LinearEos) compiled unmodified where they lie and called on plain arrays; the constructors are the two lines of
// src/ocn/Eos.cpp that allocate the coefficient scratch (Eos.cpp is not compiled: it needs Config / Field / IO).
#include "Eos.h"

namespace OMEGA {
Teos10Eos::Teos10Eos(int NVertLayers) : NVertLayers(NVertLayers) { SpecVolPCoeffs = Array2DReal("SpecVolPCoeffs", 6, VecLength); }
LinearEos::LinearEos() {}
} // namespace OMEGA

extern "C" void ref_eos_specvol(int NCells, int K, int EosType, int KDisp, const double *Ct, const double *Sa, const double *P,
                                double *SpecVol) {
   using namespace OMEGA;
   Array2DReal T("t", NCells, K), S("s", NCells, K), Pr("p", NCells, K), V("v", NCells, K);
   for (int i = 0; i < NCells; ++i)
      for (int k = 0; k < K; ++k) { T(i, k) = Ct[i * K + k]; S(i, k) = Sa[i * K + k]; Pr(i, k) = P[i * K + k]; }
   Teos10Eos Teos(K);
   LinearEos Lin;
   for (int i = 0; i < NCells; ++i)
      for (int kc = 0; kc < K / VecLength; ++kc) {
         if (EosType == 1) Teos(V, i, kc, T, S, Pr, KDisp);
         else Lin(V, i, kc, T, S);
      }
   for (int i = 0; i < NCells; ++i)
      for (int k = 0; k < K; ++k) SpecVol[i * K + k] = V(i, k);
}
