/* vertical_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C, sequential restatement of the Omega-1 column pieces the library offers next to the horizontal
 * hot path, host order arrays [cell][level], used only by tests/ as the checker.
 * Parity pin:
 *   (1) Eos: bit for bit against the reference's own functors (src/ocn/Eos.h compiled unmodified into
 *       oracle/_ref/libomega_ref.so, oracle/ref_shim/ref_vertical.cpp) and against the published check values
 *       of test/ocn/EosTest.cpp:39-47 (TEOS-10: 0.0009732819628 at SA = 30 g/kg, CT = 10 degC, P = 1000 dbar;
 *       linear: 0.0009784735812133072) -- tests/test_vertical.py;
 *   (2) VertCoord: the reference file needs Kokkos team scans (not in the host stand-in), so its four loops are
 *       restated from src/ocn/VertCoord.cpp:648-860 and checked against closed forms the way
 *       test/ocn/VertCoordTest.cpp does (uniform columns); "parity unpinned" against the reference binary;
 *   (3) tridiagonal solvers: Thomas algorithm restated from src/base/TriDiagSolvers.h:44-75 (general) and :215-265
 *       (diffusion form), checked against a dense LU solve and the stability scenario of
 *       test/base/TriDiagSolversTest.cpp:428-470.
 * Operation order follows the cited lines; build with -ffp-contract=off.
 */
#include <math.h>
#include <stddef.h>

/* VertCoord.cpp:652-690 */
void oracle_vert_pressure(int NCells, int K, const int *KMin, const int *KMax, double Rho0, const double *H,
                          const double *SurfP, double *PInt, double *PMid) {
   const double Gravity = 9.80616;
   for (int i = 0; i < NCells; ++i) {
      PInt[(size_t)i * (K + 1) + KMin[i]] = SurfP[i];
      double Accum = 0.0;
      for (int k = KMin[i]; k <= KMax[i]; ++k) {
         const double Inc = Gravity * Rho0 * H[(size_t)i * K + k];
         Accum += Inc;
         PInt[(size_t)i * (K + 1) + k + 1] = SurfP[i] + Accum;
         PMid[(size_t)i * K + k]           = SurfP[i] + Accum - 0.5 * Inc;
      }
   }
}

/* VertCoord.cpp:697-735 */
void oracle_vert_zheight(int NCells, int K, const int *KMin, const int *KMax, double Rho0, const double *BottomDepth,
                         const double *H, const double *SpecVol, double *ZInt, double *ZMid) {
   for (int i = 0; i < NCells; ++i) {
      ZInt[(size_t)i * (K + 1) + KMax[i] + 1] = -BottomDepth[i];
      double Accum = 0.0;
      for (int k = KMax[i]; k >= KMin[i]; --k) {
         const double DZ = Rho0 * SpecVol[(size_t)i * K + k] * H[(size_t)i * K + k];
         Accum += DZ;
         ZInt[(size_t)i * (K + 1) + k] = -BottomDepth[i] + Accum;
         ZMid[(size_t)i * K + k]       = -BottomDepth[i] + Accum - 0.5 * DZ;
      }
   }
}

/* VertCoord.cpp:742-778 */
void oracle_vert_geopotential(int NCells, int K, const int *KMin, const int *KMax, const double *ZMid, const double *Tidal,
                              const double *Sal, double *Geop) {
   const double Gravity = 9.80616;
   for (int i = 0; i < NCells; ++i)
      for (int k = KMin[i]; k <= KMax[i]; ++k)
         Geop[(size_t)i * K + k] = Gravity * ZMid[(size_t)i * K + k] + Tidal[i] + Sal[i];
}

/* VertCoord.cpp:786-851 */
void oracle_vert_target(int NCells, int K, const int *KMin, const int *KMax, double Rho0, const double *PInt,
                        const double *RefH, const double *Wgt, double *Target) {
   const double Gravity = 9.80616;
   for (int i = 0; i < NCells; ++i) {
      double Coeff = (PInt[(size_t)i * (K + 1) + KMax[i] + 1] - PInt[(size_t)i * (K + 1) + KMin[i]]) / (Gravity * Rho0);
      double SumWh = 0, SumRefH = 0;
      for (int k = KMin[i]; k <= KMax[i]; ++k) {
         SumWh += Wgt[k] * RefH[(size_t)i * K + k];
         SumRefH += RefH[(size_t)i * K + k];
      }
      Coeff -= SumRefH;
      for (int k = KMin[i]; k <= KMax[i]; ++k)
         Target[(size_t)i * K + k] = RefH[(size_t)i * K + k] * (1.0 + Coeff * Wgt[k] / SumWh);
   }
}

/* Eos.h:51-222: Roquet et al. (2015) specific volume, coefficient c(k, j, i) multiplies s^i t^j p^k */
static const double V[6][7][7] = {
    {{1.0769995862e-03, -3.1038981976e-04, 6.6928067038e-04, -8.5047933937e-04, 5.8086069943e-04, -2.1092370507e-04, 3.1932457305e-05},
     {-1.5649734675e-05, 3.5009599764e-05, -4.3592678561e-05, 3.4532461828e-05, -1.1959409788e-05, 1.3864594581e-06, 0},
     {2.7762106484e-05, -3.7435842344e-05, 3.5907822760e-05, -1.8698584187e-05, 3.8595339244e-06, 0, 0},
     {-1.6521159259e-05, 2.4141479483e-05, -1.4353633048e-05, 2.2863324556e-06, 0, 0, 0},
     {6.9111322702e-06, -8.7595873154e-06, 4.3703680598e-06, 0, 0, 0, 0},
     {-8.0539615540e-07, -3.3052758900e-07, 0, 0, 0, 0, 0},
     {2.0543094268e-07, 0, 0, 0, 0, 0, 0}},
    {{-1.6784136540e-05, 2.4262468747e-05, -3.4792460974e-05, 3.7470777305e-05, -1.7322218612e-05, 3.0927427253e-06, 0},
     {1.8505765429e-05, -9.5677088156e-06, 1.1100834765e-05, -9.8447117844e-06, 2.5909225260e-06, 0, 0},
     {-1.1716606853e-05, -2.3678308361e-07, 2.9283346295e-06, -4.8826139200e-07, 0, 0, 0},
     {7.9279656173e-06, -3.4558773655e-06, 3.1655306078e-07, 0, 0, 0, 0},
     {-3.4102187482e-06, 1.2956717783e-06, 0, 0, 0, 0, 0},
     {5.0736766814e-07, 0, 0, 0, 0, 0, 0},
     {0, 0, 0, 0, 0, 0, 0}},
    {{3.0623833435e-06, -5.8484432984e-07, -4.8122251597e-06, 4.9263106998e-06, -1.7811974727e-06, 0, 0},
     {-1.1736386731e-06, -5.5699154557e-06, 5.4620748834e-06, -1.3544185627e-06, 0, 0, 0},
     {2.1305028740e-06, 3.9137387080e-07, -6.5731104067e-07, 0, 0, 0, 0},
     {-4.6132540037e-07, 7.7618888092e-09, 0, 0, 0, 0, 0},
     {-6.3352916514e-08, 0, 0, 0, 0, 0, 0},
     {0, 0, 0, 0, 0, 0, 0},
     {0, 0, 0, 0, 0, 0, 0}},
    {{-3.8088938393e-07, 3.6310188515e-07, 1.6746303780e-08, 0, 0, 0, 0},
     {-3.6527006553e-07, -2.7295696237e-07, 0, 0, 0, 0, 0},
     {2.8695905159e-07, 0, 0, 0, 0, 0, 0},
     {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}},
    {{8.8302421514e-08, -1.1147125423e-07, 0, 0, 0, 0, 0},
     {3.1454099902e-07, 0, 0, 0, 0, 0, 0},
     {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}},
    {{4.2369007180e-09, 0, 0, 0, 0, 0, 0},
     {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}}};
static const double V0[6] = {-4.4015007269e-05, 6.9232335784e-06, -7.5004675975e-07, 1.7009109288e-08, -1.6884162004e-08, 1.9613503930e-09};

double oracle_teos10_specvol(double Sa, double Ct, double P) {
   const double SAu = 40.0 * 35.16504 / 35.0;
   const double Ss = sqrt((Sa + 24.0) / SAu), Tt = Ct / 40.0, Pp = P / 1e4;
   double Delta = 0.0;
   for (int k = 5; k >= 0; --k) {          /* calcDelta: Horner in Pp over the six pressure coefficients */
      /* calcPCoeffs, in the reference's association: Horner in Tt; at each step the Ss-part of the coefficient is
       * added first and its constant last, (acc * Tt + (Horner of the s^1.. terms) * Ss) + V(0, j, k) */
      const int jmax = 6 - k;
      double Ck = V[k][jmax][0];
      for (int j = jmax - 1; j >= 0; --j) {
         const int n = 6 - k - j;
         double Hs = V[k][j][n];
         for (int i = n - 1; i >= 1; --i) Hs = Hs * Ss + V[k][j][i];
         Ck = Ck * Tt + Hs * Ss + V[k][j][0];
      }
      Delta = Delta * Pp + Ck;
   }
   double Ref = 0.0;                       /* calcRefProfile */
   for (int k = 5; k >= 0; --k) Ref = Ref * Pp + V0[k];
   return Ref * Pp + Delta;
}

/* Eos.h:29-48 (TEOS-10, optional displacement), :229-247 (linear) */
void oracle_eos_specvol(int NCells, int K, int EosType, int KDisp, double DRhodT, double DRhodS, double RhoT0S0,
                        const double *Ct, const double *Sa, const double *P, double *SpecVol) {
   for (int i = 0; i < NCells; ++i)
      for (int k = 0; k < K; ++k) {
         const size_t o = (size_t)i * K + k;
         if (EosType == 0) {
            SpecVol[o] = 1.0 / (RhoT0S0 + (DRhodT * Ct[o] + DRhodS * Sa[o]));
         } else {
            int kp = k;
            if (KDisp != 0) {
               kp = k + KDisp < K - 1 ? k + KDisp : K - 1;
               kp = kp > 0 ? kp : 0;
            }
            SpecVol[o] = oracle_teos10_specvol(Sa[o], Ct[o], P[(size_t)i * K + kp]);
         }
      }
}

/* TriDiagSolvers.h:44-75 (ThomasSolver::solve), arrays [batch][row]; D and X are overwritten */
void oracle_tridiag_thomas(int NBatch, int NRow, const double *DL, double *D, const double *DU, double *X) {
   for (int b = 0; b < NBatch; ++b) {
      const double *dl = DL + (size_t)b * NRow, *du = DU + (size_t)b * NRow;
      double *d = D + (size_t)b * NRow, *x = X + (size_t)b * NRow;
      for (int k = 1; k < NRow; ++k) {
         const double W = dl[k] / d[k - 1];
         d[k] -= W * du[k - 1];
         x[k] -= W * x[k - 1];
      }
      x[NRow - 1] /= d[NRow - 1];
      for (int k = NRow - 2; k >= 0; --k) x[k] = (x[k] - du[k] * x[k + 1]) / d[k];
   }
}

/* TriDiagSolvers.h:215-265 (ThomasDiffusionSolver::solve): D(k) = H(k) + G(k-1) + G(k) is never formed; H and X are
 * overwritten, Alpha is scratch [NRow] */
void oracle_tridiag_diffusion_thomas(int NBatch, int NRow, const double *G, double *H, double *X, double *Alpha) {
   for (int b = 0; b < NBatch; ++b) {
      const double *g = G + (size_t)b * NRow;
      double *h = H + (size_t)b * NRow, *x = X + (size_t)b * NRow;
      Alpha[0] = 0;
      for (int k = 1; k < NRow; ++k)
         Alpha[k] = g[k - 1] * (h[k - 1] + Alpha[k - 1]) / (h[k - 1] + Alpha[k - 1] + g[k - 1]);
      h[0] += g[0];
      for (int k = 1; k < NRow; ++k) {
         const double AddH = Alpha[k] + g[k];
         h[k] += AddH;
         x[k] += g[k - 1] / h[k - 1] * x[k - 1];
      }
      x[NRow - 1] /= h[NRow - 1];
      for (int k = NRow - 2; k >= 0; --k) x[k] = (x[k] + g[k] * x[k + 1]) / h[k];
   }
}
