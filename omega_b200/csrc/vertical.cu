// vertical.cu -- the column-local pieces Omega-1 adds around the horizontal hot path:
//   VertCoord::computePressure / computeZHeight / computeGeopotential / computeTargetThickness
//       (components/omega/src/ocn/VertCoord.cpp:648-860),
//   Eos::computeSpecVol / computeSpecVolDisp: TEOS-10 75-term polynomial of Roquet et al. (2015) and the
//       linear equation of state (src/ocn/Eos.h:29-292, Eos.cpp),
//   TriDiagSolver / TriDiagDiffSolver (src/base/TriDiagSolvers.h): batched tridiagonal systems, one per column.
// Every array is [cell][level] with level innermost, as everywhere in this library, so a WARP owns a column:
// lanes run over levels (coalesced rows), the prefix sums of the reference's per-team parallel_scan become
// warp shuffles with a carry between 32-level chunks, and the tridiagonal systems are solved by parallel
// cyclic reduction in shared memory with one thread per row (the reference's device choice, PCRSolver).
// Sums are associated differently from the reference's sequential scans (tests hold 1e-12).
#include <cmath>

#include "kernel_util.cuh"
#include "omega_dev.h"

namespace omega {

constexpr Real kGravityVert = 9.80616; // VertCoord.cpp:659 (not the 9.80665 of the tendency terms)

__device__ __forceinline__ Real warp_scan_incl(Real v, int lane) {
#pragma unroll
   for (int d = 1; d < 32; d <<= 1) {
      const Real t = __shfl_up_sync(0xffffffffu, v, d);
      if (lane >= d) v += t;
   }
   return v;
}
__device__ __forceinline__ Real warp_sum(Real v) {
#pragma unroll
   for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
   return v;
}

// PressureInterface(i, KMin) = SurfacePressure(i); going down: + Gravity * Rho0 * h(i, k)  (VertCoord.cpp:652-690)
__global__ void __launch_bounds__(kBlock)
k_vert_pressure(int NCells, int K, const int *__restrict__ KMinC, const int *__restrict__ KMaxC, double Rho0,
                const Real *__restrict__ H, const Real *__restrict__ SurfP, Real *__restrict__ PInt,
                Real *__restrict__ PMid) {
   const int cell = (int)((blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
   if (cell >= NCells) return;
   const int kmin = KMinC[cell], kmax = KMaxC[cell];
   const Real p0 = SurfP[cell];
   if (lane == 0) PInt[(size_t)cell * (K + 1) + kmin] = p0;
   Real carry = 0.0;
   for (int k0 = kmin; k0 <= kmax; k0 += 32) {
      const int k = k0 + lane;
      const Real inc = k <= kmax ? kGravityVert * Rho0 * H[(size_t)cell * K + k] : 0.0;
      const Real acc = carry + warp_scan_incl(inc, lane);
      if (k <= kmax) {
         PInt[(size_t)cell * (K + 1) + k + 1] = p0 + acc;
         PMid[(size_t)cell * K + k]           = p0 + acc - 0.5 * inc;
      }
      carry = __shfl_sync(0xffffffffu, acc, 31);
   }
}

// ZInterface(i, KMax+1) = -BottomDepth(i); going up: + Rho0 * SpecVol * h  (VertCoord.cpp:697-735)
__global__ void __launch_bounds__(kBlock)
k_vert_zheight(int NCells, int K, const int *__restrict__ KMinC, const int *__restrict__ KMaxC, double Rho0,
               const Real *__restrict__ BottomDepth, const Real *__restrict__ H, const Real *__restrict__ SpecVol,
               Real *__restrict__ ZInt, Real *__restrict__ ZMid) {
   const int cell = (int)((blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
   if (cell >= NCells) return;
   const int kmin = KMinC[cell], kmax = KMaxC[cell];
   const Real zb = -BottomDepth[cell];
   if (lane == 0) ZInt[(size_t)cell * (K + 1) + kmax + 1] = zb;
   Real carry = 0.0;
   for (int k0 = kmax; k0 >= kmin; k0 -= 32) {
      const int k = k0 - lane;
      const size_t o = (size_t)cell * K + (k >= kmin ? k : kmin);
      const Real dz = k >= kmin ? Rho0 * SpecVol[o] * H[o] : 0.0;
      const Real acc = carry + warp_scan_incl(dz, lane);
      if (k >= kmin) {
         ZInt[(size_t)cell * (K + 1) + k] = zb + acc;
         ZMid[o]                          = zb + acc - 0.5 * dz;
      }
      carry = __shfl_sync(0xffffffffu, acc, 31);
   }
}

// GeopotentialMid = Gravity * ZMid + TidalPotential + SelfAttractionLoading on the active layers (VertCoord.cpp:742-778)
__global__ void __launch_bounds__(kBlock)
k_vert_geopotential(int NCells, int K, const int *__restrict__ KMinC, const int *__restrict__ KMaxC,
                    const Real *__restrict__ ZMid, const Real *__restrict__ Tidal, const Real *__restrict__ Sal,
                    Real *__restrict__ Geop) {
   const long long w = blockIdx.x * (long long)blockDim.x + threadIdx.x;
   if (w >= (long long)NCells * K) return;
   const int cell = (int)(w / K), k = (int)(w - (long long)cell * K);
   if (k < KMinC[cell] || k > KMaxC[cell]) return;
   Geop[w] = kGravityVert * ZMid[w] + Tidal[cell] + Sal[cell];
}

// LayerThicknessTarget = RefLayerThick * (1 + Coeff * W(k) / sum_k W(k) RefLayerThick), Coeff = column pseudo-height
// minus the reference column height (VertCoord.cpp:786-851)
__global__ void __launch_bounds__(kBlock)
k_vert_target(int NCells, int K, const int *__restrict__ KMinC, const int *__restrict__ KMaxC, double Rho0,
              const Real *__restrict__ PInt, const Real *__restrict__ RefH, const Real *__restrict__ Wgt,
              Real *__restrict__ Target) {
   const int cell = (int)((blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
   if (cell >= NCells) return;
   const int kmin = KMinC[cell], kmax = KMaxC[cell];
   Real swh = 0.0, sref = 0.0;
   for (int k = kmin + lane; k <= kmax; k += 32) {
      const Real r = RefH[(size_t)cell * K + k];
      swh += Wgt[k] * r;
      sref += r;
   }
   swh = warp_sum(swh);
   sref = warp_sum(sref);
   const Real coeff = (PInt[(size_t)cell * (K + 1) + kmax + 1] - PInt[(size_t)cell * (K + 1) + kmin]) / (kGravityVert * Rho0) - sref;
   for (int k = kmin + lane; k <= kmax; k += 32)
      Target[(size_t)cell * K + k] = RefH[(size_t)cell * K + k] * (1.0 + coeff * Wgt[k] / swh);
}

// ---- equation of state -------------------------------------------------------------------------------
// Specific volume of Roquet et al. (2015), "Accurate polynomial expressions for the density and specific volume of
// seawater using the TEOS-10 standard", Ocean Modelling 90: v = v0(p) + delta(s, t, p) with the reduced variables
// s = sqrt((SA + 24) / SAu), t = CT / 40, p = P / 1e4; delta = sum V_ijk s^i t^j p^k over i + j + k <= 6.
// The 75 published coefficients V_ijk, stored [k][j][i] (missing entries 0) and evaluated by nested Horner in the
// same order as Eos.h:51-189 (s innermost, then t, then p).
__constant__ Real cTeosV[6][7][7] = {
    // k = 0
    {{1.0769995862e-03, -3.1038981976e-04, 6.6928067038e-04, -8.5047933937e-04, 5.8086069943e-04, -2.1092370507e-04, 3.1932457305e-05},
     {-1.5649734675e-05, 3.5009599764e-05, -4.3592678561e-05, 3.4532461828e-05, -1.1959409788e-05, 1.3864594581e-06, 0},
     {2.7762106484e-05, -3.7435842344e-05, 3.5907822760e-05, -1.8698584187e-05, 3.8595339244e-06, 0, 0},
     {-1.6521159259e-05, 2.4141479483e-05, -1.4353633048e-05, 2.2863324556e-06, 0, 0, 0},
     {6.9111322702e-06, -8.7595873154e-06, 4.3703680598e-06, 0, 0, 0, 0},
     {-8.0539615540e-07, -3.3052758900e-07, 0, 0, 0, 0, 0},
     {2.0543094268e-07, 0, 0, 0, 0, 0, 0}},
    // k = 1
    {{-1.6784136540e-05, 2.4262468747e-05, -3.4792460974e-05, 3.7470777305e-05, -1.7322218612e-05, 3.0927427253e-06, 0},
     {1.8505765429e-05, -9.5677088156e-06, 1.1100834765e-05, -9.8447117844e-06, 2.5909225260e-06, 0, 0},
     {-1.1716606853e-05, -2.3678308361e-07, 2.9283346295e-06, -4.8826139200e-07, 0, 0, 0},
     {7.9279656173e-06, -3.4558773655e-06, 3.1655306078e-07, 0, 0, 0, 0},
     {-3.4102187482e-06, 1.2956717783e-06, 0, 0, 0, 0, 0},
     {5.0736766814e-07, 0, 0, 0, 0, 0, 0},
     {0, 0, 0, 0, 0, 0, 0}},
    // k = 2
    {{3.0623833435e-06, -5.8484432984e-07, -4.8122251597e-06, 4.9263106998e-06, -1.7811974727e-06, 0, 0},
     {-1.1736386731e-06, -5.5699154557e-06, 5.4620748834e-06, -1.3544185627e-06, 0, 0, 0},
     {2.1305028740e-06, 3.9137387080e-07, -6.5731104067e-07, 0, 0, 0, 0},
     {-4.6132540037e-07, 7.7618888092e-09, 0, 0, 0, 0, 0},
     {-6.3352916514e-08, 0, 0, 0, 0, 0, 0},
     {0, 0, 0, 0, 0, 0, 0},
     {0, 0, 0, 0, 0, 0, 0}},
    // k = 3
    {{-3.8088938393e-07, 3.6310188515e-07, 1.6746303780e-08, 0, 0, 0, 0},
     {-3.6527006553e-07, -2.7295696237e-07, 0, 0, 0, 0, 0},
     {2.8695905159e-07, 0, 0, 0, 0, 0, 0},
     {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}},
    // k = 4
    {{8.8302421514e-08, -1.1147125423e-07, 0, 0, 0, 0, 0},
     {3.1454099902e-07, 0, 0, 0, 0, 0, 0},
     {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}},
    // k = 5
    {{4.2369007180e-09, 0, 0, 0, 0, 0, 0},
     {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 0, 0, 0}}};
// reference profile v0(p) = p (V00 + p (V01 + ...)), Eos.h:207-220
__constant__ Real cTeosV0[6] = {-4.4015007269e-05, 6.9232335784e-06, -7.5004675975e-07, 1.7009109288e-08, -1.6884162004e-08, 1.9613503930e-09};

__device__ __forceinline__ Real teos10_specvol(Real Sa, Real Ct, Real P) {
   const Real ss = sqrt((Sa + 24.0) / (40.0 * 35.16504 / 35.0)), tt = Ct / 40.0, pp = P / 1.0e4;
   // pressure coefficient k: polynomial in (s, t) of total degree 6 - k; Horner in t, and per step the s-part of
   // the coefficient first and its constant last -- the association of Eos.h:141-189
   Real delta = 0.0;
#pragma unroll
   for (int k = 5; k >= 0; --k) {
      Real ck = cTeosV[k][6 - k][0];
#pragma unroll
      for (int j = 5 - k; j >= 0; --j) {
         const int n = 6 - k - j;
         Real hs = cTeosV[k][j][n];
#pragma unroll
         for (int i = n - 1; i >= 1; --i) hs = hs * ss + cTeosV[k][j][i];
         ck = ck * tt + hs * ss + cTeosV[k][j][0];
      }
      delta = delta * pp + ck;
   }
   Real v0 = 0.0;
#pragma unroll
   for (int k = 5; k >= 0; --k) v0 = v0 * pp + cTeosV0[k];
   return v0 * pp + delta;
}

// EosType 0: linear (Eos.h:229-247), 1: TEOS-10.  KDisp != 0: the parcel is displaced adiabatically to the pressure
// of layer clamp(k + KDisp) (Eos.h:36-48, TEOS-10 only; the linear form does not depend on pressure)
__global__ void __launch_bounds__(kBlock)
k_eos_specvol(int NCells, int K, int EosType, int KDisp, double DRhodT, double DRhodS, double RhoT0S0,
              const Real *__restrict__ Ct, const Real *__restrict__ Sa, const Real *__restrict__ P,
              Real *__restrict__ SpecVol) {
   const long long w = blockIdx.x * (long long)blockDim.x + threadIdx.x;
   if (w >= (long long)NCells * K) return;
   if (EosType == 0) {
      SpecVol[w] = 1.0 / (RhoT0S0 + (DRhodT * Ct[w] + DRhodS * Sa[w]));
      return;
   }
   const int cell = (int)(w / K), k = (int)(w - (long long)cell * K);
   int kp = k;
   if (KDisp != 0) kp = max(0, min(k + KDisp, K - 1));
   SpecVol[w] = teos10_specvol(Sa[w], Ct[w], P[(size_t)cell * K + kp]);
}

// ---- batched tridiagonal systems: parallel cyclic reduction, one CTA per system, one thread per row ----------
// DL(k) x(k-1) + D(k) x(k) + DU(k) x(k+1) = X(k).  Each reduction step eliminates the couplings at distance s and
// doubles s; after ceil(log2 NRow) steps every row stands alone.  Diffusion form (TriDiagSolvers.h:210-330): the
// system of an implicit vertical diffusion step, DL(k) = -G(k-1), DU(k) = -G(k), D(k) = H(k) + G(k-1) + G(k).
__global__ void k_tridiag_pcr(int NRow, const Real *__restrict__ DL, const Real *__restrict__ D,
                              const Real *__restrict__ DU, Real *__restrict__ X) {
   extern __shared__ Real sm[];
   Real *a = sm, *b = sm + NRow, *c = sm + 2 * NRow, *d = sm + 3 * NRow;
   const int k = threadIdx.x;
   const size_t base = (size_t)blockIdx.x * NRow;
   if (k < NRow) {
      a[k] = k > 0 ? DL[base + k] : 0.0;
      b[k] = D[base + k];
      c[k] = k + 1 < NRow ? DU[base + k] : 0.0;
      d[k] = X[base + k];
   }
   __syncthreads();
   for (int s = 1; s < NRow; s <<= 1) {
      Real na = 0, nb = 0, nc = 0, nd = 0;
      if (k < NRow) {
         const int km = k - s, kp = k + s;
         const Real al = km >= 0 ? -a[k] / b[km] : 0.0, ga = kp < NRow ? -c[k] / b[kp] : 0.0;
         nb = b[k] + (km >= 0 ? al * c[km] : 0.0) + (kp < NRow ? ga * a[kp] : 0.0);
         nd = d[k] + (km >= 0 ? al * d[km] : 0.0) + (kp < NRow ? ga * d[kp] : 0.0);
         na = km >= 0 ? al * a[km] : 0.0;
         nc = kp < NRow ? ga * c[kp] : 0.0;
      }
      __syncthreads();
      if (k < NRow) { a[k] = na; b[k] = nb; c[k] = nc; d[k] = nd; }
      __syncthreads();
   }
   if (k < NRow) X[base + k] = d[k] / b[k];
}

// The diffusion form keeps the diagonal as H plus its two couplings, D(k) = H(k) + gm(k) + gp(k), through every
// reduction step: eliminating the neighbours at distance s gives
//    H'(k) = H(k) + al H(k-s) + be H(k+s),  gm'(k) = al gm(k-s),  gp'(k) = be gp(k+s),  al = gm(k) / D(k-s), be = gp(k) / D(k+s)
// (no subtraction anywhere), which is what keeps an abrupt 1e14 jump in diffusivity finite where the general form
// produces NaNs (test/base/TriDiagSolversTest.cpp:428-470).
__global__ void k_tridiag_diff_pcr(int NRow, const Real *__restrict__ G, const Real *__restrict__ H, Real *__restrict__ X) {
   extern __shared__ Real sm[];
   Real *h = sm, *gm = sm + NRow, *gp = sm + 2 * NRow, *x = sm + 3 * NRow;
   const int k = threadIdx.x;
   const size_t base = (size_t)blockIdx.x * NRow;
   if (k < NRow) {
      h[k]  = H[base + k];
      gm[k] = k > 0 ? G[base + k - 1] : 0.0;
      gp[k] = k + 1 < NRow ? G[base + k] : 0.0;
      x[k]  = X[base + k];
   }
   __syncthreads();
   for (int s = 1; s < NRow; s <<= 1) {
      Real nh = 0, ngm = 0, ngp = 0, nx = 0;
      if (k < NRow) {
         const int km = k - s, kp = k + s;
         const Real al = km >= 0 ? gm[k] / (h[km] + gm[km] + gp[km]) : 0.0;
         const Real be = kp < NRow ? gp[k] / (h[kp] + gm[kp] + gp[kp]) : 0.0;
         nh  = h[k] + (km >= 0 ? al * h[km] : 0.0) + (kp < NRow ? be * h[kp] : 0.0);
         nx  = x[k] + (km >= 0 ? al * x[km] : 0.0) + (kp < NRow ? be * x[kp] : 0.0);
         ngm = km >= 0 ? al * gm[km] : 0.0;
         ngp = kp < NRow ? be * gp[kp] : 0.0;
      }
      __syncthreads();
      if (k < NRow) { h[k] = nh; gm[k] = ngm; gp[k] = ngp; x[k] = nx; }
      __syncthreads();
   }
   if (k < NRow) X[base + k] = x[k] / (h[k] + gm[k] + gp[k]);
}

} // namespace omega

using namespace omega;

extern "C" {

int omega_vert_compute_pressure(const OmegaMesh *M, const int *MinLayerCell, const int *MaxLayerCell, double Rho0,
                                const Real *LayerThickness, const Real *SurfacePressure, Real *PressureInterface,
                                Real *PressureMid, void *Stream) {
   OMEGA_REQUIRE(M && MinLayerCell && MaxLayerCell && LayerThickness && SurfacePressure && PressureInterface && PressureMid,
                 "omega_vert_compute_pressure: null argument");
   const int n = M->D.NCellsAll;
   if (n) k_vert_pressure<<<grid_for((long long)n * 32), kBlock, 0, (cudaStream_t)Stream>>>(
       n, M->D.K, MinLayerCell, MaxLayerCell, Rho0, LayerThickness, SurfacePressure, PressureInterface, PressureMid);
   OMEGA_CUDA_CHECK(cudaGetLastError());
   return 0;
}

int omega_vert_compute_z_height(const OmegaMesh *M, const int *MinLayerCell, const int *MaxLayerCell, double Rho0,
                                const Real *LayerThickness, const Real *SpecVol, Real *ZInterface, Real *ZMid,
                                void *Stream) {
   OMEGA_REQUIRE(M && MinLayerCell && MaxLayerCell && LayerThickness && SpecVol && ZInterface && ZMid,
                 "omega_vert_compute_z_height: null argument");
   const int n = M->D.NCellsAll;
   if (n) k_vert_zheight<<<grid_for((long long)n * 32), kBlock, 0, (cudaStream_t)Stream>>>(
       n, M->D.K, MinLayerCell, MaxLayerCell, Rho0, M->D.BottomDepth, LayerThickness, SpecVol, ZInterface, ZMid);
   OMEGA_CUDA_CHECK(cudaGetLastError());
   return 0;
}

int omega_vert_compute_geopotential(const OmegaMesh *M, const int *MinLayerCell, const int *MaxLayerCell, const Real *ZMid,
                                    const Real *TidalPotential, const Real *SelfAttractionLoading,
                                    Real *GeopotentialMid, void *Stream) {
   OMEGA_REQUIRE(M && MinLayerCell && MaxLayerCell && ZMid && TidalPotential && SelfAttractionLoading && GeopotentialMid,
                 "omega_vert_compute_geopotential: null argument");
   const long long n = (long long)M->D.NCellsAll * M->D.K;
   if (n) k_vert_geopotential<<<grid_for(n), kBlock, 0, (cudaStream_t)Stream>>>(
       M->D.NCellsAll, M->D.K, MinLayerCell, MaxLayerCell, ZMid, TidalPotential, SelfAttractionLoading, GeopotentialMid);
   OMEGA_CUDA_CHECK(cudaGetLastError());
   return 0;
}

int omega_vert_compute_target_thickness(const OmegaMesh *M, const int *MinLayerCell, const int *MaxLayerCell, double Rho0,
                                        const Real *PressureInterface, const Real *RefLayerThickness,
                                        const Real *MovementWeights, Real *LayerThicknessTarget, void *Stream) {
   OMEGA_REQUIRE(M && MinLayerCell && MaxLayerCell && PressureInterface && RefLayerThickness && MovementWeights &&
                     LayerThicknessTarget, "omega_vert_compute_target_thickness: null argument");
   const int n = M->D.NCellsAll;
   if (n) k_vert_target<<<grid_for((long long)n * 32), kBlock, 0, (cudaStream_t)Stream>>>(
       n, M->D.K, MinLayerCell, MaxLayerCell, Rho0, PressureInterface, RefLayerThickness, MovementWeights, LayerThicknessTarget);
   OMEGA_CUDA_CHECK(cudaGetLastError());
   return 0;
}

int omega_eos_compute_spec_vol(const OmegaMesh *M, int EosType, const Real *ConservTemp, const Real *AbsSalinity,
                               const Real *Pressure, int KDisp, double DRhodT, double DRhodS, double RhoT0S0,
                               Real *SpecVol, void *Stream) {
   OMEGA_REQUIRE(M && ConservTemp && AbsSalinity && SpecVol && (EosType == 0 || (EosType == 1 && Pressure)),
                 "omega_eos_compute_spec_vol: bad argument");
   const long long n = (long long)M->D.NCellsAll * M->D.K;
   if (n) k_eos_specvol<<<grid_for(n), kBlock, 0, (cudaStream_t)Stream>>>(M->D.NCellsAll, M->D.K, EosType, KDisp, DRhodT, DRhodS,
                                                                          RhoT0S0, ConservTemp, AbsSalinity, Pressure, SpecVol);
   OMEGA_CUDA_CHECK(cudaGetLastError());
   return 0;
}

int omega_tridiag_solve(int NBatch, int NRow, const Real *DL, const Real *D, const Real *DU, Real *X, void *Stream) {
   OMEGA_REQUIRE(NBatch >= 0 && NRow >= 1 && NRow <= 1024 && DL && D && DU && X, "omega_tridiag_solve: bad argument");
   if (NBatch == 0) return 0;
   const int nt = (NRow + 31) / 32 * 32;
   k_tridiag_pcr<<<NBatch, nt, 4 * NRow * sizeof(Real), (cudaStream_t)Stream>>>(NRow, DL, D, DU, X);
   OMEGA_CUDA_CHECK(cudaGetLastError());
   return 0;
}

int omega_tridiag_diffusion_solve(int NBatch, int NRow, const Real *G, const Real *H, Real *X, void *Stream) {
   OMEGA_REQUIRE(NBatch >= 0 && NRow >= 1 && NRow <= 1024 && G && H && X, "omega_tridiag_diffusion_solve: bad argument");
   if (NBatch == 0) return 0;
   const int nt = (NRow + 31) / 32 * 32;
   k_tridiag_diff_pcr<<<NBatch, nt, 4 * NRow * sizeof(Real), (cudaStream_t)Stream>>>(NRow, G, H, X);
   OMEGA_CUDA_CHECK(cudaGetLastError());
   return 0;
}

} // extern "C"
