// fused_tile.cu -- shared-memory staged cell pass of the fused RK4 pipeline.
//
// k_fCT does the work of k_fT2 + k_fC_center (fused.cu) in ONE pass and never
// materialises Del2Tracers: a CTA owns one tile (omega_dev.h: <= 64 cells and
// the two rings of neighbours around them) x 16 levels.  Rows of h and of one
// tracer after the other are gathered into shared memory with cp.async
// (LDGSTS, 16 B per lane, three tracers in flight), so
//   * every row a tile needs crosses L2 -> SM once per tile instead of once
//     per neighbour that reads it, with no index chain in front of it;
//   * the thread that owns (cell, 2 levels) keeps its per-edge coefficients in
//     registers for all tracers and addresses neighbours by precomputed
//     shared-memory offsets;
//   * Del2(phi) of the own cells AND of ring 1 is computed in shared memory
//     (ring 1 redundantly, ~0.6 extra rows per own row), which removes the
//     3 T doubles per cell-layer the Del2Tracers array costs in HBM.
// Algorithmic DRAM traffic per cell-layer and stage: 8 + 4 T doubles
// (h, hcur, hnext r/w, hprov, 3 u | phi, cur, next r/w or prov) against
// 9 + 8 T for k_fT2 + k_fC_center.
//
// Reference semantics: the same functors and update rules as k_fC_center
// (TendencyTerms.h:35-58,349-480, TracerAuxVars.h:25-91,
// LayerThicknessAuxVars.h:25-61, TimeStepper.cpp:378-524).
#include "fused_args.h"
#include "kernel_util.cuh"
#include "omega_dev.h"
#include "stepper.h"

namespace omega {

constexpr int kTKC      = 16;  // levels per tile
constexpr int kTLanes   = 8;   // threads per row (2 levels each)
constexpr int kTThreads = kTileCells * kTLanes;
constexpr int kTStages  = 3;   // tracer rows in flight

struct TileSmem {
   double H[kTileRows][kTKC];
   double Phi[kTStages][kTileRows][kTKC];
   double D2[kTileRows1][kTKC];
   double G[kTileRows1][6];
   double InvA[kTileRows1];
   unsigned short Nbl[kTileRows1][8];
   int GRow[kTileRows];
   TileHdr Hdr;
};

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async16(unsigned dst, const void *src) {
   asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ double2 lds2(const double *p) { return *reinterpret_cast<const double2 *>(p); }

template <bool HYP>
__global__ void __launch_bounds__(kTThreads, 1)
k_fCT(DMesh M, CellArgs P, int NChunks) {
   extern __shared__ __align__(16) unsigned char smem_raw[];
   TileSmem &S = *reinterpret_cast<TileSmem *>(smem_raw);
   const int tid = threadIdx.x;
   const int tile = blockIdx.x / NChunks, kc0 = (blockIdx.x % NChunks) * kTKC;
   const int K = M.K;
   const size_t CS = (size_t)M.NCellsSize * K;

   if (tid < (int)(sizeof(TileHdr) / sizeof(int)))
      reinterpret_cast<int *>(&S.Hdr)[tid] = reinterpret_cast<const int *>(M.Tiles.Hdr + tile)[tid];
   __syncthreads();
   const int c0 = S.Hdr.C0, nc = S.Hdr.NC, n1 = S.Hdr.N1, nR1 = nc + n1, nR = nR1 + S.Hdr.N2;
   for (int r = tid; r < nR; r += kTThreads) S.GRow[r] = r < nc ? c0 + r : M.Tiles.HaloRows[S.Hdr.RowOff + r - nc];
   __syncthreads();

   // ---- this thread's share of a row-set copy: items q = tid + 512 i -> row q/8, 16-byte piece q%8
   int goff[3];      // element offset of the source piece inside one [cell][level] field
   unsigned soff[3]; // byte offset of the destination piece inside one [row][16] buffer
   int ncopy = 0;
#pragma unroll
   for (int i = 0; i < 3; ++i) {
      const int q = tid + i * kTThreads, r = q >> 3, p = q & 7;
      goff[i] = 0;
      soff[i] = 0;
      if (r < nR) {
         goff[i] = S.GRow[r] * K + kc0 + 2 * p;
         soff[i] = (unsigned)(r * kTKC + 2 * p) * 8u;
         ncopy   = i + 1;
      }
   }
   auto copy_rows = [&](const double *F, double *Dst) {
      const unsigned d = smem_u32(Dst);
#pragma unroll
      for (int i = 0; i < 3; ++i)
         if (i < ncopy) cp_async16(d + soff[i], F + goff[i]);
   };
   copy_rows(P.Hs, &S.H[0][0]);
   cp_commit();
   const int T = P.NTracers;
#pragma unroll
   for (int s = 0; s < kTStages - 1; ++s) {
      if (s < T) copy_rows(P.Trs + s * CS, &S.Phi[s][0][0]);
      cp_commit();
   }

   // ---- static tables of the own + ring-1 rows
   for (int q = tid; q < nR1 * 6; q += kTThreads) {
      const int r = q / 6, j = q - r * 6;
      S.G[r][j] = M.Tiles.GCell[(size_t)S.GRow[r] * 6 + j];
   }
   for (int r = tid; r < nR1; r += kTThreads) {
      S.InvA[r] = M.Tiles.InvAreaCell[S.GRow[r]];
      *reinterpret_cast<uint4 *>(&S.Nbl[r][0]) =
          *reinterpret_cast<const uint4 *>(M.Tiles.Nbl + (size_t)(S.Hdr.NblOff + r) * 8);
   }

   // ---- own cell of this thread
   const int i = tid >> 3, p = tid & 7;
   const bool own = i < nc;
   const int c = c0 + (own ? i : 0);
   const int k = kc0 + 2 * p;
   const size_t o = (size_t)c * K + k;
   const int ne = own ? M.NEdgesOnCell[c] : 0;
   const double InvA = 1.0 / M.AreaCell[c];
   const double k2 = P.DiffOn ? P.EddyDiff2 * InvA : 0.0, k4 = HYP ? P.EddyDiff4 * InvA : 0.0;
   double us[kME][2], ut[kME][2], cfu[kME], wa[kME], cd[kME], ch[kME];
#pragma unroll
   for (int j = 0; j < kME; ++j) {
      us[j][0] = us[j][1] = ut[j][0] = ut[j][1] = 0.0;
      cfu[j] = wa[j] = cd[j] = ch[j] = 0.0;
      if (j < ne) {
         const int e       = M.EdgesOnCell[c * M.MaxEdges + j];
         const double sg   = M.EdgeSignOnCell[c * M.MaxEdges + j];
         const double mask = M.EdgeMask[e], dve = M.DvEdge[e];
         const double rdc  = dve / M.DcEdge[e];
         // s*(phi(c1)-phi(c0)) == phi(self)-phi(nbr): sign and sgn of k_fC_center cancel exactly
         cd[j]  = -k2 * (mask * (M.MeshScalingDel2[e] * rdc));
         ch[j]  = -k4 * (mask * (M.MeshScalingDel4[e] * rdc));
         cfu[j] = dve * sg;
         wa[j]  = P.AdvOn ? mask * dve * sg : 0.0;
         ldv<2>(us[j], P.Us + (size_t)e * K + k);
         if (P.UTr == P.Us) {
            ut[j][0] = us[j][0];
            ut[j][1] = us[j][1];
         } else {
            ldv<2>(ut[j], P.UTr + (size_t)e * K + k);
         }
      }
   }
   double hcur[2] = {1.0, 1.0}, hnext_old[2] = {1.0, 1.0};
   const bool hcur_is_hs = P.First && P.Hcur == P.Hs;
   if (own) {
      if (!hcur_is_hs) ldp<2>(hcur, P.Hcur + o);
      if (!P.First || P.WSel) ldp<2>(hnext_old, P.Hnext + o);
   }

   cp_wait<kTStages - 1>(); // h rows have landed (this thread's pieces)
   __syncthreads();         // ... and everybody else's; tables are visible

   unsigned nboff[kME]; // byte offset of the neighbour's piece inside a [row][16] buffer
   {
      const uint4 t = *reinterpret_cast<const uint4 *>(&S.Nbl[own ? i : 0][0]);
      const unsigned w[3] = {t.x, t.y, t.z};
#pragma unroll
      for (int j = 0; j < kME; ++j) {
         const unsigned l = (w[j >> 1] >> ((j & 1) * 16)) & 0xffffu;
         nboff[j]         = (l * kTKC + 2 * p) * 8u;
      }
   }
   const unsigned myoff = ((own ? i : 0) * kTKC + 2 * p) * 8u;
   const unsigned char *Hb = reinterpret_cast<const unsigned char *>(&S.H[0][0]);
   double E[kME][2], cfA[kME][2], S0[2] = {0, 0}, SH = 0.0, hT[2] = {0, 0};
   double hs[2];
   {
      const double2 t = lds2(reinterpret_cast<const double *>(Hb + myoff));
      hs[0] = t.x;
      hs[1] = t.y;
   }
#pragma unroll
   for (int j = 0; j < kME; ++j) {
      const double2 hn2 = lds2(reinterpret_cast<const double *>(Hb + nboff[j]));
      const double hn[2] = {hn2.x, hn2.y};
      const double g     = S.G[own ? i : 0][j];
      SH -= ch[j];
#pragma unroll
      for (int q = 0; q < 2; ++q) {
         const double mean = 0.5 * (hs[q] + hn[q]);
         hT[q] += cfu[j] * mean * us[j][q] * InvA;
         const double w = wa[j] * ut[j][q] * InvA;
         const double D = cd[j] * mean;
         E[j][q]        = 0.5 * hn[q] * w - D;
         S0[q] += 0.5 * hs[q] * w + D;
         cfA[j][q] = g * mean;
      }
   }
   if (!P.FluxDivOn) hT[0] = hT[1] = 0.0;
   if (P.MS.Enabled) {
      const double Phase = P.MS.Kx * M.XCell[c] + P.MS.Ky * M.YCell[c] - P.MS.AngFreq * P.MS.Time;
      const double src   = P.MS.Eta0 * (-P.MS.H0 * (P.MS.Kx + P.MS.Ky) * sin(Phase) - P.MS.AngFreq * cos(Phase) +
                                      P.MS.Eta0 * (P.MS.Kx + P.MS.Ky) * cos(2.0 * Phase));
      hT[0] += src;
      hT[1] += src;
   }
   // ---- thickness updates (updateThicknessByTend, TimeStepper.cpp:378-405)
   double hx[2], h2[2], rhp[2];
   {
      double hnext[2], hprov[2];
#pragma unroll
      for (int q = 0; q < 2; ++q) {
         if (hcur_is_hs) hcur[q] = hs[q];
         hnext[q] = (P.First ? hcur[q] : hnext_old[q]) + P.B * hT[q];
         hprov[q] = hcur[q] + P.A * hT[q];
         hx[q]    = P.First ? (P.WSel ? hnext_old[q] : hcur[q]) : (P.FinSel ? hcur[q] : hnext[q]);
         h2[q]    = P.H2Sel ? hnext[q] : hcur[q];
         rhp[q]   = 1.0 / hprov[q];
      }
      if (own) {
         stv<2>(P.Hnext + o, hnext);
         if (!P.Last) stv<2>(P.Hprov + o, hprov);
      }
      if (P.Last) {
         rhp[0] = 1.0 / hx[0];
         rhp[1] = 1.0 / hx[1];
      }
   }

   // ---- tracers
   const bool tcur_is_ts = P.First && P.TrCur == P.Trs;
   const bool need_cur   = !tcur_is_ts && (P.First || !P.Last);
   for (int l = 0; l < T; ++l) {
      const int buf = l % kTStages;
      cp_wait<kTStages - 2>();
      __syncthreads(); // phi_l visible; every thread is done with tracer l-1
      if (l + kTStages - 1 < T) copy_rows(P.Trs + (size_t)(l + kTStages - 1) * CS, &S.Phi[(l + kTStages - 1) % kTStages][0][0]);
      cp_commit();
      double tcur[2] = {0, 0}, acc[2] = {0, 0};
      if (own) {
         if (need_cur) ldp<2>(tcur, P.TrCur + l * CS + o);
         if (!P.First) ldp<2>(acc, P.TrNext + l * CS + o);
      }
      const unsigned char *Pb = reinterpret_cast<const unsigned char *>(&S.Phi[buf][0][0]);
      const double2 ts2 = lds2(reinterpret_cast<const double *>(Pb + myoff));
      const double ts[2] = {ts2.x, ts2.y};
      double tend[2];
      tend[0] = ts[0] * S0[0];
      tend[1] = ts[1] * S0[1];
      if (HYP) {
         // Del2(phi) of the own cell (TracerAuxVars.h:61-91) ...
         double d[2] = {0, 0};
#pragma unroll
         for (int j = 0; j < kME; ++j) {
            const double2 tn = lds2(reinterpret_cast<const double *>(Pb + nboff[j]));
            d[0] -= cfA[j][0] * (ts[0] - tn.x);
            d[1] -= cfA[j][1] * (ts[1] - tn.y);
            tend[0] += E[j][0] * tn.x;
            tend[1] += E[j][1] * tn.y;
         }
         d[0] *= InvA;
         d[1] *= InvA;
         if (own) *reinterpret_cast<double2 *>(reinterpret_cast<unsigned char *>(&S.D2[0][0]) + myoff) = make_double2(d[0], d[1]);
         // ... and of one ring-1 row piece
         for (int q = tid; q < n1 * kTLanes; q += kTThreads) {
            const int r = nc + (q >> 3);
            const unsigned ro = (r * kTKC + 2 * p) * 8u;
            const double2 hr = lds2(reinterpret_cast<const double *>(Hb + ro));
            const double2 pr = lds2(reinterpret_cast<const double *>(Pb + ro));
            const uint4 t = *reinterpret_cast<const uint4 *>(&S.Nbl[r][0]);
            const unsigned w[3] = {t.x, t.y, t.z};
            double e0 = 0, e1 = 0;
#pragma unroll
            for (int j = 0; j < kME; ++j) {
               const unsigned lo = (((w[j >> 1] >> ((j & 1) * 16)) & 0xffffu) * kTKC + 2 * p) * 8u;
               const double2 hn = lds2(reinterpret_cast<const double *>(Hb + lo));
               const double2 pn = lds2(reinterpret_cast<const double *>(Pb + lo));
               const double g   = S.G[r][j];
               e0 -= g * (0.5 * (hr.x + hn.x)) * (pr.x - pn.x);
               e1 -= g * (0.5 * (hr.y + hn.y)) * (pr.y - pn.y);
            }
            const double ia = S.InvA[r];
            *reinterpret_cast<double2 *>(reinterpret_cast<unsigned char *>(&S.D2[0][0]) + ro) = make_double2(e0 * ia, e1 * ia);
         }
         __syncthreads(); // Del2(phi) of own + ring-1 rows complete
         const unsigned char *Db = reinterpret_cast<const unsigned char *>(&S.D2[0][0]);
         const double2 ds = lds2(reinterpret_cast<const double *>(Db + myoff));
         tend[0] += ds.x * SH;
         tend[1] += ds.y * SH;
#pragma unroll
         for (int j = 0; j < kME; ++j) {
            const double2 dn = lds2(reinterpret_cast<const double *>(Db + nboff[j]));
            tend[0] += ch[j] * dn.x;
            tend[1] += ch[j] * dn.y;
         }
      } else {
#pragma unroll
         for (int j = 0; j < kME; ++j) {
            const double2 tn = lds2(reinterpret_cast<const double *>(Pb + nboff[j]));
            tend[0] += E[j][0] * tn.x;
            tend[1] += E[j][1] * tn.y;
         }
      }
      // epilogue: weight / accumulate / provisional / finalise (TimeStepper.cpp:446-524)
      if (own) {
         double out[2];
#pragma unroll
         for (int q = 0; q < 2; ++q) {
            if (tcur_is_ts) tcur[q] = ts[q];
            const double base = P.First ? tcur[q] * hx[q] : acc[q];
            out[q]            = base + P.B * tend[q];
            if (P.Last) out[q] = out[q] * rhp[q];
         }
         stv<2>(P.TrNext + l * CS + o, out);
         if (!P.Last) {
#pragma unroll
            for (int q = 0; q < 2; ++q) out[q] = (tcur[q] * h2[q] + P.A * tend[q]) * rhp[q];
            stv<2>(P.TrProv + l * CS + o, out);
         }
      }
   }
   cp_wait<0>();
}

size_t tile_smem_bytes() { return sizeof(TileSmem); }

int launch_fCT(const DMesh &D, const CellArgs &ca, cudaStream_t St) {
   static bool attr = false;
   if (!attr) {
      OMEGA_CUDA_CHECK(cudaFuncSetAttribute(k_fCT<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TileSmem)));
      OMEGA_CUDA_CHECK(cudaFuncSetAttribute(k_fCT<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TileSmem)));
      attr = true;
   }
   const int nch = D.K / kTKC;
   const unsigned grid = (unsigned)D.Tiles.NTiles * nch;
   if (ca.HypOn) k_fCT<true><<<grid, kTThreads, sizeof(TileSmem), St>>>(D, ca, nch);
   else k_fCT<false><<<grid, kTThreads, sizeof(TileSmem), St>>>(D, ca, nch);
   return 0;
}

bool tile_path_ok(const DMesh &D) { return D.Tiles.NTiles > 0 && D.K % kTKC == 0; }

} // namespace omega
