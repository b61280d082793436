// fused_tile.cu -- shared-memory staged cell pass of the fused RK4 pipeline.
//
// k_fCT does the work of k_fT2 + k_fC_center (fused.cu) in ONE pass and never
// materialises Del2Tracers: a CTA owns one tile (omega_dev.h: <= 64 cells and
// the two rings of neighbours around them) x 16 levels.  Rows of h and of one
// tracer after the other are gathered into shared memory with cp.async
// (LDGSTS, 16 B per lane, up to three tracers in flight), so
//   * every row a tile needs crosses L2 -> SM once per tile instead of once
//     per neighbour that reads it, with no index chain in front of it;
//   * per-edge coefficients stay in registers for all tracers and neighbours
//     are addressed by precomputed shared-memory offsets;
//   * Del2(phi) of the own cells AND of ring 1 is computed in shared memory
//     (ring 1 redundantly, ~0.6 extra rows per own row), which removes the
//     3 T doubles per cell-layer the Del2Tracers array costs in HBM.
// With tracer hyperdiffusion on the CTA is split into two roles of 16 warps:
//   D role: Del2(phi_m) of own + ring-1 rows into one of two D2 buffers
//   T role: issues the copies, then tendency + RK update of tracer l = m - 1
// handing buffers over with named barriers (bar.arrive / bar.sync), so the two
// halves of the stencil overlap and each role fits 64 registers per thread.
// Algorithmic DRAM traffic per cell-layer and stage: 8 + 4 T doubles
// (h, hcur, hnext r/w, hprov, 3 u | phi, cur, next r/w or prov) against
// 9 + 8 T for k_fT2 + k_fC_center.
//
// Reference semantics: the same functors and update rules as k_fC_center
// (TendencyTerms.h:35-58,349-480, TracerAuxVars.h:25-91,
// LayerThicknessAuxVars.h:25-61, TimeStepper.cpp:378-524); MeshScalingDel2/4
// are 1 (HorzMesh.cpp:604-626), which the per-cell tables assume.
#include <cuda.h> // CUtensorMap (type and enums only: the encoder is looked up through cudaGetDriverEntryPoint)

#include <cstdlib>
#include <cstring>

#include "fused_args.h"
#include "kernel_util.cuh"
#include "omega_dev.h"
#include "stepper.h"

#ifdef OMEGA_SINGLE_PRECISION
// The tile kernels are written for fp64 rows (a 16-level chunk is one 128-byte row piece, a thread owns two
// levels = one 16-byte shared-memory access).  The single-precision build steps through the untiled fused
// kernels of fused.cu and the kernel-per-term path; these stand in for the tile launchers there.
namespace omega {
bool tile_path_ok(const DMesh &) { return false; }
bool tile_partial_chunk(const DMesh &) { return false; }
bool tile_v2_enabled() { return false; }
bool tile_v2_used(const CellArgs &) { return false; }
bool tile_edge_path_ok(const DMesh &) { return false; }
size_t tile_smem_bytes() { return 0; }
int launch_fET(const DMesh &, const EdgeTileArgs &, cudaStream_t, const TileRange &, bool) {
   OMEGA_REQUIRE(false, "launch_fET: the tile kernels are not part of the single-precision build");
}
int launch_fCT(const DMesh &, const CellArgs &, cudaStream_t, const TileRange &) {
   OMEGA_REQUIRE(false, "launch_fCT: the tile kernels are not part of the single-precision build");
}
} // namespace omega
#else
namespace omega {

constexpr int kTKC     = 16;                    // levels per tile
constexpr int kTLanes  = 8;                     // threads per row (2 levels each)
constexpr int kTRole   = kTileCells * kTLanes;  // threads per role
constexpr int kTStages = 5;                     // tracer row-set buffers

// KC levels per chunk, NS tracer row-set buffers
template <int KC, int NS> struct TileSmemT {
   double H[kTileRows][KC];
   double Phi[NS][kTileRows][KC];
   double D2[2][kTileRows1][KC];
   double Cf1[kTileRing1][6][KC]; // ring-1 rows: G * mean thickness per edge (private to the thread that uses it)
   double G[kTileRows1][6];
   double Ch[kTileCells][6];
   double CfU[kTileCells][6], Wa[kTileCells][6]; // own rows: dv*sign, mask*dv*sign per edge
   double InvA[kTileRows1];
   unsigned short Nbl[kTileRows1][8];
   int GRow[kTileRows];
   int Eoc[kTileCells][6]; // own rows: EdgesOnCell (boundary slot when none)
   TileHdr Hdr;
};
using TileSmem = TileSmemT<kTKC, kTStages>;

// cudaFuncSetAttribute and the SM count are per DEVICE: a host process may drive several devices
// (one OmegaMesh each), so the one-time set-up below is keyed by the current device.
constexpr int kMaxDev = 64;
struct DevOnce {
   bool Done[kMaxDev] = {};
   int Sm[kMaxDev]    = {};
};
static int cur_dev() {
   int d = 0;
   cudaGetDevice(&d);
   return d >= 0 && d < kMaxDev ? d : 0;
}
template <class F> static int set_smem_once(DevOnce &O, F *Kernel, size_t Bytes, int *Sm = nullptr) {
   const int d = cur_dev();
   if (!O.Done[d]) {
      OMEGA_CUDA_CHECK(cudaFuncSetAttribute(Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Bytes));
      OMEGA_CUDA_CHECK(cudaDeviceGetAttribute(&O.Sm[d], cudaDevAttrMultiProcessorCount, d));
      O.Done[d] = true;
   }
   if (Sm) *Sm = O.Sm[d];
   return 0;
}
static int env_i(const char *Name, int Dflt) {
   const char *v = getenv(Name);
   return v ? atoi(v) : Dflt;
}

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async16(unsigned dst, const void *src) {
   asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
// (PTX: when the barrier completes, the prior memory accesses of every participating
// thread -- arriving or waiting -- are performed relative to all participants)
__device__ __forceinline__ void bar_arrive(int id, int n) {
   asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory");
}
__device__ __forceinline__ double2 lds2(const unsigned char *base, unsigned off) {
   return *reinterpret_cast<const double2 *>(base + off);
}
__device__ __forceinline__ void sts2(unsigned char *base, unsigned off, double a, double b) {
   *reinterpret_cast<double2 *>(base + off) = make_double2(a, b);
}
// local row of neighbour j from the packed 8 x u16 row of Nbl
__device__ __forceinline__ unsigned nbl_get(const uint4 &t, int j) {
   const unsigned w = j < 2 ? t.x : (j < 4 ? t.y : t.z);
   return (w >> ((j & 1) * 16)) & 0xffffu;
}

// named barriers (0 is __syncthreads)
__device__ __forceinline__ int bar_phi_ready(int l) { return 1 + (l & 1); } // T -> D: tracer l has landed
__device__ __forceinline__ int bar_d2_ready(int l) { return 3 + (l & 1); }  // D -> T: Del2 of tracer l is complete

// the common view of a tile a thread needs in either role
struct TileCtx {
   int c0, nc, n1, nR1, nR; // tile geometry
   int t, i, p, li;         // thread in role, own row, 16-byte piece, clamped own row
   bool own;
   unsigned myoff;          // byte offset of the own piece inside a [row][16] buffer
};

// static tables of the own + ring-1 rows (all threads of the CTA take part)
template <int NT, class SM>
__device__ __forceinline__ void load_tables(const DMesh &M, SM &S, int tid, int nR1, int nc) {
   for (int q = tid; q < nc * 6; q += NT) {
      const size_t g = (size_t)S.Hdr.C0 * 6 + q;
      (&S.Eoc[0][0])[q] = M.Tiles.EocCell[g];
      (&S.CfU[0][0])[q] = M.Tiles.CfUCell[g];
      (&S.Wa[0][0])[q]  = M.Tiles.WaCell[g];
   }
   for (int q = tid; q < nR1 * 6; q += NT) {
      const int r = q / 6, j = q - r * 6;
      S.G[r][j] = M.Tiles.GCell[(size_t)S.GRow[r] * 6 + j];
   }
   for (int r = tid; r < nR1; r += NT) {
      S.InvA[r] = M.Tiles.InvAreaCell[S.GRow[r]];
      *reinterpret_cast<uint4 *>(&S.Nbl[r][0]) =
          *reinterpret_cast<const uint4 *>(M.Tiles.Nbl + (size_t)(S.Hdr.NblOff + r) * 8);
   }
}

// One role's share of a row-set copy (h or one tracer) into a [row][16] buffer: item
// q = t + 512 i is row q/8, 16-byte piece q%8, and lands at byte 16 q of the buffer.
struct RowCopier {
   int goff[3]; // element offset of the source piece inside one [cell][level] field
   int n;
   unsigned t16;
   __device__ __forceinline__ void init(const TileSmem &S, int t, int nR, int K, int kc0) {
      n = 0;
      t16 = 16u * t;
#pragma unroll
      for (int i = 0; i < 3; ++i) {
         const int q = t + i * kTRole, r = q >> 3;
         goff[i] = 0;
         if (r < nR) {
            goff[i] = S.GRow[r] * K + kc0 + 2 * (q & 7);
            n       = i + 1;
         }
      }
   }
   __device__ __forceinline__ void copy(const double *F, const void *Dst) const {
      const unsigned d = smem_u32(Dst) + t16;
#pragma unroll
      for (int i = 0; i < 3; ++i)
         if (i < n) cp_async16(d + i * (16u * kTRole), F + goff[i]);
   }
};

// One role's share (512 threads) of a row-set copy (h or one tracer) of level chunk kc into a
// [row][16] buffer: item q = t + 512 i is row q/8, 16-byte piece q%8, and lands at byte 16 q.
// Source offsets are re-derived from GRow on every copy (three shared loads), not held in registers.
__device__ __forceinline__ void copy_rows_g(const TileSmem &S, int nR, int K, int t, const double *F, const void *Dst, int kc) {
   const unsigned d = smem_u32(Dst) + 16u * t;
#pragma unroll
   for (int i = 0; i < 3; ++i) {
      const int r = (t >> 3) + i * (kTRole / kTLanes);
      if (r < nR) cp_async16(d + i * (16u * kTRole), F + ((size_t)S.GRow[r] * K + kc + 2 * (t & 7)));
   }
}

// =================== D role: Del2(phi) of own + ring-1 rows ===================
// (TracerAuxVars.h:61-91; s*(phi(c1)-phi(c0)) == phi(self)-phi(nbr) exactly)
// With NCh > 1 the CTA walks the NCh level chunks of its tile one after the other: tables
// are loaded once, and h / the first two tracers of chunk c+1 are copied during the last
// two tracers of chunk c, so only the first chunk pays the start-up latency of a tile.
__device__ __forceinline__ void role_D(const DMesh &M, const CellArgs &P, TileSmem &S, const TileCtx &X, int kc0, int NCh,
                                       int tid) {
   const int T = P.NTracers;
   const int t = X.t, p = X.p;
   load_tables<2 * kTRole>(M, S, tid, X.nR1, X.nc);
   const bool ring = t < X.n1 * kTLanes; // this thread's ring-1 row piece (n1 <= 56 rows x 8 pieces < 512 threads)
   const int r1    = X.nc + (ring ? X.i : 0);
   const unsigned ro1 = (r1 * kTKC + 2 * p) * 8u;
   unsigned char *Cb  = reinterpret_cast<unsigned char *>(&S.Cf1[ring ? X.i : 0][0][0]) + 16u * p;
   for (int c = 0; c < NCh; ++c) {
      const int it0 = c * T;
      __syncthreads(); // h and tracer 0 of this chunk have landed (copied by the T role); tables are visible
      const unsigned char *Hb = reinterpret_cast<const unsigned char *>(&S.H[0][0]);
      // own row: -(1/A) * G * mean thickness per edge, in registers for all tracers
      double cfA[kME][2];
      {
         const double nia = -S.InvA[X.li];
         const double2 hs = lds2(Hb, X.myoff);
         const uint4 tt   = *reinterpret_cast<const uint4 *>(&S.Nbl[X.li][0]);
#pragma unroll
         for (int j = 0; j < kME; ++j) {
            const double2 hn = lds2(Hb, (nbl_get(tt, j) * kTKC + 2 * p) * 8u);
            const double g   = S.G[X.li][j] * nia;
            cfA[j][0]        = g * (0.5 * (hs.x + hn.x));
            cfA[j][1]        = g * (0.5 * (hs.y + hn.y));
         }
      }
      // ring-1 row piece: the same coefficients, kept in shared memory (private to this thread)
      if (ring) {
         const double nia = -S.InvA[r1];
         const double2 hr = lds2(Hb, ro1);
         const uint4 tt   = *reinterpret_cast<const uint4 *>(&S.Nbl[r1][0]);
#pragma unroll
         for (int j = 0; j < kME; ++j) {
            const double2 hn = lds2(Hb, (nbl_get(tt, j) * kTKC + 2 * p) * 8u);
            const double g   = S.G[r1][j] * nia;
            sts2(Cb, j * (kTKC * 8u), g * (0.5 * (hr.x + hn.x)), g * (0.5 * (hr.y + hn.y)));
         }
      }
      for (int m = 0; m < T; ++m) {
         // tracer m has landed (T role, at the top of its iteration m-1: it is then also done with
         // tracer m-2, whose D2 buffer this iteration overwrites)
         if (m > 0) bar_sync(bar_phi_ready(m), 2 * kTRole);
         const unsigned char *Pb = reinterpret_cast<const unsigned char *>(&S.Phi[(it0 + m) % kTStages][0][0]);
         unsigned char *Db       = reinterpret_cast<unsigned char *>(&S.D2[m & 1][0][0]);
         if (X.own) {
            // neighbour rows are re-read from the Nbl table every tracer: one shared load instead
            // of three registers held (and spilled) across the loop
            const double2 ts = lds2(Pb, X.myoff);
            const uint4 tt   = *reinterpret_cast<const uint4 *>(&S.Nbl[X.li][0]);
            double d0 = 0, d1 = 0;
#pragma unroll
            for (int j = 0; j < kME; ++j) {
               const double2 tn = lds2(Pb, (nbl_get(tt, j) * kTKC + 2 * p) * 8u);
               d0 += cfA[j][0] * (ts.x - tn.x);
               d1 += cfA[j][1] * (ts.y - tn.y);
            }
            sts2(Db, X.myoff, d0, d1);
         }
         if (ring) {
            const double2 pr = lds2(Pb, ro1);
            const uint4 tt   = *reinterpret_cast<const uint4 *>(&S.Nbl[r1][0]);
            double e0 = 0, e1 = 0;
#pragma unroll
            for (int j = 0; j < kME; ++j) {
               const double2 pn = lds2(Pb, (nbl_get(tt, j) * kTKC + 2 * p) * 8u);
               const double2 cf = lds2(Cb, j * (kTKC * 8u));
               e0 += cf.x * (pr.x - pn.x);
               e1 += cf.y * (pr.y - pn.y);
            }
            sts2(Db, ro1, e0, e1);
         }
         bar_arrive(bar_d2_ready(m), 2 * kTRole); // Del2 of tracer m is ready for the T role
      }
   }
}

// =================== T role: copies, thickness, tracer tendencies + RK updates ===================
template <bool HYP, int STAGE>
__device__ __forceinline__ void role_T(const DMesh &M, const CellArgs &P, TileSmem &S, const TileCtx &X, int kc0, int NCh, int tid) {
   constexpr int NT = HYP ? 2 * kTRole : kTRole;
   constexpr bool First = STAGE == 0, Last = STAGE == 2; // stage kind: first, middle, last (RK4: 0 | 1,2 | 3)
   const int K = M.K, T = P.NTracers, t = X.t, p = X.p, li = X.li;
   const size_t CS = (size_t)M.NCellsSize * K;
   const bool own  = X.own;
   RowCopier cp; // without the D role this role feeds the pipeline itself
   if (!HYP) {
      cp.init(S, t, X.nR, K, kc0);
      cp.copy(P.Hs, &S.H[0][0]);
      cp_commit();
#pragma unroll
      for (int s = 0; s < kTStages - 1; ++s) {
         if (s < T) cp.copy(P.Trs + s * CS, &S.Phi[s][0][0]);
         cp_commit();
      }
   }
   // With the D role this role feeds the pipeline too (its iterations are the shorter ones).
   // Tracer m of chunk c is item c*T + m of the copy stream and lives in buffer (c*T + m) % kTStages;
   // iteration l copies the item `lead` places ahead: a tracer of this chunk, then h and the
   // first tracers of the next chunk.
   const int lead = T >= 3 ? 3 : 2;
   if (HYP) {
      copy_rows_g(S, X.nR, K, t, P.Hs, &S.H[0][0], kc0);
      cp_commit();
      for (int s = 0; s + 1 < lead; ++s) {
         copy_rows_g(S, X.nR, K, t, P.Trs + s * CS, &S.Phi[s][0][0], kc0);
         cp_commit();
      }
   }
   load_tables<NT>(M, S, tid, X.nR1, X.nc);

   const int c = X.c0 + li;
   // with the D role the CTA walks the NCh level chunks of its tile (see role_D)
   for (int ch = 0; ch < NCh; ++ch) {
   const int k    = kc0 + ch * kTKC + 2 * p;
   const int it0  = ch * T; // stream index of this chunk's tracer 0 (phi buffer = index % kTStages)
   const size_t o = (size_t)c * K + k;
   const unsigned char *Hb = reinterpret_cast<const unsigned char *>(&S.H[0][0]);
   if (HYP) { // tracer lead-1 of this chunk (T >= 2); h and the tracers before it are already on their way
      copy_rows_g(S, X.nR, K, t, P.Trs + (size_t)(lead - 1) * CS, &S.Phi[(it0 + lead - 1) % kTStages][0][0], kc0 + ch * kTKC);
      cp_commit();
   }
   // edge indices and velocity rows of the own cell: in flight while h lands
   double us[kME][2], ut[kME][2];
   {
      int e[kME];
#pragma unroll
      for (int j = 0; j < kME; ++j) e[j] = ch == 0 ? M.Tiles.EocCell[(size_t)c * 6 + j] : S.Eoc[li][j];
#pragma unroll
      for (int j = 0; j < kME; ++j) ldv<2>(us[j], P.Us + (size_t)e[j] * K + k);
      if (P.UTr != P.Us) {
#pragma unroll
         for (int j = 0; j < kME; ++j) ldv<2>(ut[j], P.UTr + (size_t)e[j] * K + k);
      }
   }
   double hcur[2] = {0, 0}, hnext_old[2] = {1.0, 1.0};
   const bool hcur_is_hs = First && P.Hcur == P.Hs;
   if (own) {
      if (!hcur_is_hs) ldp<2>(hcur, P.Hcur + o);
      if (!First || P.WSel) ldp<2>(hnext_old, P.Hnext + o);
   }
   // h and tracer 0 have landed (this thread's pieces)
   if (!HYP) cp_wait<kTStages - 2>();
   else if (lead == 3) cp_wait<2>();
   else cp_wait<1>();
   __syncthreads(); // ... and everybody else's; tables are visible

   // byte offsets of the six neighbour pieces inside a [row][16] buffer, two per register
   unsigned nb2[3];
   {
      const uint4 tt = *reinterpret_cast<const uint4 *>(&S.Nbl[li][0]);
#pragma unroll
      for (int j = 0; j < 3; ++j)
         nb2[j] = ((nbl_get(tt, 2 * j) * kTKC + 2 * p) * 8u) | (((nbl_get(tt, 2 * j + 1) * kTKC + 2 * p) * 8u) << 16);
   }
#define NBOFF(j) (((j) & 1) ? nb2[(j) >> 1] >> 16 : nb2[(j) >> 1] & 0xffffu)
   const unsigned myoff = X.myoff;
   const double InvA    = S.InvA[li];
   const double2 hs     = lds2(Hb, myoff);
   // per-edge coefficients of the own cell (as k_fC_center, fused.cu):
   //   tend = phi(self)*S0 + sum_j E_j*phi(nbr_j) + d2(self)*SH + sum_j ch_j*d2(nbr_j)
   double E[kME][2], S0[2] = {0, 0}, SH = 0.0, hT[2] = {0, 0};
   {
      const double k2 = P.DiffOn ? P.EddyDiff2 * InvA : 0.0, k4 = HYP ? P.EddyDiff4 * InvA : 0.0;
      const bool sameU = P.UTr == P.Us;
#pragma unroll
      for (int j = 0; j < kME; ++j) {
         const double2 hn2 = lds2(Hb, NBOFF(j));
         const double hn[2] = {hn2.x, hn2.y}, hh[2] = {hs.x, hs.y};
         const double g   = S.G[li][j];
         const double cfu = S.CfU[li][j];
         const double wa  = P.AdvOn ? S.Wa[li][j] : 0.0;
         const double cd = -k2 * g, ch = -k4 * g;
         SH -= ch;
         if (HYP && own && p == j) S.Ch[li][j] = ch;
#pragma unroll
         for (int q = 0; q < 2; ++q) {
            const double mean = 0.5 * (hh[q] + hn[q]);
            // upwind options (LayerThicknessAuxVars.h:25-61, TracerAuxVars.h:25-59): the edge value is the one
            // of the cell the stage velocity comes from; cell 0 of the edge is this cell iff the sign is -1
            // (HorzMesh.cpp:536-548).  u == 0 multiplies either choice by zero.
            const bool out = (cfu < 0.0) == (us[j][q] > 0.0); // flow leaves this cell through edge j
            const double fl = P.FluxUpwind ? (out ? hh[q] : hn[q]) : mean;
            hT[q] += cfu * fl * us[j][q] * InvA;
            const double w = wa * (sameU ? us[j][q] : ut[j][q]) * InvA;
            const double D = cd * mean;
            const double as = P.TrUpwind ? (out ? 1.0 : 0.0) : 0.5;
            E[j][q]        = (P.TrUpwind ? 1.0 - as : 0.5) * hn[q] * w - D;
            S0[q] += as * hh[q] * w + D;
         }
      }
   }
   if (!P.FluxDivOn) hT[0] = hT[1] = 0.0;
   if (P.SrcH) { // manufactured thickness source (CustomTendencyTerms.cpp:107-143), evaluated per cell by k_ms_sources
      const double src = P.SrcH[c];
      hT[0] += src;
      hT[1] += src;
   }
   // ---- thickness updates (updateThicknessByTend, TimeStepper.cpp:378-405) and the
   // per-level constants of the tracer updates (TimeStepper.cpp:446-524):
   //   first stage : next = cur*w0 + B*tend          (w0 = weight thickness)
   //   other stages: next = next + B*tend, last: next = (next + B*tend)*w0 (w0 = 1/h_final)
   //   not last    : prov = cur*w1 + tend*w2         (w1 = h2/hprov, w2 = A/hprov)
   double w0[2], w1[2], w2[2];
   {
      double hnext[2], hprov[2];
      if (hcur_is_hs || !own) {
         hcur[0] = hs.x;
         hcur[1] = hs.y;
      }
#pragma unroll
      for (int q = 0; q < 2; ++q) {
         hnext[q] = (First ? hcur[q] : hnext_old[q]) + P.B * hT[q];
         hprov[q] = hcur[q] + P.A * hT[q];
         const double hx  = First ? (P.WSel ? hnext_old[q] : hcur[q]) : (P.FinSel ? hcur[q] : hnext[q]);
         const double h2  = P.H2Sel == 2 ? hprov[q] : (P.H2Sel ? hnext[q] : hcur[q]);
         const double rhp = 1.0 / (P.DivSel ? hcur[q] : hprov[q]);
         w0[q] = Last ? 1.0 / hx : hx;
         w1[q] = h2 * rhp;
         w2[q] = P.A * rhp;
      }
      if (own) {
         stv<2>(P.Hnext + o, hnext);
         if (!Last) stv<2>(P.Hprov + o, hprov);
      }
   }

   // ---- tracers
   const bool tcur_is_ts = First && P.TrCur == P.Trs;
   const bool need_cur   = !tcur_is_ts && (First || !Last);
   // cur / next of tracer l are loaded at the END of iteration l-1 (their registers are dead
   // by then), so they are in flight during the barrier wait and the stencil of iteration l
   double tcur[2] = {0, 0}, acc[2] = {0, 0};
   if (own && T > 0) {
      if (need_cur) ldp<2>(tcur, P.TrCur + o);
      if (!First) ldp<2>(acc, P.TrNext + o);
   }
   for (int l = 0; l < T; ++l) {
      if (HYP) {
         if (l + 1 < T) { // hand tracer l+1 to the D role (copied two iterations ago)
            if (lead == 3) cp_wait<1>(); else cp_wait<0>();
            bar_arrive(bar_phi_ready(l + 1), 2 * kTRole);
         }
         bar_sync(bar_d2_ready(l), 2 * kTRole); // Del2(phi_l) is complete (and the D role past its use of h)
         const int n = l + lead, kc = kc0 + ch * kTKC;
         if (n < T) copy_rows_g(S, X.nR, K, t, P.Trs + (size_t)n * CS, &S.Phi[(it0 + n) % kTStages][0][0], kc);
         else if (ch + 1 < NCh) {
            if (n == T) copy_rows_g(S, X.nR, K, t, P.Hs, &S.H[0][0], kc + kTKC);
            else copy_rows_g(S, X.nR, K, t, P.Trs + (size_t)(n - T - 1) * CS, &S.Phi[(it0 + n - 1) % kTStages][0][0], kc + kTKC);
         }
         cp_commit();
      } else {
         cp_wait<kTStages - 2>();
         __syncthreads(); // phi_l visible; every thread is done with tracer l-1
         if (l + kTStages - 1 < T) cp.copy(P.Trs + (size_t)(l + kTStages - 1) * CS, &S.Phi[(l + kTStages - 1) % kTStages][0][0]);
         cp_commit();
      }
      const unsigned char *Pb = reinterpret_cast<const unsigned char *>(&S.Phi[(it0 + l) % kTStages][0][0]);
      const double2 ts = lds2(Pb, myoff);
      double tend[2] = {ts.x * S0[0], ts.y * S0[1]};
#pragma unroll
      for (int j = 0; j < kME; ++j) {
         const double2 tn = lds2(Pb, NBOFF(j));
         tend[0] += E[j][0] * tn.x;
         tend[1] += E[j][1] * tn.y;
      }
      if (HYP) {
         const unsigned char *Db = reinterpret_cast<const unsigned char *>(&S.D2[l & 1][0][0]);
         const double2 ds = lds2(Db, myoff);
         tend[0] += ds.x * SH;
         tend[1] += ds.y * SH;
#pragma unroll
         for (int j = 0; j < kME; ++j) {
            const double2 dn = lds2(Db, NBOFF(j));
            const double ch  = S.Ch[li][j];
            tend[0] += ch * dn.x;
            tend[1] += ch * dn.y;
         }
      }
      if (own) {
         double out[2];
         if (tcur_is_ts) {
            tcur[0] = ts.x;
            tcur[1] = ts.y;
         }
#pragma unroll
         for (int q = 0; q < 2; ++q) {
            out[q] = (First ? tcur[q] * w0[q] : acc[q]) + P.B * tend[q];
            if (Last) out[q] *= w0[q];
         }
         stv<2>(P.TrNext + l * CS + o, out);
         if (!Last) {
#pragma unroll
            for (int q = 0; q < 2; ++q) out[q] = tcur[q] * w1[q] + tend[q] * w2[q];
            stv<2>(P.TrProv + l * CS + o, out);
         }
         if (l + 1 < T) {
            if (need_cur) ldp<2>(tcur, P.TrCur + (size_t)(l + 1) * CS + o);
            if (!First) ldp<2>(acc, P.TrNext + (size_t)(l + 1) * CS + o);
         }
      }
      if (HYP && l == T - 2 && ch + 1 < NCh) {
         // the next chunk's velocity / thickness lines towards L2 while this chunk finishes
#pragma unroll
         for (int j = 0; j < kME; ++j) prefetch_row(P.Us + (size_t)S.Eoc[li][j] * K + k + kTKC, 2);
         if (own) {
            if (!hcur_is_hs) prefetch_row(P.Hcur + o + kTKC, 2);
            if (!First || P.WSel) prefetch_row(P.Hnext + o + kTKC, 2);
         }
      }
   }
   } // level chunks
   cp_wait<0>();
}

#undef NBOFF

template <bool HYP, int STAGE, int RS = 0>
__global__ void __launch_bounds__(HYP ? 2 * kTRole : kTRole, 1)
k_fCT(DMesh M, CellArgs P, int NChunks, TileRange TR) {
   extern __shared__ __align__(16) unsigned char smem_raw[];
   TileSmem &S = *reinterpret_cast<TileSmem *>(smem_raw);
   constexpr int NT = HYP ? 2 * kTRole : kTRole;
   const int tid = threadIdx.x;
   // NChunks > 0: one CTA per (tile, level chunk); NChunks < 0: one CTA per tile walking -NChunks chunks
   const int NChunksPerCta = NChunks > 0 ? 1 : -NChunks;
   const int tile = TR.tile(NChunks > 0 ? blockIdx.x / NChunks : blockIdx.x), kc0 = NChunks > 0 ? (blockIdx.x % NChunks) * kTKC : 0;

   if (tid < (int)(sizeof(TileHdr) / sizeof(int)))
      reinterpret_cast<int *>(&S.Hdr)[tid] = reinterpret_cast<const int *>(M.Tiles.Hdr + tile)[tid];
   __syncthreads();
   TileCtx X;
   X.c0 = S.Hdr.C0; X.nc = S.Hdr.NC; X.n1 = S.Hdr.N1; X.nR1 = X.nc + X.n1; X.nR = X.nR1 + S.Hdr.N2;
   for (int r = tid; r < X.nR; r += NT) S.GRow[r] = r < X.nc ? X.c0 + r : M.Tiles.HaloRows[S.Hdr.RowOff + r - X.nc];
   __syncthreads();
   X.t = tid & (kTRole - 1); X.i = X.t >> 3; X.p = X.t & 7;
   X.own = X.i < X.nc; X.li = X.own ? X.i : 0;
   X.myoff = (X.li * kTKC + 2 * X.p) * 8u;
   if (HYP) {
      // 1024 threads leave 64 registers each: the T role needs more, the D role less
      if (tid >= kTRole) {
         if (RS == 0) asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
         if (RS == 1) asm volatile("setmaxnreg.dec.sync.aligned.u32 48;");
         if (RS == 2) asm volatile("setmaxnreg.dec.sync.aligned.u32 64;");
         role_D(M, P, S, X, kc0, NChunksPerCta, tid);
      } else {
         if (RS == 0) asm volatile("setmaxnreg.inc.sync.aligned.u32 72;");
         if (RS == 1) asm volatile("setmaxnreg.inc.sync.aligned.u32 80;");
         if (RS == 2) asm volatile("setmaxnreg.inc.sync.aligned.u32 64;");
         role_T<true, STAGE>(M, P, S, X, kc0, NChunksPerCta, tid);
      }
   } else {
      role_T<false, STAGE>(M, P, S, X, kc0, 1, tid);
   }
}


// ---------------------------------------------------------------------------
// k_fCT2: the cell pass with the own-row Del2 folded into the tendency threads.
// ncu on k_fCT shows its LSU data pipe 2/3 busy (45 % of peak in shared loads
// alone): it is bound by shared-memory wavefronts, and a quarter of them are the
// neighbour tracer rows read twice (T role: advection sum, D role: Del2 of the own
// rows).  Here a T thread reads each neighbour row once and feeds both sums, which
// needs both coefficient sets in registers (setmaxnreg), and the helper role R (half
// as many threads) keeps what cannot be shared: the copies and the Del2 of the
// ring-1 rows.  An iteration i of the T role runs the neighbour pass A of tracer
// i and the Del4 pass + RK update B of tracer i-1, which are independent; one
// __syncthreads per iteration is the whole protocol (everything a role writes in
// iteration i is first read in iteration i+1).
//
// Two shapes (Cfg2<KC>): 16 levels per chunk, 768 threads, one CTA per SM; and 8 levels
// per chunk, 384 threads, two CTAs per SM.  The second keeps the SM busy while one CTA
// sits in the latency of its chunk prologue or at a barrier.
// ---------------------------------------------------------------------------
template <int KC_, int RS_ = 0, int CP_ = 0, int TM_ = 0> struct Cfg2 {
   static constexpr int KC = KC_;
   static constexpr int RS = RS_;            // 0: T 96 / R 40 registers, 2: 96/48, 3: 96/40 with 384 R threads            // levels per chunk
   static constexpr int L  = KC / 2;         // threads per row piece set (2 levels each)
   static constexpr int LS = KC == 16 ? 3 : 2; // log2(L)
   static constexpr int TT = kTileCells * L; // T role threads
   static constexpr int RT = RS_ == 3 ? 3 * TT / 4 : TT / 2; // R role threads
   static constexpr int NT = TT + RT;
   static constexpr int NS = 4;              // tracer row-set buffers (copy lead 3 needs 4)
   static constexpr int CTAs = KC == 16 ? 1 : 2;
   static constexpr bool BULK = CP_ >= 1;    // row copies: 0 = cp.async 16-byte pieces (LSU), 1, 2 = one bulk copy per row (TMA unit, mbarrier)
   static constexpr bool TMAP = CP_ == 3;    // 3: tensor-map copies, FOUR rows per instruction (cp.async.bulk.tensor tile::gather4, sm_100)
   static constexpr bool ONEWAIT = CP_ == 2; // 2: ONE thread waits on the "landed" barrier, ahead of the __syncthreads that hands the buffer over
   static constexpr int TM = TM_;            // tensor memory (TMEM) for per-thread private slots: 1 = the RK weights of the T role
   static constexpr bool PAIR = false;       // one tracer per iteration (Cfg3: two)
   struct Smem : TileSmemT<KC, NS> {
      double2 W[3][TT]; // per T thread: the three per-level weights of the RK update (private slots, spares 12 registers)
      unsigned long long Bar[NS + 1]; // BULK: "landed" barriers of the tracer buffers and of h
      double SHc[kTileCells];         // own rows: -sum_j ch_j (the self coefficient of the Del4 sum)
      unsigned TmemBase;              // TM: base address of the CTA's tensor-memory allocation
      const void *MapH, *MapT;        // TMAP: tensor maps of the stage thickness [NCellsSize][K] and tracers [T NCellsSize][K]
      const double *HsBase, *TrsBase; //       and the arrays they describe (a copy's row offset follows from its source pointer)
   };
};
static_assert(sizeof(Cfg2<16>::Smem) <= 232448, "k_fCT2 shared memory");

__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
   asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
   asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// wait for the phase of parity `par`; a copy that never lands traps instead of hanging the device
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned par) {
   unsigned done = 0;
   for (int spin = 0; spin < (1 << 22); ++spin) {
      asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                   : "=r"(done)
                   : "r"(bar), "r"(par)
                   : "memory");
      if (done) return;
   }
   __trap();
}
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void *src, unsigned bytes, unsigned bar) {
   asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                "r"(bytes), "r"(bar)
                : "memory");
}

// four rows (128 bytes each: 16 levels from column Col) of a 2-D tensor map into four consecutive 128-byte rows at Dst
__device__ __forceinline__ void tma_gather4(unsigned Dst, const void *Map, int Col, int R0, int R1, int R2, int R3, unsigned Bar) {
   asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
                ::"r"(Dst), "l"(Map), "r"(Col), "r"(R0), "r"(R1), "r"(R2), "r"(R3), "r"(Bar)
                : "memory");
}

// ---- tensor memory (TMEM, 256 KB per SM: 128 lanes x 512 columns x 32 bit) as per-thread private storage.
// Nothing here is a matrix product: TMEM serves as a second on-chip scratchpad with its own datapath
// (tcgen05.st / tcgen05.ld, SASS STTM / LDTM) for values only the owning thread ever reads -- the RK weights
// of a T thread, the Del2 coefficients of an R thread -- which otherwise compete with the stencil reads for
// the shared-memory pipe this kernel is bound by.  With the 32x32b shape thread t of a warp owns lane
// 32 * (warp % 4) + t; warps that share a lane quarter use different columns.
__device__ __forceinline__ void tmem_alloc_all(unsigned *Slot) {
   asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(Slot)) : "memory");
   asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_free_all(unsigned Base) {
   asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(Base) : "memory");
}
__device__ __forceinline__ void tmem_st2(unsigned Addr, double a, double b) { // 4 columns: two doubles
   unsigned a0, a1, b0, b1;
   asm("mov.b64 {%0, %1}, %2;" : "=r"(a0), "=r"(a1) : "d"(a));
   asm("mov.b64 {%0, %1}, %2;" : "=r"(b0), "=r"(b1) : "d"(b));
   asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(Addr), "r"(a0), "r"(a1), "r"(b0), "r"(b1) : "memory");
}
__device__ __forceinline__ void tmem_ld2_issue(unsigned Addr, unsigned (&r)[4]) {
   asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(Addr) : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ double2 tmem_unpack2(const unsigned (&r)[4]) {
   double2 v;
   asm("mov.b64 %0, {%1, %2};" : "=d"(v.x) : "r"(r[0]), "r"(r[1]));
   asm("mov.b64 %0, {%1, %2};" : "=d"(v.y) : "r"(r[2]), "r"(r[3]));
   return v;
}

template <class C, bool PART = false>
__device__ __forceinline__ void copy_rows_n(const typename C::Smem &S, int nR, int K, int r, const double *F, const void *Dst, int kc) {
   // item q = r + RT i of a row-set copy is row q/L, 16-byte piece q%L, and lands at byte 16 q
   const unsigned d = smem_u32(Dst) + 16u * r;
   // last chunk of a level count that is not a multiple of KC: the pieces behind level K are not copied (their
   // shared-memory lanes hold stale values that only threads of the same, inactive lanes ever read)
   if (PART && kc + 2 * (r & (C::L - 1)) >= K) return;
#pragma unroll
   for (int i = 0; i < (kTileRows * C::L + C::RT - 1) / C::RT; ++i) {
      const int row = (r >> C::LS) + i * (C::RT / C::L);
      if (row < nR) cp_async16(d + i * (16u * C::RT), F + ((size_t)S.GRow[row] * K + kc + 2 * (r & (C::L - 1))));
   }
}

// one item of the copy stream: a row set (h or a tracer) of level chunk kc into Dst; bar = its barrier (BULK)
template <class C, bool PART = false>
__device__ __forceinline__ void issue_rows(typename C::Smem &S, int nR, int K, int r, int grow, const double *F, void *Dst, int kc,
                                           int bar) {
   if (C::TMAP) {
      // rows 4r .. 4r+3 of the tile's row list by one instruction (the last group repeats the last row: the
      // buffer holds kTileRows rows, a multiple of four)
      const unsigned b = smem_u32(&S.Bar[bar]);
      const int ng = (nR + 3) >> 2;
      if (r == 0) mbar_expect_tx(b, (unsigned)ng * (4 * C::KC * 8));
      if (r < ng) {
         const bool isH   = F == S.HsBase;
         const int base   = isH ? 0 : (int)((F - S.TrsBase) / K);
         const int4 g     = *reinterpret_cast<const int4 *>(&S.GRow[4 * r]);
         const int last   = S.GRow[nR - 1];
         const int q      = 4 * r;
         tma_gather4(smem_u32(Dst) + r * (4 * C::KC * 8), isH ? S.MapH : S.MapT, kc, base + g.x, base + (q + 1 < nR ? g.y : last),
                     base + (q + 2 < nR ? g.z : last), base + (q + 3 < nR ? g.w : last), b);
      }
   } else if (C::BULK) {
      const unsigned b = smem_u32(&S.Bar[bar]);
      if (r == 0) mbar_expect_tx(b, (unsigned)nR * (C::KC * 8));
      if (r < nR) bulk_g2s(smem_u32(Dst) + r * (C::KC * 8), F + ((size_t)grow * K + kc), C::KC * 8, b);
   } else {
      copy_rows_n<C, PART>(S, nR, K, r, F, Dst, kc);
      cp_commit();
   }
}
template <class C> __device__ __forceinline__ void wait_item(typename C::Smem &S, int item) { // tracer item of the copy stream
   if (C::BULK && !C::ONEWAIT) mbar_wait(smem_u32(&S.Bar[item % C::NS]), (item / C::NS) & 1);
}
template <class C> __device__ __forceinline__ void wait_h(typename C::Smem &S, int chunk) {
   if (C::BULK && !C::ONEWAIT) mbar_wait(smem_u32(&S.Bar[C::NS]), chunk & 1);
}
// ONEWAIT: called by the first R thread only, right before the barrier behind which the buffer is read
template <class C> __device__ __forceinline__ void wait_item_one(typename C::Smem &S, int item) {
   if (C::ONEWAIT) mbar_wait(smem_u32(&S.Bar[item % C::NS]), (item / C::NS) & 1);
}
template <class C> __device__ __forceinline__ void wait_h_one(typename C::Smem &S, int chunk) {
   if (C::ONEWAIT) mbar_wait(smem_u32(&S.Bar[C::NS]), chunk & 1);
}

__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// R role: copies (tracer i+3 while tracer i is worked on) and Del2 of the ring-1 rows
template <class C, bool PART = false>
__device__ __forceinline__ void role_R(const DMesh &M, const CellArgs &P, typename C::Smem &S, int c0, int nc, int n1, int nR,
                                       int NCh, int tid, bool Pf, int kb) {
   constexpr int KC = C::KC, L = C::L, RT = C::RT, NS = C::NS;
   const int T = P.NTracers, K = M.K, r = tid - C::TT, p = r & (L - 1);
   const size_t CS = (size_t)M.NCellsSize * K;
   static_assert(!C::BULK || C::RT >= kTileRows, "bulk copies: one row per R thread");
   const int grow = r < nR ? S.GRow[r] : 0;
   issue_rows<C, PART>(S, nR, K, r, grow, P.Hs, &S.H[0][0], kb, NS); // kb: first level of this CTA's chunks
   issue_rows<C, PART>(S, nR, K, r, grow, P.Trs, &S.Phi[0][0][0], kb, 0);
   if (T >= 2) issue_rows<C, PART>(S, nR, K, r, grow, P.Trs + CS, &S.Phi[1][0][0], kb, 1);
   // T >= 3: tracer i+3 goes out in iteration i (then h and tracers 0..2 of the next chunk).  Shorter tracer
   // lists: h and the tracers of chunk c+1 go out in iterations 0, 1, .. of chunk c (four buffers hold two chunks)
   const bool shortT = T < 3;
   load_tables<C::NT>(M, S, tid, nc + n1, nc);
   // ring-1 row pieces of this thread: items r and r + RT of n1*L (n1 <= 56)
   const int nTask = (r < n1 * L) + (r + RT < n1 * L);
   const unsigned char *Hb = reinterpret_cast<const unsigned char *>(&S.H[0][0]);
   const bool pfHcur = !(P.First && P.Hcur == P.Hs), pfHnext = !P.First || P.WSel;
   for (int c = 0; c < NCh; ++c) {
      const int it0 = c * T, kc = kb + c * KC;
      if (c == 0 && !shortT) issue_rows<C, PART>(S, nR, K, r, grow, P.Trs + 2 * CS, &S.Phi[2][0][0], kb, 2);
      if (!C::BULK) { // h and tracer 0 of this chunk have landed (this thread's pieces)
         if (!shortT) cp_wait<2>();
         else if (T == 2) cp_wait<1>();
         else cp_wait<0>();
      }
      if (r == 0) wait_h_one<C>(S, c);
      __syncthreads();            // ... and everybody else's; tables are visible; the T role is done with the previous chunk
      wait_h<C>(S, c);
      for (int u = 0; u < nTask; ++u) { // coefficients of the ring rows: -(1/A) G mean thickness, private slots in Cf1
         const int ri = (r >> C::LS) + u * (RT / L), row = nc + ri;
         unsigned char *Cb = reinterpret_cast<unsigned char *>(&S.Cf1[ri][0][0]) + 16u * p;
         const double nia = -S.InvA[row];
         const double2 hr = lds2(Hb, (row * KC + 2 * p) * 8u);
         const uint4 tt   = *reinterpret_cast<const uint4 *>(&S.Nbl[row][0]);
#pragma unroll
         for (int j = 0; j < kME; ++j) {
            const double2 hn = lds2(Hb, (nbl_get(tt, j) * KC + 2 * p) * 8u);
            const double g   = S.G[row][j] * nia;
            sts2(Cb, j * (KC * 8u), g * (0.5 * (hr.x + hn.x)), g * (0.5 * (hr.y + hn.y)));
         }
      }
      for (int i = 0; i <= T; ++i) {
         if (!C::BULK && i > 0 && i < T) { // tracer i has landed (this thread's pieces)
            if (!shortT) cp_wait<2>();
            else cp_wait<1>();
         }
         if (r == 0 && i < T) wait_item_one<C>(S, it0 + i);
         __syncthreads();                              // iteration i: the T role has finished iteration i-1
         if (shortT) {
            if (c + 1 < NCh) {
               if (i == 0) issue_rows<C, PART>(S, nR, K, r, grow, P.Hs, &S.H[0][0], kc + KC, NS);
               else issue_rows<C, PART>(S, nR, K, r, grow, P.Trs + (size_t)(i - 1) * CS, &S.Phi[(it0 + T + i - 1) % NS][0][0], kc + KC, (it0 + T + i - 1) % NS);
            } else if (!C::BULK) cp_commit();
         } else {  // next item of the copy stream: a tracer of this chunk, then h and tracers 0..2 of the next
            const int n = i + 3;
            if (n < T) issue_rows<C, PART>(S, nR, K, r, grow, P.Trs + (size_t)n * CS, &S.Phi[(it0 + n) % NS][0][0], kc, (it0 + n) % NS);
            else if (c + 1 < NCh) {
               if (n == T) issue_rows<C, PART>(S, nR, K, r, grow, P.Hs, &S.H[0][0], kc + KC, NS);
               else issue_rows<C, PART>(S, nR, K, r, grow, P.Trs + (size_t)(n - T - 1) * CS, &S.Phi[(it0 + n - 1) % NS][0][0], kc + KC, (it0 + n - 1) % NS);
            } else if (!C::BULK) cp_commit(); // keeps the group count of cp_wait<2> uniform
         }
         if (i == 1 && c + 1 < NCh && Pf && (!PART || kc + 2 * KC <= K)) { // (a partial next chunk is not prefetched)
            // pull what the T role's next chunk prologue reads from global memory into L2:
            // six velocity rows per own cell, and the thickness operands
            for (int q = r; q < nc * 8; q += RT) {
               const int cell = q >> 3, j = q & 7;
               const size_t o = (size_t)(c0 + cell) * K + kc + KC;
               if (j < 6) prefetch_l2(P.Us + (size_t)S.Eoc[cell][j] * K + kc + KC);
               else if (j == 6 ? pfHcur : pfHnext) prefetch_l2((j == 6 ? P.Hcur : P.Hnext) + o);
            }
         }
         if (i < T) {
            wait_item<C>(S, it0 + i);
            const unsigned char *Pb = reinterpret_cast<const unsigned char *>(&S.Phi[(it0 + i) % NS][0][0]);
            unsigned char *Db       = reinterpret_cast<unsigned char *>(&S.D2[i & 1][0][0]);
            for (int u = 0; u < nTask; ++u) {
               const int ri = (r >> C::LS) + u * (RT / L), row = nc + ri;
               const unsigned char *Cb = reinterpret_cast<const unsigned char *>(&S.Cf1[ri][0][0]) + 16u * p;
               const unsigned ro = (row * KC + 2 * p) * 8u;
               const double2 pr = lds2(Pb, ro);
               const uint4 tt   = *reinterpret_cast<const uint4 *>(&S.Nbl[row][0]);
               double e0 = 0, e1 = 0;
#pragma unroll
               for (int j = 0; j < kME; ++j) {
                  const double2 pn = lds2(Pb, (nbl_get(tt, j) * KC + 2 * p) * 8u);
#ifdef OMEGA_EXP_NOPRIV // timing experiment only (wrong results): private coefficient reads removed
                  const double2 cf = make_double2(1e-9 * j, 1e-9);
#else
                  const double2 cf = lds2(Cb, j * (KC * 8u));
#endif
                  e0 += cf.x * (pr.x - pn.x);
                  e1 += cf.y * (pr.y - pn.y);
               }
               sts2(Db, ro, e0, e1);
            }
         }
      }
   }
   if (!C::BULK) cp_wait<0>();
}

// ---------------------------------------------------------------------------
// k_fCT3 = k_fCT2 with TWO tracers per iteration.  k_fCT2 is not bound by shared-memory wavefronts alone
// (ncu r02a: LSU 72 % busy; removing every private coefficient read, 20 % of the wavefronts, gained 9 %):
// with one __syncthreads per tracer all 24 warps refill and drain the load pipe once per tracer, and a warp
// has one tracer's worth of independent loads to hide the shared-memory latency with.  A pair of tracers per
// barrier halves the barriers and doubles the independent work between them.  That needs four Del2 buffers
// (two pairs in flight); the room comes from moving the ring-row Del2 coefficients (Cf1, 43 KB, private to the
// R thread that uses them) out of shared memory into TENSOR MEMORY (tcgen05.st / tcgen05.ld, see above):
// an R thread loads the twelve coefficients of a ring piece once per pair instead of once per tracer.
// ---------------------------------------------------------------------------
struct Cfg3 {
   static constexpr int KC = 16, RS = 2, L = 8, LS = 3;
   static constexpr int TT = kTileCells * L, RT = TT / 2, NT = TT + RT, NS = 4, CTAs = 1;
   static constexpr bool BULK = false, ONEWAIT = false, PAIR = true, TMAP = false;
   static constexpr int TM = 2; // tensor memory holds the ring-row coefficients of the R role
   struct Smem {
      double H[kTileRows][KC];
      double Phi[NS][kTileRows][KC];
      double D2[4][kTileRows1][KC]; // [2 * (pair parity) + tracer of the pair]
      double G[kTileRows1][6];
      double Ch[kTileCells][6];
      double CfU[kTileCells][6], Wa[kTileCells][6];
      double InvA[kTileRows1];
      unsigned short Nbl[kTileRows1][8];
      int GRow[kTileRows];
      int Eoc[kTileCells][6];
      TileHdr Hdr;
      double2 W[3][TT];
      unsigned long long Bar[NS + 1]; // unused (bulk copies are a Cfg2 variant)
      double SHc[kTileCells];
      unsigned TmemBase;
   };
};
static_assert(sizeof(Cfg3::Smem) <= 232448, "k_fCT3 shared memory");

// R role of k_fCT3: copies one PAIR ahead, Del2 of the ring-1 rows of both tracers of the pair
template <class C>
__device__ __forceinline__ void role_R3(const DMesh &M, const CellArgs &P, typename C::Smem &S, int c0, int nc, int n1, int nR,
                                        int NCh, int tid, bool Pf, int kb) {
   constexpr int KC = C::KC, L = C::L, RT = C::RT, NS = C::NS;
   const int T = P.NTracers, K = M.K, r = tid - C::TT, p = r & (L - 1);
   const size_t CS = (size_t)M.NCellsSize * K;
   const int NP = (T + 1) >> 1; // pairs per chunk (the last one may hold one tracer)
   // TMEM slot: lane quarter of the warp; the two R warps of a quarter own columns [64, 112) and [112, 160):
   // 2 ring pieces x 6 coefficients x double2 = 48 columns per thread
   const int w = tid >> 5;
   const unsigned tmC = S.TmemBase + ((unsigned)(w & 3) << 21) + 64u + (unsigned)((w - C::TT / 32) >> 2) * 48u;
   // tracer l of chunk c is item c*T + l of the copy stream and lives in Phi[item % NS]
   copy_rows_n<C>(S, nR, K, r, P.Hs, &S.H[0][0], kb);
   copy_rows_n<C>(S, nR, K, r, P.Trs, &S.Phi[0][0][0], kb);
   if (T >= 2) copy_rows_n<C>(S, nR, K, r, P.Trs + CS, &S.Phi[1][0][0], kb);
   cp_commit();
   load_tables<C::NT>(M, S, tid, nc + n1, nc);
   const unsigned char *Hb = reinterpret_cast<const unsigned char *>(&S.H[0][0]);
   const bool pfHcur = !(P.First && P.Hcur == P.Hs), pfHnext = !P.First || P.WSel;
   for (int c = 0; c < NCh; ++c) {
      const int it0 = c * T, kc = kb + c * KC;
      cp_wait<0>();
      __syncthreads(); // h and the first pair of this chunk have landed; tables are visible; the T role is done with the previous chunk
      // coefficients of this thread's two ring pieces, -(1/A) G mean thickness per edge, into its TMEM slot
      // (warp-collective stores: every lane takes part, lanes without a piece store a clamped row's values)
#pragma unroll
      for (int u = 0; u < 2; ++u) {
         const int ri  = (r >> C::LS) + u * (RT / L);
         const int row = nc + (ri < n1 ? ri : 0);
         const double nia = -S.InvA[row];
         const double2 hr = lds2(Hb, (row * KC + 2 * p) * 8u);
         const uint4 tt   = *reinterpret_cast<const uint4 *>(&S.Nbl[row][0]);
#pragma unroll
         for (int j = 0; j < kME; ++j) {
            const double2 hn = lds2(Hb, (nbl_get(tt, j) * KC + 2 * p) * 8u);
            const double g   = S.G[row][j] * nia;
            tmem_st2(tmC + u * 24 + j * 4, g * (0.5 * (hr.x + hn.x)), g * (0.5 * (hr.y + hn.y)));
         }
      }
      tmem_wait_st();
      for (int pi = 0; pi <= NP; ++pi) {
         if (pi > 0) cp_wait<0>(); // the pair issued one iteration ago (this thread's pieces)
         __syncthreads();          // iteration pi: the T role has finished iteration pi - 1
         // next item of the copy stream: the next pair of this chunk, or h and the first pair of the next chunk
         if (pi + 1 < NP) {
            const int a = 2 * (pi + 1);
            copy_rows_n<C>(S, nR, K, r, P.Trs + (size_t)a * CS, &S.Phi[(it0 + a) % NS][0][0], kc);
            if (a + 1 < T) copy_rows_n<C>(S, nR, K, r, P.Trs + (size_t)(a + 1) * CS, &S.Phi[(it0 + a + 1) % NS][0][0], kc);
         } else if (pi + 1 == NP && c + 1 < NCh) {
            copy_rows_n<C>(S, nR, K, r, P.Hs, &S.H[0][0], kc + KC);
            copy_rows_n<C>(S, nR, K, r, P.Trs, &S.Phi[(it0 + T) % NS][0][0], kc + KC);
            if (T >= 2) copy_rows_n<C>(S, nR, K, r, P.Trs + CS, &S.Phi[(it0 + T + 1) % NS][0][0], kc + KC);
         }
         cp_commit();
         if (pi == 0 && c + 1 < NCh && Pf) {
            // pull what the T role's next chunk prologue reads from global memory into L2
            for (int q = r; q < nc * 8; q += RT) {
               const int cell = q >> 3, j = q & 7;
               const size_t o = (size_t)(c0 + cell) * K + kc + KC;
               if (j < 6) prefetch_l2(P.Us + (size_t)S.Eoc[cell][j] * K + kc + KC);
               else if (j == 6 ? pfHcur : pfHnext) prefetch_l2((j == 6 ? P.Hcur : P.Hnext) + o);
            }
         }
         if (pi < NP) {
            const int a = 2 * pi, nb = T - a < 2 ? T - a : 2;
#pragma unroll
            for (int u = 0; u < 2; ++u) {
               const int ri   = (r >> C::LS) + u * (RT / L);
               const bool act = ri < n1;
               const int row  = nc + (act ? ri : 0);
               const unsigned ro = (row * KC + 2 * p) * 8u;
               const uint4 tt    = *reinterpret_cast<const uint4 *>(&S.Nbl[row][0]);
               // both tracers of the pair share every coefficient load; three coefficients at a time keep the
               // role within its 48 registers (setmaxnreg: the T role's 96 come out of what this role releases)
               const unsigned char *Pb0 = reinterpret_cast<const unsigned char *>(&S.Phi[(it0 + a) % NS][0][0]);
               const unsigned char *Pb1 = reinterpret_cast<const unsigned char *>(&S.Phi[(it0 + a + (nb > 1)) % NS][0][0]);
               const double2 pr0 = lds2(Pb0, ro), pr1 = lds2(Pb1, ro);
               double e00 = 0, e01 = 0, e10 = 0, e11 = 0;
#pragma unroll
               for (int jh = 0; jh < kME; jh += 3) {
                  unsigned cw[3][4];
#pragma unroll
                  for (int j = 0; j < 3; ++j) tmem_ld2_issue(tmC + u * 24 + (jh + j) * 4, cw[j]);
                  tmem_wait_ld();
#pragma unroll
                  for (int j = 0; j < 3; ++j) {
                     const unsigned no = (nbl_get(tt, jh + j) * KC + 2 * p) * 8u;
                     const double2 cf = tmem_unpack2(cw[j]);
                     const double2 pn0 = lds2(Pb0, no), pn1 = lds2(Pb1, no);
                     e00 += cf.x * (pr0.x - pn0.x);
                     e01 += cf.y * (pr0.y - pn0.y);
                     e10 += cf.x * (pr1.x - pn1.x);
                     e11 += cf.y * (pr1.y - pn1.y);
                  }
               }
               if (act) {
                  sts2(reinterpret_cast<unsigned char *>(&S.D2[2 * (pi & 1)][0][0]), ro, e00, e01);
                  if (nb > 1) sts2(reinterpret_cast<unsigned char *>(&S.D2[2 * (pi & 1) + 1][0][0]), ro, e10, e11);
               }
            }
         }
      }
   }
   cp_wait<0>();
}

// global accesses of the T role that bypass L1 allocation: what little L1 is left beside
// 226 KB of shared memory then keeps the few spilled registers of the tracer loop
__device__ __forceinline__ void ldg_na(double (&r)[2], const double *p) {
   asm volatile("ld.global.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(r[0]), "=d"(r[1]) : "l"(p) : "memory");
}
__device__ __forceinline__ void ldg_nc_na(double (&r)[2], const double *p) {
   asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(r[0]), "=d"(r[1]) : "l"(p));
}
__device__ __forceinline__ void stg_na(double *p, const double (&r)[2]) {
   asm volatile("st.global.L1::no_allocate.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(r[0]), "d"(r[1]) : "memory");
}

// T role of k_fCT2
// OPT: the optional terms (upwind flux options, manufactured thickness source) are compiled in; the default
// instantiation keeps them out of the prologue's register budget
template <class C, int STAGE, bool OPT, bool PART = false>
__device__ __forceinline__ void role_T2(const DMesh &M, const CellArgs &P, typename C::Smem &S, const TileCtx &X, int NCh, int tid,
                                        int kb) {
   constexpr bool First = STAGE == 0, Last = STAGE == 2;
   constexpr int KC = C::KC, NS = C::NS;
   const int K = M.K, T = P.NTracers, p = X.p, li = X.li;
   const size_t CS = (size_t)M.NCellsSize * K;
   const bool own  = X.own; // (shadowed per chunk below by own-cell-and-existing-level)
   load_tables<C::NT>(M, S, tid, X.nR1, X.nc);
   const int c = X.c0 + li;
   const unsigned myoff = X.myoff;
   // TMEM address of this thread's private slot: lane quarter of the warp, 16 columns per warp of the quarter
   const unsigned tmW = C::TM == 1 ? S.TmemBase + ((unsigned)((tid >> 5) & 3) << 21) + (unsigned)(tid >> 7) * 16u : 0u;
   const unsigned char *Hb = reinterpret_cast<const unsigned char *>(&S.H[0][0]);
   const bool sameU = P.UTr == P.Us;
   const bool hcur_is_hs = First && P.Hcur == P.Hs;
   const bool tcur_is_ts = First && P.TrCur == P.Trs;
   const bool need_cur   = !tcur_is_ts && (First || !Last);
   // The global operands of a chunk's prologue (six velocity rows of the own cell, thickness
   // operands) are loaded one chunk ahead: for chunk 0 here, for chunk ch+1 in the last iteration
   // of chunk ch, where the registers of the tracer coefficients are free again.
   double us[kME][2], hcur[2] = {0, 0}, hnext_old[2] = {1.0, 1.0};
   auto load_prologue = [&](int ch) {
      // (lanes behind level K in the last chunk of a level count that is not a multiple of KC load row K-2 and
      // never store)
      const int k = PART ? min(kb + ch * KC + 2 * p, K - 2) : kb + ch * KC + 2 * p;
      int e[kME];
#pragma unroll
      for (int j = 0; j < kME; ++j) e[j] = ch == 0 ? M.Tiles.EocCell[(size_t)c * 6 + j] : S.Eoc[li][j];
#pragma unroll
      for (int j = 0; j < kME; ++j) ldg_nc_na(us[j], P.Us + (size_t)e[j] * K + k);
      if (own) {
         const size_t o = (size_t)c * K + k;
         if (!hcur_is_hs) ldg_na(hcur, P.Hcur + o);
         if (!First || P.WSel) ldg_na(hnext_old, P.Hnext + o);
      }
   };
#ifndef OMEGA_T2_HOIST
#define OMEGA_T2_HOIST 0
#endif
   if (OMEGA_T2_HOIST) load_prologue(0);
   const bool ownTile = X.own;
   for (int ch = 0; ch < NCh; ++ch) {
      const int kraw = kb + ch * KC + 2 * p, k = PART ? min(kraw, K - 2) : kraw, it0 = ch * T;
      const bool own = ownTile && (!PART || kraw < K); // own cell AND an existing level: loads its operands, stores results
      const size_t o = (size_t)c * K + k;
      if (!OMEGA_T2_HOIST) load_prologue(ch); // in flight while h lands
      __syncthreads(); // h and tracer 0 of this chunk are in shared memory; tables are visible
      wait_h<C>(S, ch);
      unsigned nb2[3]; // byte offsets of the six neighbour pieces inside a [row][16] buffer, two per register
      {
         const uint4 tt = *reinterpret_cast<const uint4 *>(&S.Nbl[li][0]);
#pragma unroll
         for (int j = 0; j < 3; ++j)
            nb2[j] = ((nbl_get(tt, 2 * j) * KC + 2 * p) * 8u) | (((nbl_get(tt, 2 * j + 1) * KC + 2 * p) * 8u) << 16);
      }
#define NBOFF(j) (((j) & 1) ? nb2[(j) >> 1] >> 16 : nb2[(j) >> 1] & 0xffffu)
      const double InvA = S.InvA[li];
      const double2 hs  = lds2(Hb, myoff);
      // per-edge coefficients of the own cell, for all tracers:
      //   A:  adv = phi(self)*S0 + sum_j E_j*phi(nbr_j)     d2(self) = sum_j cfA_j*(phi(self) - phi(nbr_j))
      //   B:  tend = adv + d2(self)*SH + sum_j ch_j*d2(nbr_j)
      double E[kME][2], cfA[kME][2], S0[2] = {0, 0}, hT[2] = {0, 0};
      { // first the sums that need the velocity rows (they retire one after the other) ...
         const double k2 = P.DiffOn ? P.EddyDiff2 * InvA : 0.0;
#pragma unroll
         for (int j = 0; j < kME; ++j) {
            const double2 hn2 = lds2(Hb, NBOFF(j));
            const double hn[2] = {hn2.x, hn2.y}, hh[2] = {hs.x, hs.y};
            const double g   = S.G[li][j];
            const double cfu = S.CfU[li][j];
            const double wa  = P.AdvOn ? S.Wa[li][j] : 0.0;
            const double cd  = -k2 * g;
            double ua[2] = {us[j][0], us[j][1]}; // velocity the tracer advection sees
            if (!sameU && P.AdvOn) {
               // stage 0 of an odd step with raw time-level indexing: tracer advection sees the
               // other velocity level (cold path)
               ldg_nc_na(ua, P.UTr + (size_t)S.Eoc[li][j] * K + k);
            }
#pragma unroll
            for (int q = 0; q < 2; ++q) {
               const double mean = 0.5 * (hh[q] + hn[q]);
               // upwind options: see role_T (the tracer edge value upwinds on the STAGE velocity, TracerAuxVars.h:25-59)
               const bool out  = (cfu < 0.0) == (us[j][q] > 0.0);
               const double fl = OPT && P.FluxUpwind ? (out ? hh[q] : hn[q]) : mean;
               const double as = OPT && P.TrUpwind ? (out ? 1.0 : 0.0) : 0.5, an = OPT && P.TrUpwind ? 1.0 - as : 0.5;
               hT[q] += cfu * fl * us[j][q] * InvA;
               const double w = wa * us[j][q] * InvA;
               const double D = cd * mean;
               E[j][q]        = an * hn[q] * w - D;
               S0[q] += as * hh[q] * w + D;
               if (!sameU && P.AdvOn) {
                  const double dw = wa * (ua[q] - us[j][q]) * InvA;
                  E[j][q] += an * hn[q] * dw;
                  S0[q] += as * hh[q] * dw;
               }
            }
         }
      }
      { // ... then the Del2 / Del4 coefficients, which need h only
         const double k4 = P.EddyDiff4 * InvA;
         double SH = 0.0;
#pragma unroll
         for (int j = 0; j < kME; ++j) {
            const double2 hn2 = lds2(Hb, NBOFF(j));
            const double g    = S.G[li][j];
            const double chj = -k4 * g, ga = -InvA * g;
            SH -= chj;
            if (ownTile && p == (j & (C::L - 1))) S.Ch[li][j] = chj; // (level independent: written by inactive lanes too)
            if (j == kME - 1 && ownTile && p == C::L - 1) S.SHc[li] = SH;
            cfA[j][0] = ga * (0.5 * (hs.x + hn2.x));
            cfA[j][1] = ga * (0.5 * (hs.y + hn2.y));
         }
      }
      if (!P.FluxDivOn) hT[0] = hT[1] = 0.0;
      if (OPT && P.SrcH) { // manufactured thickness source per cell (k_ms_sources)
         const double src = P.SrcH[c];
         hT[0] += src;
         hT[1] += src;
      }
      // thickness updates (updateThicknessByTend, TimeStepper.cpp:378-405) and the per-level
      // constants of the tracer updates (TimeStepper.cpp:446-524), as in role_T
      {
         double w0[2], w1[2], w2[2];
         double hnext[2], hprov[2];
         if (hcur_is_hs || !own) {
            hcur[0] = hs.x;
            hcur[1] = hs.y;
         }
#pragma unroll
         for (int q = 0; q < 2; ++q) {
            hnext[q] = (First ? hcur[q] : hnext_old[q]) + P.B * hT[q];
            hprov[q] = hcur[q] + P.A * hT[q];
            const double hx  = First ? (P.WSel ? hnext_old[q] : hcur[q]) : (P.FinSel ? hcur[q] : hnext[q]);
            const double h2  = P.H2Sel == 2 ? hprov[q] : (P.H2Sel ? hnext[q] : hcur[q]);
            const double rhp = 1.0 / (P.DivSel ? hcur[q] : hprov[q]);
            w0[q] = Last ? 1.0 / hx : hx;
            w1[q] = h2 * rhp;
            w2[q] = P.A * rhp;
         }
         if (own) {
            stg_na(P.Hnext + o, hnext);
            if (!Last) stg_na(P.Hprov + o, hprov);
         }
         if (C::TM == 1) { // RK weights into this thread's TMEM slot: columns {0,4,8} of its warp's 16
            if (First || Last) tmem_st2(tmW, w0[0], w0[1]);
            if (!Last) {
               tmem_st2(tmW + 4, w1[0], w1[1]);
               tmem_st2(tmW + 8, w2[0], w2[1]);
            }
            tmem_wait_st();
         } else {
            if (First || Last) S.W[0][tid] = make_double2(w0[0], w0[1]);
            if (!Last) {
               S.W[1][tid] = make_double2(w1[0], w1[1]);
               S.W[2][tid] = make_double2(w2[0], w2[1]);
            }
         }
      }
      if constexpr (C::PAIR) {
         // ---- two tracers per barrier, ONE register set for the RK operands: within an iteration the four
         // passes alternate B(a'), A(a), B(b'), A(b) (a', b' = the previous pair), so the cur / next values of
         // b' load while pass A of a runs and those of the next iteration's a' while pass A of b runs
         const int NP = (T + 1) >> 1;
         double tcur[2] = {0, 0}, acc[2] = {0, 0}, adv[2][2] = {{0, 0}, {0, 0}}, tsp[2][2] = {{0, 0}, {0, 0}};
         auto preload = [&](int l) {
            if (own && l < T) {
               if (need_cur) ldg_na(tcur, P.TrCur + (size_t)l * CS + o);
               if (!First) ldg_na(acc, P.TrNext + (size_t)l * CS + o);
            }
         };
         auto pass_A = [&](int l, int q, int db) { // every neighbour row once, two sums
            const unsigned char *Pb = reinterpret_cast<const unsigned char *>(&S.Phi[(it0 + l) % NS][0][0]);
            const double2 ts = lds2(Pb, myoff);
            double d0 = 0, d1 = 0, a0 = ts.x * S0[0], a1 = ts.y * S0[1];
#pragma unroll
            for (int j = 0; j < kME; ++j) {
               const double2 tn = lds2(Pb, NBOFF(j));
               a0 += E[j][0] * tn.x;
               a1 += E[j][1] * tn.y;
               d0 += cfA[j][0] * (ts.x - tn.x);
               d1 += cfA[j][1] * (ts.y - tn.y);
            }
            if (own) sts2(reinterpret_cast<unsigned char *>(&S.D2[db][0][0]), myoff, d0, d1);
            adv[q][0] = a0;
            adv[q][1] = a1;
            if (tcur_is_ts) { // first stage on the current state: cur is the stage value itself
               tsp[q][0] = ts.x;
               tsp[q][1] = ts.y;
            }
         };
         auto pass_B = [&](int l, int q, int db) {
            const unsigned char *Db = reinterpret_cast<const unsigned char *>(&S.D2[db][0][0]);
            const double2 ds = lds2(Db, myoff);
            const double SHl = S.SHc[li];
            double tend[2] = {adv[q][0] + ds.x * SHl, adv[q][1] + ds.y * SHl};
#pragma unroll
            for (int j = 0; j < kME; j += 2) {
               const double2 c2 = *reinterpret_cast<const double2 *>(&S.Ch[li][j]);
               const double2 da = lds2(Db, NBOFF(j)), dc = lds2(Db, NBOFF(j + 1));
               tend[0] += c2.x * da.x;
               tend[1] += c2.x * da.y;
               tend[0] += c2.y * dc.x;
               tend[1] += c2.y * dc.y;
            }
            if (own) {
               double out[2];
               double2 w0v = make_double2(0, 0);
               if (First || Last) w0v = S.W[0][tid];
               const double w0[2] = {w0v.x, w0v.y};
               const double tc[2] = {tcur_is_ts ? tsp[q][0] : tcur[0], tcur_is_ts ? tsp[q][1] : tcur[1]};
#pragma unroll
               for (int x = 0; x < 2; ++x) {
                  out[x] = (First ? tc[x] * w0[x] : acc[x]) + P.B * tend[x];
                  if (Last) out[x] *= w0[x];
               }
               stg_na(P.TrNext + l * CS + o, out);
               if (!Last) {
                  const double2 w1v = S.W[1][tid], w2v = S.W[2][tid];
                  out[0] = tc[0] * w1v.x + tend[0] * w2v.x;
                  out[1] = tc[1] * w1v.y + tend[1] * w2v.y;
                  stg_na(P.TrProv + l * CS + o, out);
               }
            }
         };
         preload(0);
         for (int pi = 0; pi <= NP; ++pi) {
            __syncthreads(); // pair pi has landed; Del2 of pair pi-1 is complete (own rows: T role, ring 1: R role)
            const int a = 2 * pi, ap = a - 2, dbp = 2 * ((pi - 1) & 1), dbn = 2 * (pi & 1);
            if (pi > 0) {
               pass_B(ap, 0, dbp);
               preload(ap + 1); // second tracer of the previous pair: lands while pass A of a runs
            }
            if (pi < NP) pass_A(a, 0, dbn);
            if (pi > 0 && ap + 1 < T) pass_B(ap + 1, 1, dbp + 1);
            if (pi < NP) {
               preload(a); // first pass B of the next iteration
               if (a + 1 < T) pass_A(a + 1, 1, dbn + 1);
            }
         }
      } else {
         double tcur[2] = {0, 0}, acc[2] = {0, 0}; // cur / next of the tracer pass B handles next
         if (own) {
            if (need_cur) ldg_na(tcur, P.TrCur + o);
            if (!First) ldg_na(acc, P.TrNext + o);
         }
         double adv[2] = {0, 0}, tsp[2] = {0, 0}; // advective part and own value of the tracer pass A handled last
         // ---- pass B of tracer l: Del4 from the Del2 rows, then the RK update
         auto pass_B = [&](int l) {
            const unsigned char *Db = reinterpret_cast<const unsigned char *>(&S.D2[l & 1][0][0]);
            const double2 ds = lds2(Db, myoff);
   #ifdef OMEGA_EXP_NOPRIV
            const double SHl = 1e-9;
   #else
            const double SHl = S.SHc[li];
   #endif
            double tend[2] = {adv[0] + ds.x * SHl, adv[1] + ds.y * SHl};
   #pragma unroll
            for (int j = 0; j < kME; j += 2) {
   #ifdef OMEGA_EXP_NOPRIV
               const double2 c2 = make_double2(1e-9 * j, 1e-9);
   #else
               const double2 c2 = *reinterpret_cast<const double2 *>(&S.Ch[li][j]);
   #endif
               const double2 da = lds2(Db, NBOFF(j)), db = lds2(Db, NBOFF(j + 1));
               tend[0] += c2.x * da.x;
               tend[1] += c2.x * da.y;
               tend[0] += c2.y * db.x;
               tend[1] += c2.y * db.y;
            }
            // TMEM loads are warp-collective (.sync.aligned): issued by every thread, before the own-cell branch
            unsigned tw0[4] = {0, 0, 0, 0}, tw1[4] = {0, 0, 0, 0}, tw2[4] = {0, 0, 0, 0};
            if (C::TM) {
               if (First || Last) tmem_ld2_issue(tmW, tw0);
               if (!Last) {
                  tmem_ld2_issue(tmW + 4, tw1);
                  tmem_ld2_issue(tmW + 8, tw2);
               }
               tmem_wait_ld();
            }
            if (own) {
               double out[2];
               double2 w0v = make_double2(0, 0);
               if (First || Last) w0v = C::TM ? tmem_unpack2(tw0) : S.W[0][tid];
               const double w0[2] = {w0v.x, w0v.y};
               if (tcur_is_ts) {
                  tcur[0] = tsp[0];
                  tcur[1] = tsp[1];
               }
   #pragma unroll
               for (int q = 0; q < 2; ++q) {
                  out[q] = (First ? tcur[q] * w0[q] : acc[q]) + P.B * tend[q];
                  if (Last) out[q] *= w0[q];
               }
               stg_na(P.TrNext + l * CS + o, out);
               if (!Last) {
   #ifdef OMEGA_EXP_NOPRIV
                  const double2 w1v = make_double2(1.0, 1.0), w2v = make_double2(P.A, P.A);
   #else
                  const double2 w1v = C::TM ? tmem_unpack2(tw1) : S.W[1][tid], w2v = C::TM ? tmem_unpack2(tw2) : S.W[2][tid];
   #endif
                  const double w1[2] = {w1v.x, w1v.y}, w2[2] = {w2v.x, w2v.y};
   #pragma unroll
                  for (int q = 0; q < 2; ++q) out[q] = tcur[q] * w1[q] + tend[q] * w2[q];
                  stg_na(P.TrProv + l * CS + o, out);
               }
               if (l + 1 < T) {
                  if (need_cur) ldg_na(tcur, P.TrCur + (size_t)(l + 1) * CS + o);
                  if (!First) ldg_na(acc, P.TrNext + (size_t)(l + 1) * CS + o);
               }
            }
         };
         for (int i = 0; i < T; ++i) {
            __syncthreads(); // tracer i has landed; Del2 of tracer i-1 is complete (own rows: T role, ring 1: R role)
            double advn[2], tsn[2];
            { // ---- pass A of tracer i: every neighbour row once, two sums
               wait_item<C>(S, it0 + i);
               const unsigned char *Pb = reinterpret_cast<const unsigned char *>(&S.Phi[(it0 + i) % NS][0][0]);
               const double2 ts = lds2(Pb, myoff);
               double d0 = 0, d1 = 0;
               advn[0] = ts.x * S0[0];
               advn[1] = ts.y * S0[1];
   #pragma unroll
               for (int j = 0; j < kME; ++j) {
                  const double2 tn = lds2(Pb, NBOFF(j));
                  advn[0] += E[j][0] * tn.x;
                  advn[1] += E[j][1] * tn.y;
                  d0 += cfA[j][0] * (ts.x - tn.x);
                  d1 += cfA[j][1] * (ts.y - tn.y);
               }
               if (own) sts2(reinterpret_cast<unsigned char *>(&S.D2[i & 1][0][0]), myoff, d0, d1);
               tsn[0] = ts.x;
               tsn[1] = ts.y;
            }
            if (i > 0) pass_B(i - 1);
            adv[0] = advn[0]; adv[1] = advn[1];
            tsp[0] = tsn[0]; tsp[1] = tsn[1];
         }
         // last iteration: only pass B of the last tracer is left, the coefficient registers are free:
         // the next chunk's prologue operands start their way from global memory now
         __syncthreads();
         if (OMEGA_T2_HOIST && ch + 1 < NCh) load_prologue(ch + 1);
         pass_B(T - 1);
      }
#undef NBOFF
   }
}

template <class C, int STAGE, bool OPT, bool PART = false>
__global__ void __launch_bounds__(C::NT, C::CTAs)
k_fCT2(DMesh M, CellArgs P, int NChAll, int Pf, int NWhole, TileRange TR, const __grid_constant__ CUtensorMap MapH,
       const __grid_constant__ CUtensorMap MapT) {
   extern __shared__ __align__(16) unsigned char smem_raw[];
   typename C::Smem &S = *reinterpret_cast<typename C::Smem *>(smem_raw);
   // A kernel queued behind this one with programmatic stream serialization (fused.cu: k_fA, or the edge pass of
   // the interior tiles -- neither reads what this pass writes) may start on the SMs this grid's last round of
   // CTAs leaves idle: every CTA has started once all have signalled.  Without such a launch this is a no-op.
   asm volatile("griddepcontrol.launch_dependents;");
   // CTAs below NWhole take a whole tile (all NChAll chunks); the tiles of the last, partial round of
   // CTAs are split into one CTA per chunk so that the tail of the grid costs a fraction of a round
   const int tid = threadIdx.x;
   int tile = blockIdx.x, kb = 0, NCh = NChAll;
   if ((int)blockIdx.x >= NWhole) {
      const int q = blockIdx.x - NWhole;
      tile = NWhole + q / NChAll;
      kb   = (q % NChAll) * C::KC;
      NCh  = 1;
   }
   tile = TR.tile(tile);
   if (tid < (int)(sizeof(TileHdr) / sizeof(int)))
      reinterpret_cast<int *>(&S.Hdr)[tid] = reinterpret_cast<const int *>(M.Tiles.Hdr + tile)[tid];
   if (C::BULK && tid == 32) {
      for (int b = 0; b <= C::NS; ++b) mbar_init(smem_u32(&S.Bar[b]), 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      if constexpr (C::TMAP) {
         S.MapH = &MapH; S.MapT = &MapT;
         S.HsBase = P.Hs; S.TrsBase = P.Trs;
      }
   }
   if (C::TM) {
      if (tid < 32) tmem_alloc_all(&S.TmemBase); // one CTA per SM: the whole tensor memory
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
   }
   __syncthreads();
   if (C::TM) asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
   TileCtx X;
   X.c0 = S.Hdr.C0; X.nc = S.Hdr.NC; X.n1 = S.Hdr.N1; X.nR1 = X.nc + X.n1; X.nR = X.nR1 + S.Hdr.N2;
   for (int r = tid; r < X.nR; r += C::NT) S.GRow[r] = r < X.nc ? X.c0 + r : M.Tiles.HaloRows[S.Hdr.RowOff + r - X.nc];
   __syncthreads();
   if (tid >= C::TT) {
      if constexpr (C::PAIR) {
         asm volatile("setmaxnreg.dec.sync.aligned.u32 48;"); // releases exactly what the T role's 96 need (768 x 80 at launch)
         role_R3<C>(M, P, S, X.c0, X.nc, X.n1, X.nR, NCh, tid, Pf != 0, kb);
      } else {
         if (C::RS == 2) asm volatile("setmaxnreg.dec.sync.aligned.u32 48;");
         else asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
         role_R<C, PART>(M, P, S, X.c0, X.nc, X.n1, X.nR, NCh, tid, Pf != 0, kb);
      }
   } else {
      asm volatile("setmaxnreg.inc.sync.aligned.u32 96;");
      X.t = tid; X.i = tid >> C::LS; X.p = tid & (C::L - 1);
      X.own = X.i < X.nc; X.li = X.own ? X.i : 0;
      X.myoff = (X.li * C::KC + 2 * X.p) * 8u;
      role_T2<C, STAGE, OPT, PART>(M, P, S, X, NCh, tid, kb);
   }
   if (C::TM) {
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncthreads();
      if (tid < 32) tmem_free_all(S.TmemBase);
   }
}

// tensor maps of the TMAP row copies: a [rows][K] fp64 array read in boxes of one row x KC levels (four such
// rows per tile::gather4 instruction).  The encoder is a driver entry point, looked up once.
static int encode_row_map(CUtensorMap *Map, const double *Base, long long Rows, int K, int KC) {
   typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
   static EncodeFn enc = nullptr;
   if (!enc) {
      cudaDriverEntryPointQueryResult qr;
      void *fn = nullptr;
      OMEGA_CUDA_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr));
      OMEGA_REQUIRE(fn && qr == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled is not available in this driver");
      enc = (EncodeFn)fn;
   }
   const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)Rows}, strides[1] = {(cuuint64_t)K * 8};
   const cuuint32_t box[2] = {(cuuint32_t)KC, 1}, es[2] = {1, 1};
   const CUresult rc = enc(Map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<double *>(Base), dims, strides, box, es,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
   OMEGA_REQUIRE(rc == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (" + std::to_string((int)rc) + ")");
   return 0;
}

template <class C, int STAGE, bool OPT> static int launch_fCT2_o(const DMesh &D, const CellArgs &ca, cudaStream_t St, const TileRange &TR) {
   static DevOnce once;
   int sm = 0;
   if (set_smem_once(once, k_fCT2<C, STAGE, OPT>, sizeof(typename C::Smem), &sm)) return 1;
   const int pf = env_i("OMEGA_TILE_PF", 1);
   // one CTA per tile; the tiles of the last partial round of CTAs are split into one CTA per chunk when
   // that shortens the tail (matters on multi-GPU partitions with ~10 rounds)
   const int split = env_i("OMEGA_TILE_SPLIT", 1);
   const int nch = (D.K + C::KC - 1) / C::KC, nt = TR.Count; // the last chunk may be partial (even level counts)
   OMEGA_REQUIRE(D.K % C::KC == 0 || (!C::BULK && !C::PAIR && !C::TM), "k_fCT2: a partial last chunk needs the default copy path");
   if (nt <= 0) return 0;
   int whole = nt;
   if (split && nch > 1) {
      const int tail = nt % sm;
      const double cost = ((tail * nch + sm - 1) / sm) * 1.15 / nch; // rounds, a single-chunk CTA costs ~15 % more per chunk
      if (tail > 0 && cost < 0.85) whole = nt - tail;
   }
   CUtensorMap mapH, mapT;
   std::memset(&mapH, 0, sizeof(mapH));
   std::memset(&mapT, 0, sizeof(mapT));
   if constexpr (C::TMAP) {
      if (encode_row_map(&mapH, ca.Hs, D.NCellsSize, D.K, C::KC)) return 1;
      if (encode_row_map(&mapT, ca.Trs, (long long)ca.NTracers * D.NCellsSize, D.K, C::KC)) return 1;
   }
   if constexpr (!C::BULK && !C::PAIR && !C::TM) {
      if (D.K % C::KC != 0) { // the last chunk is partial: the masked instantiation
         static DevOnce oncePart;
         if (set_smem_once(oncePart, k_fCT2<C, STAGE, OPT, true>, sizeof(typename C::Smem))) return 1;
         k_fCT2<C, STAGE, OPT, true><<<(unsigned)(whole + (nt - whole) * nch), C::NT, sizeof(typename C::Smem), St>>>(D, ca, nch, pf, whole, TR, mapH, mapT);
         return 0;
      }
   }
   k_fCT2<C, STAGE, OPT><<<(unsigned)(whole + (nt - whole) * nch), C::NT, sizeof(typename C::Smem), St>>>(D, ca, nch, pf, whole, TR, mapH, mapT);
   return 0;
}
template <class C, int STAGE> static int launch_fCT2_c(const DMesh &D, const CellArgs &ca, cudaStream_t St, const TileRange &TR) {
   return ca.FluxUpwind || ca.TrUpwind || ca.SrcH ? launch_fCT2_o<C, STAGE, true>(D, ca, St, TR) : launch_fCT2_o<C, STAGE, false>(D, ca, St, TR);
}
template <int STAGE> static int launch_fCT2_t(const DMesh &D, const CellArgs &ca, cudaStream_t St, const TileRange &TR) {
   // measured variants, selectable for profiling (profiles/r01e_ncu_full_summary.md): register split /
   // R-role width (OMEGA_TILE_RS = 0, 2, 3), bulk row copies (OMEGA_TILE_CP = 1, 2) and tensor-map gather4 copies
   // (OMEGA_TILE_CP = 3: 7.33 ms per launch against 5.44 -- the copy engine is slow on 128-byte row pieces).  The 8-level shape
   // Cfg2<8> (two CTAs per SM) is not instantiated: its 64-byte row pieces conflict in the banks.
   const int rs = env_i("OMEGA_TILE_RS", 2), cp = env_i("OMEGA_TILE_CP", 0), tm = env_i("OMEGA_TILE_TM", 0);
   // k_fCT3 (OMEGA_TILE_V3=1): two tracers per iteration, ring coefficients in tensor memory.  Parity-green
   // but measured slower on config 3 (5.97 ms per launch against 5.47; a first form with two register sets
   // of RK operands: 6.75, T-role spills): the R role -- half the warps, all copies, tcgen05.ld + wait in
   // front of every ring Del2 sum -- becomes the critical path (ncu r02h: 16 % of the samples are T warps at
   // the pair barrier); see profiles/r02_experiments.md
   if (env_i("OMEGA_TILE_V3", 0)) return launch_fCT2_c<Cfg3, STAGE>(D, ca, St, TR);
   if (tm == 1) return launch_fCT2_c<Cfg2<16, 2, 0, 1>, STAGE>(D, ca, St, TR);
   if (cp == 1) return launch_fCT2_c<Cfg2<16, 2, 1>, STAGE>(D, ca, St, TR);
   if (cp == 2) return launch_fCT2_c<Cfg2<16, 2, 2>, STAGE>(D, ca, St, TR);
   if (cp == 3) return launch_fCT2_c<Cfg2<16, 2, 3>, STAGE>(D, ca, St, TR);
   return rs == 3 ? launch_fCT2_c<Cfg2<16, 3>, STAGE>(D, ca, St, TR) : rs == 2 ? launch_fCT2_c<Cfg2<16, 2>, STAGE>(D, ca, St, TR) : launch_fCT2_c<Cfg2<16, 0>, STAGE>(D, ca, St, TR);
}
size_t tile_smem_bytes() { return sizeof(TileSmem); }

template <bool HYP, int STAGE, int RS> static int launch_fCT_r(const DMesh &D, const CellArgs &ca, cudaStream_t St, const TileRange &TR) {
   static DevOnce once;
   if (set_smem_once(once, k_fCT<HYP, STAGE, RS>, sizeof(TileSmem))) return 1;
   OMEGA_REQUIRE(!HYP || ca.NTracers >= 2, "k_fCT with tracer hyperdiffusion needs at least two tracers");
   OMEGA_REQUIRE(D.K % kTKC == 0, "k_fCT needs a level count that is a multiple of 16");
   const int nch = D.K / kTKC;
   // with the D role (which needs two tracers to prime its pipeline) a CTA walks all chunks of its tile
   const int persist = env_i("OMEGA_TILE_PERSIST", 1);
   const bool walk = HYP && persist;
   if (TR.Count <= 0) return 0;
   const unsigned grid = (unsigned)TR.Count * (walk ? 1 : nch);
   k_fCT<HYP, STAGE, RS><<<grid, HYP ? 2 * kTRole : kTRole, sizeof(TileSmem), St>>>(D, ca, walk ? -nch : nch, TR);
   return 0;
}
template <bool HYP, int STAGE> static int launch_fCT_t(const DMesh &D, const CellArgs &ca, cudaStream_t St, const TileRange &TR) {
   const int rs = env_i("OMEGA_TILE_REGS", 0);
   if (HYP && rs == 1) return launch_fCT_r<HYP, STAGE, HYP ? 1 : 0>(D, ca, St, TR);
   if (HYP && rs == 2) return launch_fCT_r<HYP, STAGE, HYP ? 2 : 0>(D, ca, St, TR);
   return launch_fCT_r<HYP, STAGE, 0>(D, ca, St, TR);
}

int launch_fCT(const DMesh &D, const CellArgs &ca, cudaStream_t St, const TileRange &TR) {
   OMEGA_REQUIRE(!(ca.First && ca.Last), "launch_fCT: a stage cannot be both first and last");
   const int stage = ca.First ? 0 : (ca.Last ? 2 : 1);
   if (tile_v2_used(ca))
      return stage == 0 ? launch_fCT2_t<0>(D, ca, St, TR) : stage == 1 ? launch_fCT2_t<1>(D, ca, St, TR) : launch_fCT2_t<2>(D, ca, St, TR);
   if (ca.HypOn) return stage == 0 ? launch_fCT_t<true, 0>(D, ca, St, TR) : stage == 1 ? launch_fCT_t<true, 1>(D, ca, St, TR) : launch_fCT_t<true, 2>(D, ca, St, TR);
   return stage == 0 ? launch_fCT_t<false, 0>(D, ca, St, TR) : stage == 1 ? launch_fCT_t<false, 1>(D, ca, St, TR) : launch_fCT_t<false, 2>(D, ca, St, TR);
}

// Level counts: any even K.  A K that is not a multiple of 16 ends in a partial chunk, which k_fCT2 and k_fET mask
// (tile_partial_chunk); the older k_fCT (tracer hyperdiffusion off) needs whole chunks.
bool tile_path_ok(const DMesh &D) { return D.Tiles.NTiles > 0 && D.K % 2 == 0 && D.K >= 2; }
bool tile_partial_chunk(const DMesh &D) { return D.K % kTKC != 0; }
bool tile_v2_enabled() { return env_i("OMEGA_TILE_V2", 1) != 0; }
// k_fCT2 serves every tracer count with tracer hyperdiffusion on
bool tile_v2_used(const CellArgs &ca) { return ca.HypOn && tile_v2_enabled() && ca.NTracers >= 1; }

// ---------------------------------------------------------------------------
// k_fET: shared-memory staged edge pass (the work of k_fE, fused.cu).  A CTA owns the
// edges of one tile x 16 levels.  PotentialVortHAdvOnEdge (TendencyTerms.h:81-108) sums
// w_j * F(e_j) * (q(e) + q(e_j))/2 over up to 10 neighbour edges, with
//   F(e) = FluxLayerThickEdge(e) * u(e)   (LayerThicknessAuxVars.h:25-61, centred)
//   q(e) = NormRelVortEdge + NormPlanetVortEdge   (VorticityAuxVars.h:61-76)
// each needing five gathered rows: pass 1 evaluates F and q ONCE per staged edge of the
// tile (every edge of the own + ring-1 cells) into shared memory, pass 2 does the sum for
// the own edges from shared memory through precomputed local indices.  The gradient and
// diffusion terms difference the combined fields X, Y, Z that fA / fB write (see there),
// six gathered rows instead of twelve.  updateVelocityByTend (TimeStepper.cpp:407-430) is
// the epilogue, as in k_fE.
// ---------------------------------------------------------------------------
constexpr int kEThreads = 512;

struct ETileSmem {
   double F[kTileEdges][kTKC];
   double Q[kTileEdges][kTKC];
   int SRow[kTileEdges];
   int4 CV[kTileEdges]; // cells (x,y) and vertices (z,w) of every staged edge
   unsigned short EoE[kTileEdgesOwn][10];
   TileHdr Hdr;
};

template <bool OPT, bool PART = false> // OPT: upwinded flux thickness, manufactured velocity source; PART: masked last chunk
__global__ void __launch_bounds__(kEThreads, 2)
k_fET(DMesh M, EdgeTileArgs P, int NChunks, int Pf, int Ahead, TileRange TR) {
   extern __shared__ __align__(16) unsigned char smem_raw[];
   ETileSmem &S = *reinterpret_cast<ETileSmem *>(smem_raw);
   const int tid = threadIdx.x;
   const int tix = blockIdx.x / NChunks, kc0 = (blockIdx.x % NChunks) * kTKC;
   const int tile = TR.tile(tix);
   const int K = M.K;
   if (tid < (int)(sizeof(TileHdr) / sizeof(int)))
      reinterpret_cast<int *>(&S.Hdr)[tid] = reinterpret_cast<const int *>(M.Tiles.Hdr + tile)[tid];
   __syncthreads();
   const int nOwn = S.Hdr.NEOwn, nSt = S.Hdr.NEStaged;
   for (int r = tid; r < nSt; r += kEThreads) {
      S.SRow[r] = M.Tiles.EdgeRows[S.Hdr.ERowOff + r];
      S.CV[r]   = M.Tiles.EdgeCV[S.Hdr.ERowOff + r];
   }
   {
      const unsigned *src = reinterpret_cast<const unsigned *>(M.Tiles.EoELocal + (size_t)S.Hdr.EoEOff * 10);
      unsigned *dst       = reinterpret_cast<unsigned *>(&S.EoE[0][0]);
      for (int q = tid; q < nOwn * 5; q += kEThreads) dst[q] = src[q];
   }
   __syncthreads();
   // (last chunk of a level count that is not a multiple of 16: lanes behind level K load row K-2 and store nothing)
   const int p = tid & 7, kraw = kc0 + 2 * p, k = PART ? min(kraw, K - 2) : kraw;
   const bool kin = !PART || kraw < K;
   if (Pf & 1) {
      // the eight gathered rows of pass 2a (X, Y of both cells, Z of both vertices, u_cur, u_next) of this
      // thread's own edges: one of the eight lanes of an edge each, pulled into L2 while pass 1 runs
#pragma unroll
      for (int it = 0; it < kTileEdgesOwn * kTLanes / kEThreads; ++it) {
         const int r = (tid >> 3) + it * (kEThreads / kTLanes);
         if (r < nOwn) {
            const int4 cv = S.CV[r];
            const double *q;
            switch (p) {
            case 0: q = P.X + (size_t)cv.x * K; break;
            case 1: q = P.X + (size_t)cv.y * K; break;
            case 2: q = P.Y + (size_t)cv.x * K; break;
            case 3: q = P.Y + (size_t)cv.y * K; break;
            case 4: q = P.Z + (size_t)cv.z * K; break;
            case 5: q = P.Z + (size_t)cv.w * K; break;
            case 6: q = (P.First && P.Ucur == P.Us ? P.Us : P.Ucur) + (size_t)S.SRow[r] * K; break;
            default: q = P.Unext + (size_t)S.SRow[r] * K; break;
            }
            if (p >= 2 || P.XOn) prefetch_l2(q + kc0);
         }
      }
   }
   // ---- pass 1: F and q of every staged edge, two edges per trip (ten gathered rows in flight)
   auto fq = [&](int r, double (&u)[2], double (&h0)[2], double (&h1)[2], double (&a)[2], double (&b)[2]) {
      const int e   = S.SRow[r];
      const int4 cv = S.CV[r];
      ldv<2>(u, P.Us + (size_t)e * K + k);
      ldv<2>(h0, P.Hs + (size_t)cv.x * K + k);
      ldv<2>(h1, P.Hs + (size_t)cv.y * K + k);
      ldv<2>(a, P.PVv + (size_t)cv.z * K + k);
      ldv<2>(b, P.PVv + (size_t)cv.w * K + k);
   };
   auto put = [&](int r, const double (&u)[2], const double (&h0)[2], const double (&h1)[2], const double (&a)[2],
                  const double (&b)[2]) {
      double f0 = 0.5 * (h0[0] + h1[0]), f1 = 0.5 * (h0[1] + h1[1]);
      if (OPT && P.FluxUpwind) { // u == 0: F = flux * u = 0 for either choice (the reference takes max(h0, h1))
         f0 = u[0] > 0 ? h0[0] : h1[0];
         f1 = u[1] > 0 ? h0[1] : h1[1];
      }
      *reinterpret_cast<double2 *>(&S.F[r][2 * p]) = make_double2(f0 * u[0], f1 * u[1]);
      *reinterpret_cast<double2 *>(&S.Q[r][2 * p]) = make_double2(0.5 * (a[0] + b[0]), 0.5 * (a[1] + b[1]));
   };
   for (int q = tid; q < nSt * kTLanes; q += 2 * kEThreads) {
      const int r = q >> 3, r2 = r + kEThreads / kTLanes;
      double u[2], h0[2], h1[2], a[2], b[2], u2[2], g0[2], g1[2], a2[2], b2[2];
      fq(r, u, h0, h1, a, b);
      if (r2 < nSt) fq(r2, u2, g0, g1, a2, b2);
      put(r, u, h0, h1, a, b);
      if (r2 < nSt) put(r2, u2, g0, g1, a2, b2);
   }
   // ---- pass 2a: gradient / diffusion terms and RK operands of the own edges.  A thread owns
   // the same (edge, piece) items as in pass 1, at most kTileEdgesOwn*8/512 = 4 of them; all
   // their gathered rows are in flight together, and what survives the barrier is
   //   p1 = (first stage ? u_cur : u_next) + B*m*G,   p2 = u_cur + A*m*G     (two registers pairs per item)
   constexpr int kItems = kTileEdgesOwn * kTLanes / kEThreads;
   double p1[kItems][2], p2[kItems][2];
   const bool ucur_is_us = P.First && P.Ucur == P.Us;
#pragma unroll
   for (int it = 0; it < kItems; ++it) {
      const int r = (tid >> 3) + it * (kEThreads / kTLanes);
      p1[it][0] = p1[it][1] = p2[it][0] = p2[it][1] = 0.0;
      if (r < nOwn) {
         const int e    = S.SRow[r];
         const size_t o = (size_t)e * K + k;
         const int4 cv  = S.CV[r];
         const double mask = M.EdgeMask[e];
         const double InvDc = M.InvDcEdge[e] * mask, InvDv = M.InvDvEdge[e] * mask;
         const double srcu = OPT && P.SrcU ? P.SrcU[e] : 0.0; // manufactured velocity source (not masked, CustomTendencyTerms.cpp:145-210)
         if (P.PVOn) prefetch_row(M.WeightsOnEdge + (size_t)e * M.MaxEdges2 + p, 1); // weights towards L1 for pass 2b
         double x0[2] = {0, 0}, x1[2] = {0, 0}, y0[2], y1[2], z0[2], z1[2], ucur[2] = {0, 0}, base[2] = {0, 0};
         if (P.XOn) { // off when k_fB3 already folded X into Y (W = Y - X)
            ldv<2>(x0, P.X + (size_t)cv.x * K + k);
            ldv<2>(x1, P.X + (size_t)cv.y * K + k);
         }
         ldv<2>(y0, P.Y + (size_t)cv.x * K + k);
         ldv<2>(y1, P.Y + (size_t)cv.y * K + k);
         ldv<2>(z0, P.Z + (size_t)cv.z * K + k);
         ldv<2>(z1, P.Z + (size_t)cv.w * K + k);
         if (P.First || !P.Last) ldp<2>(ucur, (ucur_is_us ? P.Us : P.Ucur) + o);
         if (!P.First) ldp<2>(base, P.Unext + o);
#pragma unroll
         for (int q2 = 0; q2 < 2; ++q2) {
            double g = (P.SY * (y1[q2] - y0[q2])) * InvDc - (P.SZ * (z1[q2] - z0[q2])) * InvDv;
            if (P.XOn) g -= (x1[q2] - x0[q2]) * InvDc;
            if (OPT) g += srcu;
            p1[it][q2] = (P.First ? ucur[q2] : base[q2]) + P.B * g;
            p2[it][q2] = ucur[q2] + P.A * g;
         }
      }
   }
   __syncthreads();
   // ---- pass 2b: neighbour sum of the own edges from shared memory (always 10 entries: the
   // weight table is zero behind the edge's count and the local index then points at the edge
   // itself), then the update
   const unsigned char *Fb = reinterpret_cast<const unsigned char *>(&S.F[0][0]);
   const unsigned char *Qb = reinterpret_cast<const unsigned char *>(&S.Q[0][0]);
   const unsigned po = (unsigned)(2 * p) * 8u;
#pragma unroll
   for (int it = 0; it < kItems; ++it) {
      const int r = (tid >> 3) + it * (kEThreads / kTLanes);
      if (r < nOwn && kin) {
         const int e    = S.SRow[r];
         const size_t o = (size_t)e * K + k;
         double acc0 = 0, acc1 = 0;
         if (P.PVOn) {
            const double2 qe = lds2(Qb, r * (kTKC * 8u) + po);
            const double2 *W = reinterpret_cast<const double2 *>(M.WeightsOnEdge + (size_t)e * M.MaxEdges2);
#pragma unroll
            for (int j2 = 0; j2 < 5; ++j2) {
               const double2 w = __ldg(W + j2);
               const unsigned ix = *reinterpret_cast<const unsigned *>(&S.EoE[r][2 * j2]);
               const unsigned la = (ix & 0xffffu) * (kTKC * 8u) + po, lb = (ix >> 16) * (kTKC * 8u) + po;
               const double2 fa = lds2(Fb, la), qa = lds2(Qb, la), fb = lds2(Fb, lb), qb = lds2(Qb, lb);
               acc0 += w.x * fa.x * ((qe.x + qa.x) * 0.5);
               acc1 += w.x * fa.y * ((qe.y + qa.y) * 0.5);
               acc0 += w.y * fb.x * ((qe.x + qb.x) * 0.5);
               acc1 += w.y * fb.y * ((qe.y + qb.y) * 0.5);
            }
            const double mask = M.EdgeMask[e];
            acc0 *= mask;
            acc1 *= mask;
         }
         double out[2] = {p1[it][0] + P.B * acc0, p1[it][1] + P.B * acc1};
         stv<2>(P.Unext + o, out);
         if (!P.Last) {
            out[0] = p2[it][0] + P.A * acc0;
            out[1] = p2[it][1] + P.A * acc1;
            stv<2>(P.Uprov + o, out);
         }
      }
   }
}

int launch_fET(const DMesh &D, const EdgeTileArgs &ea, cudaStream_t St, const TileRange &TR, bool Overlap) {
   if (TR.Count <= 0) return 0;
   static DevOnce once, onceOpt;
   int sm = 0;
   const bool opt = ea.FluxUpwind || ea.SrcU;
   if (opt ? set_smem_once(onceOpt, k_fET<true>, sizeof(ETileSmem), &sm) : set_smem_once(once, k_fET<false>, sizeof(ETileSmem), &sm)) return 1;
   const int nch = (D.K + kTKC - 1) / kTKC; // the last chunk may be partial
   // L2 prefetch of the CTA's own pass-2a rows.  (Also tried: the pass-1 rows of the tile a CTA of the NEXT wave
   // takes, requested one wave ahead -- 2.36 -> 2.85 ms on config 3, the extra requests cost more than the
   // latency they hide; removed.)
   const int pf = env_i("OMEGA_TILE_EPF", 1);
   const int ahead = env_i("OMEGA_TILE_EAHEAD", (2 * sm + nch - 1) / nch);
   if (D.K % kTKC != 0) { // the last chunk is partial: the masked instantiations
      static DevOnce oncePart, oncePartOpt;
      if (opt ? set_smem_once(oncePartOpt, k_fET<true, true>, sizeof(ETileSmem)) : set_smem_once(oncePart, k_fET<false, true>, sizeof(ETileSmem)))
         return 1;
      if (opt) k_fET<true, true><<<(unsigned)TR.Count * nch, kEThreads, sizeof(ETileSmem), St>>>(D, ea, nch, pf, ahead, TR);
      else k_fET<false, true><<<(unsigned)TR.Count * nch, kEThreads, sizeof(ETileSmem), St>>>(D, ea, nch, pf, ahead, TR);
      return 0;
   }
   if (Overlap) {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3((unsigned)TR.Count * nch);
      cfg.blockDim = dim3(kEThreads);
      cfg.dynamicSmemBytes = sizeof(ETileSmem);
      cfg.stream = St;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      at[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = at;
      cfg.numAttrs = 1;
      if (opt) OMEGA_CUDA_CHECK(cudaLaunchKernelEx(&cfg, k_fET<true>, D, ea, nch, pf, ahead, TR));
      else OMEGA_CUDA_CHECK(cudaLaunchKernelEx(&cfg, k_fET<false>, D, ea, nch, pf, ahead, TR));
      return 0;
   }
   if (opt) k_fET<true><<<(unsigned)TR.Count * nch, kEThreads, sizeof(ETileSmem), St>>>(D, ea, nch, pf, ahead, TR);
   else k_fET<false><<<(unsigned)TR.Count * nch, kEThreads, sizeof(ETileSmem), St>>>(D, ea, nch, pf, ahead, TR);
   return 0;
}

bool tile_edge_path_ok(const DMesh &D) { return tile_path_ok(D) && D.Tiles.EdgeTilesOK; }

} // namespace omega
#endif // OMEGA_SINGLE_PRECISION
