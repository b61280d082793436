// mesh.cu -- device-resident HorzMesh: space-filling-curve reordering with
// explicit permutation maps, derived fields, structure-of-arrays staging, and
// permuting host<->device field transfers.
// Reference behaviour restated: components/omega/src/ocn/HorzMesh.cpp:527-647
// (computeEdgeSign, setMasks, setMeshScaling, copyToDevice).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <numeric>

#include "omega_dev.h"

namespace omega {

static thread_local std::string g_last_error;
void set_error(const std::string &msg) { g_last_error = msg; }

// 3-D Morton key, 21 bits per axis
static inline uint64_t spread3(uint64_t x) {
   x &= 0x1fffff;
   x = (x | x << 32) & 0x1f00000000ffffULL;
   x = (x | x << 16) & 0x1f0000ff0000ffULL;
   x = (x | x << 8) & 0x100f00f00f00f00fULL;
   x = (x | x << 4) & 0x10c30c30c30c30c3ULL;
   x = (x | x << 2) & 0x1249249249249249ULL;
   return x;
}

// order [lo,hi) by key (stable), append to perm
template <class Key>
static void sort_block(std::vector<int> &perm, int lo, int hi, const std::vector<Key> &key) {
   std::vector<int> idx(hi - lo);
   std::iota(idx.begin(), idx.end(), lo);
   std::stable_sort(idx.begin(), idx.end(), [&](int a, int b) { return key[a] < key[b]; });
   perm.insert(perm.end(), idx.begin(), idx.end());
}


// Recursive coordinate bisection of the index set idx into leaves of exactly
// LeafSize elements (the last leaf takes the remainder), split on the axis of
// largest extent; leaves are emitted in tree order, so the result is a
// space-filling order whose every aligned run of LeafSize rows is a compact
// patch (mean first-ring size 37 cells for 64-cell patches on the icosahedral
// mesh, against 52 for the same run length along the Morton curve).  Inside
// a leaf rows keep Morton order.
static void rcb_rec(std::vector<int> &idx, int lo, int hi, int nleaves, int LeafSize, const double *const X[3],
                    const std::vector<uint64_t> &morton, std::vector<int> &out) {
   if (hi <= lo) return;
   if (nleaves <= 1) {
      std::stable_sort(idx.begin() + lo, idx.begin() + hi, [&](int a, int b) { return morton[a] < morton[b]; });
      out.insert(out.end(), idx.begin() + lo, idx.begin() + hi);
      return;
   }
   double mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
   for (int i = lo; i < hi; ++i)
      for (int a = 0; a < 3; ++a) {
         mn[a] = std::min(mn[a], X[a][idx[i]]);
         mx[a] = std::max(mx[a], X[a][idx[i]]);
      }
   int ax = 0;
   for (int a = 1; a < 3; ++a)
      if (mx[a] - mn[a] > mx[ax] - mn[ax]) ax = a;
   const int nl  = nleaves / 2;
   const int cut = std::min(hi, lo + nl * LeafSize);
   auto cmp = [&](int a, int b) { return X[ax][a] < X[ax][b] || (X[ax][a] == X[ax][b] && a < b); };
   std::nth_element(idx.begin() + lo, idx.begin() + cut, idx.begin() + hi, cmp);
   rcb_rec(idx, lo, cut, nl, LeafSize, X, morton, out);
   rcb_rec(idx, cut, hi, nleaves - nl, LeafSize, X, morton, out);
}
static void rcb_block(std::vector<int> &perm, int lo, int hi, int LeafSize, const double *const X[3],
                      const std::vector<uint64_t> &morton) {
   std::vector<int> idx(hi - lo);
   std::iota(idx.begin(), idx.end(), lo);
   rcb_rec(idx, 0, hi - lo, (hi - lo + LeafSize - 1) / LeafSize, LeafSize, X, morton, perm);
}

template <class T> static int to_device(OmegaMesh *M, const std::vector<T> &h, const T **out) {
   void *d = nullptr;
   size_t bytes = std::max<size_t>(h.size() * sizeof(T), 16);
   OMEGA_CUDA_CHECK(cudaMalloc(&d, bytes));
   M->Allocs.push_back(d);
   if (!h.empty()) OMEGA_CUDA_CHECK(cudaMemcpy(d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
   *out = static_cast<const T *>(d);
   return 0;
}

#ifdef OMEGA_SINGLE_PRECISION
// single-precision build: the tables are evaluated in fp64 on the host and rounded once when staged
static int to_device(OmegaMesh *M, const std::vector<double> &h, const float **out) {
   std::vector<float> f(h.begin(), h.end());
   return to_device<float>(M, f, out);
}
static int to_device(OmegaMesh *M, const std::vector<double2> &h, const float2 **out) {
   std::vector<float2> f(h.size());
   for (size_t i = 0; i < h.size(); ++i) f[i] = make_float2((float)h[i].x, (float)h[i].y);
   return to_device<float2>(M, f, out);
}
#endif

// rows of a host-order table gathered into device order
template <class T>
static std::vector<T> gather_rows(const T *src, const std::vector<int> &perm, int width) {
   std::vector<T> out(perm.size() * (size_t)width);
   for (size_t i = 0; i < perm.size(); ++i)
      std::memcpy(&out[i * width], &src[(size_t)perm[i] * width], sizeof(T) * width);
   return out;
}
// same, and translate the stored indices through iperm of the target kind
static std::vector<int> gather_index_rows(const int *src, const std::vector<int> &perm, int width,
                                          const std::vector<int> &iperm_target) {
   std::vector<int> out(perm.size() * (size_t)width);
   for (size_t i = 0; i < perm.size(); ++i)
      for (int j = 0; j < width; ++j) {
         int t = src[(size_t)perm[i] * width + j];
         if (t < 0 || t >= (int)iperm_target.size()) t = (int)iperm_target.size() - 1; // boundary slot
         out[i * width + j] = iperm_target[t];
      }
   return out;
}


// ---- tiles (omega_dev.h: DTiles) -------------------------------------------
// eoc/coe/nec are the DEVICE-order tables.  Tiles cover [0,NC) in order and do
// not straddle the owned/halo boundary.
static int build_tiles(OmegaMesh *M, int NC, int NOwned, int ME, const std::vector<int> &nec,
                       const std::vector<int> &eoc, const std::vector<int> &coe,
                       const std::vector<double> &mask, const std::vector<double> &dv,
                       const std::vector<double> &dc, const std::vector<double> &area,
                       const std::vector<double> &sign, int NE, int ME2, const std::vector<int> &neoe,
                       const std::vector<int> &eoe, const std::vector<int> &voe) {
   DTiles &T = M->D.Tiles;
   std::memset(&T, 0, sizeof(T));
   if (M->MaxEdgesUsed > 6) return 0;
   auto nbr = [&](int c, int j) {
      const int e = eoc[(size_t)c * ME + j];
      const int a = coe[(size_t)e * 2], b = coe[(size_t)e * 2 + 1];
      return a == c ? b : a;
   };
   std::vector<TileHdr> hdr;
   std::vector<int> rows;
   std::vector<unsigned short> nbl;
   std::vector<int> erows;
   std::vector<unsigned short> eoel;
   std::vector<int> eloc(NE + 1, -1);
   bool edgeOK = true;
   std::vector<int> loc(NC + 1, -1); // local row of a cell in the tile under construction
   std::vector<int> r1, r2;
   int c0 = 0;
   while (c0 < NC) {
      const int blockEnd = c0 < NOwned ? NOwned : NC;
      int nc = std::min(kTileCells, blockEnd - c0);
      for (;;) {
         r1.clear();
         r2.clear();
         for (int i = 0; i < nc; ++i) loc[c0 + i] = i;
         for (int i = 0; i < nc; ++i)
            for (int j = 0; j < nec[c0 + i]; ++j) {
               const int n = nbr(c0 + i, j);
               if (loc[n] < 0) { loc[n] = 1 << 20; r1.push_back(n); }
            }
         std::sort(r1.begin(), r1.end());
         for (int n : r1)
            if (n < NC)
               for (int j = 0; j < nec[n]; ++j) {
                  const int m = nbr(n, j);
                  if (loc[m] < 0) { loc[m] = 1 << 21; r2.push_back(m); }
               }
         std::sort(r2.begin(), r2.end());
         // edge side of the same tile: own edges (lowest-ranked valid cell is an own cell) and staged
         // edges (the own edges and the EdgesOnEdge entries they reference) must fit their caps as well
         int nOwnE = 0, nStE = 0;
         {
            std::vector<int> seen;
            for (int i = 0; i < nc; ++i) {
               const int c = c0 + i;
               for (int j = 0; j < nec[c]; ++j) {
                  const int e = eoc[(size_t)c * ME + j];
                  if (eloc[e] >= 0) continue;
                  const int a = coe[(size_t)e * 2], b = coe[(size_t)e * 2 + 1];
                  const int lo = std::min(a < NC ? a : INT32_MAX, b < NC ? b : INT32_MAX);
                  if (lo >= c0 && lo < c0 + nc) {
                     eloc[e] = 0;
                     seen.push_back(e);
                     ++nOwnE;
                  }
               }
            }
            nStE = nOwnE;
            for (int q = 0; q < nOwnE; ++q) {
               const int e = seen[q];
               for (int j = 0; j < neoe[e] && j < ME2; ++j) {
                  const int je = eoe[(size_t)e * ME2 + j];
                  if (je < NE && eloc[je] < 0) {
                     eloc[je] = 0;
                     seen.push_back(je);
                     ++nStE;
                  }
               }
            }
            for (int e : seen) eloc[e] = -1;
         }
         const bool ok = (int)r1.size() <= kTileRing1 && nc + (int)r1.size() + (int)r2.size() <= kTileRows &&
                         nOwnE <= kTileEdgesOwn && nStE <= kTileEdges;
         if (ok || nc == 1) break;
         for (int i = 0; i < nc; ++i) loc[c0 + i] = -1;
         for (int n : r1) loc[n] = -1;
         for (int n : r2) loc[n] = -1;
         nc = std::max(1, nc * 3 / 4);
      }
      if ((int)r1.size() > kTileRing1 || nc + (int)r1.size() + (int)r2.size() > kTileRows) {
         set_error("build_tiles: a single cell exceeds the tile row caps");
         return 2;
      }
      for (size_t i = 0; i < r1.size(); ++i) loc[r1[i]] = nc + (int)i;
      for (size_t i = 0; i < r2.size(); ++i) loc[r2[i]] = nc + (int)r1.size() + (int)i;
      TileHdr h;
      h.C0 = c0; h.NC = nc; h.N1 = (int)r1.size(); h.N2 = (int)r2.size();
      h.RowOff = (int)rows.size(); h.NblOff = (int)(nbl.size() / 8); h.Pad0 = h.Pad1 = 0;
      // ---- edge side: own edges = edges whose lowest-ranked valid cell is an own cell; staged =
      // the own edges and every EdgesOnEdge entry of an own edge
      h.NEOwn = h.NEStaged = 0; h.ERowOff = (int)erows.size(); h.EoEOff = (int)(eoel.size() / 10);
      if (edgeOK) {
         std::vector<int> own_e, oth_e;
         for (int i = 0; i < nc; ++i) { // an own edge always has an own cell as its lowest-ranked cell
            const int c = c0 + i;
            for (int j = 0; j < nec[c]; ++j) {
               const int e = eoc[(size_t)c * ME + j];
               if (eloc[e] >= 0) continue;
               const int a = coe[(size_t)e * 2], b = coe[(size_t)e * 2 + 1];
               const int lo = std::min(a < NC ? a : INT32_MAX, b < NC ? b : INT32_MAX);
               if (lo >= c0 && lo < c0 + nc) {
                  eloc[e] = 1 << 20;
                  own_e.push_back(e);
               }
            }
         }
         std::sort(own_e.begin(), own_e.end());
         if ((int)own_e.size() > kTileEdgesOwn) edgeOK = false;
         // the other staged edges: exactly the EdgesOnEdge entries of the own edges that are not own edges
         // themselves (about 40 % of an own edge count; staging every edge of the ring-1 cells would be 80 %)
         for (size_t i = 0; edgeOK && i < own_e.size(); ++i) {
            const int e = own_e[i];
            if (neoe[e] > 10) { // cells of at most 6 edges give at most 10 entries
               edgeOK = false;
               if (getenv("OMEGA_DEBUG")) fprintf(stderr, "edge tiles off: edge %d has %d EdgesOnEdge entries\n", e, neoe[e]);
               break;
            }
            for (int j = 0; j < neoe[e]; ++j) {
               const int je = eoe[(size_t)e * ME2 + j];
               if (je < NE && eloc[je] < 0) {
                  eloc[je] = 1 << 20;
                  oth_e.push_back(je);
               }
            }
         }
         std::sort(oth_e.begin(), oth_e.end());
         if ((int)(own_e.size() + oth_e.size()) > kTileEdges) edgeOK = false;
         for (size_t i = 0; i < own_e.size(); ++i) eloc[own_e[i]] = (int)i;
         for (size_t i = 0; i < oth_e.size(); ++i) eloc[oth_e[i]] = (int)(own_e.size() + i);
         for (size_t i = 0; edgeOK && i < own_e.size(); ++i) {
            const int e = own_e[i];
            for (int j = 0; j < 10; ++j) {
               int l = (int)i;
               if (j < neoe[e]) {
                  const int je = eoe[(size_t)e * ME2 + j];
                  if (je < NE) l = eloc[je];
               } // boundary-slot entries carry zero weight: any staged row serves
               eoel.push_back((unsigned short)l);
            }
         }
         if ((int)(own_e.size() + oth_e.size()) > kTileEdges) edgeOK = false;
         h.NEOwn = (int)own_e.size(); h.NEStaged = (int)(own_e.size() + oth_e.size());
         erows.insert(erows.end(), own_e.begin(), own_e.end());
         erows.insert(erows.end(), oth_e.begin(), oth_e.end());
         for (int e : own_e) eloc[e] = -1;
         for (int e : oth_e) eloc[e] = -1;
      }
      hdr.push_back(h);
      rows.insert(rows.end(), r1.begin(), r1.end());
      rows.insert(rows.end(), r2.begin(), r2.end());
      for (int r = 0; r < nc + (int)r1.size(); ++r) {
         const int c = r < nc ? c0 + r : r1[r - nc];
         for (int j = 0; j < 8; ++j) {
            int l = r;
            if (c < NC && j < nec[c]) l = loc[nbr(c, j)];
            nbl.push_back((unsigned short)l);
         }
      }
      for (int i = 0; i < nc; ++i) loc[c0 + i] = -1;
      for (int n : r1) loc[n] = -1;
      for (int n : r2) loc[n] = -1;
      c0 += nc;
   }
   std::vector<double> g((size_t)(NC + 1) * 6, 0.0), ia(NC + 1, 0.0), cfu(g), wa(g);
   std::vector<int> eoc6((size_t)(NC + 1) * 6, NE);
   for (int c = 0; c < NC; ++c) {
      ia[c] = 1.0 / area[c];
      for (int j = 0; j < nec[c]; ++j) {
         const int e = eoc[(size_t)c * ME + j];
         const double sg = sign[(size_t)c * ME + j];
         g[(size_t)c * 6 + j]    = mask[e] * (dv[e] / dc[e]);
         cfu[(size_t)c * 6 + j]  = dv[e] * sg;
         wa[(size_t)c * 6 + j]   = mask[e] * dv[e] * sg;
         eoc6[(size_t)c * 6 + j] = e;
      }
   }
   int rc = 0;
   rc |= to_device(M, hdr, &T.Hdr);
   rc |= to_device(M, rows, &T.HaloRows);
   rc |= to_device(M, nbl, &T.Nbl);
   rc |= to_device(M, g, &T.GCell);
   rc |= to_device(M, ia, &T.InvAreaCell);
   rc |= to_device(M, cfu, &T.CfUCell);
   rc |= to_device(M, wa, &T.WaCell);
   rc |= to_device(M, eoc6, &T.EocCell);
   rc |= to_device(M, erows, &T.EdgeRows);
   {
      std::vector<int> cv(erows.size() * 4);
      for (size_t i = 0; i < erows.size(); ++i) {
         const int e = erows[i];
         cv[4 * i] = coe[(size_t)e * 2]; cv[4 * i + 1] = coe[(size_t)e * 2 + 1];
         cv[4 * i + 2] = voe[(size_t)e * 2]; cv[4 * i + 3] = voe[(size_t)e * 2 + 1];
      }
      const int *q = nullptr;
      rc |= to_device(M, cv, &q);
      T.EdgeCV = reinterpret_cast<const int4 *>(q);
   }
   rc |= to_device(M, eoel, &T.EoELocal);
   T.EdgeTilesOK = edgeOK && !rc;
   T.NTiles = rc ? 0 : (int)hdr.size();
   if (T.NTiles) {
      M->TileC0.resize(hdr.size());
      M->TileOfCell.assign(NC + 1, -1);
      M->TileOfEdge.assign(NE + 1, -1);
      for (size_t t = 0; t < hdr.size(); ++t) {
         M->TileC0[t] = hdr[t].C0;
         for (int i = 0; i < hdr[t].NC; ++i) M->TileOfCell[hdr[t].C0 + i] = (int)t;
         if (edgeOK)
            for (int i = 0; i < hdr[t].NEOwn; ++i) M->TileOfEdge[erows[hdr[t].ERowOff + i]] = (int)t;
      }
   }
   if (getenv("OMEGA_DEBUG")) {
      int mxS = 0, mxO = 0, small = 0;
      long long sumS = 0;
      for (const TileHdr &h : hdr) {
         mxS = std::max(mxS, h.NEStaged); mxO = std::max(mxO, h.NEOwn); sumS += h.NEStaged;
         small += h.NC < kTileCells;
      }
      fprintf(stderr, "build_tiles: %d tiles (%d below %d cells), staged edges mean %.1f max %d, own max %d, edge tiles %s\n",
              (int)hdr.size(), small, kTileCells, hdr.empty() ? 0.0 : (double)sumS / hdr.size(), mxS, mxO, edgeOK ? "on" : "off");
   }
   return rc;
}

} // namespace omega

using namespace omega;

extern "C" {

const char *omega_last_error(void) { return g_last_error.c_str(); }
int omega_b200_abi_version(void) { return 1; }
int omega_b200_real_bytes(void) { return (int)sizeof(OmegaReal); }

int omega_dev_upload(void *dst, const void *src, unsigned long long bytes) {
   OMEGA_REQUIRE((dst && src) || bytes == 0, "omega_dev_upload: null argument");
   if (bytes) OMEGA_CUDA_CHECK(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice));
   return 0;
}
int omega_dev_download(void *dst, const void *src, unsigned long long bytes) {
   OMEGA_REQUIRE((dst && src) || bytes == 0, "omega_dev_download: null argument");
   if (bytes) OMEGA_CUDA_CHECK(cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost));
   return 0;
}
int omega_dev_alloc(void **ptr, unsigned long long bytes) {
   OMEGA_CUDA_CHECK(cudaMalloc(ptr, bytes ? bytes : 16));
   OMEGA_CUDA_CHECK(cudaMemset(*ptr, 0, bytes));
   return 0;
}
int omega_dev_free(void *ptr) {
   OMEGA_CUDA_CHECK(cudaFree(ptr));
   return 0;
}
int omega_dev_copy(void *dst, const void *src, unsigned long long bytes, void *stream) {
   if (bytes) OMEGA_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
   return 0;
}
int omega_dev_sync(void *stream) {
   OMEGA_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
   return 0;
}

int omega_mesh_create(const OmegaMeshDesc *D, int Reorder, OmegaMesh **Out) {
   OMEGA_REQUIRE(D && Out, "omega_mesh_create: null argument");
   OMEGA_REQUIRE(D->VertexDegree == 3, "omega_mesh_create: VertexDegree must be 3");
   OMEGA_REQUIRE(D->NCellsSize == D->NCellsAll + 1 && D->NEdgesSize == D->NEdgesAll + 1 &&
                     D->NVerticesSize == D->NVerticesAll + 1,
                 "omega_mesh_create: N*Size must equal N*All+1 (Decomp.cpp:1081-1082)");
   OmegaMesh *M = new OmegaMesh();
   M->Sizes     = *D;
   M->Reordered = Reorder;
   M->MaxEdgesUsed = 0;
   for (int c = 0; c < D->NCellsAll; ++c) M->MaxEdgesUsed = std::max(M->MaxEdgesUsed, D->NEdgesOnCell[c]);
   const int NC = D->NCellsAll, NE = D->NEdgesAll, NV = D->NVerticesAll;
   const int ME = D->MaxEdges, ME2 = D->MaxEdges2;

   // ---- derived fields in host order (HorzMesh.cpp:527-626) ----------------
   M->EdgeSignOnCellH.assign((size_t)D->NCellsSize * ME, 0.0);
   M->EdgeSignOnVertexH.assign((size_t)D->NVerticesSize * 3, 0.0);
   M->EdgeMaskH.assign(D->NEdgesSize, 1.0);
   for (int c = 0; c < NC; ++c)
      for (int i = 0; i < D->NEdgesOnCell[c]; ++i) {
         const int e = D->EdgesOnCell[c * ME + i];
         M->EdgeSignOnCellH[(size_t)c * ME + i] = (c == D->CellsOnEdge[2 * e]) ? -1.0 : 1.0;
      }
   for (int v = 0; v < NV; ++v)
      for (int i = 0; i < 3; ++i) {
         const int e = D->EdgesOnVertex[v * 3 + i];
         M->EdgeSignOnVertexH[(size_t)v * 3 + i] = (v == D->VerticesOnEdge[2 * e]) ? -1.0 : 1.0;
      }
   for (int e = 0; e < NE; ++e) {
      const int c1 = D->CellsOnEdge[2 * e], c2 = D->CellsOnEdge[2 * e + 1];
      if (!(c1 >= 0 && c1 < NC) || !(c2 >= 0 && c2 < NC)) M->EdgeMaskH[e] = 0.0;
   }
   std::vector<double> Scale2(D->NEdgesSize, 0.0), Scale4(D->NEdgesSize, 0.0);
   for (int e = 0; e < NE; ++e) Scale2[e] = Scale4[e] = 1.0;

   // ---- permutations: Perm[new] = old -----------------------------------
   std::vector<int> &PC = M->Perm[0], &PE = M->Perm[1], &PV = M->Perm[2];
   if (Reorder) {
      double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
      const double *X[3] = {D->XCell, D->YCell, D->ZCell};
      for (int a = 0; a < 3; ++a)
         for (int c = 0; c < NC; ++c) {
            lo[a] = std::min(lo[a], X[a][c]);
            hi[a] = std::max(hi[a], X[a][c]);
         }
      std::vector<uint64_t> key(NC);
      for (int c = 0; c < NC; ++c) {
         uint64_t q[3];
         for (int a = 0; a < 3; ++a) {
            const double span = hi[a] - lo[a];
            const double t    = span > 0 ? (X[a][c] - lo[a]) / span : 0.0;
            q[a]              = (uint64_t)std::min(2097151.0, std::floor(t * 2097152.0));
         }
         key[c] = spread3(q[0]) | (spread3(q[1]) << 1) | (spread3(q[2]) << 2);
      }
      const char *ord = getenv("OMEGA_ORDER");
      if (ord && std::strcmp(ord, "morton") == 0) {
         sort_block(PC, 0, D->NCellsOwned, key);
         sort_block(PC, D->NCellsOwned, NC, key);
      } else {
         rcb_block(PC, 0, D->NCellsOwned, kTileCells, X, key);
         rcb_block(PC, D->NCellsOwned, NC, kTileCells, X, key);
      }
   } else {
      PC.resize(NC);
      std::iota(PC.begin(), PC.end(), 0);
   }
   PC.push_back(NC); // boundary slot stays last
   M->IPerm[0].assign(D->NCellsSize, 0);
   for (int i = 0; i < D->NCellsSize; ++i) M->IPerm[0][PC[i]] = i;
   auto dual_perm = [&](int NAll, int NOwned, const int *CellsOnX, int width, std::vector<int> &P,
                        std::vector<int> &IP) {
      if (Reorder) {
         std::vector<int> key(NAll);
         for (int x = 0; x < NAll; ++x) {
            int best = INT32_MAX;
            for (int j = 0; j < width; ++j) {
               const int c = CellsOnX[x * width + j];
               if (c >= 0 && c < NC) best = std::min(best, M->IPerm[0][c]);
            }
            key[x] = best;
         }
         sort_block(P, 0, NOwned, key);
         sort_block(P, NOwned, NAll, key);
      } else {
         P.resize(NAll);
         std::iota(P.begin(), P.end(), 0);
      }
      P.push_back(NAll);
      IP.assign(NAll + 1, 0);
      for (int i = 0; i <= NAll; ++i) IP[P[i]] = i;
   };
   dual_perm(NE, D->NEdgesOwned, D->CellsOnEdge, 2, PE, M->IPerm[1]);
   dual_perm(NV, D->NVerticesOwned, D->CellsOnVertex, 3, PV, M->IPerm[2]);
   const std::vector<int> &IC = M->IPerm[0], &IE = M->IPerm[1], &IV = M->IPerm[2];

   // ---- device SoA in device order ---------------------------------------
   DMesh &G = M->D;
   G.NCellsAll = NC; G.NEdgesAll = NE; G.NVerticesAll = NV;
   G.NCellsOwned = D->NCellsOwned; G.NEdgesOwned = D->NEdgesOwned; G.NVerticesOwned = D->NVerticesOwned;
   G.NCellsSize = D->NCellsSize; G.NEdgesSize = D->NEdgesSize; G.NVerticesSize = D->NVerticesSize;
   G.K = D->NVertLayers; G.MaxEdges = ME; G.MaxEdges2 = ME2;
   int rc = 0;
#define STAGE(dst, vec) rc |= to_device(M, (vec), &(dst))
   STAGE(G.NEdgesOnCell, gather_rows(D->NEdgesOnCell, PC, 1));
   STAGE(G.EdgesOnCell, gather_index_rows(D->EdgesOnCell, PC, ME, IE));
   STAGE(G.EdgeSignOnCell, gather_rows(M->EdgeSignOnCellH.data(), PC, ME));
   STAGE(G.AreaCell, gather_rows(D->AreaCell, PC, 1));
   STAGE(G.BottomDepth, gather_rows(D->BottomDepth, PC, 1));
   STAGE(G.XCell, gather_rows(D->XCell, PC, 1));
   STAGE(G.YCell, gather_rows(D->YCell, PC, 1));
   std::vector<int> eoe_dev, neoe_dev; // device-order EdgesOnEdge / trimmed counts (tiles need them too)
   std::vector<int> coe = gather_index_rows(D->CellsOnEdge, PE, 2, IC);
   std::vector<int> voe = gather_index_rows(D->VerticesOnEdge, PE, 2, IV);
   {
      const int *p = nullptr;
      rc |= to_device(M, coe, &p); G.CellsOnEdge = reinterpret_cast<const int2 *>(p);
      rc |= to_device(M, voe, &p); G.VerticesOnEdge = reinterpret_cast<const int2 *>(p);
   }
   {
      // EdgesOnEdge / WeightsOnEdge; trailing boundary-slot entries (zero weight
      // padding that Decomp.cpp:2186-2200 counts) are trimmed from the count
      std::vector<int> &eoe = eoe_dev, &neoe = neoe_dev;
      eoe  = gather_index_rows(D->EdgesOnEdge, PE, ME2, IE);
      neoe = gather_rows(D->NEdgesOnEdge, PE, 1);
      std::vector<double> w = gather_rows(D->WeightsOnEdge, PE, ME2);
      for (int e = 0; e < NE; ++e) {
         int n = neoe[e];
         while (n > 0 && eoe[(size_t)e * ME2 + n - 1] == NE && w[(size_t)e * ME2 + n - 1] == 0.0) --n;
         neoe[e] = n;
      }
      STAGE(G.NEdgesOnEdge, neoe);
      STAGE(G.EdgesOnEdge, eoe);
      // zero weight behind the (trimmed) count: the tiled edge pass always sums 10 entries
      for (int e = 0; e < NE; ++e)
         for (int j = neoe[e]; j < ME2; ++j) w[(size_t)e * ME2 + j] = 0.0;
      STAGE(G.WeightsOnEdge, w);
      // flattened gather tables of the fused edge kernel: cells and vertices of every
      // EdgesOnEdge entry, so a neighbour costs one index level instead of three
      std::vector<int> ec((size_t)(NE + 1) * ME2 * 2, NC), ev((size_t)(NE + 1) * ME2 * 2, NV);
      for (int e = 0; e < NE; ++e)
         for (int j = 0; j < neoe[e]; ++j) {
            const int je = eoe[(size_t)e * ME2 + j];
            const size_t o = ((size_t)e * ME2 + j) * 2;
            ec[o] = coe[(size_t)je * 2]; ec[o + 1] = coe[(size_t)je * 2 + 1];
            ev[o] = voe[(size_t)je * 2]; ev[o + 1] = voe[(size_t)je * 2 + 1];
         }
      const int *q = nullptr;
      rc |= to_device(M, ec, &q); G.EoECells = reinterpret_cast<const int2 *>(q);
      rc |= to_device(M, ev, &q); G.EoEVerts = reinterpret_cast<const int2 *>(q);
   }
   STAGE(G.DcEdge, gather_rows(D->DcEdge, PE, 1));
   STAGE(G.DvEdge, gather_rows(D->DvEdge, PE, 1));
   STAGE(G.AngleEdge, gather_rows(D->AngleEdge, PE, 1));
   STAGE(G.FEdge, gather_rows(D->FEdge, PE, 1));
   STAGE(G.XEdge, gather_rows(D->XEdge, PE, 1));
   STAGE(G.YEdge, gather_rows(D->YEdge, PE, 1));
   STAGE(G.EdgeMask, gather_rows(M->EdgeMaskH.data(), PE, 1));
   STAGE(G.MeshScalingDel2, gather_rows(Scale2.data(), PE, 1));
   STAGE(G.MeshScalingDel4, gather_rows(Scale4.data(), PE, 1));
   STAGE(G.CellsOnVertex, gather_index_rows(D->CellsOnVertex, PV, 3, IC));
   STAGE(G.EdgesOnVertex, gather_index_rows(D->EdgesOnVertex, PV, 3, IE));
   STAGE(G.EdgeSignOnVertex, gather_rows(M->EdgeSignOnVertexH.data(), PV, 3));
   STAGE(G.KiteAreasOnVertex, gather_rows(D->KiteAreasOnVertex, PV, 3));
   STAGE(G.AreaTriangle, gather_rows(D->AreaTriangle, PV, 1));
   STAGE(G.FVertex, gather_rows(D->FVertex, PV, 1));
   {
      std::vector<double> dc = gather_rows(D->DcEdge, PE, 1), dv = gather_rows(D->DvEdge, PE, 1);
      std::vector<double> ar = gather_rows(D->AreaCell, PC, 1), at = gather_rows(D->AreaTriangle, PV, 1);
      std::vector<double> idc(dc.size(), 0.0), idv(dc.size(), 0.0), idve(dc.size(), 0.0), iar(ar.size(), 0.0),
          iat(at.size(), 0.0);
      for (int e = 0; e < NE; ++e) {
         idc[e]  = 1.0 / dc[e];
         idv[e]  = 1.0 / dv[e];
         idve[e] = 1.0 / std::max(dv[e], 0.25 * dc[e]);
      }
      for (int c = 0; c < NC; ++c) iar[c] = 1.0 / ar[c];
      for (int v = 0; v < NV; ++v) iat[v] = 1.0 / at[v];
      STAGE(G.InvDcEdge, idc);
      STAGE(G.InvDvEdge, idv);
      STAGE(G.InvDvEffEdge, idve);
      STAGE(G.InvAreaCell, iar);
      STAGE(G.InvAreaTriangle, iat);
   }
   {
      // k_fB tables (omega_dev.h).  Del2Edge(e) = mask (Div(c1)-Div(c0))/dc - (RelVort(v1)-RelVort(v0))/max(dv, dc/4)
      // (VelocityDel2AuxVars.h:21-45); Div(c1)-Div(c0) = -sign_c (Div(nbr)-Div(c)), RelVort(v1)-RelVort(v0) =
      // -sign_v (RelVort(vo)-RelVort(v)) with the edge signs of HorzMesh.cpp:527-566.
      std::vector<int> eoc = gather_index_rows(D->EdgesOnCell, PC, ME, IE), nec = gather_rows(D->NEdgesOnCell, PC, 1);
      std::vector<int> eov = gather_index_rows(D->EdgesOnVertex, PV, 3, IE);
      std::vector<double> sc = gather_rows(M->EdgeSignOnCellH.data(), PC, ME), sv = gather_rows(M->EdgeSignOnVertexH.data(), PV, 3);
      std::vector<double> dc = gather_rows(D->DcEdge, PE, 1), dv = gather_rows(D->DvEdge, PE, 1);
      std::vector<double> mk = gather_rows(M->EdgeMaskH.data(), PE, 1);
      std::vector<double> ar = gather_rows(D->AreaCell, PC, 1), at = gather_rows(D->AreaTriangle, PV, 1);
      std::vector<int> ci((size_t)(NC + 1) * 6 * 4, 0), vi((size_t)(NV + 1) * 3 * 4, 0);
      std::vector<double> cc((size_t)(NC + 1) * 6 * 2, 0.0), vc((size_t)(NV + 1) * 3 * 2, 0.0);
      for (int c = 0; c <= NC; ++c)
         for (int j = 0; j < 6; ++j) {
            const size_t o = ((size_t)c * 6 + j);
            ci[o * 4] = c; ci[o * 4 + 1] = NV; ci[o * 4 + 2] = NV;
            if (c < NC && j < nec[c] && j < ME) {
               const int e = eoc[(size_t)c * ME + j];
               const int a = coe[(size_t)e * 2], b = coe[(size_t)e * 2 + 1];
               ci[o * 4] = a == c ? b : a; ci[o * 4 + 1] = voe[(size_t)e * 2]; ci[o * 4 + 2] = voe[(size_t)e * 2 + 1];
               const double cf = dv[e] * (1.0 / ar[c]) * sc[(size_t)c * ME + j];
               cc[o * 2]     = cf * sc[(size_t)c * ME + j] * mk[e] * (1.0 / dc[e]);
               cc[o * 2 + 1] = cf * (1.0 / std::max(dv[e], 0.25 * dc[e]));
            }
         }
      for (int v = 0; v <= NV; ++v)
         for (int j = 0; j < 3; ++j) {
            const size_t o = ((size_t)v * 3 + j);
            vi[o * 4] = NC; vi[o * 4 + 1] = NC; vi[o * 4 + 2] = v;
            if (v < NV) {
               const int e = eov[(size_t)v * 3 + j];
               if (e < NE) {
                  vi[o * 4] = coe[(size_t)e * 2]; vi[o * 4 + 1] = coe[(size_t)e * 2 + 1];
                  const int a = voe[(size_t)e * 2], b = voe[(size_t)e * 2 + 1];
                  vi[o * 4 + 2] = a == v ? b : a;
                  const double cf = (1.0 / at[v]) * dc[e] * sv[(size_t)v * 3 + j];
                  vc[o * 2]     = cf * mk[e] * (1.0 / dc[e]);
                  vc[o * 2 + 1] = cf * sv[(size_t)v * 3 + j] * (1.0 / std::max(dv[e], 0.25 * dc[e]));
               }
            }
         }
      // k_fA tables (KineticAuxVars.h:20-47, VorticityAuxVars.h:24-59), coefficients evaluated in the
      // kernels' former order so the bits are unchanged
      {
         std::vector<int> cov = gather_index_rows(D->CellsOnVertex, PV, 3, IC);
         std::vector<double> kite = gather_rows(D->KiteAreasOnVertex, PV, 3);
         std::vector<int> ae((size_t)(NC + 1) * 6, NE), avi((size_t)(NV + 1) * 3 * 2, 0);
         std::vector<double> ac((size_t)(NC + 1) * 6 * 2, 0.0), avc((size_t)(NV + 1) * 3 * 2, 0.0);
         for (int c = 0; c < NC; ++c) {
            const double InvA = 1.0 / ar[c];
            for (int j = 0; j < nec[c] && j < 6 && j < ME; ++j) {
               const int e = eoc[(size_t)c * ME + j];
               ae[(size_t)c * 6 + j] = e;
               ac[((size_t)c * 6 + j) * 2]     = 0.5 * dv[e] * dc[e] * 0.5 * InvA;
               ac[((size_t)c * 6 + j) * 2 + 1] = dv[e] * InvA * sc[(size_t)c * ME + j];
            }
         }
         for (int v = 0; v <= NV; ++v)
            for (int j = 0; j < 3; ++j) {
               const size_t o = (size_t)v * 3 + j;
               avi[o * 2] = NC; avi[o * 2 + 1] = NE;
               if (v < NV) {
                  const double InvAt = 1.0 / at[v];
                  const int e = eov[o];
                  avi[o * 2] = cov[o]; avi[o * 2 + 1] = e;
                  avc[o * 2]     = InvAt * kite[o];
                  avc[o * 2 + 1] = InvAt * (e < NE ? dc[e] : 0.0) * sv[o];
               }
            }
         const int *qi2 = nullptr;
         const Real *qd2 = nullptr;
         rc |= to_device(M, ae, &G.ACellEdge);
         rc |= to_device(M, avi, &qi2); G.AVertIdx = reinterpret_cast<const int2 *>(qi2);
         rc |= to_device(M, ac, &qd2); G.ACellCf = reinterpret_cast<const Real2 *>(qd2);
         rc |= to_device(M, avc, &qd2); G.AVertCf = reinterpret_cast<const Real2 *>(qd2);
      }
      {
         // k_fB3 tables: the same sums with every gathered row read once.  Around a cell each vertex is the
         // end of two consecutive edges, around a vertex each cell borders two of its three edges, so
         //   sum_j B_j (z(v1_j) - z(v0_j)) = sum_v z(v) * (sum of +-B_j),   sum_j A'_j (D(c1_j) - D(c0_j)) = sum_c D(c) * (+-A'_j)
         // cell row: 6 neighbour cells + 6 vertices (indices [12], coefficients [12]);
         // vertex row: 4 cells + 3 opposite vertices (+ pad) (indices [8], coefficients [8])
         std::vector<int> ci3((size_t)(NC + 1) * 12, 0), vi3((size_t)(NV + 1) * 8, 0);
         std::vector<double> cc3((size_t)(NC + 1) * 12, 0.0), vc3((size_t)(NV + 1) * 8, 0.0);
         bool ok = true;
         auto merge = [](int *idx, double *cf, int cap, int skip, int i, double c, bool &good) {
            if (c == 0.0 || i == skip) return; // the boundary-slot row is zero
            for (int q = 0; q < cap; ++q) {
               if (cf[q] != 0.0 && idx[q] == i) { cf[q] += c; return; }
            }
            for (int q = 0; q < cap; ++q) {
               if (cf[q] == 0.0) { idx[q] = i; cf[q] = c; return; }
            }
            good = false;
         };
         for (int c = 0; c <= NC; ++c) {
            int *ix = &ci3[(size_t)c * 12];
            double *cf = &cc3[(size_t)c * 12];
            for (int j = 0; j < 6; ++j) { ix[j] = c; ix[6 + j] = NV; }
            for (int j = 0; j < 6; ++j) {
               const size_t o = (size_t)c * 6 + j;
               ix[j] = ci[o * 4]; cf[j] = cc[o * 2];
               merge(ix + 6, cf + 6, 6, NV, ci[o * 4 + 2], cc[o * 2 + 1], ok);
               merge(ix + 6, cf + 6, 6, NV, ci[o * 4 + 1], -cc[o * 2 + 1], ok);
            }
         }
         for (int v = 0; v <= NV; ++v) {
            int *ix = &vi3[(size_t)v * 8];
            double *cf = &vc3[(size_t)v * 8];
            for (int j = 0; j < 4; ++j) ix[j] = NC;
            for (int j = 0; j < 4; ++j) ix[4 + j] = v;
            for (int j = 0; j < 3; ++j) {
               const size_t o = (size_t)v * 3 + j;
               merge(ix, cf, 4, NC, vi[o * 4 + 1], vc[o * 2], ok);
               merge(ix, cf, 4, NC, vi[o * 4], -vc[o * 2], ok);
               ix[4 + j] = vi[o * 4 + 2]; cf[4 + j] = vc[o * 2 + 1];
            }
         }
         G.B3OK = ok ? 1 : 0;
         const int *qi3 = nullptr;
         const Real *qd3 = nullptr;
         rc |= to_device(M, ci3, &qi3); G.B3CellIdx = reinterpret_cast<const int4 *>(qi3);
         rc |= to_device(M, vi3, &qi3); G.B3VertIdx = reinterpret_cast<const int4 *>(qi3);
         rc |= to_device(M, cc3, &qd3); G.B3CellCf = reinterpret_cast<const Real2 *>(qd3);
         rc |= to_device(M, vc3, &qd3); G.B3VertCf = reinterpret_cast<const Real2 *>(qd3);
      }
      const int *qi = nullptr;
      const Real *qd = nullptr;
      rc |= to_device(M, ci, &qi); G.B2CellIdx = reinterpret_cast<const int4 *>(qi);
      rc |= to_device(M, vi, &qi); G.B2VertIdx = reinterpret_cast<const int4 *>(qi);
      rc |= to_device(M, cc, &qd); G.B2CellCf = reinterpret_cast<const Real2 *>(qd);
      rc |= to_device(M, vc, &qd); G.B2VertCf = reinterpret_cast<const Real2 *>(qd);
   }
#undef STAGE
   if (!rc)
      rc |= build_tiles(M, NC, D->NCellsOwned, ME, gather_rows(D->NEdgesOnCell, PC, 1),
                        gather_index_rows(D->EdgesOnCell, PC, ME, IE), coe,
                        gather_rows(M->EdgeMaskH.data(), PE, 1), gather_rows(D->DvEdge, PE, 1),
                        gather_rows(D->DcEdge, PE, 1), gather_rows(D->AreaCell, PC, 1),
                        gather_rows(M->EdgeSignOnCellH.data(), PC, ME), NE, ME2, neoe_dev, eoe_dev, voe);
   for (int kind = 0; kind < 3; ++kind) {
      const int *p = nullptr;
      rc |= to_device(M, M->Perm[kind], &p);
      M->DPerm[kind] = const_cast<int *>(p);
      rc |= to_device(M, M->IPerm[kind], &p);
      M->DIPerm[kind] = const_cast<int *>(p);
   }
   if (rc) {
      omega_mesh_destroy(M);
      return 1;
   }
   // scalar-only copy of the descriptor
   OmegaMeshDesc S = *D;
   std::memset(&S.NEdgesOnCell, 0, sizeof(S) - offsetof(OmegaMeshDesc, NEdgesOnCell));
   M->Sizes = S;
   *Out     = M;
   return 0;
}

int omega_mesh_destroy(OmegaMesh *M) {
   if (!M) return 0;
   for (void *p : M->Allocs) cudaFree(p);
   delete M;
   return 0;
}

int omega_mesh_perm(const OmegaMesh *M, OmegaMeshElement Elem, int *Perm, int *IPerm) {
   OMEGA_REQUIRE(M && Elem >= 0 && Elem < 3, "omega_mesh_perm: bad argument");
   if (Perm) std::memcpy(Perm, M->Perm[Elem].data(), sizeof(int) * M->Perm[Elem].size());
   if (IPerm) std::memcpy(IPerm, M->IPerm[Elem].data(), sizeof(int) * M->IPerm[Elem].size());
   return 0;
}

int omega_mesh_derived(const OmegaMesh *M, double *EdgeSignOnCell, double *EdgeSignOnVertex, double *EdgeMask) {
   OMEGA_REQUIRE(M, "omega_mesh_derived: null mesh");
   if (EdgeSignOnCell) std::memcpy(EdgeSignOnCell, M->EdgeSignOnCellH.data(), sizeof(double) * M->EdgeSignOnCellH.size());
   if (EdgeSignOnVertex) std::memcpy(EdgeSignOnVertex, M->EdgeSignOnVertexH.data(), sizeof(double) * M->EdgeSignOnVertexH.size());
   if (EdgeMask) std::memcpy(EdgeMask, M->EdgeMaskH.data(), sizeof(double) * M->EdgeMaskH.size());
   return 0;
}

// Field transfers with the permutation applied (see the three modes inside).  The slab path moves the host array
// in slabs of consecutive HOST rows through ONE fixed-size staging block (64 MB) instead of a second full-size
// copy of the field (3.4 GB for the ten tracers of config 3, 10.7 GB at 32 tracers): a slab is copied, then
// scattered to / gathered from its device rows (DIPerm: device row of a host row).
static constexpr size_t kSlabBytes = 64u << 20;

// The staging blocks come from a pool of the library's own, one per device, that keeps freed blocks (release
// threshold: never): with the default threshold a multi-GB block would go back to the driver at every
// synchronisation and be allocated again for every field of every step.  The process's default pool, which the
// host application may use itself, is left alone.
static int stage_pool(cudaMemPool_t *Pool) {
   constexpr int kMaxDevices = 64;
   static cudaMemPool_t pools[kMaxDevices] = {};
   int dev = 0;
   OMEGA_CUDA_CHECK(cudaGetDevice(&dev));
   OMEGA_REQUIRE(dev >= 0 && dev < kMaxDevices, "field transfer: device ordinal beyond 63");
   if (!pools[dev]) {
      cudaMemPoolProps props = {};
      props.allocType     = cudaMemAllocationTypePinned;
      props.handleTypes   = cudaMemHandleTypeNone;
      props.location.type = cudaMemLocationTypeDevice;
      props.location.id   = dev;
      cudaMemPool_t p;
      OMEGA_CUDA_CHECK(cudaMemPoolCreate(&p, &props));
      unsigned long long keep = ~0ull;
      OMEGA_CUDA_CHECK(cudaMemPoolSetAttribute(p, cudaMemPoolAttrReleaseThreshold, &keep));
      pools[dev] = p;
   }
   *Pool = pools[dev];
   return 0;
}

static int field_transfer(const OmegaMesh *M, OmegaMeshElement Elem, int NOuter, int NInner, Real *Dev, Real *Host,
                          cudaStream_t S, bool ToDevice) {
   const int NRows    = n_size(M, Elem);
   const size_t bytes = (size_t)NOuter * NRows * NInner * sizeof(Real);
   if (bytes == 0) return 0;
   if (!M->Reordered) {
      if (ToDevice) OMEGA_CUDA_CHECK(cudaMemcpyAsync(Dev, Host, bytes, cudaMemcpyHostToDevice, S));
      else OMEGA_CUDA_CHECK(cudaMemcpyAsync(Host, Dev, bytes, cudaMemcpyDeviceToHost, S));
      return 0;
   }
   // Three ways, measured on config 3 with the bench's three-stream e2e pipeline (uploads, steps and downloads
   // of consecutive steps overlap; 4.7 GB each way per step):
   //   full  (default)  one copy of the whole field through a full-size stream-ordered staging block + one
   //                    permute kernel: 112 ms/step.  Costs a transient second copy of the field.
   //   slab             64 MB slabs of host rows through one small block: 138 ms/step -- every slab's permute
   //                    kernel waits for SM room beside the stage kernels of the compute stream.  Taken
   //                    automatically for a field larger than an eighth of the device's memory.
   //   zerocopy         the kernel reads / writes PINNED host memory itself over PCIe: 180 ms/step.
   const char *mode = getenv("OMEGA_XFER");
   // (the device's total memory is looked up once; a per-call cudaMemGetInfo would serialise the streams)
   static size_t totalB = 0;
   if (!totalB) {
      size_t freeB = 0;
      if (cudaMemGetInfo(&freeB, &totalB) != cudaSuccess) totalB = (size_t)16 << 30;
   }
   const bool wantSlab = mode && !strcmp(mode, "slab");
   if (mode && !strcmp(mode, "zerocopy")) {
      cudaPointerAttributes at;
      const bool pinned = cudaPointerGetAttributes(&at, Host) == cudaSuccess && at.type == cudaMemoryTypeHost;
      cudaGetLastError(); // a pageable pointer is not an error
      if (pinned) {
         int dev = 0, sm = 0;
         OMEGA_CUDA_CHECK(cudaGetDevice(&dev));
         OMEGA_CUDA_CHECK(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev));
         launch_rows_hostmem(S, ToDevice, NOuter, NRows, NInner, M->DIPerm[Elem], Dev, (Real *)at.devicePointer, 2 * sm);
         OMEGA_CUDA_CHECK(cudaGetLastError());
         return 0;
      }
   }
   if (!wantSlab && bytes <= totalB / 8) {
      Real *stage = nullptr;
      cudaMemPool_t pool;
      if (int rc = stage_pool(&pool)) return rc;
      OMEGA_CUDA_CHECK(cudaMallocFromPoolAsync((void **)&stage, bytes, pool, S));
      if (ToDevice) {
         OMEGA_CUDA_CHECK(cudaMemcpyAsync(stage, Host, bytes, cudaMemcpyHostToDevice, S));
         launch_permute_rows(S, NOuter, NRows, NInner, M->DPerm[Elem], stage, Dev);   // device row r <- host row Perm[r]
      } else {
         launch_permute_rows(S, NOuter, NRows, NInner, M->DIPerm[Elem], Dev, stage);  // host row r <- device row IPerm[r]
         OMEGA_CUDA_CHECK(cudaMemcpyAsync(Host, stage, bytes, cudaMemcpyDeviceToHost, S));
      }
      OMEGA_CUDA_CHECK(cudaGetLastError());
      OMEGA_CUDA_CHECK(cudaFreeAsync(stage, S));
      return 0;
   }
   const size_t rowBytes = (size_t)NInner * sizeof(Real);
   const int slabRows    = (int)std::max<size_t>(1, std::min<size_t>(NRows, kSlabBytes / rowBytes));
   Real *stage = nullptr;
   cudaMemPool_t pool;
   if (int rc = stage_pool(&pool)) return rc;
   OMEGA_CUDA_CHECK(cudaMallocFromPoolAsync((void **)&stage, (size_t)slabRows * rowBytes, pool, S));
   for (int o = 0; o < NOuter; ++o) {
      Real *dev  = Dev + (size_t)o * NRows * NInner;
      Real *host = Host + (size_t)o * NRows * NInner;
      for (int r0 = 0; r0 < NRows; r0 += slabRows) {
         const int n = std::min(slabRows, NRows - r0);
         if (ToDevice) {
            OMEGA_CUDA_CHECK(cudaMemcpyAsync(stage, host + (size_t)r0 * NInner, (size_t)n * rowBytes, cudaMemcpyHostToDevice, S));
            launch_slab_rows(S, true, r0, n, NInner, M->DIPerm[Elem], dev, stage);
         } else {
            launch_slab_rows(S, false, r0, n, NInner, M->DIPerm[Elem], dev, stage);
            OMEGA_CUDA_CHECK(cudaMemcpyAsync(host + (size_t)r0 * NInner, stage, (size_t)n * rowBytes, cudaMemcpyDeviceToHost, S));
         }
      }
   }
   OMEGA_CUDA_CHECK(cudaGetLastError());
   OMEGA_CUDA_CHECK(cudaFreeAsync(stage, S));
   return 0;
}

int omega_field_to_device(const OmegaMesh *M, OmegaMeshElement Elem, int NOuter, int NInner,
                          const Real *Host, Real *Dev, void *Stream) {
   OMEGA_REQUIRE(M && Host && Dev && Elem >= 0 && Elem < 3, "omega_field_to_device: bad argument");
   return field_transfer(M, Elem, NOuter, NInner, Dev, const_cast<Real *>(Host), (cudaStream_t)Stream, true);
}

static int field_to_host(const OmegaMesh *M, OmegaMeshElement Elem, int NOuter, int NInner, const Real *Dev,
                         Real *Host, void *Stream, bool Sync) {
   OMEGA_REQUIRE(M && Host && Dev && Elem >= 0 && Elem < 3, "omega_field_to_host: bad argument");
   int rc = field_transfer(M, Elem, NOuter, NInner, const_cast<Real *>(Dev), Host, (cudaStream_t)Stream, false);
   if (rc) return rc;
   if (Sync) OMEGA_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)Stream));
   return 0;
}

int omega_field_to_host(const OmegaMesh *M, OmegaMeshElement Elem, int NOuter, int NInner,
                        const Real *Dev, Real *Host, void *Stream) {
   return field_to_host(M, Elem, NOuter, NInner, Dev, Host, Stream, true);
}

int omega_field_to_host_async(const OmegaMesh *M, OmegaMeshElement Elem, int NOuter, int NInner,
                              const Real *Dev, Real *Host, void *Stream) {
   return field_to_host(M, Elem, NOuter, NInner, Dev, Host, Stream, false);
}

} // extern "C"

// ---- omega_field_to_device_noalloc: the upload with its staging block in a file-scope device array -------------
// For callers that may not allocate (stream capture, serving loops, harnesses that forbid cudaMalloc).  The array is
// part of the library image: every GPU that loads the library holds kNoAllocStageBytes until the process exits.
// In host code the NAME below is a host-side stand-in, not a device address; the address a kernel or a copy may
// use comes from cudaGetSymbolAddress, looked up once per device with that device current.
namespace omega {
constexpr size_t kNoAllocStageBytes = (size_t)OMEGA_NOALLOC_MAX_BYTES;
static __device__ __align__(256) Real g_noalloc_stage[kNoAllocStageBytes / sizeof(Real)];

static int noalloc_stage_ptr(Real **Ptr) {
   constexpr int kMaxDevices = 64;
   static Real *perDevice[kMaxDevices] = {};
   int dev = 0;
   OMEGA_CUDA_CHECK(cudaGetDevice(&dev));
   OMEGA_REQUIRE(dev >= 0 && dev < kMaxDevices, "omega_field_to_device_noalloc: device ordinal beyond 63");
   if (!perDevice[dev]) {
      Real *p = nullptr;
      OMEGA_CUDA_CHECK(cudaGetSymbolAddress((void **)&p, g_noalloc_stage));
      OMEGA_REQUIRE(p, "omega_field_to_device_noalloc: cudaGetSymbolAddress returned a null address");
      perDevice[dev] = p;
   }
   *Ptr = perDevice[dev];
   return 0;
}
} // namespace omega

extern "C" int omega_field_to_device_noalloc(const OmegaMesh *M, OmegaMeshElement Elem, int NOuter, int NInner,
                                             const Real *Host, Real *Dev, void *Stream) {
   OMEGA_REQUIRE(M && Host && Dev && Elem >= 0 && Elem < 3 && NOuter >= 0 && NInner >= 0,
                 "omega_field_to_device_noalloc: bad argument");
   const int NRows    = n_size(M, Elem);
   const size_t bytes = (size_t)NOuter * NRows * NInner * sizeof(Real);
   if (bytes == 0) return 0;
   OMEGA_REQUIRE(bytes <= kNoAllocStageBytes,
                 "omega_field_to_device_noalloc: field of " + std::to_string(bytes) + " bytes exceeds the static staging block (" +
                     std::to_string(kNoAllocStageBytes) + " bytes, OMEGA_NOALLOC_MAX_BYTES); use omega_field_to_device");
   cudaStream_t S = (cudaStream_t)Stream;
   if (!M->Reordered) { // no permutation, no staging
      OMEGA_CUDA_CHECK(cudaMemcpyAsync(Dev, Host, bytes, cudaMemcpyHostToDevice, S));
      return 0;
   }
   Real *stage = nullptr;
   if (int rc = noalloc_stage_ptr(&stage)) return rc;
   // every element of stage[0, bytes) is overwritten by the copy before the kernel reads it: nothing to clear
   OMEGA_CUDA_CHECK(cudaMemcpyAsync(stage, Host, bytes, cudaMemcpyHostToDevice, S));
   launch_permute_rows(S, NOuter, NRows, NInner, M->DPerm[Elem], stage, Dev); // device row r <- host row Perm[r]
   OMEGA_CUDA_CHECK(cudaGetLastError());
   return 0;
}
