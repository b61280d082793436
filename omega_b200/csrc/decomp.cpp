// decomp.cpp -- Decomp behind the C ABI: partition, halo layers, Omega's local numbering,
// owner addresses and the halo exchange lists, for ONE rank (host code, no device work).
//
// Restates the rules of components/omega/src/base/Decomp.cpp that define the local index spaces
//  and of src/base/Halo.cpp:455-600 for the exchange lists:
//   * cells: owned in ascending global ID, then HaloWidth layers, each sorted ascending by global ID
//     (Decomp.cpp:979-1100);
//   * edges / vertices: owner = task of the first valid cell in cellsOnEdge / cellsOnVertex
//     (:1493-1502); owned ones numbered in first-encounter order walking EdgesOnCell of the owned
//     cells, the non-owned ones around owned cells in REVERSE encounter order, then one block per
//     cell-halo layer in encounter order (:1544-1609);
//   * every XxOnYy entry becomes a local index, anything not local the boundary slot N*All (:552-712).
// The reference needs MPI for two things a rank can also derive alone from the global connectivity and
// the partition: the index of a halo element on its owner (it broadcasts every task's owned list,
// :1639-1676) and the send lists (the receivers mail theirs, Halo.cpp:577-597).  Here a rank numbers
// its neighbours' partitions as well (a handful, each as cheap as its own) and reads both off them, so
// nobody builds every rank's mesh.
//
// The partitioner: recursive coordinate bisection, optionally followed by Fiduccia-Mattheyses style
// boundary refinement (move a boundary cell to the neighbouring part when that lowers the edge cut and
// keeps the balance), reported with its edge cut.  METIS_PartGraphKway (Decomp.cpp:968-971) is
// third party; results on owned elements do not depend on the partition.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <numeric>
#include <set>
#include <vector>

#include "omega_dev.h"

namespace {

struct LocalOrder {
   int Task = 0, HW = 0;
   std::vector<int> Ids[3];       // global ids in local order
   int NOwned[3] = {0, 0, 0};
   std::vector<int> NHalo[3];     // cumulative counts through layer k (cells: HW entries; edges/vertices: HW)
};

inline int first_valid_owner(const int *cells, int n, const int *task) {
   for (int j = 0; j < n; ++j)
      if (cells[j] >= 0) return task[cells[j]];
   return -1;
}

// encounter order of the entries of X-on-cell rows [r0, r1) of `cells`, skipping seen ones
void encounter(const int *XOnCell, int W, const std::vector<int> &cells, int r0, int r1, std::vector<char> &seen,
               std::vector<int> &out) {
   for (int r = r0; r < r1; ++r) {
      const int *row = XOnCell + (size_t)cells[r] * W;
      for (int j = 0; j < W; ++j) {
         const int x = row[j];
         if (x >= 0 && !seen[x]) {
            seen[x] = 1;
            out.push_back(x);
         }
      }
   }
}

void order_dual(const OmegaGlobalMesh &G, const int *XOnCell, const int *CellsOnX, int NX, int WX, const int *task,
                const LocalOrder &L, int kind, LocalOrder &Out) {
   const std::vector<int> &cells = L.Ids[0];
   std::vector<char> seen(NX, 0);
   std::vector<int> around;
   encounter(XOnCell, G.MaxEdges, cells, 0, L.NOwned[0], seen, around);
   std::vector<int> owned, halo1;
   for (int x : around) (first_valid_owner(CellsOnX + (size_t)x * WX, WX, task) == L.Task ? owned : halo1).push_back(x);
   std::reverse(halo1.begin(), halo1.end()); // stored from the end of the block inward (Decomp.cpp:1552-1582)
   std::vector<int> &ids = Out.Ids[kind];
   ids = owned;
   ids.insert(ids.end(), halo1.begin(), halo1.end());
   Out.NOwned[kind] = (int)owned.size();
   Out.NHalo[kind].assign(1, (int)around.size());
   int start = L.NOwned[0];
   for (int h = 0; h < L.HW; ++h) {
      const int end = L.NHalo[0][h];
      encounter(XOnCell, G.MaxEdges, cells, start, end, seen, ids);
      start = end;
      if (h + 1 < L.HW) Out.NHalo[kind].push_back((int)ids.size());
   }
}

LocalOrder local_order(const OmegaGlobalMesh &G, const int *task, int t, int HW, const std::vector<std::vector<int>> &owned_of) {
   LocalOrder L;
   L.Task = t; L.HW = HW;
   std::vector<int> &cells = L.Ids[0];
   cells = owned_of[t];
   L.NOwned[0] = (int)cells.size();
   std::vector<char> in(G.NCells, 0);
   for (int c : cells) in[c] = 1;
   size_t prev0 = 0, prev1 = cells.size();
   for (int h = 0; h < HW; ++h) {
      std::vector<int> nb;
      for (size_t r = prev0; r < prev1; ++r) {
         const int *row = G.CellsOnCell + (size_t)cells[r] * G.MaxEdges;
         for (int j = 0; j < G.MaxEdges; ++j) {
            const int n = row[j];
            if (n >= 0 && !in[n]) { in[n] = 1; nb.push_back(n); }
         }
      }
      std::sort(nb.begin(), nb.end());
      prev0 = cells.size();
      cells.insert(cells.end(), nb.begin(), nb.end());
      prev1 = cells.size();
      L.NHalo[0].push_back((int)cells.size());
   }
   order_dual(G, G.EdgesOnCell, G.CellsOnEdge, G.NEdges, 2, task, L, 1, L);
   order_dual(G, G.VerticesOnCell, G.CellsOnVertex, G.NVertices, G.VertexDegree, task, L, 2, L);
   return L;
}

} // namespace

struct OmegaDecomp {
   OmegaGlobalMesh G; // pointers stay the caller's
   int NTasks, MyTask, HW;
   LocalOrder L;
   std::vector<int> Loc[3];                  // [NAll+1][2]
   std::vector<int> Neighbors;
   std::vector<int> SendOff[3], SendIdx[3], RecvOff[3], RecvIdx[3];
};

using namespace omega;

extern "C" {

int omega_decomp_edge_cut(const OmegaGlobalMesh *G, const int *CellTask, long long *EdgeCut) {
   OMEGA_REQUIRE(G && CellTask && EdgeCut, "omega_decomp_edge_cut: null argument");
   long long cut = 0;
   for (int e = 0; e < G->NEdges; ++e) {
      const int a = G->CellsOnEdge[2 * e], b = G->CellsOnEdge[2 * e + 1];
      if (a >= 0 && b >= 0 && CellTask[a] != CellTask[b]) ++cut;
   }
   *EdgeCut = cut;
   return 0;
}

int omega_decomp_partition(const OmegaGlobalMesh *G, int NTasks, int Method, int *CellTask, long long *EdgeCut) {
   OMEGA_REQUIRE(G && CellTask && NTasks >= 1, "omega_decomp_partition: bad argument");
   const int N = G->NCells;
   std::vector<int> idx(N);
   std::iota(idx.begin(), idx.end(), 0);
   const double *X[3] = {G->XCell, G->YCell, G->ZCell};
   // recursive coordinate bisection (the same splits as decomp.partition_cells)
   struct Job { int lo, hi, t0, n; };
   std::vector<Job> stack{{0, N, 0, NTasks}};
   while (!stack.empty()) {
      const Job j = stack.back();
      stack.pop_back();
      if (j.n == 1) {
         for (int i = j.lo; i < j.hi; ++i) CellTask[idx[i]] = j.t0;
         continue;
      }
      const int nl = j.n / 2, m = j.hi - j.lo;
      int ax = 0;
      double best = -1;
      for (int a = 0; a < 3; ++a) {
         double lo = 1e300, hi = -1e300;
         for (int i = j.lo; i < j.hi; ++i) { lo = std::min(lo, X[a][idx[i]]); hi = std::max(hi, X[a][idx[i]]); }
         if (hi - lo > best) { best = hi - lo; ax = a; }
      }
      const int k = (int)std::nearbyint((double)m * nl / j.n); // round half to even, as Python's round()
      std::stable_sort(idx.begin() + j.lo, idx.begin() + j.hi, [&](int a, int b) { return X[ax][a] < X[ax][b]; });
      stack.push_back({j.lo + k, j.hi, j.t0 + nl, j.n - nl});
      stack.push_back({j.lo, j.lo + k, j.t0, nl});
   }
   if (Method == 1 && NTasks > 1) {
      // boundary refinement: a cell moves to the part most of its neighbours are in when that lowers the cut
      // and neither part leaves [floor(avg*0.99), ceil(avg*1.01)]
      std::vector<int> size(NTasks, 0);
      for (int c = 0; c < N; ++c) ++size[CellTask[c]];
      const double avg = (double)N / NTasks;
      const int lo = (int)std::floor(avg * 0.99), hi = (int)std::ceil(avg * 1.01);
      for (int pass = 0; pass < 8; ++pass) {
         int moved = 0;
         for (int c = 0; c < N; ++c) {
            const int t = CellTask[c];
            int cnt_own = 0, best_t = -1, best_cnt = 0;
            int nt[12], nc[12], nn = 0;
            const int *row = G->CellsOnCell + (size_t)c * G->MaxEdges;
            for (int j = 0; j < G->MaxEdges && j < 12; ++j) {
               if (row[j] < 0) continue;
               const int u = CellTask[row[j]];
               if (u == t) { ++cnt_own; continue; }
               int q = 0;
               while (q < nn && nt[q] != u) ++q;
               if (q == nn) { nt[nn] = u; nc[nn] = 0; ++nn; }
               if (++nc[q] > best_cnt) { best_cnt = nc[q]; best_t = u; }
            }
            if (best_t >= 0 && best_cnt > cnt_own && size[t] - 1 >= lo && size[best_t] + 1 <= hi) {
               CellTask[c] = best_t;
               --size[t]; ++size[best_t];
               ++moved;
            }
         }
         if (!moved) break;
      }
   }
   if (EdgeCut) return omega_decomp_edge_cut(G, CellTask, EdgeCut);
   return 0;
}

int omega_decomp_create(const OmegaGlobalMesh *G, const int *CellTask, int NTasks, int MyTask, int HaloWidth, OmegaDecomp **Out) {
   OMEGA_REQUIRE(G && CellTask && Out && NTasks >= 1 && MyTask >= 0 && MyTask < NTasks && HaloWidth >= 1,
                 "omega_decomp_create: bad argument");
   OmegaDecomp *D = new OmegaDecomp();
   D->G = *G; D->NTasks = NTasks; D->MyTask = MyTask; D->HW = HaloWidth;
   const int NC = G->NCells;
   std::vector<std::vector<int>> owned_of(NTasks);
   std::vector<int> loc_on_owner(NC);
   for (int c = 0; c < NC; ++c) {
      loc_on_owner[c] = (int)owned_of[CellTask[c]].size();
      owned_of[CellTask[c]].push_back(c);
   }
   D->L = local_order(*G, CellTask, MyTask, HaloWidth, owned_of);
   const LocalOrder &L = D->L;
   // candidate neighbours: owners of cells within HaloWidth + 2 layers (edges and vertices reach one layer further)
   std::set<int> cand;
   {
      LocalOrder W = local_order(*G, CellTask, MyTask, HaloWidth + 2, owned_of);
      for (size_t r = W.NOwned[0]; r < W.Ids[0].size(); ++r) cand.insert(CellTask[W.Ids[0][r]]);
      cand.erase(MyTask);
   }
   std::map<int, LocalOrder> nbr;
   for (int t : cand) nbr.emplace(t, local_order(*G, CellTask, t, HaloWidth, owned_of));
   // index of a global edge / vertex on its owner, for the owners we know
   const int NX[3] = {NC, G->NEdges, G->NVertices};
   std::vector<int> addr[3];
   for (int k = 1; k < 3; ++k) {
      addr[k].assign(NX[k], -1);
      for (int i = 0; i < L.NOwned[k]; ++i) addr[k][L.Ids[k][i]] = i;
      for (auto &kv : nbr)
         for (int i = 0; i < kv.second.NOwned[k]; ++i) addr[k][kv.second.Ids[k][i]] = i;
   }
   auto owner_of = [&](int k, int x) {
      if (k == 0) return CellTask[x];
      return k == 1 ? first_valid_owner(G->CellsOnEdge + (size_t)x * 2, 2, CellTask)
                    : first_valid_owner(G->CellsOnVertex + (size_t)x * G->VertexDegree, G->VertexDegree, CellTask);
   };
   std::set<int> nbset;
   std::map<int, std::vector<int>> recv[3], send[3];
   for (int k = 0; k < 3; ++k) {
      const int nAll = (int)L.Ids[k].size();
      D->Loc[k].assign((size_t)(nAll + 1) * 2, 0);
      for (int i = 0; i < nAll; ++i) {
         const int x = L.Ids[k][i], o = owner_of(k, x);
         const int a = k == 0 ? loc_on_owner[x] : addr[k][x];
         if (a < 0) {
            set_error("omega_decomp_create: owner of a halo element is outside the candidate neighbour set");
            delete D;
            return 2;
         }
         D->Loc[k][2 * i] = o; D->Loc[k][2 * i + 1] = a;
         if (i >= L.NOwned[k] && o != MyTask) { recv[k][o].push_back(i); nbset.insert(o); } // ascending local index
      }
      // boundary slot: (task, NAll) for edges and vertices; the reference never writes the cell entry (zero)
      if (k != 0) { D->Loc[k][2 * nAll] = MyTask; D->Loc[k][2 * nAll + 1] = nAll; }
      // what the others receive from me, in THEIR halo order, as MY local indices = my owned index
      for (auto &kv : nbr) {
         const LocalOrder &N = kv.second;
         for (size_t i = N.NOwned[k]; i < N.Ids[k].size(); ++i) {
            const int x = N.Ids[k][i];
            if (owner_of(k, x) == MyTask) {
               send[k][kv.first].push_back(k == 0 ? loc_on_owner[x] : addr[k][x]);
               nbset.insert(kv.first);
            }
         }
      }
   }
   D->Neighbors.assign(nbset.begin(), nbset.end());
   for (int k = 0; k < 3; ++k) {
      D->SendOff[k].assign(1, 0); D->RecvOff[k].assign(1, 0);
      for (int t : D->Neighbors) {
         const std::vector<int> &s = send[k][t], &r = recv[k][t];
         D->SendIdx[k].insert(D->SendIdx[k].end(), s.begin(), s.end());
         D->RecvIdx[k].insert(D->RecvIdx[k].end(), r.begin(), r.end());
         D->SendOff[k].push_back((int)D->SendIdx[k].size());
         D->RecvOff[k].push_back((int)D->RecvIdx[k].size());
      }
   }
   *Out = D;
   return 0;
}

int omega_decomp_destroy(OmegaDecomp *D) {
   delete D;
   return 0;
}

// Counts[k] for k = cell, edge, vertex: {NOwned, NAll, NHalo(0..HaloWidth-1)} -> 3 * (2 + HaloWidth) ints
int omega_decomp_counts(const OmegaDecomp *D, int *Counts) {
   OMEGA_REQUIRE(D && Counts, "omega_decomp_counts: null argument");
   const int w = 2 + D->HW;
   for (int k = 0; k < 3; ++k) {
      Counts[k * w] = D->L.NOwned[k];
      Counts[k * w + 1] = (int)D->L.Ids[k].size();
      for (int h = 0; h < D->HW; ++h) Counts[k * w + 2 + h] = h < (int)D->L.NHalo[k].size() ? D->L.NHalo[k][h] : 0;
   }
   return 0;
}

int omega_decomp_ids(const OmegaDecomp *D, OmegaMeshElement Elem, int *IDs) {
   OMEGA_REQUIRE(D && IDs && Elem >= 0 && Elem < 3, "omega_decomp_ids: bad argument");
   std::memcpy(IDs, D->L.Ids[Elem].data(), sizeof(int) * D->L.Ids[Elem].size());
   return 0;
}

int omega_decomp_loc(const OmegaDecomp *D, OmegaMeshElement Elem, int *Loc) {
   OMEGA_REQUIRE(D && Loc && Elem >= 0 && Elem < 3, "omega_decomp_loc: bad argument");
   std::memcpy(Loc, D->Loc[Elem].data(), sizeof(int) * D->Loc[Elem].size());
   return 0;
}

// Which: 0 CellsOnCell, 1 EdgesOnCell, 2 VerticesOnCell, 3 CellsOnEdge, 4 VerticesOnEdge, 5 EdgesOnEdge,
// 6 CellsOnVertex, 7 EdgesOnVertex.  Out has (NAll + 1) rows of the array's width; the last row and every
// entry that is not local hold the boundary slot of the target index space (Decomp.cpp:552-712).
int omega_decomp_connectivity(const OmegaDecomp *D, int Which, int *Out) {
   OMEGA_REQUIRE(D && Out && Which >= 0 && Which < 8, "omega_decomp_connectivity: bad argument");
   const OmegaGlobalMesh &G = D->G;
   const int *src[8] = {G.CellsOnCell, G.EdgesOnCell, G.VerticesOnCell, G.CellsOnEdge, G.VerticesOnEdge, G.EdgesOnEdge,
                        G.CellsOnVertex, G.EdgesOnVertex};
   const int width[8] = {G.MaxEdges, G.MaxEdges, G.MaxEdges, 2, 2, G.MaxEdges2, G.VertexDegree, G.VertexDegree};
   const int rowKind[8] = {0, 0, 0, 1, 1, 1, 2, 2}, colKind[8] = {0, 1, 2, 0, 2, 1, 0, 1};
   const int NX[3] = {G.NCells, G.NEdges, G.NVertices};
   const int rk = rowKind[Which], ck = colKind[Which], W = width[Which];
   const int nAllCol = (int)D->L.Ids[ck].size(), nAllRow = (int)D->L.Ids[rk].size();
   std::vector<int> g2l(NX[ck], nAllCol);
   for (int i = 0; i < nAllCol; ++i) g2l[D->L.Ids[ck][i]] = i;
   for (int r = 0; r < nAllRow; ++r) {
      const int *row = src[Which] + (size_t)D->L.Ids[rk][r] * W;
      for (int j = 0; j < W; ++j) Out[(size_t)r * W + j] = row[j] >= 0 ? g2l[row[j]] : nAllCol;
   }
   for (int j = 0; j < W; ++j) Out[(size_t)nAllRow * W + j] = nAllCol;
   return 0;
}

int omega_decomp_halo_neighbors(const OmegaDecomp *D, int *NNeighbors, int *Tasks) {
   OMEGA_REQUIRE(D && NNeighbors, "omega_decomp_halo_neighbors: null argument");
   *NNeighbors = (int)D->Neighbors.size();
   if (Tasks) std::memcpy(Tasks, D->Neighbors.data(), sizeof(int) * D->Neighbors.size());
   return 0;
}

// CSR lists per element kind in neighbour order (the layout of OmegaHaloDesc); pass null index
// pointers to get the offsets (and with them the sizes) first
int omega_decomp_halo_lists(const OmegaDecomp *D, OmegaMeshElement Elem, int *SendOff, int *SendIdx, int *RecvOff, int *RecvIdx) {
   OMEGA_REQUIRE(D && Elem >= 0 && Elem < 3, "omega_decomp_halo_lists: bad argument");
   if (SendOff) std::memcpy(SendOff, D->SendOff[Elem].data(), sizeof(int) * D->SendOff[Elem].size());
   if (RecvOff) std::memcpy(RecvOff, D->RecvOff[Elem].data(), sizeof(int) * D->RecvOff[Elem].size());
   if (SendIdx) std::memcpy(SendIdx, D->SendIdx[Elem].data(), sizeof(int) * D->SendIdx[Elem].size());
   if (RecvIdx) std::memcpy(RecvIdx, D->RecvIdx[Elem].data(), sizeof(int) * D->RecvIdx[Elem].size());
   return 0;
}

} // extern "C"
