// OmegaB200Decomp.h -- C++ host mirror of the reference's Decomp (src/base/Decomp.h:62-322) over the omega_decomp_*
// entry points of include/omega_b200.h.  Same public members (sizes, global IDs, owner addresses and
// local connectivity, the H-suffixed host arrays of Decomp.h:190-282), same create / get / getDefault / erase / clear
// registry (Decomp.h:285-311) and valid*ID queries (Decomp.cpp:46-54).  Host code only: the device copies the reference
// keeps next to each host array (Decomp.cpp:716-740) are the SFC-ordered tables omega_mesh_create builds from these.
// What differs, and why:
//   * the global mesh arrives as arrays (OmegaGlobalMesh, 0-based, -1 = none) instead of a file name: reading the mesh
//     file is the host's job (omega_b200/ncio.py, or Omega's own IO), this library never opens files;
//   * a rank numbers its partition alone from the global connectivity (csrc/decomp.cpp), so there is no MachEnv
//     argument: NTasks and MyTask are plain integers;
//   * the partitioner is this library's (recursive coordinate bisection, optionally refined), not METIS; a host that
//     has METIS hands its own CellTask array to the second create().
#pragma once
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "OmegaB200.h"

namespace OMEGA {

/// Decomp.h:40-45; the two methods of this library take the place of the METIS ones
enum PartMethod { PartMethodUnknown, PartMethodRCB, PartMethodRCBRefined };
inline PartMethod getPartMethodFromStr(const std::string &In) { // Decomp.cpp:1680-1700
   if (In == "RCB" || In == "rcb") return PartMethodRCB;
   if (In == "RCBRefined" || In == "rcbrefined") return PartMethodRCBRefined;
   return PartMethodUnknown;
}

/// rows x cols host array of I4 with the View call operator the reference's tests use (HostArray2DI4)
struct HostArray2DI4 {
   std::vector<I4> V;
   int N0 = 0, N1 = 0;
   void resize(int A, int B) { N0 = A; N1 = B; V.assign((size_t)A * B, 0); }
   I4 &operator()(int I, int J) { return V[(size_t)I * N1 + J]; }
   const I4 &operator()(int I, int J) const { return V[(size_t)I * N1 + J]; }
   int extent_int(int D) const { return D == 0 ? N0 : N1; }
   I4 *data() { return V.data(); }
   const I4 *data() const { return V.data(); }
};
struct HostArray1DI4 {
   std::vector<I4> V;
   void resize(int A) { V.assign((size_t)A, 0); }
   I4 &operator()(int I) { return V[I]; }
   const I4 &operator()(int I) const { return V[I]; }
   int extent_int(int) const { return (int)V.size(); }
   I4 *data() { return V.data(); }
   const I4 *data() const { return V.data(); }
};

class Decomp {
 public:
   // ---- Decomp.h:176-282 (host arrays; IDs 1-based as upstream, boundary entries NGlobal + 1) ----
   I4 HaloWidth;
   I4 NCellsGlobal, NCellsOwned, NCellsAll, NCellsSize, MaxEdges;
   HostArray1DI4 NCellsHaloH, CellIDH;
   HostArray2DI4 CellLocH;
   I4 NEdgesGlobal, NEdgesOwned, NEdgesAll, NEdgesSize, MaxCellsOnEdge;
   HostArray1DI4 NEdgesHaloH, EdgeIDH;
   HostArray2DI4 EdgeLocH;
   I4 NVerticesGlobal, NVerticesOwned, NVerticesAll, NVerticesSize, VertexDegree;
   HostArray1DI4 NVerticesHaloH, VertexIDH;
   HostArray2DI4 VertexLocH;
   HostArray2DI4 CellsOnCellH, EdgesOnCellH, VerticesOnCellH, CellsOnEdgeH, EdgesOnEdgeH, VerticesOnEdgeH, CellsOnVertexH,
       EdgesOnVertexH;
   HostArray1DI4 NEdgesOnCellH, NEdgesOnEdgeH;
   I4 MaxEdges2, NTasks, MyTask;
   long long EdgeCut = -1;     ///< cut edges of the partition used
   std::vector<I4> CellTask;   ///< owner of every global cell

   // ---- halo lists (src/base/Halo.cpp:455-600), in the layout omega_halo_create takes ----
   std::vector<I4> NeighborTasks;
   std::vector<I4> SendOffsets[3], SendIdx[3], RecvOffsets[3], RecvIdx[3];
   OmegaHaloDesc haloDesc() const {
      OmegaHaloDesc D{};
      D.MyTask = MyTask;
      D.NNeighbors = (int)NeighborTasks.size();
      D.NeighborTasks = NeighborTasks.data();
      for (int k = 0; k < 3; ++k) {
         D.SendOffsets[k] = SendOffsets[k].data(); D.SendIdx[k] = SendIdx[k].data();
         D.RecvOffsets[k] = RecvOffsets[k].data(); D.RecvIdx[k] = RecvIdx[k].data();
      }
      return D;
   }
   /// the index part of an OmegaMeshDesc (sizes and local connectivity); the geometry is HorzMesh's to fill
   void fillMeshDesc(OmegaMeshDesc &D, I4 NVertLayers) const {
      D.NCellsOwned = NCellsOwned; D.NCellsAll = NCellsAll; D.NCellsSize = NCellsSize;
      D.NEdgesOwned = NEdgesOwned; D.NEdgesAll = NEdgesAll; D.NEdgesSize = NEdgesSize;
      D.NVerticesOwned = NVerticesOwned; D.NVerticesAll = NVerticesAll; D.NVerticesSize = NVerticesSize;
      D.MaxEdges = MaxEdges; D.MaxEdges2 = MaxEdges2; D.VertexDegree = VertexDegree; D.NVertLayers = NVertLayers;
      D.NEdgesOnCell = NEdgesOnCellH.data(); D.EdgesOnCell = EdgesOnCellH.data();
      D.CellsOnEdge = CellsOnEdgeH.data(); D.VerticesOnEdge = VerticesOnEdgeH.data();
      D.NEdgesOnEdge = NEdgesOnEdgeH.data(); D.EdgesOnEdge = EdgesOnEdgeH.data();
      D.CellsOnVertex = CellsOnVertexH.data(); D.EdgesOnVertex = EdgesOnVertexH.data();
   }

   // ---- registry, Decomp.h:285-311 / Decomp.cpp:1704-1760 ----
   /// partitions with this library's partitioner
   static Decomp *create(const std::string &Name, const OmegaGlobalMesh &Mesh, I4 NParts, I4 MyTask, PartMethod Method, I4 HaloWidth) {
      if (Method != PartMethodRCB && Method != PartMethodRCBRefined) abortError("Decomp::create: unknown partition method");
      std::vector<I4> Task((size_t)Mesh.NCells);
      long long Cut = 0;
      OMEGA_B200_CHECK(omega_decomp_partition(&Mesh, NParts, Method == PartMethodRCBRefined ? 1 : 0, Task.data(), &Cut),
                       "Decomp::create: omega_decomp_partition");
      return create(Name, Mesh, Task.data(), NParts, MyTask, HaloWidth);
   }
   /// takes the host's partition (e.g. from METIS): CellTask[NCells]
   static Decomp *create(const std::string &Name, const OmegaGlobalMesh &Mesh, const I4 *CellTask, I4 NParts, I4 MyTask, I4 HaloWidth) {
      if (all().count(Name)) { // Decomp.cpp:1711-1715
         std::fprintf(stderr, "[omega_b200] Attempted to create a Decomp with name %s but it already exists\n", Name.c_str());
         return nullptr;
      }
      auto *D = new Decomp(Mesh, CellTask, NParts, MyTask, HaloWidth);
      all()[Name].reset(D);
      return D;
   }
   /// Decomp::init builds the default decomposition (Decomp.cpp:1704-1740: options from the Decomp config group)
   static void init(const OmegaGlobalMesh &Mesh, I4 NParts, I4 MyTask, I4 HaloWidth = 3, PartMethod Method = PartMethodRCBRefined) {
      defaultPtr() = create("Default", Mesh, NParts, MyTask, Method, HaloWidth);
   }
   static Decomp *getDefault() { return defaultPtr(); }
   static Decomp *get(const std::string &Name) {
      auto It = all().find(Name);
      if (It == all().end()) {
         std::fprintf(stderr, "[omega_b200] Attempt to retrieve non-existent Decomp: %s\n", Name.c_str());
         return nullptr;
      }
      return It->second.get();
   }
   static void erase(const std::string &Name) {
      if (defaultPtr() && all().count(Name) && all()[Name].get() == defaultPtr()) defaultPtr() = nullptr;
      all().erase(Name);
   }
   static void clear() {
      all().clear();
      defaultPtr() = nullptr;
   }
   bool validCellID(I4 ID) const { return ID > 0 && ID <= NCellsGlobal; }       // Decomp.cpp:46-48
   bool validEdgeID(I4 ID) const { return ID > 0 && ID <= NEdgesGlobal; }       // Decomp.cpp:49-51
   bool validVertexID(I4 ID) const { return ID > 0 && ID <= NVerticesGlobal; }  // Decomp.cpp:52-54

   Decomp(const Decomp &) = delete;
   Decomp(Decomp &&)      = delete;

 private:
   static std::map<std::string, std::unique_ptr<Decomp>> &all() {
      static std::map<std::string, std::unique_ptr<Decomp>> M;
      return M;
   }
   static Decomp *&defaultPtr() {
      static Decomp *D = nullptr;
      return D;
   }

   Decomp(const OmegaGlobalMesh &M, const I4 *Task, I4 NParts, I4 Me, I4 Width)
       : HaloWidth(Width), NCellsGlobal(M.NCells), MaxEdges(M.MaxEdges), NEdgesGlobal(M.NEdges), MaxCellsOnEdge(2),
         NVerticesGlobal(M.NVertices), VertexDegree(M.VertexDegree), MaxEdges2(M.MaxEdges2), NTasks(NParts), MyTask(Me),
         CellTask(Task, Task + M.NCells) {
      OMEGA_B200_CHECK(omega_decomp_edge_cut(&M, Task, &EdgeCut), "Decomp: omega_decomp_edge_cut");
      OmegaDecomp *H = nullptr;
      OMEGA_B200_CHECK(omega_decomp_create(&M, Task, NParts, Me, Width, &H), "Decomp: omega_decomp_create");
      std::vector<I4> Cnt(3 * (2 + (size_t)Width));
      OMEGA_B200_CHECK(omega_decomp_counts(H, Cnt.data()), "Decomp: omega_decomp_counts");
      const I4 *C = Cnt.data(), *E = C + 2 + Width, *V = E + 2 + Width;
      NCellsOwned = C[0]; NCellsAll = C[1]; NCellsSize = C[1] + 1;
      NEdgesOwned = E[0]; NEdgesAll = E[1]; NEdgesSize = E[1] + 1;
      NVerticesOwned = V[0]; NVerticesAll = V[1]; NVerticesSize = V[1] + 1;
      HostArray1DI4 *Halo[3] = {&NCellsHaloH, &NEdgesHaloH, &NVerticesHaloH};
      HostArray1DI4 *IDs[3]  = {&CellIDH, &EdgeIDH, &VertexIDH};
      HostArray2DI4 *Loc[3]  = {&CellLocH, &EdgeLocH, &VertexLocH};
      const I4 NAll[3] = {NCellsAll, NEdgesAll, NVerticesAll}, NGlob[3] = {NCellsGlobal, NEdgesGlobal, NVerticesGlobal};
      for (int k = 0; k < 3; ++k) {
         Halo[k]->resize(Width);
         for (int l = 0; l < Width; ++l) (*Halo[k])(l) = Cnt[k * (2 + Width) + 2 + l];
         IDs[k]->resize(NAll[k] + 1);
         OMEGA_B200_CHECK(omega_decomp_ids(H, (OmegaMeshElement)k, IDs[k]->data()), "Decomp: omega_decomp_ids");
         for (int i = 0; i < NAll[k]; ++i) (*IDs[k])(i) += 1;       // 1-based, as in the mesh file and upstream
         (*IDs[k])(NAll[k]) = k == 0 ? 0 : NGlob[k] + 1;            // upstream leaves the cell entry zero (Decomp.cpp:1094-1098)
         Loc[k]->resize(NAll[k] + 1, 2);
         OMEGA_B200_CHECK(omega_decomp_loc(H, (OmegaMeshElement)k, Loc[k]->data()), "Decomp: omega_decomp_loc");
      }
      HostArray2DI4 *Conn[8] = {&CellsOnCellH, &EdgesOnCellH, &VerticesOnCellH, &CellsOnEdgeH, &VerticesOnEdgeH, &EdgesOnEdgeH,
                                &CellsOnVertexH, &EdgesOnVertexH};
      const I4 Rows[8] = {NCellsSize, NCellsSize, NCellsSize, NEdgesSize, NEdgesSize, NEdgesSize, NVerticesSize, NVerticesSize};
      const I4 Cols[8] = {MaxEdges, MaxEdges, MaxEdges, 2, 2, MaxEdges2, VertexDegree, VertexDegree};
      for (int w = 0; w < 8; ++w) {
         Conn[w]->resize(Rows[w], Cols[w]);
         OMEGA_B200_CHECK(omega_decomp_connectivity(H, w, Conn[w]->data()), "Decomp: omega_decomp_connectivity");
      }
      // NEdgesOnCell: valid entries of the global row (Decomp.cpp:2043-2064); NEdgesOnEdge: entries that are valid or
      // literally 0 -- every one of a row read from a mesh file (Decomp.cpp:2186-2200); boundary rows zero
      NEdgesOnCellH.resize(NCellsSize);
      for (int i = 0; i < NCellsAll; ++i) {
         const I4 *Row = M.EdgesOnCell + (size_t)(CellIDH(i) - 1) * MaxEdges;
         I4 n = 0;
         for (int j = 0; j < MaxEdges; ++j) n += Row[j] >= 0;
         NEdgesOnCellH(i) = n;
      }
      NEdgesOnEdgeH.resize(NEdgesSize);
      for (int i = 0; i < NEdgesAll; ++i) NEdgesOnEdgeH(i) = MaxEdges2;
      // halo lists
      int NN = 0;
      OMEGA_B200_CHECK(omega_decomp_halo_neighbors(H, &NN, nullptr), "Decomp: omega_decomp_halo_neighbors");
      NeighborTasks.resize(NN);
      if (NN) OMEGA_B200_CHECK(omega_decomp_halo_neighbors(H, &NN, NeighborTasks.data()), "Decomp: omega_decomp_halo_neighbors");
      for (int k = 0; k < 3; ++k) {
         SendOffsets[k].assign(NN + 1, 0); RecvOffsets[k].assign(NN + 1, 0);
         OMEGA_B200_CHECK(omega_decomp_halo_lists(H, (OmegaMeshElement)k, SendOffsets[k].data(), nullptr, RecvOffsets[k].data(), nullptr),
                          "Decomp: omega_decomp_halo_lists");
         SendIdx[k].assign(SendOffsets[k][NN] > 0 ? SendOffsets[k][NN] : 1, 0);
         RecvIdx[k].assign(RecvOffsets[k][NN] > 0 ? RecvOffsets[k][NN] : 1, 0);
         OMEGA_B200_CHECK(omega_decomp_halo_lists(H, (OmegaMeshElement)k, nullptr, SendIdx[k].data(), nullptr, RecvIdx[k].data()),
                          "Decomp: omega_decomp_halo_lists");
      }
      omega_decomp_destroy(H);
   }
};

/// Global-order geometry arrays of the mesh file, the ones HorzMesh reads (src/ocn/HorzMesh.cpp:262-525)
struct GlobalGeometry {
   const double *AreaCell, *BottomDepth, *XCell, *YCell, *ZCell;     ///< [NCells]
   const double *DcEdge, *DvEdge, *AngleEdge, *FEdge, *XEdge, *YEdge; ///< [NEdges]
   const double *WeightsOnEdge;                                       ///< [NEdges][MaxEdges2]
   const double *AreaTriangle, *FVertex;                              ///< [NVertices]
   const double *KiteAreasOnVertex;                                   ///< [NVertices][VertexDegree]
};

/// Local[o][i][k] = Global[o][ID(i) - 1][k] for the NAll local elements of a kind, the boundary row zero: what the
/// reference's decomposed read of a mesh or state array leaves in a local array (HorzMesh.cpp:262-330)
inline void gatherField(const Decomp &D, MeshElement E, int NOuter, int NInner, const double *Global, double *Local) {
   const HostArray1DI4 &ID = E == OnCell ? D.CellIDH : E == OnEdge ? D.EdgeIDH : D.VertexIDH;
   const size_t NAll = E == OnCell ? D.NCellsAll : E == OnEdge ? D.NEdgesAll : D.NVerticesAll;
   const size_t NGlob = E == OnCell ? D.NCellsGlobal : E == OnEdge ? D.NEdgesGlobal : D.NVerticesGlobal;
   for (int o = 0; o < NOuter; ++o) {
      double *L = Local + (size_t)o * (NAll + 1) * NInner;
      const double *G = Global + (size_t)o * NGlob * NInner;
      for (size_t i = 0; i < NAll; ++i)
         for (int k = 0; k < NInner; ++k) L[i * NInner + k] = G[(size_t)(ID((int)i) - 1) * NInner + k];
      for (int k = 0; k < NInner; ++k) L[NAll * NInner + k] = 0.0;
   }
}

/// The geometry of one task's local mesh in Omega's local numbering
class LocalGeometry {
 public:
   std::vector<double> AreaCell, BottomDepth, XCell, YCell, ZCell, DcEdge, DvEdge, AngleEdge, FEdge, XEdge, YEdge, WeightsOnEdge,
       AreaTriangle, FVertex, KiteAreasOnVertex;
   LocalGeometry(const Decomp &D, const GlobalGeometry &G) {
      auto get = [&](std::vector<double> &Out, MeshElement E, int NInner, const double *Src) {
         const size_t N = E == OnCell ? D.NCellsSize : E == OnEdge ? D.NEdgesSize : D.NVerticesSize;
         Out.resize(N * NInner);
         gatherField(D, E, 1, NInner, Src, Out.data());
      };
      get(AreaCell, OnCell, 1, G.AreaCell); get(BottomDepth, OnCell, 1, G.BottomDepth);
      get(XCell, OnCell, 1, G.XCell); get(YCell, OnCell, 1, G.YCell); get(ZCell, OnCell, 1, G.ZCell);
      get(DcEdge, OnEdge, 1, G.DcEdge); get(DvEdge, OnEdge, 1, G.DvEdge); get(AngleEdge, OnEdge, 1, G.AngleEdge);
      get(FEdge, OnEdge, 1, G.FEdge); get(XEdge, OnEdge, 1, G.XEdge); get(YEdge, OnEdge, 1, G.YEdge);
      get(WeightsOnEdge, OnEdge, D.MaxEdges2, G.WeightsOnEdge);
      get(AreaTriangle, OnVertex, 1, G.AreaTriangle); get(FVertex, OnVertex, 1, G.FVertex);
      get(KiteAreasOnVertex, OnVertex, D.VertexDegree, G.KiteAreasOnVertex);
   }
   /// the geometry part of an OmegaMeshDesc (Decomp::fillMeshDesc fills the index part)
   void fillMeshDesc(OmegaMeshDesc &M) const {
      M.AreaCell = AreaCell.data(); M.BottomDepth = BottomDepth.data(); M.XCell = XCell.data(); M.YCell = YCell.data(); M.ZCell = ZCell.data();
      M.DcEdge = DcEdge.data(); M.DvEdge = DvEdge.data(); M.AngleEdge = AngleEdge.data(); M.FEdge = FEdge.data();
      M.XEdge = XEdge.data(); M.YEdge = YEdge.data(); M.WeightsOnEdge = WeightsOnEdge.data();
      M.AreaTriangle = AreaTriangle.data(); M.FVertex = FVertex.data(); M.KiteAreasOnVertex = KiteAreasOnVertex.data();
   }
};

} // namespace OMEGA
