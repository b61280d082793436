// ref_decomp_driver.cpp -- TEST INFRASTRUCTURE.  Runs the reference's own Decomp constructor
// (src/base/Decomp.cpp:445-760, compiled unmodified over decomp_shim.h) as NTasks threads of one process on a
// global mesh held in memory, with the cell partition supplied by the caller, and hands every task's local
// numbering back through a C interface (tests/test_decomp_vs_reference.py).
#define private public // the constructors are private (create() would store the objects in static maps), so are the lists
#include "Decomp.h"
#include "Halo.h"
#include "HorzMesh.h"
#undef private

#include <thread>

using namespace OMEGA;

static std::vector<std::unique_ptr<Decomp>> g_decomps;
static std::vector<std::unique_ptr<Halo>> g_halos; // src/base/Halo.cpp:150-280 on the decompositions above
static std::vector<std::unique_ptr<HorzMesh>> g_meshes; // src/ocn/HorzMesh.cpp:49-150 on the decompositions above
static const int *g_cell_task = nullptr;
static int g_ncells = 0;

extern "C" int METIS_PartGraphKway(idx_t *nvtxs, idx_t *, idx_t *, idx_t *, idx_t *, idx_t *, idx_t *, idx_t *, real_t *, real_t *,
                                   idx_t *, idx_t *edgecut, idx_t *part) {
   for (int i = 0; i < *nvtxs; ++i) part[i] = g_cell_task[i];
   *edgecut = 0;
   return METIS_OK;
}
extern "C" int ParMETIS_V3_PartKway(idx_t *, idx_t *, idx_t *, idx_t *, idx_t *, idx_t *, idx_t *, idx_t *, idx_t *, real_t *, real_t *,
                                    idx_t *, idx_t *, idx_t *, MPI_Comm *) {
   std::fprintf(stderr, "ref_decomp: the ParMETIS method is not emulated\n");
   std::abort();
}

extern "C" {

// Connectivity as a mesh file holds it: 1-based global IDs, 0 = none.
int ref_decomp_run(int NTasks, int HaloWidth, int NCells, int NEdges, int NVertices, int MaxEdges, int VertexDegree,
                   const int *CellsOnCell, const int *EdgesOnCell, const int *VerticesOnCell, const int *CellsOnEdge,
                   const int *EdgesOnEdge, const int *VerticesOnEdge, const int *CellsOnVertex, const int *EdgesOnVertex,
                   const int *CellTask) {
   static bool kokkos = false;
   if (!kokkos) {
      Kokkos::initialize();
      kokkos = true;
   }
   IO::RefMeshFile &F = IO::RefMeshFile::get();
   F.Dims = {{"NCells", NCells}, {"NEdges", NEdges}, {"NVertices", NVertices}, {"MaxEdges", MaxEdges},
             {"VertexDegree", VertexDegree}, {"MaxCellsOnEdge", 2}};
   F.Arrays = {{"CellsOnCell", CellsOnCell}, {"EdgesOnCell", EdgesOnCell}, {"VerticesOnCell", VerticesOnCell},
               {"CellsOnEdge", CellsOnEdge}, {"EdgesOnEdge", EdgesOnEdge}, {"VerticesOnEdge", VerticesOnEdge},
               {"CellsOnVertex", CellsOnVertex}, {"EdgesOnVertex", EdgesOnVertex}};
   g_cell_task = CellTask;
   g_ncells    = NCells;
   RefDecompWorld &W = RefDecompWorld::get();
   W.NumTasks = NTasks;
   W.Arrived  = 0;
   RefMpiWorld &MW = RefMpiWorld::get();
   MW.NumTasks = NTasks;
   MW.Arrived  = 0;
   MW.Box.clear();
   MW.A2A.clear();
   g_meshes.clear();
   g_halos.clear();
   g_halos.resize(NTasks);
   g_decomps.clear();
   g_decomps.resize(NTasks);
   std::vector<MachEnv> envs;
   for (int t = 0; t < NTasks; ++t) envs.emplace_back(t, NTasks);
   std::vector<std::thread> th;
   for (int t = 0; t < NTasks; ++t)
      th.emplace_back([&, t] {
         RefMpiWorld::rank() = t;
         g_decomps[t].reset(new Decomp("RefDecomp" + std::to_string(t), &envs[t], NTasks, PartMethodMetisKWay, HaloWidth, "memory"));
         g_halos[t].reset(new Halo("RefHalo" + std::to_string(t), &envs[t], g_decomps[t].get()));
      });
   for (auto &x : th) x.join();
   return 0;
}

// {NOwned, NAll, NHalo(0 .. HaloWidth-1)} for cells, edges, vertices: 3 * (2 + HaloWidth) ints
int ref_decomp_counts(int Task, int *Out) {
   const Decomp &D = *g_decomps[Task];
   const int HW = D.HaloWidth;
   int *o = Out;
   *o++ = D.NCellsOwned; *o++ = D.NCellsAll;
   for (int h = 0; h < HW; ++h) *o++ = D.NCellsHaloH(h);
   *o++ = D.NEdgesOwned; *o++ = D.NEdgesAll;
   for (int h = 0; h < HW; ++h) *o++ = D.NEdgesHaloH(h);
   *o++ = D.NVerticesOwned; *o++ = D.NVerticesAll;
   for (int h = 0; h < HW; ++h) *o++ = D.NVerticesHaloH(h);
   return 0;
}

// Which: 0 CellID, 1 EdgeID, 2 VertexID [N*Size]; 3 CellLoc, 4 EdgeLoc, 5 VertexLoc [N*Size][2];
// 10 CellsOnCell, 11 EdgesOnCell, 12 VerticesOnCell [NCellsSize][MaxEdges]; 13 CellsOnEdge, 14 VerticesOnEdge
// [NEdgesSize][2]; 15 EdgesOnEdge [NEdgesSize][2 MaxEdges]; 16 CellsOnVertex, 17 EdgesOnVertex
// [NVerticesSize][VertexDegree]; 18 NEdgesOnCell [NCellsSize]; 19 NEdgesOnEdge [NEdgesSize].  Returns the count.
int ref_decomp_get(int Task, int Which, int *Out) {
   const Decomp &D = *g_decomps[Task];
   auto c1 = [&](const HostArray1DI4 &A) { const int n = (int)A.extent(0); if (Out) for (int i = 0; i < n; ++i) Out[i] = A(i); return n; };
   auto c2 = [&](const HostArray2DI4 &A) {
      const int n = (int)A.extent(0), m = (int)A.extent(1);
      if (Out) for (int i = 0; i < n; ++i) for (int j = 0; j < m; ++j) Out[i * m + j] = A(i, j);
      return n * m;
   };
   switch (Which) {
   case 0: return c1(D.CellIDH);
   case 1: return c1(D.EdgeIDH);
   case 2: return c1(D.VertexIDH);
   case 3: return c2(D.CellLocH);
   case 4: return c2(D.EdgeLocH);
   case 5: return c2(D.VertexLocH);
   case 10: return c2(D.CellsOnCellH);
   case 11: return c2(D.EdgesOnCellH);
   case 12: return c2(D.VerticesOnCellH);
   case 13: return c2(D.CellsOnEdgeH);
   case 14: return c2(D.VerticesOnEdgeH);
   case 15: return c2(D.EdgesOnEdgeH);
   case 16: return c2(D.CellsOnVertexH);
   case 17: return c2(D.EdgesOnVertexH);
   case 18: return c1(D.NEdgesOnCellH);
   case 19: return c1(D.NEdgesOnEdgeH);
   }
   return -1;
}

// Geometry of the global mesh for HorzMesh (file variable name -> global array; a null pointer = all zeros).  Call before
// ref_horzmesh_run; the names are those of HorzMesh.cpp:420-520 ("xCell", "areaCell", "dcEdge", "weightsOnEdge", ...).
int ref_mesh_set_r8(const char *Name, const double *Ptr) {
   IO::RefMeshFile::get().ArraysR8[Name] = Ptr;
   return 0;
}
// HorzMesh::HorzMesh (reads its local geometry through the linear IO decompositions it builds from the Decomp's global
// IDs, then computeEdgeSign, setMasks, setMeshScaling) for every task of the last ref_decomp_run
int ref_horzmesh_run(int NVertLayers) {
   g_meshes.clear();
   for (size_t t = 0; t < g_decomps.size(); ++t)
      g_meshes.emplace_back(new HorzMesh("RefMesh" + std::to_string(t), g_decomps[t].get(), NVertLayers));
   return 0;
}
// Which: 0 EdgeSignOnCell [NCellsSize][MaxEdges], 1 EdgeSignOnVertex [NVerticesSize][VertexDegree], 2 EdgeMask
// [NEdgesSize][NVertLayers], 3 AreaCell, 4 DcEdge, 5 DvEdge, 6 AngleEdge, 7 FEdge, 8 FVertex, 9 AreaTriangle,
// 10 BottomDepth, 11 WeightsOnEdge [NEdgesSize][2 MaxEdges], 12 KiteAreasOnVertex [NVerticesSize][VertexDegree],
// 13 MeshScalingDel2, 14 MeshScalingDel4, 15 XCell, 16 YCell, 17 XEdge, 18 YEdge.  Returns the count.
int ref_horzmesh_get(int Task, int Which, double *Out) {
   const HorzMesh &M = *g_meshes[Task];
   auto c1 = [&](const HostArray1DReal &A) { const int n = (int)A.extent(0); if (Out) for (int i = 0; i < n; ++i) Out[i] = A(i); return n; };
   auto c2 = [&](const HostArray2DReal &A) {
      const int n = (int)A.extent(0), m = (int)A.extent(1);
      if (Out) for (int i = 0; i < n; ++i) for (int j = 0; j < m; ++j) Out[i * m + j] = A(i, j);
      return n * m;
   };
   switch (Which) {
   case 0: return c2(M.EdgeSignOnCellH);
   case 1: return c2(M.EdgeSignOnVertexH);
   case 2: return c2(createHostMirrorCopy(M.EdgeMask));
   case 3: return c1(M.AreaCellH);
   case 4: return c1(M.DcEdgeH);
   case 5: return c1(M.DvEdgeH);
   case 6: return c1(M.AngleEdgeH);
   case 7: return c1(M.FEdgeH);
   case 8: return c1(M.FVertexH);
   case 9: return c1(M.AreaTriangleH);
   case 10: return c1(M.BottomDepthH);
   case 11: return c2(M.WeightsOnEdgeH);
   case 12: return c2(M.KiteAreasOnVertexH);
   case 13: return c1(createHostMirrorCopy(M.MeshScalingDel2));
   case 14: return c1(createHostMirrorCopy(M.MeshScalingDel4));
   case 15: return c1(M.XCellH);
   case 16: return c1(M.YCellH);
   case 17: return c1(M.XEdgeH);
   case 18: return c1(M.YEdgeH);
   }
   return -1;
}

// the reference Halo's neighbour tasks (ascending); returns their number
int ref_halo_neighbors(int Task, int *Out) {
   const Halo &H = *g_halos[Task];
   if (Out) for (size_t i = 0; i < H.NeighborList.size(); ++i) Out[i] = H.NeighborList[i];
   return (int)H.NeighborList.size();
}
// exchange list of neighbour number Nbr (position in the list above), index space Elem, Dir 0 = send, 1 = receive:
// local indices concatenated over the halo layers (Halo.cpp:40-110), per-layer counts in NHalo[HaloWidth]; returns NTot
int ref_halo_list(int Task, int Nbr, int Elem, int Dir, int *Idx, int *NHalo) {
   const Halo &H = *g_halos[Task];
   const auto &L = Dir ? H.Neighbors[Nbr].RecvLists[Elem] : H.Neighbors[Nbr].SendLists[Elem];
   if (Idx) for (int i = 0; i < L.NTot; ++i) Idx[i] = L.IndexH(i);
   if (NHalo) for (size_t h = 0; h < L.NHalo.size(); ++h) NHalo[h] = L.NHalo[h];
   return L.NTot;
}

// Halo::exchangeFullArrayHalo (src/base/Halo.h:768-915) on every task at once: Arrays[t] is task t's array
// [NOuter][N*Size][NInner] (NOuter == 0: a 2-D array [N*Size][NInner]; NInner == 0 too: 1-D), double.
int ref_halo_exchange(int Elem, int NOuter, int NInner, double **Arrays) {
   const int NTasks = (int)g_halos.size();
   std::vector<std::thread> th;
   std::vector<int> err(NTasks, 0);
   for (int t = 0; t < NTasks; ++t)
      th.emplace_back([&, t] {
         RefMpiWorld::rank() = t;
         const Decomp &D = *g_decomps[t];
         const int N = Elem == 0 ? D.NCellsSize : Elem == 1 ? D.NEdgesSize : D.NVerticesSize;
         if (NOuter > 0) {
            Array3DR8 A(Arrays[t], NOuter, N, NInner);
            err[t] = g_halos[t]->exchangeFullArrayHalo(A, (MeshElement)Elem);
         } else if (NInner > 0) {
            Array2DR8 A(Arrays[t], N, NInner);
            err[t] = g_halos[t]->exchangeFullArrayHalo(A, (MeshElement)Elem);
         } else {
            Array1DR8 A(Arrays[t], N);
            err[t] = g_halos[t]->exchangeFullArrayHalo(A, (MeshElement)Elem);
         }
      });
   for (auto &x : th) x.join();
   int e = 0;
   for (int v : err) e |= v;
   return e;
}
// what task Task sent to its neighbour number Nbr in the last exchange: the packed message (Halo.h:304-502)
int ref_halo_send_buffer(int Task, int Nbr, int Count, double *Out) {
   const Halo &H = *g_halos[Task];
   const auto &B = H.Neighbors[Nbr].SendBufferH;
   const int n = std::min<int>(Count, (int)B.extent(0));
   for (int i = 0; i < n; ++i) Out[i] = B(i);
   return (int)B.extent(0);
}

} // extern "C"
