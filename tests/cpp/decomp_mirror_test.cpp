// decomp_mirror_test.cpp -- drives omega_b200/host/OmegaB200Decomp.h the way the reference's DecompTest drives Decomp
// (test/base/DecompTest.cpp): create, retrieve, read the public members.  Usage:
//   decomp_mirror_test <dir> <NTasks> <MyTask> <HaloWidth> <Method>
// <dir> holds the global mesh as raw files (written by tests/test_decomp_mirror.py); every public array is written back
// to <dir>/out_<Name>.bin and the scalars to <dir>/out_meta.txt.  Host code only: no device is touched.
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <string>
#include <vector>

#include "OmegaB200Decomp.h"

using namespace OMEGA;

static std::string Dir;

template <class T> static std::vector<T> readBin(const std::string &Name) {
   std::ifstream F(Dir + "/" + Name + ".bin", std::ios::binary | std::ios::ate);
   if (!F) { std::fprintf(stderr, "missing %s\n", Name.c_str()); std::exit(2); }
   const std::streamsize N = F.tellg();
   F.seekg(0);
   std::vector<T> V((size_t)N / sizeof(T));
   F.read(reinterpret_cast<char *>(V.data()), N);
   return V;
}
static void writeBin(const std::string &Name, const I4 *P, size_t N) {
   std::ofstream F(Dir + "/out_" + Name + ".bin", std::ios::binary);
   F.write(reinterpret_cast<const char *>(P), (std::streamsize)(N * sizeof(I4)));
}
#define REQUIRE(c) \
   do { if (!(c)) { std::fprintf(stderr, "FAIL %s:%d: %s\n", __FILE__, __LINE__, #c); return 1; } } while (0)

int main(int argc, char **argv) {
   if (argc < 6) return 2;
   Dir = argv[1];
   const int NTasks = std::atoi(argv[2]), MyTask = std::atoi(argv[3]), Width = std::atoi(argv[4]), Method = std::atoi(argv[5]);
   std::vector<int> Dims = readBin<int>("dims"); // NCells NEdges NVertices MaxEdges MaxEdges2 VertexDegree
   auto CoC = readBin<int>("cellsOnCell"), EoC = readBin<int>("edgesOnCell"), VoC = readBin<int>("verticesOnCell");
   auto CoE = readBin<int>("cellsOnEdge"), VoE = readBin<int>("verticesOnEdge"), EoE = readBin<int>("edgesOnEdge");
   auto CoV = readBin<int>("cellsOnVertex"), EoV = readBin<int>("edgesOnVertex");
   auto X = readBin<double>("xCell"), Y = readBin<double>("yCell"), Z = readBin<double>("zCell");
   OmegaGlobalMesh G{};
   G.NCells = Dims[0]; G.NEdges = Dims[1]; G.NVertices = Dims[2]; G.MaxEdges = Dims[3]; G.MaxEdges2 = Dims[4]; G.VertexDegree = Dims[5];
   G.CellsOnCell = CoC.data(); G.EdgesOnCell = EoC.data(); G.VerticesOnCell = VoC.data();
   G.CellsOnEdge = CoE.data(); G.VerticesOnEdge = VoE.data(); G.EdgesOnEdge = EoE.data();
   G.CellsOnVertex = CoV.data(); G.EdgesOnVertex = EoV.data();
   G.XCell = X.data(); G.YCell = Y.data(); G.ZCell = Z.data();

   // ---- the registry, as DecompTest.cpp:52-75,260-283 uses it ----
   REQUIRE(Decomp::getDefault() == nullptr);
   Decomp::init(G, NTasks, MyTask, Width, Method ? PartMethodRCBRefined : PartMethodRCB);
   Decomp *D = Decomp::getDefault();
   REQUIRE(D != nullptr && Decomp::get("Default") == D);
   REQUIRE(Decomp::create("Default", G, NTasks, MyTask, PartMethodRCB, Width) == nullptr); // the name is taken
   REQUIRE(Decomp::get("NoSuchDecomp") == nullptr);
   // a second decomposition from a partition the host supplies, here the first one's: same numbering
   Decomp *D2 = Decomp::create("FromHostPartition", G, D->CellTask.data(), NTasks, MyTask, Width);
   REQUIRE(D2 != nullptr && D2 != D && Decomp::get("FromHostPartition") == D2);
   REQUIRE(D2->NCellsOwned == D->NCellsOwned && D2->CellIDH.V == D->CellIDH.V && D2->EdgesOnEdgeH.V == D->EdgesOnEdgeH.V);
   Decomp::erase("FromHostPartition");
   REQUIRE(Decomp::get("FromHostPartition") == nullptr && Decomp::getDefault() == D);
   REQUIRE(getPartMethodFromStr("RCB") == PartMethodRCB && getPartMethodFromStr("MetisKWay") == PartMethodUnknown);
   REQUIRE(D->validCellID(1) && D->validCellID(G.NCells) && !D->validCellID(0) && !D->validCellID(G.NCells + 1));
   REQUIRE(D->validEdgeID(G.NEdges) && !D->validEdgeID(G.NEdges + 1) && D->validVertexID(1) && !D->validVertexID(-3));

   // ---- DecompTest.cpp:126-136: the owned IDs are this task's share of 1..N ----
   long long Sum = 0;
   for (int i = 0; i < D->NCellsOwned; ++i) {
      REQUIRE(D->validCellID(D->CellIDH(i)) && D->CellLocH(i, 0) == MyTask && D->CellLocH(i, 1) == i);
      Sum += D->CellIDH(i);
   }
   for (int i = D->NCellsOwned; i < D->NCellsAll; ++i) REQUIRE(D->CellLocH(i, 0) != MyTask);
   REQUIRE(D->NCellsHaloH(Width - 1) == D->NCellsAll && D->NCellsSize == D->NCellsAll + 1);

   // ---- the descriptors the other mirrors take ----
   OmegaMeshDesc MD{};
   D->fillMeshDesc(MD, 4);
   REQUIRE(MD.NCellsAll == D->NCellsAll && MD.EdgesOnCell == D->EdgesOnCellH.data() && MD.NVertLayers == 4);
   OmegaHaloDesc HD = D->haloDesc();
   REQUIRE(HD.MyTask == MyTask && HD.NNeighbors == (int)D->NeighborTasks.size());

   std::ofstream Meta(Dir + "/out_meta.txt");
   Meta << "NCellsOwned " << D->NCellsOwned << "\nNCellsAll " << D->NCellsAll << "\nNEdgesOwned " << D->NEdgesOwned << "\nNEdgesAll "
        << D->NEdgesAll << "\nNVerticesOwned " << D->NVerticesOwned << "\nNVerticesAll " << D->NVerticesAll << "\nEdgeCut "
        << D->EdgeCut << "\nOwnedIDSum " << Sum << "\nNNeighbors " << D->NeighborTasks.size() << "\n";
#define DUMP(A) writeBin(#A, D->A.data(), D->A.V.size())
   DUMP(NCellsHaloH); DUMP(NEdgesHaloH); DUMP(NVerticesHaloH);
   DUMP(CellIDH); DUMP(EdgeIDH); DUMP(VertexIDH); DUMP(CellLocH); DUMP(EdgeLocH); DUMP(VertexLocH);
   DUMP(CellsOnCellH); DUMP(EdgesOnCellH); DUMP(VerticesOnCellH); DUMP(CellsOnEdgeH); DUMP(VerticesOnEdgeH); DUMP(EdgesOnEdgeH);
   DUMP(CellsOnVertexH); DUMP(EdgesOnVertexH); DUMP(NEdgesOnCellH); DUMP(NEdgesOnEdgeH);
   writeBin("CellTask", D->CellTask.data(), D->CellTask.size());
   writeBin("NeighborTasks", D->NeighborTasks.data(), D->NeighborTasks.size());
   const char *Kind[3] = {"Cell", "Edge", "Vertex"};
   for (int k = 0; k < 3; ++k) {
      const int NN = (int)D->NeighborTasks.size();
      writeBin(std::string("SendOffsets") + Kind[k], D->SendOffsets[k].data(), D->SendOffsets[k].size());
      writeBin(std::string("RecvOffsets") + Kind[k], D->RecvOffsets[k].data(), D->RecvOffsets[k].size());
      writeBin(std::string("SendIdx") + Kind[k], D->SendIdx[k].data(), (size_t)D->SendOffsets[k][NN]);
      writeBin(std::string("RecvIdx") + Kind[k], D->RecvIdx[k].data(), (size_t)D->RecvOffsets[k][NN]);
   }
   Decomp::clear();
   REQUIRE(Decomp::getDefault() == nullptr && Decomp::get("Default") == nullptr);
   std::printf("decomp mirror ok: task %d of %d, %d owned cells, %d neighbours\n", MyTask, NTasks, MD.NCellsOwned, HD.NNeighbors);
   return 0;
}
