// global_mesh_test.cpp -- the whole host side in C++, from the GLOBAL mesh to stepped state, the way an Omega driver
// runs (OceanInit.cpp:50-120: Decomp::init -> Halo::init -> HorzMesh::init -> OceanState / Tracers / AuxiliaryState /
// Tendencies / TimeStepper init, then doStep): every rank reads the same global arrays, numbers its own partition
// with the Decomp mirror (OmegaB200Decomp.h), gathers its local geometry and initial state by global ID, builds the
// Halo from the Decomp's lists and steps.  Usage:
//   global_mesh_test <dir> <rank> <world> [dry]
// <dir>: global connectivity and geometry (tests/test_decomp_mirror.py write_global_mesh / write_global_case), state in
// GLOBAL order, cfg.bin, meta.txt.  Output: <dir>/r<rank>_{h,u,tr,CellID,EdgeID}.bin (local order, every local row).
// "dry": stop before anything touches a device and write the local geometry and the gathered state instead.
#include <chrono>
#include <cstdio>
#include <cuda_runtime_api.h>
#include <fstream>
#include <map>
#include <memory>
#include <thread>

#include "OmegaB200.h"
#include "OmegaB200Decomp.h"
#include "OmegaB200Stepper.h"

using namespace OMEGA;

static std::string Dir;
template <class T> static std::vector<T> load(const std::string &Name) {
   std::ifstream F(Dir + "/" + Name + ".bin", std::ios::binary | std::ios::ate);
   if (!F) { std::fprintf(stderr, "missing %s\n", Name.c_str()); std::exit(2); }
   const std::streamsize N = F.tellg();
   F.seekg(0);
   std::vector<T> V((size_t)N / sizeof(T));
   F.read(reinterpret_cast<char *>(V.data()), N);
   return V;
}
template <class T> static void save(const std::string &Name, const T *P, size_t N) {
   std::ofstream F(Dir + "/" + Name + ".bin", std::ios::binary);
   F.write(reinterpret_cast<const char *>(P), (std::streamsize)(N * sizeof(T)));
}

int main(int argc, char **argv) {
   if (argc < 4) return 2;
   Dir = argv[1];
   const int Rank = std::atoi(argv[2]), World = std::atoi(argv[3]);
   const bool Dry = argc >= 5 && std::string(argv[4]) == "dry";
   const std::string Tag = "r" + std::to_string(Rank) + "_";
   std::map<std::string, double> Meta;
   {
      std::ifstream F(Dir + "/meta.txt");
      std::string K; double V;
      while (F >> K >> V) Meta[K] = V;
   }
   const int K = (int)Meta["NVertLayers"], T = (int)Meta["NTracers"], NSteps = (int)Meta["NSteps"], Width = (int)Meta["HaloWidth"];
   const double Dt = Meta["Dt"];

   // ---- the global mesh, as every rank reads it ----
   auto Dims = load<int>("dims");
   auto CoC = load<int>("cellsOnCell"), EoC = load<int>("edgesOnCell"), VoC = load<int>("verticesOnCell");
   auto CoE = load<int>("cellsOnEdge"), VoE = load<int>("verticesOnEdge"), EoE = load<int>("edgesOnEdge");
   auto CoV = load<int>("cellsOnVertex"), EoV = load<int>("edgesOnVertex");
   auto X = load<double>("xCell"), Y = load<double>("yCell"), Z = load<double>("zCell");
   OmegaGlobalMesh G{};
   G.NCells = Dims[0]; G.NEdges = Dims[1]; G.NVertices = Dims[2]; G.MaxEdges = Dims[3]; G.MaxEdges2 = Dims[4]; G.VertexDegree = Dims[5];
   G.CellsOnCell = CoC.data(); G.EdgesOnCell = EoC.data(); G.VerticesOnCell = VoC.data();
   G.CellsOnEdge = CoE.data(); G.VerticesOnEdge = VoE.data(); G.EdgesOnEdge = EoE.data();
   G.CellsOnVertex = CoV.data(); G.EdgesOnVertex = EoV.data();
   G.XCell = X.data(); G.YCell = Y.data(); G.ZCell = Z.data();
   std::map<std::string, std::vector<double>> GA;
   GlobalGeometry GG{};
#define GEO(Member, File) GA[File] = load<double>(File); GG.Member = GA[File].data()
   GEO(AreaCell, "areaCell"); GEO(BottomDepth, "bottomDepth"); GG.XCell = X.data(); GG.YCell = Y.data(); GG.ZCell = Z.data();
   GEO(DcEdge, "dcEdge"); GEO(DvEdge, "dvEdge"); GEO(AngleEdge, "angleEdge"); GEO(FEdge, "fEdge"); GEO(XEdge, "xEdge"); GEO(YEdge, "yEdge");
   GEO(WeightsOnEdge, "weightsOnEdge"); GEO(AreaTriangle, "areaTriangle"); GEO(FVertex, "fVertex");
   GEO(KiteAreasOnVertex, "kiteAreasOnVertex");
#undef GEO

   // ---- Decomp -> local mesh ----
   Decomp::init(G, World, Rank, Width, PartMethodRCB);
   const Decomp *D = Decomp::getDefault();
   LocalGeometry Geo(*D, GG);
   OmegaMeshDesc MD{};
   D->fillMeshDesc(MD, K);
   Geo.fillMeshDesc(MD);

   // ---- the initial state: global order on file, gathered by ID ----
   auto HG = load<double>("h"), UG = load<double>("u"), TG = load<double>("tr");
   std::vector<double> H((size_t)D->NCellsSize * K), U((size_t)D->NEdgesSize * K), TR((size_t)T * D->NCellsSize * K);
   gatherField(*D, OnCell, 1, K, HG.data(), H.data());
   gatherField(*D, OnEdge, 1, K, UG.data(), U.data());
   if (T) gatherField(*D, OnCell, T, K, TG.data(), TR.data());
   save(Tag + "CellID", D->CellIDH.data(), (size_t)D->NCellsAll);
   save(Tag + "EdgeID", D->EdgeIDH.data(), (size_t)D->NEdgesAll);
   if (Dry) {
      save(Tag + "h0", H.data(), H.size()); save(Tag + "u0", U.data(), U.size()); save(Tag + "tr0", TR.data(), TR.size());
#define OUT(A) save(Tag + #A, Geo.A.data(), Geo.A.size())
      OUT(AreaCell); OUT(BottomDepth); OUT(XCell); OUT(YCell); OUT(ZCell); OUT(DcEdge); OUT(DvEdge); OUT(AngleEdge); OUT(FEdge);
      OUT(XEdge); OUT(YEdge); OUT(WeightsOnEdge); OUT(AreaTriangle); OUT(FVertex); OUT(KiteAreasOnVertex);
#undef OUT
      std::printf("global mesh test (dry): task %d of %d, %d owned of %d local cells\n", Rank, World, D->NCellsOwned, D->NCellsAll);
      return 0;
   }

   // ---- device side ----
   if (World > 1 && cudaSetDevice(Rank) != cudaSuccess) { std::fprintf(stderr, "rank %d: no device\n", Rank); return 3; }
   HorzMesh Mesh(MD, true);
   std::unique_ptr<Halo> MeshHalo;
   if (World > 1) {
      MeshHalo.reset(new Halo(&Mesh, D->haloDesc()));
      char Id[128];
      const std::string IdFile = Dir + "/nccl_id.bin"; // stands in for the MPI_Bcast of an Omega host
      if (Rank == 0) {
         Halo::uniqueId(Id);
         std::ofstream F(IdFile + ".tmp", std::ios::binary);
         F.write(Id, 128);
         F.close();
         std::rename((IdFile + ".tmp").c_str(), IdFile.c_str());
      } else {
         for (int i = 0; i < 600; ++i) {
            std::ifstream F(IdFile, std::ios::binary);
            if (F && F.read(Id, 128)) break;
            std::this_thread::sleep_for(std::chrono::milliseconds(100));
            if (i == 599) { std::fprintf(stderr, "rank %d: no NCCL id\n", Rank); return 4; }
         }
      }
      MeshHalo->initComm(Id, World, Rank);
   }
   auto CfgRaw = load<char>("cfg");
   OmegaTendConfig Cfg = *reinterpret_cast<OmegaTendConfig *>(CfgRaw.data());
   OceanState State("Default", &Mesh, MeshHalo.get(), K, 2);
   Tracers Trc(&Mesh, MeshHalo.get(), T, K, 2);
   AuxiliaryState Aux("Default", &Mesh, nullptr, K, T, Cfg);
   Tendencies Tend("Default", &Mesh, &Aux, K);
   TimeStepper Stepper(TimeStepperType::RungeKutta4, Dt, &Tend, &Aux, &Trc);
   Stepper.Fused = true;
   State.copyToDevice(0, H.data(), U.data());
   Array3DReal Cur;
   Trc.getAll(Cur, 0);
   if (T) Mesh.copyToDevice(OnCell, T, K, TR.data(), Cur.data());
   double SimTime = 0.0;
   for (int s = 0; s < NSteps; ++s) Stepper.doStep(&State, SimTime);
   State.copyToHost(0, H.data(), U.data());
   Trc.getAll(Cur, 0);
   if (T) Mesh.copyToHost(OnCell, T, K, Cur.data(), TR.data());
   save(Tag + "h", H.data(), H.size()); save(Tag + "u", U.data(), U.size()); save(Tag + "tr", TR.data(), TR.size());
   std::printf("global mesh test: task %d of %d, %d steps, SimTime %.1f\n", Rank, World, NSteps, SimTime);
   return 0;
}
