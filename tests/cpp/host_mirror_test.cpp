// host_mirror_test.cpp -- drives the C++ host mirror (omega_b200/host/OmegaB200.h)
// the way the reference's TimeStepperTest / TendenciesTest drive Omega: build
// HorzMesh, OceanState, Tracers, AuxiliaryState, Tendencies, TimeStepper, call
// computeAllTendencies and doStep.  Input arrays come from a directory of raw
// .bin files written by tests/test_host_mirror.py, results go back the same way.
#include <chrono>
#include <cstdio>
#include <cuda_runtime_api.h>
#include <fstream>
#include <thread>
#include <iostream>
#include <map>
#include <sstream>

#include <memory>

#include "OmegaB200.h"

using namespace OMEGA;

static std::string Dir;
template <class T> std::vector<T> load(const std::string &Name) {
   std::ifstream F(Dir + "/" + Name + ".bin", std::ios::binary | std::ios::ate);
   if (!F) { std::fprintf(stderr, "missing %s\n", Name.c_str()); std::exit(2); }
   const size_t Bytes = F.tellg();
   F.seekg(0);
   std::vector<T> V(Bytes / sizeof(T));
   F.read(reinterpret_cast<char *>(V.data()), Bytes);
   return V;
}
static void save(const std::string &Name, const std::vector<double> &V) {
   std::ofstream F(Dir + "/" + Name + ".bin", std::ios::binary);
   F.write(reinterpret_cast<const char *>(V.data()), V.size() * sizeof(double));
}

// usage: host_mirror_test <dir>                      one rank
//        host_mirror_test <dir> <rank> <world>       rank `rank` of `world`: <dir> holds this rank's mesh, state
//            and halo lists; the NCCL unique id travels through the file <dir>/../nccl_id.bin (the test's
//            stand-in for the MPI_Bcast an Omega host would use), one GPU per rank.
int main(int argc, char **argv) {
   if (argc < 2) return 2;
   Dir = argv[1];
   const int Rank = argc >= 4 ? std::atoi(argv[2]) : 0, World = argc >= 4 ? std::atoi(argv[3]) : 1;
   if (World > 1 && cudaSetDevice(Rank) != cudaSuccess) { std::fprintf(stderr, "rank %d: no device\n", Rank); return 3; }
   std::map<std::string, double> Meta;
   {
      std::ifstream F(Dir + "/meta.txt");
      std::string K; double V;
      while (F >> K >> V) Meta[K] = V;
   }
   OmegaMeshDesc D{};
#define I(N) D.N = (int)Meta[#N]
   I(NCellsOwned); I(NCellsAll); I(NCellsSize); I(NEdgesOwned); I(NEdgesAll); I(NEdgesSize);
   I(NVerticesOwned); I(NVerticesAll); I(NVerticesSize); I(MaxEdges); I(MaxEdges2); I(VertexDegree); I(NVertLayers);
#undef I
   std::map<std::string, std::vector<int>> IA;
   std::map<std::string, std::vector<double>> RA;
#define IP(N) IA[#N] = load<int>(#N); D.N = IA[#N].data()
#define RP(N) RA[#N] = load<double>(#N); D.N = RA[#N].data()
   IP(NEdgesOnCell); IP(EdgesOnCell); IP(CellsOnEdge); IP(VerticesOnEdge); IP(NEdgesOnEdge); IP(EdgesOnEdge);
   IP(CellsOnVertex); IP(EdgesOnVertex);
   RP(AreaCell); RP(BottomDepth); RP(XCell); RP(YCell); RP(ZCell); RP(DcEdge); RP(DvEdge); RP(AngleEdge); RP(FEdge);
   RP(XEdge); RP(YEdge); RP(WeightsOnEdge); RP(AreaTriangle); RP(FVertex); RP(KiteAreasOnVertex);
#undef IP
#undef RP
   auto CfgRaw = load<char>("cfg");
   OmegaTendConfig Cfg = *reinterpret_cast<OmegaTendConfig *>(CfgRaw.data());
   const int T = (int)Meta["NTracers"], K = D.NVertLayers, NSteps = (int)Meta["NSteps"];
   const double Dt = Meta["Dt"];
   const int Kind = (int)Meta["Stepper"]; // 0 RK4, 1 RK2, 2 FB

   HorzMesh Mesh(D, true);
   // Halo: exchange lists of this rank (Halo.cpp:455-600) + the library's NCCL transport
   std::unique_ptr<Halo> MeshHalo;
   std::map<std::string, std::vector<int>> HL;
   std::vector<int> NbrTasks;
   if (World > 1) {
      OmegaHaloDesc HD{};
      NbrTasks = load<int>("halo_neighbors");
      HD.MyTask = Rank; HD.NNeighbors = (int)NbrTasks.size(); HD.NeighborTasks = NbrTasks.data();
      const char *Kinds[3] = {"Cell", "Edge", "Vertex"};
      for (int e = 0; e < 3; ++e) {
         const std::string k = Kinds[e];
         HL["so" + k] = load<int>("halo_send_off_" + k); HL["si" + k] = load<int>("halo_send_idx_" + k);
         HL["ro" + k] = load<int>("halo_recv_off_" + k); HL["ri" + k] = load<int>("halo_recv_idx_" + k);
         HD.SendOffsets[e] = HL["so" + k].data(); HD.SendIdx[e] = HL["si" + k].data();
         HD.RecvOffsets[e] = HL["ro" + k].data(); HD.RecvIdx[e] = HL["ri" + k].data();
      }
      MeshHalo.reset(new Halo(&Mesh, HD));
      char Id[128];
      const std::string IdFile = Dir + "/../nccl_id.bin";
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
   OceanState State("Default", &Mesh, MeshHalo.get(), K, 2);
   Tracers Trc(&Mesh, MeshHalo.get(), T, K, 2);
   AuxiliaryState Aux("Default", &Mesh, nullptr, K, T, Cfg);
   Tendencies Tend("Default", &Mesh, &Aux, K);
   TimeStepper Stepper(Kind == 0 ? TimeStepperType::RungeKutta4 : Kind == 1 ? TimeStepperType::RungeKutta2
                                                                          : TimeStepperType::ForwardBackward,
                       Dt, &Tend, &Aux, &Trc);
   Stepper.Fused = Meta["Fused"] != 0;

   auto H = load<double>("h"), U = load<double>("u"), TR = load<double>("tr");
   State.copyToDevice(0, H.data(), U.data());
   Array3DReal Cur;
   Trc.getAll(Cur, 0);
   if (T) Mesh.copyToDevice(OnCell, T, K, TR.data(), Cur.data());

   // one right-hand side through the reference-named API, results read from the public arrays
   Tend.computeAllTendencies(&State, &Aux, Cur, 0, 0, 0.0);
   std::vector<double> HT((size_t)D.NCellsSize * K), UT((size_t)D.NEdgesSize * K), TT((size_t)T * D.NCellsSize * K);
   Mesh.copyToHost(OnCell, 1, K, Tend.LayerThicknessTend.data(), HT.data());
   Mesh.copyToHost(OnEdge, 1, K, Tend.NormalVelocityTend.data(), UT.data());
   if (T) Mesh.copyToHost(OnCell, T, K, Tend.TracerTend.data(), TT.data());
   save("out_htend", HT); save("out_utend", UT); save("out_trtend", TT);

   double SimTime = 0.0;
   for (int s = 0; s < NSteps; ++s) Stepper.doStep(&State, SimTime);
   State.copyToHost(0, H.data(), U.data());
   Trc.getAll(Cur, 0);
   if (T) Mesh.copyToHost(OnCell, T, K, Cur.data(), TR.data());
   save("out_h", H); save("out_u", U); save("out_tr", TR);
   std::printf("host mirror: %d steps, SimTime %.1f\n", NSteps, SimTime);
   return 0;
}
