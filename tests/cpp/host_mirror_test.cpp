// host_mirror_test.cpp -- drives the C++ host mirror (omega_b200/host/OmegaB200.h)
// the way the reference's TimeStepperTest / TendenciesTest drive Omega: build
// HorzMesh, OceanState, Tracers, AuxiliaryState, Tendencies, TimeStepper, call
// computeAllTendencies and doStep.  Input arrays come from a directory of raw
// .bin files written by tests/test_host_mirror.py, results go back the same way.
#include <chrono>
#include <cstdio>
#include <cublas_v2.h>
#include <cuda_runtime_api.h>
#include <fstream>
#include <thread>
#include <iostream>
#include <map>
#include <sstream>

#include <memory>

#include "OmegaB200.h"
#include "OmegaB200Vertical.h"

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
   AuxiliaryState Aux("Default", &Mesh, MeshHalo.get(), K, T, Cfg);
   // "Wind 1": the forcing arrays as a host reads them (owned cells and, here, the halo too), then the exchange the
   // reference's init does for them (OceanInit.cpp:188, AuxiliaryState.cpp:313-323)
   if (Meta["Wind"] != 0) {
      auto Zonal = load<double>("zonal"), Merid = load<double>("merid");
      Aux.setWindStress(Zonal.data(), Merid.data());
      if (Aux.exchangeHalo() != 0) { std::fprintf(stderr, "AuxiliaryState::exchangeHalo failed\n"); return 6; }
   }
   Tendencies Tend("Default", &Mesh, &Aux, K);
   // "Custom 1": the custom velocity tendency of the reference's TimeStepperTest (du/dt -= Coeff u,
   // test/timeStepping/TimeStepperTest.cpp:49-72; Coeff = 1e-5 / s here, the steps are minutes long) installed the
   // way that test does, as a std::function; the device work is one cuBLAS axpy on the stream the library hands over
   cublasHandle_t Blas = nullptr;
   if (Meta["Custom"] != 0) {
      if (cublasCreate(&Blas) != CUBLAS_STATUS_SUCCESS) { std::fprintf(stderr, "cublasCreate failed\n"); return 5; }
      Tend.setCustomTendencies(nullptr, [&](Array2DReal VelTend, const OceanState *Stage, const AuxiliaryState *, int, int VelLevel, double) {
         Array2DReal U;
         Stage->getNormalVelocity(U, VelLevel);
         const double Alpha = -1.0e-5;
         cublasSetStream(Blas, static_cast<cudaStream_t>(Tend.customStream()));
         if (cublasDaxpy(Blas, (int)VelTend.size(), &Alpha, U.data(), 1, VelTend.data(), 1) != CUBLAS_STATUS_SUCCESS) std::abort();
      });
   }
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
   // the Omega-1 column pieces through their reference-named mirror (OmegaB200Vertical.h), on the stepped thickness:
   // pressure, TEOS-10 specific volume (plain and displaced by one layer), z height, geopotential, target
   // thickness, and the two tridiagonal solvers on systems derived from the thickness
   if (World == 1 && K >= 2) {
      const int NC = D.NCellsSize;
      std::vector<int> Lo(NC, 0), Hi(NC, K - 1);
      for (int i = 0; i < NC; ++i) { Lo[i] = i % 3 == 0 && K > 3 ? 1 : 0; Hi[i] = i % 5 == 0 && K > 3 ? K - 2 : K - 1; }
      VertCoord VC(&Mesh, K, Lo.data(), Hi.data());
      Eos Es(&Mesh, K);
      Array2DReal HNow, Ct("ConservTemp", {NC, K}), Sa("AbsSalinity", {NC, K});
      Array1DReal Surf("SurfacePressure", {NC}), Tid("TidalPotential", {NC}), Sal("SelfAttractionLoading", {NC});
      State.getLayerThickness(HNow, 0);
      std::vector<double> CtH((size_t)NC * K), SaH((size_t)NC * K), SurfH(NC), TidH(NC), SalH(NC), Wt(K), RefH((size_t)NC * K);
      for (int i = 0; i < NC; ++i) {
         SurfH[i] = 100.0 + i % 7; TidH[i] = 0.01 * (i % 11); SalH[i] = 0.002 * (i % 13);
         for (int k = 0; k < K; ++k) {
            CtH[(size_t)i * K + k] = 2.0 + 0.01 * (i % 100) + 0.5 * k;
            SaH[(size_t)i * K + k] = 33.0 + 0.02 * (i % 50) + 0.1 * k;
            RefH[(size_t)i * K + k] = H[(size_t)i * K + k] * 0.9;
         }
      }
      for (int k = 0; k < K; ++k) Wt[k] = 1.0 + 0.1 * k;
      Mesh.copyToDevice(OnCell, 1, K, CtH.data(), Ct.data());
      Mesh.copyToDevice(OnCell, 1, K, SaH.data(), Sa.data());
      Mesh.copyToDevice(OnCell, 1, 1, SurfH.data(), Surf.data());
      Mesh.copyToDevice(OnCell, 1, 1, TidH.data(), Tid.data());
      Mesh.copyToDevice(OnCell, 1, 1, SalH.data(), Sal.data());
      Mesh.copyToDevice(OnCell, 1, K, RefH.data(), VC.RefLayerThickness.data());
      OMEGA_B200_CHECK(omega_dev_upload(VC.VertCoordMovementWeights.data(), Wt.data(), K * sizeof(double)), "weights");
      VC.computePressure(HNow, Surf);
      std::vector<double> O1((size_t)NC * K), O2((size_t)NC * (K + 1));
      // TEOS-10 takes dbar (Eos.h:36-48): Pa -> dbar on the host, as the reference's caller does
      Array2DReal Pd("PressureDbar", {NC, K});
      Mesh.copyToHost(OnCell, 1, K, VC.PressureMid.data(), O1.data());
      for (double &v : O1) v *= 1.0e-4;
      Mesh.copyToDevice(OnCell, 1, K, O1.data(), Pd.data());
      Es.computeSpecVol(Ct, Sa, Pd);
      Es.computeSpecVolDisp(Ct, Sa, Pd, 1);
      VC.computeZHeight(HNow, Es.SpecVol);
      VC.computeGeopotential(Tid, Sal);
      VC.computeTargetThickness();
      auto out2 = [&](const char *Nm, const Array2DReal &A, int W) {
         std::vector<double> &O = W == K ? O1 : O2;
         Mesh.copyToHost(OnCell, 1, W, A.data(), O.data());
         save(Nm, O);
      };
      out2("out_pint", VC.PressureInterface, K + 1); out2("out_pmid", VC.PressureMid, K);
      out2("out_specvol", Es.SpecVol, K); out2("out_specvol_disp", Es.SpecVolDisplaced, K);
      out2("out_zint", VC.ZInterface, K + 1); out2("out_zmid", VC.ZMid, K);
      out2("out_geop", VC.GeopotentialMid, K); out2("out_target", VC.LayerThicknessTarget, K);
      // tridiagonal systems, one per cell, rows = levels: diffusion form with G from the thickness, and the
      // general form of the same matrix
      const int NB = D.NCellsAll;
      Array2DReal G("G", {NB, K}), Hd("H", {NB, K}), X("X", {NB, K}), DL("DL", {NB, K}), Dd("D", {NB, K}), DU("DU", {NB, K}), X2("X2", {NB, K});
      std::vector<double> g((size_t)NB * K), hd((size_t)NB * K), x((size_t)NB * K), dl((size_t)NB * K), dd((size_t)NB * K), du((size_t)NB * K);
      for (int i = 0; i < NB; ++i)
         for (int k = 0; k < K; ++k) {
            const size_t o = (size_t)i * K + k;
            g[o] = k + 1 < K ? 2.0 / (H[o] + H[o + 1]) * 50.0 : 0.0;
            hd[o] = H[o] / 100.0;
            x[o] = 1.0 + 0.01 * k + 0.001 * (i % 17);
         }
      for (int i = 0; i < NB; ++i)
         for (int k = 0; k < K; ++k) {
            const size_t o = (size_t)i * K + k;
            const double gm = k > 0 ? g[o - 1] : 0.0, gp = k + 1 < K ? g[o] : 0.0;
            dl[o] = -gm; du[o] = -gp; dd[o] = hd[o] + gm + gp;
         }
      const unsigned long long B = (unsigned long long)NB * K * sizeof(double);
      omega_dev_upload(G.data(), g.data(), B); omega_dev_upload(Hd.data(), hd.data(), B); omega_dev_upload(X.data(), x.data(), B);
      omega_dev_upload(DL.data(), dl.data(), B); omega_dev_upload(Dd.data(), dd.data(), B); omega_dev_upload(DU.data(), du.data(), B);
      omega_dev_upload(X2.data(), x.data(), B);
      TriDiagDiffSolver::solve(G, Hd, X);
      TriDiagSolver::solve(DL, Dd, DU, X2);
      omega_dev_sync(nullptr);
      std::vector<double> xs((size_t)NB * K), xs2((size_t)NB * K);
      omega_dev_download(xs.data(), X.data(), B); omega_dev_download(xs2.data(), X2.data(), B);
      save("out_tridiag_diff", xs); save("out_tridiag", xs2);
      save("in_tridiag_g", g); save("in_tridiag_h", hd); save("in_tridiag_x", x);
   }
   if (Blas) {
      Tend.setCustomTendencies(nullptr, nullptr);
      cublasDestroy(Blas);
   }
   std::printf("host mirror: %d steps, SimTime %.1f\n", NSteps, SimTime);
   return 0;
}
