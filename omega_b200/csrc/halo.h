// halo.h -- the object behind OmegaHalo* and the internal entry points step.cu uses.
#pragma once
#include "omega_dev.h"

constexpr int kHaloMaxArrays    = 3;  // h, u, tracers of one exchange point
constexpr int kHaloMaxNeighbors = 32; // per-neighbour buffer offsets travel as kernel parameters

// launch arguments of k_halo_round: array a of neighbour n starts at Start[a][n] (doubles) of the round buffer
struct HaloRoundArgs {
   int NArr, K;
   omega::Real *Arr[kHaloMaxArrays];
   int NOuter[kHaloMaxArrays], NRows[kHaloMaxArrays];
   const int *Off[kHaloMaxArrays], *Idx[kHaloMaxArrays]; // exchange lists of the array's element kind (device order)
   long long Start[kHaloMaxArrays][kHaloMaxNeighbors];
};

struct HaloTimed {
   cudaEvent_t Start, Stop;
   long long Bytes;
};

struct OmegaHalo {
   const OmegaMesh *M = nullptr;
   int MyTask = 0, NNeighbors = 0;
   std::vector<int> NeighborTasks;
   std::vector<int> SendOff[3], RecvOff[3]; // [NNeighbors+1]
   int *DSendIdx[3] = {}, *DRecvIdx[3] = {}; // device: [offsets | indices in device order]
   // round buffers, library-owned (grown on demand)
   omega::Real *SendBuf = nullptr, *RecvBuf = nullptr;
   size_t SendCap = 0, RecvCap = 0;
   long long LastSendDoubles = 0;
   // transport
   void *Comm = nullptr; // ncclComm_t
   int OwnsComm = 0, NRanks = 1;
   // overlap with interior compute (step.cu)
   cudaStream_t CommStream = nullptr;
   cudaEvent_t EvReady = nullptr, EvDone = nullptr;
   int *DTileOrder = nullptr;   // [boundary tiles | other owned tiles | halo tiles]
   int *DTileOrderBH = nullptr; // [boundary tiles | halo tiles | other owned tiles]
   int NTilesBoundary = 0, NTilesOwned = 0, HaloCovered = 0, NumSMs = 148;
   // CUDA-event timing of the rounds (bench: ms per round, bytes against NVLink)
   int Profiling = 0;
   long long Rounds = 0;
   std::vector<HaloTimed> Events;
   std::vector<void *> Allocs;
};

namespace omega {
int halo_pack_round(OmegaHalo *H, omega::Real *const *Ptrs, const int *Elems, const int *NOuter, int N, cudaStream_t St);
int halo_unpack_round(OmegaHalo *H, omega::Real *const *Ptrs, const int *Elems, const int *NOuter, int N, cudaStream_t St);
int halo_exchange_round(OmegaHalo *H, omega::Real *const *Ptrs, const int *Elems, const int *NOuter, int N, cudaStream_t St);
} // namespace omega
