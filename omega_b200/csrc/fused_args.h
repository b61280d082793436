// fused_args.h -- argument blocks of the fused per-stage kernels (fused.cu, fused_tile.cu).
#pragma once
#include "omega_dev.h"

namespace omega {

struct CellArgs {
   const Real *Hs, *Us, *UTr, *Trs, *Del2Tr; // stage state; UTr = velocity seen by tracer advection
   const Real *Hcur;
   Real *Hnext, *Hprov;                      // Hprov == nullptr on the last stage
   const Real *TrCur;
   Real *TrNext, *TrProv;
   Real A, B;                                // a_{s+1} dt, b_s dt
   int First, Last, NTracers;
   int WSel;   // stage-0 weight thickness: 0 = Hcur, 1 = Hnext before this stage writes it
   int H2Sel;  // thickness multiplying cur tracers in the provisional update: 0 = Hcur, 1 = new Hnext, 2 = new Hprov
   int DivSel; // thickness dividing the provisional tracers: 0 = new Hprov, 1 = Hcur (RK2 with raw time-level indexing)
   int FinSel; // thickness dividing the final tracers: 0 = new Hnext, 1 = Hcur
   int FluxDivOn, AdvOn, DiffOn, HypOn, FluxUpwind, TrUpwind;
   int Prefetch; // 0 off, 1 next tracer's rows into L1, 2 into L2
   Real EddyDiff2, EddyDiff4;
   ManufacturedParams MS;
   const Real *SrcH; // tiled kernels: manufactured thickness source per cell (k_ms_sources), nullptr when off
};

constexpr int kME = 6; // max edges per cell on the fused path (checked at launch)

struct EdgeArgs {
   const Real *Hs, *Us, *URaw;                       // stage state; URaw = velocity seen by bottom drag
   const Real *KE, *Div, *RelVort, *PVv, *Del2Div, *Del2RelVort, *NormalStress;
   const Real *Ucur;
   Real *Unext, *Uprov;
   Real A, B;
   int First, Last;
   OmegaTendConfig G;
   ManufacturedParams MS;
};

// tiled edge pass: tend = m*PV term + m*[(-(X1-X0) + SY*(Y1-Y0))/dc - SZ*(Z1-Z0)/dv]
struct EdgeTileArgs {
   const Real *Us, *Hs, *PVv; // stage velocity / thickness, PV on vertices
   const Real *X, *Y, *Z;     // combined cell / cell / vertex fields written by fA and fB
   Real SY, SZ;
   const Real *Ucur;
   Real *Unext, *Uprov;
   Real A, B;
   int First, Last, PVOn, XOn;
   int FluxUpwind;     // FluxLayerThickEdge upwinded on the sign of u (LayerThicknessAuxVars.h:25-61)
   const Real *SrcU; // manufactured velocity source per edge (k_ms_sources), nullptr when off
};

// tiles Map[First .. First + Count) of the mesh (Map == nullptr: tiles First .. First + Count)
struct TileRange {
   const int *Map;
   int First, Count;
   __host__ __device__ int tile(int i) const { return Map ? Map[First + i] : First + i; }
};
inline TileRange all_tiles(const DMesh &D) { return TileRange{nullptr, 0, D.Tiles.NTiles}; }

// fused_tile.cu
bool tile_path_ok(const DMesh &D);
bool tile_partial_chunk(const DMesh &D); // K is not a multiple of 16: only k_fCT2 / k_fET take the cell / edge pass
bool tile_v2_enabled(); // k_fCT2 (any tracer count with tracer hyperdiffusion) is selectable
bool tile_v2_used(const CellArgs &ca); // launch_fCT will run k_fCT2 for these arguments
bool tile_edge_path_ok(const DMesh &D);
// Overlap: the launch does not depend on the kernel queued right before it on St (the tiled cell pass), so it
// is launched with programmatic stream serialization and its CTAs fill the SMs that kernel's tail leaves idle
int launch_fET(const DMesh &D, const EdgeTileArgs &ea, cudaStream_t St, const TileRange &TR, bool Overlap = false);
int launch_fCT(const DMesh &D, const CellArgs &ca, cudaStream_t St, const TileRange &TR);

} // namespace omega
