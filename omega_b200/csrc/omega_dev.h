// omega_dev.h -- internal declarations of libomega_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>

#include <cstdio>
#include <string>
#include <vector>

#include "omega_b200.h"
#include "real.h"

namespace omega {

void set_error(const std::string &msg);

#define OMEGA_CUDA_CHECK(expr)                                                         \
   do {                                                                                \
      cudaError_t e_ = (expr);                                                         \
      if (e_ != cudaSuccess) {                                                         \
         ::omega::set_error(std::string(#expr) + ": " + cudaGetErrorString(e_));       \
         return 1;                                                                     \
      }                                                                                \
   } while (0)

#define OMEGA_REQUIRE(cond, msg)                                                       \
   do {                                                                                \
      if (!(cond)) {                                                                   \
         ::omega::set_error(msg);                                                      \
         return 2;                                                                     \
      }                                                                                \
   } while (0)

// Cell patches ("tiles") of the shared-memory staged kernels: kTileCells
// consecutive cells of the device order (one leaf of the bisection order) plus
// the two rings of neighbours their stencils reach.  Local row numbering of a
// tile: own cells [0,nc), ring 1 [nc,nc+n1), ring 2 [nc+n1,nc+n1+n2); the
// boundary slot (row NCellsAll, all zero) appears as an ordinary ring row.
constexpr int kTileCells = 64;  // own cells per tile (fewer when a cap below would be exceeded)
constexpr int kTileRing1 = 56;  // cap on ring-1 rows
constexpr int kTileRows1 = kTileCells + kTileRing1; // cap on own + ring-1 rows
constexpr int kTileRows  = 176; // cap on own + ring-1 + ring-2 rows
constexpr int kTileEdgesOwn = 256; // cap on own edges of a tile (edges whose lowest-ranked cell is an own cell)
constexpr int kTileEdges    = 400; // cap on staged edges: the own edges first, then the EdgesOnEdge entries they reference
struct TileHdr {
   int C0, NC, N1, N2; // first own cell, counts
   int RowOff;         // offset of the tile's ring rows in HaloRows
   int NblOff;         // offset (in rows) of the tile's rows in Nbl
   int NEOwn, NEStaged; // edge counts
   int ERowOff;        // offset of the tile's staged edges in EdgeRows
   int EoEOff;         // offset (in own edges) of the tile's rows in EoELocal
   int Pad0, Pad1;
};
struct DTiles {
   int NTiles;
   const TileHdr *Hdr;
   const int *HaloRows;       // ring-1 then ring-2 rows of every tile (device cell index)
   const unsigned short *Nbl; // [own + ring-1 rows][8]: local row of the cell across edge j (self when none)
   const Real *GCell;       // [NCellsSize][6]  EdgeMask*DvEdge/DcEdge of edge j of the cell, 0 when none
   const Real *CfUCell;     // [NCellsSize][6]  DvEdge*EdgeSignOnCell
   const Real *WaCell;      // [NCellsSize][6]  EdgeMask*DvEdge*EdgeSignOnCell
   const int *EocCell;        // [NCellsSize][6]  EdgesOnCell, boundary slot when none
   // edge side (the shared-memory staged edge pass): valid when EdgeTilesOK
   int EdgeTilesOK;
   const int *EdgeRows;            // staged edges of every tile (device edge index), own edges first
   const int4 *EdgeCV;             // per staged edge: CellsOnEdge (x,y), VerticesOnEdge (z,w), same order as EdgeRows
   const unsigned short *EoELocal; // [own edges][10]: staged row of EdgesOnEdge entry j (self when none)
   const Real *InvAreaCell; // [NCellsSize], 0 in the boundary slot
};

// Device-resident HorzMesh, structure of arrays, DEVICE order.  Passed to
// kernels by value (fits the 4 KB parameter space many times over).
struct DMesh {
   int NCellsAll, NEdgesAll, NVerticesAll;
   int NCellsOwned, NEdgesOwned, NVerticesOwned;
   int NCellsSize, NEdgesSize, NVerticesSize;
   int K;        // NVertLayers
   int MaxEdges; // row stride of the per-cell tables
   int MaxEdges2;
   // cells
   const int *NEdgesOnCell;       // [C]
   const int *EdgesOnCell;        // [C][MaxEdges]
   const Real *EdgeSignOnCell;  // [C][MaxEdges]
   const Real *AreaCell, *BottomDepth, *XCell, *YCell;
   // edges
   const int2 *CellsOnEdge, *VerticesOnEdge; // [E]
   const int *NEdgesOnEdge;                  // [E] trimmed of trailing boundary-slot entries
   const int *EdgesOnEdge;                   // [E][MaxEdges2]
   const Real *WeightsOnEdge;              // [E][MaxEdges2]
   const int2 *EoECells, *EoEVerts;          // [E][MaxEdges2] cells / vertices of each EdgesOnEdge entry
   const Real *DcEdge, *DvEdge, *AngleEdge, *FEdge, *XEdge, *YEdge;
   const Real *EdgeMask, *MeshScalingDel2, *MeshScalingDel4; // [E]
   // vertices (VertexDegree == 3)
   const int *CellsOnVertex, *EdgesOnVertex;               // [V][3]
   const Real *EdgeSignOnVertex, *KiteAreasOnVertex;     // [V][3]
   const Real *AreaTriangle, *FVertex;
   // reciprocals evaluated once on the host (IEEE division, the same bits the kernels' own
   // 1.0/x gave): 1/dc, 1/dv, 1/max(dv, 0.25 dc) per edge; 1/AreaCell; 1/AreaTriangle
   const Real *InvDcEdge, *InvDvEdge, *InvDvEffEdge, *InvAreaCell, *InvAreaTriangle;
   // flattened stencils of k_fB (velocity Del2 of Del2): per (cell, edge j) the rows
   // {neighbour cell, v0, v1} and the coefficients {A, B} of
   //   Del2Div(c) = sum_j A_j (Div(nbr_j) - Div(c)) + B_j (RelVort(v1_j) - RelVort(v0_j)),
   // per (vertex, edge j) the rows {c0, c1, other vertex} and {A', B'} of
   //   Del2RelVort(v) = sum_j A'_j (Div(c1_j) - Div(c0_j)) + B'_j (RelVort(vo_j) - RelVort(v))
   const int4 *B2CellIdx, *B2VertIdx;
   const int4 *B3CellIdx, *B3VertIdx; // k_fB3: every gathered row once (cell: 6 cells + 6 vertices; vertex: 4 cells + 3 vertices)
   int B3OK;
   // flattened stencils of k_fA: per (cell, edge j) the edge row (B2... not needed: AEdge) and {KE, Div}
   // coefficients; per (vertex, j) the {cell, edge} rows and {kite/At, dc*sign/At}
   const int *ACellEdge;     // [C][6] edge row, boundary slot when none
   const Real2 *ACellCf;   // [C][6] {0.25 dv dc / A, dv sign / A}
   const int2 *AVertIdx;     // [V][3] {cell, edge}
   const Real2 *AVertCf;   // [V][3] {kite / At, dc sign / At}
   const Real2 *B2CellCf, *B2VertCf;
   const Real2 *B3CellCf, *B3VertCf;
   DTiles Tiles; // NTiles == 0 when the mesh has cells with more than 6 edges
};

} // namespace omega

struct OmegaMesh {
   omega::DMesh D;
   OmegaMeshDesc Sizes; // scalar members only (pointers nulled)
   int Reordered;
   int MaxEdgesUsed; // max NEdgesOnCell over local cells
   // host copies of the permutations (Perm[new] = old) and derived fields (host order)
   std::vector<int> Perm[3], IPerm[3];
   std::vector<double> EdgeSignOnCellH, EdgeSignOnVertexH, EdgeMaskH;
   // device permutation tables for field transfers
   int *DPerm[3], *DIPerm[3];
   std::vector<void *> Allocs; // every device allocation owned by the mesh
   // host view of the tiles (device order): first own cell of every tile, tile of a cell, tile whose OWN edge
   // list holds an edge (-1: none); omega_halo_create classifies tiles with these
   std::vector<int> TileC0, TileOfCell, TileOfEdge;
};

namespace omega {

inline int n_size(const OmegaMesh *M, int Elem) {
   return Elem == OMEGA_ON_CELL ? M->D.NCellsSize : Elem == OMEGA_ON_EDGE ? M->D.NEdgesSize : M->D.NVerticesSize;
}
inline int n_all(const OmegaMesh *M, int Elem) {
   return Elem == OMEGA_ON_CELL ? M->D.NCellsAll : Elem == OMEGA_ON_EDGE ? M->D.NEdgesAll : M->D.NVerticesAll;
}

// ---- kernels_aux.cu / kernels_tend.cu / kernels_update.cu launchers --------
struct AuxPtrs {
   Real *KE, *Div, *Flux, *Mean, *Ssh, *RelVort, *NormRelVort, *NormPlanetVort, *NormRelVortEdge,
       *NormPlanetVortEdge, *Del2Edge, *Del2Div, *Del2RelVort, *HTr, *Del2Tr, *NormalStress,
       *ZonalStress, *MeridStress;
   Real *ZonalStress2; // [NCellsSize] scratch of the fused path: kinetic energy of the bottom layer (bottom drag)
};

struct Launch {
   cudaStream_t Stream;
   long long *Count; // launch counter of the owning stepper
};

void launch_vertex_aux1(const Launch &L, const DMesh &M, const Real *H, const Real *U, const AuxPtrs &A);
void launch_cell_aux1(const Launch &L, const DMesh &M, const Real *H, const Real *U, const AuxPtrs &A);
void launch_wind_aux(const Launch &L, const DMesh &M, const AuxPtrs &A, int Aniso);
void launch_edge_aux2(const Launch &L, const DMesh &M, const Real *H, const Real *U, const AuxPtrs &A,
                      int FluxUpwind, int ThickOnly);
void launch_vertex_aux2(const Launch &L, const DMesh &M, const AuxPtrs &A);
void launch_cell_aux2(const Launch &L, const DMesh &M, const AuxPtrs &A);
void launch_tracer_aux_edge(const Launch &L, const DMesh &M, const Real *H, const Real *U,
                            const Real *Tr, int NTracers, const AuxPtrs &A, int Upwind);
void launch_tracer_aux_cell(const Launch &L, const DMesh &M, const Real *Tr, int NTracers,
                            const AuxPtrs &A);

struct ManufacturedParams {
   int Enabled;
   Real H0, Eta0, Kx, Ky, AngFreq, Time;
};

void launch_thick_tend(const Launch &L, const DMesh &M, const OmegaTendConfig &G, const AuxPtrs &A,
                       const Real *U, Real *HTend, const ManufacturedParams &MS);
void launch_vel_tend(const Launch &L, const DMesh &M, const OmegaTendConfig &G, const AuxPtrs &A,
                     const Real *U, const Real *URaw, Real *UTend, const ManufacturedParams &MS);
void launch_tracer_tend(const Launch &L, const DMesh &M, const OmegaTendConfig &G, const AuxPtrs &A,
                        const Real *URaw, const Real *Tr, int NTracers, Real *TrTend);

// Y1 = Y2 + c*T over rows [0,NAll)
void launch_axpy(const Launch &L, int NAll, int K, Real *Y1, const Real *Y2, Real C, const Real *T);
// Next = (Cur*H2 + c*T)/H1 ; Next = Cur*H ; Accum += c*T ; Next /= H   (all tracers)
void launch_update_tracers(const Launch &L, int NAll, int NSize, int K, int NTracers, Real *Next,
                           const Real *Cur, const Real *H1, const Real *H2, Real C, const Real *T);
void launch_weight_tracers(const Launch &L, int NAll, int NSize, int K, int NTracers, Real *Next,
                           const Real *Cur, const Real *H);
void launch_accumulate_tracers(const Launch &L, int NAll, int NSize, int K, int NTracers, Real *Accum,
                               Real C, const Real *T);
void launch_finalize_tracers(const Launch &L, int NAll, int NSize, int K, int NTracers, Real *Next,
                             const Real *H);

// one slab of host rows of a field between a contiguous staging block and the device-order array
void launch_slab_rows(cudaStream_t S, bool ToDevice, int Row0, int NSlab, int NInner, const int *DevRowOfHost, Real *Dev,
                      Real *Stage);
// the whole field between pinned host memory and the device-order array, no staging
void launch_rows_hostmem(cudaStream_t S, bool ToDevice, int NOuter, int NRows, int NInner, const int *DevRowOfHost, Real *Dev,
                         Real *Host, int NCtas);
// gather/scatter rows through a permutation (field transfer, halo)
void launch_permute_rows(cudaStream_t S, int NOuter, int NRows, int NInner, const int *SrcRowOfDst,
                         const Real *Src, Real *Dst);

} // namespace omega
