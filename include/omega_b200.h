/* omega_b200.h -- C ABI of the B200-native Omega hot path.
 *
 * Drop-in boundary for the per-step path of E3SM Omega (components/omega):
 * every entry point below is what the body of one reference C++ method
 * forwards to.  The reference has no FFI seam (it is plain C++ classes over
 * Kokkos Views), so the seam is "same class, same method, body calls C";
 * INTEGRATION.md shows the binding for each method.  File:line references
 * are relative to components/omega/ of the reference.
 *
 * Conventions
 *  - plain pointers and sizes only; no C++/torch types; every function
 *    returns 0 on success, non-zero on error (omega_last_error() has text);
 *    nothing throws, nothing synchronises unless the name says so;
 *  - "host order"   = Omega's local element order (Decomp.cpp local
 *    numbering: owned, then halo layers; row N*All is the boundary slot,
 *    arrays are N*Size = N*All+1 rows, src/base/Decomp.cpp:1081-1082);
 *  - "device order" = the library's space-filling-curve order of the same
 *    rows (owned block and halo block sorted separately, boundary slot
 *    last); omega_mesh_perm() returns the bit-exact maps;
 *  - fields are OmegaReal (fp64; fp32 in libomega_b200_sp.so), layout [element][level] (LayoutRight,
 *    src/base/DataTypes.h:58-61), tracers [tracer][cell][level];
 *  - device pointers handed to compute entry points are in DEVICE order and
 *    must have been allocated with row count N*Size.
 */
#ifndef OMEGA_B200_H
#define OMEGA_B200_H

#ifdef __cplusplus
extern "C" {
#endif

/* Real of the reference (src/base/DataTypes.h:30-34): fp64 unless the library is built with
 * -DOMEGA_SINGLE_PRECISION (libomega_b200_sp.so).  Every FIELD array of this header (state, auxiliary,
 * tendency, halo and transfer buffers, the column pieces) is OmegaReal; mesh geometry handed to
 * omega_mesh_create, configuration values, times and stage coefficients stay double in both builds and
 * are converted inside.  omega_b200_real_bytes() tells a host which build it has loaded. */
#ifdef OMEGA_SINGLE_PRECISION
typedef float OmegaReal;
#else
typedef double OmegaReal;
#endif

/* src/base/Halo.h:45 */
typedef enum { OMEGA_ON_CELL = 0, OMEGA_ON_EDGE = 1, OMEGA_ON_VERTEX = 2 } OmegaMeshElement;

/* Host-side description of one rank's local mesh: the public members of
 * HorzMesh the hot path reads (src/ocn/HorzMesh.h:100-265), host order,
 * 0-based local indices, boundary slot = N*All. */
typedef struct OmegaMeshDesc {
   int NCellsOwned, NCellsAll, NCellsSize;
   int NEdgesOwned, NEdgesAll, NEdgesSize;
   int NVerticesOwned, NVerticesAll, NVerticesSize;
   int MaxEdges, MaxEdges2, VertexDegree;
   int NVertLayers;
   const int *NEdgesOnCell;   /* [NCellsSize] */
   const int *EdgesOnCell;    /* [NCellsSize][MaxEdges] */
   const int *CellsOnEdge;    /* [NEdgesSize][2] */
   const int *VerticesOnEdge; /* [NEdgesSize][2] */
   const int *NEdgesOnEdge;   /* [NEdgesSize] */
   const int *EdgesOnEdge;    /* [NEdgesSize][MaxEdges2] */
   const int *CellsOnVertex;  /* [NVerticesSize][VertexDegree] */
   const int *EdgesOnVertex;  /* [NVerticesSize][VertexDegree] */
   const double *AreaCell, *BottomDepth, *XCell, *YCell, *ZCell;  /* [NCellsSize] */
   const double *DcEdge, *DvEdge, *AngleEdge, *FEdge, *XEdge, *YEdge; /* [NEdgesSize] */
   const double *WeightsOnEdge;     /* [NEdgesSize][MaxEdges2] */
   const double *AreaTriangle, *FVertex; /* [NVerticesSize] */
   const double *KiteAreasOnVertex; /* [NVerticesSize][VertexDegree] */
} OmegaMeshDesc;

/* Tendencies: + Advection: + WindStress: config groups
 * (configs/Default.yml:25-52, src/ocn/Tendencies.cpp:123-213,
 * src/ocn/AuxiliaryState.cpp:267-310) and the manufactured-solution custom
 * tendencies (src/ocn/CustomTendencyTerms.cpp:18-102). */
typedef struct OmegaTendConfig {
   int ThicknessFluxTendencyEnable, PVTendencyEnable, KETendencyEnable,
       SSHTendencyEnable, VelDiffTendencyEnable, VelHyperDiffTendencyEnable,
       WindForcingTendencyEnable, BottomDragTendencyEnable,
       TracerHorzAdvTendencyEnable, TracerDiffTendencyEnable,
       TracerHyperDiffTendencyEnable;
   int FluxThicknessType;   /* 0 Center, 1 Upwind */
   int FluxTracerType;      /* 0 Center, 1 Upwind */
   int WindStressInterpType; /* 0 Isotropic, 1 Anisotropic */
   int ManufacturedSolutionTendency; /* adds the two custom terms */
   double ViscDel2, ViscDel4, DivFactor, Density0, BottomDragCoeff,
       EddyDiff2, EddyDiff4;
   /* ManufacturedSolution: group (configs/Default.yml:143-146) */
   double ManufacturedWavelengthX, ManufacturedWavelengthY, ManufacturedAmplitude;
} OmegaTendConfig;

/* Public array members of AuxiliaryState's six aux-var groups
 * (src/ocn/auxiliaryVars/*.h).  Device order for the GPU library, host
 * order for the oracle.  Cell/edge/vertex arrays are [N*Size][NVertLayers];
 * HTracersEdge [NTracers][NEdgesSize][K]; Del2TracersCell
 * [NTracers][NCellsSize][K]; the three wind arrays are 1-D. */
typedef struct OmegaAuxArrays {
   OmegaReal *KineticEnergyCell, *VelocityDivCell;
   OmegaReal *FluxLayerThickEdge, *MeanLayerThickEdge, *SshCell;
   OmegaReal *RelVortVertex, *NormRelVortVertex, *NormPlanetVortVertex;
   OmegaReal *NormRelVortEdge, *NormPlanetVortEdge;
   OmegaReal *Del2Edge, *Del2DivCell, *Del2RelVortVertex;
   OmegaReal *HTracersEdge, *Del2TracersCell;
   OmegaReal *NormalStressEdge, *ZonalStressCell, *MeridStressCell;
} OmegaAuxArrays;

/* Tendencies::LayerThicknessTend / NormalVelocityTend / TracerTend
 * (src/ocn/Tendencies.h:55-57). */
typedef struct OmegaTendArrays {
   OmegaReal *LayerThicknessTend; /* [NCellsSize][K] */
   OmegaReal *NormalVelocityTend; /* [NEdgesSize][K] */
   OmegaReal *TracerTend;         /* [NTracers][NCellsSize][K] */
} OmegaTendArrays;

/* OceanState + Tracers time-level rings (src/ocn/OceanState.cpp:65-117,
 * 348-365,394-407; src/ocn/Tracers.cpp:269-277,457-496). */
typedef struct OmegaStateArrays {
   int NTimeLevels;      /* 2 for every stepper in the reference */
   int CurTimeIndex;     /* OceanState::CurTimeIndex */
   int TracerCurTimeIndex; /* Tracers::CurTimeIndex */
   int NTracers;
   OmegaReal *LayerThickness[2]; /* [NCellsSize][K] per ring slot */
   OmegaReal *NormalVelocity[2]; /* [NEdgesSize][K] */
   OmegaReal *Tracers[2];        /* [NTracers][NCellsSize][K] */
} OmegaStateArrays;

/* Stage coefficients already multiplied by the time step and evaluated the
 * way TimeInterval::get does (src/infra/TimeMgr.cpp:382-391), in seconds;
 * StageTime = elapsed seconds (since the manufactured-solution reference
 * time) at which stage s evaluates its tendencies. */
typedef struct OmegaStepCoeffs {
   int NStages;
   double A[4], B[4];
   double StageTime[4];
   double Dt;
} OmegaStepCoeffs;

typedef enum { OMEGA_STEPPER_RK4 = 0, OMEGA_STEPPER_RK2 = 1, OMEGA_STEPPER_FB = 2 } OmegaStepperType;

typedef struct OmegaMesh OmegaMesh;       /* device-resident HorzMesh */
typedef struct OmegaStepper OmegaStepper; /* Tendencies + AuxiliaryState + TimeStepper workspace */
typedef struct OmegaHalo OmegaHalo;       /* Halo exchange lists on device */

const char *omega_last_error(void);
int omega_b200_abi_version(void);
int omega_b200_real_bytes(void); /* sizeof(OmegaReal): 8, or 4 in the single-precision build */

/* ---- device memory (replaces Kokkos View allocation, zero-initialised) -- */
int omega_dev_alloc(void **ptr, unsigned long long bytes);
int omega_dev_free(void *ptr);
int omega_dev_copy(void *dst, const void *src, unsigned long long bytes, void *stream); /* device to device, async */
int omega_dev_sync(void *stream);
/* plain byte copies between host and device memory (Kokkos::deep_copy of index arrays and the like);
 * both return when the copy is complete */
int omega_dev_upload(void *dev_dst, const void *host_src, unsigned long long bytes);
int omega_dev_download(void *host_dst, const void *dev_src, unsigned long long bytes);

/* ---- HorzMesh (src/ocn/HorzMesh.cpp:49-150,527-647) ---------------------- */
/* Builds the SFC permutation (Reorder=1) or keeps host order (Reorder=0),
 * derives EdgeSignOnCell/EdgeSignOnVertex/EdgeMask/MeshScalingDel2,4 and
 * stages everything to the current device as structure-of-arrays. */
int omega_mesh_create(const OmegaMeshDesc *Desc, int Reorder, OmegaMesh **Mesh);
int omega_mesh_destroy(OmegaMesh *Mesh);
/* Perm[new] = old (host-order row of device row `new`), IPerm[old] = new;
 * both length N*Size for the element kind. */
int omega_mesh_perm(const OmegaMesh *Mesh, OmegaMeshElement Elem, int *Perm, int *IPerm);
/* Derived host-order arrays, for parity with HorzMesh::computeEdgeSign /
 * setMasks: EdgeSignOnCell [NCellsSize][MaxEdges], EdgeSignOnVertex
 * [NVerticesSize][VertexDegree], EdgeMask [NEdgesSize] (level-independent). */
int omega_mesh_derived(const OmegaMesh *Mesh, double *EdgeSignOnCell,
                       double *EdgeSignOnVertex, double *EdgeMask);

/* ---- host <-> device field transfer with the permutation applied -------
 * Host array [NOuter][N*Size][NInner] in host order <-> device array of the
 * same shape in device order (OceanState::copyToDevice/copyToHost,
 * src/ocn/OceanState.cpp:300-328). Host may be pageable or pinned. */
int omega_field_to_device(const OmegaMesh *Mesh, OmegaMeshElement Elem, int NOuter,
                          int NInner, const OmegaReal *Host, OmegaReal *Dev, void *Stream);
int omega_field_to_host(const OmegaMesh *Mesh, OmegaMeshElement Elem, int NOuter,
                        int NInner, const OmegaReal *Dev, OmegaReal *Host, void *Stream);
/* Same, without the final stream synchronisation: Host must be pinned and is valid
 * once the caller has synchronised Stream (or an event recorded after the call).
 * omega_field_to_device never synchronises; with a pinned Host it is asynchronous too,
 * so transfers on other streams overlap the step (bench.py e2e pipeline). */
int omega_field_to_host_async(const OmegaMesh *Mesh, OmegaMeshElement Elem, int NOuter,
                              int NInner, const OmegaReal *Dev, OmegaReal *Host, void *Stream);
/* omega_field_to_device for callers that may not allocate device memory (stream
 * capture, serving loops, harnesses that forbid cudaMalloc). It has NO counterpart in
 * the reference. Same arguments and
 * result as omega_field_to_device (bit for bit), but the staging block of the permute
 * kernel is a static device array inside the library (OMEGA_NOALLOC_MAX_BYTES per GPU,
 * held from library load to process exit) and the call makes no cudaMalloc / cudaFree /
 * stream-ordered allocation. A field larger than OMEGA_NOALLOC_MAX_BYTES
 * (NOuter * N*Size * NInner * sizeof(OmegaReal) bytes) is rejected with return code 2 and a message in
 * omega_last_error. All calls on one GPU share the block: calls that overlap on the GPU
 * (different streams, no ordering between them) race; on one stream they are safe. For
 * capture in a CUDA graph Host must be pinned. Never synchronises. */
#define OMEGA_NOALLOC_MAX_BYTES (32u << 20)
int omega_field_to_device_noalloc(const OmegaMesh *Mesh, OmegaMeshElement Elem, int NOuter,
                                  int NInner, const OmegaReal *Host, OmegaReal *Dev, void *Stream);

/* ---- Tendencies + AuxiliaryState + TimeStepper -------------------------- */
int omega_stepper_create(const OmegaMesh *Mesh, const OmegaTendConfig *Cfg, int NTracers,
                         OmegaStepper **Stepper);
int omega_stepper_destroy(OmegaStepper *Stepper);
int omega_stepper_set_config(OmegaStepper *Stepper, const OmegaTendConfig *Cfg);
/* The library owns the aux / tendency arrays (as AuxiliaryState and
 * Tendencies do); these return the device pointers so host Views can alias
 * them (src/ocn/Tendencies.h:55-57, src/ocn/AuxiliaryState.h:37-42). */
int omega_stepper_aux_arrays(OmegaStepper *Stepper, OmegaAuxArrays *Aux);
int omega_stepper_tend_arrays(OmegaStepper *Stepper, OmegaTendArrays *Tend);
/* wind stress input on cells, host order (WindForcingAuxVars.h:14-17) */
int omega_stepper_set_wind_stress(OmegaStepper *Stepper, const OmegaReal *ZonalStressCellHost,
                                  const OmegaReal *MeridStressCellHost, void *Stream);

/* AuxiliaryState::computeMomAux / computeAll (src/ocn/AuxiliaryState.cpp:60-185) */
int omega_aux_compute_mom(OmegaStepper *S, const OmegaReal *LayerThickCell,
                          const OmegaReal *NormalVelEdge, void *Stream);
int omega_aux_compute_all(OmegaStepper *S, const OmegaReal *LayerThickCell,
                          const OmegaReal *NormalVelEdge, const OmegaReal *TracerArray,
                          void *Stream);
/* Tendencies::compute*TendenciesOnly (src/ocn/Tendencies.cpp:257-486);
 * NormalVelEdgeRaw is State->NormalVelocity[VelTimeLevel] indexed RAW as the
 * reference does for bottom drag / tracer advection (:322,:445). */
int omega_tend_thickness_only(OmegaStepper *S, const OmegaReal *NormalVelEdge, double Time,
                              void *Stream);
int omega_tend_velocity_only(OmegaStepper *S, const OmegaReal *NormalVelEdge,
                             const OmegaReal *NormalVelEdgeRaw, double Time, void *Stream);
int omega_tend_tracer_only(OmegaStepper *S, const OmegaReal *NormalVelEdgeRaw,
                           const OmegaReal *TracerArray, void *Stream);
/* Tendencies::computeThickness/Velocity/Tracer/AllTendencies (:488-600) */
int omega_tend_thickness(OmegaStepper *S, const OmegaReal *LayerThickCell,
                         const OmegaReal *NormalVelEdge, double Time, void *Stream);
int omega_tend_velocity(OmegaStepper *S, const OmegaReal *LayerThickCell,
                        const OmegaReal *NormalVelEdge, const OmegaReal *NormalVelEdgeRaw,
                        double Time, void *Stream);
int omega_tend_tracer(OmegaStepper *S, const OmegaReal *LayerThickCellRaw,
                      const OmegaReal *NormalVelEdgeRaw, const OmegaReal *TracerArray,
                      void *Stream);
int omega_tend_all(OmegaStepper *S, const OmegaReal *LayerThickCell,
                   const OmegaReal *NormalVelEdge, const OmegaReal *NormalVelEdgeRaw,
                   const OmegaReal *TracerArray, double Time, void *Stream);

/* Tendencies::CustomTendencyType (src/ocn/Tendencies.h:51-53; invoked at Tendencies.cpp:288-293 for the thickness
 * and :416-421 for the velocity): a host term that runs AFTER the built-in terms of a thickness / velocity tendency
 * evaluation and adds into the tendency array.  Tend, LayerThickness and NormalVelocity are DEVICE arrays in device
 * order ([NCellsSize][K] / [NEdgesSize][K] rows permuted as omega_mesh_perm says; an element-wise term does not
 * care); LayerThickness is NULL in the ...Only evaluations, which take no thickness.  The callback enqueues its
 * work on Stream and returns 0; it runs on the calling host thread, inside omega_tend_* and inside every stage of
 * omega_stepper_do_step / step_phase.  While a custom term is installed, Fused=1 steps on the kernel-per-term
 * path (the fused stages never materialise the tendency arrays).  NULL, NULL removes both. */
typedef int (*OmegaCustomTendency)(void *User, OmegaReal *Tend, const OmegaReal *LayerThickness,
                                   const OmegaReal *NormalVelocity, double Time, void *Stream);
int omega_stepper_set_custom_tendencies(OmegaStepper *S, OmegaCustomTendency Thickness, OmegaCustomTendency Velocity,
                                        void *User);

/* TimeStepper::doStep (src/timeStepping/RungeKutta4Stepper.cpp:68-137,
 * RungeKutta2Stepper.cpp:27-73, ForwardBackwardStepper.cpp:27-82).
 * Advances State by one step and rotates both time-level rings.  Fused=1
 * runs the fused per-stage pipeline (no aux/tendency arrays materialised),
 * Fused=0 the kernel-per-term path that fills every public array.  The fused kernels are written for at most six
 * edges per cell: on a mesh with heptagons (MaxEdges >= 7 in use) Fused=1 steps on the kernel-per-term path, same
 * results (omega_stepper_last_path tells which path ran).
 * RawTimeLevelCompat=1 reproduces the reference's raw time-level indexing
 * (SURVEY.md 3.5 P1); 0 maps every level through the ring.
 * Halo may be NULL on a single rank; with a Halo (communicator initialised) the step runs its
 * exchange rounds itself, overlapped with the interior tiles on the fused tiled path.
 * Fused=1 has no place for a host custom-tendency std::function (src/ocn/Tendencies.h:51-53,
 * Tendencies.cpp:288-293,416-421): the built-in manufactured terms are the only custom terms the fused
 * kernels evaluate; with omega_stepper_set_custom_tendencies in force the step runs on the kernel-per-term path,
 * the callbacks called in every stage.
 * After one warm-up step with the same stepper and configuration the call allocates nothing and never
 * synchronises (Halo == NULL, profiling off): a caller may capture it in a CUDA graph (scratch/probe_graph.py). */
int omega_stepper_do_step(OmegaStepper *S, OmegaStepperType Type, OmegaStateArrays *State,
                          const OmegaStepCoeffs *Coeffs, OmegaHalo *Halo, int Fused,
                          int RawTimeLevelCompat, void *Stream);
/* The same step cut at its halo-exchange points, for callers that run the
 * exchange themselves (NCCL in another library, an emulated multi-rank
 * test): Phase p runs everything up to (not including) exchange point p. */
int omega_stepper_step_phase(OmegaStepper *S, OmegaStepperType Type, OmegaStateArrays *State,
                             const OmegaStepCoeffs *Coeffs, int Phase, int Fused,
                             int RawTimeLevelCompat, void *Stream);
/* Arrays to exchange after Phase p (device pointers, element kind, NOuter). */
int omega_stepper_exchange_list(OmegaStepper *S, OmegaStepperType Type,
                                const OmegaStateArrays *State, int Phase, OmegaReal **Ptrs,
                                int *Elems, int *NOuter, int *NArrays);
/* Attach the rank's Halo to the stepper (do_step with a Halo does this itself).  The fused tiled path then
 * (a) skips the halo tiles in the stages whose results the following exchange overwrites (RK4 stage 3, and
 * stage 1 unless the raw time-level indexing makes its halo thickness matter) and (b) launches the tiles that hold
 * send-list rows first, so the exchange overlaps the interior tiles.  Callers of step_phase that run
 * the exchange themselves may attach too; they must exchange after every phase. */
int omega_stepper_attach_halo(OmegaStepper *S, OmegaHalo *Halo);
/* OceanState::updateTimeLevels + Tracers::updateTimeLevels minus the exchange:
 * advance both ring indices (src/ocn/OceanState.cpp:348-365). */
int omega_state_rotate(OmegaStateArrays *State);
/* ProvisState / ProvisTracers of the RK4 stepper (RungeKutta4Stepper.cpp:43-64) */
int omega_stepper_provis_arrays(OmegaStepper *Stepper, OmegaReal **H, OmegaReal **U, OmegaReal **Tr);
/* kernels launched by this stepper since creation (bench gpu_launches) */
long long omega_stepper_launch_count(const OmegaStepper *S);
/* which kernels the last omega_stepper_step_phase / do_step ran: OMEGA_PATH_* bits (0 = the
 * kernel-per-term path).  Tests assert through this that a case exercised the tiled kernels. */
enum { OMEGA_PATH_FUSED = 1, OMEGA_PATH_TILED_CELL = 2, OMEGA_PATH_CELL_V2 = 4, OMEGA_PATH_TILED_EDGE = 8,
       OMEGA_PATH_HALO_SPLIT = 16 /* boundary tiles first, halo tiles skipped (attached halo) */ };
int omega_stepper_last_path(const OmegaStepper *S);
/* per-kernel CUDA-event timing of the fused pipeline: sums in ms and launch
 * counts per kernel kind (0 fA, 1 fB, 2 fT2, 3 fC, 4 fE) since the last call;
 * get_profile synchronises the device. */
int omega_stepper_set_profiling(OmegaStepper *S, int Enable);
int omega_stepper_get_profile(OmegaStepper *S, double *Ms, int *Counts, int NKinds);

/* ---- Halo (src/base/Halo.h:62-917, src/base/Halo.cpp:455-757) ----------- */
/* Exchange lists for one rank, host order: for each neighbour n and element
 * kind e, SendIdx/RecvIdx are the concatenation over halo layers
 * (Halo.cpp:543-597). Offsets are CSR-style [NNeighbors+1] per kind. */
typedef struct OmegaHaloDesc {
   int MyTask, NNeighbors;
   const int *NeighborTasks;            /* [NNeighbors] ascending */
   const int *SendOffsets[3], *SendIdx[3];
   const int *RecvOffsets[3], *RecvIdx[3];
} OmegaHaloDesc;
int omega_halo_create(const OmegaMesh *Mesh, const OmegaHaloDesc *Desc, OmegaHalo **Halo);
int omega_halo_destroy(OmegaHalo *Halo);

/* Transport: NCCL send/recv over NVLink, one rank per GPU, loaded at run time (libnccl.so.2).
 * The host creates the id on one rank, broadcasts its 128 bytes with whatever it has (MPI_Bcast in
 * Omega, where MachEnv owns the MPI communicator: src/base/MachEnv.h; torch.distributed in the
 * Python mirror) and every rank calls comm_init with Rank == OmegaHaloDesc.MyTask.  A host that
 * already owns an ncclComm_t over the same ranks hands it in with set_comm instead. */
int omega_halo_nccl_unique_id(char *Id128);
int omega_halo_comm_init(OmegaHalo *Halo, const char *Id128, int NRanks, int Rank);
int omega_halo_set_comm(OmegaHalo *Halo, void *NcclComm, int NRanks);

/* Halo::exchangeFullArrayHalo(Array, Elem) (src/base/Halo.h:768-915): the whole exchange of one
 * array [NOuter][N*Size][NVertLayers] (device order) on Stream: pack kernel, grouped ncclSend /
 * ncclRecv with every neighbour, unpack kernel.  Nothing synchronises the host. */
int omega_halo_exchange(OmegaHalo *Halo, OmegaReal *Array, OmegaMeshElement Elem, int NOuter, void *Stream);
/* One exchange ROUND: up to three arrays (h, u, tracers of one exchange point of the stepper) in ONE
 * pack launch, ONE message per neighbour, ONE unpack launch.  Buffers are owned by the Halo. */
int omega_halo_exchange_round(OmegaHalo *Halo, OmegaReal *const *Ptrs, const int *Elems, const int *NOuter,
                              int NArrays, void *Stream);
/* The two device halves of a round for callers that own the transport (peer copies, an in-process
 * N-rank emulation): pack_round fills the Halo's send buffer (returned; neighbour n's message is the
 * concatenation of its per-array buffers, each [outer][exchange index][level], Halo.h:304-502, and
 * messages follow each other in neighbour order; round_counts gives their sizes in doubles);
 * recv_buffer returns the buffer unpack_round scatters from, laid out the same way. */
int omega_halo_round_counts(const OmegaHalo *Halo, const int *Elems, const int *NOuter, int NArrays,
                            long long *SendCounts, long long *RecvCounts);
int omega_halo_pack_round(OmegaHalo *Halo, OmegaReal *const *Ptrs, const int *Elems, const int *NOuter,
                          int NArrays, OmegaReal **SendBuf, void *Stream);
int omega_halo_recv_buffer(OmegaHalo *Halo, const int *Elems, const int *NOuter, int NArrays, OmegaReal **RecvBuf);
int omega_halo_unpack_round(OmegaHalo *Halo, OmegaReal *const *Ptrs, const int *Elems, const int *NOuter,
                            int NArrays, void *Stream);
/* Per-array device halves with caller-owned buffers (Halo.h:304-502 buffer order; counts in doubles). */
int omega_halo_buffer_counts(const OmegaHalo *Halo, OmegaMeshElement Elem, int NOuter,
                             int NInner, long long *SendCounts, long long *RecvCounts);
int omega_halo_pack(const OmegaHalo *Halo, OmegaMeshElement Elem, int NOuter, int NInner,
                    const OmegaReal *Array, OmegaReal *SendBuf, void *Stream);
int omega_halo_unpack(const OmegaHalo *Halo, OmegaMeshElement Elem, int NOuter, int NInner,
                      const OmegaReal *RecvBuf, OmegaReal *Array, void *Stream);
/* The same exchange for arrays of any element type the reference's Halo moves (I4, I8, R4, R8; 1 to 5
 * dimensions, Halo.h:768-915, test/base/HaloTest.cpp:41-99): Array is [NOuter][N*Size][NInner] in device
 * order with NOuter the product of the dimensions in front of the element dimension and NInner the
 * dimension behind it (1 for a 1-D array); ElemBytes is 4 or 8.  pack_any / unpack_any are the device
 * halves with caller-owned buffers (omega_halo_buffer_counts entries, times ElemBytes, in bytes). */
int omega_halo_exchange_any(OmegaHalo *Halo, void *Array, OmegaMeshElement Elem, int NOuter, int NInner,
                            int ElemBytes, void *Stream);
int omega_halo_pack_any(const OmegaHalo *Halo, OmegaMeshElement Elem, int NOuter, int NInner, int ElemBytes,
                        const void *Array, void *SendBuf, void *Stream);
int omega_halo_unpack_any(const OmegaHalo *Halo, OmegaMeshElement Elem, int NOuter, int NInner, int ElemBytes,
                          const void *RecvBuf, void *Array, void *Stream);
/* Tile classes used for the overlap: tiles with send-list rows, owned tiles, all tiles. */
int omega_halo_tile_classes(const OmegaHalo *Halo, int *NBoundary, int *NOwned, int *NTiles);
/* CUDA-event timing of the exchange rounds since the last call (pack .. unpack on the exchange
 * stream): total ms, rounds, bytes sent + received; get_profile synchronises the device. */
int omega_halo_set_profiling(OmegaHalo *Halo, int Enable);
int omega_halo_get_profile(OmegaHalo *Halo, double *Ms, long long *Rounds, long long *Bytes);

/* ---- Omega-1 column pieces around the horizontal path ------------
 * Device arrays in device order, [NCellsSize][NVertLayers] unless noted; MinLayerCell / MaxLayerCell
 * [NCellsSize] int (device order): the active layer range of a column (VertCoord.h).
 * VertCoord::computePressure / computeZHeight / computeGeopotential / computeTargetThickness
 * (src/ocn/VertCoord.cpp:648-860): interfaces are [NCellsSize][NVertLayers + 1]; surface pressure, tidal
 * potential, self attraction and loading [NCellsSize]; MovementWeights [NVertLayers]; BottomDepth is the mesh's. */
int omega_vert_compute_pressure(const OmegaMesh *Mesh, const int *MinLayerCell, const int *MaxLayerCell, double Rho0,
                                const OmegaReal *LayerThickness, const OmegaReal *SurfacePressure,
                                OmegaReal *PressureInterface, OmegaReal *PressureMid, void *Stream);
int omega_vert_compute_z_height(const OmegaMesh *Mesh, const int *MinLayerCell, const int *MaxLayerCell, double Rho0,
                                const OmegaReal *LayerThickness, const OmegaReal *SpecVol, OmegaReal *ZInterface,
                                OmegaReal *ZMid, void *Stream);
int omega_vert_compute_geopotential(const OmegaMesh *Mesh, const int *MinLayerCell, const int *MaxLayerCell,
                                    const OmegaReal *ZMid, const OmegaReal *TidalPotential,
                                    const OmegaReal *SelfAttractionLoading, OmegaReal *GeopotentialMid, void *Stream);
int omega_vert_compute_target_thickness(const OmegaMesh *Mesh, const int *MinLayerCell, const int *MaxLayerCell,
                                        double Rho0, const OmegaReal *PressureInterface, const OmegaReal *RefLayerThickness,
                                        const OmegaReal *MovementWeights, OmegaReal *LayerThicknessTarget, void *Stream);
/* Eos::computeSpecVol / computeSpecVolDisp (src/ocn/Eos.h:29-292).  EosType 0: linear, 1 / (RhoT0S0 + DRhodT CT +
 * DRhodS SA); 1: TEOS-10 75-term polynomial of Roquet et al. (2015), pressure in dbar.  KDisp != 0: specific volume
 * of the parcel displaced to the pressure of layer clamp(k + KDisp) (TEOS-10). */
int omega_eos_compute_spec_vol(const OmegaMesh *Mesh, int EosType, const OmegaReal *ConservTemp, const OmegaReal *AbsSalinity,
                               const OmegaReal *Pressure, int KDisp, double DRhodT, double DRhodS, double RhoT0S0,
                               OmegaReal *SpecVol, void *Stream);
/* TriDiagSolver / TriDiagDiffSolver (src/base/TriDiagSolvers.h): NBatch independent systems of NRow <= 1024 rows,
 * arrays [NBatch][NRow] on the device, X holds the right-hand side on entry and the solution on return.
 * General form: DL(k) x(k-1) + D(k) x(k) + DU(k) x(k+1) = X(k).  Diffusion form: D(k) = H(k) + G(k-1) + G(k),
 * DL(k) = -G(k-1), DU(k) = -G(k), solved without ever forming D (stable for abrupt jumps in G). */
int omega_tridiag_solve(int NBatch, int NRow, const OmegaReal *DL, const OmegaReal *D, const OmegaReal *DU, OmegaReal *X, void *Stream);
int omega_tridiag_diffusion_solve(int NBatch, int NRow, const OmegaReal *G, const OmegaReal *H, OmegaReal *X, void *Stream);

/* ---- Decomp (src/base/Decomp.cpp) -- host code, no device work ------------------
 * One rank's partition of the GLOBAL mesh in Omega's local numbering, the owner
 * address of every local element (CellLoc / EdgeLoc / VertexLoc, Decomp.cpp:1094-1098,1639-1676) and the
 * halo exchange lists (src/base/Halo.cpp:455-600), computed by the rank alone from the global
 * connectivity and the partition: it also numbers its neighbours' partitions, which stands in for the
 * reference's two MPI exchanges (broadcast of owned lists, mailing of the receive lists).
 * Global arrays are 0-based with -1 for "none" (file values minus one). */
typedef struct OmegaGlobalMesh {
   int NCells, NEdges, NVertices, MaxEdges, MaxEdges2, VertexDegree;
   const int *CellsOnCell, *EdgesOnCell, *VerticesOnCell; /* [NCells][MaxEdges] */
   const int *CellsOnEdge, *VerticesOnEdge;               /* [NEdges][2] */
   const int *EdgesOnEdge;                                /* [NEdges][MaxEdges2] */
   const int *CellsOnVertex, *EdgesOnVertex;              /* [NVertices][VertexDegree] */
   const double *XCell, *YCell, *ZCell;                   /* partitioner input */
} OmegaGlobalMesh;
typedef struct OmegaDecomp OmegaDecomp;
/* Method 0: recursive coordinate bisection; 1: the same followed by boundary refinement of the edge cut
 * (balance kept within 1 %).  The reference calls METIS_PartGraphKway (Decomp.cpp:968-971, third party);
 * results on owned elements do not depend on the partition.  EdgeCut may be NULL. */
int omega_decomp_partition(const OmegaGlobalMesh *Mesh, int NTasks, int Method, int *CellTask, long long *EdgeCut);
int omega_decomp_edge_cut(const OmegaGlobalMesh *Mesh, const int *CellTask, long long *EdgeCut);
int omega_decomp_create(const OmegaGlobalMesh *Mesh, const int *CellTask, int NTasks, int MyTask, int HaloWidth,
                        OmegaDecomp **Decomp);
int omega_decomp_destroy(OmegaDecomp *Decomp);
/* per element kind (cell, edge, vertex): {NOwned, NAll, NHalo(0 .. HaloWidth-1)} = 3 * (2 + HaloWidth) ints */
int omega_decomp_counts(const OmegaDecomp *Decomp, int *Counts);
/* global 0-based ids of the local elements in local order [NAll]; owner address [NAll+1][2] */
int omega_decomp_ids(const OmegaDecomp *Decomp, OmegaMeshElement Elem, int *IDs);
int omega_decomp_loc(const OmegaDecomp *Decomp, OmegaMeshElement Elem, int *Loc);
/* local connectivity, boundary slot for anything not local (Decomp.cpp:552-712).  Which: 0 CellsOnCell,
 * 1 EdgesOnCell, 2 VerticesOnCell, 3 CellsOnEdge, 4 VerticesOnEdge, 5 EdgesOnEdge, 6 CellsOnVertex,
 * 7 EdgesOnVertex; Out has NAll + 1 rows. */
int omega_decomp_connectivity(const OmegaDecomp *Decomp, int Which, int *Out);
/* halo lists in the layout of OmegaHaloDesc: neighbours ascending, CSR offsets [NNeighbors+1] per kind */
int omega_decomp_halo_neighbors(const OmegaDecomp *Decomp, int *NNeighbors, int *Tasks);
int omega_decomp_halo_lists(const OmegaDecomp *Decomp, OmegaMeshElement Elem, int *SendOff, int *SendIdx,
                            int *RecvOff, int *RecvIdx);

#ifdef __cplusplus
}
#endif
#endif /* OMEGA_B200_H */
