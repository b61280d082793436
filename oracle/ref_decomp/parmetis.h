// parmetis.h -- TEST INFRASTRUCTURE stand-in: the types and the two entry points Decomp.cpp names.  The "partitioner"
// returns the partition the test driver supplies (ref_decomp_driver.cpp).
#pragma once
typedef int idx_t;
typedef float real_t;
#define METIS_OK 1
extern "C" int METIS_PartGraphKway(idx_t *nvtxs, idx_t *ncon, idx_t *xadj, idx_t *adjncy, idx_t *vwgt, idx_t *vsize, idx_t *adjwgt,
                                   idx_t *nparts, real_t *tpwgts, real_t *ubvec, idx_t *options, idx_t *edgecut, idx_t *part);
extern "C" int ParMETIS_V3_PartKway(idx_t *vtxdist, idx_t *xadj, idx_t *adjncy, idx_t *vwgt, idx_t *adjwgt, idx_t *wgtflag,
                                    idx_t *numflag, idx_t *ncon, idx_t *nparts, real_t *tpwgts, real_t *ubvec, idx_t *options,
                                    idx_t *edgecut, idx_t *part, MPI_Comm *comm);
