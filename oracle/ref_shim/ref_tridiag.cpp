// ref_tridiag.cpp -- TEST INFRASTRUCTURE.  The reference's own src/base/TriDiagSolvers.h, compiled unmodified over the
// Kokkos stand-in (one-thread teams, team scratch = a bump allocator): the array forms of the CPU solvers,
// ThomasSolver::solve and ThomasDiffusionSolver::solve (TriDiagSolvers.h:97-135, 324-360), on plain arrays
// [NBatch][NRow].  Pins oracle_tridiag_thomas / oracle_tridiag_diffusion_thomas (tests/test_vertical.py).
#include "TriDiagSolvers.h"

using namespace OMEGA;

extern "C" int ref_tridiag_thomas(int NBatch, int NRow, const double *DL, const double *D, const double *DU, double *X) {
   Array2DReal a("DL", NBatch, NRow), b("D", NBatch, NRow), c("DU", NBatch, NRow), x("X", NBatch, NRow);
   for (int i = 0; i < NBatch * NRow; ++i) { a.data()[i] = DL[i]; b.data()[i] = D[i]; c.data()[i] = DU[i]; x.data()[i] = X[i]; }
   ThomasSolver::solve(a, b, c, x);
   for (int i = 0; i < NBatch * NRow; ++i) X[i] = x.data()[i];
   return 0;
}
extern "C" int ref_tridiag_diffusion_thomas(int NBatch, int NRow, const double *G, const double *H, double *X) {
   Array2DReal g("G", NBatch, NRow), h("H", NBatch, NRow), x("X", NBatch, NRow);
   for (int i = 0; i < NBatch * NRow; ++i) { g.data()[i] = G[i]; h.data()[i] = H[i]; x.data()[i] = X[i]; }
   ThomasDiffusionSolver::solve(g, h, x);
   for (int i = 0; i < NBatch * NRow; ++i) X[i] = x.data()[i];
   return 0;
}
