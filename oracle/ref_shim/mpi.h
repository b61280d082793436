// mpi.h -- TEST INFRASTRUCTURE stand-in: the compiled reference sources only name MPI_Comm.
#pragma once
typedef int MPI_Comm;
#define MPI_COMM_WORLD 0
