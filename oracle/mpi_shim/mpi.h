/* TEST INFRASTRUCTURE ONLY -- stand-in <mpi.h> for building the reference CPU solver as an oracle.
 *
 * The image has no MPI. The reference solver TUs (source/sartsolver.cpp, source/sartsolver_cuda.cpp)
 * need only a handful of names: MPI_Comm / MPI_Datatype / MPI_Op, MPI_COMM_WORLD, MPI_DOUBLE,
 * MPI_FLOAT, MPI_SUM, MPI_MAX and MPI_Allreduce (reference call sites: source/sartsolver.cpp:47,
 * 158, 164, 206, 222, 260, 268, 312, 329 and source/sartsolver_cuda.cpp:149, 156, 174, 243, 255).
 * "Ranks" are threads of one process; the implementation of MPI_Allreduce lives in
 * oracle/ref_harness.cpp (barrier + rank-ordered sum, which is bit-identical to any real MPI for
 * two ranks because a+b is commutative).
 *
 * Nothing under oracle/ is part of the product: only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may use it.
 */
#pragma once
#include <cstddef>

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;

#define MPI_COMM_WORLD 0
#define MPI_SUCCESS 0

#define MPI_DOUBLE 1
#define MPI_FLOAT 2

#define MPI_SUM 1
#define MPI_MAX 2

int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype datatype, MPI_Op op,
                  MPI_Comm comm);
