/* Single-rank stand-in for <mpi.h>, just wide enough for the reference's bench drivers
 * (bench_mass.cpp:53-54,92,97-101,135-137,153,187,198).  Used only to build those drivers
 * unchanged on hosts without MPI; with a real MPI installation use that instead. */
#ifndef FEMTC_MPI_STUB_H
#define FEMTC_MPI_STUB_H
#include <string.h>
#include <time.h>
typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
#define MPI_COMM_WORLD 0
#define MPI_DOUBLE 8
#define MPI_SUM 1
#define MPI_SUCCESS 0
static inline int MPI_Init(int* argc, char*** argv) { (void)argc; (void)argv; return MPI_SUCCESS; }
static inline int MPI_Finalize(void) { return MPI_SUCCESS; }
static inline int MPI_Comm_rank(MPI_Comm c, int* rank) { (void)c; *rank = 0; return MPI_SUCCESS; }
static inline int MPI_Comm_size(MPI_Comm c, int* size) { (void)c; *size = 1; return MPI_SUCCESS; }
static inline int MPI_Barrier(MPI_Comm c) { (void)c; return MPI_SUCCESS; }
static inline double MPI_Wtime(void)
{
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}
static inline int MPI_Reduce(const void* send, void* recv, int count, MPI_Datatype t, MPI_Op op, int root, MPI_Comm c)
{
  (void)op; (void)root; (void)c;
  memcpy(recv, send, (size_t)count * (size_t)t);
  return MPI_SUCCESS;
}
#endif
