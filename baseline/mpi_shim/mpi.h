/* mpi.h - the subset of MPI that the reference (Youhe-Jiang/AllReduce-Over-MPI) uses, implemented over
 * shared memory between processes of ONE host (baseline/mpi_shim/mpi_shim.cc + ftmpirun).
 *
 * Why this exists: the image has no MPI installation and no network to fetch one, but the baseline
 * must be the reference's OWN code. With this header + glog_shim the *unmodified* reference sources
 * (allreduce_over_mpi/mpi_mod.hpp, benchmark.cpp) compile and run. Only the 13 functions and the
 * constants the reference touches are provided. Not a general MPI.
 */
#ifndef FT_MPI_SHIM_H_
#define FT_MPI_SHIM_H_
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef struct { int kind; int peer; unsigned long long seq; void* buf; size_t bytes; } MPI_Request;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;

#define MPI_SUCCESS 0
#define MPI_COMM_WORLD 0
#define MPI_IN_PLACE ((void*)-1)
#define MPI_THREAD_MULTIPLE 3
#define MPI_STATUSES_IGNORE ((MPI_Status*)0)

enum { MPI_SUM = 1, MPI_BAND = 2, MPI_MAX = 3, MPI_MIN = 4 };
enum {
  MPI_UINT8_T = 1, MPI_INT8_T, MPI_UINT16_T, MPI_INT16_T, MPI_INT32_T, MPI_INT64_T, MPI_FLOAT, MPI_DOUBLE,
  MPI_C_BOOL, MPI_LONG_LONG_INT, MPI_UINT32_T, MPI_UINT64_T
};
#define MPI_LONG_LONG MPI_LONG_LONG_INT
#define MPI_INT MPI_INT32_T

int MPI_Init_thread(int* argc, char*** argv, int required, int* provided);
int MPI_Finalize(void);
int MPI_Comm_size(MPI_Comm comm, int* size);
int MPI_Comm_rank(MPI_Comm comm, int* rank);
int MPI_Comm_split(MPI_Comm comm, int color, int key, MPI_Comm* newcomm);
int MPI_Type_size(MPI_Datatype dt, int* size);
int MPI_Type_get_name(MPI_Datatype dt, char* name, int* len);
int MPI_Isend(const void* buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request* req);
int MPI_Irecv(void* buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Request* req);
int MPI_Waitall(int count, MPI_Request* reqs, MPI_Status* statuses);
int MPI_Barrier(MPI_Comm comm);
double MPI_Wtime(void);
int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype dt, MPI_Op op, MPI_Comm comm);

#ifdef __cplusplus
}
#endif
#endif
