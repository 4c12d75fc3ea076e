/* oracle/shim/mpi.h -- single-rank stand-in for MPI.  TEST INFRASTRUCTURE.
 *
 * Exists only so that the UNMODIFIED reference headers under /root/reference
 * compile into oracle/_ref/ (see oracle/Makefile, DESIGN.md "Oracle").  MPI is
 * absent from this image; with one rank every collective is a copy and
 * point-to-point messages go through an in-process queue. */
#ifndef PF_SHIM_MPI_H
#define PF_SHIM_MPI_H
#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Request;
typedef int MPI_Info;
typedef struct MPI_Status { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;

#define MPI_COMM_WORLD 0
#define MPI_SUCCESS 0
#define MPI_INFO_NULL 0
#define MPI_REQUEST_NULL (-1)
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
#define MPI_STATUSES_IGNORE ((MPI_Status*)0)

/* datatype = size in bytes */
#define MPI_CHAR 1
#define MPI_INT 4
#define MPI_UNSIGNED 4
#define MPI_FLOAT 4
#define MPI_DOUBLE 8
#define MPI_LONG_DOUBLE 16
#define MPI_SUM 1
#define MPI_MIN 2
#define MPI_MAX 3

int MPI_Init(int*, char***);
int MPI_Finalize(void);
int MPI_Comm_rank(MPI_Comm, int*);
int MPI_Comm_size(MPI_Comm, int*);
int MPI_Barrier(MPI_Comm);
int MPI_Allreduce(const void*, void*, int, MPI_Datatype, MPI_Op, MPI_Comm);
int MPI_Isend(const void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Request*);
int MPI_Recv(void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Status*);
int MPI_Wait(MPI_Request*, MPI_Status*);
int MPI_Waitall(int, MPI_Request*, MPI_Status*);

#ifdef __cplusplus
}
#endif
#endif
