/* standins/mpi.h -- single-process stand-in for the MPI calls the reference's
 * HOST headers make (fieldtraits.hh, stochastic.hh, randomfield.hh).  MPI is
 * absent from this image; the CUDA path is driven by one process per plan, so
 * every collective is a copy and a message to oneself goes through a queue. */
#ifndef PF_STANDIN_MPI_H
#define PF_STANDIN_MPI_H
#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype; /* = size of one element in bytes */
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
