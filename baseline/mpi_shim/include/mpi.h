/* mpi.h of the single-node MPI stand-in used ONLY to build and run the reference
 * (sandialabs/tenzing) as the timed baseline.  Nothing in tenzing_b200 includes or links this.
 *
 * It implements the subset of MPI the reference calls (see baseline/README.md): one process per
 * rank on one host, control messages and host buffers through POSIX shared memory, device
 * buffers "CUDA-aware": the receiver maps the sender's allocation with CUDA IPC (handles cached,
 * as UCX's cuda_ipc transport does) and pulls the payload with cudaMemcpyAsync over NVLink on
 * one of several internal streams, completion by CUDA event.
 */
#ifndef TZ_MPI_SHIM_H
#define TZ_MPI_SHIM_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MPI_VERSION 3
#define MPI_SUBVERSION 1
#define TZ_MPI_SHIM 1

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Request;
typedef long MPI_Aint;
typedef long long MPI_Offset;
typedef long long MPI_Count;

typedef struct MPI_Status {
  int MPI_SOURCE;
  int MPI_TAG;
  int MPI_ERROR;
  long long tz_bytes; /* payload size of the matched message */
} MPI_Status;

#define MPI_SUCCESS 0
#define MPI_ERR_OTHER 15
#define MPI_COMM_NULL ((MPI_Comm)0)
#define MPI_COMM_WORLD ((MPI_Comm)1)
#define MPI_COMM_SELF ((MPI_Comm)2)
#define MPI_REQUEST_NULL ((MPI_Request)-1)
#define MPI_STATUS_IGNORE ((MPI_Status *)0)
#define MPI_STATUSES_IGNORE ((MPI_Status *)0)
#define MPI_IN_PLACE ((void *)-1)
#define MPI_ANY_SOURCE (-1)
#define MPI_ANY_TAG (-1)
#define MPI_PROC_NULL (-2)
#define MPI_UNDEFINED (-32766)
#define MPI_MAX_PROCESSOR_NAME 256
#define MPI_MAX_ERROR_STRING 256
#define MPI_THREAD_SINGLE 0
#define MPI_THREAD_FUNNELED 1
#define MPI_THREAD_SERIALIZED 2
#define MPI_THREAD_MULTIPLE 3

/* datatypes: the value is an index into the shim's size table */
#define MPI_DATATYPE_NULL ((MPI_Datatype)0)
#define MPI_CHAR ((MPI_Datatype)1)
#define MPI_BYTE ((MPI_Datatype)2)
#define MPI_INT ((MPI_Datatype)3)
#define MPI_FLOAT ((MPI_Datatype)4)
#define MPI_DOUBLE ((MPI_Datatype)5)
#define MPI_UNSIGNED ((MPI_Datatype)6)
#define MPI_LONG ((MPI_Datatype)7)
#define MPI_UNSIGNED_LONG ((MPI_Datatype)8)
#define MPI_LONG_LONG ((MPI_Datatype)9)
#define MPI_LONG_LONG_INT MPI_LONG_LONG
#define MPI_UNSIGNED_LONG_LONG ((MPI_Datatype)10)
#define MPI_SHORT ((MPI_Datatype)11)
#define MPI_UNSIGNED_CHAR ((MPI_Datatype)12)
#define MPI_INT64_T ((MPI_Datatype)13)
#define MPI_UINT64_T ((MPI_Datatype)14)
#define MPI_INT32_T ((MPI_Datatype)15)
#define MPI_UINT32_T ((MPI_Datatype)16)
#define MPI_C_BOOL ((MPI_Datatype)17)

#define MPI_OP_NULL ((MPI_Op)0)
#define MPI_MAX ((MPI_Op)1)
#define MPI_MIN ((MPI_Op)2)
#define MPI_SUM ((MPI_Op)3)

int MPI_Init(int *argc, char ***argv);
int MPI_Init_thread(int *argc, char ***argv, int required, int *provided);
int MPI_Finalize(void);
int MPI_Initialized(int *flag);
int MPI_Finalized(int *flag);
int MPI_Abort(MPI_Comm comm, int errorcode);
int MPI_Comm_rank(MPI_Comm comm, int *rank);
int MPI_Comm_size(MPI_Comm comm, int *size);
int MPI_Get_processor_name(char *name, int *resultlen);
double MPI_Wtime(void);
double MPI_Wtick(void);

int MPI_Barrier(MPI_Comm comm);
int MPI_Bcast(void *buffer, int count, MPI_Datatype datatype, int root, MPI_Comm comm);
int MPI_Allreduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype datatype, MPI_Op op,
                  MPI_Comm comm);

int MPI_Isend(const void *buf, int count, MPI_Datatype datatype, int dest, int tag, MPI_Comm comm,
              MPI_Request *request);
int MPI_Irecv(void *buf, int count, MPI_Datatype datatype, int source, int tag, MPI_Comm comm,
              MPI_Request *request);
int MPI_Send(const void *buf, int count, MPI_Datatype datatype, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void *buf, int count, MPI_Datatype datatype, int source, int tag, MPI_Comm comm,
             MPI_Status *status);
int MPI_Wait(MPI_Request *request, MPI_Status *status);
int MPI_Waitall(int count, MPI_Request array_of_requests[], MPI_Status array_of_statuses[]);
int MPI_Test(MPI_Request *request, int *flag, MPI_Status *status);
int MPI_Probe(int source, int tag, MPI_Comm comm, MPI_Status *status);
int MPI_Iprobe(int source, int tag, MPI_Comm comm, int *flag, MPI_Status *status);
int MPI_Get_count(const MPI_Status *status, MPI_Datatype datatype, int *count);
int MPI_Ialltoallv(const void *sendbuf, const int sendcounts[], const int sdispls[],
                   MPI_Datatype sendtype, void *recvbuf, const int recvcounts[], const int rdispls[],
                   MPI_Datatype recvtype, MPI_Comm comm, MPI_Request *request);
int MPI_Type_size(MPI_Datatype datatype, int *size);
int MPI_Error_string(int errorcode, char *string, int *resultlen);

#ifdef __cplusplus
}
#endif
#endif /* TZ_MPI_SHIM_H */
