/* Minimal MPI-1 subset over Unix socketpairs (no MPI exists on the build image or the GPU box).
 * Enough to (a) build yusupov1alik/MPI-Jordan-crazy-acceleration's main.cpp UNMODIFIED — the 16
 * calls listed in SURVEY.md Appendix D, including the MPI-1-only MPI_Address / MPI_Type_struct and
 * a rank-ordered, non-commutative user-op MPI_Allreduce — and (b) bootstrap the ranks of bin/jordan.
 * Launch with bin/jb_mpirun -np P prog args...; without the launcher a program runs as 1 rank. */
#ifndef JB_MPI_SHIM_H
#define JB_MPI_SHIM_H
#include <cstdio> /* the reference gets printf/fopen/fscanf through mpi.h */
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef long MPI_Aint;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR, count_bytes; } MPI_Status;
typedef void(MPI_User_function)(void* invec, void* inoutvec, int* len, MPI_Datatype* datatype);

#define MPI_COMM_WORLD 0
#define MPI_SUCCESS 0
#define MPI_ANY_TAG (-1)
#define MPI_STATUS_IGNORE ((MPI_Status*)0)

#define MPI_BYTE 1
#define MPI_CHAR 2
#define MPI_INT 3
#define MPI_UNSIGNED 4
#define MPI_LONG_LONG 5
#define MPI_DOUBLE 6
#define MPI_FLOAT 7

#define MPI_SUM 1
#define MPI_MAX 2
#define MPI_MIN 3

int MPI_Init(int* argc, char*** argv);
int MPI_Finalize(void);
int MPI_Abort(MPI_Comm comm, int code);
int MPI_Comm_size(MPI_Comm comm, int* size);
int MPI_Comm_rank(MPI_Comm comm, int* rank);
double MPI_Wtime(void);
int MPI_Send(const void* buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void* buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm,
             MPI_Status* status);
int MPI_Sendrecv_replace(void* buf, int count, MPI_Datatype type, int dest, int sendtag, int source,
                         int recvtag, MPI_Comm comm, MPI_Status* status);
int MPI_Bcast(void* buf, int count, MPI_Datatype type, int root, MPI_Comm comm);
int MPI_Barrier(MPI_Comm comm);
int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op,
                  MPI_Comm comm);
int MPI_Allgather(const void* sendbuf, int sendcount, MPI_Datatype sendtype, void* recvbuf,
                  int recvcount, MPI_Datatype recvtype, MPI_Comm comm);
int MPI_Gather(const void* sendbuf, int sendcount, MPI_Datatype sendtype, void* recvbuf,
               int recvcount, MPI_Datatype recvtype, int root, MPI_Comm comm);
/* MPI-1 names removed in MPI-3.0 but used by the reference (main.cpp:1008-1020) */
int MPI_Address(void* location, MPI_Aint* address);
int MPI_Get_address(const void* location, MPI_Aint* address);
int MPI_Type_struct(int count, int* blocklens, MPI_Aint* displs, MPI_Datatype* types,
                    MPI_Datatype* newtype);
int MPI_Type_commit(MPI_Datatype* type);
int MPI_Type_free(MPI_Datatype* type);
int MPI_Op_create(MPI_User_function* fn, int commute, MPI_Op* op);
int MPI_Op_free(MPI_Op* op);

/* Extension of the shim (no MPI counterpart): hand an open file descriptor to another rank on the
 * same host (SCM_RIGHTS over the pair's socket).  Returns 0 / the received descriptor, -1 on error.
 * A build against a real MPI has no such channel: test JB_MPI_SHIM_HAS_FD_PASSING. */
#define JB_MPI_SHIM_HAS_FD_PASSING 1
int jb_mpi_send_fd(int fd, int dest);
int jb_mpi_recv_fd(int src);

#ifdef __cplusplus
}
#endif
#endif
