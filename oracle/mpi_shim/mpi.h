/*
 * mpi.h -- thread-backed stand-in for the handful of MPI calls the reference's
 * hot path makes (src/MpiNaiveSimulation.cpp:21-23, 33-35, 112-116 and the
 * MPI_Finalize calls in src/CheckpointManager.cpp).  TEST INFRASTRUCTURE: it
 * exists so that MpiNaiveSimulation.cpp can be compiled UNMODIFIED (-DUSE_MPI)
 * in a container with no MPI toolchain, with each "rank" a host thread.
 * Implemented in oracle/ref_shim.cpp.  Not an MPI implementation.
 */
#pragma once
#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef struct { int unused; } MPI_Status;

#define MPI_COMM_WORLD 0
#define MPI_INT 1
#define MPI_BYTE 2
#define MPI_FLOAT 3
#define MPI_SUCCESS 0
#define MPI_STATUS_IGNORE ((MPI_Status *)0)

int MPI_Comm_rank(MPI_Comm comm, int *rank);
int MPI_Comm_size(MPI_Comm comm, int *size);
int MPI_Allgather(const void *sendbuf, int sendcount, MPI_Datatype sendtype,
                  void *recvbuf, int recvcount, MPI_Datatype recvtype, MPI_Comm comm);
int MPI_Sendrecv(const void *sendbuf, int sendcount, MPI_Datatype sendtype, int dest, int sendtag,
                 void *recvbuf, int recvcount, MPI_Datatype recvtype, int source, int recvtag,
                 MPI_Comm comm, MPI_Status *status);
int MPI_Finalize(void);

#ifdef __cplusplus
}
#endif
