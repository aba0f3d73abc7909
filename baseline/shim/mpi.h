// mpi.h — single-host stand-in for the six MPI calls the reference makes
// (/root/reference/namegensf.cu:362-364, :615, :636, :889).  This image has no
// MPI; ranks are plain processes (RANK / WORLD_SIZE from the launcher) that
// meet in one POSIX shared-memory segment.  Root-0 host-memory semantics are
// kept: Scatter/Gather stage through the segment exactly once per call.
#pragma once
#include <cstddef>

typedef int MPI_Comm;
typedef int MPI_Datatype;
#define MPI_COMM_WORLD 0
#define MPI_FLOAT 4
#define MPI_CHAR 1
#define MPI_DOUBLE 8
#define MPI_MAX_PROCESSOR_NAME 256
#define MPI_SUCCESS 0

int MPI_Init(int *argc, char ***argv);
int MPI_Finalize(void);
int MPI_Get_processor_name(char *name, int *resultlen);
int MPI_Comm_rank(MPI_Comm comm, int *rank);
int MPI_Comm_size(MPI_Comm comm, int *size);
int MPI_Barrier(MPI_Comm comm);
int MPI_Scatter(const void *sendbuf, int sendcount, MPI_Datatype sendtype,
                void *recvbuf, int recvcount, MPI_Datatype recvtype, int root,
                MPI_Comm comm);
int MPI_Gather(const void *sendbuf, int sendcount, MPI_Datatype sendtype,
               void *recvbuf, int recvcount, MPI_Datatype recvtype, int root,
               MPI_Comm comm);
// harness-only helper (not used by the reference): max over ranks
double shim_allreduce_max(double v);
