// Stand-in for <mpi.h>: the handful of MPI calls the reference's fpcode makes, implemented in
// mpi_shim.cpp over POSIX shared memory between the torchrun-launched processes of one node.
// Baseline use only.
#ifndef CME213B_MPI_SHIM_
#define CME213B_MPI_SHIM_
#ifdef __cplusplus
extern "C" {
#endif
typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
#define MPI_COMM_WORLD 0
#define MPI_SUCCESS 0
#define MPI_INT 1
#define MPI_DOUBLE 2
#define MPI_SUM 1
int MPI_Init(int* argc, char*** argv);
int MPI_Finalize(void);
int MPI_Comm_size(MPI_Comm comm, int* size);
int MPI_Comm_rank(MPI_Comm comm, int* rank);
int MPI_Barrier(MPI_Comm comm);
int MPI_Bcast(void* buf, int count, MPI_Datatype type, int root, MPI_Comm comm);
int MPI_Scatter(const void* sendbuf, int sendcount, MPI_Datatype sendtype, void* recvbuf, int recvcount,
                MPI_Datatype recvtype, int root, MPI_Comm comm);
int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm);
#ifdef __cplusplus
}
#endif
#endif
