/* Minimal single-host MPI stand-in — BASELINE TOOLING ONLY, not part of the product.
 *
 * The image and the GPU box have no MPI at all (no mpi.h, libmpi, mpicxx, mpiexec), so the
 * reference (which needs exactly these seven entry points, main.c:62-64,149-152,174,195-197,242)
 * could not be built or run. This header + mpi_shim.cpp + mpiexec.cpp provide just that subset
 * over a shared file mapping with a process-shared barrier, so that the UNMODIFIED reference
 * sources compile and run with `mpiexec -np P ./final < input`.
 */
#ifndef SA_MPI_SHIM_H_
#define SA_MPI_SHIM_H_

#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;

#define MPI_COMM_WORLD 0
#define MPI_CHAR 1
#define MPI_INT 4
#define MPI_SUCCESS 0

int MPI_Init(int* argc, char*** argv);
int MPI_Comm_rank(MPI_Comm comm, int* rank);
int MPI_Comm_size(MPI_Comm comm, int* size);
int MPI_Bcast(void* buffer, int count, MPI_Datatype datatype, int root, MPI_Comm comm);
int MPI_Scatter(const void* sendbuf, int sendcount, MPI_Datatype sendtype, void* recvbuf, int recvcount,
                MPI_Datatype recvtype, int root, MPI_Comm comm);
int MPI_Gather(const void* sendbuf, int sendcount, MPI_Datatype sendtype, void* recvbuf, int recvcount,
               MPI_Datatype recvtype, int root, MPI_Comm comm);
int MPI_Finalize(void);
double MPI_Wtime(void);

#ifdef __cplusplus
}
#endif
#endif
