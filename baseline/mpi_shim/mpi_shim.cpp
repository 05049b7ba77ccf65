// Single-host MPI subset over a shared file mapping (baseline tooling; see mpi.h).
#include "mpi.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "shim_common.h"

namespace {
int g_rank = 0, g_size = 1;
ShimControl* g_ctl = nullptr;
unsigned char* g_data = nullptr;
size_t g_map_bytes = 0;

size_t type_size(MPI_Datatype t) { return t == MPI_INT ? sizeof(int) : 1; }
void barrier() {
    if (g_ctl) pthread_barrier_wait(&g_ctl->barrier);
}
void need(size_t bytes, const char* what) {
    if (g_ctl && bytes > g_ctl->data_bytes) {
        std::fprintf(stderr, "mpi_shim: %s needs %zu bytes, mapping has %zu (set SHIM_BYTES)\n", what, bytes,
                     g_ctl->data_bytes);
        std::abort();
    }
}
}  // namespace

extern "C" {

int MPI_Init(int*, char***) {
    const char* r = std::getenv("SHIM_RANK");
    const char* s = std::getenv("SHIM_SIZE");
    const char* path = std::getenv("SHIM_FILE");
    if (!r || !s || !path) {  // run without the launcher: a world of one
        g_rank = 0;
        g_size = 1;
        return MPI_SUCCESS;
    }
    g_rank = std::atoi(r);
    g_size = std::atoi(s);
    const int fd = open(path, O_RDWR);
    if (fd < 0) {
        std::perror("mpi_shim: open mapping");
        std::abort();
    }
    struct stat st;
    fstat(fd, &st);
    g_map_bytes = (size_t)st.st_size;
    void* m = mmap(nullptr, g_map_bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (m == MAP_FAILED) {
        std::perror("mpi_shim: mmap");
        std::abort();
    }
    g_ctl = static_cast<ShimControl*>(m);
    g_data = static_cast<unsigned char*>(m) + SHIM_DATA_OFFSET;
    barrier();
    return MPI_SUCCESS;
}

int MPI_Comm_rank(MPI_Comm, int* rank) {
    *rank = g_rank;
    return MPI_SUCCESS;
}
int MPI_Comm_size(MPI_Comm, int* size) {
    *size = g_size;
    return MPI_SUCCESS;
}

int MPI_Bcast(void* buffer, int count, MPI_Datatype datatype, int root, MPI_Comm) {
    if (g_size == 1) return MPI_SUCCESS;
    const size_t bytes = (size_t)count * type_size(datatype);
    need(bytes, "MPI_Bcast");
    if (g_rank == root) std::memcpy(g_data, buffer, bytes);
    barrier();
    if (g_rank != root) std::memcpy(buffer, g_data, bytes);
    barrier();
    return MPI_SUCCESS;
}

int MPI_Scatter(const void* sendbuf, int sendcount, MPI_Datatype sendtype, void* recvbuf, int recvcount,
                MPI_Datatype recvtype, int root, MPI_Comm) {
    const size_t sbytes = (size_t)sendcount * type_size(sendtype);
    const size_t rbytes = (size_t)recvcount * type_size(recvtype);
    if (g_size == 1) {
        std::memcpy(recvbuf, sendbuf, rbytes);
        return MPI_SUCCESS;
    }
    need(sbytes * (size_t)g_size, "MPI_Scatter");
    if (g_rank == root) std::memcpy(g_data, sendbuf, sbytes * (size_t)g_size);
    barrier();
    std::memcpy(recvbuf, g_data + sbytes * (size_t)g_rank, rbytes);
    barrier();
    return MPI_SUCCESS;
}

int MPI_Gather(const void* sendbuf, int sendcount, MPI_Datatype sendtype, void* recvbuf, int recvcount,
               MPI_Datatype recvtype, int root, MPI_Comm) {
    const size_t sbytes = (size_t)sendcount * type_size(sendtype);
    const size_t rbytes = (size_t)recvcount * type_size(recvtype);
    if (g_size == 1) {
        std::memcpy(recvbuf, sendbuf, sbytes);
        return MPI_SUCCESS;
    }
    need(sbytes * (size_t)g_size, "MPI_Gather");
    std::memcpy(g_data + sbytes * (size_t)g_rank, sendbuf, sbytes);
    barrier();
    if (g_rank == root) std::memcpy(recvbuf, g_data, rbytes * (size_t)g_size);
    barrier();
    return MPI_SUCCESS;
}

int MPI_Finalize(void) {
    barrier();
    if (g_ctl) munmap(g_ctl, g_map_bytes);
    g_ctl = nullptr;
    return MPI_SUCCESS;
}

double MPI_Wtime(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

}  // extern "C"
