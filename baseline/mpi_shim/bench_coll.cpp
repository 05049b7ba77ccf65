// Latency/bandwidth of the MPI stand-in's Bcast and Gather (host memory, like the reference's
// buffers: main.c:149-152,195-197), 1 KB .. 1 GB. Baseline tooling only.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "mpi.h"

int main(int argc, char** argv) {
    MPI_Init(&argc, &argv);
    int rank, size;
    MPI_Comm_rank(MPI_COMM_WORLD, &rank);
    MPI_Comm_size(MPI_COMM_WORLD, &size);
    const size_t max_bytes = argc > 1 ? std::strtoull(argv[1], nullptr, 10) : (1ull << 30);
    std::vector<char> buf(max_bytes, (char)rank), gat(rank == 0 ? max_bytes : 1);
    int token = 0;
    for (size_t bytes = 1024; bytes <= max_bytes; bytes *= 4) {
        double best_b = 1e30, best_g = 1e30;
        const int per = (int)(bytes / (size_t)size / sizeof(int));
        for (int it = 0; it < 5; ++it) {
            MPI_Bcast(&token, 1, MPI_INT, 0, MPI_COMM_WORLD);
            double t0 = MPI_Wtime();
            MPI_Bcast(buf.data(), (int)bytes, MPI_CHAR, 0, MPI_COMM_WORLD);
            double t1 = MPI_Wtime();
            if (t1 - t0 < best_b) best_b = t1 - t0;
            MPI_Bcast(&token, 1, MPI_INT, 0, MPI_COMM_WORLD);
            t0 = MPI_Wtime();
            if (per > 0) MPI_Gather(buf.data(), per, MPI_INT, gat.data(), per, MPI_INT, 0, MPI_COMM_WORLD);
            t1 = MPI_Wtime();
            if (t1 - t0 < best_g) best_g = t1 - t0;
        }
        if (rank == 0)
            std::printf("{\"impl\": \"mpi-standin-host\", \"world\": %d, \"bytes\": %zu, \"bcast_us\": %.2f, \"gather_us\": %.2f}\n",
                        size, bytes, best_b * 1e6, best_g * 1e6);
    }
    MPI_Finalize();
    return 0;
}
