// Self-test of the MPI stand-in: the same collective sequence the reference issues
// (4x Bcast, 1x Scatter, 3x Gather; main.c:149-152,174,195-197) with checkable payloads.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "mpi.h"

int main(int argc, char** argv) {
    MPI_Init(&argc, &argv);
    int rank = -1, size = -1;
    MPI_Comm_rank(MPI_COMM_WORLD, &rank);
    MPI_Comm_size(MPI_COMM_WORLD, &size);
    const int rows = 3, stride = 2000;
    std::vector<char> seq1(3000, 0);
    int weights[16] = {0};
    int my_size = 0, nrows = 0;
    std::vector<char> all;
    if (rank == 0) {
        std::snprintf(seq1.data(), seq1.size(), "HELLOWORLD");
        for (int i = 0; i < 4; ++i) weights[i] = 10 + i;
        my_size = rows * stride;
        nrows = rows;
        all.assign((size_t)size * my_size, 0);
        for (int r = 0; r < size; ++r)
            for (int j = 0; j < rows; ++j) std::snprintf(&all[(size_t)r * my_size + (size_t)j * stride], stride, "R%dROW%d", r, j);
        char line[64] = {0};
        if (std::fgets(line, sizeof(line), stdin)) std::printf("rank0 stdin: %s", line);
    } else {
        char line[64] = {0};
        if (std::fgets(line, sizeof(line), stdin)) std::printf("rank%d unexpectedly read stdin\n", rank);
    }
    MPI_Bcast(seq1.data(), 3000, MPI_CHAR, 0, MPI_COMM_WORLD);
    MPI_Bcast(weights, 4, MPI_INT, 0, MPI_COMM_WORLD);
    MPI_Bcast(&my_size, 1, MPI_INT, 0, MPI_COMM_WORLD);
    MPI_Bcast(&nrows, 1, MPI_INT, 0, MPI_COMM_WORLD);
    std::vector<char> mine((size_t)my_size);
    MPI_Scatter(all.data(), my_size, MPI_CHAR, mine.data(), my_size, MPI_CHAR, 0, MPI_COMM_WORLD);
    int ok = std::strcmp(seq1.data(), "HELLOWORLD") == 0 && weights[3] == 13 && nrows == rows;
    std::vector<int> a(rows), b(rows), c(rows);
    for (int j = 0; j < rows; ++j) {
        char expect[32];
        std::snprintf(expect, sizeof(expect), "R%dROW%d", rank, j);
        ok = ok && std::strcmp(&mine[(size_t)j * stride], expect) == 0;
        a[j] = rank * 100 + j;
        b[j] = -a[j];
        c[j] = ok;
    }
    std::vector<int> ga((size_t)size * rows), gb((size_t)size * rows), gc((size_t)size * rows);
    MPI_Gather(a.data(), rows, MPI_INT, ga.data(), rows, MPI_INT, 0, MPI_COMM_WORLD);
    MPI_Gather(b.data(), rows, MPI_INT, gb.data(), rows, MPI_INT, 0, MPI_COMM_WORLD);
    MPI_Gather(c.data(), rows, MPI_INT, gc.data(), rows, MPI_INT, 0, MPI_COMM_WORLD);
    if (rank == 0) {
        int good = 1;
        for (int r = 0; r < size; ++r)
            for (int j = 0; j < rows; ++j) {
                const size_t k = (size_t)r * rows + j;
                good = good && ga[k] == r * 100 + j && gb[k] == -(r * 100 + j) && gc[k] == 1;
            }
        std::printf("world %d: %s\n", size, good ? "PASS" : "FAIL");
    }
    MPI_Finalize();
    return 0;
}
