// Minimal C++ use of the GPU session (needs a B200):
//   g++ -std=c++17 -Iinclude examples/solve_gpu.cpp -Lmpi_openmp_cuda_b200/_lib -lseqalign_cuda -lseqalign_core \
//       -Wl,-rpath,$PWD/mpi_openmp_cuda_b200/_lib -o solve_gpu && ./solve_gpu < data/input3.txt
#include <cstdio>
#include <iostream>
#include <iterator>
#include <string>
#include <vector>

#include "seqalign.h"

int main() {
    const std::string text((std::istreambuf_iterator<char>(std::cin)), std::istreambuf_iterator<char>());
    sa_problem* p = nullptr;
    if (sa_parse(text.data(), text.size(), &p) != SA_OK) {
        std::fprintf(stderr, "parse failed: %s\n", sa_last_error());
        return 1;
    }
    sa_session* s = nullptr;
    if (sa_session_create(0, &s) != SA_OK) {
        std::fprintf(stderr, "no session: %s\n", sa_last_error());
        return 1;
    }
    const int n = sa_problem_nseq(p);
    std::vector<int32_t> res((size_t)n * 3 + 3);
    int rc = sa_session_solve(s, p, res.data());          // plan + pack + H2D + kernel + results, one call
    if (rc == SA_ERR_UNSUPPORTED) rc = sa_solve_cpu_fast(p, res.data(), 0);  // outside the GPU limits
    if (rc != SA_OK) {
        std::fprintf(stderr, "solve failed: %s\n", sa_last_error());
        return 1;
    }
    std::string out(sa_format(res.data(), n, nullptr, 0), '\0');
    sa_format(res.data(), n, out.data(), out.size());
    std::fwrite(out.data(), 1, out.size(), stdout);
    sa_session_destroy(s);
    sa_problem_free(p);
    return 0;
}
