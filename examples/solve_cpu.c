/* Minimal C use of the core library (no CUDA needed):
 *   gcc -Iinclude examples/solve_cpu.c -Lmpi_openmp_cuda_b200/_lib -lseqalign_core -Wl,-rpath,$PWD/mpi_openmp_cuda_b200/_lib -o solve_cpu
 *   ./solve_cpu < data/input1.txt                                                                       */
#include <stdio.h>
#include <stdlib.h>

#include "seqalign.h"

int main(void) {
    size_t cap = 1 << 16, len = 0;
    char* text = malloc(cap);
    for (int c; (c = getchar()) != EOF;) {
        if (len + 1 >= cap) text = realloc(text, cap *= 2);
        text[len++] = (char)c;
    }
    sa_problem* p = NULL;
    if (sa_parse(text, len, &p) != SA_OK) {
        fprintf(stderr, "parse failed: %s\n", sa_last_error());
        return 1;
    }
    const int n = sa_problem_nseq(p);
    int32_t* res = malloc(sizeof(int32_t) * 3 * (size_t)(n > 0 ? n : 1));
    if (sa_solve_cpu_fast(p, res, 0) != SA_OK) {
        fprintf(stderr, "solve failed: %s\n", sa_last_error());
        return 1;
    }
    const size_t need = sa_format(res, n, NULL, 0);
    char* out = malloc(need + 1);
    sa_format(res, n, out, need);
    fwrite(out, 1, need, stdout);
    sa_problem_free(p);
    free(out);
    free(res);
    free(text);
    return 0;
}
