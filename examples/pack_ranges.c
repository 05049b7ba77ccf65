/* The planner / packer layer on its own (no CUDA needed): the problem is cut into ranges of sequences, every
 * range is planned and packed into its own blob (sa_plan_pack_range — what sa_session_solve does for large
 * batches so that packing overlaps the copy and the kernel of the previous range), and each blob is walked on
 * the CPU with the kernels' own tiling (sa_blob_execute_cpu). Row i of a range's result is sequence first + i.
 *   gcc -std=c99 -Iinclude examples/pack_ranges.c -Lmpi_openmp_cuda_b200/_lib -lseqalign_core \
 *       -Wl,-rpath,$PWD/mpi_openmp_cuda_b200/_lib -o pack_ranges
 *   ./pack_ranges 3 < data/input3.txt          (3 ranges; same bytes on stdout as ./final)             */
#include <stdio.h>
#include <stdlib.h>

#include "seqalign.h"

int main(int argc, char** argv) {
    const int nranges = argc > 1 && atoi(argv[1]) > 0 ? atoi(argv[1]) : 2;
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
    const uint32_t n = (uint32_t)sa_problem_nseq(p);
    int32_t* res3 = malloc(sizeof(int32_t) * 3 * (size_t)(n ? n : 1));
    for (int k = 0; k < nranges; ++k) {
        const uint32_t first = (uint32_t)((uint64_t)n * (uint64_t)k / (uint64_t)nranges);
        const uint32_t count = (uint32_t)((uint64_t)n * (uint64_t)(k + 1) / (uint64_t)nranges) - first;
        size_t need = 0, got = 0;
        if (sa_plan_pack_range(p, first, count, 1, SA_MODE_AUTO, NULL, 0, &need) != SA_OK) {
            fprintf(stderr, "planning range %d failed: %s\n", k, sa_last_error());
            return 1;
        }
        void* blob = malloc(need ? need : 1);
        int32_t* out4 = malloc(sizeof(int32_t) * 4 * (size_t)(count ? count : 1));
        if (sa_plan_pack_range(p, first, count, 1, SA_MODE_AUTO, blob, need, &got) != SA_OK ||
            sa_blob_execute_cpu(blob, got, out4) != SA_OK) {
            fprintf(stderr, "range %d failed: %s\n", k, sa_last_error());
            return 1;
        }
        const sa_blob_header* h = (const sa_blob_header*)blob;
        fprintf(stderr, "range %d: sequences [%u, %u), %zu bytes, %u work items, mode %s\n", k, first, first + count, got, h->nitems,
                (h->flags & SA_FLAG_STREAM_MODE) ? "stream" : (h->flags & SA_FLAG_PAIR_MODE) ? "pair" : "tile");
        for (uint32_t i = 0; i < count; ++i)
            for (int c = 0; c < 3; ++c) res3[3 * (size_t)(first + i) + c] = out4[4 * (size_t)i + c];
        free(out4);
        free(blob);
    }
    const size_t need = sa_format(res3, (int32_t)n, NULL, 0);
    char* out = malloc(need + 1);
    sa_format(res3, (int32_t)n, out, need);
    fwrite(out, 1, need, stdout);
    sa_problem_free(p);
    free(out);
    free(res3);
    free(text);
    return 0;
}
