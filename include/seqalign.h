/*
 * seqalign.h — C ABI of the B200-native exhaustive sequence-alignment solver.
 *
 * The problem solved is the one the reference course project solves with
 * MPI + OpenMP + CUDA (reference: main.c:46-244, cudaFunctions.cu:63-176):
 * given weights w1..w4, a long sequence Seq1 and N shorter sequences Seq2_i,
 * find for every Seq2_i the offset n and single-gap ("mutant") position k that
 * maximise  w1*#identical - w2*#conservative - w3*#semi-conservative - w4*#other.
 *
 * Layers exposed here (all `extern "C"`, usable from C, C++, ctypes):
 *   1. parsing / formatting   (reference: main.c:76-108, main.c:199-211)
 *   2. pair-value table        (reference: main.c:14-44, main.c:59-60)
 *   3. CPU oracle              (race-free restatement of cudaFunctions.cu:63-176)
 *   4. planner + packer        (reference: main.c:110-121 work split; cudaFunctions.cu:35-61 uploads)
 *   5. GPU executor            (reference: cudaFunctions.cu:178-242 + kernel :63-176)
 *
 * Everything in 1-4 lives in libseqalign_core.so (no CUDA dependency);
 * 5 lives in libseqalign_cuda.so (sm_100a only).
 */
#ifndef SEQALIGN_H_
#define SEQALIGN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ limits */

/* Buffer caps of the reference (myProto.h:3-4). Inputs inside these caps are
 * guaranteed to take the GPU path; the parser itself accepts longer tokens. */
#define SA_REF_MAX_SEQ1 2999
#define SA_REF_MAX_SEQ2 1999

/* Hard limits of the GPU executor (longer inputs fall back to the CPU oracle). */
#define SA_GPU_MAX_SEQ1 (1u << 20)
#define SA_GPU_MAX_SEQ2 8192u

/* Alphabet: code 0 = padding (scores 0 against everything), 1..26 = 'A'..'Z',
 * 27..31 = up to five distinct non-letter bytes seen in the input. */
#define SA_ALPHABET 32

/* ------------------------------------------------------------- error codes */

enum sa_status {
    SA_OK = 0,
    SA_ERR_PARSE = 1,       /* malformed input (missing tokens, bad integers) */
    SA_ERR_ARG = 2,         /* bad argument */
    SA_ERR_UNSUPPORTED = 3, /* input outside the GPU executor's limits */
    SA_ERR_CUDA = 4,        /* a CUDA runtime call failed (see sa_last_error) */
    SA_ERR_NOMEM = 5,
    SA_ERR_TOO_LARGE = 6,   /* exceeds a fixed-capacity entry point's maximum */
    SA_ERR_SMALL_BUFFER = 7 /* caller buffer too small; the needed size is returned */
};

/* Thread-local human-readable description of the last failure. */
const char* sa_last_error(void);
/* Used by the sibling libraries to report through the same channel. */
void sa_set_last_error(const char* msg);

/* ----------------------------------------------------------------- problem */

typedef struct sa_problem sa_problem;

/* Parse the reference's stdin grammar (main.c:76-108): whitespace-separated
 * tokens `w1 w2 w3 w4  Seq1  N  Seq2_0 .. Seq2_{N-1}`; ASCII a-z are
 * upper-cased; CRLF, repeated blanks and a missing final newline are fine.
 * Missing tokens are an error (the reference reads garbage there). */
int sa_parse(const char* text, size_t len, sa_problem** out);

/* Build a problem from already-tokenised parts (sequences are upper-cased). */
int sa_problem_create(const int32_t weights[4], const char* seq1, size_t len1,
                      int32_t nseq, const char* const* seq2, const size_t* len2,
                      sa_problem** out);
void sa_problem_free(sa_problem* p);

int32_t sa_problem_nseq(const sa_problem* p);
void sa_problem_weights(const sa_problem* p, int32_t w[4]);
size_t sa_problem_len1(const sa_problem* p);
const char* sa_problem_seq1(const sa_problem* p);
size_t sa_problem_len2(const sa_problem* p, int32_t i);
const char* sa_problem_seq2(const sa_problem* p, int32_t i);
/* Number of (n,k) candidates of sequence i: (len1-len2)*len2, 1 if the lengths
 * are equal, 0 if len2 > len1 (cudaFunctions.cu:74,116-118). */
int64_t sa_problem_candidates(const sa_problem* p, int32_t i);
int64_t sa_problem_total_candidates(const sa_problem* p);
/* Serialise back to the input grammar (used by the synthetic generator). */
size_t sa_problem_serialize(const sa_problem* p, char* buf, size_t cap);

/* --------------------------------------------------------------- pair table */

/* class(a,b) for two raw (upper-cased) bytes: 0 identical, 1 same type-1
 * group, 2 same type-2 group, 3 otherwise (main.c:59-60; cudaFunctions.cu:134-141). */
int sa_pair_class(unsigned char a, unsigned char b);

/* ------------------------------------------------------------------- oracle */

/* Result triple per sequence: out[3*i+0] = score, +1 = n, +2 = k.
 * `sa_solve_cpu_naive` is the literal O(len1*len2^2) restatement of the
 * reference kernel's intended semantics; `sa_solve_cpu_fast` is an independent
 * O(len1*len2) formulation; `threads` <= 0 means hardware concurrency. */
int sa_solve_cpu_naive(const sa_problem* p, int32_t* out, int threads);
int sa_solve_cpu_fast(const sa_problem* p, int32_t* out, int threads);

/* ---------------------------------------------------------------- formatter */

/* `#i: score: S, n: N, k: K\n` per sequence (main.c:204,209). Returns the number
 * of bytes needed (excluding NUL); writes at most cap bytes. */
size_t sa_format(const int32_t* results, int32_t nseq, char* buf, size_t cap);

/* ------------------------------------------------------------ planner/packer */

/* Static geometry shared by the packer and the kernel. */
#define SA_CTA_THREADS 512        /* 16 warps per CTA                          */
#define SA_CTA_WARPS 16
#define SA_DIAG_PER_LANE 8        /* R: consecutive offsets per thread          */
#define SA_DIAG_PER_WARP 256      /* 32 * R                                     */
#define SA_ROW_UNROLL 8           /* U: rows per inner-loop trip                */
#define SA_BLOB_MAGIC 0x31424153u /* "SAB1" */
#define SA_BLOB_VERSION 3u
#define SA_PLAN_SMS 148u          /* SMs per GPU the planner balances for (B200)       */
#define SA_SEQ1_TAIL_PAD 4608u    /* zero bytes after Seq1 so every window is in range */

#define SA_MAX_WORLD 16

/* Execution modes of a packed problem (sa_blob_header.flags bit 0).
 *  TILE: item = (sequence, offset tile), int32 lanes, pair values looked up in a 32x32 table;
 *        any input, any number of GPUs, best latency for small batches.
 *  PAIR: item = two sequences of similar length scored together as u16x2 lanes of one 32-bit
 *        register against a per-pair profile in shared memory; the CTA owns both sequences end
 *        to end (no atomics). Throughput mode for large single-GPU batches.
 *  STREAM: as PAIR (u16x2 lanes), but one WARP owns a (pair, 256-offset tile) and walks all rows,
 *        rebuilding a private 32-row profile block as it goes: no block barriers, no row-chunk
 *        merge; per-sequence keys are combined with atomics as in TILE. Preferred by AUTO when
 *        rows * value_range fits 16 bits. */
#define SA_MODE_AUTO 0
#define SA_MODE_TILE 1
#define SA_MODE_PAIR 2
#define SA_MODE_STREAM 3
/* OR into the mode: deal TILE items to the GPUs tile by tile even when sequences are plentiful, so that
 * one sequence's offset range is split over GPUs and reduced with the cross-GPU argmax (also SA_SPLIT_TILES=1). */
#define SA_MODE_SPLIT_TILES 0x100
#define SA_FLAG_PAIR_MODE 1u
#define SA_FLAG_STREAM_MODE 2u
#define SA_STREAM_MIN_ITEMS_PER_GPU 16384u /* AUTO uses STREAM only with several items per resident warp */
#define SA_STREAM_ROW_BLOCK 32u
#define SA_STREAM_MAX_SEQ1 (24u << 10)  /* padded Seq1 bytes that may stay resident in shared memory */
#define SA_NO_SEQ 0xFFFFFFFFu
#define SA_PAIR_MIN_SEQS 1024u          /* AUTO switches to PAIR from this many scorable sequences */
#define SA_PAIR_FIXED_SMEM (44u << 10)  /* table + merge arrays + bookkeeping */
#define SA_PAIR_MAX_SMEM (227u << 10)

typedef struct sa_blob_header {
    uint32_t magic, version, total_bytes, flags;
    int32_t weights[4];
    uint32_t len1, len1_padded, nseq, nitems;
    uint32_t off_lut;    /* int32[32*32]: value(seq2 code a, seq1 code b) at a*32+b */
    uint32_t off_seq1;   /* uint8[len1_padded]: code*4 (pre-scaled LUT byte offset) */
    uint32_t off_seqtab; /* sa_seq_desc[nseq]                                       */
    uint32_t off_items;  /* sa_item[nitems], grouped by rank, costliest first       */
    uint32_t off_seq2;   /* uint8 codes, grouped by rank, each sequence 16-byte aligned, zero padded */
    uint32_t max_rows_padded;  /* largest padded row count of any sequence          */
    uint32_t max_window_bytes; /* largest Seq1 window any item stages (16-B multiple) */
    uint32_t chunk_bytes; /* fabric staging granularity (one ready flag per chunk)  */
    uint32_t nchunks;
    uint32_t pad0;
    uint64_t total_cells;  /* sum over sequences of candidates                 */
    uint64_t padded_cells; /* cells the tiling actually evaluates              */
    uint32_t world;        /* number of GPUs the item list is partitioned for  */
    uint32_t rank_first_item[SA_MAX_WORLD + 1]; /* rank r owns items [r], [r+1]) */
    int32_t value_min;     /* min over {0, w1, -w2, -w3, -w4}: bias of the u16 pair lanes */
    uint32_t value_range;  /* max - min of the same set                               */
    uint32_t off_pairs;    /* STREAM: sa_pair_rec[npairs], grouped by rank            */
    uint32_t npairs;
    uint32_t off_tile_items; /* TILE: sa_tile_item[nitems], same order as the item list */
    /* ---- what each rank of a multi-GPU solve needs (the analogue of MPI_Bcast + MPI_Scatter,
     * main.c:149-152,174): everyone the shared part [0, off_shared_end) = header + table + Seq1,
     * rank r its slice of the item list, of the pair / tile-item records, and the letters of the
     * sequences it scores, [rank_seq2_lo[r], rank_seq2_hi[r]). */
    uint32_t off_shared_end;
    uint32_t rank_first_pair[SA_MAX_WORLD + 1];
    uint32_t rank_seq2_lo[SA_MAX_WORLD];
    uint32_t rank_seq2_hi[SA_MAX_WORLD];
    uint32_t reserved[157];
} sa_blob_header; /* 1024 bytes */

typedef struct sa_seq_desc {
    uint32_t off;      /* byte offset of the sequence's codes inside the blob  */
    uint16_t len;      /* len2 (<= SA_GPU_MAX_SEQ2)                            */
    uint16_t crow;     /* rows per warp chunk (multiple of SA_ROW_UNROLL)      */
    int32_t ndiag;     /* len1-len2 ; 0 = equal length ; <0 = no candidates    */
    uint16_t ntiles;   /* items of this sequence                               */
    uint8_t wd_log2;   /* log2(warps laid across offsets); rows get 16>>wd_log2 */
    uint8_t home;      /* rank holding this sequence's key + arrival counter   */
} sa_seq_desc; /* 16 bytes */

/* TILE mode: everything a CTA needs to start an item in ONE 32-byte record (the item list and the
 * sequence table stay, for tools and for the other modes): one dependent read of cold memory
 * instead of two before the first TMA copy can be issued. */
typedef struct sa_tile_item {
    uint32_t seq, tile;
    uint32_t off;     /* = seqtab[seq].off   */
    int32_t ndiag;    /* = seqtab[seq].ndiag */
    uint16_t len, crow, ntiles;
    uint8_t wd_log2, home;
    uint8_t narrow;   /* a lane owns SA_DIAG_PER_LANE >> narrow offsets: tile = (256 >> narrow) << wd_log2 offsets.
                         Non-zero only when the whole problem is fewer items than SMs (latency mode): the longest
                         items are cut into narrower tiles so that more SMs share them. */
    uint8_t pad8[3];
    uint32_t pad;
} sa_tile_item; /* 32 bytes */
#define SA_MAX_NARROW 2

/* STREAM mode: one self-contained record per pair of sequences scored as u16x2 lanes (what
 * sa_tile_item is for TILE mode): a warp needs the item and this record, nothing else. */
typedef struct sa_pair_rec {
    uint32_t a, b;         /* sequence indices (= result rows); b == SA_NO_SEQ: unpaired   */
    uint32_t off_a, off_b; /* blob offsets of the two sequences' codes                     */
    uint16_t len_a, len_b;
    uint16_t crow;         /* rows every tile of the pair walks (multiple of 32)           */
    uint16_t ntiles;
    uint8_t home;          /* rank holding the pair's keys + arrival counter               */
    uint8_t pad8[3];
    uint32_t pad;
} sa_pair_rec; /* 32 bytes */

typedef struct sa_item {
    uint32_t seq;  /* TILE: sequence          PAIR: first sequence     STREAM: pair index    */
    uint32_t tile; /* TILE: offset tile index PAIR: second / SA_NO_SEQ STREAM: offset tile   */
} sa_item;

/* Plan (choose per-sequence tiling, order items by cost) and pack everything
 * the GPU needs into ONE contiguous blob (one H2D copy / one fabric broadcast).
 * `*blob` is malloc'ed; free with sa_blob_free. Returns SA_ERR_UNSUPPORTED when
 * the input is outside the GPU executor's limits. */
int sa_plan_pack(const sa_problem* p, int world, void** blob, size_t* blob_bytes);
/* Same, into caller memory (e.g. pinned): returns needed size in *blob_bytes
 * (call with dst == NULL to size the buffer). */
int sa_plan_pack_into(const sa_problem* p, int world, void* dst, size_t cap, size_t* blob_bytes);
/* As sa_plan_pack_into with an explicit SA_MODE_*. */
int sa_plan_pack_mode(const sa_problem* p, int world, int mode, void* dst, size_t cap, size_t* blob_bytes);
/* The same for sequences [first, first + count) only: row i of that launch's result table is sequence
 * first + i. sa_session_solve packs a large batch range by range so that packing overlaps the copy and
 * the kernel of the previous range (the reference prepares all rows, then scores them one by one:
 * main.c:96-121, 178-183). */
int sa_plan_pack_range(const sa_problem* p, uint32_t first, uint32_t count, int world, int mode, void* dst, size_t cap,
                       size_t* blob_bytes);
void sa_blob_free(void* blob);

/* CPU executor of a packed blob: walks the item list with exactly the tile /
 * row-chunk / combine / packed-key / first-k logic of the GPU kernel. Used to
 * validate planner+packer without a GPU (out: int32[4*nseq] = score,n,k,0). */
int sa_blob_execute_cpu(const void* blob, size_t blob_bytes, int32_t* out4);
/* The two halves of the walk, for the multi-process CPU configuration (BASELINE.json
 * configs[0]): each rank scores its own item range into packed keys (keys[nseq], max-merged,
 * start from zero), the keys are max-reduced across ranks, then rank 0 finalises. */
int sa_blob_score_items_cpu(const void* blob, size_t blob_bytes, uint32_t first_item,
                            uint32_t last_item, uint64_t* keys);
int sa_blob_finalize_cpu(const void* blob, size_t blob_bytes, const uint64_t* keys, int32_t* out4);
/* The byte ranges of a blob that rank `rank` of a multi-GPU solve reads: [0] the shared part
 * (header + table + Seq1: the MPI_Bcast of main.c:149-152), [1] its slice of the item list, [2] of the
 * tile-item / pair records, [3] the letters of the sequences it scores (the MPI_Scatter of main.c:174).
 * The fabric kernel pulls exactly these (rounded out to chunks); the CPU world scatters exactly these. */
int sa_blob_rank_ranges(const sa_blob_header* h, uint32_t rank, uint32_t lo[4], uint32_t hi[4]);

/* The work partition over `world` GPUs (new-build analogue of main.c:110-121) is part
 * of the plan: whole sequences are dealt to the least-loaded rank, costliest first;
 * when there are fewer than 2*world sequences the deal happens per tile instead, so a
 * single long sequence still spreads over all GPUs. */

/* ------------------------------------------------------------- GPU executor */

/* Bytes of device scratch needed for a blob with `nseq` sequences
 * (packed argmax keys, per-sequence arrival counters, control words). */
size_t sa_cuda_scratch_bytes(uint32_t nseq);

/* Zero a freshly allocated scratch area (the kernel leaves it zeroed again). */
int sa_cuda_scratch_init(void* d_scratch, size_t bytes, void* stream);

/* Solve a packed blob that already sits in device memory.
 *   d_blob     device pointer to the blob (16-byte aligned)
 *   header     HOST copy of the blob's header (sizeof(sa_blob_header) bytes) (sizes the launch; no D2H needed)
 *   d_out      device int32[4*nseq]: score, n, k, 0 per sequence
 *   d_scratch  device scratch of sa_cuda_scratch_bytes(nseq), zeroed once with
 *              sa_cuda_scratch_init (the kernel leaves it zeroed again)
 *   stream     cudaStream_t (NULL = default stream)
 * One kernel launch on the current device, no allocation, no synchronisation.
 * Replaces cudaFunctions.cu:178-242 (per-row malloc+launch+sync) and :63-176. */
int sa_cuda_solve_blob(const void* d_blob, const sa_blob_header* header, int32_t* d_out,
                       void* d_scratch, void* stream);

/* Convenience: plan+pack `p`, upload, solve, download (allocates; synchronous).
 * out: int32[3*nseq]. */
int sa_cuda_solve_problem(const sa_problem* p, int32_t* out, int device);

/* Largest problem (in sequences) the allocation-free entry point below accepts. */
#define SA_NOALLOC_MAX_SEQ 4096u

/* Same solve as sa_cuda_solve_blob, for callers that cannot allocate device memory.
 * (No counterpart in the reference.)
 *
 * Meant for code that runs where HBM already belongs to someone else's pool — a plug-in or
 * custom op executing inside a server or framework that owns the allocator — or where the
 * device-wide stall of cudaMalloc/cudaFree cannot be afforded. You pass the blob and the
 * result table as usual; the solver's scratch is a 49,408-byte array compiled into
 * libseqalign_cuda.so (each GPU that loads the library carries one copy until exit).
 *
 *  - header->nseq must be <= SA_NOALLOC_MAX_SEQ, otherwise SA_ERR_TOO_LARGE is returned and
 *    nothing is launched.
 *  - Makes no cudaMalloc / cudaFree / allocating library call. It enqueues one small memset
 *    and one kernel on `stream`, on the current device.
 *  - The first call on each device looks up the array's device address
 *    (cudaGetSymbolAddress); make that first call outside any stream capture.
 *  - All calls on a device share the one array, and it keeps the previous call's contents;
 *    every call therefore clears the part it uses first. Two calls whose GPU work overlaps
 *    (different streams, no dependency) race on it — serialise them, or give the second one
 *    its own scratch through sa_cuda_solve_blob. */
int sa_cuda_solve_blob_noalloc(const void* d_blob, const sa_blob_header* header, int32_t* d_out,
                               void* stream);

/* Solver session: owns a stream, pinned staging and device buffers on one GPU, all reused and
 * only ever grown. sa_session_solve is the complete per-problem host path in one call (plan +
 * pack into pinned memory, H2D, one launch, D2H, stream synchronise) — what
 * send_divided_Seq2_To_Cuda (cudaFunctions.cu:178-242) does with 4+N cudaMallocs, N device
 * synchronisations and 5 cudaFrees per call. out3: int32[3*nseq]. Not thread-safe per session. */
typedef struct sa_session sa_session;
int sa_session_create(int device, sa_session** out);
void sa_session_destroy(sa_session* s);
int sa_session_solve(sa_session* s, const sa_problem* p, int32_t* out3);
/* Transfer policy. Blobs of at most zero_copy_max_bytes are read by the kernel in place from
 * pinned host memory (no DMA copy; 0 = always copy, the default — measured on B200 the in-place
 * read is 1.5-2.5x slower end to end because every CTA re-reads the table over PCIe); with direct_out != 0
 * (default) result rows are posted by the kernel straight into pinned host memory instead of
 * being copied back afterwards. Either way every input byte crosses PCIe once per solve. */
void sa_session_configure(sa_session* s, size_t zero_copy_max_bytes, int direct_out);
/* Bytes moved over PCIe by the last solve (for benchmark reporting). */
void sa_session_last_bytes(const sa_session* s, uint64_t* h2d, uint64_t* d2h);
/* Batches of at least min_seq sequences (default 16384; 0 = never) are packed, copied and launched in
 * ranges of min_seq/2 or more sequences, so that packing range k+1 on the host overlaps the copy and the
 * kernel of range k; results are identical. sa_session_last_launches: kernel launches of the last solve. */
void sa_session_set_pipeline(sa_session* s, size_t min_seq);
uint32_t sa_session_last_launches(const sa_session* s);

/* ------------------------------------------------------------------- fabric */

/* Multi-GPU execution: one process per GPU, every process launches ONE kernel per solve.
 * Replaces the reference's MPI data path (main.c:149-152 Bcast x4, :174 Scatter, :195-197
 * Gather x3): rank 0's kernel broadcasts the packed blob into every peer's buffer (NVLS
 * multicast store, or per-peer P2P stores) while it already scores; peers score as soon as
 * the chunks they need have landed; per-sequence argmax keys are reduced with system-scope
 * atomics in the home rank's memory; results are written straight into rank 0's table and the
 * last CTA of each rank signals rank 0. The host only exchanges buffer mappings once
 * (bootstrap) — no MPI/NCCL call on the data path. */
#define SA_FABRIC_MAX_CHUNKS 16384u

typedef struct sa_fabric_layout_t {
    uint32_t off_ctrl, off_hdr_ll, off_key, off_cnt, off_out, off_out_ll, off_blob;
    uint32_t max_seq;
    uint64_t max_blob_bytes;
    uint64_t total_bytes; /* size of the symmetric buffer every rank must allocate */
} sa_fabric_layout_t;

/* Offsets of the regions inside each rank's symmetric buffer. */
int sa_fabric_layout(uint32_t max_seq, size_t max_blob_bytes, sa_fabric_layout_t* out);

typedef struct sa_fabric_desc {
    uint32_t rank, world;
    uint32_t epoch;         /* 1, 2, 3, ... — the same on every rank for the same solve      */
    uint32_t use_multicast; /* broadcast with multimem.st through mc_base (else P2P stores)  */
    uint32_t timeout_ms;    /* bound on every cross-GPU wait (0 = 10 s)                       */
    uint32_t pad0;          /* bit 0: record a globaltimer timeline in control words 16..31 (debug);
                               bit 1: re-solve the blob of the previous launch without re-broadcasting it
                               (pass the previous epoch again) */
    void* peer_base[SA_MAX_WORLD]; /* rank r's symmetric buffer as mapped in THIS process     */
    void* mc_base;                 /* multicast mapping of the buffer, or NULL                */
    sa_fabric_layout_t layout;
} sa_fabric_desc;

/* Launch this rank's share of a solve on the current device/stream. Rank 0 must have copied the
 * blob (planned with world == f->world) to peer_base[0] + layout.off_blob and passes the host
 * copy of its header; other ranks pass NULL. The symmetric buffers must be zero before the
 * first solve. After the stream has drained, rank 0 finds int32[4*nseq] at
 * peer_base[0] + layout.off_out, and every rank finds a non-zero word at
 * peer_base[rank] + layout.off_ctrl + 12 if one of its waits timed out. */
int sa_cuda_fabric_solve(const sa_fabric_desc* f, const sa_blob_header* header, void* stream);

/* The two fabric primitives in isolation, for the 1 KB - 1 GB sweep against NCCL and the MPI path
 * (BASELINE.json configs[4]). Both use the same symmetric buffer, flags and epoch discipline as
 * sa_cuda_fabric_solve; every rank calls them once per operation.
 *  bcast : rank 0's blob region [0, bytes) is replicated into every peer's blob region (rank 0 pushes).
 *  pull  : the same result, every other rank copying the range out of rank 0's buffer itself (what the
 *          solver does since blob version 3); the caller makes sure rank 0's range is in place first.
 *  argmax: every rank's nkeys packed u64 keys are max-folded into rank 0's blob region (which rank 0
 *          zeroes beforehand); rank 0's launch ends when all ranks have arrived. */
int sa_cuda_fabric_bcast(const sa_fabric_desc* f, uint64_t bytes, int ctas, void* stream);
int sa_cuda_fabric_pull(const sa_fabric_desc* f, uint64_t bytes, int ctas, void* stream);
int sa_cuda_fabric_argmax(const sa_fabric_desc* f, const void* d_keys, uint64_t nkeys, int ctas, void* stream);

/* One worker of a native multi-process solve (`final --gpus N`; the reference's `mpiexec -np N ./final`,
 * makefile:11, with MPI_Init's wire-up, main.c:62-64): allocates this rank's symmetric buffer with the
 * CUDA virtual-memory API, trades POSIX descriptors of the buffers (and of one multicast object when
 * the GPUs support it) with the other workers through the launcher's socket `sock`
 * (csrc/host/native_wire.h), maps them, and runs sa_cuda_fabric_solve. No Python, torch or NCCL.
 * Every worker passes the same world, max_seq, max_blob_bytes, timeout_ms, use_multicast and repeat. */
typedef struct sa_native_args {
    int rank, world, sock, device;
    const void* blob;             /* rank 0: the problem packed for `world` GPUs; others: NULL        */
    const sa_blob_header* header; /* rank 0: host copy of the blob's header; others: NULL             */
    uint32_t max_seq;             /* capacity of the result table (>= nseq)                           */
    size_t max_blob_bytes;        /* capacity of the blob area (>= total_bytes)                       */
    int32_t* results;             /* rank 0: int32[4*nseq] (score, n, k, 0)                           */
    uint32_t timeout_ms;          /* bound on every cross-GPU wait (0 = 10 s)                         */
    int use_multicast;            /* 0 = peer stores only; otherwise multicast when every GPU can     */
    int repeat;                   /* solves in a row (>= 1); kernel_ms is the fastest                 */
    int used_multicast;           /* out */
    float kernel_ms;              /* out: device time of this rank's kernel (CUDA events)             */
    double times[4];              /* out: seconds since entry at: CUDA up, buffers mapped, ready to launch, results on the host */
} sa_native_args;
int sa_native_worker(sa_native_args* a);

/* Number of kernel launches issued by this library in this process. */
uint64_t sa_cuda_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* SEQALIGN_H_ */
