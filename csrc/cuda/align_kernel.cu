// sm_100a scoring kernel + host launch path.
//
// Replaces, in the reference:
//   * cudaFunctions.cu:63-176  `calc_result` — one thread per Seq2 letter, the (n,k) double loop
//     inside the kernel with 3 block barriers and 2 levels of atomics per candidate;
//   * cudaFunctions.cu:178-242 — per-row cudaMalloc + cudaGetDeviceProperties + launch +
//     cudaDeviceSynchronize;
//   * cudaFunctions.cu:9-13,35-61 — five __constant__ symbols filled by separate copies.
//
// Design (not a port): ONE persistent launch per problem. A work item is (sequence,
// offset tile). Each thread owns R=8 consecutive offsets n and walks rows i, keeping the
// running prefixes P_n[k] of 9 neighbouring diagonals and the running maxima of
// P_n[k]-P_{n+1}[k] in registers (score(n,k) = P_{n+1}[m] + P_n[k] - P_{n+1}[k], see plan.cpp).
// The 16 warps of a CTA split an item as wd offset groups x wr row chunks; chunk results
// are merged in row order through shared memory, folded into one packed 64-bit
// (score, ~n) key per tile and atomicMax'ed into the sequence's slot. The CTA that delivers a
// sequence's last tile rescans the winning diagonal for the first k and writes the result.
// Inputs are staged with TMA bulk copies (cp.async.bulk + mbarrier); the pair-value table
// sits in shared memory with one bank per letter, so the per-cell lookup is conflict free.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <climits>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "ptx_helpers.cuh"
#include "seqalign.h"

namespace {

constexpr int kThreads = SA_CTA_THREADS;
constexpr int kWarps = SA_CTA_WARPS;
constexpr int kR = SA_DIAG_PER_LANE;
constexpr int kU = SA_ROW_UNROLL;
constexpr int kNeg = -(1 << 30);
constexpr uint32_t kLutBytes = SA_ALPHABET * SA_ALPHABET * 4;
constexpr uint32_t kSumWords = kWarps * (kR + 1) * 32;
constexpr uint32_t kMaxWords = kWarps * kR * 32;
static_assert(kR == 8 && kU == 8, "the inner loop is written for R = U = 8");

struct KParams {
    const unsigned char* blob;  // packed problem (sa_plan_pack)
    int4* out;                  // (score, n, k, 0) per sequence
    unsigned long long* key;    // packed argmax key per sequence, zero between launches
    unsigned int* cnt;          // tiles delivered per sequence, zero between launches
    unsigned int* ctrl;         // [0] next item, [1] CTAs finished; zero between launches
    uint32_t item_begin, item_end;
    uint32_t s2_cap, win_cap;   // shared-memory capacities (bytes, multiples of 16)
    uint32_t write_empty;       // 1: this launch writes the rows of candidate-free sequences
    uint32_t h_off_lut, h_off_seq1, h_off_seqtab, h_off_items, h_nseq, h_off_tile_items;  // header fields the host already knows (non-fabric)
    // ---- fabric (multi-GPU) launches only; see sa_cuda_fabric_solve
    uint32_t rank, world, epoch, use_mc;
    unsigned char* peer[SA_MAX_WORLD];  // every rank's symmetric buffer as mapped into this process
    unsigned char* mc;                  // NVLS multicast alias of the same buffer, or null
    uint32_t off_ctrl, off_hdr_ll, off_key, off_cnt, off_out, off_out_ll, off_blob;
    unsigned long long timeout_ns;
    uint32_t timeline;  // 1: record the globaltimer timeline in the control block
    uint32_t skip_push; // 1: re-solve of the blob every rank already holds (resident re-solve)
    uint32_t* done_flag;  // single-GPU launches: pinned host word that receives done_epoch once every result row is visible
    uint32_t done_epoch;  // to the host (the session waits on it instead of on the stream: several microseconds sooner)
    uint32_t dbg_src;   // debug (SA_DEBUG_SRC): bit 0 read Seq2 letters, bit 1 read items/records from rank 0's copy instead of the staged one
};

// Words of the control block at the start of every rank's symmetric buffer (all local to the rank
// that owns the buffer: nothing in here is written by another GPU).
constexpr uint32_t kCtlNext = 0, kCtlCtasDone = 1, kCtlSharedNext = 2, kCtlError = 3, kCtlAllHere = 4, kCtlCopied = 5, kCtlFlags = 64;
constexpr uint32_t kMaxChunks = SA_FABRIC_MAX_CHUNKS;
constexpr uint32_t kHdrWords = sizeof(sa_blob_header) / 4;
// Optional timeline (globaltimer ns) in control words 16..63 as 64-bit slots:
// 0 earliest CTA start, 1 header known (latest CTA), 2 last item finished, 4 every result row validated (rank 0),
// 5 all of this rank's chunks staged, 6 items that had to wait for a chunk, 7 ns spent so, 8..10 item-queue quartiles,
// 11..15 latest CTA to reach each phase of its first item, 16 last CTA exit, 17 first copy item done.
constexpr uint32_t kCtlTimes = 16;

using namespace sa_ptx;  // csrc/cuda/ptx_helpers.cuh

__device__ __forceinline__ unsigned long long pack_key(int score, uint32_t n) {
    return ((unsigned long long)((uint32_t)score ^ 0x80000000u) << 32) | (unsigned long long)(0xFFFFFFFFu - n);
}
__device__ __forceinline__ unsigned long long umax64(unsigned long long a, unsigned long long b) {
    return a > b ? a : b;
}
__device__ __forceinline__ uint32_t byte_of(uint32_t lo, uint32_t hi, int k) {  // k in 0..7, compile time
    return ((k < 4 ? lo : hi) >> (8 * (k & 3))) & 0xFFu;
}
// Same, as one PRMT instead of shift + mask: in the u16x2 kernels the ALU pipe is the busiest one
// (measured: STREAM +6 %, PAIR +2 %; the int32 TILE kernel is 7 % slower with it, so it keeps byte_of).
__device__ __forceinline__ uint32_t byte_of_prmt(uint32_t lo, uint32_t hi, int k) {
    return __byte_perm(k < 4 ? lo : hi, 0u, 0x4440u | (uint32_t)(k & 3));
}


// ------------------------------------------------------------------ the scan step ("difference form")
//
// What a lane carries for its R = 8 offsets is NOT the nine running prefixes P[0..8] of the
// neighbouring diagonals but their eight neighbour differences D[r] = P[r] - P[r+1] — the quantity
// whose running maximum is wanted — plus ONE running total (P8, the ninth diagonal). Per cell that
// is one 3-input add (D += v[r] - v[r+1], a single IADD3) and half a 3-input max (two rows per
// VIMNMX3) instead of add + subtract + max: the ALU pipe issues one instruction every other clock
// per SM sub-partition (profiles/MEASURED_PIPES.json), and it was the binding unit of the P-form
// loop. The nine totals are rebuilt once per item from P8 and the final differences (to_prefixes).
// u16x2: both halves are biased to stay non-negative (D starts at BIG per half, values are
// v - vmin >= 0), so the 32-bit adds never carry or borrow between the halves as long as
// rows * value_range <= 32767 — the planner's limit; the CPU walk re-checks every step.
constexpr uint32_t kBig = 32767u, kBig2 = kBig | (kBig << 16);

// Eight rows of one lane: A[k] points at the profile word of row 0 for the letter under diagonal k.
__device__ __forceinline__ void scan8_u16x2(const unsigned char* const (&A)[2 * kU], uint32_t (&D)[kR], uint32_t (&best)[kR],
                                            uint32_t& P8) {
#pragma unroll
    for (int u = 0; u < kU; u += 2) {
        uint32_t v[kR + 1], w[kR + 1];
#pragma unroll
        for (int r = 0; r <= kR; ++r) v[r] = *reinterpret_cast<const uint32_t*>(A[u + r] + u * (SA_ALPHABET * 4));
#pragma unroll
        for (int r = 0; r <= kR; ++r) w[r] = *reinterpret_cast<const uint32_t*>(A[u + 1 + r] + (u + 1) * (SA_ALPHABET * 4));
#pragma unroll
        for (int r = 0; r < kR; ++r) {
            const uint32_t d0 = D[r] + v[r] - v[r + 1];
            D[r] = d0 + w[r] - w[r + 1];
            best[r] = __vimax3_u16x2(best[r], d0, D[r]);
        }
        P8 += v[kR] + w[kR];
    }
}
// P[r] = P[r+1] + D[r] (bias removed): the nine biased prefix totals of the lane's diagonals.
__device__ __forceinline__ void to_prefixes_u16x2(const uint32_t (&D)[kR], uint32_t P8, uint32_t (&P)[kR + 1]) {
    P[kR] = P8;
#pragma unroll
    for (int r = kR - 1; r >= 0; --r) P[r] = P[r + 1] + D[r] - kBig2;
}

// First k in [0, m) whose candidate (n fixed) reaches `bestv` — the reference's strict-'>' update
// keeps the first maximum (cudaFunctions.cu:118,161). Whole CTA; `c1` = Seq1 codes (x4) at the
// winning offset (shared or global), `s2` = the sequence's codes in shared memory.
// scan[kWarps], tmp[kWarps] and *slot are shared scratch; *slot must hold 0x7FFFFFFF on entry.
__device__ __forceinline__ int first_k_on_diagonal(const int* lut_i, const unsigned char* s2, const unsigned char* c1,
                                                   uint32_t m, int bestv, int* scan, int* tmp, uint32_t* slot) {
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t per = (m + kThreads - 1) / kThreads;
    const uint32_t i0 = min(tid * per, m), i1 = min(i0 + per, m);
    int da = 0, sb = 0;
    for (uint32_t i = i0; i < i1; ++i) {
        const int* lrow = lut_i + (uint32_t)s2[i] * SA_ALPHABET;
        const int a = lrow[c1[i] >> 2], b = lrow[c1[i + 1] >> 2];
        da += a - b;
        sb += b;
    }
    // block-wide exclusive scan of da, block sums of da and sb
    int inc = da, tsb = sb;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int x = __shfl_up_sync(0xFFFFFFFFu, inc, o);
        if (lane >= (uint32_t)o) inc += x;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tsb += __shfl_xor_sync(0xFFFFFFFFu, tsb, o);
    if (lane == 31) scan[warp] = inc;
    if (lane == 0) tmp[warp] = tsb;
    __syncthreads();
    int before = 0, tot_da = 0, T1 = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
        const int x = scan[w];
        if ((uint32_t)w < warp) before += x;
        tot_da += x;
        T1 += tmp[w];
    }
    const int T0 = T1 + tot_da;
    if (T0 == bestv) return 0;  // k = 0 is tried first and wins the tie
    int S = before + inc - da;
    for (uint32_t i = i0; i < i1; ++i) {
        const int* lrow = lut_i + (uint32_t)s2[i] * SA_ALPHABET;
        S += lrow[c1[i] >> 2] - lrow[c1[i + 1] >> 2];
        if (i + 1 < m && T1 + S == bestv) {
            atomicMin(slot, i + 1u);
            break;
        }
    }
    __syncthreads();
    return (int)*slot;
}

// ------------------------------------------------------------------ the kernel

// Shared by every fabric launch: set when one of this CTA's bounded cross-GPU waits timed out.
__shared__ uint32_t s_abort;
// Set once this CTA has seen that all of its rank's chunks are staged: no more per-item waits.
__shared__ uint32_t s_all_here;
// The blob header as every CTA of a fabric launch sees it: kernel parameter on rank 0, received
// through the flag-in-data header packet on the others.
__shared__ __align__(16) sa_blob_header s_hdr;

// ---- how the data moves in a multi-GPU solve (replaces main.c:149-152 Bcast x4, :174 Scatter, :195-197 Gather x3)
//
//  * Broadcast of what is shared. The 1 KB header (weights, sizes, the whole partition) travels as a
//    "flag in data" packet: rank 0's first CTA stores, for every 32-bit header word, the 64-bit value
//    (epoch << 32 | word) into EVERY rank's packet area — one multimem.st per word through the
//    NVSwitch (or one store per peer without multicast). 8-byte stores are single-copy atomic, so a
//    receiver that sees the epoch in the upper half has the word: no fence on the sender, no separate
//    flag, one NVLink one-way trip (the packet also says "rank 0's kernel is running", i.e. its H2D
//    copy is complete). The table and Seq1 follow by the pull below.
//  * Scatter. Every rank other than 0 PULLS what it needs from rank 0's buffer into its own: the
//    shared part [0, off_shared_end) and its own slices of the item list, the record table and the
//    Seq2 letters (the header says where they are) — ingress per GPU is its own share, rank 0's SMs
//    do not copy anything. The first items of the rank's work queue are these copies, one 8 KB chunk
//    each (LDG.128 from the peer mapping -> STG.128 local -> gpu fence -> local chunk flag); scoring
//    items follow and wait, per item, only for the chunks they read.
//  * Gather. Whoever finishes a sequence stores its row straight into rank 0's table as two
//    self-validating 8-byte words (score | epoch, n | k | epoch): no fence, no arrival counter. Rank 0's
//    CTAs, once out of items, check every row's tags (bounded) and rewrite the row as
//    (score, n, k, 0); rank 0's kernel completing therefore means the table is complete.
//  * Reduce (a sequence whose tiles were dealt to several ranks): packed 64-bit (score, ~n) keys and
//    the arrival counter live on the sequence's home rank; tiles fold in with atomicMax.sys /
//    atomicAdd.sys over NVLink and the last arriver finishes the sequence.
struct PullPlan {
    uint32_t first[4], count[4], total;
};
__device__ __forceinline__ PullPlan make_pull_plan(const sa_blob_header& h, uint32_t rank) {
    PullPlan pp;
    const bool stream = (h.flags & SA_FLAG_STREAM_MODE) != 0u;
    const uint32_t i0 = h.rank_first_item[rank], i1 = h.rank_first_item[rank + 1];
    uint32_t lo[4], hi[4];
    lo[0] = 0u, hi[0] = h.off_shared_end;
    lo[1] = h.off_items + 8u * i0, hi[1] = h.off_items + 8u * i1;
    if (stream) lo[2] = h.off_pairs + 32u * h.rank_first_pair[rank], hi[2] = h.off_pairs + 32u * h.rank_first_pair[rank + 1];
    else lo[2] = h.off_tile_items + 32u * i0, hi[2] = h.off_tile_items + 32u * i1;
    lo[3] = h.rank_seq2_lo[rank], hi[3] = h.rank_seq2_hi[rank];
    const uint32_t cb = h.chunk_bytes;
    int prev_last = -1;
    pp.total = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        pp.first[i] = 0u, pp.count[i] = 0u;
        if (i1 <= i0 || hi[i] <= lo[i]) continue;  // a rank without items stages nothing
        int f = (int)(lo[i] / cb);
        const int l = (int)((hi[i] - 1u) / cb);
        if (f <= prev_last) f = prev_last + 1;  // the boundary chunk of the previous range is already on the list
        if (l >= f) pp.first[i] = (uint32_t)f, pp.count[i] = (uint32_t)(l - f + 1);
        if (l > prev_last) prev_last = l;
        pp.total += pp.count[i];
    }
    return pp;
}
__device__ __forceinline__ uint32_t pull_chunk_of(const PullPlan& pp, uint32_t j) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (j < pp.count[i]) return pp.first[i] + j;
        j -= pp.count[i];
    }
    return 0u;
}

// Copy item `j` of this rank: stage one chunk of rank 0's blob locally. `t`/`nt`: this thread's index and
// the number of threads cooperating (a CTA in TILE mode, a warp in STREAM mode). The caller synchronises
// the cooperating threads afterwards and lets ONE of them call pull_publish.
__device__ __forceinline__ void pull_copy(const KParams& p, const sa_blob_header& h, uint32_t chunk, uint32_t t, uint32_t nt) {
    if (p.dbg_src & 4u) return;  // debug: pretend-copy (flags and queue behave as usual, no bytes move)
    const uint32_t cb = h.chunk_bytes, base = chunk * cb;
    const uint32_t n16 = (min(cb, h.total_bytes - base) + 15u) >> 4;
    const int4* src = reinterpret_cast<const int4*>(p.peer[0] + p.off_blob + base);
    int4* dst = reinterpret_cast<int4*>(p.peer[p.rank] + p.off_blob + base);
    uint32_t i = t;
    for (; i + 3u * nt < n16; i += 4u * nt) {  // four loads in flight per thread: an NVLink read is ~2 us away
        const int4 a = src[i], b = src[i + nt], c = src[i + 2u * nt], d = src[i + 3u * nt];
        dst[i] = a, dst[i + nt] = b, dst[i + 2u * nt] = c, dst[i + 3u * nt] = d;
    }
    for (; i < n16; i += nt) dst[i] = src[i];
}
__device__ __forceinline__ void pull_publish(const KParams& p, uint32_t* my_ctl, uint32_t chunk, uint32_t n_copy) {
    unsigned long long* const times = reinterpret_cast<unsigned long long*>(my_ctl + kCtlTimes);
    __threadfence();  // the cooperating threads' stores (ordered before this by their barrier) are visible GPU-wide
    st_relaxed_gpu(my_ctl + kCtlFlags + chunk, p.epoch);
    if (p.timeline) atomicMin(&times[17], global_ns());
    if (atomicAdd(&my_ctl[kCtlCopied], 1u) + 1u == n_copy) {  // every pusher fenced before it counted itself
        my_ctl[kCtlCopied] = 0u;
        st_relaxed_gpu(my_ctl + kCtlAllHere, p.epoch);
        if (p.timeline) times[5] = global_ns();
    }
}

// Bounded wait for a LOCAL flag raised by another CTA of this GPU (relaxed polls, one fence by the caller).
__device__ __forceinline__ bool wait_local_flag(const uint32_t* flag, uint32_t epoch, unsigned long long timeout_ns) {
    if ((int)(ld_relaxed_gpu(flag) - epoch) >= 0) return true;
    const unsigned long long t0 = global_ns();
    for (;;) {
        if ((int)(ld_relaxed_gpu(flag) - epoch) >= 0) return true;
        __nanosleep(32);
        if (global_ns() - t0 > timeout_ns) return false;
    }
}

// True when the blob bytes [lo, hi) are staged in this rank's buffer (always, on rank 0).
// Fast path: the cached / single "all here" flag. Slow path: the flags of the covering chunks.
// May be called by ONE thread (TILE: thread 0, the block barrier that follows re-joins the CTA) or by ALL
// lanes of a warp with identical arguments (STREAM). A single lane of a warp must NOT call it on its own:
// the wait loop is a divergent region that ptxas re-joins only at the end of the enclosing item, so the
// warp then walks the whole item twice, once as lane 0 and once as lanes 1..31 (measured: the items that
// had waited took 2x as long; ncu showed 29 active threads per instruction instead of 32).
__device__ __forceinline__ bool fabric_bytes_ready(const KParams& p, uint32_t* my_ctl, uint32_t lo, uint32_t hi, bool leader = true) {
    volatile uint32_t* const cached = &s_all_here;
    if (*cached) return true;
    unsigned long long* const times = reinterpret_cast<unsigned long long*>(my_ctl + kCtlTimes);
    if ((int)(ld_relaxed_gpu(my_ctl + kCtlAllHere) - p.epoch) >= 0) {
        __threadfence();  // acquire: everything staged before the flag is visible from here on
        *cached = 1u;
        return true;
    }
    const unsigned long long t0 = p.timeline ? global_ns() : 0ull;
    const uint32_t cb = s_hdr.chunk_bytes;
    for (uint32_t c = lo / cb; c <= (hi - 1u) / cb; ++c)
        if (!wait_local_flag(my_ctl + kCtlFlags + c, p.epoch, p.timeout_ns)) return false;
    __threadfence();  // one acquire fence for all the chunk flags just observed
    if (p.timeline && leader) {
        atomicAdd(&times[6], 1ull);
        atomicAdd(&times[7], global_ns() - t0);
    }
    return true;
}

// Fabric prologue, once per CTA (requires a preceding __syncthreads after s_abort = 0): rank 0 sends the
// header packet (its first CTA only) and takes the header from its kernel parameter; every other rank
// receives it. On return s_hdr is valid in every CTA, or s_abort is set.
__device__ __forceinline__ void fabric_prologue(const KParams& p, const sa_blob_header& hdr0) {
    const uint32_t tid = threadIdx.x;
    uint32_t* const my_ctl = reinterpret_cast<uint32_t*>(p.peer[p.rank] + p.off_ctrl);
    uint32_t* const dst = reinterpret_cast<uint32_t*>(&s_hdr);
    if (p.rank == 0) {
        const uint32_t* const src = reinterpret_cast<const uint32_t*>(&hdr0);
        if (tid < kHdrWords) {
            const uint32_t w = src[tid];
            dst[tid] = w;
            if (blockIdx.x == 0 && p.world > 1 && !p.skip_push) {
                const unsigned long long pkt = ((unsigned long long)p.epoch << 32) | w;
                if (p.use_mc) {
                    multimem_st_u64(p.mc + p.off_hdr_ll + 8u * tid, pkt);
                } else {
                    for (uint32_t r = 1; r < p.world; ++r)
                        st_relaxed_sys_u64(reinterpret_cast<unsigned long long*>(p.peer[r] + p.off_hdr_ll) + tid, pkt);
                }
            }
        }
    } else {
        const unsigned long long* const pkt = reinterpret_cast<const unsigned long long*>(p.peer[p.rank] + p.off_hdr_ll) + tid;
        if (tid < kHdrWords) {
            unsigned long long v = ld_relaxed_sys_u64(pkt);
            if ((uint32_t)(v >> 32) != p.epoch) {
                const unsigned long long t0 = global_ns();
                for (;;) {
                    v = ld_relaxed_sys_u64(pkt);
                    if ((uint32_t)(v >> 32) == p.epoch) break;
                    __nanosleep(32);
                    if (global_ns() - t0 > p.timeout_ns) {
                        s_abort = 1u;
                        atomicExch(my_ctl + kCtlError, 1u);
                        break;
                    }
                }
            }
            dst[tid] = (uint32_t)v;
        }
    }
    __syncthreads();
    if (p.rank != 0 && s_abort == 0u && (s_hdr.magic != SA_BLOB_MAGIC || s_hdr.world != p.world)) {  // never walk a header that is not one
        if (tid == 0) {
            s_abort = 1u;
            atomicExch(my_ctl + kCtlError, 3u);
        }
        __syncthreads();
    }
    if (tid == 0 && (p.rank == 0 || p.skip_push)) s_all_here = 1u;  // nothing to stage
    __syncthreads();
}

// A finished sequence's row, stored into rank 0's table as two self-validating words.
__device__ __forceinline__ void post_result(const KParams& p, uint32_t seq, int score, uint32_t n, uint32_t k) {
    unsigned long long* const row = reinterpret_cast<unsigned long long*>(p.peer[0] + p.off_out_ll) + 2u * seq;
    st_relaxed_sys_u64(row, ((unsigned long long)p.epoch << 32) | (uint32_t)score);
    st_relaxed_sys_u64(row + 1, ((unsigned long long)(p.epoch & 0x7FFFFFFFu) << 33) | ((unsigned long long)(k & 0x1FFFu) << 20) | (n & 0xFFFFFu));
}

// Rank 0, once a CTA is out of items: validate a share of the result rows (bounded) and rewrite them in
// the host's (score, n, k, 0) format. Rows of candidate-free sequences were written directly.
__device__ __forceinline__ void collect_results(const KParams& p, uint32_t* my_ctl) {
    const sa_seq_desc* const seqtab = reinterpret_cast<const sa_seq_desc*>(p.blob + s_hdr.off_seqtab);
    const unsigned long long* const ll = reinterpret_cast<const unsigned long long*>(p.peer[0] + p.off_out_ll);
    int4* const out = reinterpret_cast<int4*>(p.peer[0] + p.off_out);
    const unsigned long long tag1 = (unsigned long long)(p.epoch & 0x7FFFFFFFu);
    for (uint32_t s = blockIdx.x * kThreads + threadIdx.x; s < s_hdr.nseq; s += gridDim.x * kThreads) {
        if (seqtab[s].ntiles == 0) continue;
        unsigned long long a = ld_relaxed_sys_u64(ll + 2u * s), b = ld_relaxed_sys_u64(ll + 2u * s + 1u);
        if ((uint32_t)(a >> 32) != p.epoch || (b >> 33) != tag1) {
            const unsigned long long t0 = global_ns();
            for (;;) {
                a = ld_relaxed_sys_u64(ll + 2u * s), b = ld_relaxed_sys_u64(ll + 2u * s + 1u);
                if ((uint32_t)(a >> 32) == p.epoch && (b >> 33) == tag1) break;
                __nanosleep(64);
                if (global_ns() - t0 > p.timeout_ns) {
                    atomicExch(my_ctl + kCtlError, 2u);
                    return;
                }
            }
        }
        out[s] = make_int4((int)(uint32_t)a, (int)(b & 0xFFFFFu), (int)((b >> 20) & 0x1FFFu), 0);
    }
}

// Launch epilogue, once per CTA after its last item: the last CTA of the launch re-arms the work queue.
template <bool kFabric>
__device__ __forceinline__ void launch_epilogue(const KParams& p) {
    const uint32_t tid = threadIdx.x;
    uint32_t* const my_ctl = kFabric ? reinterpret_cast<uint32_t*>(p.peer[p.rank] + p.off_ctrl) : p.ctrl;
    unsigned long long* const times = reinterpret_cast<unsigned long long*>(my_ctl + kCtlTimes);
    if (kFabric && p.rank == 0 && s_abort == 0u) {
        collect_results(p, my_ctl);
        if (p.timeline) {
            __syncthreads();  // every thread of this CTA has validated its rows
            if (tid == 0) atomicMax(&times[4], global_ns());
        }
    }
    if (tid == 0) {
        const bool tell_host = !kFabric && p.done_flag != nullptr;
        // result rows written by this CTA sit in pinned host memory: with a host that polls the flag below
        // instead of waiting for the launch to end, they must be visible system-wide before the count
        if (tell_host) __threadfence_system(); else __threadfence();
        const unsigned int done = atomicAdd(&my_ctl[kCtlCtasDone], 1u);
        if (p.timeline) atomicMax(&times[16], global_ns());
        if (done + 1u == gridDim.x) {
            my_ctl[kCtlNext] = 0u;
            my_ctl[kCtlCtasDone] = 0u;
            my_ctl[kCtlSharedNext] = 0u;
            if (tell_host) {  // the last store of the launch
                __threadfence_system();
                st_relaxed_sys(p.done_flag, p.done_epoch);
            }
        }
    }
}

// kFabric = false: one GPU owns everything (p.key/p.cnt/p.out/p.ctrl are local).
// kFabric = true : one of `world` cooperating launches, one per GPU (see sa_cuda_fabric_solve and the
//   data-movement notes above fabric_prologue); the caller has run fabric_prologue, so s_hdr is valid.
// TILE scan of one warp: R consecutive offsets per lane (8, or 4 / 2 for the narrow tiles of latency mode), rows
// [0, crow) of the warp's row chunk. Difference form (see scan8_u16x2): D[r] = P[r] - P[r+1] is carried, the
// running maximum takes two rows per VIMNMX3. Writes the warp's R+1 prefixes and R maxima for the merge.
template <int R>
__device__ __forceinline__ void tile_scan(const unsigned char* lut_b, const unsigned char* wp, const unsigned char* sp, uint32_t crow,
                                          int* csum, int* cmax, uint32_t warp, uint32_t lane) {
    int D[R], best[R], PR = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) D[r] = 0, best[r] = kNeg;
    // letters of columns (row + d0 + k): wp is aligned to R bytes (8, 4 or 2)
    auto load8 = [](const unsigned char* a) -> uint2 {
        if (R == kR) return *reinterpret_cast<const uint2*>(a);
        if (R == 4) return make_uint2(*reinterpret_cast<const uint32_t*>(a), *reinterpret_cast<const uint32_t*>(a + 4));
        const unsigned short* h = reinterpret_cast<const unsigned short*>(a);
        return make_uint2((uint32_t)h[0] | ((uint32_t)h[1] << 16), (uint32_t)h[2] | ((uint32_t)h[3] << 16));
    };
    uint2 lw0 = load8(wp);
    for (uint32_t row = 0; row < crow; row += kU) {
        const uint2 lw1 = load8(wp + row + kU);
        const uint2 sc = *reinterpret_cast<const uint2*>(sp + row);
        uint32_t L[2 * kU];
#pragma unroll
        for (int k = 0; k < kU; ++k) {
            L[k] = byte_of(lw0.x, lw0.y, k);
            L[kU + k] = byte_of(lw1.x, lw1.y, k);
        }
#pragma unroll
        for (int u = 0; u < kU; u += 2) {
            const unsigned char* lrow0 = lut_b + byte_of(sc.x, sc.y, u) * (SA_ALPHABET * 4u);
            const unsigned char* lrow1 = lut_b + byte_of(sc.x, sc.y, u + 1) * (SA_ALPHABET * 4u);
            int v[R + 1], w[R + 1];
#pragma unroll
            for (int r = 0; r <= R; ++r) v[r] = *reinterpret_cast<const int*>(lrow0 + L[u + r]);
#pragma unroll
            for (int r = 0; r <= R; ++r) w[r] = *reinterpret_cast<const int*>(lrow1 + L[u + 1 + r]);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int d0 = D[r] + v[r] - v[r + 1];
                D[r] = d0 + w[r] - w[r + 1];
                best[r] = __vimax3_s32(best[r], d0, D[r]);
            }
            PR += v[R] + w[R];
        }
        lw0 = lw1;
    }
    int P[R + 1];
    P[R] = PR;
#pragma unroll
    for (int r = R - 1; r >= 0; --r) P[r] = P[r + 1] + D[r];
#pragma unroll
    for (int r = 0; r <= R; ++r) csum[(warp * (kR + 1) + r) * 32 + lane] = P[r];
#pragma unroll
    for (int r = 0; r < R; ++r) cmax[(warp * kR + r) * 32 + lane] = best[r];
}

template <bool kFabric>
__device__ __forceinline__ void score_body(const KParams& p, const bool epilogue_later = false) {
    extern __shared__ __align__(16) unsigned char smem[];
    unsigned char* const lut_b = smem;
    unsigned char* const s2 = lut_b + kLutBytes;
    unsigned char* const win = s2 + p.s2_cap;
    int* const csum = reinterpret_cast<int*>(win + p.win_cap);
    int* const cmax = csum + kSumWords;
    unsigned long long* const red = reinterpret_cast<unsigned long long*>(cmax + kMaxWords);  // [kWarps]
    unsigned long long* const mbar = red + kWarps;
    int* const scan = reinterpret_cast<int*>(mbar + 1);  // [kWarps]
    uint32_t* const ctl = reinterpret_cast<uint32_t*>(scan + kWarps);  // [0] item, [1] last flag, [2] first k

    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const sa_blob_header* const hdr = kFabric ? &s_hdr : reinterpret_cast<const sa_blob_header*>(p.blob);
    const uint32_t bar = smem_u32(mbar);
    uint32_t parity = 0;
    uint32_t* const my_ctl = kFabric ? reinterpret_cast<uint32_t*>(p.peer[p.rank] + p.off_ctrl) : p.ctrl;
    unsigned long long* const times = reinterpret_cast<unsigned long long*>(my_ctl + kCtlTimes);
    const bool timeline = p.timeline != 0u && !kFabric;
    const bool tl_any = p.timeline != 0u;  // slots 11..15: latest CTA to reach each phase of its FIRST item
    if (timeline && tid == 0) atomicMin(&times[0], global_ns());

    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    const bool dead_on_arrival = kFabric && s_abort != 0u;

    // A single-GPU launch gets the header's offsets as kernel parameters (the host has the header):
    // no dependent read of cold global memory before the first item can be fetched.
    const uint32_t off_seqtab = kFabric ? (dead_on_arrival ? 0u : hdr->off_seqtab) : p.h_off_seqtab;
    const uint32_t off_seq1 = kFabric ? (dead_on_arrival ? 0u : hdr->off_seq1) : p.h_off_seq1;
    const uint32_t off_lut = kFabric ? (dead_on_arrival ? 0u : hdr->off_lut) : p.h_off_lut;
    const uint32_t nseq_all = kFabric ? (dead_on_arrival ? 0u : hdr->nseq) : p.h_nseq;
    const uint32_t off_fat = kFabric ? (dead_on_arrival ? 0u : hdr->off_tile_items) : p.h_off_tile_items;
    const sa_seq_desc* const seqtab = reinterpret_cast<const sa_seq_desc*>(p.blob + off_seqtab);
    const sa_tile_item* const fat = reinterpret_cast<const sa_tile_item*>(p.blob + off_fat);
    const unsigned char* const seq1g = p.blob + off_seq1;
    const uint32_t item_begin = kFabric ? (dead_on_arrival ? 0u : hdr->rank_first_item[p.rank]) : p.item_begin;
    const uint32_t item_end = kFabric ? (dead_on_arrival ? 0u : hdr->rank_first_item[p.rank + 1]) : p.item_end;
    int4* const out = kFabric ? reinterpret_cast<int4*>(p.peer[0] + p.off_out) : p.out;
    // Ranks other than 0 first stage their share of rank 0's blob: the first n_copy entries of the work
    // queue are copy items (one chunk each), scoring items follow.
    PullPlan pp{};
    if (kFabric && p.rank != 0 && !p.skip_push && !dead_on_arrival) pp = make_pull_plan(*hdr, p.rank);
    const uint32_t n_copy = pp.total;

    // Sequences longer than Seq1 have no candidates: (INT_MIN, 0, 0) as the reference's
    // untouched `max` (cudaFunctions.cu:113,170-172).
    if (p.write_empty && !dead_on_arrival) {
        for (uint32_t s = blockIdx.x * kThreads + tid; s < nseq_all; s += gridDim.x * kThreads)
            if (seqtab[s].ntiles == 0) out[s] = make_int4(INT_MIN, 0, 0, 0);
    }

    // CTA-uniform: the first queue entry is blockIdx.x (no round trip to the queue) — unless copy items exist:
    // those must only ever be held by CTAs that are running, or a waiting CTA could wait for one that has not started
    bool fresh = n_copy == 0u;
    const uint32_t q_base = fresh ? gridDim.x : 0u;
    bool first = true;  // CTA-uniform: the first scoring item brings the table along
    for (;;) {
        if (tid == 0) {
            uint32_t it = fresh ? blockIdx.x : q_base + atomicAdd(&my_ctl[kCtlNext], 1u);
            if (!kFabric && fresh && hdr->magic != SA_BLOB_MAGIC) it = 0xFFFFFFF0u;  // never walk memory that is not a packed problem
            if (kFabric && it >= n_copy && it - n_copy < item_end - item_begin && p.rank != 0) {
                // this item's record and letters (and, once per CTA, the shared part) may still be on their way
                const uint32_t idx = item_begin + (it - n_copy);
                const uint32_t rec = off_fat + 32u * idx;
                bool ok = fabric_bytes_ready(p, my_ctl, rec, rec + 32u);
                if (ok) {
                    const sa_tile_item d = fat[idx];
                    const uint32_t bytes = ((uint32_t)d.crow * ((uint32_t)kWarps >> d.wd_log2) + 15u) & ~15u;
                    ok = fabric_bytes_ready(p, my_ctl, d.off, d.off + bytes);
                }
                if (ok && first) ok = fabric_bytes_ready(p, my_ctl, 0u, hdr->off_shared_end);
                if (!ok) {
                    s_abort = 1u;
                    atomicExch(my_ctl + kCtlError, 1u);
                }
                fence_proxy_async();
            }
            ctl[0] = it;
            if (tl_any && first) atomicMax(&times[11], global_ns());
        }
        __syncthreads();  // also: every read of s2/win/csum of the previous item is done
        fresh = false;
        const uint32_t raw = ctl[0];
        if (kFabric && s_abort != 0u) break;
        if (raw < n_copy) {  // copy item: the whole CTA stages one chunk, one thread publishes it
            const uint32_t chunk = pull_chunk_of(pp, raw);
            pull_copy(p, *hdr, chunk, tid, kThreads);
            __syncthreads();
            if (tid == 0) pull_publish(p, my_ctl, chunk, n_copy);
            continue;
        }
        if (raw - n_copy >= item_end - item_begin) break;
        const uint32_t item = item_begin + (raw - n_copy);

        const sa_tile_item sd = fat[item];  // one self-contained 32-byte record per item
        const uint32_t seq = sd.seq, tile = sd.tile;
        const uint32_t wd_log2 = sd.wd_log2, wd = 1u << wd_log2, wr = (uint32_t)kWarps >> wd_log2;
        const uint32_t crow = sd.crow, rows_padded = crow * wr;
        const uint32_t rl = 3u - sd.narrow;  // log2(offsets per lane)
        const uint32_t tn = (32u << rl) << wd_log2, n0 = tile * tn;

        unsigned long long* const keyp =
            (kFabric ? reinterpret_cast<unsigned long long*>(p.peer[sd.home] + p.off_key) : p.key) + seq;
        unsigned int* const cntp = (kFabric ? reinterpret_cast<unsigned int*>(p.peer[sd.home] + p.off_cnt) : p.cnt) + seq;

        if (tid == 0) {
            const uint32_t s2_bytes = (rows_padded + 15u) & ~15u;
            const uint32_t win_bytes = (rows_padded + tn + 15u) & ~15u;
            mbar_expect_tx(bar, s2_bytes + win_bytes + (first ? kLutBytes : 0u));
            if (first) bulk_g2s(smem_u32(lut_b), p.blob + off_lut, kLutBytes, bar);  // pair-value table, once per CTA
            bulk_g2s(smem_u32(s2), p.blob + sd.off, s2_bytes, bar);
            bulk_g2s(smem_u32(win), seq1g + n0, win_bytes, bar);
        }
        mbar_wait(bar, parity);
        parity ^= 1u;
        const bool was_first = first;
        first = false;
        if (tl_any && was_first && tid == 0) atomicMax(&times[12], global_ns());

        // ---- scan: this warp = offset group g, row chunk q
        {
            const uint32_t g = warp & (wd - 1u), q = warp >> wd_log2;
            const unsigned char* wp = win + ((g * 32u + lane) << rl) + q * crow;
            if (sd.narrow == 0) tile_scan<kR>(lut_b, wp, s2 + q * crow, crow, csum, cmax, warp, lane);
            else if (sd.narrow == 1) tile_scan<kR / 2>(lut_b, wp, s2 + q * crow, crow, csum, cmax, warp, lane);
            else tile_scan<kR / 4>(lut_b, wp, s2 + q * crow, crow, csum, cmax, warp, lane);
        }
        __syncthreads();
        if (tl_any && was_first && tid == 0) atomicMax(&times[13], global_ns());

        // ---- merge row chunks in order, score every offset of the tile, keep the best
        unsigned long long mykey = 0ull;
        for (uint32_t t = tid; t < tn; t += kThreads) {
            const uint32_t l = t & 31u, r = (t >> 5) & ((1u << rl) - 1u), g = t >> (5u + rl);
            int A = 0, T0 = 0, T1 = 0, M = kNeg;
            for (uint32_t q = 0; q < wr; ++q) {
                const uint32_t w = (q << wd_log2) + g;
                const int s0 = csum[(w * (kR + 1) + r) * 32 + l];
                const int s1 = csum[(w * (kR + 1) + r + 1) * 32 + l];
                const int mx = cmax[(w * kR + r) * 32 + l];
                M = max(M, A + mx);
                A += s0 - s1;
                T0 += s0;
                T1 += s1;
            }
            const uint32_t n = n0 + (((g << 5) + l) << rl) + r;
            if (sd.ndiag == 0) {
                if (n == 0) mykey = pack_key(T0, 0u);  // equal lengths: one candidate (cudaFunctions.cu:74-106)
            } else if (n < (uint32_t)sd.ndiag) {
                mykey = umax64(mykey, pack_key(T1 + M, n));
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mykey = umax64(mykey, __shfl_xor_sync(0xFFFFFFFFu, mykey, o));
        if (lane == 0) red[warp] = mykey;
        __syncthreads();
        if (warp == 0) {
            unsigned long long v = lane < (uint32_t)kWarps ? red[lane] : 0ull;
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) v = umax64(v, __shfl_xor_sync(0xFFFFFFFFu, v, o));
            if (lane == 0 && sd.ntiles == 1) {  // the sequence's only tile: nothing to combine, no atomics
                red[0] = v;
                ctl[1] = 1u;
                ctl[2] = 0x7FFFFFFFu;
            } else if (lane == 0) {
                unsigned int old;
                if (kFabric) {
                    atomicMax_system(keyp, v);
                    // both atomics resolve in the home rank's L2; a device-scope fence orders them there
                    if (sd.home == p.rank) __threadfence(); else __threadfence_system();
                    old = atomicAdd_system(cntp, 1u);
                } else {
                    atomicMax(keyp, v);
                    __threadfence();
                    old = atomicAdd(cntp, 1u);
                }
                ctl[1] = (old + 1u == sd.ntiles) ? 1u : 0u;
                ctl[2] = 0x7FFFFFFFu;
            }
        }
        __syncthreads();

        if (tl_any && was_first && tid == 0) atomicMax(&times[14], global_ns());
        // ---- last tile of this sequence: find the first k on the winning diagonal
        if (ctl[1]) {
            if (tid == 0 && sd.ntiles != 1) {
                if (kFabric) {
                    if (sd.home == p.rank) __threadfence(); else __threadfence_system();
                    red[0] = atomicMax_system(keyp, 0ull);  // coherent read of the final key
                } else {
                    __threadfence();
                    red[0] = atomicMax(keyp, 0ull);
                }
            }
            __syncthreads();
            const unsigned long long fk = red[0];
            const int bestv = (int)((uint32_t)(fk >> 32) ^ 0x80000000u);
            const uint32_t nwin = 0xFFFFFFFFu - (uint32_t)(fk & 0xFFFFFFFFull);
            const int kres = sd.ndiag == 0 ? 0 : first_k_on_diagonal(reinterpret_cast<const int*>(lut_b), s2, seq1g + nwin,
                                                                      sd.len, bestv, scan, csum, &ctl[2]);
            if (tid == 0) {
                if (sd.ntiles != 1) {
                    *keyp = 0ull;  // leave the scratch zeroed for the next launch ...
                    *cntp = 0u;
                    // ... and make sure it IS zero before anybody can learn that this solve is over: the row
                    // below is what lets rank 0's launch end and the next solve's atomics start
                    if (kFabric) { if (sd.home == p.rank) __threadfence(); else __threadfence_system(); }
                }
                if (kFabric) post_result(p, seq, bestv, nwin, (uint32_t)kres);
                else out[seq] = make_int4(bestv, (int)nwin, kres, 0);
                if (tl_any && was_first) atomicMax(&times[15], global_ns());
            }
        }
    }
    if (tl_any && tid == 0) atomicMax(&times[2], global_ns());

    if (!epilogue_later) launch_epilogue<kFabric>(p);
}

__global__ void __launch_bounds__(kThreads, 2) sa_score_kernel(const KParams p) { score_body<false>(p); }

// ------------------------------------------------------------------ PAIR mode kernel
//
// Throughput mode for large batches (SA_MODE_PAIR). One CTA owns two sequences of similar length
// end to end:
//   * their biased pair values are laid out as ONE profile in shared memory,
//     prof[row][letter] = (valueA - vmin) | (valueB - vmin) << 16, so a single conflict-free LDS
//     feeds both sequences and its address is `letter register + immediate` (no per-cell math);
//   * prefixes and running maxima are u16x2: P += v is a plain IADD (halves never carry: the
//     planner bounds chunk_rows * value_range by 32767), the neighbour difference is one IADD3
//     `P[r] + BIG - P[r+1]` (never borrows) and the running maximum is one DPX VIMNMX.U16x2;
//   * Seq1 stays resident in shared memory for the whole kernel; no atomics, no fences: the tile
//     loop keeps the two best keys in registers and the CTA finishes both sequences itself.
__global__ void __launch_bounds__(kThreads, 1) sa_score_pair_kernel(const KParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    const sa_blob_header* const hdr = reinterpret_cast<const sa_blob_header*>(p.blob);
    if (hdr->magic != SA_BLOB_MAGIC) return;  // never walk memory that is not a packed problem
    const uint32_t seq1_bytes = (hdr->len1_padded + 15u) & ~15u;
    unsigned char* const lut_b = smem;
    unsigned char* const seq1s = lut_b + kLutBytes;
    unsigned char* const sA = seq1s + seq1_bytes;
    unsigned char* const sB = sA + p.s2_cap;
    uint32_t* const prof = reinterpret_cast<uint32_t*>(sB + p.s2_cap);
    uint32_t* const csum = prof + (size_t)p.s2_cap * SA_ALPHABET;
    uint32_t* const cmax = csum + kSumWords;
    unsigned long long* const red = reinterpret_cast<unsigned long long*>(cmax + kMaxWords);  // [2 * kWarps]
    unsigned long long* const mbar = red + 2 * kWarps;
    int* const scan = reinterpret_cast<int*>(mbar + 1);               // [kWarps]
    int* const tmp = scan + kWarps;                                   // [kWarps]
    uint32_t* const ctl = reinterpret_cast<uint32_t*>(tmp + kWarps);  // [0] item, [2] first k

    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const sa_seq_desc* const seqtab = reinterpret_cast<const sa_seq_desc*>(p.blob + hdr->off_seqtab);
    const sa_item* const items = reinterpret_cast<const sa_item*>(p.blob + hdr->off_items);
    const int* const lut_i = reinterpret_cast<const int*>(lut_b);
    const int vmin = hdr->value_min;
    constexpr uint32_t BIG = kBig;
    const uint32_t bar = smem_u32(mbar);
    uint32_t parity = 0;

    if (p.write_empty) {
        for (uint32_t s = blockIdx.x * kThreads + tid; s < hdr->nseq; s += gridDim.x * kThreads)
            if (seqtab[s].ntiles == 0) p.out[s] = make_int4(INT_MIN, 0, 0, 0);
    }
    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (tid == 0) {  // table + the whole of Seq1, once per CTA
        mbar_expect_tx(bar, kLutBytes + seq1_bytes);
        bulk_g2s(smem_u32(lut_b), p.blob + hdr->off_lut, kLutBytes, bar);
        bulk_g2s(smem_u32(seq1s), p.blob + hdr->off_seq1, seq1_bytes, bar);
    }
    mbar_wait(bar, parity);
    parity ^= 1u;

    for (;;) {
        if (tid == 0) ctl[0] = p.item_begin + atomicAdd(&p.ctrl[kCtlNext], 1u);
        __syncthreads();  // also: the previous pair's shared state is no longer read
        const uint32_t item = ctl[0];
        if (item >= p.item_end) break;
        const uint32_t ia = items[item].seq, ib = items[item].tile;
        const sa_seq_desc da = seqtab[ia];
        const bool has_b = ib != SA_NO_SEQ;
        sa_seq_desc db = da;
        if (has_b) db = seqtab[ib];
        const uint32_t wd_log2 = da.wd_log2, wd = 1u << wd_log2, wr = (uint32_t)kWarps >> wd_log2;
        const uint32_t crow = da.crow, rows = crow * wr, tn = (uint32_t)SA_DIAG_PER_WARP << wd_log2;
        const uint32_t s2_bytes = (rows + 15u) & ~15u;

        if (tid == 0) {
            mbar_expect_tx(bar, has_b ? 2u * s2_bytes : s2_bytes);
            bulk_g2s(smem_u32(sA), p.blob + da.off, s2_bytes, bar);
            if (has_b) bulk_g2s(smem_u32(sB), p.blob + db.off, s2_bytes, bar);
        }
        if (!has_b)
            for (uint32_t i = tid; i < s2_bytes; i += kThreads) sB[i] = 0;  // an all-padding partner
        mbar_wait(bar, parity);
        parity ^= 1u;
        __syncthreads();

        // ---- pair profile: one warp per row, lane = letter
        for (uint32_t idx = tid; idx < rows * SA_ALPHABET; idx += kThreads) {
            const uint32_t row = idx >> 5, c = idx & 31u;
            const uint32_t ea = (uint32_t)(lut_i[(uint32_t)sA[row] * SA_ALPHABET + c] - vmin);
            const uint32_t eb = (uint32_t)(lut_i[(uint32_t)sB[row] * SA_ALPHABET + c] - vmin);
            prof[idx] = ea | (eb << 16);
        }
        __syncthreads();

        unsigned long long keyA = 0ull, keyB = 0ull;
        const uint32_t g = warp & (wd - 1u), q = warp >> wd_log2;
        const uint32_t d0 = g * SA_DIAG_PER_WARP + lane * kR;
        for (uint32_t tile = 0; tile < da.ntiles; ++tile) {
            const uint32_t n0 = tile * tn;
            // ---- scan: this warp = offset group g, row chunk q
            {
                const unsigned char* wp = seq1s + n0 + d0 + q * crow;
                const unsigned char* pb = reinterpret_cast<const unsigned char*>(prof) + (size_t)q * crow * (SA_ALPHABET * 4u);
                uint32_t D[kR], best[kR], P8 = 0u;
#pragma unroll
                for (int r = 0; r < kR; ++r) D[r] = kBig2, best[r] = 0u;
                uint2 lw0 = *reinterpret_cast<const uint2*>(wp);
                for (uint32_t row = 0; row < crow; row += kU) {
                    const uint2 lw1 = *reinterpret_cast<const uint2*>(wp + row + kU);
                    const unsigned char* A[2 * kU];
#pragma unroll
                    for (int k = 0; k < kU; ++k) {
                        A[k] = pb + byte_of_prmt(lw0.x, lw0.y, k);
                        A[kU + k] = pb + byte_of_prmt(lw1.x, lw1.y, k);
                    }
                    scan8_u16x2(A, D, best, P8);
                    lw0 = lw1;
                    pb += kU * (SA_ALPHABET * 4u);
                }
                uint32_t P[kR + 1];
                to_prefixes_u16x2(D, P8, P);
#pragma unroll
                for (int r = 0; r <= kR; ++r) csum[(warp * (kR + 1) + r) * 32 + lane] = P[r];
#pragma unroll
                for (int r = 0; r < kR; ++r) cmax[(warp * kR + r) * 32 + lane] = best[r];
            }
            __syncthreads();
            // ---- merge row chunks in order for both sequences, keep the best offsets so far
            const int cbias = (int)crow * vmin;
            for (uint32_t t = tid; t < tn; t += kThreads) {
                const uint32_t l = t & 31u, r = (t >> 5) & 7u, gg = t >> 8;
                int A0 = 0, T00 = 0, T10 = 0, M0 = kNeg, A1 = 0, T01 = 0, T11 = 0, M1 = kNeg;
                for (uint32_t qq = 0; qq < wr; ++qq) {
                    const uint32_t w = (qq << wd_log2) + gg;
                    const uint32_t s0 = csum[(w * (kR + 1) + r) * 32 + l];
                    const uint32_t s1 = csum[(w * (kR + 1) + r + 1) * 32 + l];
                    const uint32_t mx = cmax[(w * kR + r) * 32 + l];
                    const int s0a = (int)(s0 & 0xFFFFu) + cbias, s1a = (int)(s1 & 0xFFFFu) + cbias;
                    const int s0b = (int)(s0 >> 16) + cbias, s1b = (int)(s1 >> 16) + cbias;
                    M0 = max(M0, A0 + (int)(mx & 0xFFFFu) - (int)BIG);
                    M1 = max(M1, A1 + (int)(mx >> 16) - (int)BIG);
                    A0 += s0a - s1a;
                    A1 += s0b - s1b;
                    T00 += s0a;
                    T10 += s1a;
                    T01 += s0b;
                    T11 += s1b;
                }
                const uint32_t n = n0 + gg * SA_DIAG_PER_WARP + l * kR + r;
                if (da.ndiag == 0) {
                    if (n == 0) keyA = pack_key(T00, 0u);
                } else if (n < (uint32_t)da.ndiag) {
                    keyA = umax64(keyA, pack_key(T10 + M0, n));
                }
                if (has_b) {
                    if (db.ndiag == 0) {
                        if (n == 0) keyB = pack_key(T01, 0u);
                    } else if (n < (uint32_t)db.ndiag) {
                        keyB = umax64(keyB, pack_key(T11 + M1, n));
                    }
                }
            }
            __syncthreads();
        }
        // ---- block argmax for both sequences
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            keyA = umax64(keyA, __shfl_xor_sync(0xFFFFFFFFu, keyA, o));
            keyB = umax64(keyB, __shfl_xor_sync(0xFFFFFFFFu, keyB, o));
        }
        if (lane == 0) {
            red[warp] = keyA;
            red[kWarps + warp] = keyB;
        }
        __syncthreads();
        unsigned long long fa = 0ull, fb = 0ull;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) {
            fa = umax64(fa, red[w]);
            fb = umax64(fb, red[kWarps + w]);
        }
        // ---- first k of each winner, results out
        for (int half = 0; half < (has_b ? 2 : 1); ++half) {
            const sa_seq_desc& sd = half == 0 ? da : db;
            const unsigned long long fk = half == 0 ? fa : fb;
            const int bestv = (int)((uint32_t)(fk >> 32) ^ 0x80000000u);
            const uint32_t nwin = 0xFFFFFFFFu - (uint32_t)(fk & 0xFFFFFFFFull);
            if (tid == 0) ctl[2] = 0x7FFFFFFFu;
            __syncthreads();
            const int kres = sd.ndiag == 0 ? 0 : first_k_on_diagonal(lut_i, half == 0 ? sA : sB, seq1s + nwin, sd.len, bestv, scan, tmp, &ctl[2]);
            if (tid == 0) p.out[half == 0 ? ia : ib] = make_int4(bestv, (int)nwin, kres, 0);
            __syncthreads();
        }
    }
    if (tid == 0) {
        __threadfence();
        const unsigned int done = atomicAdd(&p.ctrl[kCtlCtasDone], 1u);
        if (done + 1u == gridDim.x) {
            p.ctrl[kCtlNext] = 0u;
            p.ctrl[kCtlCtasDone] = 0u;
        }
    }
}

// ------------------------------------------------------------------ STREAM mode kernel
//
// SA_MODE_STREAM: the u16x2 arithmetic of PAIR mode without any block-level phase. Every WARP is
// autonomous: it takes a (pair, 256-offset tile) from the queue and walks ALL rows of the pair,
// 32 at a time; for each block of 32 rows it rebuilds a private 4 KB profile block
// (lane = Seq1 letter, one conflict-free row per Seq2 position) and scans it with the same
// LDS + IADD + IADD3 + VIMNMX.U16x2 inner loop. There is no row-chunk merge and no
// __syncthreads after start-up; per-sequence keys are combined with atomics and the warp that
// delivers a pair's last tile finishes both sequences on its own.
__device__ __forceinline__ int warp_first_k(const uint32_t* lutlo, const unsigned char* g2, const unsigned char* c1,
                                            uint32_t m, int bestv, int vmin, uint32_t lane) {
    const uint32_t per = (m + 31u) >> 5;
    const uint32_t i0 = min(lane * per, m), i1 = min(i0 + per, m);
    int da = 0, sb = 0;
    for (uint32_t i = i0; i < i1; ++i) {
        const uint32_t* lrow = lutlo + (uint32_t)g2[i] * SA_ALPHABET;
        const int a = (int)lrow[c1[i] >> 2], b = (int)lrow[c1[i + 1] >> 2];  // both biased: the bias cancels in a - b
        da += a - b;
        sb += b;
    }
    int inc = da, tsb = sb;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int x = __shfl_up_sync(0xFFFFFFFFu, inc, o);
        if (lane >= (uint32_t)o) inc += x;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tsb += __shfl_xor_sync(0xFFFFFFFFu, tsb, o);
    const int T1 = tsb + (int)m * vmin;
    const int T0 = T1 + __shfl_sync(0xFFFFFFFFu, inc, 31);
    if (T0 == bestv) return 0;  // k = 0 is tried first and wins the tie
    int S = inc - da;
    uint32_t mine = 0x7FFFFFFFu;
    for (uint32_t i = i0; i < i1; ++i) {
        const uint32_t* lrow = lutlo + (uint32_t)g2[i] * SA_ALPHABET;
        S += (int)lrow[c1[i] >> 2] - (int)lrow[c1[i + 1] >> 2];
        if (i + 1 < m && T1 + S == bestv) {
            mine = i + 1u;
            break;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mine = min(mine, __shfl_xor_sync(0xFFFFFFFFu, mine, o));
    return (int)mine;
}

template <bool kFabric>
__device__ __forceinline__ void stream_body(const KParams& p, const bool epilogue_later = false) {
    extern __shared__ __align__(16) unsigned char smem[];
    const sa_blob_header* const hdr = kFabric ? &s_hdr : reinterpret_cast<const sa_blob_header*>(p.blob);
    if (!kFabric && hdr->magic != SA_BLOB_MAGIC) return;  // never walk memory that is not a packed problem
    const uint32_t seq1_bytes = (hdr->len1_padded + 15u) & ~15u;
    uint32_t* const lutlo = reinterpret_cast<uint32_t*>(smem);
    // lutd[code][l] = the biased values of letters 2*(l & 15) and 2*(l & 15) + 1 as two 16-bit halves; each row holds
    // them twice (l = 0..15 and 16..31), so that the two half-warps of the profile build below read different banks
    uint32_t* const lutd = lutlo + SA_ALPHABET * SA_ALPHABET;
    unsigned char* const seq1s = reinterpret_cast<unsigned char*>(lutd + SA_ALPHABET * SA_ALPHABET);
    uint32_t* const prof = reinterpret_cast<uint32_t*>(seq1s + seq1_bytes);
    unsigned long long* const mbar = reinterpret_cast<unsigned long long*>(prof + kWarps * SA_STREAM_ROW_BLOCK * SA_ALPHABET);

    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const sa_seq_desc* const seqtab = reinterpret_cast<const sa_seq_desc*>(p.blob + hdr->off_seqtab);
    const unsigned char* const desc_base = (kFabric && (p.dbg_src & 2u)) ? p.peer[0] + p.off_blob : p.blob;
    const unsigned char* const code_base = (kFabric && (p.dbg_src & 1u)) ? p.peer[0] + p.off_blob : p.blob;
    const sa_item* const items = reinterpret_cast<const sa_item*>(desc_base + hdr->off_items);
    const sa_pair_rec* const pairs = reinterpret_cast<const sa_pair_rec*>(desc_base + hdr->off_pairs);
    const int vmin = hdr->value_min;
    const int len1 = (int)hdr->len1;
    constexpr uint32_t BIG = kBig;
    const uint32_t bar = smem_u32(mbar);
    uint32_t* const my_ctl = kFabric ? reinterpret_cast<uint32_t*>(p.peer[p.rank] + p.off_ctrl) : p.ctrl;
    int4* const out = kFabric ? reinterpret_cast<int4*>(p.peer[0] + p.off_out) : p.out;
    const uint32_t item_begin = kFabric ? hdr->rank_first_item[p.rank] : p.item_begin;
    const uint32_t item_end = kFabric ? hdr->rank_first_item[p.rank + 1] : p.item_end;
    // Ranks other than 0 stage their share of rank 0's blob first (see fabric_prologue's notes): the
    // shared part by whole CTAs right here (the CTA-wide start-up below reads it), the rest by single
    // warps as the first entries of the warp-level work queue.
    PullPlan pp{};
    if (kFabric && p.rank != 0 && !p.skip_push) pp = make_pull_plan(*hdr, p.rank);
    const uint32_t n_copy = pp.total, n_shared = pp.count[0], n_warp_copy = n_copy - n_shared;
    if (kFabric && n_shared) {
        __shared__ uint32_t s_chunk;
        for (;;) {
            if (tid == 0) s_chunk = atomicAdd(&my_ctl[kCtlSharedNext], 1u);
            __syncthreads();
            const uint32_t j = s_chunk;
            if (j >= n_shared) break;
            pull_copy(p, *hdr, pp.first[0] + j, tid, kThreads);
            __syncthreads();
            if (tid == 0) pull_publish(p, my_ctl, pp.first[0] + j, n_copy);
        }
        if (tid == 0) {
            if (!fabric_bytes_ready(p, my_ctl, 0u, hdr->off_shared_end)) {
                s_abort = 1u;
                atomicExch(my_ctl + kCtlError, 1u);
            }
            fence_proxy_async();
        }
        __syncthreads();
        if (s_abort != 0u) {
            if (!epilogue_later) launch_epilogue<kFabric>(p);
            return;
        }
    }

    if (p.write_empty) {
        for (uint32_t s = blockIdx.x * kThreads + tid; s < hdr->nseq; s += gridDim.x * kThreads)
            if (seqtab[s].ntiles == 0) out[s] = make_int4(INT_MIN, 0, 0, 0);
    }
    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (tid == 0) {  // raw table + the whole of Seq1, once per CTA
        mbar_expect_tx(bar, kLutBytes + seq1_bytes);
        bulk_g2s(smem_u32(lutlo), p.blob + hdr->off_lut, kLutBytes, bar);
        bulk_g2s(smem_u32(seq1s), p.blob + hdr->off_seq1, seq1_bytes, bar);
    }
    mbar_wait(bar, 0u);
    // biased table (first-k rescan) and its letter-paired form (profile build)
    for (uint32_t i = tid; i < SA_ALPHABET * SA_ALPHABET; i += kThreads)
        lutlo[i] = (uint32_t)(reinterpret_cast<const int*>(lutlo)[i] - vmin);
    __syncthreads();
    for (uint32_t i = tid; i < SA_ALPHABET * SA_ALPHABET; i += kThreads) {
        const uint32_t row = i & ~(SA_ALPHABET - 1u), c = 2u * (i & 15u);
        lutd[i] = lutlo[row + c] | (lutlo[row + c + 1u] << 16);
    }
    __syncthreads();  // the last block-wide barrier of this kernel

    uint32_t* const pw = prof + warp * (SA_STREAM_ROW_BLOCK * SA_ALPHABET);
    const uint32_t d0 = lane * kR;
    unsigned long long dbg_copied = 0ull;
    for (;;) {
        uint32_t raw = 0;
        if (lane == 0) raw = atomicAdd(&my_ctl[kCtlNext], 1u);
        raw = __shfl_sync(0xFFFFFFFFu, raw, 0);
        if (kFabric && *reinterpret_cast<volatile uint32_t*>(&s_abort) != 0u) break;
        if (raw < n_warp_copy) {  // copy item: this warp stages one chunk of its rank's share
            const uint32_t chunk = pull_chunk_of(pp, n_shared + raw);
            pull_copy(p, *hdr, chunk, lane, 32u);
            __syncwarp();
            if (lane == 0) pull_publish(p, my_ctl, chunk, n_copy);
            dbg_copied = 1u;
            continue;
        }
        if (raw - n_warp_copy >= item_end - item_begin) break;
        const uint32_t item = item_begin + (raw - n_warp_copy);
        const unsigned long long t_item0 = (kFabric && p.timeline) ? global_ns() : 0ull;
        if (kFabric && p.timeline && lane == 0) {  // when were 25 / 50 / 75 % of this rank's items handed out?
            const uint32_t n_it = item_end - item_begin, k = item - item_begin;
            unsigned long long* const times = reinterpret_cast<unsigned long long*>(my_ctl + kCtlTimes);
            if (k == n_it / 4) times[8] = global_ns();
            if (k == n_it / 2) times[9] = global_ns();
            if (k == (n_it / 4) * 3) times[10] = global_ns();
        }
        if (kFabric && p.rank != 0) {  // the item and its pair record may still be on their way (warp-uniform wait)
            const uint32_t io = hdr->off_items + 8u * item;
            bool ok = fabric_bytes_ready(p, my_ctl, io, io + 8u, lane == 0);
            if (ok) {
                const uint32_t po = hdr->off_pairs + 32u * items[item].seq;
                ok = fabric_bytes_ready(p, my_ctl, po, po + 32u, lane == 0);
            }
            if (!ok) {
                *reinterpret_cast<volatile uint32_t*>(&s_abort) = 1u;
                if (lane == 0) atomicExch(my_ctl + kCtlError, 1u);
                break;
            }
        }
        const uint32_t pid = items[item].seq, tile = items[item].tile;
        const sa_pair_rec pr = pairs[pid];
        const uint32_t ia = pr.a, ib = pr.b;
        const bool has_b = ib != SA_NO_SEQ;
        const int ndiag_a = len1 - (int)pr.len_a, ndiag_b = len1 - (int)pr.len_b;
        const uint32_t rows = pr.crow, n0 = tile * SA_DIAG_PER_WARP;
        const unsigned char* const gA = code_base + pr.off_a;
        const unsigned char* const gB = code_base + pr.off_b;
        unsigned long long* const keyA_p = (kFabric ? reinterpret_cast<unsigned long long*>(p.peer[pr.home] + p.off_key) : p.key) + ia;
        unsigned long long* const keyB_p = (kFabric ? reinterpret_cast<unsigned long long*>(p.peer[pr.home] + p.off_key) : p.key) + (has_b ? ib : ia);
        unsigned int* const cnt_p = (kFabric ? reinterpret_cast<unsigned int*>(p.peer[pr.home] + p.off_cnt) : p.cnt) + ia;
        if (kFabric && p.rank != 0) {  // the two sequences' letters may still be on their way (warp-uniform wait)
            bool ok = fabric_bytes_ready(p, my_ctl, pr.off_a, pr.off_a + rows, lane == 0);
            if (ok && has_b) ok = fabric_bytes_ready(p, my_ctl, pr.off_b, pr.off_b + rows, lane == 0);
            if (!ok) {
                *reinterpret_cast<volatile uint32_t*>(&s_abort) = 1u;
                if (lane == 0) atomicExch(my_ctl + kCtlError, 1u);
                break;
            }
        }

        uint32_t D[kR], best[kR], P8 = 0u;
#pragma unroll
        for (int r = 0; r < kR; ++r) D[r] = kBig2, best[r] = 0u;
        const unsigned char* wp = seq1s + n0 + d0;
        uint2 lw0 = *reinterpret_cast<const uint2*>(wp);
        for (uint32_t rb = 0; rb < rows; rb += SA_STREAM_ROW_BLOCK) {
            // ---- private profile block. Every lane fetches the block's 32+32 codes itself (4 broadcast
            // LDG.128 served by L1) and picks its row's bytes with ALU ops: the load/store pipe is the
            // busiest unit of this kernel (93 % of its wavefront peak under ncu), so no SHFL here.
            uint32_t wa[8], wb[8];
            {
                const uint4 a0 = *reinterpret_cast<const uint4*>(gA + rb), a1 = *reinterpret_cast<const uint4*>(gA + rb + 16);
                wa[0] = a0.x, wa[1] = a0.y, wa[2] = a0.z, wa[3] = a0.w, wa[4] = a1.x, wa[5] = a1.y, wa[6] = a1.z, wa[7] = a1.w;
                if (has_b) {
                    const uint4 b0 = *reinterpret_cast<const uint4*>(gB + rb), b1 = *reinterpret_cast<const uint4*>(gB + rb + 16);
                    wb[0] = b0.x, wb[1] = b0.y, wb[2] = b0.z, wb[3] = b0.w, wb[4] = b1.x, wb[5] = b1.y, wb[6] = b1.z, wb[7] = b1.w;
                } else {
#pragma unroll
                    for (int k = 0; k < 8; ++k) wb[k] = 0u;
                }
            }
            // Two rows per step: half-warp h builds row 2j + h, a lane the letters 2c and 2c + 1 of it (c = lane & 15):
            // one LDS per sequence (a wavefront each: the halves read different banks) and one 8-byte store,
            // 64 wavefronts per block where a row-per-step build needs 96.
            {
                const uint32_t h = lane >> 4;
                const uint32_t sel0 = 0x4440u | h, sel1 = 0x4440u | (2u + h);  // byte (2j + h) & 3 of word j >> 1
                uint2* const pw2 = reinterpret_cast<uint2*>(pw);
#pragma unroll
                for (int j = 0; j < (int)SA_STREAM_ROW_BLOCK / 2; ++j) {
                    const uint32_t sel = (j & 1) ? sel1 : sel0;
                    const uint32_t va = lutd[__byte_perm(wa[j >> 1], 0u, sel) * SA_ALPHABET + lane];
                    const uint32_t vb = lutd[__byte_perm(wb[j >> 1], 0u, sel) * SA_ALPHABET + lane];
                    // row 2j + h, words 2c and 2c + 1 = 8-byte slot (2j + h) * 16 + c = 32 j + lane
                    pw2[j * 32 + lane] = make_uint2(__byte_perm(va, vb, 0x5410u), __byte_perm(va, vb, 0x7632u));
                }
            }
            __syncwarp();
            // ---- scan the block: 4 trips of 8 rows
            const unsigned char* pb = reinterpret_cast<const unsigned char*>(pw);
#pragma unroll 1
            for (uint32_t row = 0; row < SA_STREAM_ROW_BLOCK; row += kU) {
                const uint2 lw1 = *reinterpret_cast<const uint2*>(wp + rb + row + kU);
                const unsigned char* A[2 * kU];
#pragma unroll
                for (int k = 0; k < kU; ++k) {
                    A[k] = pb + byte_of_prmt(lw0.x, lw0.y, k);
                    A[kU + k] = pb + byte_of_prmt(lw1.x, lw1.y, k);
                }
                scan8_u16x2(A, D, best, P8);
                lw0 = lw1;
                pb += kU * (SA_ALPHABET * 4u);
            }
            __syncwarp();
        }
        // ---- scores of this warp's 256 offsets for both sequences
        uint32_t P[kR + 1];
        to_prefixes_u16x2(D, P8, P);
        const int vb = (int)rows * vmin;
        unsigned long long keyA = 0ull, keyB = 0ull;
#pragma unroll
        for (int r = 0; r < kR; ++r) {
            const uint32_t n = n0 + d0 + r;
            const int t0a = (int)(P[r] & 0xFFFFu) + vb, t1a = (int)(P[r + 1] & 0xFFFFu) + vb;
            const int t0b = (int)(P[r] >> 16) + vb, t1b = (int)(P[r + 1] >> 16) + vb;
            const int ma = (int)(best[r] & 0xFFFFu) - (int)BIG, mb = (int)(best[r] >> 16) - (int)BIG;
            if (ndiag_a == 0) {
                if (n == 0) keyA = pack_key(t0a, 0u);
            } else if (n < (uint32_t)ndiag_a) {
                keyA = umax64(keyA, pack_key(t1a + ma, n));
            }
            if (has_b) {
                if (ndiag_b == 0) {
                    if (n == 0) keyB = pack_key(t0b, 0u);
                } else if (n < (uint32_t)ndiag_b) {
                    keyB = umax64(keyB, pack_key(t1b + mb, n));
                }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            keyA = umax64(keyA, __shfl_xor_sync(0xFFFFFFFFu, keyA, o));
            keyB = umax64(keyB, __shfl_xor_sync(0xFFFFFFFFu, keyB, o));
        }
        uint32_t last = 0;
        const bool home_is_me = !kFabric || pr.home == p.rank;
        if (lane == 0) {
            if (pr.ntiles == 1) {  // the pair's only tile: nothing to combine, no atomics
                last = 1u;
            } else if (kFabric) {
                atomicMax_system(keyA_p, keyA);
                if (has_b) atomicMax_system(keyB_p, keyB);
                if (home_is_me) __threadfence(); else __threadfence_system();
                last = (atomicAdd_system(cnt_p, 1u) + 1u == pr.ntiles) ? 1u : 0u;
            } else {
                atomicMax(keyA_p, keyA);
                if (has_b) atomicMax(keyB_p, keyB);
                __threadfence();
                last = (atomicAdd(cnt_p, 1u) + 1u == pr.ntiles) ? 1u : 0u;
            }
        }
        last = __shfl_sync(0xFFFFFFFFu, last, 0);
        if (last) {  // this warp delivered the pair's last tile: finish both sequences
            unsigned long long fka = keyA, fkb = keyB;
            if (pr.ntiles != 1) {
                if (lane == 0) {
                    if (kFabric) {
                        if (home_is_me) __threadfence(); else __threadfence_system();
                        fka = atomicMax_system(keyA_p, 0ull);
                        if (has_b) fkb = atomicMax_system(keyB_p, 0ull);
                    } else {
                        __threadfence();
                        fka = atomicMax(keyA_p, 0ull);
                        if (has_b) fkb = atomicMax(keyB_p, 0ull);
                    }
                    // leave the scratch zeroed for the next launch, and make sure it IS zero before the rows
                    // below can tell anybody that this solve is over
                    *keyA_p = 0ull;
                    if (has_b) *keyB_p = 0ull;
                    *cnt_p = 0u;
                    if (kFabric) { if (home_is_me) __threadfence(); else __threadfence_system(); }
                }
                fka = __shfl_sync(0xFFFFFFFFu, fka, 0);
                fkb = __shfl_sync(0xFFFFFFFFu, fkb, 0);
            }
            for (int half = 0; half < (has_b ? 2 : 1); ++half) {
                const uint32_t sidx = half == 0 ? ia : ib;
                const unsigned long long fk = half == 0 ? fka : fkb;
                const int nd = half == 0 ? ndiag_a : ndiag_b;
                const int bestv = (int)((uint32_t)(fk >> 32) ^ 0x80000000u);
                const uint32_t nwin = 0xFFFFFFFFu - (uint32_t)(fk & 0xFFFFFFFFull);
                const int kres = nd == 0 ? 0 : warp_first_k(lutlo, half == 0 ? gA : gB, seq1s + nwin, half == 0 ? pr.len_a : pr.len_b, bestv, vmin, lane);
                if (lane == 0) {
                    if (kFabric) post_result(p, sidx, bestv, nwin, (uint32_t)kres);
                    else out[sidx] = make_int4(bestv, (int)nwin, kres, 0);
                }
            }
        }
        if (kFabric && p.timeline && lane == 0) {  // debug: longest item overall (20) and of the last quarter (21), ns
            unsigned long long* const times = reinterpret_cast<unsigned long long*>(my_ctl + kCtlTimes);
            const unsigned long long dt = global_ns() - t_item0;
            if (atomicMax(&times[20], dt) < dt)  // (racy, debug only) who ran the longest item: start us | item | warp | did a copy before
                times[23] = ((t_item0 - times[0]) / 1000ull) << 40 | (unsigned long long)(item - item_begin) << 16 | (unsigned long long)warp << 8 | (blockIdx.x & 1u) << 1 | dbg_copied;
            if (item - item_begin >= ((item_end - item_begin) / 4u) * 3u) atomicMax(&times[21], dt);
            atomicAdd(&times[3], 1ull);                          // items executed
            atomicAdd(&times[22], (unsigned long long)rows);     // rows walked
        }
    }
    if (kFabric && p.timeline && lane == 0) {  // debug: first (18) and last (19) warp to run out of items
        unsigned long long* const times = reinterpret_cast<unsigned long long*>(my_ctl + kCtlTimes);
        const unsigned long long t = global_ns();
        atomicMin(&times[18], t);
        atomicMax(&times[19], t);
    }
    __syncthreads();
    if (!epilogue_later) launch_epilogue<kFabric>(p);
}

__global__ void __launch_bounds__(kThreads, 2) sa_score_stream_kernel(const KParams p) { stream_body<false>(p); }

// One fabric kernel for both modes: ranks other than 0 launch before they know anything about
// the problem, so the mode is read from the header packet. hdr0: rank 0's host copy of the header.
__global__ void __launch_bounds__(kThreads, 2) sa_score_fabric_kernel(const __grid_constant__ KParams p, const __grid_constant__ sa_blob_header hdr0) {
    unsigned long long* const times = reinterpret_cast<unsigned long long*>(reinterpret_cast<uint32_t*>(p.peer[p.rank] + p.off_ctrl) + kCtlTimes);
    if (threadIdx.x == 0) {
        s_abort = 0u;
        s_all_here = 0u;
        if (p.timeline) atomicMin(&times[0], global_ns());
    }
    __syncthreads();
    fabric_prologue(p, hdr0);
    if (p.timeline && threadIdx.x == 0) atomicMax(&times[1], global_ns());
    if (s_abort == 0u && (s_hdr.flags & SA_FLAG_STREAM_MODE))
        stream_body<true>(p, true);
    else
        score_body<true>(p, true);
    launch_epilogue<true>(p);
}

// ------------------------------------------------------------------ host side

std::atomic<uint64_t> g_launches{0};

struct DeviceInfo {
    bool ready = false;
    int sm_count = 0;
    int max_optin_smem = 0;
    int smem_set = 0;
    int smem_fabric_set = 0;
    int smem_pair_set = 0;
    int smem_stream_set = 0;
};
std::mutex g_mu;
DeviceInfo g_dev[64];

int fail(cudaError_t e, const char* what) {
    std::string msg = std::string(what) + ": " + cudaGetErrorString(e);
    sa_set_last_error(msg.c_str());
    return SA_ERR_CUDA;
}

#define SA_CUDA_TRY(expr)                                  \
    do {                                                   \
        cudaError_t e__ = (expr);                          \
        if (e__ != cudaSuccess) return fail(e__, #expr);   \
    } while (0)

inline size_t round_up_sz(size_t v, size_t m) { return (v + m - 1) / m * m; }

size_t key_bytes(uint32_t nseq) { return round_up_sz((size_t)nseq * 8, 256); }
size_t cnt_bytes(uint32_t nseq) { return round_up_sz((size_t)nseq * 4, 256); }

uint32_t smem_bytes_for(uint32_t s2_cap, uint32_t win_cap) {
    return kLutBytes + s2_cap + win_cap + (kSumWords + kMaxWords) * 4 + kWarps * 8 + 8 + kWarps * 4 + 16;
}

int device_info(int dev, DeviceInfo** out) {
    if (dev < 0 || dev >= 64) {
        sa_set_last_error("device index out of range");
        return SA_ERR_ARG;
    }
    std::lock_guard<std::mutex> lock(g_mu);
    DeviceInfo& d = g_dev[dev];
    if (!d.ready) {
        SA_CUDA_TRY(cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, dev));
        SA_CUDA_TRY(cudaDeviceGetAttribute(&d.max_optin_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
        d.ready = true;
    }
    *out = &d;
    return SA_OK;
}

}  // namespace

extern "C" {

size_t sa_cuda_scratch_bytes(uint32_t nseq) { return key_bytes(nseq) + cnt_bytes(nseq) + 256; }

int sa_cuda_scratch_init(void* d_scratch, size_t bytes, void* stream) {
    SA_CUDA_TRY(cudaMemsetAsync(d_scratch, 0, bytes, static_cast<cudaStream_t>(stream)));
    return SA_OK;
}

namespace {
int solve_blob_impl(const void* d_blob, const sa_blob_header* h, int32_t* d_out, void* d_scratch, void* stream, uint32_t* done_flag,
                    uint32_t done_epoch);
}

int sa_cuda_solve_blob(const void* d_blob, const sa_blob_header* h, int32_t* d_out, void* d_scratch, void* stream) {
    return solve_blob_impl(d_blob, h, d_out, d_scratch, stream, nullptr, 0u);
}

namespace {
// done_flag: see KParams (TILE and STREAM launches only; the PAIR kernel ignores it, so callers pass none there)
int solve_blob_impl(const void* d_blob, const sa_blob_header* h, int32_t* d_out, void* d_scratch, void* stream, uint32_t* done_flag,
                    uint32_t done_epoch) {
    if (!d_blob || !h || !d_out || !d_scratch) {
        sa_set_last_error("sa_cuda_solve_blob: null argument");
        return SA_ERR_ARG;
    }
    if (h->magic != SA_BLOB_MAGIC || h->version != SA_BLOB_VERSION) {
        sa_set_last_error("sa_cuda_solve_blob: header is not a blob of this version");
        return SA_ERR_ARG;
    }
    if (h->nseq == 0) return SA_OK;
    int dev = 0;
    SA_CUDA_TRY(cudaGetDevice(&dev));
    DeviceInfo* di = nullptr;
    int rc = device_info(dev, &di);
    if (rc != SA_OK) return rc;

    KParams p{};
    p.blob = static_cast<const unsigned char*>(d_blob);
    p.out = reinterpret_cast<int4*>(d_out);
    unsigned char* sc = static_cast<unsigned char*>(d_scratch);
    p.key = reinterpret_cast<unsigned long long*>(sc);
    p.cnt = reinterpret_cast<unsigned int*>(sc + key_bytes(h->nseq));
    p.ctrl = reinterpret_cast<unsigned int*>(sc + key_bytes(h->nseq) + cnt_bytes(h->nseq));
    p.item_begin = 0;
    p.item_end = h->nitems;
    p.h_off_lut = h->off_lut;
    p.h_off_seq1 = h->off_seq1;
    p.h_off_seqtab = h->off_seqtab;
    p.h_off_items = h->off_items;
    p.h_nseq = h->nseq;
    p.h_off_tile_items = h->off_tile_items;
    p.s2_cap = (h->max_rows_padded + 15u) & ~15u;
    p.win_cap = (h->max_window_bytes + 15u) & ~15u;
    p.write_empty = 1;
    static const bool env_timeline = std::getenv("SA_TIMELINE") != nullptr;  // debug: phase timestamps in the control words
    p.timeline = env_timeline ? 1u : 0u;
    p.done_flag = done_flag;
    p.done_epoch = done_epoch;
    if (h->flags & SA_FLAG_STREAM_MODE) {
        const uint32_t smem = 2u * kLutBytes + ((h->len1_padded + 15u) & ~15u) +
                              (uint32_t)kWarps * SA_STREAM_ROW_BLOCK * SA_ALPHABET * 4u + 64u;
        if ((int)smem > di->max_optin_smem) {
            sa_set_last_error("sa_cuda_solve_blob: Seq1 too long for the stream kernel's shared memory");
            return SA_ERR_UNSUPPORTED;
        }
        {
            std::lock_guard<std::mutex> lock(g_mu);
            if ((int)smem > di->smem_stream_set) {
                SA_CUDA_TRY(cudaFuncSetAttribute(sa_score_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                di->smem_stream_set = (int)smem;
            }
        }
        static const uint32_t env_cps = [] { const char* e = std::getenv("SA_STREAM_CTAS_PER_SM"); return e ? (uint32_t)std::atoi(e) : 2u; }();
        uint32_t grid = (uint32_t)di->sm_count * (env_cps == 1u ? 1u : 2u);
        const uint32_t need = (h->nitems + kWarps - 1u) / kWarps;
        if (need < grid) grid = need ? need : 1u;
        sa_score_stream_kernel<<<grid, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(p);
        SA_CUDA_TRY(cudaGetLastError());
        g_launches.fetch_add(1, std::memory_order_relaxed);
        return SA_OK;
    }
    if (h->flags & SA_FLAG_PAIR_MODE) {
        const uint32_t smem = kLutBytes + ((h->len1_padded + 15u) & ~15u) + 2u * p.s2_cap + p.s2_cap * SA_ALPHABET * 4u +
                              (kSumWords + kMaxWords) * 4u + 2u * kWarps * 8u + 8u + 2u * kWarps * 4u + 16u;
        if ((int)smem > di->max_optin_smem) {
            sa_set_last_error("sa_cuda_solve_blob: pair profile needs more shared memory than the device has");
            return SA_ERR_UNSUPPORTED;
        }
        {
            std::lock_guard<std::mutex> lock(g_mu);
            if ((int)smem > di->smem_pair_set) {
                SA_CUDA_TRY(cudaFuncSetAttribute(sa_score_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                di->smem_pair_set = (int)smem;
            }
        }
        const uint32_t per_sm = smem + 1024u <= (uint32_t)di->max_optin_smem / 2u ? 2u : 1u;
        uint32_t grid = (uint32_t)di->sm_count * per_sm;
        if (h->nitems < grid) grid = h->nitems ? h->nitems : 1u;
        sa_score_pair_kernel<<<grid, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(p);
        SA_CUDA_TRY(cudaGetLastError());
        g_launches.fetch_add(1, std::memory_order_relaxed);
        return SA_OK;
    }
    const uint32_t smem = smem_bytes_for(p.s2_cap, p.win_cap);
    if ((int)smem > di->max_optin_smem) {
        sa_set_last_error("sa_cuda_solve_blob: problem needs more shared memory than the device has");
        return SA_ERR_UNSUPPORTED;
    }
    {
        std::lock_guard<std::mutex> lock(g_mu);
        if ((int)smem > di->smem_set) {
            SA_CUDA_TRY(cudaFuncSetAttribute(sa_score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            di->smem_set = (int)smem;
        }
    }
    uint32_t grid = (uint32_t)di->sm_count * 2u;
    if (h->nitems < grid) grid = h->nitems ? h->nitems : 1u;
    sa_score_kernel<<<grid, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(p);
    SA_CUDA_TRY(cudaGetLastError());
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return SA_OK;
}
}  // namespace

int sa_cuda_solve_problem(const sa_problem* prob, int32_t* out, int device) {
    if (!prob || !out) {
        sa_set_last_error("sa_cuda_solve_problem: null argument");
        return SA_ERR_ARG;
    }
    void* blob = nullptr;
    size_t bytes = 0;
    int rc = sa_plan_pack(prob, 1, &blob, &bytes);
    if (rc != SA_OK) return rc;
    const sa_blob_header* h = static_cast<const sa_blob_header*>(blob);
    const uint32_t nseq = h->nseq;
    if (nseq == 0) {
        sa_blob_free(blob);
        return SA_OK;
    }
    void *d_blob = nullptr, *d_out = nullptr, *d_scratch = nullptr;
    int4* h_out = nullptr;
    auto cleanup = [&]() {
        cudaFree(d_blob);
        cudaFree(d_out);
        cudaFree(d_scratch);
        free(h_out);
        sa_blob_free(blob);
    };
    cudaError_t e;
#define SA_TRY_CLEAN(expr)                  \
    do {                                    \
        e = (expr);                         \
        if (e != cudaSuccess) {             \
            cleanup();                      \
            return fail(e, #expr);          \
        }                                   \
    } while (0)
    SA_TRY_CLEAN(cudaSetDevice(device));
    const size_t scratch = sa_cuda_scratch_bytes(nseq);
    SA_TRY_CLEAN(cudaMalloc(&d_blob, bytes));
    SA_TRY_CLEAN(cudaMalloc(&d_out, (size_t)nseq * 16));
    SA_TRY_CLEAN(cudaMalloc(&d_scratch, scratch));
    SA_TRY_CLEAN(cudaMemcpy(d_blob, blob, bytes, cudaMemcpyHostToDevice));
    SA_TRY_CLEAN(cudaMemset(d_scratch, 0, scratch));
    rc = sa_cuda_solve_blob(d_blob, h, static_cast<int32_t*>(d_out), d_scratch, nullptr);
    if (rc != SA_OK) {
        cleanup();
        return rc;
    }
    h_out = static_cast<int4*>(malloc((size_t)nseq * 16));
    if (!h_out) {
        cleanup();
        return SA_ERR_NOMEM;
    }
    SA_TRY_CLEAN(cudaMemcpy(h_out, d_out, (size_t)nseq * 16, cudaMemcpyDeviceToHost));
#undef SA_TRY_CLEAN
    for (uint32_t s = 0; s < nseq; ++s) {
        out[3 * s] = h_out[s].x;
        out[3 * s + 1] = h_out[s].y;
        out[3 * s + 2] = h_out[s].z;
    }
    cleanup();
    return SA_OK;
}

uint64_t sa_cuda_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

// ------------------------------------------------------------------ native solver session
//
// The whole per-solve host path in one native call: plan + pack straight into pinned memory,
// one async H2D, one launch, one async D2H, one stream synchronise. This is the reference's
// `send_divided_Seq2_To_Cuda` (cudaFunctions.cu:178-242) minus its 4+N cudaMallocs, N device
// syncs and 5 cudaFrees per call: all buffers belong to the session and only ever grow.
struct sa_session {
    int device = 0;
    cudaStream_t stream = nullptr;
    unsigned char* h_in = nullptr;   // pinned
    size_t h_in_cap = 0;
    int4* h_out = nullptr;           // pinned
    size_t h_out_cap = 0;            // sequences
    unsigned char* d_blob = nullptr;
    size_t d_blob_cap = 0;
    int4* d_out = nullptr;
    size_t d_out_cap = 0;
    unsigned char* d_scratch = nullptr;
    size_t d_scratch_cap = 0;        // sequences
    uint64_t h2d_bytes = 0, d2h_bytes = 0;
    size_t zero_copy_max = 0;          // blobs up to this size are read in place from pinned memory (measured slower: off)
    bool direct_out = true;            // result rows are written straight into pinned memory
    size_t pipe_min_seq = 16384;       // batches at least this large are packed and launched in ranges of >= half of it
                                       // (8192 sequences keep a range above AUTO's threshold for the STREAM kernel)
    uint32_t last_launches = 0;
    uint32_t* h_flag = nullptr;          // pinned: the kernel's "every row is visible" word (see KParams::done_flag)
    uint32_t flag_epoch = 0;
    size_t flag_max_seq = 4096;          // problems up to this size are waited for by polling the flag
    cudaStream_t copy_stream = nullptr;  // pipelined path: range copies run beside the previous range's kernel
    cudaEvent_t copied[8] = {};
};

int sa_session_create(int device, sa_session** out) {
    if (!out) {
        sa_set_last_error("sa_session_create: null argument");
        return SA_ERR_ARG;
    }
    SA_CUDA_TRY(cudaSetDevice(device));
    sa_session* s = new (std::nothrow) sa_session();
    if (!s) return SA_ERR_NOMEM;
    s->device = device;
    if (const char* v = std::getenv("SA_PIPE_MIN_SEQ")) s->pipe_min_seq = (size_t)std::atoll(v);  // 0 = never pipeline
    if (const char* v = std::getenv("SA_FLAG_MAX_SEQ")) s->flag_max_seq = (size_t)std::atoll(v);  // 0 = always wait on the stream
    if (cudaMallocHost(reinterpret_cast<void**>(&s->h_flag), 64) == cudaSuccess) *s->h_flag = 0u;
    else s->h_flag = nullptr, cudaGetLastError();
    cudaError_t e = cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        delete s;
        return fail(e, "cudaStreamCreateWithFlags");
    }
    *out = s;
    return SA_OK;
}

void sa_session_destroy(sa_session* s) {
    if (!s) return;
    cudaSetDevice(s->device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    cudaFreeHost(s->h_in);
    cudaFreeHost(s->h_out);
    cudaFreeHost(s->h_flag);
    cudaFree(s->d_blob);
    cudaFree(s->d_out);
    cudaFree(s->d_scratch);
    for (cudaEvent_t e : s->copied)
        if (e) cudaEventDestroy(e);
    if (s->copy_stream) cudaStreamDestroy(s->copy_stream);
    if (s->stream) cudaStreamDestroy(s->stream);
    delete s;
}

namespace {
int session_reserve(sa_session* s, size_t blob_bytes, size_t nseq) {
    if (blob_bytes > s->h_in_cap) {
        const size_t cap = std::max(blob_bytes, std::max<size_t>(2 * s->h_in_cap, 1 << 16));
        cudaFreeHost(s->h_in);
        s->h_in = nullptr;
        s->h_in_cap = 0;
        SA_CUDA_TRY(cudaMallocHost(reinterpret_cast<void**>(&s->h_in), cap));
        s->h_in_cap = cap;
    }
    if (blob_bytes > s->d_blob_cap) {
        const size_t cap = std::max(blob_bytes, std::max<size_t>(2 * s->d_blob_cap, 1 << 16));
        cudaFree(s->d_blob);
        s->d_blob = nullptr;
        s->d_blob_cap = 0;
        SA_CUDA_TRY(cudaMalloc(reinterpret_cast<void**>(&s->d_blob), cap));
        s->d_blob_cap = cap;
    }
    if (nseq > s->h_out_cap) {
        const size_t cap = std::max(nseq, std::max<size_t>(2 * s->h_out_cap, 64));
        cudaFreeHost(s->h_out);
        s->h_out = nullptr;
        s->h_out_cap = 0;
        SA_CUDA_TRY(cudaMallocHost(reinterpret_cast<void**>(&s->h_out), cap * sizeof(int4)));
        s->h_out_cap = cap;
    }
    if (nseq > s->d_out_cap) {
        const size_t cap = std::max(nseq, std::max<size_t>(2 * s->d_out_cap, 64));
        cudaFree(s->d_out);
        s->d_out = nullptr;
        s->d_out_cap = 0;
        SA_CUDA_TRY(cudaMalloc(reinterpret_cast<void**>(&s->d_out), cap * sizeof(int4)));
        s->d_out_cap = cap;
    }
    if (nseq > s->d_scratch_cap) {
        const size_t cap = std::max(nseq, std::max<size_t>(2 * s->d_scratch_cap, 64));
        cudaFree(s->d_scratch);
        s->d_scratch = nullptr;
        s->d_scratch_cap = 0;
        const size_t bytes = sa_cuda_scratch_bytes((uint32_t)cap);
        SA_CUDA_TRY(cudaMalloc(reinterpret_cast<void**>(&s->d_scratch), bytes));
        SA_CUDA_TRY(cudaMemsetAsync(s->d_scratch, 0, bytes, s->stream));
        s->d_scratch_cap = cap;
    }
    return SA_OK;
}

// Large batches: the host needs longer to pack the problem than the GPU needs to score it (input3 x 1024:
// 4.3 ms against 1.3 ms on a B200 box, and most of the 4.3 is single-threaded planning), so the batch is cut
// into ranges of sequences. A few host threads plan and pack one range each, concurrently, every range into
// its own slot of the pinned and of the device buffer; whoever finishes a range issues its copy and its
// launch on the session's stream (in whatever order: rows are disjoint, the self-cleaning scratch is
// shared by launches that run one after the other). Packing overlaps packing, copies and kernels.
int session_solve_pipelined(sa_session* s, const sa_problem* prob, int32_t* out3, size_t nseq) {
    const size_t nranges = std::min<size_t>(8, std::max<size_t>(2, nseq / std::max<size_t>(1, s->pipe_min_seq / 2)));
    const unsigned nthreads = (unsigned)std::min<size_t>(nranges, std::max(1u, std::min(4u, std::thread::hardware_concurrency() / 2)));
    int rc = session_reserve(s, s->h_in_cap, nseq);  // result table and scratch for the whole batch
    if (rc != SA_OK) return rc;
    if (!s->copy_stream) SA_CUDA_TRY(cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking));
    for (size_t k = 0; k < nranges; ++k)
        if (!s->copied[k]) SA_CUDA_TRY(cudaEventCreateWithFlags(&s->copied[k], cudaEventDisableTiming));
    auto rezero = [&]() {
        cudaStreamSynchronize(s->copy_stream);
        cudaStreamSynchronize(s->stream);
        cudaMemsetAsync(s->d_scratch, 0, sa_cuda_scratch_bytes((uint32_t)s->d_scratch_cap), s->stream);
    };
    int32_t* const out_dev = reinterpret_cast<int32_t*>(s->direct_out ? s->h_out : s->d_out);
    for (int attempt = 0; attempt < 2; ++attempt) {
        const size_t slot = (std::min(s->h_in_cap, s->d_blob_cap) / nranges) & ~(size_t)255;
        std::atomic<size_t> next{0};
        std::atomic<size_t> need_max{0};  // a range that did not fit its slot: the largest size asked for
        std::atomic<int> failed{SA_OK};
        std::mutex launch_mutex, err_mutex;
        std::string err_text;
        uint64_t h2d = 0;
        uint32_t launches = 0;
        auto worker = [&]() {
            if (cudaSetDevice(s->device) != cudaSuccess) {
                failed.store(SA_ERR_CUDA);
                return;
            }
            for (;;) {
                const size_t k = next.fetch_add(1);
                if (k >= nranges || failed.load() != SA_OK) return;
                const uint32_t first = (uint32_t)(nseq * k / nranges), count = (uint32_t)(nseq * (k + 1) / nranges) - first;
                unsigned char* const h_blob = s->h_in + k * slot;
                size_t got = 0;
                int r = sa_plan_pack_range(prob, first, count, 1, SA_MODE_AUTO, slot ? h_blob : nullptr, slot, &got);
                if (r == SA_OK && !slot) r = SA_ERR_SMALL_BUFFER;
                if (r == SA_ERR_SMALL_BUFFER) {
                    size_t cur = need_max.load();
                    while (got > cur && !need_max.compare_exchange_weak(cur, got)) {
                    }
                    continue;  // keep sizing the other ranges; nothing more is launched in this attempt
                }
                if (r == SA_OK && need_max.load() != 0) continue;
                if (r == SA_OK) {
                    std::lock_guard<std::mutex> g(launch_mutex);
                    unsigned char* const d_blob = s->d_blob + k * slot;
                    // the copy goes to its own stream, so that it runs beside the kernel of an earlier range
                    cudaError_t e = cudaMemcpyAsync(d_blob, h_blob, got, cudaMemcpyHostToDevice, s->copy_stream);
                    if (e == cudaSuccess) e = cudaEventRecord(s->copied[k], s->copy_stream);
                    if (e == cudaSuccess) e = cudaStreamWaitEvent(s->stream, s->copied[k], 0);
                    if (e != cudaSuccess) {
                        r = fail(e, "cudaMemcpyAsync(range)");
                    } else {
                        h2d += got;
                        const sa_blob_header* h = reinterpret_cast<const sa_blob_header*>(h_blob);  // read at launch time only
                        if (h->nseq) {
                            r = sa_cuda_solve_blob(d_blob, h, out_dev + 4 * (size_t)first, s->d_scratch, s->stream);
                            if (r == SA_OK) ++launches;
                        }
                    }
                }
                if (r != SA_OK) {
                    std::lock_guard<std::mutex> g(err_mutex);
                    if (failed.load() == SA_OK) {
                        err_text = sa_last_error();  // the message lives in this thread: hand it to the caller's
                        failed.store(r);
                    }
                    return;
                }
            }
        };
        std::vector<std::thread> pool;
        for (unsigned t = 1; t < nthreads; ++t) pool.emplace_back(worker);
        worker();
        for (auto& th : pool) th.join();
        if (failed.load() != SA_OK) {
            rezero();
            sa_set_last_error(err_text.c_str());
            return failed.load();  // SA_ERR_UNSUPPORTED: the caller falls back to the CPU for the whole batch
        }
        if (need_max.load() != 0) {
            // first large batch of this session (or a larger one than before): grow the buffers — which frees
            // them, so nothing may be in flight — and start over
            cudaError_t e = cudaStreamSynchronize(s->copy_stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
            if (e != cudaSuccess) return fail(e, "cudaStreamSynchronize");
            rc = session_reserve(s, nranges * ((need_max.load() + need_max.load() / 8 + 511) & ~(size_t)255), nseq);
            if (rc != SA_OK) return rc;
            continue;
        }
        cudaError_t e = cudaSuccess;
        if (!s->direct_out) e = cudaMemcpyAsync(s->h_out, s->d_out, nseq * sizeof(int4), cudaMemcpyDeviceToHost, s->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
        if (e != cudaSuccess) {
            rezero();
            return fail(e, "cudaStreamSynchronize");
        }
        s->h2d_bytes = h2d;
        s->d2h_bytes = nseq * sizeof(int4);
        s->last_launches = launches;
        for (size_t i = 0; i < nseq; ++i) {
            out3[3 * i] = s->h_out[i].x;
            out3[3 * i + 1] = s->h_out[i].y;
            out3[3 * i + 2] = s->h_out[i].z;
        }
        return SA_OK;
    }
    sa_set_last_error("sa_session_solve: a range did not fit the buffers sized for it");
    return SA_ERR_NOMEM;
}
}  // namespace

int sa_session_solve(sa_session* s, const sa_problem* prob, int32_t* out3) {
    if (!s || !prob || !out3) {
        sa_set_last_error("sa_session_solve: null argument");
        return SA_ERR_ARG;
    }
    SA_CUDA_TRY(cudaSetDevice(s->device));
    const size_t nseq = (size_t)sa_problem_nseq(prob);
    if (nseq == 0) return SA_OK;
    if (s->pipe_min_seq && nseq >= s->pipe_min_seq) return session_solve_pipelined(s, prob, out3, nseq);
    s->last_launches = 1;
    // plan + pack once, straight into the pinned buffer; only a buffer that turns out too small
    // costs a second pass
    size_t got = 0;
    int rc = s->h_in ? sa_plan_pack_into(prob, 1, s->h_in, s->h_in_cap, &got) : SA_ERR_SMALL_BUFFER;
    if (rc == SA_ERR_SMALL_BUFFER) {
        if (!s->h_in) {
            rc = sa_plan_pack_into(prob, 1, nullptr, 0, &got);
            if (rc != SA_OK) return rc;
        }
        rc = session_reserve(s, got, nseq);
        if (rc != SA_OK) return rc;
        rc = sa_plan_pack_into(prob, 1, s->h_in, s->h_in_cap, &got);
    }
    if (rc != SA_OK) return rc;
    rc = session_reserve(s, got, nseq);
    if (rc != SA_OK) return rc;
    const sa_blob_header* h = reinterpret_cast<const sa_blob_header*>(s->h_in);
    // One DMA copy in; the finalising CTAs post their result rows straight into pinned host
    // memory (no copy back: -16 us on input1, -2 us on input3 on B200). Reading the blob in place
    // from pinned memory is possible (zero_copy_max) but measured 1.5-2.5x slower, so it is off.
    const unsigned char* blob_dev = s->h_in;
    if (got > s->zero_copy_max) {
        SA_CUDA_TRY(cudaMemcpyAsync(s->d_blob, s->h_in, got, cudaMemcpyHostToDevice, s->stream));
        blob_dev = s->d_blob;
    }
    int32_t* out_dev = reinterpret_cast<int32_t*>(s->direct_out ? s->h_out : s->d_out);
    // The scratch is carved by the CURRENT nseq and is only correct while the whole buffer is zero between
    // solves; a launch that did not run to completion leaves dirty words, so any failure re-zeroes it.
    auto rezero = [&]() { cudaMemsetAsync(s->d_scratch, 0, sa_cuda_scratch_bytes((uint32_t)s->d_scratch_cap), s->stream); };
    // Small problems: the kernel's last CTA tells the host through a pinned word that every row is visible, and
    // the host polls that word — it sees it several microseconds before cudaStreamSynchronize would return.
    const bool use_flag = s->direct_out && s->h_flag && nseq <= s->flag_max_seq && !(h->flags & SA_FLAG_PAIR_MODE);
    const uint32_t epoch = use_flag ? ++s->flag_epoch : 0u;
    rc = solve_blob_impl(blob_dev, h, out_dev, s->d_scratch, s->stream, use_flag ? s->h_flag : nullptr, epoch);
    if (rc != SA_OK) {
        rezero();
        return rc;
    }
    if (!s->direct_out) {
        const cudaError_t e = cudaMemcpyAsync(s->h_out, s->d_out, nseq * sizeof(int4), cudaMemcpyDeviceToHost, s->stream);
        if (e != cudaSuccess) {
            rezero();
            return fail(e, "cudaMemcpyAsync(results)");
        }
    }
    bool seen = false;
    if (use_flag) {
        const volatile uint32_t* flag = s->h_flag;
        const auto t0 = std::chrono::steady_clock::now();
        for (uint32_t spins = 1; !(seen = (*flag == epoch)); ++spins) {
            __builtin_ia32_pause();
            if ((spins & 2047u) == 0u) {  // every few tens of microseconds: did the launch die, or is this taking long?
                const cudaError_t q = cudaStreamQuery(s->stream);
                if (q != cudaErrorNotReady) break;  // finished (flag is set by now) or failed: the synchronize below tells
                if (std::chrono::steady_clock::now() - t0 > std::chrono::milliseconds(2)) break;  // not a small problem after all
            }
        }
    }
    if (!seen) {
        const cudaError_t e = cudaStreamSynchronize(s->stream);
        if (e != cudaSuccess) {
            rezero();
            return fail(e, "cudaStreamSynchronize");
        }
    }
    s->h2d_bytes = got;
    s->d2h_bytes = nseq * sizeof(int4);
    for (size_t i = 0; i < nseq; ++i) {
        out3[3 * i] = s->h_out[i].x;
        out3[3 * i + 1] = s->h_out[i].y;
        out3[3 * i + 2] = s->h_out[i].z;
    }
    return SA_OK;
}

void sa_session_configure(sa_session* s, size_t zero_copy_max_bytes, int direct_out) {
    if (!s) return;
    s->zero_copy_max = zero_copy_max_bytes;
    s->direct_out = direct_out != 0;
}

void sa_session_set_pipeline(sa_session* s, size_t min_seq) {
    if (s) s->pipe_min_seq = min_seq;
}

uint32_t sa_session_last_launches(const sa_session* s) { return s ? s->last_launches : 0; }

void sa_session_last_bytes(const sa_session* s, uint64_t* h2d, uint64_t* d2h) {
    if (h2d) *h2d = s ? s->h2d_bytes : 0;
    if (d2h) *d2h = s ? s->d2h_bytes : 0;
}

// ------------------------------------------------------------------ fabric (multi-GPU) launch

int sa_fabric_layout(uint32_t max_seq, size_t max_blob_bytes, sa_fabric_layout_t* out) {
    if (!out || max_seq == 0) {
        sa_set_last_error("sa_fabric_layout: bad argument");
        return SA_ERR_ARG;
    }
    size_t off = 0;
    out->off_ctrl = (uint32_t)off;
    off += round_up_sz((kCtlFlags + kMaxChunks) * 4, 256);
    out->off_hdr_ll = (uint32_t)off;  // header packet: one (epoch | word) pair per header word
    off += round_up_sz(kHdrWords * 8, 256);
    out->off_key = (uint32_t)off;
    off += key_bytes(max_seq);
    out->off_cnt = (uint32_t)off;
    off += cnt_bytes(max_seq);
    out->off_out = (uint32_t)off;
    off += round_up_sz((size_t)max_seq * 16, 256);
    out->off_out_ll = (uint32_t)off;  // result rows as they arrive: two self-validating words each
    off += round_up_sz((size_t)max_seq * 16, 256);
    out->off_blob = (uint32_t)off;
    off += round_up_sz(max_blob_bytes, 256);
    if (off >= 0xFFFFFFFFull) {
        sa_set_last_error("sa_fabric_layout: symmetric buffer would exceed 4 GiB");
        return SA_ERR_UNSUPPORTED;
    }
    out->total_bytes = off;
    out->max_seq = max_seq;
    out->max_blob_bytes = max_blob_bytes;
    return SA_OK;
}

int sa_cuda_fabric_solve(const sa_fabric_desc* f, const sa_blob_header* header, void* stream) {
    if (!f || f->world < 1 || f->world > SA_MAX_WORLD || f->rank >= f->world) {
        sa_set_last_error("sa_cuda_fabric_solve: bad descriptor");
        return SA_ERR_ARG;
    }
    if (f->rank == 0) {
        if (!header || header->magic != SA_BLOB_MAGIC || header->version != SA_BLOB_VERSION) {
            sa_set_last_error("sa_cuda_fabric_solve: rank 0 must pass the host copy of the blob header");
            return SA_ERR_ARG;
        }
        if (header->world != f->world) {
            sa_set_last_error("sa_cuda_fabric_solve: blob was planned for a different number of GPUs");
            return SA_ERR_ARG;
        }
        if (header->flags & SA_FLAG_PAIR_MODE) {
            sa_set_last_error("sa_cuda_fabric_solve: PAIR-mode blobs are single-GPU; plan with SA_MODE_AUTO, TILE or STREAM");
            return SA_ERR_ARG;
        }
        if (header->nseq > f->layout.max_seq || header->total_bytes > f->layout.max_blob_bytes ||
            header->nchunks > kMaxChunks || header->max_rows_padded > SA_GPU_MAX_SEQ2 + 128u ||
            ((header->flags & SA_FLAG_STREAM_MODE) && ((header->len1_padded + 15u) & ~15u) > SA_STREAM_MAX_SEQ1)) {
            sa_set_last_error("sa_cuda_fabric_solve: problem exceeds the capacity this fabric was created with");
            return SA_ERR_TOO_LARGE;
        }
    }
    int dev = 0;
    SA_CUDA_TRY(cudaGetDevice(&dev));
    DeviceInfo* di = nullptr;
    int rc = device_info(dev, &di);
    if (rc != SA_OK) return rc;

    KParams p{};
    for (uint32_t r = 0; r < f->world; ++r) {
        if (!f->peer_base[r]) {
            sa_set_last_error("sa_cuda_fabric_solve: missing peer mapping");
            return SA_ERR_ARG;
        }
        p.peer[r] = static_cast<unsigned char*>(f->peer_base[r]);
    }
    p.mc = static_cast<unsigned char*>(f->mc_base);
    p.use_mc = (f->use_multicast && f->mc_base) ? 1u : 0u;
    p.rank = f->rank;
    p.world = f->world;
    p.epoch = f->epoch;
    p.off_ctrl = f->layout.off_ctrl;
    p.off_hdr_ll = f->layout.off_hdr_ll;
    p.off_out_ll = f->layout.off_out_ll;
    p.off_key = f->layout.off_key;
    p.off_cnt = f->layout.off_cnt;
    p.off_out = f->layout.off_out;
    p.off_blob = f->layout.off_blob;
    p.timeout_ns = (unsigned long long)(f->timeout_ms ? f->timeout_ms : 10000u) * 1000000ull;
    p.timeline = f->pad0 & 1u;
    p.skip_push = (f->pad0 >> 1) & 1u;
    static const uint32_t env_dbg_src = [] { const char* e = std::getenv("SA_DEBUG_SRC"); return e ? (uint32_t)std::atoi(e) : 0u; }();
    p.dbg_src = env_dbg_src;
    p.blob = p.peer[f->rank] + p.off_blob;
    p.write_empty = f->rank == 0 ? 1u : 0u;
    // Ranks other than 0 know nothing about the problem at launch time, so every rank launches the
    // same input-independent shape: full persistent grid, shared memory for the largest sequence.
    p.s2_cap = (SA_GPU_MAX_SEQ2 + 128u + 15u) & ~15u;
    p.win_cap = (SA_GPU_MAX_SEQ2 + 128u + 4096u + 16u + 15u) & ~15u;
    const uint32_t smem_stream = 2u * kLutBytes + SA_STREAM_MAX_SEQ1 + (uint32_t)kWarps * SA_STREAM_ROW_BLOCK * SA_ALPHABET * 4u + 64u;
    const uint32_t smem = std::max(smem_bytes_for(p.s2_cap, p.win_cap), smem_stream);
    {
        std::lock_guard<std::mutex> lock(g_mu);
        if ((int)smem > di->smem_fabric_set) {
            SA_CUDA_TRY(cudaFuncSetAttribute(sa_score_fabric_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            di->smem_fabric_set = (int)smem;
        }
    }
    static const uint32_t env_cps = [] { const char* e = std::getenv("SA_STREAM_CTAS_PER_SM"); return e ? (uint32_t)std::atoi(e) : 2u; }();
    const uint32_t grid = (uint32_t)di->sm_count * (env_cps == 1u ? 1u : 2u);
    static const sa_blob_header no_header{};  // ranks other than 0 receive the header through the fabric
    sa_score_fabric_kernel<<<grid, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(p, f->rank == 0 ? *header : no_header);
    SA_CUDA_TRY(cudaGetLastError());
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return SA_OK;
}

// ---------------------------------------------------------------------------------------
// Fixed-capacity variant for callers that must not allocate device memory.
//
// Who that is: code that runs after the host application has handed every byte of HBM to
// its own pool (an inference server's plug-in hook, a framework custom op executing under a
// caching allocator it must not bypass) and code on a latency-critical path where the
// device-wide synchronisation hidden inside cudaMalloc/cudaFree is not acceptable. Such a
// caller can still provide the problem blob and the result table (they are its inputs and
// outputs), but it has nowhere to get the solver's private scratch from.
//
// So the scratch (argmax keys, arrival counters, work-queue words) lives in the library
// image itself: a file-scope device array, one copy per GPU that loads this library, held
// until the process exits. It is sized for SA_NOALLOC_MAX_SEQ sequences — 49,408 bytes.
//
// The array's name below is only a host-side handle. Its device address is looked up once
// per device with cudaGetSymbolAddress while that device is current, and that address is what
// the kernel receives.
constexpr size_t kStaticScratchBytes =
    ((size_t)SA_NOALLOC_MAX_SEQ * 8 + 255) / 256 * 256 + ((size_t)SA_NOALLOC_MAX_SEQ * 4 + 255) / 256 * 256 + 256;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

namespace {
std::mutex g_static_mu;
void* g_static_scratch_addr[64] = {nullptr};
}  // namespace

int sa_cuda_solve_blob_noalloc(const void* d_blob, const sa_blob_header* header, int32_t* d_out, void* stream) {
    if (!d_blob || !header || !d_out) {
        sa_set_last_error("sa_cuda_solve_blob_noalloc: null argument");
        return SA_ERR_ARG;
    }
    // Validate the header before anything is enqueued: a garbage header must not be reported as "too large",
    // and nothing (not even the memset) may be launched for it.
    if (header->magic != SA_BLOB_MAGIC || header->version != SA_BLOB_VERSION) {
        sa_set_last_error("sa_cuda_solve_blob_noalloc: header is not a blob of this version");
        return SA_ERR_ARG;
    }
    if (header->nseq > SA_NOALLOC_MAX_SEQ) {
        char msg[160];
        std::snprintf(msg, sizeof(msg),
                      "sa_cuda_solve_blob_noalloc: %u sequences exceed the built-in scratch capacity of %u; "
                      "use sa_cuda_solve_blob with caller-provided scratch",
                      header->nseq, (unsigned)SA_NOALLOC_MAX_SEQ);
        sa_set_last_error(msg);
        return SA_ERR_TOO_LARGE;
    }
    int dev = 0;
    SA_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) {
        sa_set_last_error("sa_cuda_solve_blob_noalloc: device index out of range");
        return SA_ERR_ARG;
    }
    void* scratch = nullptr;
    {
        // First call on this device: resolve the array's device address (`dev` is current).
        std::lock_guard<std::mutex> lock(g_static_mu);
        if (!g_static_scratch_addr[dev]) {
            void* ptr = nullptr;
            SA_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_static_scratch));
            g_static_scratch_addr[dev] = ptr;
        }
        scratch = g_static_scratch_addr[dev];
    }
    // The array keeps whatever the previous call left. A call that ran to completion leaves
    // zeros, one that was torn down half-way does not, so clear the part this launch reads.
    const size_t used = sa_cuda_scratch_bytes(header->nseq);
    SA_CUDA_TRY(cudaMemsetAsync(scratch, 0, used, static_cast<cudaStream_t>(stream)));
    return sa_cuda_solve_blob(d_blob, header, d_out, scratch, stream);
}

}  // extern "C"
