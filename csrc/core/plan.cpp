// Planner + packer + CPU blob executor.
//
// What this replaces in the reference:
//   * main.c:110-121,184-185 — static "rows = N / P, remainder on root" split. Here the unit
//     of work is a (sequence, offset-tile) item whose cost is its padded cell count; whole
//     sequences (or, when sequences are scarce, single tiles) are dealt to the least-loaded
//     GPU, costliest first.
//   * cudaFunctions.cu:35-61 — four separate constant-memory uploads. Here weights, pair
//     table, Seq1, the per-sequence descriptors, the item list and every Seq2 are packed into
//     ONE blob: one H2D copy on a single GPU, one in-kernel broadcast on several.
//
// Algorithm the tiling serves: with V[i][j] = value(seq2[i], seq1[j]) and
// P_n[k] = sum_{i<k} V[i][i+n],
//     score(n,0) = P_n[m]                       (no gap)
//     score(n,k) = P_{n+1}[m] + (P_n[k] - P_{n+1}[k])   for 1 <= k < m
// so per offset n the best k is a running max of the prefix difference of two neighbouring
// diagonals — O(len1*len2) instead of the reference's O(len1*len2^2).
#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <cstdio>
#include <climits>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <thread>
#include <pthread.h>

#include "core.hpp"

static_assert(sizeof(sa_blob_header) == 1024, "blob header must be 1024 bytes");
static_assert(sizeof(sa_pair_rec) == 32, "pair record must be 32 bytes");
static_assert(sizeof(sa_seq_desc) == 16, "seq desc must be 16 bytes");
static_assert(sizeof(sa_item) == 8, "item must be 8 bytes");
static_assert(sizeof(sa_tile_item) == 32, "tile item must be 32 bytes");

namespace sa {
namespace {

constexpr uint32_t kR = SA_DIAG_PER_LANE;
constexpr uint32_t kWarpDiag = SA_DIAG_PER_WARP;
constexpr uint32_t kWarps = SA_CTA_WARPS;
constexpr uint32_t kU = SA_ROW_UNROLL;
constexpr int32_t kNeg = -(1 << 30);  // "minus infinity" that survives one addition
// Cost of merging one (offset, row chunk) partial, in units of one scanned cell (from the SASS
// instruction counts: ~15 vs 5.4 instructions in TILE mode, ~25 vs 2.45 in PAIR mode).
constexpr uint32_t kMergeCellsTile = 3, kMergeCellsPair = 10;

inline uint32_t round_up(uint32_t v, uint32_t m) { return (v + m - 1) / m * m; }

// Host threads for the packer's bulk loops (the reference's only OpenMP use is also input preparation,
// main.c:82,96 — there it races on stdin). The workers are started once and parked on a condition
// variable: a large batch is packed range by range (sa_plan_pack_range), and starting 15 threads per range
// would cost more than the range's share of the work. Several callers may have a job open at once (the
// ranges of one batch are packed concurrently): a job is a counter of chunks, workers take chunks from
// whichever job has some left, and every caller works on its own job too.
class Pool {
   public:
    static Pool* get() {
        static std::once_flag once;
        std::call_once(once, [] { pthread_atfork(nullptr, nullptr, [] { instance().store(nullptr); }); });
        Pool* p = instance().load(std::memory_order_acquire);
        if (!p) {
            static std::mutex make;
            std::lock_guard<std::mutex> g(make);
            p = instance().load(std::memory_order_acquire);
            if (!p) {
                p = new Pool(std::min(16u, std::max(1u, std::thread::hardware_concurrency())));  // never freed: workers outlive main()
                instance().store(p, std::memory_order_release);
            }
        }
        return p;
    }
    unsigned size() const { return nworkers_ + 1; }
    // fn(c) for every chunk c in [0, nchunks); returns when all of them have run
    template <typename F>
    void run(unsigned nchunks, F&& fn) {
        Job job;
        job.fn = [&fn](unsigned c) { fn(c); };
        job.n = nchunks;
        job.left = nchunks;
        {
            std::lock_guard<std::mutex> g(m_);
            jobs_.push_back(&job);
        }
        wake_.notify_all();
        for (;;) {  // the caller takes chunks of its own job
            const unsigned c = job.next.fetch_add(1, std::memory_order_relaxed);
            if (c >= nchunks) break;
            fn(c);
            job.left.fetch_sub(1, std::memory_order_acq_rel);
        }
        std::unique_lock<std::mutex> g(m_);
        jobs_.erase(std::find(jobs_.begin(), jobs_.end(), &job));  // no new worker can pick it up any more
        done_.wait(g, [&] { return job.left.load(std::memory_order_acquire) == 0 && job.users == 0; });
    }

   private:
    struct Job {
        std::function<void(unsigned)> fn;
        unsigned n = 0;
        std::atomic<unsigned> next{0}, left{0};
        unsigned users = 0;  // workers currently inside this job (guarded by m_)
    };
    static std::atomic<Pool*>& instance() {
        static std::atomic<Pool*> p{nullptr};
        return p;
    }
    explicit Pool(unsigned n) : nworkers_(n > 1 ? n - 1 : 0) {
        for (unsigned i = 0; i < nworkers_; ++i) std::thread([this] { loop(); }).detach();
    }
    Job* pick() {  // m_ held
        for (Job* j : jobs_)
            if (j->next.load(std::memory_order_relaxed) < j->n) return j;
        return nullptr;
    }
    void loop() {
        std::unique_lock<std::mutex> g(m_);
        for (;;) {
            Job* j = nullptr;
            wake_.wait(g, [&] { return (j = pick()) != nullptr; });
            ++j->users;
            g.unlock();
            for (;;) {
                const unsigned c = j->next.fetch_add(1, std::memory_order_relaxed);
                if (c >= j->n) break;
                j->fn(c);
                j->left.fetch_sub(1, std::memory_order_acq_rel);
            }
            g.lock();
            --j->users;
            done_.notify_all();
        }
    }
    const unsigned nworkers_;
    std::mutex m_;
    std::condition_variable wake_, done_;
    std::vector<Job*> jobs_;
};

// Split of [0, n) over host threads when the job is big enough to pay for them.
template <typename F>
void parallel_for(uint32_t n, uint64_t work_bytes, F fn) {
    static const bool no_pool = std::getenv("SA_PACK_POOL") && std::getenv("SA_PACK_POOL")[0] == '0';  // A/B switch
    unsigned nc = (work_bytes < (1u << 20) || no_pool) ? 1u : 2u * Pool::get()->size();
    if (nc > n) nc = n ? n : 1u;
    if (nc <= 1) {
        fn(0u, n);
        return;
    }
    Pool::get()->run(nc, [&](unsigned c) {
        const uint32_t a = (uint32_t)((uint64_t)n * c / nc), b = (uint32_t)((uint64_t)n * (c + 1) / nc);
        fn(a, b);
    });
}

struct SeqPlan {
    uint32_t len = 0;
    int32_t ndiag = 0;
    uint32_t ntiles = 0;
    uint32_t wd_log2 = 0;
    uint32_t narrow = 0;     // TILE latency mode: a lane owns 8 >> narrow offsets, a warp 256 >> narrow (sa_tile_item.narrow)
    uint32_t crow = kU;
    uint64_t item_cost = 0;  // padded cells per tile
    bool stream = false;     // STREAM mode: one warp walks all rows (no row chunks)
    uint32_t rows_padded() const { return stream ? crow : crow * (kWarps >> wd_log2); }
    uint32_t tile_offsets() const { return (kWarpDiag >> narrow) << wd_log2; }
};

SeqPlan plan_sequence(uint32_t len1, uint32_t len2) {
    SeqPlan sp;
    sp.len = len2;
    sp.ndiag = (int32_t)len1 - (int32_t)len2;
    if (sp.ndiag < 0 || (len2 == 0 && len1 > 0)) {  // no candidates (cudaFunctions.cu:116-118 loops never run)
        sp.ntiles = 0;
        return sp;
    }
    const uint32_t d = sp.ndiag == 0 ? 1u : (uint32_t)sp.ndiag;  // offsets to score
    uint64_t best_cost = ~0ull;
    for (uint32_t l = 0; (1u << l) <= kWarps; ++l) {
        const uint32_t wd = 1u << l, wr = kWarps / wd;
        const uint32_t crow = round_up((len2 + wr - 1) / wr, kU);
        const uint32_t tn = kWarpDiag * wd;
        const uint32_t ntiles = (d + tn - 1) / tn;
        // scan cells + the row-chunk merge (about kMergeCellsTile cell-equivalents per offset and chunk)
        const uint64_t cost = (uint64_t)ntiles * tn * wr * (crow + kMergeCellsTile);
        if (cost < best_cost) {
            best_cost = cost;
            sp.wd_log2 = l;
            sp.crow = crow;
            sp.ntiles = ntiles;
            sp.item_cost = (uint64_t)tn * wr * crow;
        }
    }
    return sp;
}

// dst[i] = src[i] & 31 (the code of an upper-case letter); returns non-zero when some byte is not 'A'..'Z'.
// Restrict-qualified and branch-free so that the compiler vectorises it: this loop is the whole cost of packing.
unsigned encode_letters(const unsigned char* __restrict src, unsigned char* __restrict dst, size_t n) {
    unsigned char notletter = 0;
    for (size_t i = 0; i < n; ++i) {
        const unsigned char c = src[i];
        dst[i] = (unsigned char)(c & 31u);
        notletter |= (unsigned char)((unsigned char)(c - 'A') > 25u);
    }
    return notletter;
}

inline uint64_t pack_key(int32_t score, uint32_t n) {
    return ((uint64_t)((uint32_t)score ^ 0x80000000u) << 32) | (uint64_t)(0xFFFFFFFFu - n);
}

}  // namespace
}  // namespace sa

extern "C" {

// mode: SA_MODE_AUTO picks SA_MODE_PAIR for large single-GPU batches when its limits hold,
// SA_MODE_TILE otherwise. Forcing SA_MODE_PAIR on an input outside its limits is an error.
// Plans and packs sequences [first, first + count) of the problem (their result rows are 0..count-1 of the
// launch). The whole problem is the range [0, nseq); large batches are packed range by range so that
// packing one range overlaps the copy and the kernel of the previous one (sa_session_solve).
int sa_plan_pack_range(const sa_problem* p, uint32_t first, uint32_t count, int world, int mode, void* dst, size_t cap, size_t* blob_bytes) {
    using namespace sa;
    const bool split_flag = (mode & SA_MODE_SPLIT_TILES) != 0;
    mode &= ~SA_MODE_SPLIT_TILES;
    if (!p || !blob_bytes || world < 1 || world > SA_MAX_WORLD || mode < SA_MODE_AUTO || mode > SA_MODE_STREAM ||
        (uint64_t)first + count > p->seq2.size()) {
        set_error("sa_plan_pack: bad argument");
        return SA_ERR_ARG;
    }
    const uint32_t len1 = (uint32_t)p->seq1.size();
    const uint32_t nseq = count;
    const std::string* const seq2 = p->seq2.data() + first;
    // SA_PACK_TIMING=1: phase times of this call on stderr (profiling aid; the printing itself is not counted)
    static const bool timing = std::getenv("SA_PACK_TIMING") != nullptr;
    auto t_prev = std::chrono::steady_clock::now();
    auto mark = [&](const char* what) {
        if (!timing || !dst) return;
        const auto t = std::chrono::steady_clock::now();
        std::fprintf(stderr, "pack %-10s %8.1f us\n", what, std::chrono::duration<double, std::micro>(t - t_prev).count());
        t_prev = std::chrono::steady_clock::now();
    };
    if (p->seq1.size() > SA_GPU_MAX_SEQ1) {
        set_error("Seq1 longer than the GPU executor supports");
        return SA_ERR_UNSUPPORTED;
    }
    int64_t maxw = 0;
    for (int i = 0; i < 4; ++i) maxw = std::max<int64_t>(maxw, std::llabs((int64_t)p->weights[i]));
    // Optimistic alphabet: inputs are almost always pure letters, so the letters are coded without
    // looking at the data; the first non-letter byte met while filling triggers one full scan
    // (build_alphabet) and a refill. Saves a pass over every byte on the latency-critical path.
    Alphabet al = letters_only_alphabet();
    uint32_t max_len2 = 0, nvalid = 0;
    std::vector<uint32_t> len2(nseq);  // the planning passes below walk this array, not the strings
    for (uint32_t s = 0; s < nseq; ++s) {
        const size_t l2 = seq2[s].size();
        len2[s] = (uint32_t)l2;
        if (l2 > SA_GPU_MAX_SEQ2) {
            set_error("a Seq2 is longer than the GPU executor supports");
            return SA_ERR_UNSUPPORTED;
        }
        // every partial sum must stay inside +-2^29 so that kNeg + sum never wraps
        if (maxw * (int64_t)(l2 + 256) * 4 >= (1ll << 31)) {
            set_error("weights too large for the int32 GPU path");
            return SA_ERR_UNSUPPORTED;
        }
        max_len2 = std::max<uint32_t>(max_len2, (uint32_t)l2);
        if (l2 <= len1 && (l2 > 0 || len1 == 0)) ++nvalid;
    }

    // ---- pair mode (two sequences per 32-bit lane as u16x2, per-pair profile in shared memory)
    const int64_t vals[5] = {0, p->weights[0], -(int64_t)p->weights[1], -(int64_t)p->weights[2], -(int64_t)p->weights[3]};
    const int64_t wmin = *std::min_element(vals, vals + 5), wmax = *std::max_element(vals, vals + 5);
    const int64_t wrange = wmax - wmin;
    struct Pair {
        uint32_t a, b;  // b == SA_NO_SEQ: unpaired
        SeqPlan g;
    };
    std::vector<Pair> pairs;
    bool pair_ok = wrange <= 32767 && nvalid > 0 && (world == 1 || mode != SA_MODE_PAIR);
    std::string pair_why = (world != 1 && mode == SA_MODE_PAIR) ? "PAIR mode is single-GPU (STREAM and TILE run on the fabric)"
                                                                 : (wrange > 32767 ? "weight range exceeds 16 bits" : "no sequence has candidates");
    uint32_t pair_cols_needed = 0;
    const bool want_pairs = mode == SA_MODE_PAIR || mode == SA_MODE_STREAM || (mode == SA_MODE_AUTO && nvalid >= SA_PAIR_MIN_SEQS);
    bool stream_ok = pair_ok && mode != SA_MODE_PAIR;
    std::string stream_why = pair_why;
    std::vector<uint32_t> order;  // scorable sequences, longest first; neighbours become pairs
    if (pair_ok && want_pairs) {
        // stable counting sort by length, longest first (lengths are bounded by SA_GPU_MAX_SEQ2)
        std::vector<uint32_t> start((size_t)max_len2 + 2, 0);
        auto scorable = [&](uint32_t s) { return len2[s] <= len1 && (len2[s] > 0 || len1 == 0); };
        for (uint32_t s = 0; s < nseq; ++s)
            if (scorable(s)) ++start[(size_t)(max_len2 - len2[s]) + 1];
        for (size_t i = 1; i < start.size(); ++i) start[i] += start[i - 1];
        order.resize(start.back());
        for (uint32_t s = 0; s < nseq; ++s)
            if (scorable(s)) order[start[(size_t)(max_len2 - len2[s])]++] = s;
    }
    // Fills `pairs` with the STREAM (one warp per tile, all rows) or PAIR (CTA per pair) geometry.
    auto plan_pairs = [&](bool as_stream, std::string& why) -> bool {
        pairs.clear();
        pairs.reserve(order.size() / 2 + 1);
        pair_cols_needed = 0;
        for (size_t i = 0; i < order.size(); i += 2) {
            Pair pr;
            pr.a = order[i];
            pr.b = i + 1 < order.size() ? order[i + 1] : SA_NO_SEQ;
            const uint32_t ma = len2[pr.a];
            const uint32_t mb = pr.b == SA_NO_SEQ ? ma : len2[pr.b];
            const uint32_t rows = std::max(ma, mb), dmax = len1 - std::min(ma, mb);
            SeqPlan g;
            g.len = rows;
            g.ndiag = (int32_t)dmax;
            const uint32_t d = dmax == 0 ? 1u : dmax;
            if (as_stream) {
                g.stream = true;
                g.wd_log2 = 0;
                g.crow = round_up(rows, SA_STREAM_ROW_BLOCK);
                if ((int64_t)g.crow * wrange > 32767) {
                    why = "a whole sequence would overflow 16 bits";
                    return false;
                }
                g.ntiles = (d + kWarpDiag - 1) / kWarpDiag;
                g.item_cost = (uint64_t)kWarpDiag * g.crow;  // per tile
            } else {
                uint64_t best_cost = ~0ull;
                for (uint32_t l = 0; (1u << l) <= kWarps; ++l) {
                    const uint32_t wd = 1u << l, wr = kWarps / wd;
                    const uint32_t crow = round_up((rows + wr - 1) / wr, kU);
                    if ((int64_t)crow * wrange > 32767) continue;
                    const uint32_t tn = kWarpDiag * wd, ntiles = (d + tn - 1) / tn;
                    const uint64_t cost = (uint64_t)ntiles * tn * wr * (crow + kMergeCellsPair);
                    if (cost < best_cost) {
                        best_cost = cost;
                        g.wd_log2 = l;
                        g.crow = crow;
                        g.ntiles = ntiles;
                        g.item_cost = (uint64_t)ntiles * tn * wr * crow;  // whole pair, scan cells only
                    }
                }
                if (best_cost == ~0ull) {
                    why = "a row chunk would overflow 16 bits";
                    return false;
                }
            }
            pr.g = g;
            pair_cols_needed = std::max(pair_cols_needed, g.ntiles * (kWarpDiag << g.wd_log2) + g.rows_padded());
            pairs.push_back(pr);
        }
        if (as_stream) {
            if (round_up(std::max(len1 + SA_SEQ1_TAIL_PAD, pair_cols_needed + 16u), 16) > SA_STREAM_MAX_SEQ1) {
                why = "Seq1 does not fit in shared memory next to the profile blocks";
                return false;
            }
        } else {
            uint32_t mr = kU;
            for (const Pair& pr : pairs) mr = std::max(mr, pr.g.rows_padded());
            const uint64_t smem = (uint64_t)mr * 128 + 2ull * round_up(mr, 16) +
                                  round_up(std::max(len1 + SA_SEQ1_TAIL_PAD, pair_cols_needed + 16u), 16) + SA_PAIR_FIXED_SMEM;
            if (smem > SA_PAIR_MAX_SMEM) {
                why = "pair profile does not fit in shared memory";
                return false;
            }
        }
        return true;
    };
    if (pair_ok && want_pairs) {
        if (stream_ok) stream_ok = plan_pairs(true, stream_why);
        if (stream_ok && mode == SA_MODE_AUTO) {
            // A STREAM item keeps one warp busy for a long time; with fewer than a few items per resident
            // warp the tail dominates (measured: 2048 sequences per GPU run 25% slower than TILE).
            uint64_t n_items = 0;
            for (const Pair& pr : pairs) n_items += pr.g.ntiles;
            if (n_items < (uint64_t)SA_STREAM_MIN_ITEMS_PER_GPU * (uint64_t)world) {
                stream_ok = false;
                stream_why = "too few items per GPU for warp-autonomous tiles";
            }
        }
        if (!stream_ok) {
            if (world == 1 && mode != SA_MODE_STREAM) pair_ok = plan_pairs(false, pair_why);
            else pair_ok = false;  // the fabric kernel runs TILE or STREAM items only
        }
    } else if (pair_ok) {
        pair_ok = false;
        stream_ok = false;
        pair_why = stream_why = "batch too small for pair mode";
    }
    if (!pair_ok) stream_ok = false;
    if (mode == SA_MODE_PAIR && !pair_ok) {
        set_error("pair mode not applicable: " + pair_why);
        return SA_ERR_UNSUPPORTED;
    }
    if (mode == SA_MODE_STREAM && !stream_ok) {
        set_error("stream mode not applicable: " + stream_why);
        return SA_ERR_UNSUPPORTED;
    }
    const bool use_pairs = pair_ok && mode != SA_MODE_TILE;
    const bool use_stream = use_pairs && stream_ok;
    mark("pairs");

    // ---- per-sequence plans
    std::vector<SeqPlan> plans(nseq);
    uint32_t max_rows = kU, max_window = 16;
    uint64_t total_cells = 0, padded_cells = 0, nitems64 = 0;
    for (uint32_t s = 0; s < nseq; ++s) {
        if (use_pairs) {  // the pair's geometry replaces the sequence's own below: only len / ndiag are kept
            plans[s].len = len2[s];
            plans[s].ndiag = (int32_t)len1 - (int32_t)len2[s];
            plans[s].ntiles = 0;
        } else {
            plans[s] = plan_sequence(len1, len2[s]);
        }
        total_cells += len2[s] > len1 ? 0u : len2[s] == len1 ? 1u : (uint64_t)(len1 - len2[s]) * len2[s];
    }
    if (!use_pairs) {
        // Latency mode. plan_sequence minimises each sequence's TOTAL padded cost, which is right when
        // items outnumber SMs. When the whole problem is fewer items than SMs, the solve time is the
        // LONGEST item (an SM scans one row of a 16-warp CTA per ~144 shared-memory wavefronts, so an
        // item lasts ~crow row-steps plus ~2 steps per row chunk it merges) while most SMs idle: split
        // the longest item's rows over more warps (narrower offset tiles, more of them) until the SMs
        // are used up. The reference scores a whole row in ONE block (cudaFunctions.cu:210-216).
        const uint64_t slots = (uint64_t)SA_PLAN_SMS * (uint64_t)world;
        // an item's duration in ninths of a row-step: a row costs (R + 1) table lookups per lane, the merge
        // about two row-steps per row chunk
        auto steps = [](const SeqPlan& sp) { return (uint64_t)sp.crow * ((kR >> sp.narrow) + 1u) + 18ull * (kWarps >> sp.wd_log2); };
        static const uint32_t max_narrow = [] {
            const char* e = std::getenv("SA_PLAN_MAX_NARROW");  // experiments: 0 switches narrow tiles off
            return e ? (uint32_t)std::min(SA_MAX_NARROW, std::max(0, std::atoi(e))) : (uint32_t)SA_MAX_NARROW;
        }();
        uint64_t nit = 0;
        for (uint32_t s = 0; s < nseq; ++s) nit += plans[s].ntiles;
        if (nit < slots) {
            // longest item first: a max-heap over (steps, sequence)
            std::vector<std::pair<uint64_t, uint32_t>> heap;
            heap.reserve(nseq);
            for (uint32_t s = 0; s < nseq; ++s)
                if (plans[s].ntiles) heap.emplace_back(steps(plans[s]), s);
            std::make_heap(heap.begin(), heap.end());
            while (nit < slots && !heap.empty()) {
                std::pop_heap(heap.begin(), heap.end());
                const uint32_t worst = heap.back().second;
                heap.pop_back();
                SeqPlan sp = plans[worst];
                // first more row chunks over narrower tiles; when all 16 warps already are row chunks, fewer
                // offsets per lane (a 16-warp CTA then covers 128 or 64 offsets instead of 256)
                if (sp.wd_log2 > 0) sp.wd_log2 -= 1;
                else if (sp.narrow < max_narrow) sp.narrow += 1;
                else break;  // the longest item cannot be split further
                const uint32_t wr = kWarps >> sp.wd_log2, tn = sp.tile_offsets();
                const uint32_t d = sp.ndiag == 0 ? 1u : (uint32_t)sp.ndiag;
                sp.crow = round_up((sp.len + wr - 1) / wr, kU);
                sp.ntiles = (d + tn - 1) / tn;
                sp.item_cost = (uint64_t)tn * wr * sp.crow;
                if (steps(sp) >= steps(plans[worst]) || nit - plans[worst].ntiles + sp.ntiles > slots) break;
                nit += sp.ntiles - plans[worst].ntiles;
                plans[worst] = sp;
                heap.emplace_back(steps(sp), worst);
                std::push_heap(heap.begin(), heap.end());
            }
        }
    }
    if (use_pairs) {
        for (const Pair& pr : pairs) {
            for (uint32_t s : {pr.a, pr.b}) {
                if (s == SA_NO_SEQ) continue;
                const int32_t nd = plans[s].ndiag;
                const uint32_t len = plans[s].len;
                plans[s] = pr.g;  // both members share the pair's geometry
                plans[s].ndiag = nd;
                plans[s].len = len;
            }
            max_rows = std::max(max_rows, pr.g.rows_padded());
            padded_cells += 2 * pr.g.item_cost * (use_stream ? pr.g.ntiles : 1u);
            if (use_stream) nitems64 += pr.g.ntiles;
        }
        if (!use_stream) nitems64 = pairs.size();
    } else {
        for (uint32_t s = 0; s < nseq; ++s) {
            max_rows = std::max(max_rows, plans[s].rows_padded());
            if (plans[s].ntiles)
                max_window = std::max(max_window, round_up(plans[s].rows_padded() + plans[s].tile_offsets(), 16));
            padded_cells += plans[s].item_cost * plans[s].ntiles;
            nitems64 += plans[s].ntiles;
        }
    }
    if (nitems64 > 0x7FFFFFFFull) {
        set_error("too many work items");
        return SA_ERR_UNSUPPORTED;
    }
    const uint32_t nitems = (uint32_t)nitems64;
    mark("plans");

    // ---- partition: deal sequences (or tiles, when sequences are scarce) to ranks
    struct RankItem {
        uint32_t seq, tile;
    };
    std::vector<std::vector<RankItem>> per_rank((size_t)world);
    std::vector<uint8_t> home(nseq, 0);
    std::vector<uint32_t> rank_first_pair(SA_MAX_WORLD + 1, 0);
    bool by_tile = false;  // tiles of one sequence dealt to different ranks (cross-GPU argmax reduce)
    // SA_SPLIT_TILES=1 forces the per-tile deal (tests and the strong-scaling benchmark line)
    static const bool split_env = [] { const char* e = std::getenv("SA_SPLIT_TILES"); return e && *e && *e != '0'; }();
    if (use_pairs) {
        // STREAM: a tile costs 256 x crow and the pairs were formed longest first, so they already are
        // costliest first; PAIR geometries differ per pair and need the sort
        if (!use_stream)
            std::stable_sort(pairs.begin(), pairs.end(), [](const Pair& x, const Pair& y) { return x.g.item_cost > y.g.item_cost; });
        if (use_stream && world == 1) {
            rank_first_pair[0] = 0;
            for (size_t r = 1; r <= SA_MAX_WORLD; ++r) rank_first_pair[r] = (uint32_t)pairs.size();
            per_rank[0].reserve(nitems);
            for (uint32_t pid = 0; pid < pairs.size(); ++pid)
                for (uint32_t t = 0; t < pairs[pid].g.ntiles; ++t) per_rank[0].push_back({pid, t});
        } else if (use_stream) {
            // whole pairs are dealt to the least-loaded rank, costliest first: every atomic stays local
            std::vector<uint64_t> load((size_t)world, 0);
            std::vector<uint8_t> pair_rank(pairs.size(), 0);
            for (uint32_t pid = 0; pid < pairs.size(); ++pid) {
                const int r = (int)(std::min_element(load.begin(), load.end()) - load.begin());
                pair_rank[pid] = (uint8_t)r;
                home[pairs[pid].a] = (uint8_t)r;
                if (pairs[pid].b != SA_NO_SEQ) home[pairs[pid].b] = (uint8_t)r;
                load[(size_t)r] += pairs[pid].g.item_cost * pairs[pid].g.ntiles;
            }
            // renumber so that a rank's pairs are contiguous: a rank then needs ONE slice of the record table
            std::vector<Pair> grouped;
            grouped.reserve(pairs.size());
            for (int r = 0; r < world; ++r) {
                rank_first_pair[(size_t)r] = (uint32_t)grouped.size();
                per_rank[(size_t)r].reserve((size_t)(nitems / (uint32_t)world) + 64);
                for (uint32_t pid = 0; pid < pairs.size(); ++pid) {
                    if (pair_rank[pid] != r) continue;
                    const uint32_t nid = (uint32_t)grouped.size();
                    grouped.push_back(pairs[pid]);
                    for (uint32_t t = 0; t < pairs[pid].g.ntiles; ++t) per_rank[(size_t)r].push_back({nid, t});
                }
            }
            pairs.swap(grouped);
            for (size_t r = (size_t)world; r <= SA_MAX_WORLD; ++r) rank_first_pair[r] = (uint32_t)pairs.size();
        } else {
            for (const Pair& pr : pairs) per_rank[0].push_back({pr.a, pr.b});
        }
    } else {
        std::vector<uint32_t> by_cost(nseq);
        std::iota(by_cost.begin(), by_cost.end(), 0u);
        std::stable_sort(by_cost.begin(), by_cost.end(),
                         [&](uint32_t a, uint32_t b) { return plans[a].item_cost > plans[b].item_cost; });
        std::vector<uint64_t> load((size_t)world, 0);
        by_tile = world > 1 && (nseq < 2u * (uint32_t)world || split_env || split_flag);
        auto lightest = [&]() { return (int)(std::min_element(load.begin(), load.end()) - load.begin()); };
        for (uint32_t s : by_cost) {
            const SeqPlan& sp = plans[s];
            if (sp.ntiles == 0) continue;
            if (!by_tile) {
                const int r = lightest();
                home[s] = (uint8_t)r;
                for (uint32_t t = 0; t < sp.ntiles; ++t) per_rank[(size_t)r].push_back({s, t});
                load[(size_t)r] += sp.item_cost * sp.ntiles;
            } else {
                for (uint32_t t = 0; t < sp.ntiles; ++t) {
                    const int r = lightest();
                    if (t == 0) home[s] = (uint8_t)r;
                    per_rank[(size_t)r].push_back({s, t});
                    load[(size_t)r] += sp.item_cost;
                }
            }
        }
    }

    mark("partition");
    // ---- layout
    uint32_t len1_padded = round_up(len1 + SA_SEQ1_TAIL_PAD, 16);
    if (use_pairs) len1_padded = std::max(len1_padded, round_up(pair_cols_needed + 16u, 16));
    uint64_t off = sizeof(sa_blob_header);
    const uint32_t off_lut = (uint32_t)off;
    off += sizeof(int32_t) * SA_ALPHABET * SA_ALPHABET;
    const uint32_t off_seq1 = (uint32_t)off;
    off += len1_padded;
    const uint32_t off_seqtab = (uint32_t)off;
    off += (uint64_t)sizeof(sa_seq_desc) * nseq;
    off = (off + 15) / 16 * 16;
    const uint32_t off_items = (uint32_t)off;
    off += (uint64_t)sizeof(sa_item) * nitems;
    off = (off + 15) / 16 * 16;
    const uint32_t off_pairs = (uint32_t)off;
    if (use_stream) off += (uint64_t)sizeof(sa_pair_rec) * pairs.size();
    off = (off + 15) / 16 * 16;
    const uint32_t off_tile_items = (uint32_t)off;
    if (!use_pairs) off += (uint64_t)sizeof(sa_tile_item) * nitems;
    off = (off + 31) / 32 * 32;
    const uint32_t off_seq2 = (uint32_t)off;
    // Seq2 letters grouped by the rank that scores them, so that a rank's share is ONE byte range it can
    // pull from rank 0 (the MPI_Scatter of main.c:174). With tiles of one sequence dealt to several
    // ranks (by_tile: few sequences) every rank needs every sequence: one group, every range = all.
    std::vector<uint32_t> seq_off(nseq);
    std::vector<uint32_t> seq2_lo((size_t)world, 0), seq2_hi((size_t)world, 0);
    // Inside a rank's group the sequences follow the rank's item order (costliest first): the chunks a
    // rank stages first are the ones its first items read.
    std::vector<std::vector<uint32_t>> group((size_t)world);
    {
        std::vector<uint8_t> placed(nseq, 0);
        for (int r = 0; r < world; ++r) group[(size_t)r].reserve(by_tile ? (r == 0 ? nseq : 0u) : nseq / (uint32_t)world + 16u);
        if (use_stream) {
            // a rank's items are its pairs in order, all tiles of a pair in a row: first-seen order = pair order
            for (int r = 0; r < world; ++r)
                for (uint32_t pid = rank_first_pair[(size_t)r]; pid < rank_first_pair[(size_t)r + 1]; ++pid)
                    for (uint32_t s : {pairs[pid].a, pairs[pid].b}) {
                        if (s == SA_NO_SEQ || placed[s]) continue;
                        placed[s] = 1;
                        group[(size_t)home[s]].push_back(s);
                    }
        }
        for (int r = 0; r < world && !use_stream; ++r)
            for (const RankItem& it : per_rank[(size_t)r]) {
                uint32_t members[2] = {it.seq, SA_NO_SEQ};
                if (use_stream) members[0] = pairs[it.seq].a, members[1] = pairs[it.seq].b;
                else if (use_pairs) members[1] = it.tile;  // PAIR items are (first, second)
                for (uint32_t s : members) {
                    if (s == SA_NO_SEQ || placed[s]) continue;
                    placed[s] = 1;
                    group[by_tile ? 0u : (size_t)home[s]].push_back(s);
                }
            }
        for (uint32_t s = 0; s < nseq; ++s)
            if (!placed[s]) group[0].push_back(s);  // candidate-free sequences: never read by a kernel
    }
    for (int r = 0; r < world; ++r) {
        seq2_lo[(size_t)r] = (uint32_t)off;
        for (uint32_t s : group[(size_t)r]) {
            seq_off[s] = (uint32_t)off;
            off += round_up(std::max(plans[s].rows_padded(), len2[s]), 16);
            if (off >= 0xFFFFF000ull) {
                set_error("blob larger than 4 GiB");
                return SA_ERR_UNSUPPORTED;
            }
        }
        seq2_hi[(size_t)r] = (uint32_t)off;
    }
    if (by_tile)
        for (int r = 0; r < world; ++r) seq2_lo[(size_t)r] = off_seq2, seq2_hi[(size_t)r] = (uint32_t)off;
    const size_t total = (size_t)off;
    *blob_bytes = total;
    if (!dst) return SA_OK;
    if (cap < total) {
        set_error("sa_plan_pack_into: destination too small");
        return SA_ERR_SMALL_BUFFER;
    }

    mark("layout");
    // ---- fill
    unsigned char* base = static_cast<unsigned char*>(dst);
    std::memset(base, 0, off_seq2);  // header + tables + Seq1; every Seq2 slot zeroes its own padding below
    sa_blob_header* h = reinterpret_cast<sa_blob_header*>(base);
    h->magic = SA_BLOB_MAGIC;
    h->version = SA_BLOB_VERSION;
    h->total_bytes = (uint32_t)total;
    h->flags = use_stream ? SA_FLAG_STREAM_MODE : (use_pairs ? SA_FLAG_PAIR_MODE : 0u);
    h->off_pairs = use_stream ? off_pairs : 0u;
    h->npairs = use_stream ? (uint32_t)pairs.size() : 0u;
    std::memcpy(h->weights, p->weights, sizeof(h->weights));
    h->len1 = len1;
    h->len1_padded = len1_padded;
    h->nseq = nseq;
    h->nitems = nitems;
    h->off_lut = off_lut;
    h->off_seq1 = off_seq1;
    h->off_seqtab = off_seqtab;
    h->off_items = off_items;
    h->off_seq2 = off_seq2;
    h->max_rows_padded = max_rows;
    h->max_window_bytes = max_window;
    // fabric staging granularity: 8 KiB (one ready flag per chunk, one chunk per copy item: even a
    // 30 KB problem is pulled by several CTAs at once), grown so that a blob never needs more flags than exist
    h->chunk_bytes = std::max<uint32_t>(1u << 13, round_up((uint32_t)((total + SA_FABRIC_MAX_CHUNKS - 1) / SA_FABRIC_MAX_CHUNKS), 1u << 13));
    h->nchunks = (uint32_t)((total + h->chunk_bytes - 1) / h->chunk_bytes);
    h->total_cells = total_cells;
    h->padded_cells = padded_cells;
    h->world = (uint32_t)world;
    h->value_min = (int32_t)wmin;
    h->value_range = (uint32_t)wrange;
    sa_seq_desc* tab = reinterpret_cast<sa_seq_desc*>(base + off_seqtab);
    for (int attempt = 0; attempt < 2; ++attempt) {
        std::atomic<int> unknown_byte{0};
        build_lut(al, p->weights, reinterpret_cast<int32_t*>(base + off_lut));
        if (attempt == 0) {  // letters only: code = low five bits, stored pre-scaled by 4 (a vectorised pass)
            const unsigned char* __restrict src = reinterpret_cast<const unsigned char*>(p->seq1.data());
            unsigned char* __restrict d1 = base + off_seq1;
            unsigned char notletter = 0;
            for (uint32_t j = 0; j < len1; ++j) {
                const unsigned char c = src[j];
                d1[j] = (unsigned char)((c & 31u) << 2);
                notletter |= (unsigned char)((unsigned char)(c - 'A') > 25u);
            }
            if (notletter) unknown_byte.store(1);
        } else {
            int bad = 0;
            for (uint32_t j = 0; j < len1; ++j) {
                const int c = al.code[(unsigned char)p->seq1[j]];
                bad |= c;
                base[off_seq1 + j] = (unsigned char)(c * 4);
            }
            if (bad < 0) unknown_byte.store(1);
        }
        parallel_for(nseq, total, [&](uint32_t s0, uint32_t s1) {
            int bad = 0;
            for (uint32_t s = s0; s < s1; ++s) {
                const SeqPlan& sp = plans[s];
                tab[s].off = seq_off[s];
                tab[s].len = (uint16_t)sp.len;
                tab[s].ndiag = sp.ndiag;
                tab[s].ntiles = (uint16_t)sp.ntiles;
                tab[s].wd_log2 = (uint8_t)sp.wd_log2;
                tab[s].home = home[s];
                tab[s].crow = (uint16_t)sp.crow;
                const std::string& q = seq2[s];
                unsigned char* dstq = base + seq_off[s];
                if (attempt == 0) {
                    // letters-only first pass: the code of an upper-case letter is its low five bits; anything
                    // else sets `bad` and triggers the table-driven refill (a branch-free loop the compiler vectorises)
                    const unsigned notletter = encode_letters(reinterpret_cast<const unsigned char*>(q.data()), dstq, q.size());
                    if (notletter) bad = -1;
                } else {
                    for (size_t i = 0; i < q.size(); ++i) {
                        const int c = al.code[(unsigned char)q[i]];
                        bad |= c;
                        dstq[i] = (unsigned char)c;
                    }
                }
                const size_t slot = round_up(std::max(sp.rows_padded(), (uint32_t)q.size()), 16);
                std::memset(dstq + q.size(), 0, slot - q.size());
            }
            if (bad < 0) unknown_byte.store(1);
        });
        if (!unknown_byte.load()) break;
        // a byte outside A-Z: assign the spare codes from a full scan and fill once more
        al = build_alphabet(*p);
        if (!al.ok || attempt == 1) {
            set_error("more than five distinct non-letter bytes in the input");
            return SA_ERR_UNSUPPORTED;
        }
    }
    mark("letters");
    sa_item* items = reinterpret_cast<sa_item*>(base + off_items);
    uint32_t w = 0;
    for (int r = 0; r < world; ++r) {
        h->rank_first_item[r] = w;
        for (const RankItem& it : per_rank[(size_t)r]) items[w++] = sa_item{it.seq, it.tile};
    }
    for (int r = world; r <= SA_MAX_WORLD; ++r) h->rank_first_item[r] = w;
    h->off_shared_end = off_seqtab;
    for (int r = 0; r <= SA_MAX_WORLD; ++r) h->rank_first_pair[r] = use_stream ? rank_first_pair[(size_t)r] : 0u;
    for (int r = 0; r < SA_MAX_WORLD; ++r) {
        h->rank_seq2_lo[r] = r < world ? seq2_lo[(size_t)r] : (uint32_t)total;
        h->rank_seq2_hi[r] = r < world ? seq2_hi[(size_t)r] : (uint32_t)total;
    }
    if (use_stream) {
        sa_pair_rec* pt = reinterpret_cast<sa_pair_rec*>(base + off_pairs);
        for (uint32_t pid = 0; pid < pairs.size(); ++pid) {
            const Pair& pr = pairs[pid];
            sa_pair_rec rec{};
            rec.a = pr.a;
            rec.b = pr.b;
            rec.off_a = tab[pr.a].off;
            rec.len_a = tab[pr.a].len;
            rec.off_b = pr.b == SA_NO_SEQ ? rec.off_a : tab[pr.b].off;
            rec.len_b = pr.b == SA_NO_SEQ ? rec.len_a : tab[pr.b].len;
            rec.crow = (uint16_t)pr.g.crow;
            rec.ntiles = (uint16_t)pr.g.ntiles;
            rec.home = tab[pr.a].home;
            pt[pid] = rec;
        }
    }
    h->off_tile_items = use_pairs ? 0u : off_tile_items;
    if (!use_pairs) {
        sa_tile_item* fat = reinterpret_cast<sa_tile_item*>(base + off_tile_items);
        for (uint32_t i = 0; i < nitems; ++i) {
            const sa_seq_desc& d = tab[items[i].seq];
            fat[i].seq = items[i].seq;
            fat[i].tile = items[i].tile;
            fat[i].off = d.off;
            fat[i].ndiag = d.ndiag;
            fat[i].len = d.len;
            fat[i].crow = d.crow;
            fat[i].ntiles = d.ntiles;
            fat[i].wd_log2 = d.wd_log2;
            fat[i].narrow = (uint8_t)plans[items[i].seq].narrow;
            fat[i].home = d.home;
        }
    }
    mark("records");
    return SA_OK;
}

int sa_plan_pack_mode(const sa_problem* p, int world, int mode, void* dst, size_t cap, size_t* blob_bytes) {
    if (!p) {
        sa::set_error("sa_plan_pack: bad argument");
        return SA_ERR_ARG;
    }
    return sa_plan_pack_range(p, 0, (uint32_t)p->seq2.size(), world, mode, dst, cap, blob_bytes);
}

int sa_plan_pack_into(const sa_problem* p, int world, void* dst, size_t cap, size_t* blob_bytes) {
    return sa_plan_pack_mode(p, world, SA_MODE_AUTO, dst, cap, blob_bytes);
}

int sa_plan_pack(const sa_problem* p, int world, void** blob, size_t* blob_bytes) {
    if (!blob || !blob_bytes) {
        sa::set_error("sa_plan_pack: null argument");
        return SA_ERR_ARG;
    }
    size_t need = 0;
    int rc = sa_plan_pack_into(p, world, nullptr, 0, &need);
    if (rc != SA_OK) return rc;
    void* mem = nullptr;
    if (posix_memalign(&mem, 256, need ? need : 256) != 0) return SA_ERR_NOMEM;
    rc = sa_plan_pack_into(p, world, mem, need, blob_bytes);
    if (rc != SA_OK) {
        std::free(mem);
        return rc;
    }
    *blob = mem;
    return SA_OK;
}

void sa_blob_free(void* blob) { std::free(blob); }

// CPU walk of the blob with the kernel's exact decomposition (see align_kernel.cu):
// per item, 16 "warps" = wd offset groups x wr row chunks; per lane R+1 running prefixes and R
// running maxima; per-chunk results combined in row order; best (score, smallest n) folded
// into a packed 64-bit key. `sa_blob_score_items_cpu` does that for a range of items (one rank's
// share) and max-merges into keys[nseq]; `sa_blob_finalize_cpu` turns complete keys into
// (score, n, k) by rescanning each winning diagonal for the first k.
namespace {

struct BlobView {
    const unsigned char* base;
    const sa_blob_header* h;
    const int32_t* lut;
    const unsigned char* seq1;
    const sa_seq_desc* tab;
    const sa_item* items;
    int32_t value(unsigned code2, unsigned c4) const { return lut[code2 * SA_ALPHABET + (c4 >> 2)]; }
};

int open_blob(const void* blob, size_t blob_bytes, BlobView* v) {
    if (!blob || blob_bytes < sizeof(sa_blob_header)) {
        sa::set_error("blob: bad argument");
        return SA_ERR_ARG;
    }
    v->base = static_cast<const unsigned char*>(blob);
    v->h = reinterpret_cast<const sa_blob_header*>(v->base);
    if (v->h->magic != SA_BLOB_MAGIC || v->h->version != SA_BLOB_VERSION || v->h->total_bytes > blob_bytes) {
        sa::set_error("blob: not a blob of this version");
        return SA_ERR_ARG;
    }
    v->lut = reinterpret_cast<const int32_t*>(v->base + v->h->off_lut);
    v->seq1 = v->base + v->h->off_seq1;
    v->tab = reinterpret_cast<const sa_seq_desc*>(v->base + v->h->off_seqtab);
    v->items = reinterpret_cast<const sa_item*>(v->base + v->h->off_items);
    return SA_OK;
}

}  // namespace

namespace {

// PAIR mode walk: bit-exact emulation of the kernel's u16x2 arithmetic (plain 32-bit adds on
// biased, non-negative halves; per-half unsigned max), with checks that no half ever carries or
// borrows into its neighbour — the property the planner's limits are there to guarantee.
int score_pair_cpu(const BlobView& b, uint32_t it, uint64_t* keys) {
    using namespace sa;
    // PAIR: the item is the pair, all its tiles, 16 warps = wd x wr.
    // STREAM: the item is (pair, one tile), a single warp walks all rows (wd = wr = 1).
    const bool stream = (b.h->flags & SA_FLAG_STREAM_MODE) != 0;
    const sa_pair_rec* pairs = reinterpret_cast<const sa_pair_rec*>(b.base + b.h->off_pairs);
    const uint32_t ia = stream ? pairs[b.items[it].seq].a : b.items[it].seq;
    const uint32_t ib = stream ? pairs[b.items[it].seq].b : b.items[it].tile;
    // STREAM: the kernel reads only the self-contained pair record, and so does this walk (a rank of a
    // multi-GPU solve holds no sequence table); where the table IS present the two must agree.
    sa_seq_desc da_rec{}, db_rec{};
    if (stream) {
        const sa_pair_rec& pr = pairs[b.items[it].seq];
        const sa_seq_desc& ta = b.tab[pr.a];
        const bool have_tab = ta.off != 0u;
        const bool ok_b = pr.b == SA_NO_SEQ || !have_tab || (pr.off_b == b.tab[pr.b].off && pr.len_b == b.tab[pr.b].len);
        if (have_tab && (pr.off_a != ta.off || pr.len_a != ta.len || pr.crow != ta.crow || pr.ntiles != ta.ntiles || pr.home != ta.home || !ok_b)) {
            set_error("pair walk: pair record disagrees with the sequence table");
            return SA_ERR_ARG;
        }
        da_rec.off = pr.off_a, da_rec.len = pr.len_a, da_rec.crow = pr.crow, da_rec.ntiles = pr.ntiles, da_rec.home = pr.home;
        da_rec.ndiag = (int32_t)b.h->len1 - (int32_t)pr.len_a;
        db_rec = da_rec;
        db_rec.off = pr.off_b, db_rec.len = pr.len_b, db_rec.ndiag = (int32_t)b.h->len1 - (int32_t)pr.len_b;
    }
    const sa_seq_desc& da = stream ? da_rec : b.tab[ia];
    const sa_seq_desc* db = ib == SA_NO_SEQ ? nullptr : (stream ? &db_rec : &b.tab[ib]);
    const uint32_t nwarps = stream ? 1u : kWarps;
    const uint32_t wd = 1u << da.wd_log2, wr = nwarps / wd, crow = da.crow, rows = crow * wr;
    const uint32_t tile_begin = stream ? b.items[it].tile : 0u, tile_end = stream ? tile_begin + 1u : da.ntiles;
    const uint32_t tn = kWarpDiag * wd;
    const int32_t wmin = b.h->value_min;
    const uint32_t BIG = 32767u, BIG2 = BIG | (BIG << 16);
    const unsigned char* sA = b.base + da.off;
    std::vector<unsigned char> zeros(rows, 0);
    const unsigned char* sB = db ? b.base + db->off : zeros.data();
    std::vector<uint32_t> prof((size_t)rows * SA_ALPHABET);
    for (uint32_t row = 0; row < rows; ++row)
        for (uint32_t c = 0; c < SA_ALPHABET; ++c) {
            const int64_t ea = (int64_t)b.lut[sA[row] * SA_ALPHABET + c] - wmin;
            const int64_t eb = (int64_t)b.lut[sB[row] * SA_ALPHABET + c] - wmin;
            if (ea < 0 || eb < 0 || ea > 65535 || eb > 65535) {
                set_error("pair walk: biased value outside u16");
                return SA_ERR_ARG;
            }
            prof[(size_t)row * SA_ALPHABET + c] = (uint32_t)ea | ((uint32_t)eb << 16);
        }
    std::vector<uint32_t> csum((size_t)kWarps * (kR + 1) * 32), cmax((size_t)kWarps * kR * 32);
    uint64_t keyA = 0, keyB = 0;
    for (uint32_t tile = tile_begin; tile < tile_end; ++tile) {
        const uint32_t n0 = tile * tn;
        if ((uint64_t)n0 + rows + tn > b.h->len1_padded) {
            set_error("pair walk: window outside padded Seq1");
            return SA_ERR_ARG;
        }
        for (uint32_t w = 0; w < nwarps; ++w) {
            const uint32_t g = w & (wd - 1), q = w >> da.wd_log2;
            for (uint32_t lane = 0; lane < 32; ++lane) {
                const uint32_t d0 = g * kWarpDiag + lane * kR;
                uint32_t P[kR + 1], best[kR];
                for (uint32_t r = 0; r <= kR; ++r) P[r] = 0;
                for (uint32_t r = 0; r < kR; ++r) best[r] = 0;
                for (uint32_t i = q * crow; i < (q + 1) * crow; ++i) {
                    for (uint32_t r = 0; r <= kR; ++r) {
                        const uint32_t v = prof[(size_t)i * SA_ALPHABET + (b.seq1[n0 + i + d0 + r] >> 2)];
                        if ((P[r] & 0xFFFFu) + (v & 0xFFFFu) > BIG || (P[r] >> 16) + (v >> 16) > BIG) {
                            set_error("pair walk: prefix overflows its 16-bit lane");
                            return SA_ERR_ARG;
                        }
                        P[r] += v;
                    }
                    for (uint32_t r = 0; r < kR; ++r) {
                        const uint32_t t = P[r] + BIG2 - P[r + 1];  // no borrow: every half of P <= BIG
                        const uint32_t lo = std::max(best[r] & 0xFFFFu, t & 0xFFFFu);
                        const uint32_t hi = std::max(best[r] >> 16, t >> 16);
                        best[r] = lo | (hi << 16);
                    }
                }
                for (uint32_t r = 0; r <= kR; ++r) csum[((size_t)w * (kR + 1) + r) * 32 + lane] = P[r];
                for (uint32_t r = 0; r < kR; ++r) cmax[((size_t)w * kR + r) * 32 + lane] = best[r];
            }
        }
        for (uint32_t d = 0; d < tn; ++d) {
            const uint32_t g = d / kWarpDiag, lane = (d % kWarpDiag) / kR, r = d % kR;
            for (int half = 0; half < 2; ++half) {
                const sa_seq_desc* sd = half == 0 ? &da : db;
                if (!sd) continue;
                const int sh = half * 16;
                int32_t A = 0, T0 = 0, T1 = 0, M = kNeg;
                for (uint32_t q = 0; q < wr; ++q) {
                    const uint32_t w = q * wd + g;
                    const int32_t s0 = (int32_t)((csum[((size_t)w * (kR + 1) + r) * 32 + lane] >> sh) & 0xFFFFu) + (int32_t)crow * wmin;
                    const int32_t s1 = (int32_t)((csum[((size_t)w * (kR + 1) + r + 1) * 32 + lane] >> sh) & 0xFFFFu) + (int32_t)crow * wmin;
                    const int32_t mx = (int32_t)((cmax[((size_t)w * kR + r) * 32 + lane] >> sh) & 0xFFFFu) - (int32_t)BIG;
                    M = std::max(M, A + mx);
                    A += s0 - s1;
                    T0 += s0;
                    T1 += s1;
                }
                const uint32_t n = n0 + d;
                uint64_t& key = half == 0 ? keyA : keyB;
                if (sd->ndiag == 0) {
                    if (n == 0) key = std::max(key, pack_key(T0, 0));
                } else if (n < (uint32_t)sd->ndiag) {
                    key = std::max(key, pack_key(T1 + M, n));
                }
            }
        }
    }
    keys[ia] = std::max(keys[ia], keyA);
    if (db) keys[ib] = std::max(keys[ib], keyB);
    return SA_OK;
}

}  // namespace

int sa_blob_score_items_cpu(const void* blob, size_t blob_bytes, uint32_t first_item, uint32_t last_item,
                            uint64_t* keys) {
    using namespace sa;
    BlobView b;
    int rc = open_blob(blob, blob_bytes, &b);
    if (rc != SA_OK) return rc;
    if (!keys || first_item > last_item || last_item > b.h->nitems) {
        set_error("sa_blob_score_items_cpu: bad item range");
        return SA_ERR_ARG;
    }
    if (b.h->flags & (SA_FLAG_PAIR_MODE | SA_FLAG_STREAM_MODE)) {
        for (uint32_t it = first_item; it < last_item; ++it) {
            rc = score_pair_cpu(b, it, keys);
            if (rc != SA_OK) return rc;
        }
        return SA_OK;
    }
    std::vector<int32_t> csum((size_t)kWarps * (kR + 1) * 32), cmax((size_t)kWarps * kR * 32);
    const sa_tile_item* fat = reinterpret_cast<const sa_tile_item*>(b.base + b.h->off_tile_items);
    for (uint32_t it = first_item; it < last_item; ++it) {
        const uint32_t s = b.items[it].seq, tile = b.items[it].tile;
        // the kernel reads only the self-contained record, and so does this walk; where the sequence
        // table is present (it is not on a rank that pulled just its own share) the two must agree
        const sa_seq_desc& tabd = b.tab[s];
        if (fat[it].seq != s || fat[it].tile != tile ||
            (tabd.off != 0u && (fat[it].off != tabd.off || fat[it].ndiag != tabd.ndiag || fat[it].len != tabd.len || fat[it].crow != tabd.crow ||
                                fat[it].ntiles != tabd.ntiles || fat[it].wd_log2 != tabd.wd_log2 || fat[it].home != tabd.home))) {
            set_error("sa_blob_score_items_cpu: tile item record disagrees with the item list / sequence table");
            return SA_ERR_ARG;
        }
        sa_seq_desc sd{};
        sd.off = fat[it].off, sd.len = fat[it].len, sd.crow = fat[it].crow, sd.ndiag = fat[it].ndiag;
        sd.ntiles = fat[it].ntiles, sd.wd_log2 = fat[it].wd_log2, sd.home = fat[it].home;
        const unsigned char* s2 = b.base + sd.off;
        const uint32_t wd = 1u << sd.wd_log2, wr = kWarps / wd, crow = sd.crow;
        if (fat[it].narrow > SA_MAX_NARROW) {
            set_error("sa_blob_score_items_cpu: bad tile geometry");
            return SA_ERR_ARG;
        }
        const uint32_t R = kR >> fat[it].narrow, warp_diag = 32u * R;  // offsets per lane / per warp of this item
        const uint32_t tn = warp_diag * wd, n0 = tile * tn;
        if ((uint64_t)n0 + wr * crow + tn > b.h->len1_padded) {
            set_error("sa_blob_score_items_cpu: window outside padded Seq1");
            return SA_ERR_ARG;
        }
        for (uint32_t w = 0; w < kWarps; ++w) {
            const uint32_t g = w & (wd - 1), q = w >> sd.wd_log2;
            for (uint32_t lane = 0; lane < 32; ++lane) {
                const uint32_t d0 = g * warp_diag + lane * R;
                int32_t P[kR + 1], best[kR];
                for (uint32_t r = 0; r <= R; ++r) P[r] = 0;
                for (uint32_t r = 0; r < R; ++r) best[r] = kNeg;
                for (uint32_t i = q * crow; i < (q + 1) * crow; ++i) {
                    const unsigned code2 = s2[i];
                    for (uint32_t r = 0; r <= R; ++r) P[r] += b.value(code2, b.seq1[n0 + i + d0 + r]);
                    for (uint32_t r = 0; r < R; ++r) best[r] = std::max(best[r], P[r] - P[r + 1]);
                }
                for (uint32_t r = 0; r <= R; ++r) csum[((size_t)w * (kR + 1) + r) * 32 + lane] = P[r];
                for (uint32_t r = 0; r < R; ++r) cmax[((size_t)w * kR + r) * 32 + lane] = best[r];
            }
        }
        uint64_t tile_key = 0;
        for (uint32_t d = 0; d < tn; ++d) {
            const uint32_t g = d / warp_diag, lane = (d % warp_diag) / R, r = d % R;
            int32_t A = 0, T0 = 0, T1 = 0, M = kNeg;
            for (uint32_t q = 0; q < wr; ++q) {
                const uint32_t w = q * wd + g;
                const int32_t s0 = csum[((size_t)w * (kR + 1) + r) * 32 + lane];
                const int32_t s1 = csum[((size_t)w * (kR + 1) + r + 1) * 32 + lane];
                const int32_t mx = cmax[((size_t)w * kR + r) * 32 + lane];
                M = std::max(M, A + mx);
                A += s0 - s1;
                T0 += s0;
                T1 += s1;
            }
            const uint32_t n = n0 + d;
            if (sd.ndiag == 0) {
                if (n == 0) tile_key = std::max(tile_key, pack_key(T0, 0));
            } else if (n < (uint32_t)sd.ndiag) {
                tile_key = std::max(tile_key, pack_key(T1 + M, n));
            }
        }
        keys[s] = std::max(keys[s], tile_key);
    }
    return SA_OK;
}

int sa_blob_rank_ranges(const sa_blob_header* h, uint32_t rank, uint32_t lo[4], uint32_t hi[4]) {
    if (!h || !lo || !hi || h->magic != SA_BLOB_MAGIC || h->version != SA_BLOB_VERSION || rank >= h->world) {
        sa::set_error("sa_blob_rank_ranges: bad argument");
        return SA_ERR_ARG;
    }
    const bool stream = (h->flags & SA_FLAG_STREAM_MODE) != 0;
    const uint32_t i0 = h->rank_first_item[rank], i1 = h->rank_first_item[rank + 1];
    lo[0] = 0, hi[0] = h->off_shared_end;
    lo[1] = h->off_items + 8u * i0, hi[1] = h->off_items + 8u * i1;
    if (stream) lo[2] = h->off_pairs + 32u * h->rank_first_pair[rank], hi[2] = h->off_pairs + 32u * h->rank_first_pair[rank + 1];
    else lo[2] = h->off_tile_items + 32u * i0, hi[2] = h->off_tile_items + 32u * i1;
    lo[3] = h->rank_seq2_lo[rank], hi[3] = h->rank_seq2_hi[rank];
    if (i1 <= i0)
        for (int i = 0; i < 4; ++i) lo[i] = hi[i] = 0;  // a rank without items needs nothing
    return SA_OK;
}

int sa_blob_finalize_cpu(const void* blob, size_t blob_bytes, const uint64_t* keys, int32_t* out4) {
    using namespace sa;
    BlobView b;
    int rc = open_blob(blob, blob_bytes, &b);
    if (rc != SA_OK) return rc;
    if (!keys || !out4) {
        set_error("sa_blob_finalize_cpu: null argument");
        return SA_ERR_ARG;
    }
    for (uint32_t s = 0; s < b.h->nseq; ++s) {
        const sa_seq_desc& sd = b.tab[s];
        if (sd.ntiles == 0) {
            out4[4 * s] = INT_MIN;
            out4[4 * s + 1] = out4[4 * s + 2] = out4[4 * s + 3] = 0;
            continue;
        }
        const unsigned char* s2 = b.base + sd.off;
        const int32_t best = (int32_t)((uint32_t)(keys[s] >> 32) ^ 0x80000000u);
        const uint32_t n = 0xFFFFFFFFu - (uint32_t)(keys[s] & 0xFFFFFFFFu);
        int32_t k = 0;
        if (sd.ndiag != 0) {
            if (n >= (uint32_t)sd.ndiag) {
                set_error("sa_blob_finalize_cpu: key holds no valid offset (were all items scored?)");
                return SA_ERR_ARG;
            }
            int32_t T0 = 0, T1 = 0;
            for (uint32_t i = 0; i < sd.len; ++i) {
                T0 += b.value(s2[i], b.seq1[n + i]);
                T1 += b.value(s2[i], b.seq1[n + i + 1]);
            }
            if (T0 != best) {
                int32_t S = 0;
                k = -1;
                for (uint32_t i = 0; i + 1 < sd.len; ++i) {
                    S += b.value(s2[i], b.seq1[n + i]) - b.value(s2[i], b.seq1[n + i + 1]);
                    if (T1 + S == best) {
                        k = (int32_t)(i + 1);
                        break;
                    }
                }
                if (k < 0) {
                    set_error("sa_blob_finalize_cpu: winning diagonal does not reproduce the best score");
                    return SA_ERR_ARG;
                }
            }
        }
        out4[4 * s] = best;
        out4[4 * s + 1] = (int32_t)n;
        out4[4 * s + 2] = k;
        out4[4 * s + 3] = 0;
    }
    return SA_OK;
}

int sa_blob_execute_cpu(const void* blob, size_t blob_bytes, int32_t* out4) {
    if (!blob || !out4 || blob_bytes < sizeof(sa_blob_header)) {
        sa::set_error("sa_blob_execute_cpu: bad argument");
        return SA_ERR_ARG;
    }
    const sa_blob_header* h = static_cast<const sa_blob_header*>(blob);
    std::vector<uint64_t> keys(h->nseq ? h->nseq : 1, 0);
    int rc = sa_blob_score_items_cpu(blob, blob_bytes, 0, h->nitems, keys.data());
    if (rc != SA_OK) return rc;
    return sa_blob_finalize_cpu(blob, blob_bytes, keys.data(), out4);
}

}  // extern "C"
