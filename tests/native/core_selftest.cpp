// Native self-test of libseqalign_core, meant to be built with -fsanitize=address,undefined (and,
// separately, -fsanitize=thread for the multi-threaded packer / oracle). The reference has no
// sanitizer story at all while containing real races and overruns; this is
// the host-side half of ours (the device side is scripts/gpu_sanitize.sh).
//
//   core_selftest <input files...>   exits 0 when every check passes.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <random>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

#include "seqalign.h"

static int g_fail = 0;
#define CHECK(cond, what)                                                     \
    do {                                                                      \
        if (!(cond)) {                                                        \
            std::fprintf(stderr, "FAIL %s:%d %s (%s)\n", __FILE__, __LINE__, what, sa_last_error()); \
            ++g_fail;                                                         \
        }                                                                     \
    } while (0)

static std::string slurp(const char* path) {
    std::ifstream f(path, std::ios::binary);
    std::stringstream ss;
    ss << f.rdbuf();
    return ss.str();
}

static void check_problem(sa_problem* p, bool run_naive) {
    const int n = sa_problem_nseq(p);
    std::vector<int32_t> fast((size_t)n * 3 + 3), naive((size_t)n * 3 + 3), walk((size_t)n * 4 + 4);
    CHECK(sa_solve_cpu_fast(p, fast.data(), 4) == SA_OK, "fast oracle");
    if (run_naive) {
        CHECK(sa_solve_cpu_naive(p, naive.data(), 4) == SA_OK, "naive oracle");
        CHECK(std::memcmp(fast.data(), naive.data(), (size_t)n * 12) == 0, "naive == fast");
    }
    for (int world : {1, 3}) {
        for (int mode : {SA_MODE_AUTO, SA_MODE_TILE, SA_MODE_PAIR, SA_MODE_STREAM}) {
            if (mode == SA_MODE_PAIR && world != 1) continue;
            size_t need = 0;
            int rc = sa_plan_pack_mode(p, world, mode, nullptr, 0, &need);
            if (rc == SA_ERR_UNSUPPORTED) continue;  // mode not applicable to this input: fine
            CHECK(rc == SA_OK, "size pass");
            std::vector<unsigned char> blob(need);
            size_t got = 0;
            rc = sa_plan_pack_mode(p, world, mode, blob.data(), blob.size(), &got);
            if (rc == SA_ERR_UNSUPPORTED) continue;
            CHECK(rc == SA_OK && got == need, "fill pass");
            if (need > 16) {  // a buffer that is too small must be refused, not overrun
                size_t g2 = 0;
                CHECK(sa_plan_pack_mode(p, world, mode, blob.data(), need - 16, &g2) == SA_ERR_SMALL_BUFFER, "small buffer");
            }
            CHECK(sa_blob_execute_cpu(blob.data(), got, walk.data()) == SA_OK, "blob walk");
            for (int i = 0; i < n; ++i)
                CHECK(walk[4 * i] == fast[3 * i] && walk[4 * i + 1] == fast[3 * i + 1] && walk[4 * i + 2] == fast[3 * i + 2], "walk == oracle");
        }
    }
    const size_t need = sa_format(fast.data(), n, nullptr, 0);
    std::string out(need, '\0');
    CHECK(sa_format(fast.data(), n, out.data(), need) == need, "format");
    const size_t ser = sa_problem_serialize(p, nullptr, 0);
    std::string text(ser, '\0');
    sa_problem_serialize(p, text.data(), ser);
    sa_problem* q = nullptr;
    CHECK(sa_parse(text.data(), text.size(), &q) == SA_OK && sa_problem_nseq(q) == n, "serialize round trip");
    sa_problem_free(q);
}

// Ranges of one batch packed by several threads at once (what sa_session_solve does for large batches): every
// range's blob, walked on the CPU, must give exactly the oracle's rows of that range. Under TSan this is the
// test of the packer's shared thread pool with several jobs open at the same time.
static void check_concurrent_ranges(const sa_problem* p) {
    const uint32_t n = (uint32_t)sa_problem_nseq(p);
    std::vector<int32_t> want((size_t)n * 3);
    CHECK(sa_solve_cpu_fast(p, want.data(), 0) == SA_OK, "oracle");
    const uint32_t nranges = 4;
    std::vector<int> ok(nranges, 0);
    std::vector<std::thread> th;
    for (uint32_t k = 0; k < nranges; ++k)
        th.emplace_back([&, k] {
            const uint32_t first = n * k / nranges, count = n * (k + 1) / nranges - first;
            size_t need = 0, got = 0;
            if (sa_plan_pack_range(p, first, count, 1, SA_MODE_AUTO, nullptr, 0, &need) != SA_OK) return;
            std::vector<unsigned char> blob(need);
            if (sa_plan_pack_range(p, first, count, 1, SA_MODE_AUTO, blob.data(), need, &got) != SA_OK || got != need) return;
            std::vector<int32_t> out((size_t)count * 4 + 4);
            if (sa_blob_execute_cpu(blob.data(), got, out.data()) != SA_OK) return;
            for (uint32_t i = 0; i < count; ++i)
                for (int c = 0; c < 3; ++c)
                    if (out[(size_t)i * 4 + c] != want[(size_t)(first + i) * 3 + c]) return;
            ok[k] = 1;
        });
    for (auto& t : th) t.join();
    for (uint32_t k = 0; k < nranges; ++k) CHECK(ok[k] == 1, "concurrently packed range differs from the oracle");
}

int main(int argc, char** argv) {
    for (int i = 1; i < argc; ++i) {
        const std::string text = slurp(argv[i]);
        sa_problem* p = nullptr;
        CHECK(sa_parse(text.data(), text.size(), &p) == SA_OK, argv[i]);
        if (!p) continue;
        check_problem(p, sa_problem_total_candidates(p) < 200000);
        sa_problem_free(p);
    }
    // malformed and hostile inputs: must be errors (or clean parses), never memory errors
    const char* bad[] = {"", " ", "1 2 3", "1 2 3 x", "1 2 3 4", "1 2 3 4 ABC", "1 2 3 4 ABC 2 AB", "1 2 3 4 ABC -1",
                         "99999999999999999999 1 1 1 A 1 A", "1 1 1 1 A 2147483647 A", "1 1 1 1 \xff\xfe\x80 1 \x01\x02",
                         "1 1 1 1 ABC 1 ABCDEFGHIJKLMNOPQRSTUVWXYZ", "-5 -6 -7 -8 abc 2 a bc", "1 1 1 1 A-*+/= 1 B!?"};
    for (const char* t : bad) {
        sa_problem* p = nullptr;
        const int rc = sa_parse(t, std::strlen(t), &p);
        if (rc == SA_OK && p) {
            if (sa_problem_nseq(p) < 1000) check_problem(p, true);
            sa_problem_free(p);
        }
    }
    // random problems, including a batch big enough to use the multi-threaded packer and STREAM planning
    std::mt19937 rng(7);
    const int reps = std::getenv("SA_SELFTEST_QUICK") ? 5 : 6;  // quick mode skips the 20000-sequence batch
    for (int rep = 0; rep < reps; ++rep) {
        const int nseq = rep == 5 ? 20000 : 1 + (int)(rng() % 40);
        const size_t len1 = rep == 5 ? 500 : 1 + rng() % 700;
        std::string s1(len1, 'A');
        for (auto& c : s1) c = (char)('A' + rng() % 26);
        std::vector<std::string> seqs((size_t)nseq);
        std::vector<const char*> ptr((size_t)nseq);
        std::vector<size_t> len((size_t)nseq);
        for (int i = 0; i < nseq; ++i) {
            const size_t l = 1 + rng() % (rep == 5 ? 300 : len1 + 3);
            seqs[(size_t)i].assign(l, 'A');
            for (auto& c : seqs[(size_t)i]) c = (char)((rng() % 50 == 0 ? 'a' : 'A') + rng() % 26);
            ptr[(size_t)i] = seqs[(size_t)i].data();
            len[(size_t)i] = l;
        }
        const int32_t w[4] = {(int32_t)(rng() % 30), (int32_t)(rng() % 9), (int32_t)(rng() % 9), (int32_t)(rng() % 9)};
        sa_problem* p = nullptr;
        CHECK(sa_problem_create(w, s1.data(), s1.size(), nseq, ptr.data(), len.data(), &p) == SA_OK, "create");
        if (p) {
            check_problem(p, rep < 3);
            if (rep == 5) check_concurrent_ranges(p);
            sa_problem_free(p);
        }
    }
    std::printf("core_selftest: %s (%d failures)\n", g_fail ? "FAILED" : "ok", g_fail);
    return g_fail ? 1 : 0;
}
