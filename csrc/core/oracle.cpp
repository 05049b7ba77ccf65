// CPU oracle: the intended, race-free semantics of the reference kernel
// (cudaFunctions.cu:63-176 as specified in SURVEY.md §8.2), stated twice:
//   * solve_one_naive — literal: every (n,k) candidate re-counts all len2 pairs,
//     O(len1*len2^2), same iteration order and strict-'>' update as cudaFunctions.cu:116-166.
//   * solve_one_fast  — independent O(len1*len2): prefix of offset n + suffix of offset n+1.
// The reference ships no CPU scoring code, so this is both the test oracle and the timed
// CPU baseline.
#include <atomic>
#include <climits>
#include <thread>

#include "core.hpp"

namespace sa {

namespace {

inline int32_t wrap32(int64_t v) { return (int32_t)(uint32_t)(uint64_t)v; }

struct PairValues {
    // value[a][b] for raw bytes; built per call (256x256 would be wasteful: use class table
    // over bytes lazily through sa_pair_class for non-letters, dense 26x26 for letters).
    int32_t letter[26][26];
    int32_t w[4];
    explicit PairValues(const int32_t weights[4]) {
        for (int i = 0; i < 4; ++i) w[i] = weights[i];
        for (int a = 0; a < 26; ++a)
            for (int b = 0; b < 26; ++b)
                letter[a][b] = class_value(sa_pair_class((unsigned char)('A' + a), (unsigned char)('A' + b)), w);
    }
    inline int32_t operator()(unsigned char a, unsigned char b) const {
        const unsigned ia = (unsigned)a - 'A', ib = (unsigned)b - 'A';
        if (ia < 26u && ib < 26u) return letter[ia][ib];
        return class_value(a == b ? 0 : 3, w);
    }
};

}  // namespace

Triple solve_one_naive(const int32_t w[4], const std::string& s1, const std::string& s2) {
    const int64_t l1 = (int64_t)s1.size(), l2 = (int64_t)s2.size();
    Triple best{INT_MIN, 0, 0};
    if (l2 > l1) return best;  // no candidates: the loops at cudaFunctions.cu:116 never run
    auto score_of = [&](int64_t n, int64_t k) {
        int64_t c[4] = {0, 0, 0, 0};
        for (int64_t i = 0; i < l2; ++i) {
            const int64_t j = (k == 0 || i < k) ? i + n : i + n + 1;  // cudaFunctions.cu:132-153
            ++c[sa_pair_class((unsigned char)s2[(size_t)i], (unsigned char)s1[(size_t)j])];
        }
        return wrap32((int64_t)w[0] * c[0] - (int64_t)w[1] * c[1] - (int64_t)w[2] * c[2] - (int64_t)w[3] * c[3]);
    };
    if (l2 == l1) return Triple{score_of(0, 0), 0, 0};  // cudaFunctions.cu:74-106
    for (int64_t n = 0; n < l1 - l2; ++n)
        for (int64_t k = 0; k < l2; ++k) {
            const int32_t s = score_of(n, k);
            if (best.score < s) best = Triple{s, (int32_t)n, (int32_t)k};
        }
    return best;
}

Triple solve_one_fast(const int32_t w[4], const std::string& s1, const std::string& s2) {
    const int64_t l1 = (int64_t)s1.size(), l2 = (int64_t)s2.size();
    Triple best{INT_MIN, 0, 0};
    if (l2 > l1) return best;
    if (l2 == 0 && l1 > 0) return best;  // an empty Seq2 has no (n,k) candidate at all (the k loop never runs)
    const PairValues val(w);
    auto diag_total = [&](int64_t n) {
        int64_t t = 0;
        for (int64_t i = 0; i < l2; ++i) t += val((unsigned char)s2[(size_t)i], (unsigned char)s1[(size_t)(i + n)]);
        return t;
    };
    if (l2 == l1) return Triple{wrap32(diag_total(0)), 0, 0};
    std::vector<int64_t> suf((size_t)l2 + 1);
    for (int64_t n = 0; n < l1 - l2; ++n) {
        // suf[k] = sum_{i>=k} value(s2[i], s1[i+n+1]) ; pre runs along offset n.
        suf[(size_t)l2] = 0;
        for (int64_t i = l2 - 1; i >= 0; --i)
            suf[(size_t)i] = suf[(size_t)i + 1] + val((unsigned char)s2[(size_t)i], (unsigned char)s1[(size_t)(i + n + 1)]);
        const int32_t s0 = wrap32(diag_total(n));
        if (best.score < s0) best = Triple{s0, (int32_t)n, 0};
        int64_t pre = 0;
        for (int64_t k = 1; k < l2; ++k) {
            pre += val((unsigned char)s2[(size_t)k - 1], (unsigned char)s1[(size_t)(k - 1 + n)]);
            const int32_t s = wrap32(pre + suf[(size_t)k]);
            if (best.score < s) best = Triple{s, (int32_t)n, (int32_t)k};
        }
    }
    return best;
}

namespace {

template <typename F>
int solve_all(const sa_problem* p, int32_t* out, int threads, F one) {
    if (!p || !out) {
        set_error("solve: null argument");
        return SA_ERR_ARG;
    }
    const int n = (int)p->seq2.size();
    int nt = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
    if (nt < 1) nt = 1;
    if (nt > n) nt = n > 0 ? n : 1;
    std::atomic<int> next{0};
    auto work = [&]() {
        for (;;) {
            const int i = next.fetch_add(1);
            if (i >= n) break;
            const Triple t = one(p->weights, p->seq1, p->seq2[(size_t)i]);
            out[3 * i] = t.score;
            out[3 * i + 1] = t.n;
            out[3 * i + 2] = t.k;
        }
    };
    if (nt == 1) {
        work();
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < nt; ++t) pool.emplace_back(work);
        for (auto& th : pool) th.join();
    }
    return SA_OK;
}

}  // namespace
}  // namespace sa

extern "C" {

int sa_solve_cpu_naive(const sa_problem* p, int32_t* out, int threads) {
    return sa::solve_all(p, out, threads, sa::solve_one_naive);
}
int sa_solve_cpu_fast(const sa_problem* p, int32_t* out, int threads) {
    return sa::solve_all(p, out, threads, sa::solve_one_fast);
}

}  // extern "C"
