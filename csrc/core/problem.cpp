// Input grammar, problem container, pair classes and the result formatter.
//
// Reference behaviour being matched (not its code):
//   * main.c:76-108  — fscanf("%d %d %d %d"), "%s", "%d", N x "%s" on stdin; a-z upper-cased.
//   * main.c:59-60   — the two group lists; main.c:14-44 turns them into 27x27 tables.
//   * main.c:204,209 — "#%d: score: %d, n: %d, k: %d\n".
// Deliberate differences: the table is fully zero-initialised (the reference clears only
// 313 of 729 bytes, main.c:20-26), missing tokens are an error instead of garbage, and
// token lengths are not capped at 2999/1999 (myProto.h:3-4).
#include <cstdio>
#include <cstring>
#include <limits>
#include <algorithm>
#include <array>
#include <new>
#include <thread>
#include <vector>

#include "core.hpp"

namespace sa {

static thread_local std::string g_last_error;
void set_error(const std::string& msg) { g_last_error = msg; }

namespace {

// scanf's notion of white space (isspace in the "C" locale).
inline bool is_space(unsigned char c) {
    return c == ' ' || c == '\t' || c == '\n' || c == '\v' || c == '\f' || c == '\r';
}

struct Cursor {
    const unsigned char* p;
    const unsigned char* end;
    void skip_ws() {
        while (p < end && is_space(*p)) ++p;
    }
    // "%d": optional sign, then decimal digits; stops at the first other byte.
    bool read_int(int32_t* out) {
        skip_ws();
        const unsigned char* q = p;
        bool neg = false;
        if (q < end && (*q == '+' || *q == '-')) {
            neg = (*q == '-');
            ++q;
        }
        if (q >= end || *q < '0' || *q > '9') return false;
        int64_t v = 0;
        while (q < end && *q >= '0' && *q <= '9') {
            v = v * 10 + (*q - '0');
            if (v > (int64_t)std::numeric_limits<int32_t>::max() + 1) v = (int64_t)std::numeric_limits<int32_t>::max() + 1;
            ++q;
        }
        if (neg) v = -v;
        if (v > std::numeric_limits<int32_t>::max()) v = std::numeric_limits<int32_t>::max();
        if (v < std::numeric_limits<int32_t>::min()) v = std::numeric_limits<int32_t>::min();
        *out = (int32_t)v;
        p = q;
        return true;
    }
    // "%s": the next run of non-white-space bytes.
    bool read_token(std::string* out) {
        skip_ws();
        if (p >= end) return false;
        const unsigned char* q = p;
        while (q < end && !is_space(*q)) ++q;
        out->assign(reinterpret_cast<const char*>(p), q - p);
        p = q;
        return true;
    }
};

inline void upper_ascii(std::string& s) {
    for (char& c : s)
        if (c >= 'a' && c <= 'z') c = (char)(c - 'a' + 'A');
}

// Group membership bitmasks per letter, built once. Bit g of g1[c] is set when letter c
// belongs to type-1 group g; two letters share a group iff the masks intersect.
struct Groups {
    uint32_t g1[26] = {0};
    uint32_t g2[26] = {0};
    Groups() {
        static const char* const type1[] = {"NDEQ", "MILV", "FYW", "NEQK", "QHRK", "HY", "STA", "NHQK", "MILF"};
        static const char* const type2[] = {"SAG", "SGND", "NEQHRK", "HFY", "ATV", "STPA",
                                            "NDEQHK", "FVLIM", "CSA", "STNK", "SNDEQK"};
        for (unsigned g = 0; g < sizeof(type1) / sizeof(type1[0]); ++g)
            for (const char* c = type1[g]; *c; ++c) g1[*c - 'A'] |= 1u << g;
        for (unsigned g = 0; g < sizeof(type2) / sizeof(type2[0]); ++g)
            for (const char* c = type2[g]; *c; ++c) g2[*c - 'A'] |= 1u << g;
    }
};
const Groups& groups() {
    static const Groups g;
    return g;
}

}  // namespace

const Alphabet& letters_only_alphabet() {
    static const Alphabet letters = [] {
        Alphabet al;
        for (int i = 0; i < 256; ++i) al.code[i] = -2;  // any non-letter byte is "unknown" (negative)
        std::memset(al.byte_of, 0, sizeof(al.byte_of));
        for (int c = 0; c < 26; ++c) {
            al.code['A' + c] = (int16_t)(1 + c);
            al.byte_of[1 + c] = (unsigned char)('A' + c);
        }
        return al;
    }();
    return letters;
}

Alphabet build_alphabet(const sa_problem& p) {
    Alphabet al;
    for (int i = 0; i < 256; ++i) al.code[i] = -2;  // -2 = does not occur
    std::memset(al.byte_of, 0, sizeof(al.byte_of));
    for (int c = 0; c < 26; ++c) {
        al.code['A' + c] = (int16_t)(1 + c);
        al.byte_of[1 + c] = (unsigned char)('A' + c);
    }
    // which bytes occur at all (threads for big inputs; the answer does not depend on the split)
    const size_t n = p.seq2.size();
    size_t bytes = p.seq1.size();
    for (const auto& s : p.seq2) bytes += s.size();
    unsigned nt = bytes < (4u << 20) ? 1u : std::min(16u, std::max(1u, std::thread::hardware_concurrency()));
    std::vector<std::array<bool, 256>> seen(nt);
    for (auto& a : seen) a.fill(false);
    auto scan = [&](unsigned t) {
        auto& mine = seen[t];
        if (t == 0)
            for (unsigned char b : p.seq1) mine[b] = true;
        for (size_t i = n * t / nt; i < n * (t + 1) / nt; ++i)
            for (unsigned char b : p.seq2[i]) mine[b] = true;
    };
    if (nt == 1) {
        scan(0);
    } else {
        std::vector<std::thread> pool;
        for (unsigned t = 0; t < nt; ++t) pool.emplace_back(scan, t);
        for (auto& th : pool) th.join();
    }
    // non-letter bytes get the spare codes 27..31 in increasing byte order
    for (int b = 0; b < 256; ++b) {
        bool present = false;
        for (unsigned t = 0; t < nt; ++t) present = present || seen[t][(size_t)b];
        if (!present || al.code[b] != -2) continue;
        if (al.used < SA_ALPHABET) {
            al.code[b] = (int16_t)al.used;
            al.byte_of[al.used] = (unsigned char)b;
            ++al.used;
        } else {
            al.code[b] = -1;
            al.ok = false;
        }
    }
    return al;
}

void build_lut(const Alphabet& al, const int32_t w[4], int32_t lut[SA_ALPHABET * SA_ALPHABET]) {
    // Codes 1..26 are always 'A'..'Z', 27.. the problem's non-letter bytes (class 3 against everything but
    // themselves), so the class of a code pair does not depend on the problem: one table, built once. This runs
    // on every solve and was a third of the packing time of a 30-sequence problem when it asked sa_pair_class.
    static const struct Classes {
        unsigned char c[SA_ALPHABET][SA_ALPHABET];
        Classes() {
            for (int a = 0; a < SA_ALPHABET; ++a)
                for (int b = 0; b < SA_ALPHABET; ++b)
                    c[a][b] = (unsigned char)(a == b ? 0 : (a > 26 || b > 26) ? 3 : sa_pair_class((unsigned char)('A' + a - 1), (unsigned char)('A' + b - 1)));
        }
    } classes;
    const int32_t value[4] = {w[0], -w[1], -w[2], -w[3]};
    std::memset(lut, 0, sizeof(int32_t) * SA_ALPHABET * SA_ALPHABET);
    for (int a = 1; a < al.used; ++a)
        for (int b = 1; b < al.used; ++b) lut[a * SA_ALPHABET + b] = value[classes.c[a][b]];
}

}  // namespace sa

extern "C" {

const char* sa_last_error(void) { return sa::g_last_error.c_str(); }
void sa_set_last_error(const char* msg) { sa::g_last_error = msg ? msg : ""; }

int sa_pair_class(unsigned char a, unsigned char b) {
    if (a == b) return 0;
    if (a < 'A' || a > 'Z' || b < 'A' || b > 'Z') return 3;
    const sa::Groups& g = sa::groups();
    if (g.g1[a - 'A'] & g.g1[b - 'A']) return 1;
    if (g.g2[a - 'A'] & g.g2[b - 'A']) return 2;
    return 3;
}

int sa_parse(const char* text, size_t len, sa_problem** out) {
    if (!text || !out) {
        sa::set_error("sa_parse: null argument");
        return SA_ERR_ARG;
    }
    sa_problem* p = new (std::nothrow) sa_problem();
    if (!p) return SA_ERR_NOMEM;
    sa::Cursor cur{reinterpret_cast<const unsigned char*>(text), reinterpret_cast<const unsigned char*>(text) + len};
    for (int i = 0; i < 4; ++i) {
        if (!cur.read_int(&p->weights[i])) {
            sa::set_error("parse error: expected four integer weights");
            delete p;
            return SA_ERR_PARSE;
        }
    }
    if (!cur.read_token(&p->seq1)) {
        sa::set_error("parse error: missing Seq1");
        delete p;
        return SA_ERR_PARSE;
    }
    sa::upper_ascii(p->seq1);
    int32_t n = 0;
    if (!cur.read_int(&n) || n < 0) {
        sa::set_error("parse error: missing or negative sequence count");
        delete p;
        return SA_ERR_PARSE;
    }
    // every sequence needs at least one byte and one separator: a count larger than that can allow is
    // a malformed input, and must not be allowed to size an allocation (found by `make asan`)
    if ((size_t)n > (size_t)(cur.end - cur.p + 1) / 2 + 1) {
        sa::set_error("parse error: expected " + std::to_string(n) + " sequences, the input is too short to hold them");
        delete p;
        return SA_ERR_PARSE;
    }
    p->seq2.resize((size_t)n);
    for (int32_t i = 0; i < n; ++i) {
        if (!cur.read_token(&p->seq2[(size_t)i])) {
            sa::set_error("parse error: expected " + std::to_string(n) + " sequences, found " + std::to_string(i));
            delete p;
            return SA_ERR_PARSE;
        }
        sa::upper_ascii(p->seq2[(size_t)i]);
    }
    *out = p;
    return SA_OK;
}

int sa_problem_create(const int32_t weights[4], const char* seq1, size_t len1, int32_t nseq,
                      const char* const* seq2, const size_t* len2, sa_problem** out) {
    if (!weights || !seq1 || !out || nseq < 0 || (nseq > 0 && (!seq2 || !len2))) {
        sa::set_error("sa_problem_create: bad argument");
        return SA_ERR_ARG;
    }
    sa_problem* p = new (std::nothrow) sa_problem();
    if (!p) return SA_ERR_NOMEM;
    std::memcpy(p->weights, weights, sizeof(p->weights));
    p->seq1.assign(seq1, len1);
    sa::upper_ascii(p->seq1);
    p->seq2.resize((size_t)nseq);
    for (int32_t i = 0; i < nseq; ++i) {
        p->seq2[(size_t)i].assign(seq2[i], len2[i]);
        sa::upper_ascii(p->seq2[(size_t)i]);
    }
    *out = p;
    return SA_OK;
}

void sa_problem_free(sa_problem* p) { delete p; }

int32_t sa_problem_nseq(const sa_problem* p) { return (int32_t)p->seq2.size(); }
void sa_problem_weights(const sa_problem* p, int32_t w[4]) { std::memcpy(w, p->weights, sizeof(p->weights)); }
size_t sa_problem_len1(const sa_problem* p) { return p->seq1.size(); }
const char* sa_problem_seq1(const sa_problem* p) { return p->seq1.c_str(); }
size_t sa_problem_len2(const sa_problem* p, int32_t i) { return p->seq2[(size_t)i].size(); }
const char* sa_problem_seq2(const sa_problem* p, int32_t i) { return p->seq2[(size_t)i].c_str(); }

int64_t sa_problem_candidates(const sa_problem* p, int32_t i) {
    const int64_t l1 = (int64_t)p->seq1.size(), l2 = (int64_t)p->seq2[(size_t)i].size();
    if (l2 > l1) return 0;
    if (l2 == l1) return 1;
    return (l1 - l2) * l2;
}

int64_t sa_problem_total_candidates(const sa_problem* p) {
    int64_t t = 0;
    for (int32_t i = 0; i < (int32_t)p->seq2.size(); ++i) t += sa_problem_candidates(p, i);
    return t;
}

size_t sa_problem_serialize(const sa_problem* p, char* buf, size_t cap) {
    std::string s = std::to_string(p->weights[0]) + " " + std::to_string(p->weights[1]) + " " +
                    std::to_string(p->weights[2]) + " " + std::to_string(p->weights[3]) + "\n";
    s += p->seq1;
    s += "\n";
    s += std::to_string(p->seq2.size());
    s += "\n";
    for (const auto& q : p->seq2) {
        s += q;
        s += "\n";
    }
    if (buf && cap) {
        const size_t n = s.size() < cap ? s.size() : cap;
        std::memcpy(buf, s.data(), n);
    }
    return s.size();
}

size_t sa_format(const int32_t* results, int32_t nseq, char* buf, size_t cap) {
    size_t need = 0;
    char line[96];
    for (int32_t i = 0; i < nseq; ++i) {
        const int n = std::snprintf(line, sizeof(line), "#%d: score: %d, n: %d, k: %d\n", i, results[3 * i],
                                    results[3 * i + 1], results[3 * i + 2]);
        if (buf && need < cap) {
            const size_t room = cap - need;
            std::memcpy(buf + need, line, (size_t)n < room ? (size_t)n : room);
        }
        need += (size_t)n;
    }
    return need;
}

}  // extern "C"
