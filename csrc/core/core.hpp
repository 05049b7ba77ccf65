// Internal declarations shared by the core translation units.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "seqalign.h"

struct sa_problem {
    int32_t weights[4] = {0, 0, 0, 0};
    std::string seq1;               // upper-cased
    std::vector<std::string> seq2;  // upper-cased
};

namespace sa {

void set_error(const std::string& msg);

// Pair values with the weights folded in: +w1 / -w2 / -w3 / -w4 for the four
// classes (cudaFunctions.cu:103,161 applies the weights after counting; folding
// them per pair is the same sum).
inline int32_t class_value(int cls, const int32_t w[4]) {
    switch (cls) {
        case 0: return w[0];
        case 1: return -w[1];
        case 2: return -w[2];
        default: return -w[3];
    }
}

// Byte -> alphabet code map for one problem. code[b] in 0..31, or -1 when the
// problem uses more than five distinct non-letter bytes.
struct Alphabet {
    int16_t code[256];
    unsigned char byte_of[SA_ALPHABET];  // inverse (0 for pad / unused)
    int used = 27;
    bool ok = true;
};
Alphabet build_alphabet(const sa_problem& p);
const Alphabet& letters_only_alphabet();  // A-Z only; every other byte maps to a negative code

// value(seq2 code a, seq1 code b) laid out lut[a*32+b]; pad code scores 0.
void build_lut(const Alphabet& al, const int32_t w[4], int32_t lut[SA_ALPHABET * SA_ALPHABET]);

struct Triple {
    int32_t score, n, k;
};
Triple solve_one_naive(const int32_t w[4], const std::string& s1, const std::string& s2);
Triple solve_one_fast(const int32_t w[4], const std::string& s1, const std::string& s2);

}  // namespace sa
