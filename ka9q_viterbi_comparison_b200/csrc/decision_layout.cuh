// decision_layout.cuh -- how each kernel family stores one step's S decision bits, and how the chainback
// kernels (and the host-side canonicaliser in api.cu) find the bit of a given state.
#pragma once
#include "kernels.h"

namespace b200v {

// state n -> (word, bit) inside a ballot-packed row: butterfly j = n>>1 wrote bit (j & 31) of word
// 2*(j>>5) + (n & 1)
__host__ __device__ __forceinline__ void ballot_position(unsigned n, unsigned& word, unsigned& bit) {
    const unsigned j = n >> 1;
    word = ((j >> 5) << 1) | (n & 1u);
    bit = j & 31u;
}

template <int LAYOUT> struct DecisionReader;

template <> struct DecisionReader<DL_BALLOT_FRAME_MAJOR> {
    const uint32_t* base;
    int row_words;
    __device__ DecisionReader(const ChainbackArgs& a, int frame)
        : base((const uint32_t*)a.dec + (size_t)frame * a.dec_rows * a.row_words), row_words(a.row_words) {}
    __device__ __forceinline__ unsigned bit(unsigned row, unsigned state) const {
        unsigned w, b;
        ballot_position(state, w, b);
        return (__ldg(base + (size_t)row * row_words + w) >> b) & 1u;
    }
};

}  // namespace b200v
