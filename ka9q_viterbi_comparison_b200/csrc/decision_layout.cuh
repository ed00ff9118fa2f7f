// decision_layout.cuh -- how each kernel family stores one step's S decision bits, and how the chainback
// kernels (and the host-side canonicaliser in api.cu) find the bit of a given state.
#pragma once
#include "kernels.h"
#include "acs_k7.cuh"
#include "acs_grp.cuh"

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


// register-resident K=7 family: one u64 per (row, frame), frame-interleaved; the bit of state n depends on the
// phase of the step that wrote the row (acs_k7.cuh: decision_bit_of). `tab` is a [6][64] byte table in shared
// memory filled by fill_k7_table().
template <> struct DecisionReader<DL_K7_INTERLEAVED> {
    const unsigned long long* base;   // + frame
    size_t stride;                    // n_frames
    const uint8_t* row_phase;
    const uint8_t* tab;
    __device__ DecisionReader(const ChainbackArgs& a, int frame, const uint8_t* table)
        : base((const unsigned long long*)a.dec + frame), stride((size_t)a.n_frames), row_phase(a.row_phase),
          tab(table) {}
    __device__ __forceinline__ unsigned long long load(unsigned row) const { return __ldg(base + (size_t)row * stride); }
    __device__ __forceinline__ unsigned bit_of(unsigned long long word, unsigned row, unsigned state) const {
        return (unsigned)(word >> tab[row_phase[row] * k7::S + state]) & 1u;
    }
};

template <> struct DecisionReader<DL_SMALLK_INTERLEAVED> {
    const uint32_t* base;             // + frame
    size_t stride;                    // n_frames
    __device__ DecisionReader(const ChainbackArgs& a, int frame)
        : base((const uint32_t*)a.dec + frame), stride((size_t)a.n_frames) {}
    __device__ __forceinline__ unsigned bit(unsigned row, unsigned state) const {
        return (__ldg(base + (size_t)row * stride) >> state) & 1u;
    }
};

template <> struct DecisionReader<DL_CANONICAL> {
    const uint32_t* base;
    int row_words;
    __device__ DecisionReader(const ChainbackArgs& a, int frame)
        : base((const uint32_t*)a.dec + (size_t)frame * a.dec_rows * a.row_words), row_words(a.row_words) {}
    __device__ __forceinline__ unsigned bit(unsigned row, unsigned state) const {
        return (__ldg(base + (size_t)row * row_words + (state >> 5)) >> (state & 31u)) & 1u;
    }
};

// group family: one u64 per (row, frame, tx); the (tx, bit) of state n depends on the row's phase.
template <> struct DecisionReader<DL_GRP_INTERLEAVED> {
    const unsigned long long* base;   // + frame * TPF
    size_t stride;                    // n_frames * TPF
    const uint8_t* row_phase;
    int nxb;
    __device__ DecisionReader(const ChainbackArgs& a, int frame)
        : base((const unsigned long long*)a.dec + (size_t)frame * (a.c.S >> 6)),
          stride((size_t)a.n_frames * (a.c.S >> 6)), row_phase(a.row_phase), nxb(a.c.K - 1 - grp::NL) {}
    __device__ __forceinline__ unsigned bit(unsigned row, unsigned state) const {
        unsigned tx, b;
        grp::decision_pos(nxb, row_phase[row], state, tx, b);
        const uint32_t* p = reinterpret_cast<const uint32_t*>(base + (size_t)row * stride + tx);
        return (__ldg(p + (b >> 5)) >> (b & 31u)) & 1u;
    }
};

__device__ __forceinline__ void fill_k7_table(uint8_t* tab, int policy) {
    const bool bytewise = k7::decisions_bytewise(policy);
    for (int i = threadIdx.x; i < k7::NB * k7::S; i += blockDim.x)
        tab[i] = (uint8_t)k7::decision_bit_of(i / k7::S, i % k7::S, bytewise);
    __syncthreads();
}

}  // namespace b200v
