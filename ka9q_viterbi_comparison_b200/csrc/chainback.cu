// chainback.cu -- traceback kernels: sequential within a frame, parallel across frames (one thread per
// frame). Takes the role of chainback_viterbiXX_sse2 (viterbi27_sse2.cpp:78-105, viterbi615_sse2.cpp:65-91),
// chainback_spiralXX (spiral27.cpp:81-120) and ViterbiDecoder_Core::chainback
// (viterbi_decoder_core.h:214-236), which all walk the same register:
//     k = decision bit `state` of row n+K-1;  reg = (reg >> 1) | (k << top);  byte n/8 = reg >> sub
// and of chainback_viterbi224_sse2 (viterbi224_sse2.cpp:79-121), reproduced literally.
#include "kernels.h"
#include "decision_layout.cuh"

namespace b200v {

namespace {

template <int LAYOUT>
__global__ void chainback_kernel(const ChainbackArgs a) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= a.count) return;
    const int f = a.frame0 + idx;
    const int K = a.c.K;
    const DecisionReader<LAYOUT> rd(a, f);   // frame-major layouts only
    uint8_t* out = a.out + (size_t)f * a.out_stride;

    if (a.c.policy == P_KA9Q_S16_224) {
        // no look-past-tail; emits (state & 1) into a byte that shifts right; stored when n % 8 == 0
        unsigned state = (a.endstates ? a.endstates[f] : a.endstate) & ((1u << (K - 1)) - 1u);
        unsigned dbyte = 0;
        for (unsigned n = a.nbits; n-- > 0;) {
            dbyte = ((state & 1u) << 7) | (dbyte >> 1);
            if ((n & 7u) == 0) out[n >> 3] = (uint8_t)dbyte;
            const unsigned bit = rd.bit(n, state);
            state = (bit << (K - 2)) | (state >> 1);
        }
        if (a.results) a.results[f] = 0;
        return;
    }

    const int add = (K - 1 < 8) ? 8 - (K - 1) : 0;
    const int sub = (K - 1 > 8) ? (K - 1) - 8 : 0;
    unsigned reg = (a.endstates ? a.endstates[f] : a.endstate) % (unsigned)a.c.S;
    if (a.results) {
        int r = 0;
        if (a.c.policy == P_KA9Q_S16_615) r = (int)(int16_t)a.metrics[(size_t)f * a.c.S + reg];
        a.results[f] = r;
    }
    reg <<= add;
    const unsigned top = (unsigned)(K - 2 + add);
    for (unsigned n = a.nbits; n-- > 0;) {
        const unsigned k = rd.bit(n + (unsigned)(K - 1), reg >> add);
        reg = (reg >> 1) | (k << top);
        // the reference stores every iteration; the value that sticks is the one written at the lowest
        // n of each byte, so store only there
        if ((n & 7u) == 0) out[n >> 3] = (uint8_t)(reg >> sub);
    }
}

// K=7 register-resident family. The decision word of a row does not depend on the traced state, so the
// loads are issued CHUNK rows ahead of the dependent bit extraction: the kernel streams the decision buffer
// at memory bandwidth instead of paying one DRAM latency per decoded bit.
__global__ void __launch_bounds__(128) chainback_k7_kernel(const ChainbackArgs a) {
    __shared__ uint8_t tab[k7::NB * k7::S];
    fill_k7_table(tab, a.c.policy);
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= a.count) return;
    const int f = a.frame0 + idx;
    constexpr int K = 7, CHUNK = 16;
    const DecisionReader<DL_K7_INTERLEAVED> rd(a, f, tab);
    uint8_t* out = a.out + (size_t)f * a.out_stride;
    // K-1 = 6 < 8: the register keeps the state in its top 6 bits of a byte (ADDSHIFT = 2)
    unsigned reg = ((a.endstates ? a.endstates[f] : a.endstate) % 64u) << 2;
    if (a.results) a.results[f] = 0;
    unsigned n = a.nbits;
    while (n > 0) {
        const unsigned cnt = n < (unsigned)CHUNK ? n : (unsigned)CHUNK;
        unsigned long long wbuf[CHUNK];
#pragma unroll
        for (int i = 0; i < CHUNK; i++)
            if ((unsigned)i < cnt) wbuf[i] = rd.load(n - 1 - i + (K - 1));
#pragma unroll
        for (int i = 0; i < CHUNK; i++) {
            if ((unsigned)i < cnt) {
                const unsigned row = n - 1 - i + (K - 1);
                const unsigned k = rd.bit_of(wbuf[i], row, reg >> 2);
                reg = (reg >> 1) | (k << 7);
                if (((n - 1 - i) & 7u) == 0) out[(n - 1 - i) >> 3] = (uint8_t)reg;
            }
        }
        n -= cnt;
    }
}

// smallest path metric of every frame: one block per frame, packed (key << 32 | state) minimum
__global__ void best_state_kernel(const CodecParams c, const uint16_t* metrics, unsigned* best) {
    const int f = blockIdx.x;
    const uint16_t* m = metrics + (size_t)f * c.S;
    const int m0 = (int)m[0];
    unsigned long long b = ~0ull;
    for (int i = threadIdx.x; i < c.S; i += blockDim.x) {
        unsigned key;
        if (c.policy == P_KA9Q_U8_MOD) key = (unsigned)((int)(int8_t)(m[i] - m0) + 128);
        else if (c.policy == P_KA9Q_S16_615 || c.policy == P_KA9Q_S16_224) key = (unsigned)((int)(int16_t)m[i] + 32768);
        else key = m[i];
        const unsigned long long v = ((unsigned long long)key << 32) | (unsigned)i;
        b = v < b ? v : b;
    }
    __shared__ unsigned long long s[32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long t = __shfl_xor_sync(0xffffffffu, b, o);
        b = t < b ? t : b;
    }
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = b;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); w++) b = s[w] < b ? s[w] : b;
        best[f] = (unsigned)b;
    }
}

}  // namespace

cudaError_t launch_best_state(const CodecParams& c, const uint16_t* metrics, unsigned* best, int n_frames, cudaStream_t st) {
    if (n_frames <= 0) return cudaSuccess;
    const int threads = c.S >= 256 ? 256 : (c.S < 32 ? 32 : c.S);
    best_state_kernel<<<n_frames, threads, 0, st>>>(c, metrics, best);
    return cudaGetLastError();
}

cudaError_t launch_chainback(const ChainbackArgs& a, cudaStream_t st) {
    if (a.count <= 0) return cudaSuccess;
    const int threads = 128;
    const int blocks = (a.count + threads - 1) / threads;
    switch (a.layout) {
    case DL_BALLOT_FRAME_MAJOR: chainback_kernel<DL_BALLOT_FRAME_MAJOR><<<blocks, threads, 0, st>>>(a); break;
    case DL_K7_INTERLEAVED: chainback_k7_kernel<<<blocks, threads, 0, st>>>(a); break;
    case DL_GRP_INTERLEAVED: chainback_kernel<DL_GRP_INTERLEAVED><<<blocks, threads, 0, st>>>(a); break;
    case DL_CANONICAL: chainback_kernel<DL_CANONICAL><<<blocks, threads, 0, st>>>(a); break;
    case DL_SMALLK_INTERLEAVED: chainback_kernel<DL_SMALLK_INTERLEAVED><<<blocks, threads, 0, st>>>(a); break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

}  // namespace b200v
