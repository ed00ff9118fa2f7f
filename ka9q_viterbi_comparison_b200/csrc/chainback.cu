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
__global__ void __launch_bounds__(256) chainback_k7_kernel(const ChainbackArgs a) {
    __shared__ uint8_t tab[k7::NB * k7::S];
    fill_k7_table(tab, a.c.policy);
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= a.count) return;
    const int f = a.frame0 + idx;
    constexpr int K = 7, CHUNK = 8;
    const DecisionReader<DL_K7_INTERLEAVED> rd(a, f, tab);
    uint8_t* out = a.out + (size_t)f * a.out_stride;
    // K-1 = 6 < 8: the register keeps the state in its top 6 bits of a byte (ADDSHIFT = 2)
    unsigned reg = ((a.endstates ? a.endstates[f] : a.endstate) % 64u) << 2;
    if (a.results) a.results[f] = 0;
    unsigned n = a.nbits;
    // ragged top (nbits not a multiple of 8): guarded, bit by bit, down to the next byte boundary
    while (n & 7u) {
        const unsigned row = n - 1 + (K - 1);
        const unsigned k = rd.bit_of(rd.load(row), row, reg >> 2);
        reg = (reg >> 1) | (k << 7);
        n--;
        if ((n & 7u) == 0) out[n >> 3] = (uint8_t)reg;     // the value written at the lowest n of the byte sticks
    }
    // whole bytes: 8 rows in flight, no guards; the table row (phase of the step that wrote the decision row) is the
    // same for every frame, the bit position inside the 64-bit word comes from shared memory
    const unsigned long long* base = rd.base;
    const size_t stride = rd.stride;
    while (n > 0) {
        unsigned long long wbuf[CHUNK];
        unsigned ph[CHUNK];
#pragma unroll
        for (int i = 0; i < CHUNK; i++) {
            const unsigned row = n - 1 - i + (K - 1);
            wbuf[i] = __ldg(base + (size_t)row * stride);
            ph[i] = (unsigned)__ldg(a.row_phase + row) * k7::S;
        }
#pragma unroll
        for (int i = 0; i < CHUNK; i++) {
            const unsigned sh = tab[ph[i] + (reg >> 2)];
            const unsigned k = (unsigned)(wbuf[i] >> sh) & 1u;
            reg = (reg >> 1) | (k << 7);
        }
        n -= CHUNK;
        out[n >> 3] = (uint8_t)reg;
    }
}

// ---- speculative block-parallel traceback (small and medium batches) -----------------------------------------
// The traceback is a pointer chase: one dependent decision-bit load per decoded bit, ~100-270 ns each when a
// frame has the GPU to itself (8192 bits: ~0.9 ms through main.cpp's chainback call; the reference CPU: 12 us).
// It parallelises exactly: the state after K-1 backward steps consists of the last K-1 decision bits only, so two
// traces that ever hold the same state at the same row stay identical. One CUDA block per frame; thread b decodes
// bits [b*L, (b+1)*L). It starts D rows further up from a guessed state (0) -- or from the true end state when that
// point is the end of the frame -- and walks down; after the D warm-up rows its state has merged with the true
// survivor with overwhelming probability. The guess is then VERIFIED: thread b's state on entering its own bits
// must equal the state thread b+1 left its bits with. The topmost thread is exact by construction, so if every
// boundary matches, every thread traced the reference's path (induction downwards) and the output is bit-exact. On
// a mismatch the failing threads re-trace their own bits from the state the thread above handed down, and the
// check repeats: each round makes at least the highest failing block exact, so it terminates (worst case: as many
// dependent steps as the sequential kernel).
template <int LAYOUT> struct ParReader {
    DecisionReader<LAYOUT> rd;
    __device__ ParReader(const ChainbackArgs& a, int f, const uint8_t*) : rd(a, f) {}
    __device__ __forceinline__ unsigned bit(unsigned row, unsigned state) const { return rd.bit(row, state); }
};
template <> struct ParReader<DL_K7_INTERLEAVED> {
    DecisionReader<DL_K7_INTERLEAVED> rd;
    __device__ ParReader(const ChainbackArgs& a, int f, const uint8_t* tab) : rd(a, f, tab) {}
    __device__ __forceinline__ unsigned bit(unsigned row, unsigned state) const { return rd.bit_of(rd.load(row), row, state); }
};

template <int LAYOUT>
__global__ void __launch_bounds__(1024) chainback_par_kernel(const ChainbackArgs a, const unsigned L, const unsigned D) {
    extern __shared__ unsigned s_exit[];                 // state each thread left its bits with
    __shared__ uint8_t tab[LAYOUT == DL_K7_INTERLEAVED ? k7::NB * k7::S : 1];
    if (LAYOUT == DL_K7_INTERLEAVED) fill_k7_table(tab, a.c.policy);
    const unsigned B = blockDim.x, b = threadIdx.x;
    const int f = a.frame0 + blockIdx.x;
    const unsigned K1 = (unsigned)(a.c.K - 1), N = a.nbits;
    const ParReader<LAYOUT> rd(a, f, tab);
    uint8_t* out = a.out + (size_t)f * a.out_stride;
    const unsigned end_state = (a.endstates ? a.endstates[f] : a.endstate) % (unsigned)a.c.S;
    if (b == 0) {
        if (a.results) a.results[f] = a.c.policy == P_KA9Q_S16_615 ? (int)(int16_t)a.metrics[(size_t)f * a.c.S + end_state] : 0;
    }
    const unsigned lo = b * L, hi = min(N, lo + L);      // this thread's decoded bits (lo < N for every thread)
    const unsigned n0 = min(N, hi + D);
    unsigned s = n0 == N ? end_state : 0u;
    for (unsigned n = n0; n-- > hi;) s = (s >> 1) | (rd.bit(n + K1, s) << (K1 - 1));
    unsigned entry = s;
    bool redo = true;
    for (;;) {
        if (redo) {
            unsigned cur = 0;
            for (unsigned n = hi; n-- > lo;) {
                const unsigned k = rd.bit(n + K1, s);
                s = (s >> 1) | (k << (K1 - 1));
                cur = (cur >> 1) | (k << 7);
                if ((n & 7u) == 0) out[n >> 3] = (uint8_t)cur;
            }
            s_exit[b] = s;
        }
        __syncthreads();
        // verification: the state this thread entered its bits with must be the state the thread above left with.
        // The topmost thread is exact, so the highest failing thread is repaired by one re-trace from that state:
        // every round extends the exact prefix by at least one block (usually all of them), at most B rounds.
        const unsigned want = b + 1 < B ? s_exit[b + 1] : entry;
        redo = want != entry;
        if (!__syncthreads_or(redo)) break;              // also orders this round's reads before the next writes
        s = entry = want;
    }
}

// geometry of the block-parallel traceback for `count` frames of `nbits` bits, or false: use the per-frame kernels
static bool par_geometry(const ChainbackArgs& a, unsigned& L, unsigned& D, unsigned& B) {
    if (a.par_mode == 1 || a.c.policy == P_KA9Q_S16_224 || (a.nbits & 7u) || a.nbits < 64u) return false;
    const bool forced = a.par_mode == 2;
    unsigned budget = min(1024u, (262144u / (unsigned)a.count) & ~31u);            // threads per frame
    if (forced) budget = max(budget, 32u);
    if (budget < 32u) return false;          // more than 8192 frames per launch: enough independent chases in flight already
    D = a.par_warmup > 0 ? (unsigned)a.par_warmup : max(64u, 12u * (unsigned)(a.c.K - 1));
    D = (D + 7u) & ~7u;
    const unsigned Lmin = forced ? 8u : 32u;
    L = max(Lmin, ((a.nbits + budget - 1) / budget + 7u) & ~7u);
    // Automatic choice. Up to ~16 Mi decoded bits per launch a frame-at-a-time traceback is purely latency-bound and
    // the warm-up rows cost nothing (measured, 2048-bit frames: K=15 1024 frames 1.57 -> 0.35 ms, K=9 4096 frames
    // 1.10 -> 0.74 ms). Larger launches are limited by 32-byte sectors fetched per decision bit, so there the bits
    // of a block are at least as many as its warm-up rows (at most 2x the sectors of the per-frame kernels):
    // 16384-bit frames, K=15 x 2048 frames 12.3 -> 1.9 ms, K=9 x 4096 frames 9.4 -> 2.6 ms.
    // The K=7 register layout keeps its per-frame kernel there: it streams whole rows, coalesced (4096 frames x
    // 16384 bits: 2.17 ms against 2.36 ms block-parallel).
    // Round 2, K=7 register layout re-measured by frame count (device time of the traceback, per-frame / block-parallel):
    // 2048 x 65536 bits 10.06 / 4.13 ms, 4096 x 16384 2.50 / 2.33, 8192 x 16384 2.34 / 3.67: the streaming kernel needs
    // thousands of frames in flight, so it takes over above 4096 frames only.
    if (!forced && (unsigned long long)a.count * a.nbits > (1ull << 24)) {
        if (a.layout == DL_K7_INTERLEAVED && a.count > 4096) return false;
        L = max(L, D);
    }
    B = (a.nbits + L - 1) / L;
    return forced ? B >= 2u : (B >= 4u && 4u * (L + D) <= a.nbits);
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

// ---- sliding-window traceback: one thread per stream walks the ring from the stream's end state ----
template <int LAYOUT>
__global__ void __launch_bounds__(128) stream_traceback_kernel(const StreamTracebackArgs s) {
    __shared__ uint8_t tab[LAYOUT == DL_K7_INTERLEAVED ? k7::NB * k7::S : 1];
    if (LAYOUT == DL_K7_INTERLEAVED) fill_k7_table(tab, s.a.c.policy);
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= s.a.count) return;
    const int f = s.a.frame0 + idx;
    const ParReader<LAYOUT> rd(s.a, f, tab);
    const unsigned K1 = (unsigned)(s.a.c.K - 1), W = s.ring_rows;
    uint8_t* out = s.a.out + (size_t)f * s.a.out_stride;
    unsigned st = (s.a.endstates ? s.a.endstates[f] : s.a.endstate) % (unsigned)s.a.c.S;
    const unsigned long long e0 = s.first_bit, e1 = s.first_bit + s.a.nbits;     // emitted bits [e0, e1)
    unsigned cur = 0;
    auto emit = [&](unsigned long long b, unsigned k) {
        if (b >= e1) return;
        cur = (cur >> 1) | (k << 7);
        if (((b - e0) & 7ull) == 0) { out[(b - e0) >> 3] = (uint8_t)cur; cur = 0; }
    };
    unsigned long long b = s.top_row + 1;           // one past the newest input bit known so far
    if (s.state_bits)                                // bit j of the state = input bit of step top_row - j
        for (unsigned j = 0; j < K1 && b > e0; j++) { b--; emit(b, (st >> j) & 1u); }
    else
        b = s.top_row + 1 - K1;
    // row r decides input bit r - (K-1); the state before the walk step is the one after row r
    unsigned long long r = s.top_row;
    unsigned p = (unsigned)(r % W);
    while (b > e0) {
        b--;
        const unsigned k = rd.bit(p, st);
        st = (st >> 1) | (k << (K1 - 1));
        emit(b, k);
        p = p == 0 ? W - 1 : p - 1;
    }
}

}  // namespace

cudaError_t launch_stream_traceback(const StreamTracebackArgs& s, cudaStream_t st) {
    if (s.a.count <= 0 || s.a.nbits == 0) return cudaSuccess;
    const int threads = 128, blocks = (s.a.count + threads - 1) / threads;
    switch (s.a.layout) {
    case DL_BALLOT_FRAME_MAJOR: stream_traceback_kernel<DL_BALLOT_FRAME_MAJOR><<<blocks, threads, 0, st>>>(s); break;
    case DL_K7_INTERLEAVED: stream_traceback_kernel<DL_K7_INTERLEAVED><<<blocks, threads, 0, st>>>(s); break;
    case DL_GRP_INTERLEAVED: stream_traceback_kernel<DL_GRP_INTERLEAVED><<<blocks, threads, 0, st>>>(s); break;
    case DL_SMALLK_INTERLEAVED: stream_traceback_kernel<DL_SMALLK_INTERLEAVED><<<blocks, threads, 0, st>>>(s); break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

cudaError_t launch_best_state(const CodecParams& c, const uint16_t* metrics, unsigned* best, int n_frames, cudaStream_t st) {
    if (n_frames <= 0) return cudaSuccess;
    const int threads = c.S >= 256 ? 256 : (c.S < 32 ? 32 : c.S);
    best_state_kernel<<<n_frames, threads, 0, st>>>(c, metrics, best);
    return cudaGetLastError();
}

cudaError_t launch_chainback(const ChainbackArgs& a, cudaStream_t st) {
    if (a.count <= 0) return cudaSuccess;
    unsigned L, D, B;
    if (par_geometry(a, L, D, B)) {
        const size_t sm = (size_t)B * sizeof(unsigned);
        switch (a.layout) {
        case DL_BALLOT_FRAME_MAJOR: chainback_par_kernel<DL_BALLOT_FRAME_MAJOR><<<a.count, B, sm, st>>>(a, L, D); break;
        case DL_K7_INTERLEAVED: chainback_par_kernel<DL_K7_INTERLEAVED><<<a.count, B, sm, st>>>(a, L, D); break;
        case DL_GRP_INTERLEAVED: chainback_par_kernel<DL_GRP_INTERLEAVED><<<a.count, B, sm, st>>>(a, L, D); break;
        case DL_CANONICAL: chainback_par_kernel<DL_CANONICAL><<<a.count, B, sm, st>>>(a, L, D); break;
        case DL_SMALLK_INTERLEAVED: chainback_par_kernel<DL_SMALLK_INTERLEAVED><<<a.count, B, sm, st>>>(a, L, D); break;
        default: return cudaErrorInvalidValue;
        }
        return cudaGetLastError();
    }
    const int threads = 128;
    const int blocks = (a.count + threads - 1) / threads;
    switch (a.layout) {
    case DL_BALLOT_FRAME_MAJOR: chainback_kernel<DL_BALLOT_FRAME_MAJOR><<<blocks, threads, 0, st>>>(a); break;
    case DL_K7_INTERLEAVED: chainback_k7_kernel<<<(a.count + 255) / 256, 256, 0, st>>>(a); break;   // 256 threads, 8 rows ahead: 3.46 -> 3.27 ms per 1 Mi frames (128 / 16: 3.46, 128 / 8: 3.32, 128 / 32: 4.40, 512 / 8: 3.25)
    case DL_GRP_INTERLEAVED: chainback_kernel<DL_GRP_INTERLEAVED><<<blocks, threads, 0, st>>>(a); break;
    case DL_CANONICAL: chainback_kernel<DL_CANONICAL><<<blocks, threads, 0, st>>>(a); break;
    case DL_SMALLK_INTERLEAVED: chainback_kernel<DL_SMALLK_INTERLEAVED><<<blocks, threads, 0, st>>>(a); break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

}  // namespace b200v
