// acs_k7_warp.cu -- the OTHER K=7 design BASELINE.json's configs[1] names: one WARP per frame, butterflies across
// lanes with __shfl_sync, decisions packed with __ballot_sync. Kernel family 3: any polynomial set, any rate, every
// policy a K=7 codec can have. Two roles:
//   * the measured comparison that backs the thread-per-frame register kernel (acs_k7.cu): a warp spends ~30
//     instructions per trellis step on ONE frame, the register kernel 242 on 32 frames
//     (profiles/r01_k7_warp_vs_thread.json: 78.0 vs 15.9 ms per 1 Mi frames);
//   * the automatic family for K=7 codes WITHOUT a compile-time instance of the register kernel (other
//     polynomials): ~4x faster than the generic shared-memory kernel.
//
// Lane j holds old[j] and old[j+32] = both inputs of butterfly j; the per-policy arithmetic is policy.cuh's (the
// same functions the generic kernel uses: viterbi27_sse2.cpp:137-158, spiral27.cpp:164-254, spiral47.cpp:143-219,
// sse_u8.h:92-154, sse_u16.h:93-177). New state n lives in lane n & 31, slot n >> 5, so lane r takes new[r] from
// butterfly r >> 1 and new[r + 32] from butterfly 16 + (r >> 1). The two ballots of a step are exactly the generic
// family's decision row (DL_BALLOT_FRAME_MAJOR, row_words = 2), so chainback is shared. Renormalisation: new[0] is
// lane 0's first slot (one shuffle), the minimum over the 64 states one REDUX.
#include "kernels.h"
#include "policy.cuh"

namespace b200v {

namespace {

// policy.cuh's branch_metric with the loop over the R symbols unrolled (RT > 0: compile-time rate), so that the
// symbols stay in registers
template <int POLICY, int RT>
__device__ __forceinline__ int branch_metric_rt(const CodecParams& c, unsigned cw, const int* sym) {
    constexpr int RMAX = RT ? RT : MAX_R;
    const int R = RT ? RT : c.R;
    if (POLICY == P_WY_U8 || POLICY == P_WY_U16) {
        int tot = 0;
#pragma unroll
        for (int i = 0; i < RMAX; i++) {
            if (i < R) {
                const int bt = ((cw >> i) & 1u) ? c.soft_hi : c.soft_lo;
                int e = bt - sym[i];
                if (POLICY == P_WY_U8) { e = clamp_i(e, -128, 127); e = abs(e); tot = min(tot + e, 255); }
                else { e = clamp_i(e, -32768, 32767); e = abs(e); tot = min(tot + e, 65535); }
            }
        }
        return tot;
    }
    if (POLICY == P_KA9Q_U8_MOD || POLICY == P_SPIRAL_R2) return branch_metric<POLICY>(c, cw, sym);
    int s = 0;
#pragma unroll
    for (int i = 0; i < RMAX; i++) {
        if (i < R) {
            const int x = ((cw >> i) & 1u) ? (sym[i] ^ 255) : sym[i];
            s = min(s + ((x >> 2) & 63), 255);
        }
    }
    return (s >> 2) & 63;
}

// ---- latency helpers (round 2) ------------------------------------------------------------------------------------
// r = 1/2 codes: the lane that fetched a step's symbols also evaluates the step's FOUR branch metrics (code words 0..3)
// and packs them into one word (two for 16-bit metrics); a step then broadcasts that word and every lane picks its own
// code word's field with a shift -- instead of broadcasting the symbols and evaluating the metric in all 32 lanes (~9
// instructions per step off every lane's stream).
#ifndef WARP_BM_PRE
#define WARP_BM_PRE 1
#endif
template <int POLICY> struct BmPack {
    static constexpr bool wide = (POLICY == P_WY_U16);
    static constexpr int WORDS = wide ? 2 : 1;
    __device__ __forceinline__ static void pack(const CodecParams& c, const int* sym, uint32_t* out) {
        const uint32_t b0 = (uint32_t)branch_metric<POLICY>(c, 0u, sym), b1 = (uint32_t)branch_metric<POLICY>(c, 1u, sym);
        const uint32_t b2 = (uint32_t)branch_metric<POLICY>(c, 2u, sym), b3 = (uint32_t)branch_metric<POLICY>(c, 3u, sym);
        if (wide) { out[0] = b0 | (b1 << 16); out[1] = b2 | (b3 << 16); }
        else out[0] = b0 | (b1 << 8) | (b2 << 16) | (b3 << 24);
    }
    // field of code word cw out of the broadcast word(s)
    __device__ __forceinline__ static int pick(const uint32_t* w, unsigned cw) {
        if (wide) return (int)(((cw & 2u) ? w[1] : w[0]) >> ((cw & 1u) * 16) & 0xffffu);
        return (int)((w[0] >> (cw * 8u)) & 0xffu);
    }
};
// ka9q modular policy on UNWRAPPED metrics (reference metric = U mod 256): the candidate difference stays below 128
// (bound: acs_k7.cu header; K=9: the CPU suite), so the reference's (int8)(m0 - m1) > 0 is the plain compare and the
// mod-256 masking leaves the dependent chain: add -> min instead of add -> and -> sub -> sign-extend -> compare -> select.
__device__ __forceinline__ void acs_mod_unwrapped(int lo, int hi, int bm, int& s0, int& s1, int& d0, int& d1) {
    const int cm = 15 - bm;
    const int m0 = lo + bm, m1 = hi + cm, m2 = lo + cm, m3 = hi + bm;
    d0 = m0 > m1; d1 = m2 > m3;
    s0 = min(m0, m1); s1 = min(m2, m3);
}
#ifndef WARP_MOD_UNWRAPPED
#define WARP_MOD_UNWRAPPED 1
#endif
#ifndef WARP_PACK_SURVIVORS
#define WARP_PACK_SURVIVORS 0      // K=7: both survivors of a butterfly in one shuffled word (2 shuffles instead of 4 per step): measured no faster for one frame (416 us either way) and slower for 1024 (1.75 vs 1.61 ms); off
#endif

// RT = code rate known at compile time (2, 3, 4) or 0: run-time c.R (up to MAX_R).
// Latency matters here as much as throughput -- this is the family a single frame through the reference's
// create/init/update/chainback API runs on (src/main.cpp:264-278) -- so nothing in the step loop waits on global
// memory: the symbols of 32 steps are fetched at once (lane l holds step t0 + l, the next 32 steps are in flight
// while these run) and a step broadcasts its words with one shuffle each.
template <int POLICY, int RT>
__global__ void __launch_bounds__(128) acs_k7_warp_kernel(const GenericArgs a) {
    const int lane = threadIdx.x & 31;
    const int fidx = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (fidx >= a.count) return;                                   // whole warps leave together
    const int frame = a.frame0 + fidx;
    const CodecParams& c = a.c;
    const int R = RT ? RT : c.R;
    constexpr int SYMB = (POLICY == P_WY_U16) ? 2 : 1;
    constexpr int WPS = RT ? (RT * SYMB + 3) / 4 : (MAX_R * SYMB + 3) / 4;     // 32-bit words per step
    const int step_bytes = R * SYMB;
    uint16_t* gm = a.metrics + (size_t)frame * 64;
    int lo = gm[lane], hi = gm[lane + 32];
    if (PolicyTraits<POLICY>::is_signed16) { lo = (int16_t)lo; hi = (int16_t)hi; }
    constexpr bool UNWRAP = WARP_MOD_UNWRAPPED && POLICY == P_KA9Q_U8_MOD;
    constexpr bool PRE = WARP_BM_PRE && RT == 2;
    constexpr int CW = PRE ? BmPack<POLICY>::WORDS : WPS;          // words a step broadcasts
    if (UNWRAP) {      // around state 0 (lane 0's lo): U = 256 + m0 + (int8)(m - m0)
        const int m0 = __shfl_sync(0xffffffffu, lo, 0);
        lo = 256 + m0 + (int)(int8_t)(lo - m0);
        hi = 256 + m0 + (int)(int8_t)(hi - m0);
    }
    const unsigned cw = a.cw[lane];                                // code word of butterfly `lane` (host table)
    const unsigned char* sp = (const unsigned char*)a.syms + (size_t)frame * a.sym_stride;
    unsigned long long rsum = 0;     // minima subtracted by this launch's renormalisations (uniform over the warp)
    uint32_t* drow = a.dec + ((size_t)frame * a.dec_rows + a.step0) * 2;
    const int src0 = lane >> 1, src1 = 16 + (lane >> 1);
    const bool odd = lane & 1;

    auto symbols_of = [&](const uint32_t* sw, int* sym) {
#pragma unroll
        for (int r = 0; r < (RT ? RT : MAX_R); r++) {
            if (r < R) {
                const int bpos = r * SYMB;
                const uint32_t v = sw[bpos >> 2] >> ((bpos & 3) * 8);
                sym[r] = (POLICY == P_WY_U16) ? (int)(int16_t)(v & 0xffffu)
                       : (POLICY == P_WY_U8) ? (int)(int8_t)(v & 0xffu) : (int)(v & 0xffu);
            }
        }
    };
    auto fetch = [&](uint32_t (&buf)[WPS], int t0) {              // this lane's step of the 32 starting at t0
        const int t = t0 + lane;
        uint32_t raw[WPS];
#pragma unroll
        for (int w = 0; w < WPS; w++) raw[w] = 0;
        if (t < a.nsteps) {
            const unsigned char* p = sp + (size_t)t * step_bytes;
#pragma unroll
            for (int j = 0; j < WPS * 4; j++)
                if (j < step_bytes) raw[j >> 2] |= (uint32_t)__ldg(p + j) << ((j & 3) * 8);
        }
#pragma unroll
        for (int w = 0; w < WPS; w++) buf[w] = raw[w];
    };
    // raw symbol words of the lane's step -> the words a step broadcasts. Called when a block of 32 steps is ENTERED, a
    // whole block after the loads were issued: evaluating the metrics right behind the loads would expose the global
    // round trip every 32 steps (measured: wy-u8 K9R2 single frame 638 -> 759 us that way).
    auto prepare = [&](const uint32_t (&raw)[WPS], uint32_t (&out)[CW]) {
        if constexpr (PRE) {
            int sym[RT ? RT : MAX_R];
            symbols_of(raw, sym);
            BmPack<POLICY>::pack(c, sym, out);
        } else {
#pragma unroll
            for (int w = 0; w < WPS; w++) out[w] = raw[w];
        }
    };
    uint32_t cur[CW], nxt[WPS];
    fetch(nxt, 0);
    for (int t0 = 0; t0 < a.nsteps; t0 += 32) {
        prepare(nxt, cur);
        fetch(nxt, t0 + 32);
        const int nhere = min(32, a.nsteps - t0);
        if (UNWRAP && WARP_PACK_SURVIVORS) {      // 16-bit halves: bring the unwrapped values down by a multiple of 256
            const int ref = __shfl_sync(0xffffffffu, lo, 0);
            const int sub = ref >= 512 ? ((ref - 256) & ~255) : 0;
            lo -= sub; hi -= sub;
        }
        for (int i = 0; i < nhere; i++) {
            uint32_t sw[CW];
#pragma unroll
            for (int w = 0; w < CW; w++) sw[w] = __shfl_sync(0xffffffffu, cur[w], i);
            int bm;
            if constexpr (PRE) bm = BmPack<POLICY>::pick(sw, cw);
            else {
                int sym[RT ? RT : MAX_R];
                symbols_of(sw, sym);
                bm = branch_metric_rt<POLICY, RT>(c, cw, sym);
            }
            int v0, v1, d0, d1;
            if constexpr (UNWRAP) acs_mod_unwrapped(lo, hi, bm, v0, v1, d0, d1);
            else acs<POLICY>(c.max_error, lo, hi, bm, v0, v1, d0, d1);     // new[2j], new[2j+1]
            const unsigned w0 = __ballot_sync(0xffffffffu, d0), w1 = __ballot_sync(0xffffffffu, d1);
            if (WARP_PACK_SURVIVORS) {      // both survivors of a butterfly travel as one u16x2 word: two shuffles per step
                const uint32_t pk = ((uint32_t)v0 & 0xffffu) | ((uint32_t)v1 << 16);
                const uint32_t pa = __shfl_sync(0xffffffffu, pk, src0), pc = __shfl_sync(0xffffffffu, pk, src1);
                lo = (int)(odd ? (pa >> 16) : (pa & 0xffffu));
                hi = (int)(odd ? (pc >> 16) : (pc & 0xffffu));
            } else {
                const int a0 = __shfl_sync(0xffffffffu, v0, src0), a1 = __shfl_sync(0xffffffffu, v1, src0);
                const int c0 = __shfl_sync(0xffffffffu, v0, src1), c1 = __shfl_sync(0xffffffffu, v1, src1);
                lo = odd ? a1 : a0;
                hi = odd ? c1 : c0;
            }
            if (lane == 0) *reinterpret_cast<uint2*>(drow + 2 * (size_t)(t0 + i)) = make_uint2(w0, w1);
            if (PolicyTraits<POLICY>::has_renorm) {
                const int m0 = __shfl_sync(0xffffffffu, lo, 0);
                if (renorm_needed<POLICY>(c, m0)) {                     // uniform over the warp
                    const int mn = __reduce_min_sync(0xffffffffu, min(lo, hi));
                    rsum += renorm_amount<POLICY>(mn);
                    lo = renorm_apply<POLICY>(lo, mn);
                    hi = renorm_apply<POLICY>(hi, mn);
                }
            }
        }
    }
    if (lane == 0 && a.err && rsum) a.err[frame] += rsum;
    if (UNWRAP) { lo &= 255; hi &= 255; }
    gm[lane] = (uint16_t)lo;
    gm[lane + 32] = (uint16_t)hi;
}

// ---- K=9: the same design with 8 states (4 butterflies) per lane --------------------------------------------
// Lane l holds old[l + 32 i], i = 0..7: butterfly j = l + 32 b (b = 0..3) finds both inputs, old[j] (slot b) and
// old[j + 128] (slot b + 4), in its own lane. New state n = r + 32 s (lane r, slot s) comes from butterfly
// (r >> 1) + 16 s = lane (r >> 1) + 16 (s & 1), butterfly slot s >> 1, output r & 1: the two survivors of a butterfly
// travel as one packed u16x2 word, so a step costs 8 shuffles. The eight ballots of a step are the generic family's
// decision row (DL_BALLOT_FRAME_MAJOR, row_words = 8): word 2b + (n & 1), bit l.
template <int POLICY, int RT>
__global__ void __launch_bounds__(128) acs_k9_warp_kernel(const GenericArgs a) {
    constexpr int NBF = 4, NS = 8;
    const int lane = threadIdx.x & 31;
    const int fidx = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (fidx >= a.count) return;
    const int frame = a.frame0 + fidx;
    const CodecParams& c = a.c;
    const int R = RT ? RT : c.R;
    constexpr int SYMB = (POLICY == P_WY_U16) ? 2 : 1;
    constexpr int WPS = RT ? (RT * SYMB + 3) / 4 : (MAX_R * SYMB + 3) / 4;
    const int step_bytes = R * SYMB;
    uint16_t* gm = a.metrics + (size_t)frame * 256;
    int m[NS];
    unsigned cw[NBF];
#pragma unroll
    for (int i = 0; i < NS; i++) m[i] = gm[lane + 32 * i];
#pragma unroll
    for (int b = 0; b < NBF; b++) cw[b] = a.cw[lane + 32 * b];
    constexpr bool UNWRAP = WARP_MOD_UNWRAPPED && POLICY == P_KA9Q_U8_MOD;
    // (the williamyang policies measured slower with the packed metrics here -- wy-u8 K9R2 single frame 638 -> 759 us in
    // the same run -- while ka9q29 / spiral29 gained 27 % / 20 %: four butterflies per lane share one symbol decode)
    constexpr bool PRE = WARP_BM_PRE && RT == 2 && !PolicyTraits<POLICY>::is_wy;
    constexpr int CW = PRE ? BmPack<POLICY>::WORDS : WPS;          // words a step broadcasts
    if (UNWRAP) {      // around state 0 (lane 0, slot 0): U = 256 + m0 + (int8)(m - m0)
        const int m0 = __shfl_sync(0xffffffffu, m[0], 0);
#pragma unroll
        for (int i = 0; i < NS; i++) m[i] = 256 + m0 + (int)(int8_t)(m[i] - m0);
    }
    const unsigned char* sp = (const unsigned char*)a.syms + (size_t)frame * a.sym_stride;
    unsigned long long rsum = 0;     // minima subtracted by this launch's renormalisations (uniform over the warp)
    uint32_t* drow = a.dec + ((size_t)frame * a.dec_rows + a.step0) * 8;
    const int srcA = lane >> 1, srcB = 16 + (lane >> 1);
    const bool odd = lane & 1;

    auto symbols_of = [&](const uint32_t* sw, int* sym) {
#pragma unroll
        for (int r = 0; r < (RT ? RT : MAX_R); r++) {
            if (r < R) {
                const int bpos = r * SYMB;
                const uint32_t v = sw[bpos >> 2] >> ((bpos & 3) * 8);
                sym[r] = (POLICY == P_WY_U16) ? (int)(int16_t)(v & 0xffffu)
                       : (POLICY == P_WY_U8) ? (int)(int8_t)(v & 0xffu) : (int)(v & 0xffu);
            }
        }
    };
    auto fetch = [&](uint32_t (&buf)[WPS], int t0) {
        const int t = t0 + lane;
        uint32_t raw[WPS];
#pragma unroll
        for (int w = 0; w < WPS; w++) raw[w] = 0;
        if (t < a.nsteps) {
            const unsigned char* p = sp + (size_t)t * step_bytes;
#pragma unroll
            for (int j = 0; j < WPS * 4; j++)
                if (j < step_bytes) raw[j >> 2] |= (uint32_t)__ldg(p + j) << ((j & 3) * 8);
        }
#pragma unroll
        for (int w = 0; w < WPS; w++) buf[w] = raw[w];
    };
    // raw symbol words of the lane's step -> the words a step broadcasts. Called when a block of 32 steps is ENTERED, a
    // whole block after the loads were issued: evaluating the metrics right behind the loads would expose the global
    // round trip every 32 steps (measured: wy-u8 K9R2 single frame 638 -> 759 us that way).
    auto prepare = [&](const uint32_t (&raw)[WPS], uint32_t (&out)[CW]) {
        if constexpr (PRE) {
            int sym[RT ? RT : MAX_R];
            symbols_of(raw, sym);
            BmPack<POLICY>::pack(c, sym, out);
        } else {
#pragma unroll
            for (int w = 0; w < WPS; w++) out[w] = raw[w];
        }
    };
    uint32_t cur[CW], nxt[WPS];
    fetch(nxt, 0);
    for (int t0 = 0; t0 < a.nsteps; t0 += 32) {
        prepare(nxt, cur);
        fetch(nxt, t0 + 32);
        const int nhere = min(32, a.nsteps - t0);
        if (UNWRAP) {      // survivors travel as 16-bit halves: bring the unwrapped values down by a multiple of 256
            const int ref = __shfl_sync(0xffffffffu, m[0], 0);
            const int sub = ref >= 512 ? ((ref - 256) & ~255) : 0;
#pragma unroll
            for (int s = 0; s < NS; s++) m[s] -= sub;
        }
        for (int i = 0; i < nhere; i++) {
            uint32_t sw[CW];
#pragma unroll
            for (int w = 0; w < CW; w++) sw[w] = __shfl_sync(0xffffffffu, cur[w], i);
            int sym[RT ? RT : MAX_R];
            if constexpr (!PRE) symbols_of(sw, sym);
            uint32_t pk[NBF], row[2 * NBF];
#pragma unroll
            for (int b = 0; b < NBF; b++) {
                int bm;
                if constexpr (PRE) bm = BmPack<POLICY>::pick(sw, cw[b]);
                else bm = branch_metric_rt<POLICY, RT>(c, cw[b], sym);
                int v0, v1, d0, d1;
                if constexpr (UNWRAP) acs_mod_unwrapped(m[b], m[b + NBF], bm, v0, v1, d0, d1);
                else acs<POLICY>(c.max_error, m[b], m[b + NBF], bm, v0, v1, d0, d1);     // new[2j], new[2j+1]
                row[2 * b] = __ballot_sync(0xffffffffu, d0);
                row[2 * b + 1] = __ballot_sync(0xffffffffu, d1);
                pk[b] = ((uint32_t)v0 & 0xffffu) | ((uint32_t)v1 << 16);
            }
#pragma unroll
            for (int s = 0; s < NS; s++) {
                const uint32_t t = __shfl_sync(0xffffffffu, pk[s >> 1], (s & 1) ? srcB : srcA);
                m[s] = (int)(odd ? (t >> 16) : (t & 0xffffu));
            }
            if (lane == 0) {
                uint4* d = reinterpret_cast<uint4*>(drow + 8 * (size_t)(t0 + i));
                d[0] = make_uint4(row[0], row[1], row[2], row[3]);
                d[1] = make_uint4(row[4], row[5], row[6], row[7]);
            }
            if (PolicyTraits<POLICY>::has_renorm) {
                const int m0 = __shfl_sync(0xffffffffu, m[0], 0);
                if (renorm_needed<POLICY>(c, m0)) {                     // uniform over the warp
                    int lm = m[0];
#pragma unroll
                    for (int s = 1; s < NS; s++) lm = min(lm, m[s]);
                    const int mn = __reduce_min_sync(0xffffffffu, lm);
                    rsum += renorm_amount<POLICY>(mn);
#pragma unroll
                    for (int s = 0; s < NS; s++) m[s] = renorm_apply<POLICY>(m[s], mn);
                }
            }
        }
    }
    if (lane == 0 && a.err && rsum) a.err[frame] += rsum;
#pragma unroll
    for (int i = 0; i < NS; i++) gm[lane + 32 * i] = (uint16_t)(UNWRAP ? (m[i] & 255) : m[i]);
}

template <int POLICY>
cudaError_t launch_policy(const GenericArgs& a, cudaStream_t st) {
    const int warps_per_block = 4;
    const int blocks = (a.count + warps_per_block - 1) / warps_per_block;
    const dim3 g(blocks), b(warps_per_block * 32);
    if (a.c.K == 9) {
        switch (a.c.R) {
        case 2: acs_k9_warp_kernel<POLICY, 2><<<g, b, 0, st>>>(a); break;
        case 4: acs_k9_warp_kernel<POLICY, 4><<<g, b, 0, st>>>(a); break;
        default: acs_k9_warp_kernel<POLICY, 0><<<g, b, 0, st>>>(a); break;
        }
        return cudaGetLastError();
    }
    switch (a.c.R) {
    case 2: acs_k7_warp_kernel<POLICY, 2><<<g, b, 0, st>>>(a); break;
    case 3: acs_k7_warp_kernel<POLICY, 3><<<g, b, 0, st>>>(a); break;
    case 4: acs_k7_warp_kernel<POLICY, 4><<<g, b, 0, st>>>(a); break;
    default: acs_k7_warp_kernel<POLICY, 0><<<g, b, 0, st>>>(a); break;
    }
    return cudaGetLastError();
}

}  // namespace

bool k7_warp_supported(const CodecParams& c) {
    return (c.K == 7 || c.K == 9) && c.R <= MAX_R && c.policy != P_KA9Q_S16_615 && c.policy != P_KA9Q_S16_224 && c.policy != P_SPIRAL_615;
}

cudaError_t launch_acs_k7_warp(const GenericArgs& a, cudaStream_t st) {
    if (a.count <= 0 || a.nsteps <= 0) return cudaSuccess;
    switch (a.c.policy) {
    case P_KA9Q_U8_MOD: return launch_policy<P_KA9Q_U8_MOD>(a, st);
    case P_SPIRAL_R2: return launch_policy<P_SPIRAL_R2>(a, st);
    case P_SPIRAL_R4: return launch_policy<P_SPIRAL_R4>(a, st);
    case P_WY_U8: return launch_policy<P_WY_U8>(a, st);
    case P_WY_U16: return launch_policy<P_WY_U16>(a, st);
    default: return cudaErrorInvalidValue;
    }
}

}  // namespace b200v
