// acs_k7.cu -- register-resident add-compare-select for K=7 (64 states): one thread per frame, metrics as
// u16x2 pairs in 32 registers updated in place (see acs_k7.cuh for the index algebra), DPX instructions
// (VIADDMNMX.U16x2 / VIMNMX.U16x2 with predicate outputs) for saturating add and compare-select, decisions
// of a step packed into one u64 and written frame-interleaved (dec[row][frame]) so that both this kernel's
// stores and the chainback kernel's loads are coalesced.
//
// Specialised at compile time on (policy, R, polynomials) so that every register index and every
// branch-metric selection is static. Takes the role of update_viterbi27_blk_sse2
// (ka9q_libfec_port/viterbi27_sse2.cpp:119-175), update_spiral27/47 (spiral/spiral27.cpp:130-386,
// spiral47.cpp) and ViterbiDecoder_SSE_u8<7,R>::update (viterbi/x86/viterbi_decoder_sse_u8.h:45-154).
//
// Bit-exactness notes
//  * saturating policies (spiral, wy-u8): registers hold true_metric + off, where `off` is a per-frame
//    offset; the reference's "subtract the minimum when new[0] crosses the threshold" becomes off = min(w),
//    and saturation at 255 becomes a clamp at 255 + off, so no per-frame divergent subtract pass is needed.
//  * ka9q modular policy: registers hold the unwrapped metric U (reference metric = U mod 256). The
//    reference decides with (int8)(m0 - m1) > 0; for K=7 the difference of the two candidates is bounded by
//    113 < 128 for every possible input (max metric <= 63 + 7t for t <= 6 because the two branches into a
//    state cost bm and 15-bm; spread <= 90 afterwards), so the modular compare equals the plain compare.
#include "kernels.h"
#include "acs_k7.cuh"
#include "dpx.cuh"
#include "policy.cuh"

namespace b200v {
namespace k7 {

template <int R_, int P0, int P1, int P2, int P3>
struct Code {
    static constexpr int R = R_;
    __host__ __device__ static constexpr int poly(int i) { return i == 0 ? P0 : i == 1 ? P1 : i == 2 ? P2 : P3; }
    __host__ __device__ static constexpr int cw(int j) {
        int c = 0;
        for (int i = 0; i < R; i++) c |= parity_of((unsigned)(2 * j) & (unsigned)poly(i)) << i;
        return c;
    }
};

template <int POLICY>
struct Arith {
    static constexpr bool modular = (POLICY == P_KA9Q_U8_MOD);
    static constexpr bool wide = (POLICY == P_WY_U16);          // u16 metrics: offset may be negative, rebase by minimum
    static constexpr int TOP = wide ? 65535 : 255;
    static constexpr int SB = wide ? 2 : 1;                     // bytes per input symbol
    __device__ __forceinline__ static int bm_cap(int C) { return wide ? 2 * C : 255; }
    // candidate = old + branch metric (saturating at the clamp for the u8-saturating policies)
    __device__ __forceinline__ static uint32_t add(uint32_t a, uint32_t w, uint32_t clamp) {
        if (modular) return a + w;                 // lanes never overflow (rebased every 64 periods)
        return __viaddmin_u16x2(a, w, clamp);
    }
};

// scalar branch metric and complement for code word CW (compile-time), symbols sym[R]
template <int POLICY, int R, int CW>
__device__ __forceinline__ void bm_pair(const int* sym, int C, int soft_hi, int soft_lo, uint32_t& bm, uint32_t& cm) {
    int v;
    if (POLICY == P_WY_U8) {
        int tot = 0;
#pragma unroll
        for (int i = 0; i < R; i++) {
            const int bt = ((CW >> i) & 1) ? soft_hi : soft_lo;
            int e = clamp_i(bt - sym[i], -128, 127);
            tot = min(tot + abs(e), 255);
        }
        v = tot;
    } else if (POLICY == P_WY_U16) {
        int tot = 0;
#pragma unroll
        for (int i = 0; i < R; i++) {
            const int bt = ((CW >> i) & 1) ? soft_hi : soft_lo;
            tot = min(tot + abs(clamp_i(bt - sym[i], -32768, 32767)), 65535);
        }
        v = tot;
    } else if (POLICY == P_KA9Q_U8_MOD || POLICY == P_SPIRAL_R2) {
        const int x0 = (CW & 1) ? (sym[0] ^ 255) : sym[0];
        const int x1 = (CW & 2) ? (sym[1] ^ 255) : sym[1];
        const int avg = (x0 + x1 + 1) >> 1;
        v = POLICY == P_KA9Q_U8_MOD ? (avg >> 4) : ((avg >> 2) & 63);
    } else {   // P_SPIRAL_R4
        int s = 0;
#pragma unroll
        for (int i = 0; i < R; i++) {
            const int x = ((CW >> i) & 1) ? (sym[i] ^ 255) : sym[i];
            s = min(s + ((x >> 2) & 63), 255);
        }
        v = (s >> 2) & 63;
    }
    bm = (uint32_t)v;
    cm = (uint32_t)max(C - v, 0);
}

template <int POLICY, class CODE, int CW, int NCW>
struct BmFill {
    __device__ __forceinline__ static void run(const int* sym, int C, int hi, int lo, uint32_t* bm, uint32_t* cm) {
        bm_pair<POLICY, CODE::R, CW>(sym, C, hi, lo, bm[CW], cm[CW]);
        BmFill<POLICY, CODE, CW + 1, NCW>::run(sym, C, hi, lo, bm, cm);
    }
};
template <int POLICY, class CODE, int NCW>
struct BmFill<POLICY, CODE, NCW, NCW> {
    __device__ __forceinline__ static void run(const int*, int, int, int, uint32_t*, uint32_t*) {}
};

struct K7Args {
    const unsigned char* syms; size_t sym_stride;
    int nsteps, step0;
    unsigned long long* dec;       // dec[row * n_frames + frame]
    uint8_t* row_phase;            // phase of the step that produced each row
    uint16_t* metrics;             // canonical [n_frames][64] (reference metric values)
    int n_frames;
    int frame0, count;             // frames [frame0, frame0 + count) of the n_frames-wide layout
    int C, thr, soft_hi, soft_lo;
    int one, neg1;                 // 1 and -1, opaque to ptxas: multipliers that keep plain adds on the FMA pipe
    int zero;                      // 0, opaque to ptxas: third operand that keeps an add on the ALU pipe (IADD3)
    int two;                       // 2, opaque to ptxas: multiplier of the decision fold (keeps it an IMAD on the FMA pipe)
    // depuncturing fused into the symbol fetch (PUNCT kernels): syms holds only the TRANSMITTED symbols of a frame;
    // pidx[k] = index of full-rate symbol k (counted from this launch's first step) in that stream, or < 0: not
    // transmitted, the decoder sees `fill` (puncture_code_helpers.h:17-55)
    const int* pidx; int fill;
    unsigned long long* err;       // [n_frames] or nullptr: += the minima this launch's renormalisations subtract
};

// ---- unrolled butterfly loops (template recursion keeps every register index and bit position static) ----
template <int POLICY, class CODE, int P, int KK>
struct PairLoop {
    __device__ __forceinline__ static void run(uint32_t (&w)[NW], const uint32_t* W, const uint32_t* Wc,
                                               uint32_t clamp, const K7Args& a, uint32_t& acc0, uint32_t& acc1) {
        constexpr int XB = NB - 2 - P;                         // register-index bit that holds the state MSB
        constexpr int iA = expand_bit(KK, XB), iB = iA | (1 << XB);
        constexpr int c = CODE::cw(logical_of(P, iA << 1));
        const uint32_t A = w[iA], B = w[iB];
        if (Arith<POLICY>::modular) {      // decision = (lo path > hi path), tie -> 0 (viterbi27_sse2.cpp:155-158)
#ifndef K7_ALU_ADDS
#define K7_ALU_ADDS 32
#endif
            // pipe balance: ptxas emits two-input adds as VIADD / IMAD.IADD, which ncu attributes to the FMA-heavy
            // pipe (82 % busy, the binding pipe of this kernel: profiles/r01b_k7_ka9q27.txt); the lo-path adds are
            // forced onto the ALU pipe as IADD3 (K7_ALU_ADDS of the 64 adds of a step; measured 0 / 16 / 32:
            // 16.75 / 16.07 / 15.88 ms per 1 Mi frames). Keep the statement order: ptxas spills predicates otherwise.
            const uint32_t m0 = (KK < K7_ALU_ADDS) ? add_alu(A, W[c], (uint32_t)a.zero) : A + W[c];
            const uint32_t m1 = B + Wc[c];
            const uint32_t m2 = (KK + 16 < K7_ALU_ADDS) ? add_alu(A, Wc[c], (uint32_t)a.zero) : A + Wc[c];
            const uint32_t m3 = B + W[c];
            w[iA] = min_acc<KK, 16 + KK, true>(m0, m1, acc0);
            w[iB] = min_acc<KK, 16 + KK, true>(m2, m3, acc1);
        } else {
            // Saturating policies. Only the hi-path candidates need the clamp (the survivor is min(lo, hi) <= hi
            // <= clamp, and a doubly saturated butterfly still decides "hi" because hi == clamp); the lo-path adds
            // are plain adds on the FMA pipe. decision = (min == hi path) = (lo - hi >= 0), tie -> 1
            // (spiral27.cpp:178-181): bit 15 of each lane of lo + 0x8000 - hi (|lo - hi| < 2^15), also on the FMA
            // pipe. No predicates: ptxas spilled them in the predicated form of this loop.
            const uint32_t m0 = add_fma(A, (uint32_t)a.one, W[c]), m1 = __viaddmin_u16x2(B, Wc[c], clamp);
            const uint32_t m2 = add_fma(A, (uint32_t)a.one, Wc[c]), m3 = __viaddmin_u16x2(B, W[c], clamp);
            const uint32_t x0 = add_fma(m1, (uint32_t)a.neg1, add_fma(m0, (uint32_t)a.one, 0x80008000u));
            const uint32_t x1 = add_fma(m3, (uint32_t)a.neg1, add_fma(m2, (uint32_t)a.one, 0x80008000u));
            w[iA] = __vminu2(m1, m0);
            w[iB] = __vminu2(m3, m2);
            // four decisions (A lane 0, A lane 1, B lane 0, B lane 1) as 0x00 / 0xff bytes: PRMT replicates the sign
            // bit of bytes 1 and 3 of x0 and x1; one LOP3 masks bit KK & 7 of each byte into the word of group KK
            // (acs_k7.cuh: bytewise packing). Two ALU instructions per four decisions.
            const uint32_t r = sign_bytes(x0, x1);
            uint32_t& acc = (KK < 8) ? acc0 : acc1;
            acc |= r & (0x01010101u << (KK & 7));
        }
        PairLoop<POLICY, CODE, P, KK + 1>::run(w, W, Wc, clamp, a, acc0, acc1);
    }
};
template <int POLICY, class CODE, int P>
struct PairLoop<POLICY, CODE, P, 16> {
    __device__ __forceinline__ static void run(uint32_t (&)[NW], const uint32_t*, const uint32_t*, uint32_t,
                                               const K7Args&, uint32_t&, uint32_t&) {}
};

// ---- mask-decision form of the same butterflies (dpx.cuh, "mask decisions") ------------------------------------------
// Per register of two new states: 1 add (lo-path candidate t), 1 VIADDMNMX (n = min(hi + metric, t)), 1 HSET2 (lane mask of
// the decision), 1 IMAD (fold): 4 instructions where the predicate form needs 5 (2 adds, VIMNMX, 2 predicated ORs), and
// no predicate registers. KK runs 15 ... 0 so that pair KK lands at bit KK / 16 + KK of the folded word (the bitwise
// packing of acs_k7.cuh). Window note: the unwrapped lanes of the modular policy stay below 512 + 128 + 96 * 15 + 15
// (rebased every 16 periods), far inside [0, 0x7c00].
#ifndef K7_MASK_DECISIONS
#define K7_MASK_DECISIONS 1
#endif
template <int POLICY, class CODE, int P, int KK>
struct PairLoopM {
    __device__ __forceinline__ static void run(uint32_t (&w)[NW], const uint32_t* W, const uint32_t* Wc,
                                               const K7Args& a, uint32_t& h0, uint32_t& h1) {
        constexpr int XB = NB - 2 - P;
        constexpr int iA = expand_bit(KK, XB), iB = iA | (1 << XB);
        constexpr int c = CODE::cw(logical_of(P, iA << 1));
        const uint32_t A = w[iA], B = w[iB];
        // decision = (lo path > hi path), tie -> 0 (viterbi27_sse2.cpp:155-158): the survivor differs from the lo path
        // (measured on the 1 Mi-frame batch, update kernel: predicate form 14.05 ms; this form 13.09 ms with plain adds
        // -- ptxas emits them as IMAD.IADD on the FMA pipe, which balances the two ALU instructions; forcing 8 or 16 of
        // the adds or folds onto the ALU pipe costs 13.8 / 14.3 ms, __launch_bounds__(128, 7) spills: 13.7 ms)
        const uint32_t t0 = A + W[c];
        const uint32_t n0 = __viaddmin_u16x2(B, Wc[c], t0);
        h0 = horner_mask(h0, (uint32_t)a.two, ne_mask_u16x2(n0, t0));
        const uint32_t t1 = A + Wc[c];
        const uint32_t n1 = __viaddmin_u16x2(B, W[c], t1);
        h1 = horner_mask(h1, (uint32_t)a.two, ne_mask_u16x2(n1, t1));
        w[iA] = n0;
        w[iB] = n1;
        PairLoopM<POLICY, CODE, P, KK - 1>::run(w, W, Wc, a, h0, h1);
    }
};
template <int POLICY, class CODE, int P>
struct PairLoopM<POLICY, CODE, P, -1> {
    __device__ __forceinline__ static void run(uint32_t (&)[NW], const uint32_t*, const uint32_t*, const K7Args&,
                                               uint32_t&, uint32_t&) {}
};
// last phase: register i = (old[i], old[i+32]) -> (new[2i], new[2i+1]); BC[c] = (bm, cm), CB[c] = (cm, bm)
template <int POLICY, class CODE, int I>
struct InRegLoopM {
    __device__ __forceinline__ static void run(uint32_t (&w)[NW], const uint32_t* BC, const uint32_t* CB,
                                               const K7Args& a, uint32_t& h0, uint32_t& h1) {
        constexpr int c = CODE::cw(I);
        const uint32_t x = w[I];
        const uint32_t t = __byte_perm(x, x, 0x1010) + BC[c];                          // (lo + bm, lo + cm)
        const uint32_t n = __viaddmin_u16x2(__byte_perm(x, x, 0x3232), CB[c], t);      // min with (hi + cm, hi + bm)
        uint32_t& h = (I < 16) ? h0 : h1;
        h = horner_mask(h, (uint32_t)a.two, ne_mask_u16x2(n, t));
        w[I] = n;
        InRegLoopM<POLICY, CODE, I - 1>::run(w, BC, CB, a, h0, h1);
    }
};
template <int POLICY, class CODE>
struct InRegLoopM<POLICY, CODE, -1> {
    __device__ __forceinline__ static void run(uint32_t (&)[NW], const uint32_t*, const uint32_t*, const K7Args&,
                                               uint32_t&, uint32_t&) {}
};

// last phase: register i = (old[i], old[i+32]) -> (new[2i], new[2i+1])
template <int POLICY, class CODE, int I>
struct InRegLoop {
    __device__ __forceinline__ static void run(uint32_t (&w)[NW], const uint32_t* bm, const uint32_t* cm,
                                               uint32_t clamp, uint32_t& acc0, uint32_t& acc1, uint32_t xe = 0) {
        constexpr int c = CODE::cw(I);
        const uint32_t x = w[I];
        const uint32_t ma = Arith<POLICY>::add(x, bm[c] * 0x10001u, clamp);   // (m0, m3)
        const uint32_t mb = Arith<POLICY>::add(x, cm[c] * 0x10001u, clamp);   // (m2, m1)
        const uint32_t lo_paths = __byte_perm(ma, mb, 0x5410);               // (m0, m2)
        const uint32_t hi_paths = __byte_perm(mb, ma, 0x7632);               // (m1, m3)
        uint32_t& acc = (I < 16) ? acc0 : acc1;
        if (Arith<POLICY>::modular) w[I] = min_acc<(I & 15), 16 + (I & 15), true>(lo_paths, hi_paths, acc);
        else {      // same arithmetic decision as PairLoop; registers (I, I + 1) share one PRMT + LOP3 (bytewise packing)
            w[I] = __vminu2(hi_paths, lo_paths);
            const uint32_t x = lo_paths + 0x80008000u - hi_paths;
            if (I & 1) acc |= sign_bytes(xe, x) & (0x01010101u << ((I >> 1) & 7));
            else xe = x;
        }
        InRegLoop<POLICY, CODE, I + 1>::run(w, bm, cm, clamp, acc0, acc1, xe);
    }
};
template <int POLICY, class CODE>
struct InRegLoop<POLICY, CODE, NW> {
    __device__ __forceinline__ static void run(uint32_t (&)[NW], const uint32_t*, const uint32_t*, uint32_t,
                                               uint32_t&, uint32_t&, uint32_t) {}
};

// ---- one trellis step in phase P -------------------------------------------------------------------
template <int POLICY, class CODE, int P>
__device__ __forceinline__ void step(uint32_t (&w)[NW], const int* sym, const K7Args& a, uint32_t& off,
                                     unsigned long long* dec_out, uint32_t& rsum) {
    constexpr int R = CODE::R, NCW = 1 << R;
    uint32_t bm[NCW], cm[NCW];
    BmFill<POLICY, CODE, 0, NCW>::run(sym, a.C, a.soft_hi, a.soft_lo, bm, cm);
    // registers hold true + off (off is a wrapping 32-bit integer, negative for the u16 policy); the clamp is exact
    // whenever it can bind and otherwise sits above every reachable lane value
    const uint32_t clamp = (uint32_t)min(Arith<POLICY>::TOP + (int)off, 65535 - Arith<POLICY>::bm_cap(a.C)) * 0x10001u;
    uint32_t acc0 = 0, acc1 = 0;

    if constexpr (P < NB - 1) {
        constexpr int DELTA = CODE::cw(1 << P);            // lane 1 holds butterfly j | (1 << P)
        uint32_t W[NCW], Wc[NCW];
        // ka9q / spiral: the complement is exactly C - bm (bm <= C), so the packed complement is one packed subtract
        constexpr bool EXACT_COMPLEMENT = (POLICY == P_KA9Q_U8_MOD || POLICY == P_SPIRAL_R2 || POLICY == P_SPIRAL_R4);
#pragma unroll
        for (int c = 0; c < NCW; c++) {
            W[c] = bm[c] + bm[c ^ DELTA] * 65536u;
            Wc[c] = EXACT_COMPLEMENT ? (uint32_t)a.C * 0x10001u - W[c] : cm[c] + cm[c ^ DELTA] * 65536u;
        }
        if constexpr (K7_MASK_DECISIONS && Arith<POLICY>::modular) {
            PairLoopM<POLICY, CODE, P, 15>::run(w, W, Wc, a, acc0, acc1);
            acc0 = horner_decode(acc0); acc1 = horner_decode(acc1);
        } else {
            PairLoop<POLICY, CODE, P, 0>::run(w, W, Wc, clamp, a, acc0, acc1);
        }
    } else if constexpr (K7_MASK_DECISIONS && Arith<POLICY>::modular) {
        uint32_t BC[NCW], CB[NCW];
#pragma unroll
        for (int c = 0; c < NCW; c++) { BC[c] = pack2(bm[c], cm[c]); CB[c] = pack2(cm[c], bm[c]); }
        InRegLoopM<POLICY, CODE, NW - 1>::run(w, BC, CB, a, acc0, acc1);
        acc0 = horner_decode(acc0); acc1 = horner_decode(acc1);
    } else {
        // in-register butterflies: register i = (old[i], old[i+32]) -> (new[2i], new[2i+1])
        InRegLoop<POLICY, CODE, 0>::run(w, bm, cm, clamp, acc0, acc1);
    }
    *dec_out = (unsigned long long)acc0 | ((unsigned long long)acc1 << 32);

    // renormalisation (saturating policies): new[0] is always register 0, lane 0
    if (!Arith<POLICY>::modular) {
        uint32_t mn = __vimin3_u16x2(w[0], w[1], w[2]);
#pragma unroll
        for (int i = 3; i + 1 < NW; i += 2) mn = __vimin3_u16x2(mn, w[i], w[i + 1]);
        mn = __vminu2(mn, w[NW - 1]);
        const uint32_t m = min(mn & 0xffffu, mn >> 16);
        const int true0 = (int)(w[0] & 0xffffu) - (int)off;
        const bool need = (POLICY == P_WY_U8 || POLICY == P_WY_U16) ? (true0 >= a.thr) : (true0 > a.thr);
        if (need) { rsum += m - off; off = m; }      // the reference subtracts the true minimum m - off
        if (Arith<POLICY>::wide && m > 4096u) {      // keep the u16 window at the bottom of the lanes
            const uint32_t sub = m * 0x10001u;
#pragma unroll
            for (int i = 0; i < NW; i++) w[i] -= sub;
            off -= m;
        }
    }
}

// ---- symbol fetch: one period (6 steps * R bytes) per frame into registers ---------------------------
template <int R, int SB> struct PeriodBuf {
    static constexpr int BYTES = NB * R * SB;
    static constexpr int WORDS = (BYTES + 3) / 4;
    uint32_t w[WORDS];
    __device__ __forceinline__ void load(const unsigned char* p) {
        if ((BYTES % 4 == 0) && ((reinterpret_cast<uintptr_t>(p) & 3) == 0)) {
#pragma unroll
            for (int i = 0; i < WORDS; i++) w[i] = __ldg(reinterpret_cast<const uint32_t*>(p) + i);
        } else {
#pragma unroll
            for (int i = 0; i < WORDS; i++) {
                uint32_t v = 0;
#pragma unroll
                for (int b = 0; b < 4; b++)
                    if (i * 4 + b < BYTES) v |= (uint32_t)__ldg(p + i * 4 + b) << (8 * b);
                w[i] = v;
            }
        }
    }
    // the same NSYM <= NB * R symbols gathered through the puncturing table (one index load, uniform over the warp, and
    // one symbol load each)
    __device__ __forceinline__ void load_punctured(const unsigned char* in, const int* idx, uint32_t fill, int nsym = NB * R) {
#pragma unroll
        for (int i = 0; i < WORDS; i++) w[i] = 0;
#pragma unroll
        for (int k = 0; k < NB * R; k++) {
            if (k < nsym) {
                const int ix = __ldg(idx + k);
                uint32_t v = fill;
                if (ix >= 0) v = SB == 2 ? (uint32_t)__ldg(reinterpret_cast<const unsigned short*>(in) + ix) : (uint32_t)__ldg(in + ix);
                w[(k * SB) >> 2] |= v << (((k * SB) & 3) * 8);
            }
        }
    }
    template <int POLICY>
    __device__ __forceinline__ void symbols(int step, int* sym) const {
#pragma unroll
        for (int i = 0; i < R; i++) {
            const int b = (step * R + i) * SB;
            const uint32_t v = (w[b >> 2] >> ((b & 3) * 8)) & (SB == 2 ? 0xffffu : 0xffu);
            sym[i] = (SB == 2) ? (int)(int16_t)v : (POLICY == P_WY_U8) ? (int)(int8_t)v : (int)v;
        }
    }
};

template <int POLICY, class CODE, int P>
__device__ __forceinline__ void run_step(uint32_t (&w)[NW], const PeriodBuf<CODE::R, Arith<POLICY>::SB>& buf, const K7Args& a,
                                         uint32_t& off, unsigned long long* dec_row, uint32_t& rsum) {
    int sym[CODE::R];
    buf.template symbols<POLICY>(P, sym);
    step<POLICY, CODE, P>(w, sym, a, off, dec_row, rsum);
}

// (__launch_bounds__(128, 6) for the saturating policies -- 80 registers, 8-16 bytes of spills -- measured no faster:
// spiral27 21.6 ms either way, spiral47 23.0 -> 23.7 ms. Round 2: the group kernels' decision collection for these
// policies -- VIMNMX with predicate outputs + predicated FADDs -- takes 127 registers here (4 resident blocks): 23.9 ms.)
template <int POLICY, class CODE, bool PUNCT>
__global__ void __launch_bounds__(128) acs_k7_kernel(const K7Args a) {
    constexpr int R = CODE::R;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int f = a.frame0 + idx;
    if (blockIdx.x == 0) {   // phase of every decision row this launch writes (read by chainback)
        for (int t = threadIdx.x; t < a.nsteps; t += blockDim.x) a.row_phase[a.step0 + t] = (uint8_t)(t % NB);
    }
    if (idx >= a.count) return;

    // ---- load canonical metrics into the phase-0 layout: register i = states (2i, 2i+1) ----
    uint32_t w[NW];
    uint32_t off = 0;
    uint32_t rsum = 0;          // minima subtracted in the current stretch of periods (flushed into rtot every 16 periods)
    unsigned long long rtot = 0;
    {
        const uint32_t* g = reinterpret_cast<const uint32_t*>(a.metrics + (size_t)f * S);
#pragma unroll
        for (int i = 0; i < NW; i++) w[i] = g[i];
        if (Arith<POLICY>::modular) {
            // unwrap the reference's mod-256 metrics around state 0 (spread < 128): U = 256 + m0 + (int8)(m - m0)
            const int m0 = (int)(w[0] & 0xffu);
#pragma unroll
            for (int i = 0; i < NW; i++) {
                const int lo = 256 + m0 + (int)(int8_t)((w[i] & 0xffu) - m0);
                const int hi = 256 + m0 + (int)(int8_t)(((w[i] >> 16) & 0xffu) - m0);
                w[i] = pack2((uint32_t)lo, (uint32_t)hi);
            }
        }
    }

    if (Arith<POLICY>::wide) {     // u16 metrics may start anywhere in [0, 65535]: bring the window down first
        uint32_t mn = w[0];
#pragma unroll
        for (int i = 1; i < NW; i++) mn = __vminu2(mn, w[i]);
        const uint32_t m = min(mn & 0xffffu, mn >> 16);
#pragma unroll
        for (int i = 0; i < NW; i++) w[i] -= m * 0x10001u;
        off = 0u - m;
    }

    const unsigned char* sp = a.syms + (size_t)f * a.sym_stride;
    unsigned long long* drow = a.dec + (size_t)a.step0 * a.n_frames + f;
    const size_t dstride = (size_t)a.n_frames;

    const int periods = a.nsteps / NB;
    constexpr int SB = Arith<POLICY>::SB;
    using PBuf = PeriodBuf<R, SB>;
    PBuf cur, nxt;
    const uint32_t pfill = (uint32_t)a.fill & (SB == 2 ? 0xffffu : 0xffu);
    if (periods > 0) {
        if (PUNCT) cur.load_punctured(sp, a.pidx, pfill);
        else cur.load(sp);
    }
    // Two nested loops: the rebase (every 16 periods) sits BETWEEN runs of the inner loop, not behind an `if` inside it.
    // A conditional block in the hot loop ends in a join where ptxas brings the 32 renamed metric registers back to
    // common places on every iteration (27 IMAD.MOV per period on the FMA-heavy pipe, the binding one: round 2, ncu
    // source page of the ka9q27 kernel).
#ifndef K7_NESTED_LOOP
#define K7_NESTED_LOOP 1
#endif
    int pd = 0;
#pragma unroll 1
    while (pd < periods) {
        const int pend = K7_NESTED_LOOP ? min(periods, (pd | 15) + 1) : pd + 1;
#ifndef K7_PERIOD_UNROLL
#define K7_PERIOD_UNROLL 1
#endif
        constexpr int PERIOD_UNROLL = K7_PERIOD_UNROLL;
#pragma unroll PERIOD_UNROLL
        for (; pd < pend; pd++) {
            if (pd + 1 < periods) {
                if (PUNCT) nxt.load_punctured(sp, a.pidx + (size_t)(pd + 1) * NB * R, pfill);
                else nxt.load(sp + (size_t)(pd + 1) * NB * R * SB);
            }
            unsigned long long d0, d1, d2, d3, d4, d5;
            run_step<POLICY, CODE, 0>(w, cur, a, off, &d0, rsum);
            run_step<POLICY, CODE, 1>(w, cur, a, off, &d1, rsum);
            run_step<POLICY, CODE, 2>(w, cur, a, off, &d2, rsum);
            run_step<POLICY, CODE, 3>(w, cur, a, off, &d3, rsum);
            run_step<POLICY, CODE, 4>(w, cur, a, off, &d4, rsum);
            run_step<POLICY, CODE, 5>(w, cur, a, off, &d5, rsum);
            drow[0] = d0; drow[dstride] = d1; drow[2 * dstride] = d2;
            drow[3 * dstride] = d3; drow[4 * dstride] = d4; drow[5 * dstride] = d5;
            drow += 6 * dstride;
            cur = nxt;
        }
        // keep the 16-bit lanes from overflowing: after every 16th period rebase (uniform across the warp)
        if ((pd & 15) == 0) {
            rtot += rsum; rsum = 0;
            if (Arith<POLICY>::modular) {
                // subtract a multiple of 256 (keeps U mod 256): floor(min/256)*256 - 256
                uint32_t mn = w[0];
#pragma unroll
                for (int i = 1; i < NW; i++) mn = __vminu2(mn, w[i]);
                const uint32_t m = min(mn & 0xffffu, mn >> 16);
                const uint32_t sub = (m & ~255u) >= 256u ? (m & ~255u) - 256u : 0u;
#pragma unroll
                for (int i = 0; i < NW; i++) w[i] -= sub * 0x10001u;
            } else if (!Arith<POLICY>::wide) {
#pragma unroll
                for (int i = 0; i < NW; i++) w[i] -= off * 0x10001u;    // every lane >= off: no borrow
                off = 0;
            }
        }
    }

    // ---- tail: fewer than 6 steps left ----
    const int rem = a.nsteps - periods * NB;
    if (rem > 0) {
        PBuf tb;
        const unsigned char* tp = sp + (size_t)periods * NB * R * SB;
        if (PUNCT) tb.load_punctured(sp, a.pidx + (size_t)periods * NB * R, pfill, rem * R);
        else {
#pragma unroll
            for (int i = 0; i < PBuf::WORDS; i++) {
                uint32_t v = 0;
#pragma unroll
                for (int b = 0; b < 4; b++)
                    if (i * 4 + b < rem * R * SB) v |= (uint32_t)__ldg(tp + i * 4 + b) << (8 * b);
                tb.w[i] = v;
            }
        }
        unsigned long long d;
        if (rem > 0) { run_step<POLICY, CODE, 0>(w, tb, a, off, &d, rsum); drow[0] = d; }
        if (rem > 1) { run_step<POLICY, CODE, 1>(w, tb, a, off, &d, rsum); drow[dstride] = d; }
        if (rem > 2) { run_step<POLICY, CODE, 2>(w, tb, a, off, &d, rsum); drow[2 * dstride] = d; }
        if (rem > 3) { run_step<POLICY, CODE, 3>(w, tb, a, off, &d, rsum); drow[3 * dstride] = d; }
        if (rem > 4) { run_step<POLICY, CODE, 4>(w, tb, a, off, &d, rsum); drow[4 * dstride] = d; }
    }

    if (!Arith<POLICY>::modular && a.err) {
        rtot += rsum;
        if (rtot) a.err[f] += rtot;
    }

    // ---- store canonical metrics: logical state n sits at physical_of(rem, n) ----
    uint32_t tmp[NW];
#pragma unroll
    for (int i = 0; i < NW; i++) tmp[i] = w[i];
    uint16_t* gm = a.metrics + (size_t)f * S;
    for (int n = 0; n < S; n++) {
        const int q = physical_of(rem, n);
        const uint32_t v = (tmp[q >> 1] >> ((q & 1) * 16)) & 0xffffu;
        gm[n] = (uint16_t)(Arith<POLICY>::modular ? (v & 0xffu) : (v - off));
    }
}


template <int POLICY, class CODE>
static cudaError_t launch_one(const K7Args& a, cudaStream_t st) {
    const int threads = 128;
    const int blocks = (a.count + threads - 1) / threads;
    if (a.pidx) acs_k7_kernel<POLICY, CODE, true><<<blocks, threads, 0, st>>>(a);
    else acs_k7_kernel<POLICY, CODE, false><<<blocks, threads, 0, st>>>(a);
    return cudaGetLastError();
}

using CodeK7R2 = Code<2, 0x6d, 0x4f, 0, 0>;       // src/main.cpp:367 (Voyager)
using CodeK7R4 = Code<4, 121, 117, 91, 111>;      // src/main.cpp:377
using CodeK7R3 = Code<3, 91, 117, 121, 0>;        // williamyang_viterbi_lib/examples/helpers/common_codes.h:26 (LTE)
using CodeDAB = Code<4, 109, 79, 83, 109>;        // common_codes.h:27 (DAB Radio; run_punctured_decoder.cpp's code)

// index of the instantiation matching (policy, R, poly), or -1
static int find_instance(const CodecParams& c) {
    if (c.K != 7) return -1;
    if (c.R == 2 && c.poly[0] == 0x6d && c.poly[1] == 0x4f) {
        if (c.policy == P_KA9Q_U8_MOD) return 0;
        if (c.policy == P_SPIRAL_R2) return 1;
        if (c.policy == P_WY_U8) return 2;
        if (c.policy == P_WY_U16) return 6;
    }
    if (c.R == 4 && c.poly[0] == 121 && c.poly[1] == 117 && c.poly[2] == 91 && c.poly[3] == 111) {
        if (c.policy == P_SPIRAL_R4) return 3;
        if (c.policy == P_WY_U8) return 4;
        if (c.policy == P_WY_U16) return 7;
    }
    if (c.R == 4 && c.poly[0] == 109 && c.poly[1] == 79 && c.poly[2] == 83 && c.poly[3] == 109) {
        if (c.policy == P_WY_U8) return 9;
        if (c.policy == P_WY_U16) return 10;
    }
    if (c.R == 3 && c.poly[0] == 91 && c.poly[1] == 117 && c.poly[2] == 121 && c.policy == P_WY_U8) return 5;
    if (c.R == 3 && c.poly[0] == 91 && c.poly[1] == 117 && c.poly[2] == 121 && c.policy == P_WY_U16) return 8;
    return -1;
}

}  // namespace k7

bool k7_supported(const CodecParams& c) {
    if (k7::find_instance(c) < 0) return false;
    // wy-u8: the offset representation needs headroom above the saturation point
    if (c.policy == P_WY_U8 && (c.max_error > 255 || c.renorm_threshold > 255 || c.init_other > 255)) return false;
    if (c.policy == P_WY_U16) {     // the lane window: rebase threshold + metric spread + two capped branch metrics
        if (c.max_error <= 0 || c.init_start < 0 || c.init_other < 0 || c.renorm_threshold < 0 ||
            c.init_start > 65535 || c.init_other > 65535 || c.renorm_threshold > 65535) return false;
        const long long cap = 2LL * c.max_error;
        long long spread = (long long)(c.K - 1) * cap;
        const long long init_spread = (long long)c.init_other - c.init_start;
        if (init_spread > spread) spread = init_spread;
        if (-init_spread > spread) spread = -init_spread;
        if (4096 + spread + 2 * cap > 32000) return false;   // the arithmetic decision needs |lo - hi| < 2^15
    }
    return true;
}

cudaError_t launch_acs_k7(const CodecParams& c, const void* syms, size_t sym_stride, int nsteps, int step0,
                          unsigned long long* dec, uint8_t* row_phase, uint16_t* metrics, int n_frames,
                          int frame0, int count, cudaStream_t st, const int* pidx, int fill, unsigned long long* err) {
    if (count <= 0 || nsteps <= 0) return cudaSuccess;
    k7::K7Args a;
    a.pidx = pidx; a.fill = fill; a.err = err;
    a.syms = (const unsigned char*)syms; a.sym_stride = sym_stride; a.nsteps = nsteps; a.step0 = step0;
    a.dec = dec; a.row_phase = row_phase; a.metrics = metrics; a.n_frames = n_frames;
    a.frame0 = frame0; a.count = count;
    a.C = c.max_error; a.thr = c.renorm_threshold; a.soft_hi = c.soft_hi; a.soft_lo = c.soft_lo;
    a.one = 1; a.neg1 = -1; a.zero = 0; a.two = 2;
    switch (k7::find_instance(c)) {
    case 0: return k7::launch_one<P_KA9Q_U8_MOD, k7::CodeK7R2>(a, st);
    case 1: return k7::launch_one<P_SPIRAL_R2, k7::CodeK7R2>(a, st);
    case 2: return k7::launch_one<P_WY_U8, k7::CodeK7R2>(a, st);
    case 3: return k7::launch_one<P_SPIRAL_R4, k7::CodeK7R4>(a, st);
    case 4: return k7::launch_one<P_WY_U8, k7::CodeK7R4>(a, st);
    case 5: return k7::launch_one<P_WY_U8, k7::CodeK7R3>(a, st);
    case 6: return k7::launch_one<P_WY_U16, k7::CodeK7R2>(a, st);
    case 7: return k7::launch_one<P_WY_U16, k7::CodeK7R4>(a, st);
    case 8: return k7::launch_one<P_WY_U16, k7::CodeK7R3>(a, st);
    case 9: return k7::launch_one<P_WY_U8, k7::CodeDAB>(a, st);
    case 10: return k7::launch_one<P_WY_U16, k7::CodeDAB>(a, st);
    default: return cudaErrorInvalidValue;
    }
}

}  // namespace b200v
