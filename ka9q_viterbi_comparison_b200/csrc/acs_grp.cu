// acs_grp.cu -- "group" add-compare-select kernels for K = 9 and K = 15: TPF = 2^(K-7) threads per frame, each
// holding 64 path metrics in 32 registers of u16x2 (index algebra: acs_grp.cuh). Five trellis steps run
// entirely in registers (in-place rotating layout, static register indices); every fifth step the threads of
// a frame exchange metrics through shared memory (one transposition, swizzled, conflict-free for K=15).
//   K=15: 256 threads = one frame per block;  K=9: 4 threads per frame, 32 frames per 128-thread block.
// Per pair of registers (2 butterflies, 4 new states): 1 LDS.64 of the packed branch metrics
// {(bm[c], bm[c^D]), (cm[c], cm[c^D])} from a per-step table in shared memory (indexed by the static code word
// of the register pair XOR the thread's own code-word contribution), 4 VIADDMNMX.U16x2 (saturating add as
// add+min), 2 VIMNMX.U16x2 with predicate outputs (compare-select + both decisions), 4 predicated ORs.
//
// Takes the role of update_viterbi615_blk_sse2 (ka9q_libfec_port/viterbi615_sse2.cpp:104-191),
// update_spiral29/49/615 (spiral/spiral29.cpp, spiral49.cpp, spiral615.cpp:130-415) and
// ViterbiDecoder_SSE_u8<K,R>::update (viterbi/x86/viterbi_decoder_sse_u8.h:45-154) for a batch of frames.
//
// Bit-exactness: lanes hold w = tb - base, tb = the reference's metric (s16 metrics biased by +32768 so that
// signed min/saturation become unsigned), base = a per-frame integer every thread of the frame tracks.
//  * saturation at TOP (255 / 65535) is the clamp operand TOP - base of VIADDMNMX (exact whenever it can bind);
//  * "if new[0] crosses the threshold subtract the minimum of all states" (viterbi615_sse2.cpp:160-183,
//    spiral615.cpp:31-40, sse_u8.h:127-154) becomes base = -min(w): the frame-wide minimum is reduced after
//    EVERY step (REDUX + one barrier for K=15, two shuffles for K=9), so the trigger is evaluated at exactly
//    the step the reference evaluates it;
//  * lanes are kept inside 16 bits by rebasing (w -= min, base += min) whenever min(w) > 16384.
#include "kernels.h"
#include "acs_grp.cuh"
#include "dpx.cuh"
#include "policy.cuh"

#include <type_traits>

// This file is compiled twice: as itself (the kernels that read full-rate symbols, and the family's host functions) and
// through acs_grp_punct.cu with GRP_PUNCT_TU = 1 (the kernels whose symbol fetch gathers through a puncturing table,
// launch_acs_grp_punct only). PUNCT is a template parameter of everything that differs, so the two sets of
// instantiations have different names.
#ifndef GRP_PUNCT_TU
#define GRP_PUNCT_TU 0
#endif

namespace b200v {
namespace grp {

template <int K_, int R_, int P0, int P1, int P2, int P3, int P4, int P5>
struct Code {
    static constexpr int K = K_, R = R_;
    __host__ __device__ static constexpr unsigned poly(int i) {
        return (unsigned)(i == 0 ? P0 : i == 1 ? P1 : i == 2 ? P2 : i == 3 ? P3 : i == 4 ? P4 : P5);
    }
    // code word of the butterfly whose doubled index 2j is x: bit i = parity(x & poly_i)
    __host__ __device__ static constexpr unsigned cw2(unsigned x) {
        unsigned c = 0;
        for (int i = 0; i < R; i++) c |= (unsigned)k7::parity_of(x & poly(i)) << i;
        return c;
    }
};

template <int POLICY, class CODE>
struct Cfg {
    static constexpr int K = CODE::K, R = CODE::R, NB = K - 1, NXB = NB - NL, S = 1 << NB, NCW = 1 << R;
    static constexpr int TPF = 1 << (NXB - 1);                  // threads per frame
    static constexpr int FPB = TPF >= 32 ? 1 : 128 / TPF;       // frames per block
    static constexpr int THREADS = TPF * FPB;
    static constexpr int MET_STRIDE = S * 2 + (TPF < 32 ? 16 : 0);   // bytes of exchange buffer per frame
    static constexpr bool modular = (POLICY == P_KA9Q_U8_MOD);
    static constexpr int ESHIFT = 3;                            // table entry: {W, Wc}
    static constexpr int SUBTAB = NCW << ESHIFT;                // one step's table
    static constexpr int TAB_BYTES = 2 * NL * SUBTAB;           // two groups (double buffer) x five steps
    // Several frames per warp (K=9) read their own tables with one LDS: frame f's tables must sit in other banks
    // than frame f+1's. Step tables narrower than 128 B: pad the per-frame stride so that consecutive frames are
    // one step-table apart modulo the 32 banks. 128-byte step tables fill all banks: frames permute their entries
    // instead (frame_perm below).
    static constexpr int tab_stride() {
        if (FPB == 1 || SUBTAB >= 128) return TAB_BYTES;
        int st = TAB_BYTES;
        while ((st / 4) % 32 != (SUBTAB / 4) % 32) st += SUBTAB;
        return st;
    }
    static constexpr int TAB_STRIDE = tab_stride();
    static constexpr bool PERMUTE = (FPB > 1 && SUBTAB >= 128);
    static constexpr int RED_WORDS = 2 * 16;                    // block reduction scratch (TPF >= 32)
    static constexpr int SMEM = FPB * TAB_STRIDE + FPB * MET_STRIDE + RED_WORDS * 4;
    static constexpr bool s16 = (POLICY == P_KA9Q_S16_615);
    static constexpr bool strict = PolicyTraits<POLICY>::renorm_strict;
    static constexpr int TOP = (s16 || POLICY == P_WY_U16) ? 65535 : 255;
    static constexpr int SB = (POLICY == P_WY_U16) ? 2 : 1;     // bytes per input symbol
    // rebase the lanes when their minimum exceeds TH: window = TH + metric spread + two branch metrics < 2^16
    // mask decisions (dpx.cuh): compare-select as one VIADDMNMX, decision read back with HSET2 -- needs every lane inside
    // [0, 0x7c00]. The modular policy always uses them (ka9q29 update 15.7 -> 14.3 ms per 256 Ki frames together with the
    // unguarded whole groups of the main loop). For the SATURATING policies the form exists (GRP_MASK_SAT=1, bit-exact:
    // 8-bit policies: lanes <= TH + 255 + two branch metrics; ka9q s16: TH + (skip + 1) branch metrics + spread
    // (K-1) * 1530 + one candidate, fits with TH = 512 and at most three deferred steps; wy-u16 does not fit) but it
    // measured SLOWER: the clamp add, the fused select and the HSET2 are three ALU-pipe instructions per register where
    // the predicate form has two (its decisions are predicated FADDs on the otherwise idle FMA-lite pipe) --
    // spiral29 18.8 -> 20.9 ms, wy-u8 K9R2 19.9 -> 22.2, ka9q615 37.0 -> 43.7, spiral615 43.4 -> 47.2 (profiles/r02_ab_mask_sat.txt).
#ifndef GRP_MASK_SAT
#define GRP_MASK_SAT 0
#endif
    static constexpr bool MASK = modular || (GRP_MASK_SAT && POLICY != P_WY_U16);
    static constexpr int WINDOW = MASK ? 0x7c00 : 65535;
    static constexpr unsigned TH = (TOP == 65535) ? ((MASK && s16) ? 512u : 4096u) : 16384u;
    // 16-bit policies trigger their renormalisation rarely (threshold >> one branch metric): after a checked step
    // the next `skip` steps provably cannot trigger (state 0 gains at most one branch metric per step), so those
    // steps run without the frame-wide minimum and its barrier. Block-uniform only when a frame owns the block.
    static constexpr bool DEFER = (TOP == 65535) && TPF >= 32;
    // 8-bit policies trigger every few steps, but the trigger itself needs only state 0's metric: a frame that owns
    // its block broadcasts that one lane per step and reduces the frame-wide minimum only on the steps that
    // renormalise (block-uniform branch). Lanes cannot overflow in between: w = tb - base <= 255 - base.
    // Measured, K=15 at sigma = 1: wy-u8 (threshold 183, triggers every ~20 steps) 43.9 -> 39.0 ms per 8192 frames;
    // spiral615 (threshold 74, triggers every few steps, two barriers on those) 43.5 -> 46.1 ms: stays eager.
    static constexpr bool LAZY = (POLICY == P_WY_U8) && TPF >= 32;
    static_assert(NXB >= 3, "at least 4 threads per frame");
    static_assert(!modular || TPF < 32, "the modular prologue unwraps around state 0 with a warp shuffle");
    static_assert(TPF <= 4 || NXB + 1 >= 10, "exchange swizzle is written for K=9 and K=15");
};

struct GrpArgs {
    const unsigned char* syms; size_t sym_stride;
    int nsteps, step0;
    unsigned long long* dec;       // dec[((step0 + t) * n_frames + frame) * TPF + tx]
    uint8_t* row_phase;
    uint16_t* metrics;             // canonical [n_frames][S]
    int n_frames;
    int frame0, count;             // frames [frame0, frame0 + count) of the n_frames-wide layout
    int C, thr, soft_hi, soft_lo;
    int one;                       // 1 (a multiplier ptxas cannot fold: keeps the lo-path adds on the FMA pipe)
    int two;                       // 2 (multiplier of the decision fold, opaque to ptxas: keeps it an IMAD on the FMA pipe)
    unsigned long long* err;       // [n_frames] or nullptr: += the minima this launch's renormalisations subtract
    const int* pidx; int fill;     // fused depuncturing (kernels.h): gather table of this launch's full-rate symbols, erasure value
};

// symbol i of a step (u8 for ka9q/spiral, int8 / int16 for the williamyang policies)
template <int POLICY>
__device__ __forceinline__ int ld_symbol(const unsigned char* step_ptr, int i) {
    if (POLICY == P_WY_U16) return (int)(int16_t)__ldg(reinterpret_cast<const unsigned short*>(step_ptr) + i);
    if (POLICY == P_WY_U8) return (int)(int8_t)__ldg(step_ptr + i);
    return (int)__ldg(step_ptr + i);
}
// symbol i of step t of the launch: straight from the full-rate stream, or (PUNCT kernels, instantiated in
// acs_grp_punct.cu) gathered from the transmitted stream through the puncturing table; `sp` is the frame's symbol base in
// either case. A run-time switch here cost the r = 1/4 kernels 17 % (spiral49 27.4 -> 33.8 ms): compile-time.
template <int POLICY, int R, int SB, bool PUNCT>
__device__ __forceinline__ int ld_step_symbol(const GrpArgs& a, const unsigned char* sp, int t, int i) {
    if (PUNCT) {
        const int ix = __ldg(a.pidx + t * R + i);
        if (ix < 0) return POLICY == P_WY_U16 ? (int)(int16_t)a.fill : POLICY == P_WY_U8 ? (int)(int8_t)a.fill : (a.fill & 255);
        return ld_symbol<POLICY>(sp, ix);
    }
    return ld_symbol<POLICY>(sp + (size_t)t * R * SB, i);
}
// largest single branch metric the lane window is sized for (wy-u16: api.cu checks the symbols of every update
// against the range that guarantees it, and routes other inputs to the generic kernel)
template <int POLICY, int TOP>
__device__ __forceinline__ int bm_cap(int C) { return TOP == 255 ? 255 : (POLICY == P_WY_U16 ? 2 * C : C); }

// scalar branch metric / complement of code word cw for one step's symbols (same formulas as policy.cuh)
template <int POLICY, int R>
__device__ __forceinline__ void bm_eval(unsigned cw, const int* sym, const GrpArgs& a, uint32_t& bm, uint32_t& cm) {
    int v;
    if (POLICY == P_WY_U8) {
        int tot = 0;
#pragma unroll
        for (int i = 0; i < R; i++) {
            const int bt = ((cw >> i) & 1u) ? a.soft_hi : a.soft_lo;
            tot = min(tot + abs(clamp_i(bt - sym[i], -128, 127)), 255);
        }
        v = tot;
    } else if (POLICY == P_WY_U16) {
        int tot = 0;
#pragma unroll
        for (int i = 0; i < R; i++) {
            const int bt = ((cw >> i) & 1u) ? a.soft_hi : a.soft_lo;
            tot = min(tot + abs(clamp_i(bt - sym[i], -32768, 32767)), 65535);
        }
        v = tot;
    } else if (POLICY == P_KA9Q_S16_615) {
        int s = 0;
#pragma unroll
        for (int i = 0; i < R; i++) s += ((cw >> i) & 1u) ? (sym[i] ^ 255) : sym[i];
        v = s;
    } else if (POLICY == P_SPIRAL_R2 || POLICY == P_KA9Q_U8_MOD) {
        const int x0 = (cw & 1u) ? (sym[0] ^ 255) : sym[0];
        const int x1 = (cw & 2u) ? (sym[1] ^ 255) : sym[1];
        const int avg = (x0 + x1 + 1) >> 1;
        v = POLICY == P_KA9Q_U8_MOD ? (avg >> 4) : ((avg >> 2) & 63);
    } else {   // P_SPIRAL_R4, P_SPIRAL_615
        int s = 0;
#pragma unroll
        for (int i = 0; i < R; i++) {
            const int x = ((cw >> i) & 1u) ? (sym[i] ^ 255) : sym[i];
            s = min(s + ((x >> 2) & 63), 255);
        }
        v = (s >> 2) & 63;
    }
    bm = (uint32_t)v;
    cm = (uint32_t)max(a.C - v, 0);
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void lds64(uint32_t addr, uint32_t& x, uint32_t& y) {
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(x), "=r"(y) : "r"(addr));
}
__device__ __forceinline__ void sts64(uint32_t addr, uint32_t x, uint32_t y) {
    asm volatile("st.shared.v2.u32 [%0], {%1, %2};" :: "r"(addr), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void lds128(uint32_t addr, uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d) {
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(addr));
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" :: "r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
// 16-byte-chunk swizzle of the exchange buffer: chunk index ^= low bits of the 128-byte line index
__device__ __forceinline__ uint32_t swz(uint32_t byte_off) { return byte_off ^ (((byte_off >> 7) & 7u) << 4); }

// ---- frame-wide minimum (and state 0's metric) after a step ----------------------------------------------
// lm: the thread's minimum over its 64 lanes (u16 value); w0: the thread's register 0 (state 0 lives in lane 0
// of register 0 of thread tx = 0 in every phase). TPF >= 32: one REDUX per warp, one barrier per step, scratch
// double-buffered by `rpar` so that a single barrier per step suffices. TPF < 32: shuffles inside the warp.
template <class CF>
__device__ __forceinline__ void frame_min(uint32_t lm, uint32_t w0, uint32_t* s_red, int& rpar, int tid,
                                          uint32_t& gmin, uint32_t& g0) {
    if (CF::TPF >= 32) {
        constexpr int NWARP = CF::TPF / 32;
        const uint32_t wm = __reduce_min_sync(0xffffffffu, lm);
        uint32_t* r = s_red + rpar * 16;
        if ((tid & 31) == 0) r[tid >> 5] = wm;
        if (tid == 0) r[NWARP] = w0 & 0xffffu;
        __syncthreads();
        uint32_t m = r[0];
#pragma unroll
        for (int i = 1; i < NWARP; i++) m = min(m, r[i]);
        gmin = m;
        g0 = r[NWARP];
        rpar ^= 1;
    } else {
#pragma unroll
        for (int o = 1; o < CF::TPF; o <<= 1) lm = min(lm, __shfl_xor_sync(0xffffffffu, lm, o));
        gmin = lm;
        g0 = __shfl_sync(0xffffffffu, w0, (tid & 31) & ~(CF::TPF - 1)) & 0xffffu;
    }
}

__device__ __forceinline__ uint32_t local_min(const uint32_t (&w)[NREG]) {
    uint32_t mn = __vimin3_u16x2(w[0], w[1], w[2]);
#pragma unroll
    for (int i = 3; i + 1 < NREG; i += 2) mn = __vimin3_u16x2(mn, w[i], w[i + 1]);
    mn = __vminu2(mn, w[NREG - 1]);
    return min(mn & 0xffffu, mn >> 16);
}

// per-frame running state every thread of the frame keeps identically
struct FrameState {
    int base;          // tb = w + base
    uint32_t clamp;    // packed clamp operand of the saturating adds
    int skip;          // DEFER: steps that need no trigger check (uniform over the frame)
    int skip_cap;      // DEFER: bound that keeps min(w) + spread inside the 16-bit window between rebases
    // minima subtracted by the renormalisations of this launch (reference: update()'s return value): 8-bit policies sum
    // in 32 bits (255 per renormalisation: exact for 16 M steps per launch), 16-bit policies in 64
    uint32_t rsum32;
    unsigned long long rsum;
};

template <int POLICY, class CF>
__device__ __forceinline__ void set_clamp(FrameState& fs, int bm_max) {
    const int c = min(CF::TOP - fs.base, CF::WINDOW - bm_max);
    fs.clamp = (uint32_t)c * 0x10001u;
}

// trigger + rebase after a step, given the frame-wide minimum and w[0] of state 0
template <int POLICY, class CF>
__device__ __forceinline__ void renorm(uint32_t (&w)[NREG], FrameState& fs, uint32_t gmin, uint32_t g0, const GrpArgs& a) {
    const int tb0 = (int)g0 + fs.base;
    const int thr = a.thr + (CF::s16 ? 32768 : 0);
    const bool need = CF::strict ? (tb0 > thr) : (tb0 >= thr);
    if (need) {      // true minimum = gmin + base
        if (CF::TOP == 255) fs.rsum32 += (unsigned)((int)gmin + fs.base);
        else fs.rsum += (unsigned)((int)gmin + fs.base);
        fs.base = -(int)gmin;
    }
    if (gmin > CF::TH) {       // uniform over the frame
        const uint32_t sub = gmin * 0x10001u;
#pragma unroll
        for (int i = 0; i < NREG; i++) w[i] -= sub;      // every lane >= gmin: no borrow between lanes
        fs.base += (int)gmin;
    }
    const int bmx = bm_cap<POLICY, CF::TOP>(a.C);
    set_clamp<POLICY, CF>(fs, bmx);
    if (CF::DEFER) {
        // new[0] <= old[0] + bm_max: no trigger for the next (thr - tb0) / bm_max steps (">=" trigger: thr - 1)
        const int tb0n = need ? (int)g0 - (int)gmin : tb0;
        const int room = thr - (CF::strict ? 0 : 1) - tb0n;
        fs.skip = room > 0 ? min(room / bmx, fs.skip_cap) : 0;
    }
}

// ---- butterflies of one step in phase P -------------------------------------------------------------------
// Pipe balance (ncu: the ALU pipe binds): only the hi-path candidates need the saturating VIADDMNMX -- the
// survivor is min(lo path, hi path) <= hi path <= clamp, and when both saturate the reference's tie -> hi
// decision is reproduced because the hi path equals the clamp -- so the lo-path adds are plain adds issued on
// the FMA pipe (IMAD), and the decisions are accumulated with predicated FADDs (dpx.cuh: min_facc).
struct DecAcc { float al, ah, bl, bh; };

// Where a register pair finds its packed branch metrics {W, Wc} of (static) code word c:
//  * TabLds: one LDS.64 from the step's shared-memory table at entry c ^ (the thread's own code-word contribution);
//  * TabReg (step tables of four entries, r = 1/2): the thread fetches its permutation of the whole table once per step
//    (4 LDS.64) and every pair reads registers: 12 LDS.64 and 12 address LOP3s fewer per step.
struct TabLds {
    uint32_t A;
    template <unsigned C> __device__ __forceinline__ void get(uint32_t& W, uint32_t& Wc) const { lds64(A ^ (C << 3), W, Wc); }
};
struct TabReg {
    uint32_t W[4], Wc[4];
    __device__ __forceinline__ explicit TabReg(uint32_t A) {
#pragma unroll
        for (unsigned c = 0; c < 4; c++) lds64(A ^ (c << 3), W[c], Wc[c]);
    }
    template <unsigned C> __device__ __forceinline__ void get(uint32_t& w, uint32_t& wc) const { w = W[C]; wc = Wc[C]; }
};

template <int POLICY, class CODE, int P, int KK>
struct PairLoop {
    template <class TAB>
    __device__ __forceinline__ static void run(uint32_t (&w)[NREG], const TAB& tab, uint32_t clamp, uint32_t one, DecAcc& d) {
        constexpr int NXB = CODE::K - 1 - NL;
        constexpr int XB = NL - 1 - P;
        constexpr int iA = k7::expand_bit(KK, XB), iB = iA | (1 << XB);
        constexpr unsigned c = CODE::cw2(state_of(NXB, P, iA, 0) << 1);
        uint32_t W, Wc;
        tab.template get<c>(W, Wc);
        const uint32_t m0 = add_fma(w[iA], one, W), m1 = __viaddmin_u16x2(w[iB], Wc, clamp);
        const uint32_t m2 = add_fma(w[iA], one, Wc), m3 = __viaddmin_u16x2(w[iB], W, clamp);
        // decision = (min == hi path): tie -> 1 (viterbi615_sse2.cpp:145-148, spiral615.cpp:247-250, sse_u8.h:111-114)
        w[iA] = min_facc<KK>(m1, m0, d.al, d.ah);
        w[iB] = min_facc<KK>(m3, m2, d.bl, d.bh);
        PairLoop<POLICY, CODE, P, KK + 1>::run(w, tab, clamp, one, d);
    }
};
template <int POLICY, class CODE, int P>
struct PairLoop<POLICY, CODE, P, 16> {
    template <class TAB>
    __device__ __forceinline__ static void run(uint32_t (&)[NREG], const TAB&, uint32_t, uint32_t, DecAcc&) {}
};

// ka9q modular policy (viterbi29_sse2.cpp:137-158), mask decisions (dpx.cuh): the lanes hold the UNWRAPPED metric U
// (reference metric = U mod 256; the kernel prologue unwraps around state 0; the spread stays below 128 -- bound in the
// file header of acs_k7.cu, pinned by the CPU test suite -- so the reference's (int8)(m0 - m1) > 0 is the plain U compare):
// survivor = min, decision = (lo path > hi path), tie -> 0 = (survivor != lo path). No saturation, no renormalisation.
// Per register: plain add (lo path), VIADDMNMX (min with hi path + metric), HSET2 (lane mask), IMAD (fold): with the table
// fetch 10 instructions per register pair instead of 12, and no predicates. KK runs 15 ... 0 so that register pair KK
// lands at bit KK / 16 + KK of the folded words (the layout of acs_grp.cuh). Window note: U < 512 + 128 + 1280 * 15 + 15
// between the rebases (every 256 groups), inside [0, 0x7c00].
template <int POLICY, class CODE, int P, int KK>
struct PairLoopM {
    template <class TAB>
    __device__ __forceinline__ static void run(uint32_t (&w)[NREG], const TAB& tab, uint32_t two, uint32_t& h0, uint32_t& h1) {
        constexpr int NXB = CODE::K - 1 - NL;
        constexpr int XB = NL - 1 - P;
        constexpr int iA = k7::expand_bit(KK, XB), iB = iA | (1 << XB);
        constexpr unsigned c = CODE::cw2(state_of(NXB, P, iA, 0) << 1);
        uint32_t W, Wc;
        tab.template get<c>(W, Wc);
        const uint32_t t0 = w[iA] + W;
        const uint32_t n0 = __viaddmin_u16x2(w[iB], Wc, t0);
        h0 = horner_mask(h0, two, ne_mask_u16x2(n0, t0));
        const uint32_t t1 = w[iA] + Wc;
        const uint32_t n1 = __viaddmin_u16x2(w[iB], W, t1);
        h1 = horner_mask(h1, two, ne_mask_u16x2(n1, t1));
        w[iA] = n0;
        w[iB] = n1;
        PairLoopM<POLICY, CODE, P, KK - 1>::run(w, tab, two, h0, h1);
    }
};
template <int POLICY, class CODE, int P>
struct PairLoopM<POLICY, CODE, P, -1> {
    template <class TAB>
    __device__ __forceinline__ static void run(uint32_t (&)[NREG], const TAB&, uint32_t, uint32_t&, uint32_t&) {}
};

// saturating policies, mask decisions: hi path with the saturating clamp, survivor = min(lo + metric, hi path) as the second
// VIADDMNMX, decision = (survivor == hi path): tie -> 1 (viterbi615_sse2.cpp:145-148, spiral615.cpp:247-250,
// sse_u8.h:111-114), also when both paths saturate (the hi path then equals the clamp and the lo path exceeds it).
template <int POLICY, class CODE, int P, int KK>
struct PairLoopS {
    template <class TAB>
    __device__ __forceinline__ static void run(uint32_t (&w)[NREG], const TAB& tab, uint32_t clamp, uint32_t two,
                                               uint32_t& h0, uint32_t& h1) {
        constexpr int NXB = CODE::K - 1 - NL;
        constexpr int XB = NL - 1 - P;
        constexpr int iA = k7::expand_bit(KK, XB), iB = iA | (1 << XB);
        constexpr unsigned c = CODE::cw2(state_of(NXB, P, iA, 0) << 1);
        uint32_t W, Wc;
        tab.template get<c>(W, Wc);
        const uint32_t m1 = __viaddmin_u16x2(w[iB], Wc, clamp);
        const uint32_t n0 = __viaddmin_u16x2(w[iA], W, m1);
        h0 = horner_mask(h0, two, eq_mask_u16x2(n0, m1));
        const uint32_t m3 = __viaddmin_u16x2(w[iB], W, clamp);
        const uint32_t n1 = __viaddmin_u16x2(w[iA], Wc, m3);
        h1 = horner_mask(h1, two, eq_mask_u16x2(n1, m3));
        w[iA] = n0;
        w[iB] = n1;
        PairLoopS<POLICY, CODE, P, KK - 1>::run(w, tab, clamp, two, h0, h1);
    }
};
template <int POLICY, class CODE, int P>
struct PairLoopS<POLICY, CODE, P, -1> {
    template <class TAB>
    __device__ __forceinline__ static void run(uint32_t (&)[NREG], const TAB&, uint32_t, uint32_t, uint32_t&, uint32_t&) {}
};

// exchange after a whole group: phase-5 layout (register q, lane, tx) = state (2tx+lane)*32 + q  ->  shared
template <class CF>
__device__ __forceinline__ void exchange_store(const uint32_t (&w)[NREG], uint32_t met_addr, int tx) {
    const uint32_t line = met_addr + (uint32_t)tx * 128u;
    const uint32_t key = ((uint32_t)tx & 7u) << 4;
#pragma unroll
    for (int ch = 0; ch < 8; ch++) {
        const int q0 = (ch & 3) * 8;
        const uint32_t sel = (ch < 4) ? 0x5410u : 0x7632u;     // lane 0 halves / lane 1 halves of (w[q], w[q+1])
        sts128(line + (((uint32_t)ch << 4) ^ key),
               __byte_perm(w[q0], w[q0 + 1], sel), __byte_perm(w[q0 + 2], w[q0 + 3], sel),
               __byte_perm(w[q0 + 4], w[q0 + 5], sel), __byte_perm(w[q0 + 6], w[q0 + 7], sel));
    }
}
// shared -> phase-0 layout: register q = states (q * 2^NXB + 2tx, +1)
template <class CF>
__device__ __forceinline__ void exchange_load(uint32_t (&w)[NREG], uint32_t met_addr, int tx) {
    if (CF::NXB + 1 >= 10) {
        const uint32_t t = met_addr + swz(4u * (uint32_t)tx);
#pragma unroll
        for (int q = 0; q < NREG; q++) w[q] = lds32(t + ((uint32_t)q << (CF::NXB + 1)));
    } else {
        const uint32_t t = met_addr + 4u * (uint32_t)tx;
#pragma unroll
        for (int q = 0; q < NREG; q++) w[q] = lds32(t + swz((uint32_t)q << (CF::NXB + 1)));
    }
}

// ---- per-frame entry permutation of 128-byte step tables (K=9 r=1/4) -----------------------------------------
// The four threads of a frame read entries c ^ cx(tx), a coset of V_p = span(cw2(1 << (p+2)), cw2(1 << (p+3))).
// Frame f of a half-warp (4 frames = one LDS.64 transaction) stores entry c at position c ^ H_p(f & 3), where
// H_p(.) runs over coset representatives of V_p, so the 16 lanes of the transaction hit 16 different entries.
__host__ __device__ constexpr unsigned gf2_reduce(unsigned v, unsigned b0, unsigned b1, unsigned b2, unsigned b3) {
    for (int it = 0; it < 8; it++) {
        if (b0 && (v ^ b0) < v) v ^= b0;
        if (b1 && (v ^ b1) < v) v ^= b1;
        if (b2 && (v ^ b2) < v) v ^= b2;
        if (b3 && (v ^ b3) < v) v ^= b3;
    }
    return v;
}
// which-th (0 or 1) unit vector that extends {g1, g2} towards a basis of GF(2)^R
__host__ __device__ constexpr unsigned gf2_complement(unsigned g1, unsigned g2, int R, int which) {
    unsigned b0 = g1, b1 = gf2_reduce(g2, g1, 0, 0, 0), b2 = 0, b3 = 0;
    int found = 0;
    for (int k = 0; k < R; k++) {
        const unsigned r = gf2_reduce(1u << k, b0, b1, b2, b3);
        if (r) {
            if (found == which) return 1u << k;
            if (found == 0) b2 = r; else b3 = r;
            found++;
        }
    }
    return 0;
}
template <class CODE>
__host__ __device__ constexpr unsigned frame_perm(int p, int f) {
    const unsigned g1 = CODE::cw2(1u << (p + 2)), g2 = CODE::cw2(1u << (p + 3));
    return ((f & 1) ? gf2_complement(g1, g2, CODE::R, 0) : 0u) ^ ((f & 2) ? gf2_complement(g1, g2, CODE::R, 1) : 0u);
}

// ---- per-symbol costs: every policy's branch metric is post(sum_i v_i[c_i]) (same formulas as bm_eval) -----
template <int POLICY>
__device__ __forceinline__ void symbol_costs(int sym, const GrpArgs& a, int& v0, int& v1) {
    if (POLICY == P_WY_U8) {
        v0 = abs(clamp_i(a.soft_lo - sym, -128, 127));
        v1 = abs(clamp_i(a.soft_hi - sym, -128, 127));
    } else if (POLICY == P_WY_U16) {
        v0 = abs(clamp_i(a.soft_lo - sym, -32768, 32767));
        v1 = abs(clamp_i(a.soft_hi - sym, -32768, 32767));
    } else if (POLICY == P_SPIRAL_R4 || POLICY == P_SPIRAL_615) {
        v0 = (sym >> 2) & 63;
        v1 = ((sym ^ 255) >> 2) & 63;
    } else {
        v0 = sym;
        v1 = sym ^ 255;
    }
}
template <int POLICY>
__device__ __forceinline__ int bm_post(int sum) {
    if (POLICY == P_WY_U8) return min(sum, 255);
    if (POLICY == P_WY_U16) return min(sum, 65535);
    if (POLICY == P_SPIRAL_R4 || POLICY == P_SPIRAL_615) return (min(sum, 255) >> 2) & 63;
    if (POLICY == P_SPIRAL_R2) return (((sum + 1) >> 1) >> 2) & 63;
    if (POLICY == P_KA9Q_U8_MOD) return ((sum + 1) >> 1) >> 4;
    return sum;
}

// The symbols of one group of five steps in registers (K=9 kernels, GRP_SYM_PREFETCH): loaded one group AHEAD of the table
// build that consumes them, so the build does not wait on global memory (ncu, ka9q29: 0.98 long-scoreboard stall cycles
// per issue with the loads inside the build). An experiment that did not pay: see the kernel.
template <int R, int SB> struct GroupSyms {
    static constexpr int BYTES = NL * R * SB, WORDS = (BYTES + 3) / 4;
    uint32_t w[WORDS];
    __device__ __forceinline__ void load(const unsigned char* p, int valid_bytes) {
#pragma unroll
        for (int i = 0; i < WORDS; i++) {
            uint32_t v = 0;
#pragma unroll
            for (int b = 0; b < 4; b++)
                if (i * 4 + b < BYTES && i * 4 + b < valid_bytes) v |= (uint32_t)__ldg(p + i * 4 + b) << (8 * b);
            w[i] = v;
        }
    }
    template <int POLICY>
    __device__ __forceinline__ int symbol(int s, int i) const {      // s, i compile-time after unrolling
        const int b = (s * R + i) * SB;
        const uint32_t v = (w[b >> 2] >> ((b & 3) * 8)) & (SB == 2 ? 0xffffu : 0xffu);
        return (SB == 2) ? (int)(int16_t)v : (POLICY == P_WY_U8) ? (int)(int8_t)v : (int)v;
    }
};

// Cooperative table build for TPF = 4 (K=9): per step, each of the frame's four threads evaluates NCW/4 code
// words as subset sums of the per-symbol costs, fetches the lane-1 partner (code word ^ D_s) with one shuffle
// and writes its NCW/4 entries. ~4x fewer instructions than one bm_eval per code word and task.
template <int POLICY, class CODE, class CF, bool PUNCT>
__device__ __forceinline__ void fill_tables_coop(const GrpArgs& a, const unsigned char* sp, int g, uint32_t tab_addr, int tx, int fl,
                                                 const GroupSyms<CF::R, CF::SB>* gs = nullptr) {
    constexpr int R = CF::R, NPT = CF::NCW / 4, LOGN = (NPT == 1 ? 0 : NPT == 2 ? 1 : NPT == 4 ? 2 : NPT == 8 ? 3 : 4);
    static_assert(CF::TPF == 4 && CF::NCW >= 4 && (1 << LOGN) == NPT, "four threads per frame, NCW/4 code words each");
    const uint32_t buf = tab_addr + (uint32_t)(g & 1) * (NL * CF::SUBTAB);
#pragma unroll
    for (int s = 0; s < NL; s++) {
        const int t = g * NL + s;
        if (t >= a.nsteps) break;                       // uniform over the block
        constexpr int DUMMY = 0; (void)DUMMY;
        const unsigned D = CODE::cw2(1u << (s + 1));    // compile-time after unrolling
        int v0[R], v1[R];
#pragma unroll
        for (int i = 0; i < R; i++) {
            const int sym = (!PUNCT && gs) ? gs->template symbol<POLICY>(s, i) : ld_step_symbol<POLICY, R, CF::SB, PUNCT>(a, sp, t, i);
            symbol_costs<POLICY>(sym, a, v0[i], v1[i]);
        }
        // code words of this thread: c = tx * NPT + j; the two top bits of c come from tx
        int top = ((tx & 1) ? v1[LOGN] : v0[LOGN]) + ((tx & 2) ? v1[LOGN + 1] : v0[LOGN + 1]);
        uint32_t X[NPT];                                // (bm, cm) packed
#pragma unroll
        for (int j = 0; j < NPT; j++) {
            int sum = top;
#pragma unroll
            for (int i = 0; i < LOGN; i++) sum += ((j >> i) & 1) ? v1[i] : v0[i];
            const int bm = bm_post<POLICY>(sum), cm = max(a.C - bm, 0);
            X[j] = pack2((uint32_t)bm, (uint32_t)cm);
        }
        const unsigned fp = CF::PERMUTE ? (((fl & 1) ? frame_perm<CODE>(s, 1) : 0u) ^ ((fl & 2) ? frame_perm<CODE>(s, 2) : 0u)) : 0u;
        const uint32_t st = buf + (uint32_t)s * CF::SUBTAB;
        const unsigned c0 = (unsigned)(tx * NPT);
#pragma unroll
        for (int j = 0; j < NPT; j++) {
            const int jp = j ^ (int)(D & (NPT - 1));
            const uint32_t Xd = (D >> LOGN) ? __shfl_xor_sync(0xffffffffu, X[jp], (int)(D >> LOGN)) : X[jp];
            const uint32_t W = __byte_perm(X[j], Xd, 0x5410), Wc = __byte_perm(X[j], Xd, 0x7632);
                sts64(st + (((c0 + j) ^ fp) << 3), W, Wc);
        }
    }
}

// ---- branch-metric tables of one group of five steps ---------------------------------------------------
// entry (s, c) = { pack(bm[c], bm[c ^ D_s]), pack(cm[c], cm[c ^ D_s]) }, D_s = code-word difference between the
// two lanes of a register in phase s. One task computes bm of c and c ^ D_s once and writes both entries.
template <int POLICY, class CODE, class CF, bool PUNCT>
__device__ __forceinline__ void fill_tables(const GrpArgs& a, const unsigned char* sp, int g, uint32_t tab_addr, int tx, int fl,
                                            const GroupSyms<CF::R, CF::SB>* gs = nullptr) {
    constexpr int R = CF::R, HALF = CF::NCW / 2;
    if constexpr (CF::TPF == 4 && CF::NCW >= 4) {
        fill_tables_coop<POLICY, CODE, CF, PUNCT>(a, sp, g, tab_addr, tx, fl, gs);
        return;
    }
    const uint32_t buf = tab_addr + (uint32_t)(g & 1) * (NL * CF::SUBTAB);
    for (int e = tx; e < NL * HALF; e += CF::TPF) {
        const int s = e / HALF, idx = e - s * HALF;
        const int t = g * NL + s;
        if (t >= a.nsteps) break;
        const unsigned D = CODE::cw2(1u << (s + 1));
        unsigned c1, c2;
        if (D) { c1 = (unsigned)k7::expand_bit(idx, 31 - __clz(D)); c2 = c1 ^ D; }
        else { c1 = (unsigned)idx; c2 = (unsigned)(idx + HALF); }
        int sym[R];
#pragma unroll
        for (int i = 0; i < R; i++) sym[i] = ld_step_symbol<POLICY, R, CF::SB, PUNCT>(a, sp, t, i);
        uint32_t b1, m1, b2, m2;
        bm_eval<POLICY, R>(c1, sym, a, b1, m1);
        bm_eval<POLICY, R>(c2, sym, a, b2, m2);
        const uint32_t st = buf + (uint32_t)s * CF::SUBTAB;
        if (D) {
            sts64(st + c1 * 8, pack2(b1, b2), pack2(m1, m2));
            sts64(st + c2 * 8, pack2(b2, b1), pack2(m2, m1));
        } else {
            sts64(st + c1 * 8, pack2(b1, b1), pack2(m1, m1));
            sts64(st + c2 * 8, pack2(b2, b2), pack2(m2, m2));
        }
    }
}

// ---- one trellis step in phase P -------------------------------------------------------------------------
// EXCH: this is the fifth step of a whole group: store the metrics for the exchange before the step's barrier
// and reload them in the phase-0 layout after it.
template <int POLICY, class CODE, class CF, int P, bool EXCH>
__device__ __forceinline__ void step(uint32_t (&w)[NREG], FrameState& fs, const GrpArgs& a, uint32_t sub_addr,
                                     uint32_t cx8, unsigned long long* dec_out, bool live, uint32_t met_addr,
                                     uint32_t* s_red, int& rpar, int tid, int tx) {
#ifndef GRP_TABREG
#define GRP_TABREG 1
#endif
    using TAB = typename std::conditional<(GRP_TABREG && CF::NCW == 4), TabReg, TabLds>::type;
    const TAB tab{ sub_addr + cx8 };
    if constexpr (CF::modular) {
        uint32_t h0 = 0, h1 = 0;
        PairLoopM<POLICY, CODE, P, 15>::run(w, tab, fs.clamp, h0, h1);     // modular policy: the clamp slot carries 2
        if (live) *dec_out = (unsigned long long)horner_decode(h0) | ((unsigned long long)horner_decode(h1) << 32);
    } else if constexpr (CF::MASK) {
        uint32_t h0 = 0, h1 = 0;
        PairLoopS<POLICY, CODE, P, 15>::run(w, tab, fs.clamp, (uint32_t)a.two, h0, h1);
        if (live) *dec_out = (unsigned long long)horner_decode(h0) | ((unsigned long long)horner_decode(h1) << 32);
    } else {
        DecAcc d = { FDEC_ZERO, FDEC_ZERO, FDEC_ZERO, FDEC_ZERO };
        PairLoop<POLICY, CODE, P, 0>::run(w, tab, fs.clamp, (uint32_t)a.one, d);
        if (live) *dec_out = (unsigned long long)fdec_pack(d.al, d.ah) | ((unsigned long long)fdec_pack(d.bl, d.bh) << 32);
    }
    if constexpr (CF::modular) {
        if (EXCH) {
            exchange_store<CF>(w, met_addr, tx);
            if (CF::TPF < 32) __syncwarp(); else __syncthreads();
            exchange_load<CF>(w, met_addr, tx);
        }
        return;
    }
    if constexpr (CF::DEFER) {
        if (fs.skip > 0) {
            fs.skip--;
            if (EXCH) {
                exchange_store<CF>(w, met_addr, tx);
                __syncthreads();
                exchange_load<CF>(w, met_addr, tx);
            }
            return;
        }
    }
    if constexpr (CF::LAZY) {
        uint32_t* r = s_red + rpar * 16;
        if (tid == 0) r[15] = w[0] & 0xffffu;           // state 0: lane 0 of register 0 of thread 0 in every phase
        if (EXCH) exchange_store<CF>(w, met_addr, tx);
        __syncthreads();
        const uint32_t g0 = r[15];
        rpar ^= 1;
        if (EXCH) exchange_load<CF>(w, met_addr, tx);
        const int tb0 = (int)g0 + fs.base;
        if (CF::strict ? (tb0 > a.thr) : (tb0 >= a.thr)) {      // uniform over the block
            uint32_t gmin, g0b;
            frame_min<CF>(local_min(w), 0u, s_red, rpar, tid, gmin, g0b);
            renorm<POLICY, CF>(w, fs, gmin, g0, a);
        }
        return;
    }
    const uint32_t lm = local_min(w);
    const uint32_t w0 = w[0];
    if (EXCH) {
        exchange_store<CF>(w, met_addr, tx);
        if (CF::TPF < 32) __syncwarp();
    }
    uint32_t gmin, g0;
    frame_min<CF>(lm, w0, s_red, rpar, tid, gmin, g0);
    if (EXCH) exchange_load<CF>(w, met_addr, tx);
    renorm<POLICY, CF>(w, fs, gmin, g0, a);
}

#ifndef GRP_MINB
#define GRP_MINB 4
#endif
#ifndef GRP_SKIP_CAP
#define GRP_SKIP_CAP 22      // DEFER: most steps without a trigger check in a row (the 16-bit window allows 23 for ka9q615; 10 -> 22: 37.4 -> 37.0 ms)
#endif
template <int POLICY, class CODE, bool PUNCT>
__global__ void __launch_bounds__((Cfg<POLICY, CODE>::THREADS), (Cfg<POLICY, CODE>::TPF >= 32 ? 3 : GRP_MINB))
acs_grp_kernel(const GrpArgs a) {
    using CF = Cfg<POLICY, CODE>;
    extern __shared__ __align__(1024) unsigned char smem[];
    const int tid = threadIdx.x;
    const int fl = tid / CF::TPF, tx = tid % CF::TPF;
    const int fidx = blockIdx.x * CF::FPB + fl;
    const bool live = fidx < a.count;
    const int frame = a.frame0 + (live ? fidx : a.count - 1);

    const uint32_t tab_addr = smem_u32(smem) + (uint32_t)fl * CF::TAB_STRIDE;
    unsigned char* met_ptr = smem + CF::FPB * CF::TAB_STRIDE + (size_t)fl * CF::MET_STRIDE;
    const uint32_t met_addr = smem_u32(met_ptr);
    uint32_t* s_red = reinterpret_cast<uint32_t*>(smem + CF::FPB * CF::TAB_STRIDE + CF::FPB * CF::MET_STRIDE);
    int rpar = 0;

    if (blockIdx.x == 0)
        for (int t = tid; t < a.nsteps; t += blockDim.x) a.row_phase[a.step0 + t] = (uint8_t)(t % NL);

    // ---- canonical metrics -> phase-0 layout: register q = states (q * 2^NXB + 2tx, +1) ----
    uint32_t w[NREG];
    uint16_t* gm = a.metrics + (size_t)frame * CF::S;
    {
        const uint32_t* g32 = reinterpret_cast<const uint32_t*>(gm);
#pragma unroll
        for (int q = 0; q < NREG; q++) {
            w[q] = g32[q * CF::TPF + tx];
            if (CF::s16) w[q] ^= 0x80008000u;          // s16 -> biased u16
        }
    }
    if (CF::modular) {
        // unwrap the reference's mod-256 metrics around state 0 (spread < 128): U = 256 + m0 + (int8)(m - m0).
        // State 0 is lane 0 of register 0 of the frame's thread 0 in the phase-0 layout.
        const int m0 = (int)(__shfl_sync(0xffffffffu, w[0], (tid & 31) & ~(CF::TPF - 1)) & 0xffu);
#pragma unroll
        for (int q = 0; q < NREG; q++) {
            const int lo = 256 + m0 + (int)(int8_t)((w[q] & 0xffu) - m0);
            const int hi = 256 + m0 + (int)(int8_t)(((w[q] >> 16) & 0xffu) - m0);
            w[q] = pack2((uint32_t)lo, (uint32_t)hi);
        }
    }
    FrameState fs;
    fs.base = 0;
    fs.rsum = 0;
    fs.rsum32 = 0;
    fs.clamp = CF::modular ? (uint32_t)a.two : 0u;      // modular policy: no clamp; the slot carries the fold's multiplier 2
    fs.skip = 0;
    {   // min(w) <= TH + skip * bm_max between rebases; window: that + (K + 1) branch metrics of spread and headroom
        const int bmx = max(bm_cap<POLICY, CF::TOP>(a.C), 1);
        fs.skip_cap = max(0, min(GRP_SKIP_CAP, (CF::WINDOW - (int)CF::TH - (CF::K + 1) * bmx) / bmx - 1));
    }
    if (!CF::modular) {   // bring the lanes down to the bottom of the 16-bit window
        uint32_t gmin, g0;
        frame_min<CF>(local_min(w), w[0], s_red, rpar, tid, gmin, g0);
        const uint32_t sub = gmin * 0x10001u;
#pragma unroll
        for (int i = 0; i < NREG; i++) w[i] -= sub;
        fs.base = (int)gmin;
        set_clamp<POLICY, CF>(fs, bm_cap<POLICY, CF::TOP>(a.C));
    }

    // the thread's own contribution to the code word of its butterflies, per phase (byte offset into a table)
    uint32_t cx8[NL];
#pragma unroll
    for (int p = 0; p < NL; p++) {
        unsigned c = 0;
#pragma unroll
        for (int i = 0; i < CF::R; i++) c |= (unsigned)(__popc(((unsigned)tx << (p + 2)) & CODE::poly(i)) & 1) << i;
        if (CF::PERMUTE) c ^= ((fl & 1) ? frame_perm<CODE>(p, 1) : 0u) ^ ((fl & 2) ? frame_perm<CODE>(p, 2) : 0u);
        cx8[p] = c << CF::ESHIFT;
    }

    const unsigned char* sp = a.syms + (size_t)frame * a.sym_stride;
    const size_t dstride = (size_t)a.n_frames * CF::TPF;
    unsigned long long* drow = a.dec + ((size_t)a.step0 * a.n_frames + frame) * CF::TPF + tx;

    const int ngroups = (a.nsteps + NL - 1) / NL;
    // Measured and left off: with the prefetch the K=9 kernels take 7-9 more registers (ka9q29 96 -> 103: one resident
    // block fewer) and run SLOWER -- ka9q29 14.27 -> 15.08 ms, spiral29 18.83 -> 19.48, spiral49 27.78 -> 28.46, wy-u8 K9R4
    // 30.17 -> 32.11 per 256 Ki frames; capping the registers (5 blocks) spills: 14.93 / 19.80 / 29.92 / 33.70.
#ifndef GRP_SYM_PREFETCH
#define GRP_SYM_PREFETCH 0
#endif
    constexpr bool PREFETCH = GRP_SYM_PREFETCH && CF::TPF == 4 && CF::NCW >= 4 && !PUNCT;
    using GS = GroupSyms<CF::R, CF::SB>;
    GS gs;                                   // symbols of the group whose tables are built next
    auto load_group = [&](int g) {
        if (PREFETCH && g < ngroups) gs.load(sp + (size_t)g * GS::BYTES, (a.nsteps - g * NL) * CF::R * CF::SB);
    };
    load_group(0);
    if (ngroups > 0) fill_tables<POLICY, CODE, CF, PUNCT>(a, sp, 0, tab_addr, tx, fl, PREFETCH ? &gs : nullptr);
    load_group(1);
    if (CF::TPF >= 32) __syncthreads(); else __syncwarp();

    // Whole groups run their five steps unconditionally: a step guarded by `if (nsub > s)` ends in a join where ptxas
    // must bring the 32 renamed metric registers of the taken path back to the places of the skipped one (measured on
    // ka9q29: 24 IMAD.MOV per step, 12 % of the instruction stream). Only the last, partial group keeps the guards.
    const int nfull = a.nsteps / NL;
    auto modular_rebase = [&](int g) {
        if (CF::modular && (g & 255) == 255) {
            // unwrapped lanes gain at most 15 per step: every 1280 steps bring them down by a multiple of 256
            // (keeps U mod 256), leaving the minimum in [256, 512); the lanes stay inside the window of the mask compare
            uint32_t gmin, g0;
            frame_min<CF>(local_min(w), w[0], s_red, rpar, tid, gmin, g0);
            const uint32_t sub = (gmin & ~255u) >= 256u ? (gmin & ~255u) - 256u : 0u;
#pragma unroll
            for (int i = 0; i < NREG; i++) w[i] -= sub * 0x10001u;
        }
    };
#ifndef GRP_UNCOND_ALL
#define GRP_UNCOND_ALL 0
#endif
    constexpr bool UNCOND = CF::modular || GRP_UNCOND_ALL;
    const int nmain = UNCOND ? nfull : 0;
#pragma unroll 1
    for (int g = 0; g < nmain; g++) {
        {
            const GS cur = gs;
            load_group(g + 2);
            if (g + 1 < ngroups) fill_tables<POLICY, CODE, CF, PUNCT>(a, sp, g + 1, tab_addr, tx, fl, PREFETCH ? &cur : nullptr);
        }
        const uint32_t gt = tab_addr + (uint32_t)(g & 1) * (NL * CF::SUBTAB);
        step<POLICY, CODE, CF, 0, false>(w, fs, a, gt, cx8[0], drow, live, met_addr, s_red, rpar, tid, tx);
        step<POLICY, CODE, CF, 1, false>(w, fs, a, gt + CF::SUBTAB, cx8[1], drow + dstride, live, met_addr, s_red, rpar, tid, tx);
        step<POLICY, CODE, CF, 2, false>(w, fs, a, gt + 2 * CF::SUBTAB, cx8[2], drow + 2 * dstride, live, met_addr, s_red, rpar, tid, tx);
        step<POLICY, CODE, CF, 3, false>(w, fs, a, gt + 3 * CF::SUBTAB, cx8[3], drow + 3 * dstride, live, met_addr, s_red, rpar, tid, tx);
        step<POLICY, CODE, CF, 4, true>(w, fs, a, gt + 4 * CF::SUBTAB, cx8[4], drow + 4 * dstride, live, met_addr, s_red, rpar, tid, tx);
        drow += NL * dstride;
        modular_rebase(g);
    }
    // the last, partial group (and, for the kernels that measured slower with unguarded steps, every group)
#pragma unroll 1
    for (int g = nmain; g < ngroups; g++) {
        const int nsub = min(NL, a.nsteps - g * NL);
        {
            const GS cur = gs;
            load_group(g + 2);
            if (g + 1 < ngroups) fill_tables<POLICY, CODE, CF, PUNCT>(a, sp, g + 1, tab_addr, tx, fl, PREFETCH ? &cur : nullptr);
        }
        const uint32_t gt = tab_addr + (uint32_t)(g & 1) * (NL * CF::SUBTAB);
        step<POLICY, CODE, CF, 0, false>(w, fs, a, gt, cx8[0], drow, live, met_addr, s_red, rpar, tid, tx);
        if (nsub > 1) step<POLICY, CODE, CF, 1, false>(w, fs, a, gt + CF::SUBTAB, cx8[1], drow + dstride, live, met_addr, s_red, rpar, tid, tx);
        if (nsub > 2) step<POLICY, CODE, CF, 2, false>(w, fs, a, gt + 2 * CF::SUBTAB, cx8[2], drow + 2 * dstride, live, met_addr, s_red, rpar, tid, tx);
        if (nsub > 3) step<POLICY, CODE, CF, 3, false>(w, fs, a, gt + 3 * CF::SUBTAB, cx8[3], drow + 3 * dstride, live, met_addr, s_red, rpar, tid, tx);
        if (nsub > 4) step<POLICY, CODE, CF, 4, true>(w, fs, a, gt + 4 * CF::SUBTAB, cx8[4], drow + 4 * dstride, live, met_addr, s_red, rpar, tid, tx);
        drow += NL * dstride;
        modular_rebase(g);
    }

    // ---- registers (phase pf layout) -> canonical metrics, through shared memory ----
    const int pf = a.nsteps % NL;
    if (CF::TPF >= 32) __syncthreads(); else __syncwarp();
#pragma unroll
    for (int q = 0; q < NREG; q++)
        asm volatile("st.shared.u32 [%0], %1;" :: "r"(met_addr + (uint32_t)(q * CF::TPF + tx) * 4u), "r"(w[q]) : "memory");
    if (CF::TPF >= 32) __syncthreads(); else __syncwarp();
    if (live && tx == 0 && a.err && (fs.rsum | fs.rsum32)) a.err[frame] += fs.rsum + fs.rsum32;
    if (live) {
        const uint16_t* raw = reinterpret_cast<const uint16_t*>(met_ptr);
        for (unsigned n = (unsigned)tx; n < (unsigned)CF::S; n += CF::TPF) {
            int q;
            unsigned xx;
            locate_state(CF::NXB, pf, n, q, xx);
            const int rv = (int)raw[(q * CF::TPF + (int)(xx >> 1)) * 2 + (int)(xx & 1u)];
            const int tb = rv + fs.base;
            gm[n] = (uint16_t)(CF::modular ? (rv & 0xff) : CF::s16 ? (tb ^ 0x8000) : tb);
        }
    }
}

template <int POLICY, class CODE>
static cudaError_t launch_one(const GrpArgs& a, cudaStream_t st) {
    using CF = Cfg<POLICY, CODE>;
    auto k = acs_grp_kernel<POLICY, CODE, GRP_PUNCT_TU != 0>;
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, CF::SMEM);
    if (e != cudaSuccess) return e;
    const int blocks = (a.count + CF::FPB - 1) / CF::FPB;
    k<<<blocks, CF::THREADS, CF::SMEM, st>>>(a);
    return cudaGetLastError();
}

using CodeK9R2 = Code<9, 2, 0x1af, 0x11d, 0, 0, 0, 0>;                                   // src/main.cpp:386
using CodeK9R4 = Code<9, 4, 501, 441, 331, 315, 0, 0>;                                   // src/main.cpp:396
using CodeIS95 = Code<9, 2, 491, 369, 0, 0, 0, 0>;       // williamyang_viterbi_lib/examples/helpers/common_codes.h:28 (CDMA IS-95A)
using CodeK15R6 = Code<15, 6, 042631, 047245, 056507, 073363, 077267, 064537>;           // src/main.cpp:405

template <class CODE>
static bool same_code(const CodecParams& c) {
    if (c.K != CODE::K || c.R != CODE::R) return false;
    for (int i = 0; i < CODE::R; i++)
        if (c.poly[i] != (int)CODE::poly(i)) return false;
    return true;
}

enum Instance { I_NONE = -1, I_KA9Q29, I_KA9Q615, I_SPIRAL615, I_WY8_K15R6, I_SPIRAL29, I_WY8_K9R2, I_SPIRAL49, I_WY8_K9R4,
                I_WY16_K15R6, I_WY16_K9R2, I_WY16_K9R4, I_WY8_IS95, I_WY16_IS95 };

static int find_instance(const CodecParams& c) {
    if (same_code<CodeK15R6>(c)) {
        if (c.policy == P_KA9Q_S16_615) return I_KA9Q615;
        if (c.policy == P_SPIRAL_615) return I_SPIRAL615;
        if (c.policy == P_WY_U8) return I_WY8_K15R6;
        if (c.policy == P_WY_U16) return I_WY16_K15R6;
    } else if (same_code<CodeK9R2>(c)) {
        if (c.policy == P_KA9Q_U8_MOD) return I_KA9Q29;
        if (c.policy == P_SPIRAL_R2) return I_SPIRAL29;
        if (c.policy == P_WY_U8) return I_WY8_K9R2;
        if (c.policy == P_WY_U16) return I_WY16_K9R2;
    } else if (same_code<CodeIS95>(c)) {
        if (c.policy == P_WY_U8) return I_WY8_IS95;
        if (c.policy == P_WY_U16) return I_WY16_IS95;
    } else if (same_code<CodeK9R4>(c)) {
        if (c.policy == P_SPIRAL_R4) return I_SPIRAL49;
        if (c.policy == P_WY_U8) return I_WY8_K9R4;
        if (c.policy == P_WY_U16) return I_WY16_K9R4;
    }
    return I_NONE;
}

}  // namespace grp

#if !GRP_PUNCT_TU
bool grp_supported(const CodecParams& c) {
    if (grp::find_instance(c) < 0) return false;
    // wy-u8: the clamp arithmetic assumes the reference's u8 ranges
    if (c.policy == P_WY_U8 && (c.max_error > 255 || c.renorm_threshold > 255 || c.init_other > 255 ||
                                c.init_start > 255 || c.max_error < 0 || c.init_other < 0 || c.init_start < 0))
        return false;
    // wy-u16: the 16-bit lane window must hold rebase threshold + metric spread + two (capped) branch metrics
    if (c.policy == P_WY_U16) {
        if (c.max_error <= 0 || c.init_start < 0 || c.init_other < 0 || c.renorm_threshold < 0 ||
            c.init_start > 65535 || c.init_other > 65535 || c.renorm_threshold > 65535) return false;
        const long long cap = 2LL * c.max_error;
        long long spread = (long long)(c.K - 1) * cap;
        const long long init_spread = (long long)c.init_other - c.init_start;
        if (init_spread > spread) spread = init_spread;
        if (-init_spread > spread) spread = -init_spread;
        if (4096 + spread + 2 * cap > 64000) return false;
    }
    return true;
}

int grp_threads_per_frame(const CodecParams& c) { return 1 << (c.K - 1 - grp::NL - 1); }

cudaError_t launch_acs_grp_punct(const CodecParams& c, const void* syms, size_t sym_stride, int nsteps, int step0,
                                 unsigned long long* dec, uint8_t* row_phase, uint16_t* metrics, int n_frames,
                                 int frame0, int count, cudaStream_t st, const int* pidx, int fill, unsigned long long* err);
#endif

#if GRP_PUNCT_TU
cudaError_t launch_acs_grp_punct(
#else
cudaError_t launch_acs_grp(
#endif
                           const CodecParams& c, const void* syms, size_t sym_stride, int nsteps, int step0,
                           unsigned long long* dec, uint8_t* row_phase, uint16_t* metrics, int n_frames,
                           int frame0, int count, cudaStream_t st, const int* pidx, int fill, unsigned long long* err) {
    if (count <= 0 || nsteps <= 0) return cudaSuccess;
#if !GRP_PUNCT_TU
    if (pidx) return launch_acs_grp_punct(c, syms, sym_stride, nsteps, step0, dec, row_phase, metrics, n_frames, frame0, count,
                                          st, pidx, fill, err);
#endif
    grp::GrpArgs a;
    a.pidx = pidx; a.fill = fill; a.err = err;
    a.syms = (const unsigned char*)syms; a.sym_stride = sym_stride; a.nsteps = nsteps; a.step0 = step0;
    a.dec = dec; a.row_phase = row_phase; a.metrics = metrics; a.n_frames = n_frames;
    a.frame0 = frame0; a.count = count;
    a.C = c.max_error; a.thr = c.renorm_threshold; a.soft_hi = c.soft_hi; a.soft_lo = c.soft_lo; a.one = 1; a.two = 2;
    switch (grp::find_instance(c)) {
    case grp::I_WY8_IS95: return grp::launch_one<P_WY_U8, grp::CodeIS95>(a, st);
    case grp::I_WY16_IS95: return grp::launch_one<P_WY_U16, grp::CodeIS95>(a, st);
    case grp::I_KA9Q29: return grp::launch_one<P_KA9Q_U8_MOD, grp::CodeK9R2>(a, st);
    case grp::I_KA9Q615: return grp::launch_one<P_KA9Q_S16_615, grp::CodeK15R6>(a, st);
    case grp::I_SPIRAL615: return grp::launch_one<P_SPIRAL_615, grp::CodeK15R6>(a, st);
    case grp::I_WY8_K15R6: return grp::launch_one<P_WY_U8, grp::CodeK15R6>(a, st);
    case grp::I_SPIRAL29: return grp::launch_one<P_SPIRAL_R2, grp::CodeK9R2>(a, st);
    case grp::I_WY8_K9R2: return grp::launch_one<P_WY_U8, grp::CodeK9R2>(a, st);
    case grp::I_SPIRAL49: return grp::launch_one<P_SPIRAL_R4, grp::CodeK9R4>(a, st);
    case grp::I_WY8_K9R4: return grp::launch_one<P_WY_U8, grp::CodeK9R4>(a, st);
    case grp::I_WY16_K15R6: return grp::launch_one<P_WY_U16, grp::CodeK15R6>(a, st);
    case grp::I_WY16_K9R2: return grp::launch_one<P_WY_U16, grp::CodeK9R2>(a, st);
    case grp::I_WY16_K9R4: return grp::launch_one<P_WY_U16, grp::CodeK9R4>(a, st);
    default: return cudaErrorInvalidValue;
    }
}

}  // namespace b200v
