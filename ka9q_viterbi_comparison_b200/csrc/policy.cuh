// policy.cuh -- per-policy scalar arithmetic (int32 domain) used by the generic kernel.
// Each function reproduces one row of the policy table in SURVEY.md 8(a); the packed (u16x2 / DPX)
// forms used by the register-resident kernels live beside those kernels.
#pragma once
#include "codec.h"

namespace b200v {

template <int POLICY> struct PolicyTraits {
    static constexpr bool is_signed16 = (POLICY == P_KA9Q_S16_615 || POLICY == P_KA9Q_S16_224);
    static constexpr bool is_wy = (POLICY == P_WY_U8 || POLICY == P_WY_U16);
    static constexpr bool has_renorm = (POLICY != P_KA9Q_U8_MOD);
    // renormalisation compares new[0] with ">" (spiral*.cpp) or ">=" (everything else)
    static constexpr bool renorm_strict = (POLICY == P_SPIRAL_R2 || POLICY == P_SPIRAL_R4 || POLICY == P_SPIRAL_615);
};

__device__ __forceinline__ int clamp_i(int x, int lo, int hi) { return min(max(x, lo), hi); }

// Load symbol i of the current step as an integer (u8 for ka9q/spiral, int8 / int16 for wy).
template <int POLICY>
__device__ __forceinline__ int load_symbol(const void* p, int i) {
    if (POLICY == P_WY_U8) return (int)((const int8_t*)p)[i];
    if (POLICY == P_WY_U16) return (int)((const int16_t*)p)[i];
    return (int)((const uint8_t*)p)[i];
}

// Branch metric of a butterfly whose expected code word is cw (bit i = parity((2j) & poly_i)).
//   ka9q-u8-mod : viterbi27_sse2.cpp:137-146   avg_epu8(x0,x1) >> 4
//   ka9q-s16    : viterbi615_sse2.cpp:132-135, viterbi224_sse2.cpp:159   sum of x_i
//   spiral r=1/2: spiral27.cpp:164-170          (avg_epu8(x0,x1) >> 2) & 63
//   spiral r=1/4, 615: spiral47.cpp:143-181, spiral615.cpp:157-238  ((sat255 sum (x_i>>2)) >> 2) & 63
//   wy u8 / u16 : sse_u8.h:92-97, sse_u16.h:93-97   sat sum |sat(bt_i - s_i)|
template <int POLICY>
__device__ __forceinline__ int branch_metric(const CodecParams& c, unsigned cw, const int* sym) {
    if (POLICY == P_WY_U8 || POLICY == P_WY_U16) {
        int tot = 0;
        for (int i = 0; i < c.R; i++) {
            const int bt = ((cw >> i) & 1u) ? c.soft_hi : c.soft_lo;
            int e = bt - sym[i];
            if (POLICY == P_WY_U8) { e = clamp_i(e, -128, 127); e = abs(e); tot = min(tot + e, 255); }
            else { e = clamp_i(e, -32768, 32767); e = abs(e); tot = min(tot + e, 65535); }
        }
        return tot;
    }
    int s = 0;
    if (POLICY == P_KA9Q_U8_MOD || POLICY == P_SPIRAL_R2) {
        const int x0 = (cw & 1u) ? (sym[0] ^ 255) : sym[0];
        const int x1 = (cw & 2u) ? (sym[1] ^ 255) : sym[1];
        const int avg = (x0 + x1 + 1) >> 1;
        return POLICY == P_KA9Q_U8_MOD ? (avg >> 4) : ((avg >> 2) & 63);
    }
    for (int i = 0; i < c.R; i++) {
        const int x = ((cw >> i) & 1u) ? (sym[i] ^ 255) : sym[i];
        if (POLICY == P_KA9Q_S16_615 || POLICY == P_KA9Q_S16_224) s += x;
        else s = min(s + ((x >> 2) & 63), 255);
    }
    if (POLICY == P_KA9Q_S16_615 || POLICY == P_KA9Q_S16_224) return s;
    return (s >> 2) & 63;
}

// Add-compare-select of one butterfly. lo = old[j], hi = old[j+S/2], bm = branch metric, C = c.max_error.
// Outputs survivors for new states 2j, 2j+1 and their decision bits (1 = the hi predecessor survived).
template <int POLICY>
__device__ __forceinline__ void acs(int C, int lo, int hi, int bm, int& s0, int& s1, int& d0, int& d1) {
    if (POLICY == P_KA9Q_U8_MOD) {
        // viterbi27_sse2.cpp:148-158: wrapping u8 adds; decision = (int8)(a-b) > 0; tie -> 0
        const int cm = 15 - bm;
        const int m0 = (lo + bm) & 255, m1 = (hi + cm) & 255, m2 = (lo + cm) & 255, m3 = (hi + bm) & 255;
        d0 = ((int)(int8_t)(m0 - m1)) > 0;
        d1 = ((int)(int8_t)(m2 - m3)) > 0;
        s0 = d0 ? m1 : m0;
        s1 = d1 ? m3 : m2;
    } else if (POLICY == P_KA9Q_S16_615 || POLICY == P_KA9Q_S16_224) {
        // viterbi615_sse2.cpp:136-148 (min, cmpeq: tie -> 1); viterbi224_sse2.cpp:160-194 (cmpgt: tie -> 0)
        const int cm = C - bm;
        const int m0 = clamp_i(lo + bm, -32768, 32767), m1 = clamp_i(hi + cm, -32768, 32767);
        const int m2 = clamp_i(lo + cm, -32768, 32767), m3 = clamp_i(hi + bm, -32768, 32767);
        s0 = min(m0, m1);
        s1 = min(m2, m3);
        if (POLICY == P_KA9Q_S16_615) { d0 = (s0 == m1); d1 = (s1 == m3); }
        else { d0 = (m0 > m1); d1 = (m2 > m3); }
    } else {
        // spiral27.cpp:171-181, spiral615.cpp:240-250, sse_u8.h:105-114, sse_u16.h:105-115:
        // complement = subs(C, BM), saturating adds, min, decision = (min == hi path): tie -> 1
        const int top = (POLICY == P_WY_U16) ? 65535 : 255;
        const int cm = max(C - bm, 0);
        const int m0 = min(lo + bm, top), m1 = min(hi + cm, top), m2 = min(lo + cm, top), m3 = min(hi + bm, top);
        s0 = min(m0, m1);
        s1 = min(m2, m3);
        d0 = (s0 == m1);
        d1 = (s1 == m3);
    }
}

// Renormalisation trigger on new[0] (checked after every step).
template <int POLICY>
__device__ __forceinline__ bool renorm_needed(const CodecParams& c, int m0) {
    if (!PolicyTraits<POLICY>::has_renorm) return false;
    return PolicyTraits<POLICY>::renorm_strict ? (m0 > c.renorm_threshold) : (m0 >= c.renorm_threshold);
}

// Apply renormalisation to one metric given the minimum over all states.
//   s16: subtract (min - SHRT_MIN) with wrapping sub_epi16 (viterbi615_sse2.cpp:166-183)
//   u8/u16: subtract min (spiral27.cpp:236-254, sse_u8.h:127-154)
template <int POLICY>
__device__ __forceinline__ int renorm_apply(int m, int mn) {
    if (PolicyTraits<POLICY>::is_signed16) return (int)(int16_t)(uint16_t)(m - (mn + 32768));
    return m - mn;
}

// What one renormalisation subtracts from every metric, given the minimum: the value the reference's update() sums up
// and returns (sse_u8.h:57-69: total_error += renormalise(new_metrics)).
template <int POLICY>
__device__ __forceinline__ unsigned renorm_amount(int mn) {
    return (unsigned)(PolicyTraits<POLICY>::is_signed16 ? mn + 32768 : mn);
}

}  // namespace b200v
