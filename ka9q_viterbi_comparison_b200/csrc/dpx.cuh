// dpx.cuh -- packed 16-bit helpers shared by the register-resident ACS kernels (acs_k7.cu, acs_grp.cu).
#pragma once
#include <stdint.h>

#ifndef DPX_VOLATILE
#define DPX_VOLATILE volatile
#endif

namespace b200v {

__device__ __forceinline__ uint32_t pack2(uint32_t lo, uint32_t hi) { return lo | (hi << 16); }

// s = min(a, b) per u16 lane (VIMNMX.U16x2 with predicate outputs); decision of a lane = (a <= b), or its
// negation when NEG; a set decision ORs the lane's bit into acc with one predicated LOP3. Written as one PTX
// block so that ptxas keeps the predicates (the C++ form costs a clear + a set per decision).
template <int BIT_LO, int BIT_HI, bool NEG>
__device__ __forceinline__ uint32_t min_acc(uint32_t a, uint32_t b, uint32_t& acc) {
    uint32_t s;
    if (NEG) {
        asm DPX_VOLATILE("{\n\t.reg .pred pu, pv;\n\t.reg .u16 r0, r1, r2, r3;\n\t"
            "min.u16x2 %0, %2, %3;\n\t"
            "mov.b32 {r0, r1}, %0;\n\tmov.b32 {r2, r3}, %2;\n\t"
            "setp.eq.u16 pv, r0, r2;\n\tsetp.eq.u16 pu, r1, r3;\n\t"
            "@!pv or.b32 %1, %1, %4;\n\t@!pu or.b32 %1, %1, %5;\n\t}"
            : "=&r"(s), "+r"(acc) : "r"(a), "r"(b), "n"(1u << BIT_LO), "n"(1u << BIT_HI));
    } else {
        asm DPX_VOLATILE("{\n\t.reg .pred pu, pv;\n\t.reg .u16 r0, r1, r2, r3;\n\t"
            "min.u16x2 %0, %2, %3;\n\t"
            "mov.b32 {r0, r1}, %0;\n\tmov.b32 {r2, r3}, %2;\n\t"
            "setp.eq.u16 pv, r0, r2;\n\tsetp.eq.u16 pu, r1, r3;\n\t"
            "@pv or.b32 %1, %1, %4;\n\t@pu or.b32 %1, %1, %5;\n\t}"
            : "=&r"(s), "+r"(acc) : "r"(a), "r"(b), "n"(1u << BIT_LO), "n"(1u << BIT_HI));
    }
    return s;
}

// Same compare-select, but the two decisions are accumulated on the FMA pipe: a set decision adds 2^BIT to an
// fp32 accumulator that starts at 2^23 (exact: at most 16 distinct powers of two below 2^16), leaving the ALU
// pipe to the packed min/add instructions. fdec_pack() extracts the two 16-bit fields.
template <int BIT>
__device__ __forceinline__ uint32_t min_facc(uint32_t a, uint32_t b, float& acc_lo, float& acc_hi) {
    uint32_t s;
    asm DPX_VOLATILE("{\n\t.reg .pred pu, pv;\n\t.reg .u16 r0, r1, r2, r3;\n\t"
        "min.u16x2 %0, %3, %4;\n\t"
        "mov.b32 {r0, r1}, %0;\n\tmov.b32 {r2, r3}, %3;\n\t"
        "setp.eq.u16 pv, r0, r2;\n\tsetp.eq.u16 pu, r1, r3;\n\t"
        "@pv add.rn.f32 %1, %1, %5;\n\t@pu add.rn.f32 %2, %2, %5;\n\t}"
        : "=&r"(s), "+f"(acc_lo), "+f"(acc_hi) : "r"(a), "r"(b), "f"((float)(1u << BIT)));
    return s;
}
// min(a, b) per u16 lane with the opposite tie-break: decision = (min != a), i.e. b survived strictly (tie -> 0)
template <int BIT>
__device__ __forceinline__ uint32_t min_facc_strict(uint32_t a, uint32_t b, float& acc_lo, float& acc_hi) {
    uint32_t s;
    asm DPX_VOLATILE("{\n\t.reg .pred pu, pv;\n\t.reg .u16 r0, r1, r2, r3;\n\t"
        "min.u16x2 %0, %3, %4;\n\t"
        "mov.b32 {r0, r1}, %0;\n\tmov.b32 {r2, r3}, %3;\n\t"
        "setp.eq.u16 pv, r0, r2;\n\tsetp.eq.u16 pu, r1, r3;\n\t"
        "@!pv add.rn.f32 %1, %1, %5;\n\t@!pu add.rn.f32 %2, %2, %5;\n\t}"
        : "=&r"(s), "+f"(acc_lo), "+f"(acc_hi) : "r"(a), "r"(b), "f"((float)(1u << BIT)));
    return s;
}
// z = min(t, 0) per s16 lane; decision of a lane = (t > 0), accumulated like min_facc
template <int BIT>
__device__ __forceinline__ uint32_t relu_neg_facc(uint32_t t, float& acc_lo, float& acc_hi) {
    uint32_t z;
    asm DPX_VOLATILE("{\n\t.reg .pred pu, pv;\n\t.reg .s16 r0, r1, r2, r3;\n\t"
        "min.s16x2 %0, %3, %4;\n\t"
        "mov.b32 {r0, r1}, %0;\n\tmov.b32 {r2, r3}, %3;\n\t"
        "setp.eq.s16 pv, r0, r2;\n\tsetp.eq.s16 pu, r1, r3;\n\t"
        "@!pv add.rn.f32 %1, %1, %5;\n\t@!pu add.rn.f32 %2, %2, %5;\n\t}"
        : "=&r"(z), "+f"(acc_lo), "+f"(acc_hi) : "r"(t), "r"(0u), "f"((float)(1u << BIT)));
    return z;
}
constexpr float FDEC_ZERO = 8388608.0f;   // 2^23
__device__ __forceinline__ uint32_t fdec_pack(float acc_lo, float acc_hi) {
    return __byte_perm(__float_as_uint(acc_lo), __float_as_uint(acc_hi), 0x5410);
}
// a + b on the FMA pipe (IMAD with a multiplier ptxas cannot fold); packed lanes must not overflow
__device__ __forceinline__ uint32_t add_fma(uint32_t a, uint32_t one, uint32_t b) {
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(one), "r"(b));
    return d;
}
// PRMT in its generic mode: selector nibbles with bit 3 set replicate the sign bit of the selected byte
// (__byte_perm masks that bit away). sign_bytes(x0, x1) = {msb(x0.b1), msb(x0.b3), msb(x1.b1), msb(x1.b3)} as
// 0x00 / 0xff bytes: the four bit-15 / bit-31 decisions of two packed differences.
__device__ __forceinline__ uint32_t sign_bytes(uint32_t x0, uint32_t x1) {
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, 0xfdb9;" : "=r"(d) : "r"(x0), "r"(x1));
    return d;
}
// ---- mask decisions (round 2): compare-select fused into one VIADDMNMX, decision read back as a lane mask --------------
// n = min(hi + Wc, t) is ONE VIADDMNMX.U16x2 when t (the lo-path candidate) is its third operand; the decision of a lane
// is then (n != t) [tie -> 0 policies] or (n == hi-path candidate) [tie -> 1 policies]. HSET2 compares the two u16 lanes
// of its operands as f16 values and writes 0xffff / 0x0000 per lane: for lane values in [0, 0x7c00] (every bit pattern
// there is a distinct, non-NaN f16 -- subnormals included, the instruction does not flush them) float equality IS
// integer equality. The kernels keep their lanes inside that window (see each kernel's window note).
__device__ __forceinline__ uint32_t ne_mask_u16x2(uint32_t a, uint32_t b) {
    uint32_t d;
    asm("set.ne.u32.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
__device__ __forceinline__ uint32_t eq_mask_u16x2(uint32_t a, uint32_t b) {
    uint32_t d;
    asm("set.eq.u32.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
// Sixteen lane masks m_15 ... m_0 (in that order) are folded into one word with h = 2 h + m on the FMA pipe (IMAD; `two`
// is a register holding 2 that ptxas cannot fold). Per lane 0xffff = -1 (mod 2^16), so h = -(DL + 2^16 (DH - DL)) with
// DL / DH the 16-bit words of lane-0 / lane-1 decisions (bit k = mask k); one multiply undoes the lane interplay and the
// sign:  DL | DH << 16 = h * -(2^16 + 1)  (mod 2^32).
__device__ __forceinline__ uint32_t horner_mask(uint32_t h, uint32_t two, uint32_t m) {
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(h), "r"(two), "r"(m));
    return d;
}
__device__ __forceinline__ uint32_t horner_decode(uint32_t h) { return h * 0xFFFEFFFFu; }

// a + b on the ALU pipe: a three-input add with an opaque zero stays an IADD3 (ptxas turns two-input adds into
// VIADD / IMAD.IADD, both of which ncu attributes to the FMA-heavy pipe on B200)
__device__ __forceinline__ uint32_t add_alu(uint32_t a, uint32_t b, uint32_t zero) {
    uint32_t d;
    asm("{\n\t.reg .u32 t;\n\tadd.u32 t, %1, %2;\n\tadd.u32 %0, t, %3;\n\t}" : "=r"(d) : "r"(a), "r"(b), "r"(zero));
    return d;
}

}  // namespace b200v
