// dpx.cuh -- packed 16-bit helpers shared by the register-resident ACS kernels (acs_k7.cu, acs_grp.cu).
#pragma once
#include <stdint.h>

namespace b200v {

__device__ __forceinline__ uint32_t pack2(uint32_t lo, uint32_t hi) { return lo | (hi << 16); }

// s = min(a, b) per u16 lane (VIMNMX.U16x2 with predicate outputs); decision of a lane = (a <= b), or its
// negation when NEG; a set decision ORs the lane's bit into acc with one predicated LOP3. Written as one PTX
// block so that ptxas keeps the predicates (the C++ form costs a clear + a set per decision).
template <int BIT_LO, int BIT_HI, bool NEG>
__device__ __forceinline__ uint32_t min_acc(uint32_t a, uint32_t b, uint32_t& acc) {
    uint32_t s;
    if (NEG) {
        asm("{\n\t.reg .pred pu, pv;\n\t.reg .u16 r0, r1, r2, r3;\n\t"
            "min.u16x2 %0, %2, %3;\n\t"
            "mov.b32 {r0, r1}, %0;\n\tmov.b32 {r2, r3}, %2;\n\t"
            "setp.eq.u16 pv, r0, r2;\n\tsetp.eq.u16 pu, r1, r3;\n\t"
            "@!pv or.b32 %1, %1, %4;\n\t@!pu or.b32 %1, %1, %5;\n\t}"
            : "=&r"(s), "+r"(acc) : "r"(a), "r"(b), "n"(1u << BIT_LO), "n"(1u << BIT_HI));
    } else {
        asm("{\n\t.reg .pred pu, pv;\n\t.reg .u16 r0, r1, r2, r3;\n\t"
            "min.u16x2 %0, %2, %3;\n\t"
            "mov.b32 {r0, r1}, %0;\n\tmov.b32 {r2, r3}, %2;\n\t"
            "setp.eq.u16 pv, r0, r2;\n\tsetp.eq.u16 pu, r1, r3;\n\t"
            "@pv or.b32 %1, %1, %4;\n\t@pu or.b32 %1, %1, %5;\n\t}"
            : "=&r"(s), "+r"(acc) : "r"(a), "r"(b), "n"(1u << BIT_LO), "n"(1u << BIT_HI));
    }
    return s;
}

}  // namespace b200v
