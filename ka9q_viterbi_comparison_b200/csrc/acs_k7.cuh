// acs_k7.cuh -- index algebra of the register-resident K=7 kernels (shared by the kernel, the chainback
// reader and the host-side canonicaliser).
//
// One thread decodes one frame. The 64 path metrics live in 32 registers as u16x2 pairs and are updated IN
// PLACE: a butterfly reads two register lanes (old states j and j+32) and writes its two results (new states
// 2j and 2j+1) back to the same two lanes. The logical state held by a physical lane therefore rotates by one
// bit per step; after NB = K-1 = 6 steps the layout is back where it started, so the kernel body is six
// unrolled "phases" with compile-time register indices.
//
//   physical address q: bit 0 = lane within the register, bits 1..5 = register index
//   phase p: physical bit x holds logical state bit (x + p) mod 6
//   the logical MSB (the bit that distinguishes a butterfly's two inputs) sits at physical bit 5-p:
//     p < 5 : register-index bit 4-p  -> partner registers A (bit = 0) and B (bit = 1), two butterflies per pair
//     p = 5 : the lane bit            -> both inputs of a butterfly sit in one register
#pragma once
#include <stdint.h>

namespace b200v {
namespace k7 {

constexpr int NB = 6;           // state bits
constexpr int NW = 32;          // metric registers per frame
constexpr int S = 64;

// logical state stored at physical address q in phase p
__host__ __device__ constexpr int logical_of(int p, int q) {
    int n = 0;
    for (int x = 0; x < NB; x++)
        if ((q >> x) & 1) n |= 1 << ((x + p) % NB);
    return n;
}
// physical address of logical state n in phase p
__host__ __device__ constexpr int physical_of(int p, int n) {
    int q = 0;
    for (int b = 0; b < NB; b++)
        if ((n >> b) & 1) q |= 1 << (((b - p) % NB + NB) % NB);
    return q;
}
// insert a zero bit at position pos of k
__host__ __device__ constexpr int expand_bit(int k, int pos) {
    return ((k >> pos) << (pos + 1)) | (k & ((1 << pos) - 1));
}
// remove bit pos of i
__host__ __device__ constexpr int compress_bit(int i, int pos) {
    return ((i >> (pos + 1)) << pos) | (i & ((1 << pos) - 1));
}

// Bit position (0..63) inside the per-step decision word of the decision for NEW logical state n produced by
// a step executed in phase p. Decision word = acc0 | acc1 << 32. Two packings:
//  * bitwise (ka9q modular policy: each decision is a predicated OR of one constant bit): for p < 5, acc0 collects
//    the A registers (new even states) and acc1 the B registers (new odd states), k-th pair at bit k (lane 0) and
//    16+k (lane 1); for p = 5, acc0 collects registers 0..15 and acc1 registers 16..31.
//  * bytewise (saturating policies: four decisions at a time leave a PRMT as 0x00/0xff bytes and are masked into
//    the word with one LOP3): group g = pair k (p < 5) or register pair i >> 1 (p = 5) lands in word g >> 3 at
//    bit g & 7 of byte {A lane 0, A lane 1, B lane 0, B lane 1} (p < 5) / {even register lane 0, lane 1, odd
//    register lane 0, lane 1} (p = 5).
__host__ __device__ constexpr int decision_bit_of(int p, int n, bool bytewise) {
    const int q = physical_of((p + 1) % NB, n);   // layout after the step
    const int lane = q & 1, i = q >> 1;
    if (p < NB - 1) {
        const int xb = NB - 2 - p;
        const int is_b = (i >> xb) & 1, k = compress_bit(i, xb);
        if (bytewise) return (k >> 3) * 32 + (is_b * 2 + lane) * 8 + (k & 7);
        return is_b * 32 + lane * 16 + k;
    }
    if (bytewise) return (i >> 4) * 32 + ((i & 1) * 2 + lane) * 8 + ((i >> 1) & 7);
    return (i >> 4) * 32 + lane * 16 + (i & 15);
}
// which packing a policy's kernel writes (policy ids: codec.h)
__host__ __device__ constexpr bool decisions_bytewise(int policy) { return policy != 0 /* P_KA9Q_U8_MOD */; }

// code word of butterfly j: bit i = parity((2j) & poly_i)
__host__ __device__ constexpr int parity_of(unsigned x) {
    x ^= x >> 16; x ^= x >> 8; x ^= x >> 4; x ^= x >> 2; x ^= x >> 1;
    return (int)(x & 1u);
}

}  // namespace k7
}  // namespace b200v
