// acs_grp.cuh -- index algebra of the "group" ACS family (K = 9 ... 15): several threads per frame, 64 path
// metrics per thread in 32 registers of u16x2, five trellis steps between shared-memory exchanges.
// Shared by the kernel (acs_grp.cu), the chainback reader and the host-side canonicaliser (api.cu).
//
// State bits (NB = K-1) at the start of a group of five steps:   n = [ i : 5 bits ][ xx : NXB = NB-5 bits ]
//   i   -> register index (the five bits the next five steps shift out, so all butterflies are thread-local)
//   xx  -> bit 0 = lane inside the u16x2 register, bits 1.. = thread index tx inside the frame (TPF = 2^(NXB-1))
// After s steps the state held by (register q, xx) is  [ i low 5-s bits ][ xx ][ s new bits ]; the five "local"
// bits (i low bits, new bits) form a 5-bit logical index l that evolves like a K=6 trellis and is kept IN
// PLACE: phase p stores logical bit (x + p) mod 5 at physical register-index bit x. The state MSB sits at
// physical bit 4-p, so the butterfly partners of phase p are registers A (bit 4-p = 0) and B (bit = 1); both
// lanes of a register pair are two independent butterflies.
//
// Decisions of one step = one u64 per thread: bit kk (lane 0) / 16+kk (lane 1) for the kk-th A register (new
// even local index), +32 for the B register; stored frame-interleaved as dec[(row*n_frames + frame)*TPF + tx].
#pragma once
#include <stdint.h>
#include "acs_k7.cuh"   // expand_bit / compress_bit / parity_of

namespace b200v {
namespace grp {

constexpr int NL = 5;        // local state bits = steps per group
constexpr int NREG = 32;     // metric registers per thread

__host__ __device__ constexpr int logical_of(int p, int q) {
    int l = 0;
    for (int x = 0; x < NL; x++)
        if ((q >> x) & 1) l |= 1 << ((x + p) % NL);
    return l;
}
__host__ __device__ constexpr int physical_of(int p, int l) {
    int q = 0;
    for (int b = 0; b < NL; b++)
        if ((l >> b) & 1) q |= 1 << (((b - p) % NL + NL) % NL);
    return q;
}

// state held by register q / x-value xx BEFORE the step of phase p (p = 0..5; p = 5 is the layout after a
// whole group, equal to phase 0 of the next group once the threads have exchanged)
__host__ __device__ constexpr unsigned state_of(int nxb, int p, int q, unsigned xx) {
    const unsigned l = (unsigned)logical_of(p % NL, q);
    return ((l >> p) << (nxb + p)) | (xx << p) | (l & ((1u << p) - 1u));
}
// inverse: where state n lives in phase p
__host__ __device__ inline void locate_state(int nxb, int p, unsigned n, int& q, unsigned& xx) {
    const unsigned l = ((n >> (nxb + p)) << p) | (n & ((1u << p) - 1u));
    xx = (n >> p) & ((1u << nxb) - 1u);
    q = physical_of(p % NL, (int)l);
}
// decision of NEW state n in a row written by a step of phase p: which thread's u64, and which bit
__host__ __device__ inline void decision_pos(int nxb, int p, unsigned n, unsigned& tx, unsigned& bit) {
    int q;
    unsigned xx;
    locate_state(nxb, p + 1, n, q, xx);
    const int xb = NL - 1 - p;
    tx = xx >> 1;
    bit = (unsigned)(((q >> xb) & 1) * 32 + (int)(xx & 1u) * 16 + k7::compress_bit(q, xb));
}

}  // namespace grp
}  // namespace b200v
