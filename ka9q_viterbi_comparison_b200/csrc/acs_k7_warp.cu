// acs_k7_warp.cu -- the OTHER K=7 design BASELINE.json's configs[1] names: one WARP per frame, butterflies across
// lanes with __shfl_sync, decisions packed with __ballot_sync. Kept as kernel family 3 (ka9q modular policy, any
// polynomial pair) so that the choice of the thread-per-frame register kernel (acs_k7.cu) is backed by a
// measurement on the same inputs: a warp spends ~30 instructions per trellis step on ONE frame, the register
// kernel 253 on 32 frames (profiles/r01_k7_warp_vs_thread.json).
//
// Lane j holds old[j] and old[j+32] = both inputs of butterfly j (viterbi27_sse2.cpp:148-158: wrapping u8 adds,
// decision = (int8)(m0 - m1) > 0, tie -> 0). New state n lives in lane n & 31, slot n >> 5, so lane r takes
// new[r] from butterfly r >> 1 and new[r + 32] from butterfly 16 + (r >> 1). The two ballots of a step are exactly
// the generic family's decision row (DL_BALLOT_FRAME_MAJOR, row_words = 2), so chainback is shared.
#include "kernels.h"

namespace b200v {

namespace {

__global__ void __launch_bounds__(128) acs_k7_warp_kernel(const GenericArgs a) {
    const int lane = threadIdx.x & 31;
    const int fidx = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (fidx >= a.count) return;                                   // whole warps leave together
    const int frame = a.frame0 + fidx;
    uint16_t* gm = a.metrics + (size_t)frame * 64;
    unsigned lo = gm[lane] & 0xffu, hi = gm[lane + 32] & 0xffu;
    // code word of butterfly `lane`: bit i = parity((2 * lane) & poly_i)
    const unsigned b0 = __popc((2u * lane) & (unsigned)a.c.poly[0]) & 1u ? 255u : 0u;
    const unsigned b1 = __popc((2u * lane) & (unsigned)a.c.poly[1]) & 1u ? 255u : 0u;
    const unsigned char* sp = (const unsigned char*)a.syms + (size_t)frame * a.sym_stride;
    uint32_t* drow = a.dec + ((size_t)frame * a.dec_rows + a.step0) * 2;
    const int src0 = lane >> 1, src1 = 16 + (lane >> 1);
    const bool odd = lane & 1;
    for (int t = 0; t < a.nsteps; t++) {
        const unsigned s0 = __ldg(sp + 2 * t), s1 = __ldg(sp + 2 * t + 1);
        const unsigned bm = ((((s0 ^ b0) + (s1 ^ b1) + 1u) >> 1) >> 4), cm = 15u - bm;
        const unsigned m0 = (lo + bm) & 255u, m1 = (hi + cm) & 255u, m2 = (lo + cm) & 255u, m3 = (hi + bm) & 255u;
        const bool d0 = (int)(int8_t)(m0 - m1) > 0, d1 = (int)(int8_t)(m2 - m3) > 0;
        const unsigned v0 = d0 ? m1 : m0, v1 = d1 ? m3 : m2;       // new[2j], new[2j+1]
        const unsigned w0 = __ballot_sync(0xffffffffu, d0), w1 = __ballot_sync(0xffffffffu, d1);
        const unsigned a0 = __shfl_sync(0xffffffffu, v0, src0), a1 = __shfl_sync(0xffffffffu, v1, src0);
        const unsigned c0 = __shfl_sync(0xffffffffu, v0, src1), c1 = __shfl_sync(0xffffffffu, v1, src1);
        lo = odd ? a1 : a0;
        hi = odd ? c1 : c0;
        if (lane == 0) *reinterpret_cast<uint2*>(drow + 2 * (size_t)t) = make_uint2(w0, w1);
    }
    gm[lane] = (uint16_t)lo;
    gm[lane + 32] = (uint16_t)hi;
}

}  // namespace

bool k7_warp_supported(const CodecParams& c) { return c.K == 7 && c.R == 2 && c.policy == P_KA9Q_U8_MOD; }

cudaError_t launch_acs_k7_warp(const GenericArgs& a, cudaStream_t st) {
    if (a.count <= 0 || a.nsteps <= 0) return cudaSuccess;
    const int warps_per_block = 4;
    const int blocks = (a.count + warps_per_block - 1) / warps_per_block;
    acs_k7_warp_kernel<<<blocks, warps_per_block * 32, 0, st>>>(a);
    return cudaGetLastError();
}

}  // namespace b200v
