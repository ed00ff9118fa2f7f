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

template <int POLICY>
__global__ void __launch_bounds__(128) acs_k7_warp_kernel(const GenericArgs a) {
    const int lane = threadIdx.x & 31;
    const int fidx = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (fidx >= a.count) return;                                   // whole warps leave together
    const int frame = a.frame0 + fidx;
    const CodecParams& c = a.c;
    const int R = c.R;
    uint16_t* gm = a.metrics + (size_t)frame * 64;
    int lo = gm[lane], hi = gm[lane + 32];
    if (PolicyTraits<POLICY>::is_signed16) { lo = (int16_t)lo; hi = (int16_t)hi; }
    const unsigned cw = a.cw[lane];                                // code word of butterfly `lane` (host table)
    const unsigned char* sp = (const unsigned char*)a.syms + (size_t)frame * a.sym_stride;
    const int step_bytes = R * c.sym_bytes;
    uint32_t* drow = a.dec + ((size_t)frame * a.dec_rows + a.step0) * 2;
    const int src0 = lane >> 1, src1 = 16 + (lane >> 1);
    const bool odd = lane & 1;
    for (int t = 0; t < a.nsteps; t++) {
        int sym[MAX_R];
        const void* p = sp + (size_t)t * step_bytes;
        for (int i = 0; i < R; i++) sym[i] = load_symbol<POLICY>(p, i);
        const int bm = branch_metric<POLICY>(c, cw, sym);
        int v0, v1, d0, d1;
        acs<POLICY>(c.max_error, lo, hi, bm, v0, v1, d0, d1);     // new[2j], new[2j+1]
        const unsigned w0 = __ballot_sync(0xffffffffu, d0), w1 = __ballot_sync(0xffffffffu, d1);
        const int a0 = __shfl_sync(0xffffffffu, v0, src0), a1 = __shfl_sync(0xffffffffu, v1, src0);
        const int c0 = __shfl_sync(0xffffffffu, v0, src1), c1 = __shfl_sync(0xffffffffu, v1, src1);
        lo = odd ? a1 : a0;
        hi = odd ? c1 : c0;
        if (lane == 0) *reinterpret_cast<uint2*>(drow + 2 * (size_t)t) = make_uint2(w0, w1);
        if (PolicyTraits<POLICY>::has_renorm) {
            const int m0 = __shfl_sync(0xffffffffu, lo, 0);
            if (renorm_needed<POLICY>(c, m0)) {                     // uniform over the warp
                const int mn = __reduce_min_sync(0xffffffffu, min(lo, hi));
                lo = renorm_apply<POLICY>(lo, mn);
                hi = renorm_apply<POLICY>(hi, mn);
            }
        }
    }
    gm[lane] = (uint16_t)lo;
    gm[lane + 32] = (uint16_t)hi;
}

template <int POLICY>
cudaError_t launch_policy(const GenericArgs& a, cudaStream_t st) {
    const int warps_per_block = 4;
    const int blocks = (a.count + warps_per_block - 1) / warps_per_block;
    acs_k7_warp_kernel<POLICY><<<blocks, warps_per_block * 32, 0, st>>>(a);
    return cudaGetLastError();
}

}  // namespace

bool k7_warp_supported(const CodecParams& c) {
    return c.K == 7 && c.R <= MAX_R && c.policy != P_KA9Q_S16_615 && c.policy != P_KA9Q_S16_224 && c.policy != P_SPIRAL_615;
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
