// acs_smallk.cu -- register-resident ACS for K = 3 ... 6 (S = 4 ... 32 states): one thread per frame, all path
// metrics of the frame in registers, one 32-bit decision word per step (bit n = new state n, the canonical
// vector itself), stored frame-interleaved so that a warp writes 128 contiguous bytes per step. Serves the short
// codes of the williamyang library's list (examples/helpers/common_codes.h:23-24: K=3 and K=5 r=1/2), any
// polynomial set, williamyang u8 / u16 policies (sse_u8.h:45-154, sse_u16.h:45-177 through policy.cuh's scalar
// restatement: saturating adds, min, decision = (min == hi path), renormalisation on new[0]).
// The trellis is fully unrolled with compile-time state indices; the code word of a butterfly is a run-time value
// (polynomials are per instance), so its branch metric is evaluated from the step's symbols directly.
#include "kernels.h"
#include "policy.cuh"

namespace b200v {

namespace {

// RT = compile-time rate (symbols per step) or 0: run-time rate (symbols indexed in local memory, slower)
template <int POLICY, int K, int RT>
__global__ void __launch_bounds__(128) acs_smallk_kernel(const GenericArgs a) {
    constexpr int S = 1 << (K - 1), H = S / 2;
    constexpr int TOP = (POLICY == P_WY_U16) ? 65535 : 255;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= a.count) return;
    const int frame = a.frame0 + idx;
    const CodecParams& c = a.c;
    const int R = RT ? RT : c.R;
    uint16_t* gm = a.metrics + (size_t)frame * S;
    int m[S];
#pragma unroll
    for (int i = 0; i < S; i++) m[i] = gm[i];
    unsigned cw[H];
#pragma unroll
    for (int j = 0; j < H; j++) cw[j] = a.cw[j];
    const unsigned char* sp = (const unsigned char*)a.syms + (size_t)frame * a.sym_stride;
    const int step_bytes = R * c.sym_bytes;
    uint32_t* drow = a.dec + (size_t)a.step0 * a.n_frames + frame;
    const int lim = (POLICY == P_WY_U8) ? 127 : 32767;
    unsigned long long rsum = 0;     // minima subtracted by this launch's renormalisations
#pragma unroll 1
    for (int t = 0; t < a.nsteps; t++) {
        const void* p = sp + (size_t)t * step_bytes;
        int n[S];
        uint32_t dec = 0;
        if (RT) {
            // per-symbol costs once per step: |sat(bt - s)| for bt = soft_lo / soft_hi (sse_u8.h:92-97); all costs are
            // >= 0, so the reference's running saturating sum equals one saturation of the plain sum
            int v0[RT ? RT : 1], v1[RT ? RT : 1];
#pragma unroll
            for (int i = 0; i < RT; i++) {
                const int s = load_symbol<POLICY>(p, i);
                v0[i] = abs(clamp_i(c.soft_lo - s, -lim - 1, lim));
                v1[i] = abs(clamp_i(c.soft_hi - s, -lim - 1, lim));
                if (POLICY == P_WY_U8) { v0[i] = min(v0[i], 255); v1[i] = min(v1[i], 255); }
            }
#pragma unroll
            for (int j = 0; j < H; j++) {
                int bm = 0;
#pragma unroll
                for (int i = 0; i < RT; i++) bm += ((cw[j] >> i) & 1u) ? v1[i] : v0[i];
                bm = min(bm, TOP);
                int d0, d1;
                acs<POLICY>(c.max_error, m[j], m[j + H], bm, n[2 * j], n[2 * j + 1], d0, d1);
                dec |= ((uint32_t)d0 << (2 * j)) | ((uint32_t)d1 << (2 * j + 1));
            }
        } else {
            int sym[MAX_R];
            for (int i = 0; i < R; i++) sym[i] = load_symbol<POLICY>(p, i);
#pragma unroll
            for (int j = 0; j < H; j++) {
                const int bm = branch_metric<POLICY>(c, cw[j], sym);
                int d0, d1;
                acs<POLICY>(c.max_error, m[j], m[j + H], bm, n[2 * j], n[2 * j + 1], d0, d1);
                dec |= ((uint32_t)d0 << (2 * j)) | ((uint32_t)d1 << (2 * j + 1));
            }
        }
        drow[(size_t)t * a.n_frames] = dec;
        if (renorm_needed<POLICY>(c, n[0])) {
            int mn = n[0];
#pragma unroll
            for (int i = 1; i < S; i++) mn = min(mn, n[i]);
            rsum += renorm_amount<POLICY>(mn);
#pragma unroll
            for (int i = 0; i < S; i++) n[i] = renorm_apply<POLICY>(n[i], mn);
        }
#pragma unroll
        for (int i = 0; i < S; i++) m[i] = n[i];
    }
#pragma unroll
    for (int i = 0; i < S; i++) gm[i] = (uint16_t)m[i];
    if (a.err && rsum) a.err[frame] += rsum;
}

template <int POLICY, int K>
cudaError_t launch_rate(const GenericArgs& a, int blocks, int threads, cudaStream_t st) {
    switch (a.c.R) {
    case 2: acs_smallk_kernel<POLICY, K, 2><<<blocks, threads, 0, st>>>(a); break;
    case 3: acs_smallk_kernel<POLICY, K, 3><<<blocks, threads, 0, st>>>(a); break;
    case 4: acs_smallk_kernel<POLICY, K, 4><<<blocks, threads, 0, st>>>(a); break;
    default: acs_smallk_kernel<POLICY, K, 0><<<blocks, threads, 0, st>>>(a); break;
    }
    return cudaGetLastError();
}

template <int POLICY>
cudaError_t launch_policy(const GenericArgs& a, cudaStream_t st) {
    const int threads = 128;
    const int blocks = (a.count + threads - 1) / threads;
    switch (a.c.K) {
    case 3: return launch_rate<POLICY, 3>(a, blocks, threads, st);
    case 4: return launch_rate<POLICY, 4>(a, blocks, threads, st);
    case 5: return launch_rate<POLICY, 5>(a, blocks, threads, st);
    case 6: return launch_rate<POLICY, 6>(a, blocks, threads, st);
    default: return cudaErrorInvalidValue;
    }
}

}  // namespace

bool smallk_supported(const CodecParams& c) {
    return c.K >= 3 && c.K <= 6 && c.R <= MAX_R && (c.policy == P_WY_U8 || c.policy == P_WY_U16);
}

// decisions: dec[(step0 + t) * n_frames + frame] (u32), layout DL_SMALLK_INTERLEAVED
cudaError_t launch_acs_smallk(const GenericArgs& a, cudaStream_t st) {
    if (a.count <= 0 || a.nsteps <= 0) return cudaSuccess;
    return a.c.policy == P_WY_U8 ? launch_policy<P_WY_U8>(a, st) : launch_policy<P_WY_U16>(a, st);
}

}  // namespace b200v
