// acs_generic.cu -- generic add-compare-select kernel: any K <= 16, any R <= 8, any polynomial set, all
// arithmetic policies. One frame per group of `tpf` threads; path metrics ping-pong in shared memory as
// 16-bit words; the 2^R branch metrics of a step are computed once into shared memory and selected per
// butterfly through a code-word table; decisions are packed with __ballot_sync.
//
// This family is (a) the fallback for polynomials the specialised kernels are not instantiated for and
// (b) the checker the specialised kernels are tested against on the GPU. It takes the role of
// update_viterbiXX_blk_sse2 (viterbi27_sse2.cpp:119-175), FULL_SPIRAL (spiral27.cpp:130-381) and
// ViterbiDecoder_SSE_u8::update (sse_u8.h:45-70) for a whole batch of frames.
#include "kernels.h"
#include "policy.cuh"

namespace b200v {

namespace {

template <bool WARP_FRAME>
__device__ __forceinline__ void frame_sync() {
    if (WARP_FRAME) __syncwarp(); else __syncthreads();
}

// min over the `n` 16-bit metrics of one frame. All tpf threads of the frame call this.
template <int POLICY, bool WARP_FRAME>
__device__ __forceinline__ int frame_min(const uint16_t* m, int n, int lt, int tpf, int* scratch) {
    int mn = PolicyTraits<POLICY>::is_signed16 ? 32767 : 65535;
    for (int i = lt; i < n; i += tpf) {
        const int v = PolicyTraits<POLICY>::is_signed16 ? (int)(int16_t)m[i] : (int)m[i];
        mn = min(mn, v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    if (WARP_FRAME) return mn;
    const int warp = lt >> 5, nwarps = tpf >> 5;
    if ((lt & 31) == 0) scratch[warp] = mn;
    __syncthreads();
    int r = scratch[0];
    for (int w = 1; w < nwarps; w++) r = min(r, scratch[w]);
    __syncthreads();   // scratch may be reused by the next reduction
    return r;
}

template <int POLICY, bool WARP_FRAME>
__global__ void acs_generic_kernel(const GenericArgs a, const int tpf, const int fpb) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const CodecParams& c = a.c;
    const int S = c.S, H = S >> 1, R = c.R, nbm = 1 << R;
    const int fl = threadIdx.x / tpf, lt = threadIdx.x - fl * tpf;
    const int fidx = blockIdx.x * fpb + fl;
    const bool live = fidx < a.count;
    const int frame = a.frame0 + (live ? fidx : a.count - 1);

    // ---- shared memory carve-up ----
    // [cw: H bytes, padded to 16] then per frame slot: met[2][S] u16 | bm[2][nbm] int | decrow[2][row_words] u32 | scratch[32]
    uint8_t* s_cw = smem_raw;
    const int cw_bytes = (H + 15) & ~15;
    const int slot_bytes = 2 * S * 2 + 2 * nbm * 4 + 2 * a.row_words * 4 + 32 * 4;
    unsigned char* slot = smem_raw + cw_bytes + (size_t)fl * slot_bytes;
    uint16_t* s_met = (uint16_t*)slot;
    int* s_bm = (int*)(slot + 2 * S * 2);
    uint32_t* s_dec = (uint32_t*)(slot + 2 * S * 2 + 2 * nbm * 4);
    int* s_scratch = (int*)(slot + 2 * S * 2 + 2 * nbm * 4 + 2 * a.row_words * 4);

    for (int i = threadIdx.x; i < H; i += blockDim.x) s_cw[i] = a.cw[i];
    uint16_t* g_met = a.metrics + (size_t)frame * S;
    for (int i = lt; i < S; i += tpf) s_met[i] = g_met[i];

    const unsigned char* sym_base = (const unsigned char*)a.syms + (size_t)frame * a.sym_stride;
    const int step_bytes = R * c.sym_bytes;
    uint32_t* g_dec = a.dec + ((size_t)frame * a.dec_rows + a.step0) * a.row_words;

    auto compute_bm = [&](int t, int buf) {
        int sym[MAX_R];
        const void* p = sym_base + (size_t)t * step_bytes;
        for (int i = 0; i < R; i++) sym[i] = load_symbol<POLICY>(p, i);
        for (int w = lt; w < nbm; w += tpf) s_bm[buf * nbm + w] = branch_metric<POLICY>(c, (unsigned)w, sym);
    };

    if (a.nsteps > 0 && lt < nbm) compute_bm(0, 0);
    __syncthreads();

    const int iters = (H + tpf - 1) / tpf;
    unsigned long long rsum = 0;   // minima subtracted by this launch's renormalisations (thread 0 of the frame)
    int cur = 0;   // which metric buffer is "old"
    for (int t = 0; t < a.nsteps; t++) {
        const int pb = t & 1;
        const uint16_t* oldm = s_met + cur * S;
        uint32_t* new32 = (uint32_t*)(s_met + (cur ^ 1) * S);
        const int* bm = s_bm + pb * nbm;
        uint32_t* drow = s_dec + pb * a.row_words;

        // A. butterflies of step t
        for (int it = 0; it < iters; it++) {
            const int j = it * tpf + lt;
            int d0 = 0, d1 = 0;
            if (j < H) {
                int lo, hi;
                if (PolicyTraits<POLICY>::is_signed16) { lo = (int16_t)oldm[j]; hi = (int16_t)oldm[j + H]; }
                else { lo = oldm[j]; hi = oldm[j + H]; }
                int s0, s1;
                acs<POLICY>(c.max_error, lo, hi, bm[s_cw[j]], s0, s1, d0, d1);
                new32[j] = (uint32_t)(s0 & 0xffff) | ((uint32_t)s1 << 16);
            }
            const unsigned b0 = __ballot_sync(0xffffffffu, d0);
            const unsigned b1 = __ballot_sync(0xffffffffu, d1);
            if ((lt & 31) == 0) {
                const int g = (it * tpf + lt) >> 5;
                drow[2 * g] = b0;
                drow[2 * g + 1] = b1;
            }
        }
        // B. branch metrics of step t+1 into the other table
        if (t + 1 < a.nsteps && lt < nbm) compute_bm(t + 1, pb ^ 1);
        // C. decision row of step t-1 -> global (coalesced)
        if (t > 0 && live) {
            const uint32_t* prow = s_dec + (pb ^ 1) * a.row_words;
            uint32_t* grow = g_dec + (size_t)(t - 1) * a.row_words;
            for (int w = lt; w < a.row_words; w += tpf) grow[w] = prow[w];
        }
        frame_sync<WARP_FRAME>();

        // D. renormalise when new[0] crosses the threshold (viterbi615_sse2.cpp:160, spiral27.cpp:236, sse_u8.h:63)
        if (PolicyTraits<POLICY>::has_renorm) {
            uint16_t* newm = s_met + (cur ^ 1) * S;
            const int m0 = PolicyTraits<POLICY>::is_signed16 ? (int)(int16_t)newm[0] : (int)newm[0];
            if (renorm_needed<POLICY>(c, m0)) {
                const int mn = frame_min<POLICY, WARP_FRAME>(newm, S, lt, tpf, s_scratch);
                rsum += renorm_amount<POLICY>(mn);
                for (int i = lt; i < S; i += tpf) {
                    const int v = PolicyTraits<POLICY>::is_signed16 ? (int)(int16_t)newm[i] : (int)newm[i];
                    newm[i] = (uint16_t)renorm_apply<POLICY>(v, mn);
                }
                frame_sync<WARP_FRAME>();
            }
        }
        cur ^= 1;
    }

    // epilogue: last decision row, final metrics
    if (a.nsteps > 0 && live) {
        const uint32_t* prow = s_dec + ((a.nsteps - 1) & 1) * a.row_words;
        uint32_t* grow = g_dec + (size_t)(a.nsteps - 1) * a.row_words;
        for (int w = lt; w < a.row_words; w += tpf) grow[w] = prow[w];
    }
    if (live) {
        const uint16_t* fin = s_met + cur * S;
        for (int i = lt; i < S; i += tpf) g_met[i] = fin[i];
        if (lt == 0 && a.err && rsum) a.err[frame] += rsum;
    }
}

__global__ void init_metrics_kernel(uint16_t* metrics, int n_frames, int S, int init_other, int init_start,
                                    int starting_state) {
    const size_t total = (size_t)n_frames * S;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int s = (int)(i % (size_t)S);
        metrics[i] = (uint16_t)(s == (starting_state & (S - 1)) ? init_start : init_other);
    }
}

template <int POLICY>
cudaError_t launch_policy(const GenericArgs& a, const GenericPlan& p, cudaStream_t st) {
    const int blocks = (a.count + p.frames_per_block - 1) / p.frames_per_block;
    const int threads = p.threads_per_frame * p.frames_per_block;
    cudaError_t e;
    if (p.threads_per_frame == 32) {
        auto k = acs_generic_kernel<POLICY, true>;
        e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, p.smem_bytes);
        if (e != cudaSuccess) return e;
        k<<<blocks, threads, p.smem_bytes, st>>>(a, p.threads_per_frame, p.frames_per_block);
    } else {
        auto k = acs_generic_kernel<POLICY, false>;
        e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, p.smem_bytes);
        if (e != cudaSuccess) return e;
        k<<<blocks, threads, p.smem_bytes, st>>>(a, p.threads_per_frame, p.frames_per_block);
    }
    return cudaGetLastError();
}

}  // namespace

GenericPlan generic_plan(const CodecParams& c) {
    GenericPlan p;
    const int H = c.S / 2;
    int tpf = H < 32 ? 32 : H;
    if (tpf > 512) tpf = 512;
    p.threads_per_frame = tpf;
    p.frames_per_block = tpf == 32 ? 4 : 1;
    const int row_words = c.S / 32 < 2 ? 2 : c.S / 32;
    const int slot = 2 * c.S * 2 + 2 * (1 << c.R) * 4 + 2 * row_words * 4 + 32 * 4;
    p.smem_bytes = ((H + 15) & ~15) + p.frames_per_block * slot;
    return p;
}

cudaError_t launch_acs_generic(const GenericArgs& a, cudaStream_t st) {
    if (a.count <= 0) return cudaSuccess;
    const GenericPlan p = generic_plan(a.c);
    switch (a.c.policy) {
    case P_KA9Q_U8_MOD: return launch_policy<P_KA9Q_U8_MOD>(a, p, st);
    case P_KA9Q_S16_615: return launch_policy<P_KA9Q_S16_615>(a, p, st);
    case P_KA9Q_S16_224: return launch_policy<P_KA9Q_S16_224>(a, p, st);
    case P_SPIRAL_R2: return launch_policy<P_SPIRAL_R2>(a, p, st);
    case P_SPIRAL_R4: return launch_policy<P_SPIRAL_R4>(a, p, st);
    case P_SPIRAL_615: return launch_policy<P_SPIRAL_615>(a, p, st);
    case P_WY_U8: return launch_policy<P_WY_U8>(a, p, st);
    case P_WY_U16: return launch_policy<P_WY_U16>(a, p, st);
    default: return cudaErrorInvalidValue;
    }
}

cudaError_t launch_init_metrics(uint16_t* metrics, int n_frames, const CodecParams& c, int starting_state,
                                cudaStream_t st) {
    if (n_frames <= 0) return cudaSuccess;
    const size_t total = (size_t)n_frames * c.S;
    int blocks = (int)((total + 255) / 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    init_metrics_kernel<<<blocks, 256, 0, st>>>(metrics, n_frames, c.S, c.init_other, c.init_start, starting_state);
    return cudaGetLastError();
}

}  // namespace b200v
