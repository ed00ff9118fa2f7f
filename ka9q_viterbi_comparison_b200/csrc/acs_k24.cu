// acs_k24.cu -- add-compare-select for K = 24 (2^23 states, ka9q viterbi224 policy): the whole GPU works on one
// trellis step of a frame; one launch per step (a step's inputs are every output of the previous step).
// Path metrics ping-pong between two 16 MiB buffers per frame (32 MiB: resident in the 126 MB L2 for small
// batches), decisions are written as the canonical S-bit vector of the step (1 MiB per step per frame, bit n =
// new state n, as viterbi224_sse2.cpp:201 lays them out), so the kernel is bound by L2/HBM traffic
// (33 MiB per step per frame), not by arithmetic: plain 32-bit integer ops, no packing tricks.
//
// Takes the role of update_viterbi224_blk_sse2 (ka9q_libfec_port/viterbi224_sse2.cpp:135-258):
//   branch metric x0 + x1, complement 510 - BM (:159-160); signed saturating adds (:163-166); decision =
//   (lo path > hi path), tie -> 0, survivor = min (:190-194); after EVERY step, if new[0] >= 25000 the minimum
//   over all states is brought down to SHRT_MIN with a wrapping subtract (:226-246).
// The renormalisation is folded into the next step: every block reduces the minimum of the metrics it writes
// into one atomicMin slot per (row, frame); the next launch reads new[0] and that minimum, decides the trigger
// exactly as the reference does, and subtracts while loading. A pending subtraction after the last step of an
// update call is applied by renorm_k24_kernel so that the stored metrics are always the reference's values.
#include "kernels.h"
#include "policy.cuh"

namespace b200v {

namespace {

constexpr int K24_BPT = 8;           // butterflies per thread: 16 B + 16 B in, 32 B + 16 decision bits out
constexpr int K24_THREADS = 256;

struct K24Args {
    const uint16_t* old_m;          // [n_frames][S] raw 16-bit patterns (s16 for ka9q224, u16 / u8 values for wy)
    uint16_t* new_m;
    const unsigned char* syms; size_t sym_stride;   // symbols of THIS step start at syms + step*2*sym_bytes
    int step;                        // step index inside this update call
    uint16_t* dec;                   // canonical row of this step: dec + frame * dec_frame_stride (u16 units)
    size_t dec_frame_stride;
    int* gmin_prev;                  // [n_frames] minimum of old_m, or nullptr when nothing is pending
    int* gmin_cur;                   // [n_frames] slot this step reduces into (pre-set to INT_MAX)
    unsigned poly0, poly1;
    int S, thr;
    int C, soft_hi, soft_lo;         // williamyang policies: max_error and the two soft-decision levels
    unsigned long long* err;         // [n_frames] or nullptr: += every adjustment applied (the reference's update() return value)
};

// metric as an integer: signed for the ka9q s16 policy, unsigned for the williamyang policies
template <int POLICY> __device__ __forceinline__ int k24_value(unsigned raw16) {
    return POLICY == P_KA9Q_S16_224 ? (int)(int16_t)(uint16_t)raw16 : (int)(raw16 & 0xffffu);
}

// pending renormalisation of the buffer a launch reads, as the amount to subtract (mod 2^16) from every metric:
// ka9q224: min - SHRT_MIN if old[0] >= thr (viterbi224_sse2.cpp:226-246); williamyang: min if old[0] >= thr
// (sse_u8.h:63-65,127-154, sse_u16.h:152-177)
template <int POLICY>
__device__ __forceinline__ int pending_adjust(const uint16_t* old_frame, const int* gmin_prev, int frame, int thr) {
    if (!gmin_prev) return 0;
    const int m0 = k24_value<POLICY>(old_frame[0]);
    if (m0 < thr) return 0;
    return POLICY == P_KA9Q_S16_224 ? gmin_prev[frame] + 32768 : gmin_prev[frame];
}

template <int POLICY>
__global__ void __launch_bounds__(K24_THREADS) acs_k24_kernel(const K24Args a) {
    constexpr bool WY = (POLICY != P_KA9Q_S16_224);
    constexpr int TOP = POLICY == P_WY_U8 ? 255 : POLICY == P_WY_U16 ? 65535 : 32767;
    const int frame = blockIdx.y;
    const int H = a.S >> 1;
    const int j0 = (blockIdx.x * K24_THREADS + threadIdx.x) * K24_BPT;     // first butterfly of this thread
    const uint16_t* old_f = a.old_m + (size_t)frame * a.S;
    uint16_t* new_f = a.new_m + (size_t)frame * a.S;
    const int adj = pending_adjust<POLICY>(old_f, a.gmin_prev, frame, a.thr);
    if (adj != 0 && a.err && blockIdx.x == 0 && threadIdx.x == 0) a.err[frame] += (unsigned)adj;

    int s0, s1;
    int e00 = 0, e01 = 0, e10 = 0, e11 = 0;          // williamyang: |sat(bt - s_i)| for bt = lo / hi, symbol i
    if (POLICY == P_WY_U16) {
        const int16_t* sp = reinterpret_cast<const int16_t*>(a.syms + (size_t)frame * a.sym_stride) + (size_t)a.step * 2;
        s0 = sp[0]; s1 = sp[1];
        e00 = abs(clamp_i(a.soft_lo - s0, -32768, 32767)); e01 = abs(clamp_i(a.soft_hi - s0, -32768, 32767));
        e10 = abs(clamp_i(a.soft_lo - s1, -32768, 32767)); e11 = abs(clamp_i(a.soft_hi - s1, -32768, 32767));
    } else if (POLICY == P_WY_U8) {
        const int8_t* sp = reinterpret_cast<const int8_t*>(a.syms + (size_t)frame * a.sym_stride) + (size_t)a.step * 2;
        s0 = sp[0]; s1 = sp[1];
        e00 = abs(clamp_i(a.soft_lo - s0, -128, 127)); e01 = abs(clamp_i(a.soft_hi - s0, -128, 127));
        e10 = abs(clamp_i(a.soft_lo - s1, -128, 127)); e11 = abs(clamp_i(a.soft_hi - s1, -128, 127));
    } else {
        const unsigned char* sp = a.syms + (size_t)frame * a.sym_stride + (size_t)a.step * 2;
        s0 = sp[0]; s1 = sp[1];
    }

    const uint4 lo4 = *reinterpret_cast<const uint4*>(old_f + j0);
    const uint4 hi4 = *reinterpret_cast<const uint4*>(old_f + H + j0);
    const uint32_t lw[4] = { lo4.x, lo4.y, lo4.z, lo4.w }, hw[4] = { hi4.x, hi4.y, hi4.z, hi4.w };
    uint32_t out[8];
    unsigned dbits = 0;
    int mn = TOP;
#pragma unroll
    for (int k = 0; k < K24_BPT; k++) {
        const int j = j0 + k;
        // wrapping 16-bit subtract of the pending adjustment (viterbi224_sse2.cpp:243-244; the williamyang
        // subtract cannot go below zero: every metric >= the minimum)
        const int lo = k24_value<POLICY>(((lw[k >> 1] >> ((k & 1) * 16)) & 0xffffu) - (unsigned)adj);
        const int hi = k24_value<POLICY>(((hw[k >> 1] >> ((k & 1) * 16)) & 0xffffu) - (unsigned)adj);
        const bool b0 = __popc((2u * (unsigned)j) & a.poly0) & 1, b1 = __popc((2u * (unsigned)j) & a.poly1) & 1;
        int bm, cm, d0, d1, v0, v1;
        if (WY) {       // sse_u8.h:92-114: saturating error sum, complement = subs(C, BM), tie -> 1
            bm = min((b0 ? e01 : e00) + (b1 ? e11 : e10), TOP);
            cm = max(a.C - bm, 0);
            const int m0 = min(lo + bm, TOP), m1 = min(hi + cm, TOP), m2 = min(lo + cm, TOP), m3 = min(hi + bm, TOP);
            v0 = min(m0, m1); v1 = min(m2, m3);
            d0 = (v0 == m1); d1 = (v1 == m3);
        } else {        // viterbi224_sse2.cpp:159-194: tie -> 0
            bm = (b0 ? (s0 ^ 255) : s0) + (b1 ? (s1 ^ 255) : s1);
            cm = 510 - bm;
            const int m0 = min(lo + bm, 32767), m1 = min(hi + cm, 32767);
            const int m2 = min(lo + cm, 32767), m3 = min(hi + bm, 32767);
            d0 = m0 > m1; d1 = m2 > m3;
            v0 = min(m0, m1); v1 = min(m2, m3);
        }
        out[k] = (uint32_t)(v0 & 0xffff) | ((uint32_t)(v1 & 0xffff) << 16);
        dbits |= (unsigned)(d0 | (d1 << 1)) << (2 * k);
        mn = min(mn, min(v0, v1));
    }
    uint4* np = reinterpret_cast<uint4*>(new_f + 2 * (size_t)j0);
    np[0] = make_uint4(out[0], out[1], out[2], out[3]);
    np[1] = make_uint4(out[4], out[5], out[6], out[7]);
    (a.dec + (size_t)frame * a.dec_frame_stride)[j0 / K24_BPT] = (uint16_t)dbits;

    // block minimum -> one atomic per block
    __shared__ int s_min[K24_THREADS / 32];
    mn = __reduce_min_sync(0xffffffffu, mn);
    if ((threadIdx.x & 31) == 0) s_min[threadIdx.x >> 5] = mn;
    __syncthreads();
    if (threadIdx.x == 0) {
        int m = s_min[0];
#pragma unroll
        for (int i = 1; i < K24_THREADS / 32; i++) m = min(m, s_min[i]);
        atomicMin(a.gmin_cur + frame, m);
    }
}

// apply a pending renormalisation in place (after the last step of an update call). Each frame's trigger is
// evaluated from a snapshot taken BEFORE any element is modified (snapshot_k24_kernel), so blocks do not race.
template <int POLICY>
__global__ void snapshot_k24_kernel(const uint16_t* m, const int* gmin_prev, int* adj_out, int S, int thr, int n_frames,
                                    unsigned long long* err) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f < n_frames) {
        const int adj = pending_adjust<POLICY>(m + (size_t)f * S, gmin_prev, f, thr);
        adj_out[f] = adj;
        if (adj != 0 && err) err[f] += (unsigned)adj;
    }
}
__global__ void renorm_k24_kernel(uint16_t* m, const int* adj_in, int S) {
    const int frame = blockIdx.y;
    const int adj = adj_in[frame];
    if (adj == 0) return;
    uint32_t* p = reinterpret_cast<uint32_t*>(m + (size_t)frame * S);
    const uint32_t sub = (uint32_t)(adj & 0xffff);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < S / 2; i += gridDim.x * blockDim.x) {
        const uint32_t v = p[i];
        p[i] = ((v - sub) & 0xffffu) | (((v >> 16) - sub) << 16);
    }
}

__global__ void fill_int_kernel(int* p, int n, int v) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

}  // namespace

// Enqueues the launches of `nsteps` steps on `st`. metrics[0] / metrics[1] are the two ping-pong buffers, *cur says
// which one holds the current metrics (updated on return). gmin: [dec_rows + 1][n_frames] ints of scratch,
// adj: [n_frames] ints.
template <int POLICY>
static cudaError_t enqueue_k24_policy(const CodecParams& c, const void* syms, size_t sym_stride, int nsteps, int step0,
                               uint32_t* dec, int dec_rows, uint16_t* metrics0, uint16_t* metrics1, int* cur,
                               int* gmin, int* adj, int n_frames, long long* launches, cudaStream_t st,
                               unsigned long long* err) {
    const int S = c.S;
    const size_t row_u16 = (size_t)S / 16;
    const int total = nsteps * n_frames;
    fill_int_kernel<<<(total + 255) / 256, 256, 0, st>>>(gmin, total, 0x7fffffff);
    (*launches)++;
    uint16_t* bufs[2] = { metrics0, metrics1 };
    const dim3 grid((unsigned)(S / 2 / K24_BPT / K24_THREADS), (unsigned)n_frames);
    for (int t = 0; t < nsteps; t++) {
        K24Args a;
        a.old_m = bufs[*cur]; a.new_m = bufs[*cur ^ 1];
        a.syms = (const unsigned char*)syms; a.sym_stride = sym_stride; a.step = t;
        a.dec = reinterpret_cast<uint16_t*>(dec) + (size_t)(step0 + t) * row_u16;
        a.dec_frame_stride = (size_t)dec_rows * row_u16;
        a.gmin_prev = t > 0 ? gmin + (size_t)(t - 1) * n_frames : nullptr;
        a.gmin_cur = gmin + (size_t)t * n_frames;
        a.poly0 = (unsigned)c.poly[0]; a.poly1 = (unsigned)c.poly[1];
        a.S = S; a.thr = c.renorm_threshold;
        a.C = c.max_error; a.soft_hi = c.soft_hi; a.soft_lo = c.soft_lo; a.err = err;
        acs_k24_kernel<POLICY><<<grid, K24_THREADS, 0, st>>>(a);
        (*launches)++;
        *cur ^= 1;
    }
    // pending renormalisation of the last step
    snapshot_k24_kernel<POLICY><<<(n_frames + 127) / 128, 128, 0, st>>>(bufs[*cur], gmin + (size_t)(nsteps - 1) * n_frames, adj,
                                                                 S, c.renorm_threshold, n_frames, err);
    renorm_k24_kernel<<<dim3(148 * 4, (unsigned)n_frames), 256, 0, st>>>(bufs[*cur], adj, S);
    (*launches) += 2;
    return cudaGetLastError();
}

static cudaError_t enqueue_k24(const CodecParams& c, const void* syms, size_t sym_stride, int nsteps, int step0,
                               uint32_t* dec, int dec_rows, uint16_t* metrics0, uint16_t* metrics1, int* cur,
                               int* gmin, int* adj, int n_frames, long long* launches, cudaStream_t st,
                               unsigned long long* err) {
    switch (c.policy) {
    case P_KA9Q_S16_224:
        return enqueue_k24_policy<P_KA9Q_S16_224>(c, syms, sym_stride, nsteps, step0, dec, dec_rows, metrics0, metrics1,
                                                  cur, gmin, adj, n_frames, launches, st, err);
    case P_WY_U8:
        return enqueue_k24_policy<P_WY_U8>(c, syms, sym_stride, nsteps, step0, dec, dec_rows, metrics0, metrics1, cur,
                                           gmin, adj, n_frames, launches, st, err);
    case P_WY_U16:
        return enqueue_k24_policy<P_WY_U16>(c, syms, sym_stride, nsteps, step0, dec, dec_rows, metrics0, metrics1, cur,
                                            gmin, adj, n_frames, launches, st, err);
    default: return cudaErrorInvalidValue;
    }
}

bool k24_supported(const CodecParams& c) {
    return c.K > 16 && c.K <= 24 && c.R == 2 &&
           (c.policy == P_KA9Q_S16_224 || c.policy == P_WY_U8 || c.policy == P_WY_U16);
}

// The step loop is launch-bound for small batches (one frame: ~7 us of L2/HBM traffic per launch, ~3.5 us of launch
// gap), so the whole sequence of an update call is captured once into a CUDA graph and replayed while the call's
// arguments repeat (the reference's benchmark loop decodes from the same buffers every iteration: main.cpp:264-278).
// `cache` may be nullptr (no graph). Falls back to plain launches if capture is not possible on this stream.
cudaError_t launch_acs_k24(const CodecParams& c, const void* syms, size_t sym_stride, int nsteps, int step0,
                           uint32_t* dec, int dec_rows, uint16_t* metrics0, uint16_t* metrics1, int* cur,
                           int* gmin, int* adj, int n_frames, long long* launches, K24GraphCache* cache,
                           cudaStream_t st, unsigned long long* err) {
    if (n_frames <= 0 || nsteps <= 0) return cudaSuccess;
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (!cache || cudaStreamIsCapturing(st, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone)
        return enqueue_k24(c, syms, sym_stride, nsteps, step0, dec, dec_rows, metrics0, metrics1, cur, gmin, adj,
                           n_frames, launches, st, err);
    const bool hit = cache->exec && cache->syms == syms && cache->sym_stride == sym_stride && cache->nsteps == nsteps &&
                     cache->step0 == step0 && cache->cur == *cur && cache->n_frames == n_frames && cache->stream == st;
    if (!hit) {
        if (cache->exec) { cudaGraphExecDestroy((cudaGraphExec_t)cache->exec); cache->exec = nullptr; }
        int cur_after = *cur;
        long long counted = 0;
        cudaGraph_t graph = nullptr;
        if (cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
            cudaGetLastError();
            return enqueue_k24(c, syms, sym_stride, nsteps, step0, dec, dec_rows, metrics0, metrics1, cur, gmin, adj,
                               n_frames, launches, st, err);
        }
        const cudaError_t e1 = enqueue_k24(c, syms, sym_stride, nsteps, step0, dec, dec_rows, metrics0, metrics1,
                                           &cur_after, gmin, adj, n_frames, &counted, st, err);
        const cudaError_t e2 = cudaStreamEndCapture(st, &graph);
        cudaGraphExec_t exec = nullptr;
        if (e1 != cudaSuccess || e2 != cudaSuccess || !graph ||
            cudaGraphInstantiate(&exec, graph, 0) != cudaSuccess) {
            if (graph) cudaGraphDestroy(graph);
            cudaGetLastError();
            return enqueue_k24(c, syms, sym_stride, nsteps, step0, dec, dec_rows, metrics0, metrics1, cur, gmin, adj,
                               n_frames, launches, st, err);
        }
        cudaGraphDestroy(graph);
        cache->exec = exec; cache->syms = syms; cache->sym_stride = sym_stride; cache->nsteps = nsteps;
        cache->step0 = step0; cache->cur = *cur; cache->n_frames = n_frames; cache->stream = st;
        cache->cur_after = cur_after; cache->kernels = counted;
    }
    const cudaError_t e = cudaGraphLaunch((cudaGraphExec_t)cache->exec, st);
    if (e != cudaSuccess) return e;
    *cur = cache->cur_after;
    *launches += cache->kernels;
    return cudaSuccess;
}

void k24_graph_cache_release(K24GraphCache* cache) {
    if (cache && cache->exec) { cudaGraphExecDestroy((cudaGraphExec_t)cache->exec); cache->exec = nullptr; }
}

}  // namespace b200v
