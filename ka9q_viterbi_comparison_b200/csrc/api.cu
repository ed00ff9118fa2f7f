// api.cu -- the C-ABI of include/b200_viterbi.h: instance state, buffer management, kernel dispatch.
// Host code only; every compute call ends in a kernel launch from acs_*.cu / chainback.cu (no CPU path).
#include "../../include/b200_viterbi.h"
#include "kernels.h"
#include "decision_layout.cuh"
#include "acs_grp.cuh"
#include "api_internal.h"

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

using namespace b200v;

// ---------------------------------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

static int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
#define CUDA_TRY(expr)                                                                                   \
    do {                                                                                                 \
        cudaError_t e_ = (expr);                                                                         \
        if (e_ != cudaSuccess)                                                                           \
            return fail(e_ == cudaErrorMemoryAllocation ? B200V_ERR_NOMEM : B200V_ERR_CUDA,              \
                        "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__);    \
    } while (0)

extern "C" const char* b200v_last_error(void) { return g_err; }
extern "C" int b200v_version(void) { return B200V_VERSION; }
extern "C" int b200v_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

// ---------------------------------------------------------------------------------------------------
// codec table
// ---------------------------------------------------------------------------------------------------
struct CodecInfo { int policy, K, R; };
static bool codec_info(int codec, CodecInfo& ci) {
    switch (codec) {
    case B200V_KA9Q27: ci = { P_KA9Q_U8_MOD, 7, 2 }; return true;
    case B200V_KA9Q29: ci = { P_KA9Q_U8_MOD, 9, 2 }; return true;
    case B200V_KA9Q615: ci = { P_KA9Q_S16_615, 15, 6 }; return true;
    case B200V_KA9Q224: ci = { P_KA9Q_S16_224, 24, 2 }; return true;
    case B200V_SPIRAL27: ci = { P_SPIRAL_R2, 7, 2 }; return true;
    case B200V_SPIRAL29: ci = { P_SPIRAL_R2, 9, 2 }; return true;
    case B200V_SPIRAL47: ci = { P_SPIRAL_R4, 7, 4 }; return true;
    case B200V_SPIRAL49: ci = { P_SPIRAL_R4, 9, 4 }; return true;
    case B200V_SPIRAL615: ci = { P_SPIRAL_615, 15, 6 }; return true;
    case B200V_WY_U8: ci = { P_WY_U8, 0, 0 }; return true;
    case B200V_WY_U16: ci = { P_WY_U16, 0, 0 }; return true;
    default: return false;
    }
}

extern "C" int b200v_codec_shape(int codec, int* K, int* R) {
    CodecInfo ci;
    if (!codec_info(codec, ci)) return fail(B200V_ERR_ARG, "unknown codec %d", codec);
    if (K) *K = ci.K;
    if (R) *R = ci.R;
    return B200V_OK;
}
extern "C" int b200v_codec_symbol_bytes(int codec) { return codec == B200V_WY_U16 ? 2 : 1; }
extern "C" int b200v_codec_metric_bytes(int codec) {
    return (codec == B200V_KA9Q615 || codec == B200V_KA9Q224 || codec == B200V_WY_U16) ? 2 : 1;
}

// Policy constants; each cites where the reference fixes the value.
static int fill_params(CodecParams& c, int codec, int K, int R, const int* poly, const b200v_wy_config* wy) {
    CodecInfo ci;
    if (!codec_info(codec, ci)) return fail(B200V_ERR_ARG, "unknown codec %d", codec);
    if (ci.K) {
        if ((K && K != ci.K) || (R && R != ci.R))
            return fail(B200V_ERR_ARG, "codec %d is K=%d R=%d, got K=%d R=%d", codec, ci.K, ci.R, K, R);
        K = ci.K; R = ci.R;
    }
    if (K < 3 || K > 24) return fail(B200V_ERR_ARG, "constraint length K=%d out of range [3,24]", K);
    if (R < 2 || R > MAX_R) return fail(B200V_ERR_ARG, "code rate 1/%d out of range [2,%d]", R, MAX_R);
    if (!poly) return fail(B200V_ERR_ARG, "poly is NULL");
    memset(&c, 0, sizeof(c));
    c.policy = ci.policy; c.K = K; c.R = R; c.S = 1 << (K - 1);
    for (int i = 0; i < R; i++) c.poly[i] = poly[i];
    c.sym_bytes = b200v_codec_symbol_bytes(codec);
    c.metric_bytes = b200v_codec_metric_bytes(codec);
    c.soft_hi = 255; c.soft_lo = 0;
    switch (ci.policy) {
    case P_KA9Q_U8_MOD:   // viterbi27_sse2.cpp:46-52
        c.init_other = 63; c.init_start = 0; c.max_error = 15; break;
    case P_KA9Q_S16_615:  // viterbi615_sse2.cpp:33-39, :160
        c.init_other = -32768 + 1000; c.init_start = -32768; c.max_error = 255 * R;
        c.renorm_threshold = 32767 - 12750; break;
    case P_KA9Q_S16_224:  // viterbi224_sse2.cpp:39-45, :226
        c.init_other = -32768 + 5000; c.init_start = -32768; c.max_error = 510; c.renorm_threshold = 25000; break;
    case P_SPIRAL_R2:     // spiral27.cpp:54-57, :236; spiral29.cpp:507
        c.init_other = 63; c.max_error = 63; c.renorm_threshold = 210; break;
    case P_SPIRAL_R4:     // spiral47.cpp:313; spiral49.cpp:790
        c.init_other = 63; c.max_error = 63; c.renorm_threshold = (K == 9) ? 103 : 126; break;
    case P_SPIRAL_615:    // spiral615.cpp:240, :269
        c.init_other = 63; c.max_error = 94; c.renorm_threshold = 74; break;
    case P_WY_U8:         // src/viterbi_configs.h:37-50 (soft8)
        c.soft_hi = 3; c.soft_lo = -3; c.max_error = (6 * R) & 0xff; c.init_other = (12 * R) & 0xff;
        c.renorm_threshold = 255 - ((12 * R) & 0xff); break;
    case P_WY_U16:        // src/viterbi_configs.h:22-35 (soft16)
        c.soft_hi = 127; c.soft_lo = -127; c.max_error = 254 * R; c.init_other = 1270 * R;
        c.renorm_threshold = 65535 - 1270 * R; break;
    }
    if (wy && policy_is_wy(ci.policy)) {
        c.soft_hi = wy->soft_decision_high; c.soft_lo = wy->soft_decision_low;
        c.max_error = (int)wy->soft_decision_max_error;
        c.init_start = (int)wy->initial_start_error;
        c.init_other = (int)wy->initial_non_start_error;
        c.renorm_threshold = (int)wy->renormalisation_threshold;
    }
    return B200V_OK;
}

// the same two services for the library's other host files (api_internal.h)
int b200v::api_fail(int code, const char* msg) { return fail(code, "%s", msg); }
int b200v::api_fill_params(CodecParams& c, int codec, int K, int R, const int* poly, const b200v_wy_config* wy) {
    return fill_params(c, codec, K, R, poly, wy);
}

// ---------------------------------------------------------------------------------------------------
// instance
// ---------------------------------------------------------------------------------------------------
struct b200v_batch {
    int codec = 0;
    CodecParams c{};
    int n_frames = 0, frame_steps = 0, dec_rows = 0, row_words = 0;
    int device = 0;
    int cursor = 0;              // trellis steps consumed since init
    int variant = 1;             // kernel family in use (1 generic, 2 register-resident)
    int layout = DL_BALLOT_FRAME_MAJOR;
    bool has_fast = false;
    bool generic_ok = true;      // the generic shared-memory family exists for this K
    int fast_layout = DL_BALLOT_FRAME_MAJOR;   // decision layout of the specialised family, if any
    cudaStream_t own_stream = nullptr, stream = nullptr;
// streams of the chunk pipeline. Measured e2e, 256 Ki K=9 r=1/2 frames (copy time ~ kernel time, so the overlap shows):
// 2 streams 11.4, 4 streams 17.5, 8 streams 24.2, 16 streams 23.6 Gbit/s; 16 / 24 / 32 / 64 chunks with 8 streams:
// 23.9 / 23.9 / 24.2 / 20.8. The PCIe-bound K=7 batch is indifferent (27.1 Gbit/s throughout).
#ifndef B200V_NPIPE
#define B200V_NPIPE 8
#endif
#ifndef B200V_MAXCHUNKS
#define B200V_MAXCHUNKS 32
#endif
    static constexpr int NPIPE = B200V_NPIPE;
    cudaStream_t pipe[NPIPE] = {};   // b200v_batch_decode's chunk pipeline
    cudaEvent_t ev_pipe = nullptr;
    cudaEvent_t ev_u0 = nullptr, ev_u1 = nullptr, ev_c0 = nullptr, ev_c1 = nullptr;
    bool timed_update = false, timed_chainback = false;
    long long launches = 0;
    // device buffers
    void* d_dec = nullptr;       size_t dec_bytes = 0;
    uint16_t* d_metrics = nullptr;
    // K=24 family: second ping-pong metric buffer, which buffer is current, per-step minima, pending adjustments
    uint16_t* d_metrics2 = nullptr; int cur_metrics = 0; int* d_gmin = nullptr; int* d_adj = nullptr;
    uint8_t* d_cw = nullptr;
    uint8_t* d_row_phase = nullptr;   // K=7 register family: phase of the step that wrote each decision row
    int* d_results = nullptr;
    unsigned long long* d_err = nullptr;   // per frame: sum of the minima subtracted by renormalisations since init
    unsigned* d_best = nullptr;  // per-frame best end states (b200v_batch_chainback_best_dev)
    int* d_scan = nullptr;       // wy-u16 symbol range check
    int sym_bound = 0;           // b200v_batch_set_symbol_bound: the caller's promise max |symbol| (0 = none: scan every update)
    K24GraphCache k24_graph;     // K=24: captured step loop of the last update call
    bool user_variant = false;   // set_variant() was called: no automatic family changes
    int tb_mode = 0, tb_warmup = 0;   // b200v_batch_set_traceback
    // puncturing pattern of a frame (b200v_batch_set_puncturing): gather table over the full-rate symbols
    int* d_pidx = nullptr; int punct_steps = 0, punct_tx = 0, punct_fill = 0;
    void* d_punct_stage = nullptr; size_t punct_stage_bytes = 0;
    // lazily allocated staging for the host-pointer entry points
    void* d_syms_stage = nullptr; size_t syms_stage_bytes = 0;
    uint8_t* d_out_stage = nullptr; size_t out_stage_bytes = 0;
};

static void free_all(b200v_batch* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->own_stream) cudaStreamSynchronize(h->own_stream);
    k24_graph_cache_release(&h->k24_graph);
    cudaFree(h->d_dec); cudaFree(h->d_metrics); cudaFree(h->d_metrics2); cudaFree(h->d_gmin); cudaFree(h->d_adj); cudaFree(h->d_cw); cudaFree(h->d_results); cudaFree(h->d_row_phase); cudaFree(h->d_scan); cudaFree(h->d_best);
    cudaFree(h->d_syms_stage); cudaFree(h->d_out_stage); cudaFree(h->d_pidx); cudaFree(h->d_punct_stage); cudaFree(h->d_err);
    if (h->ev_u0) cudaEventDestroy(h->ev_u0);
    if (h->ev_u1) cudaEventDestroy(h->ev_u1);
    if (h->ev_c0) cudaEventDestroy(h->ev_c0);
    if (h->ev_c1) cudaEventDestroy(h->ev_c1);
    for (int i = 0; i < b200v_batch::NPIPE; i++) if (h->pipe[i]) cudaStreamDestroy(h->pipe[i]);
    if (h->ev_pipe) cudaEventDestroy(h->ev_pipe);
    if (h->own_stream) cudaStreamDestroy(h->own_stream);
    delete h;
}

static int popcount32(uint32_t x) { return __builtin_popcount(x); }

static int create_impl(b200v_batch* h, int starting_state);

// automatic kernel family: the specialised one where it exists; otherwise warp-per-frame for K=7 (any polynomial
// set, ~4x the generic kernel), else the generic shared-memory kernel. Families 1 and 3 share the decision layout.
// Small K=7 batches -- down to the single frame of the reference's own benchmark loop (src/main.cpp:264-278) -- are
// latency-bound: a frame that owns a warp advances a step in ~70 ns against ~200 ns for a frame that owns one thread
// of the register kernel (measured, 8198-step frames: 1 frame 0.58 vs 1.63 ms, 1024 frames 2.7 vs 3.6 ms including
// the upload, equal at 4096 frames), so up to 2048 frames the warp-per-frame family is the automatic choice.
// Round 2 (faster warp-per-frame step, r = 1/2 codes): update + chainback through the host-buffer calls, 8198- / 4104-step
// frames, warp-per-frame against the register / group family: K=7 4096 frames 9.97 vs 11.54 ms (8192: 20.4 vs 19.8),
// K=9 2048 frames 3.57 vs 3.76 ms (4096: 6.44 vs 6.35) -> the limits move from 2048 / 1024 to 4096 / 2048 frames.
constexpr int K7_WARP_MAX_FRAMES = 4096;
constexpr int K9_WARP_MAX_FRAMES = 2048;   // K=9 (8 states per lane)
static int fallback_variant(const b200v_batch* h) {
    if (k7_warp_supported(h->c) &&
        (!h->has_fast || h->n_frames <= (h->c.K == 7 ? K7_WARP_MAX_FRAMES : K9_WARP_MAX_FRAMES))) return 3;
    return h->has_fast ? 2 : 1;
}
static int layout_of_variant(const b200v_batch* h, int variant) { return variant == 2 ? h->fast_layout : DL_BALLOT_FRAME_MAJOR; }
static int exact_variant(const b200v_batch* h) { return k7_warp_supported(h->c) ? 3 : 1; }

extern "C" b200v_batch* b200v_batch_create(int codec, int K, int R, const int* poly, int frame_steps,
                                           int n_frames, int device, const b200v_wy_config* wy_config) {
    CodecParams c;
    if (fill_params(c, codec, K, R, poly, wy_config) != B200V_OK) return nullptr;
    if (frame_steps <= 0 || n_frames <= 0) {
        fail(B200V_ERR_ARG, "frame_steps=%d n_frames=%d must be positive", frame_steps, n_frames);
        return nullptr;
    }
    if (c.K > 16 && !k24_supported(c)) {
        fail(B200V_ERR_ARG, "K=%d r=1/%d: constraint lengths above 16 exist for r=1/2 with the ka9q viterbi224 and the "
             "williamyang policies", c.K, c.R);
        return nullptr;
    }
    b200v_batch* h = new (std::nothrow) b200v_batch();
    if (!h) { fail(B200V_ERR_NOMEM, "host allocation failed"); return nullptr; }
    h->codec = codec; h->c = c; h->n_frames = n_frames; h->frame_steps = frame_steps; h->device = device;
    if (const char* e = getenv("B200V_TRACEBACK")) {     // initial b200v_batch_set_traceback mode (A/B measurements)
        const int m = atoi(e);
        if (m >= 0 && m <= 2) h->tb_mode = m;
    }
    if (create_impl(h, 0) != B200V_OK) { free_all(h); return nullptr; }
    return h;
}

extern "C" void b200v_batch_destroy(b200v_batch* h) { free_all(h); }

static int create_impl(b200v_batch* h, int starting_state) {
    const CodecParams& c = h->c;
    CUDA_TRY(cudaSetDevice(h->device));
    CUDA_TRY(cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking));
    h->stream = h->own_stream;
    CUDA_TRY(cudaEventCreate(&h->ev_u0)); CUDA_TRY(cudaEventCreate(&h->ev_u1));
    CUDA_TRY(cudaEventCreate(&h->ev_c0)); CUDA_TRY(cudaEventCreate(&h->ev_c1));

    // the reference allocates len + K-1 decision rows (viterbi27_sse2.cpp:72)
    h->dec_rows = h->frame_steps + c.K - 1;
    h->row_words = c.S / 32 < 2 ? 2 : c.S / 32;
    h->dec_bytes = (size_t)h->n_frames * h->dec_rows * h->row_words * sizeof(uint32_t);
    CUDA_TRY(cudaMalloc(&h->d_dec, h->dec_bytes));
    CUDA_TRY(cudaMalloc(&h->d_metrics, (size_t)h->n_frames * c.S * sizeof(uint16_t)));
    CUDA_TRY(cudaMalloc(&h->d_results, (size_t)h->n_frames * sizeof(int)));
    CUDA_TRY(cudaMalloc(&h->d_err, (size_t)h->n_frames * sizeof(unsigned long long)));
    CUDA_TRY(cudaMemsetAsync(h->d_results, 0, (size_t)h->n_frames * sizeof(int), h->stream));

    // code word of butterfly j: bit i = parity((2j) & poly_i) (viterbi27_sse2.cpp:61-70,
    // viterbi_branch_table.h:44-54); spiral: a negative polynomial inverts the bit (spiral27.cpp:69)
    const int H = c.K > 16 ? 0 : c.S / 2;
    std::vector<uint8_t> cw((size_t)H, 0);
    for (int i = 0; i < c.R; i++) {
        const int p = c.poly[i];
        const bool inv = policy_is_spiral(c.policy) && p < 0;
        const uint32_t ap = (uint32_t)(inv ? -p : p);
        for (int j = 0; j < H; j++)
            cw[j] |= (uint8_t)((((unsigned)popcount32((2u * (uint32_t)j) & ap) & 1u) ^ (inv ? 1u : 0u)) << i);
    }
    CUDA_TRY(cudaMalloc(&h->d_cw, (size_t)(H ? H : 1)));
    if (H) CUDA_TRY(cudaMemcpyAsync(h->d_cw, cw.data(), (size_t)H, cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));   // cw is a local vector

    if (c.K > 16) {     // K=24 family: no shared-memory kernel, no code-word table (computed from the polynomials)
        CUDA_TRY(cudaMalloc(&h->d_metrics2, (size_t)h->n_frames * c.S * sizeof(uint16_t)));
        CUDA_TRY(cudaMalloc(&h->d_gmin, (size_t)h->n_frames * h->dec_rows * sizeof(int)));
        CUDA_TRY(cudaMalloc(&h->d_adj, (size_t)h->n_frames * sizeof(int)));
        CUDA_TRY(cudaMalloc(&h->d_row_phase, (size_t)h->dec_rows));
        h->fast_layout = DL_CANONICAL; h->has_fast = true; h->generic_ok = false;
        h->variant = 2; h->layout = DL_CANONICAL;
        return b200v_batch_init(h, starting_state);
    }
    const GenericPlan plan = generic_plan(c);
    if (plan.smem_bytes > 227 * 1024)
        return fail(B200V_ERR_ARG, "K=%d needs %d bytes of shared memory per block", c.K, plan.smem_bytes);
    CUDA_TRY(cudaMalloc(&h->d_row_phase, (size_t)h->dec_rows));
    CUDA_TRY(cudaMemsetAsync(h->d_row_phase, 0, (size_t)h->dec_rows, h->stream));
    // the register-resident family exists for the polynomial sets it was instantiated for; both families
    // use 8 bytes per (row, frame) for K=7, so they share the decision buffer
    h->fast_layout = k7_supported(c) ? DL_K7_INTERLEAVED : grp_supported(c) ? DL_GRP_INTERLEAVED :
                     smallk_supported(c) ? DL_SMALLK_INTERLEAVED : DL_BALLOT_FRAME_MAJOR;
    h->has_fast = h->fast_layout != DL_BALLOT_FRAME_MAJOR;
    h->variant = fallback_variant(h);
    h->layout = layout_of_variant(h, h->variant);
    return b200v_batch_init(h, starting_state);
}

extern "C" int b200v_batch_set_stream(b200v_batch* h, void* cuda_stream) {
    if (!h) return fail(B200V_ERR_ARG, "NULL instance");
    h->stream = cuda_stream ? (cudaStream_t)cuda_stream : h->own_stream;
    return B200V_OK;
}

extern "C" int b200v_batch_set_variant(b200v_batch* h, int variant) {
    if (!h) return fail(B200V_ERR_ARG, "NULL instance");
    h->user_variant = variant != 0;
    if (variant == 0) variant = fallback_variant(h);
    if (variant != h->variant && h->cursor != 0)
        return fail(B200V_ERR_STATE, "the kernel family can only change right after init (decision layouts differ)");
    if (variant == 1 && h->generic_ok) { h->variant = 1; h->layout = DL_BALLOT_FRAME_MAJOR; return B200V_OK; }
    if (variant == 2 && h->has_fast) { h->variant = 2; h->layout = h->fast_layout; return B200V_OK; }
    if (variant == 3 && k7_warp_supported(h->c)) { h->variant = 3; h->layout = DL_BALLOT_FRAME_MAJOR; return B200V_OK; }
    return fail(B200V_ERR_ARG, "kernel family %d is not available for this instance", variant);
}
extern "C" int b200v_batch_get_variant(const b200v_batch* h) { return h ? h->variant : 0; }

extern "C" int b200v_batch_set_symbol_bound(b200v_batch* h, int max_abs_symbol) {
    if (!h) return fail(B200V_ERR_ARG, "NULL instance");
    if (max_abs_symbol < 0) return fail(B200V_ERR_ARG, "negative symbol bound");
    h->sym_bound = max_abs_symbol;
    return B200V_OK;
}

extern "C" int b200v_batch_set_traceback(b200v_batch* h, int mode, int warmup_rows) {
    if (!h) return fail(B200V_ERR_ARG, "NULL instance");
    if (mode < 0 || mode > 2 || warmup_rows < 0) return fail(B200V_ERR_ARG, "traceback mode %d / warm-up %d", mode, warmup_rows);
    h->tb_mode = mode; h->tb_warmup = warmup_rows;
    return B200V_OK;
}

extern "C" int b200v_batch_init(b200v_batch* h, int starting_state) {
    if (!h) return fail(B200V_ERR_ARG, "NULL instance");
    CUDA_TRY(cudaSetDevice(h->device));
    CUDA_TRY(launch_init_metrics(h->d_metrics, h->n_frames, h->c, starting_state, h->stream));
    CUDA_TRY(cudaMemsetAsync(h->d_err, 0, (size_t)h->n_frames * sizeof(unsigned long long), h->stream));
    h->launches++;
    h->cursor = 0;
    h->cur_metrics = 0;
    if (!h->user_variant && h->layout != DL_CANONICAL) {     // back to the automatic family (wy-u16 may have left it)
        h->variant = fallback_variant(h);
        h->layout = layout_of_variant(h, h->variant);
    }
    return B200V_OK;
}

// ---- wy-u16: symbol range check ---------------------------------------------------------------------------
// The specialised u16 kernels size their 16-bit lane window for branch metrics up to 2 * soft_decision_max_error
// (true for every symbol within about three times the soft-decision range). int16 symbols outside that range are
// legal input for the reference, so every update of a wy-u16 instance first reduces max |symbol| on the device;
// a batch that exceeds the range is decoded by the generic kernel (exact for any input) instead.
__global__ void max_abs_s16_kernel(const int16_t* syms, size_t stride_elems, int per_frame, int n_frames, int* out) {
    int m = 0;
    const size_t total = (size_t)per_frame * n_frames;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t f = i / per_frame, k = i - f * per_frame;
        const int v = (int)syms[f * stride_elems + k];
        m = max(m, v < 0 ? -v : v);
    }
    m = __reduce_max_sync(0xffffffffu, m);
    if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(out, m);
}

static int wy16_symbol_limit(const CodecParams& c) {
    const int bt = c.soft_hi > -c.soft_lo ? (c.soft_hi > 0 ? c.soft_hi : -c.soft_hi) : (c.soft_lo < 0 ? -c.soft_lo : c.soft_lo);
    return 2 * c.max_error / c.R - bt;      // R * (|bt| + |s|) <= 2 * max_error
}

// Returns 1 if every symbol of this update is inside the specialised kernels' range, 0 if not, < 0 on error.
// With a symbol bound promised by the caller (b200v_batch_set_symbol_bound) nothing is scanned and nothing synchronises.
static int wy16_symbols_in_range(b200v_batch* h, const void* d_syms, size_t stride_bytes, int per_frame) {
    if (h->sym_bound > 0) return h->sym_bound <= wy16_symbol_limit(h->c) ? 1 : 0;
    if (!h->d_scan) CUDA_TRY(cudaMalloc(&h->d_scan, sizeof(int)));
    CUDA_TRY(cudaMemsetAsync(h->d_scan, 0, sizeof(int), h->stream));
    const size_t total = (size_t)per_frame * h->n_frames;
    int blocks = (int)((total + 1023) / 1024);
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    max_abs_s16_kernel<<<blocks, 256, 0, h->stream>>>((const int16_t*)d_syms, stride_bytes / 2, per_frame, h->n_frames, h->d_scan);
    CUDA_TRY(cudaGetLastError());
    h->launches++;
    int m = 0;
    CUDA_TRY(cudaMemcpyAsync(&m, h->d_scan, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    return m <= wy16_symbol_limit(h->c) ? 1 : 0;
}

// Launch the ACS kernel of the instance's family for frames [f0, f0 + cnt) on stream st.
static int ensure_stage(void** p, size_t* have, size_t need);

// pidx != nullptr: d_syms holds only the transmitted symbols (stride = bytes per frame of that stream) and pidx is the
// gather table of this launch's steps. The K=7 register kernels and the group kernels fetch through the table; the other
// families get the frames expanded into the full-rate staging buffer first.
static int launch_update_range(b200v_batch* h, const void* d_syms, size_t stride, int nsteps, int step0, int f0, int cnt,
                               cudaStream_t st, const int* pidx = nullptr) {
    const CodecParams& c = h->c;
    if (nsteps <= 0 || cnt <= 0) return B200V_OK;
    const bool fused = h->variant == 2 && (h->layout == DL_GRP_INTERLEAVED || h->layout == DL_K7_INTERLEAVED);
    if (pidx && !fused) {
        const size_t srow = (size_t)nsteps * c.R * c.sym_bytes;
        if (h->syms_stage_bytes < srow * h->n_frames)
            return fail(B200V_ERR_STATE, "full-rate staging buffer missing for the depuncturing expansion");
        CUDA_TRY(launch_depuncture_gather(d_syms, stride, h->d_syms_stage, srow, pidx, nsteps * c.R, h->punct_fill,
                                          c.sym_bytes, f0, cnt, st));
        h->launches++;
        d_syms = h->d_syms_stage; stride = srow; pidx = nullptr;
    }
    const int pfill = h->punct_fill;
    if (h->layout == DL_CANONICAL) {
        if (f0 != 0 || cnt != h->n_frames) return fail(B200V_ERR_STATE, "the K=24 family works on whole batches");
        CUDA_TRY(launch_acs_k24(c, d_syms, stride, nsteps, step0, (uint32_t*)h->d_dec, h->dec_rows,
                                h->d_metrics, h->d_metrics2, &h->cur_metrics, h->d_gmin, h->d_adj, h->n_frames,
                                &h->launches, &h->k24_graph, st, h->d_err));
    } else if (h->variant == 2 && h->layout == DL_GRP_INTERLEAVED) {
        CUDA_TRY(launch_acs_grp(c, d_syms, stride, nsteps, step0, (unsigned long long*)h->d_dec,
                                h->d_row_phase, h->d_metrics, h->n_frames, f0, cnt, st, pidx, pfill, h->d_err));
        h->launches++;
    } else if (h->variant == 2 && h->layout == DL_SMALLK_INTERLEAVED) {
        GenericArgs a{};
        a.c = c; a.syms = d_syms; a.sym_stride = stride; a.nsteps = nsteps; a.step0 = step0;
        a.dec = (uint32_t*)h->d_dec; a.dec_rows = h->dec_rows; a.row_words = 1;
        a.metrics = h->d_metrics; a.cw = h->d_cw; a.n_frames = h->n_frames; a.frame0 = f0; a.count = cnt; a.err = h->d_err;
        CUDA_TRY(launch_acs_smallk(a, st));
        h->launches++;
    } else if (h->variant == 2) {
        CUDA_TRY(launch_acs_k7(c, d_syms, stride, nsteps, step0, (unsigned long long*)h->d_dec,
                               h->d_row_phase, h->d_metrics, h->n_frames, f0, cnt, st, pidx, pfill, h->d_err));
        h->launches++;
    } else {
        GenericArgs a{};
        a.c = c; a.syms = d_syms; a.sym_stride = stride; a.nsteps = nsteps; a.step0 = step0;
        a.dec = (uint32_t*)h->d_dec; a.dec_rows = h->dec_rows; a.row_words = h->row_words;
        a.metrics = h->d_metrics; a.cw = h->d_cw; a.n_frames = h->n_frames; a.frame0 = f0; a.count = cnt; a.err = h->d_err;
        if (h->variant == 3) CUDA_TRY(launch_acs_k7_warp(a, st));
        else CUDA_TRY(launch_acs_generic(a, st));
        h->launches++;
    }
    return B200V_OK;
}

extern "C" int b200v_batch_update_dev(b200v_batch* h, const void* d_syms, size_t frame_stride_bytes, int nsteps) {
    if (!h) return fail(B200V_ERR_ARG, "NULL instance");
    if (nsteps < 0 || (!d_syms && nsteps > 0)) return fail(B200V_ERR_ARG, "bad symbol buffer / step count");
    const CodecParams& c = h->c;
    if (policy_is_spiral(c.policy)) {   // spiral27.cpp:383-386: restart at row 0, nbits/2 double-steps
        h->cursor = 0;
        nsteps = (nsteps / 2) * 2;
    }
    if (h->cursor + nsteps > h->dec_rows)
        return fail(B200V_ERR_STATE, "update of %d steps at step %d exceeds the %d decision rows allocated",
                    nsteps, h->cursor, h->dec_rows);
    if (nsteps > 0 && frame_stride_bytes < (size_t)nsteps * c.R * c.sym_bytes && h->n_frames > 1)
        return fail(B200V_ERR_ARG, "frame stride %zu smaller than one frame's symbols", frame_stride_bytes);
    CUDA_TRY(cudaSetDevice(h->device));
    if (c.policy == P_WY_U16 && h->variant == 2 && h->layout != DL_SMALLK_INTERLEAVED && h->layout != DL_CANONICAL &&
        nsteps > 0) {   // the small-K and K=24 kernels compute in int32: exact for any input
        const int ok = wy16_symbols_in_range(h, d_syms, frame_stride_bytes, nsteps * c.R);
        if (ok < 0) return ok;
        if (!ok) {
            if (h->cursor != 0)
                return fail(B200V_ERR_ARG, "int16 symbols beyond +-%d in a continued update: the specialised u16 kernels "
                            "cannot take over mid-frame; select the generic family (variant 1) after init",
                            wy16_symbol_limit(c));
            h->variant = exact_variant(h); h->layout = DL_BALLOT_FRAME_MAJOR;      // exact for any input
        }
    }
    CUDA_TRY(cudaEventRecord(h->ev_u0, h->stream));
    int rc = launch_update_range(h, d_syms, frame_stride_bytes, nsteps, h->cursor, 0, h->n_frames, h->stream);
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(h->ev_u1, h->stream));
    h->timed_update = true;
    h->cursor += nsteps;
    return B200V_OK;
}

static int ensure_stage(void** p, size_t* have, size_t need) {
    if (*have >= need) return B200V_OK;
    if (*p) { CUDA_TRY(cudaFree(*p)); *p = nullptr; *have = 0; }
    CUDA_TRY(cudaMalloc(p, need));
    *have = need;
    return B200V_OK;
}

extern "C" int b200v_batch_update(b200v_batch* h, const void* syms, size_t frame_stride_bytes, int nsteps) {
    if (!h) return fail(B200V_ERR_ARG, "NULL instance");
    if (nsteps < 0 || (!syms && nsteps > 0)) return fail(B200V_ERR_ARG, "bad symbol buffer / step count");
    CUDA_TRY(cudaSetDevice(h->device));
    const size_t row = (size_t)nsteps * h->c.R * h->c.sym_bytes;
    if (h->n_frames == 1) frame_stride_bytes = row;
    if (row > 0) {
        int rc = ensure_stage(&h->d_syms_stage, &h->syms_stage_bytes, row * h->n_frames);
        if (rc) return rc;
        CUDA_TRY(cudaMemcpy2DAsync(h->d_syms_stage, row, syms, frame_stride_bytes, row, (size_t)h->n_frames,
                                   cudaMemcpyHostToDevice, h->stream));
    }
    int rc = b200v_batch_update_dev(h, h->d_syms_stage, row, nsteps);
    if (rc) return rc;
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    return B200V_OK;
}

static int launch_chainback_range(b200v_batch* h, uint8_t* d_data, size_t stride, unsigned nbits, unsigned endstate,
                                  int f0, int cnt, cudaStream_t st, const unsigned* endstates = nullptr,
                                  bool pipelined = false) {
    ChainbackArgs a{};
    a.c = h->c; a.layout = h->layout; a.dec = h->d_dec; a.dec_rows = h->dec_rows; a.row_words = h->row_words;
    a.row_phase = h->d_row_phase;
    a.n_frames = h->n_frames; a.frame0 = f0; a.count = cnt;
    a.metrics = h->cur_metrics ? h->d_metrics2 : h->d_metrics; a.out = d_data; a.out_stride = stride;
    a.nbits = nbits; a.endstate = endstate; a.endstates = endstates; a.results = h->d_results;
    // chunks of the pipelined host-buffer decode run beside other chunks' kernels and copies: throughput, not
    // latency, so they keep the per-frame kernels (measured: K=7 65536 frames e2e 17.4 vs 10.3 Gbit/s)
    // (block-parallel traceback for the chunks of the other layouts measured slower too: K=9 e2e 17.5 -> 13.3 Gbit/s)
    a.par_mode = (pipelined && h->tb_mode == 0) ? 1 : h->tb_mode; a.par_warmup = h->tb_warmup;
    CUDA_TRY(launch_chainback(a, st));
    h->launches++;
    return B200V_OK;
}

extern "C" int b200v_batch_chainback_dev(b200v_batch* h, uint8_t* d_data, size_t frame_stride_bytes,
                                         unsigned nbits, unsigned endstate) {
    if (!h) return fail(B200V_ERR_ARG, "NULL instance");
    if (!d_data && nbits) return fail(B200V_ERR_ARG, "NULL output buffer");
    const CodecParams& c = h->c;
    const unsigned rows_needed = nbits + (c.policy == P_KA9Q_S16_224 ? 0u : (unsigned)(c.K - 1));
    if (rows_needed > (unsigned)h->dec_rows)
        return fail(B200V_ERR_STATE, "chainback of %u bits needs %u decision rows, %d allocated", nbits,
                    rows_needed, h->dec_rows);
    CUDA_TRY(cudaSetDevice(h->device));
    CUDA_TRY(cudaEventRecord(h->ev_c0, h->stream));
    int rc = launch_chainback_range(h, d_data, frame_stride_bytes, nbits, endstate, 0, h->n_frames, h->stream);
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(h->ev_c1, h->stream));
    h->timed_chainback = true;
    return B200V_OK;
}

// Traceback of an UNTERMINATED stream: every frame is traced from its own best state (smallest path
// metric after the steps consumed so far) instead of a common end state; the caller of a sliding-window decoder
// keeps the first nbits - depth bits and re-decodes the rest when more symbols have arrived.
extern "C" int b200v_batch_chainback_best_dev(b200v_batch* h, uint8_t* d_data, size_t frame_stride_bytes, unsigned nbits,
                                              unsigned* d_best_states) {
    if (!h) return fail(B200V_ERR_ARG, "NULL instance");
    if (!d_data && nbits) return fail(B200V_ERR_ARG, "NULL output buffer");
    const CodecParams& c = h->c;
    const unsigned rows_needed = nbits + (c.policy == P_KA9Q_S16_224 ? 0u : (unsigned)(c.K - 1));
    if (rows_needed != (unsigned)h->cursor)      // the best state belongs to the last row written
        return fail(B200V_ERR_STATE, "best-state traceback covers all %d rows written since init: nbits must be %d, got %u",
                    h->cursor, h->cursor - (int)(rows_needed - nbits), nbits);
    CUDA_TRY(cudaSetDevice(h->device));
    if (!h->d_best) CUDA_TRY(cudaMalloc(&h->d_best, (size_t)h->n_frames * sizeof(unsigned)));
    CUDA_TRY(launch_best_state(c, h->cur_metrics ? h->d_metrics2 : h->d_metrics, h->d_best, h->n_frames, h->stream));
    h->launches++;
    if (d_best_states)
        CUDA_TRY(cudaMemcpyAsync(d_best_states, h->d_best, (size_t)h->n_frames * sizeof(unsigned), cudaMemcpyDeviceToDevice,
                                 h->stream));
    return launch_chainback_range(h, d_data, frame_stride_bytes, nbits, 0, 0, h->n_frames, h->stream, h->d_best);
}

extern "C" int b200v_batch_chainback(b200v_batch* h, uint8_t* data, size_t frame_stride_bytes, unsigned nbits,
                                     unsigned endstate) {
    if (!h) return fail(B200V_ERR_ARG, "NULL instance");
    if (!data && nbits) return fail(B200V_ERR_ARG, "NULL output buffer");
    CUDA_TRY(cudaSetDevice(h->device));
    const size_t row = (nbits + 7) / 8;
    if (h->n_frames == 1) frame_stride_bytes = row;
    if (row == 0) return B200V_OK;
    int rc = ensure_stage((void**)&h->d_out_stage, &h->out_stage_bytes, row * h->n_frames);
    if (rc) return rc;
    rc = b200v_batch_chainback_dev(h, h->d_out_stage, row, nbits, endstate);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpy2DAsync(data, frame_stride_bytes, h->d_out_stage, row, row, (size_t)h->n_frames,
                               cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    return B200V_OK;
}

extern "C" int b200v_batch_chainback_results(b200v_batch* h, int* results) {
    if (!h || !results) return fail(B200V_ERR_ARG, "NULL argument");
    CUDA_TRY(cudaSetDevice(h->device));
    CUDA_TRY(cudaMemcpyAsync(results, h->d_results, (size_t)h->n_frames * sizeof(int), cudaMemcpyDeviceToHost,
                             h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    return B200V_OK;
}

// Batch entry point. Large batches are cut into chunks of frames that flow through NPIPE streams, each chunk
// doing H2D copy -> update -> chainback -> D2H copy on its stream, so that the copies of one chunk overlap the
// kernels of another (the path is PCIe-bound end to end: R*T symbol bytes in per decoded bit).
static int decode_host(b200v_batch* h, const void* syms, uint8_t* data, unsigned nbits, bool punct) {
    if (!h) return fail(B200V_ERR_ARG, "NULL instance");
    if (!syms || (!data && nbits)) return fail(B200V_ERR_ARG, "NULL buffer");
    const CodecParams& c = h->c;
    if (punct && (!h->d_pidx || h->punct_steps != h->frame_steps))
        return fail(B200V_ERR_STATE, "b200v_batch_decode_punctured needs a puncturing pattern that covers the %d steps of a frame "
                    "(b200v_batch_set_puncturing; the installed one covers %d)", h->frame_steps, h->punct_steps);
    const size_t srow = (size_t)h->frame_steps * c.R * c.sym_bytes;      // full-rate bytes per frame
    const size_t irow = punct ? (size_t)h->punct_tx * c.sym_bytes : srow; // bytes per frame that cross PCIe
    const size_t orow = (nbits + 7) / 8;
    const int min_chunk = 2048;
    // wy-u16: the specialised kernels need in-range symbols; host symbols can only be scanned after the copy, so the
    // chunk pipeline is open to wy-u16 instances whose caller promised a bound (b200v_batch_set_symbol_bound)
    const bool wy16_unchecked = c.policy == P_WY_U16 && !(h->sym_bound > 0 && h->sym_bound <= wy16_symbol_limit(c) &&
                                                          (!punct || abs(h->punct_fill) <= wy16_symbol_limit(c)));
    if (h->layout == DL_CANONICAL || wy16_unchecked || h->n_frames < 2 * min_chunk || orow == 0) {
        int rc = b200v_batch_init(h, 0);
        if (rc) return rc;
        if (punct) {
            CUDA_TRY(cudaSetDevice(h->device));
            rc = ensure_stage(&h->d_punct_stage, &h->punct_stage_bytes, irow * h->n_frames);
            if (rc) return rc;
            CUDA_TRY(cudaMemcpyAsync(h->d_punct_stage, syms, irow * h->n_frames, cudaMemcpyHostToDevice, h->stream));
            rc = b200v_batch_update_punctured_dev(h, h->d_punct_stage, irow);
        } else {
            rc = b200v_batch_update(h, syms, srow, h->frame_steps);
        }
        if (rc) return rc;
        return b200v_batch_chainback(h, data, orow, nbits, 0);
    }
    const unsigned rows_needed = nbits + (unsigned)(c.K - 1);
    if (rows_needed > (unsigned)h->dec_rows)
        return fail(B200V_ERR_STATE, "chainback of %u bits needs %u decision rows, %d allocated", nbits,
                    rows_needed, h->dec_rows);
    CUDA_TRY(cudaSetDevice(h->device));
    const bool fused = h->variant == 2 && (h->layout == DL_GRP_INTERLEAVED || h->layout == DL_K7_INTERLEAVED);
    int rc = B200V_OK;
    if (!punct || !fused) rc = ensure_stage(&h->d_syms_stage, &h->syms_stage_bytes, srow * h->n_frames);
    if (rc) return rc;
    if (punct) rc = ensure_stage(&h->d_punct_stage, &h->punct_stage_bytes, irow * h->n_frames);
    if (rc) return rc;
    char* d_in = (char*)(punct ? h->d_punct_stage : h->d_syms_stage);
    rc = ensure_stage((void**)&h->d_out_stage, &h->out_stage_bytes, orow * h->n_frames);
    if (rc) return rc;
    for (int i = 0; i < b200v_batch::NPIPE; i++)
        if (!h->pipe[i]) CUDA_TRY(cudaStreamCreateWithFlags(&h->pipe[i], cudaStreamNonBlocking));
    if (!h->ev_pipe) CUDA_TRY(cudaEventCreateWithFlags(&h->ev_pipe, cudaEventDisableTiming));

    rc = b200v_batch_init(h, 0);
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(h->ev_pipe, h->stream));
    const int nsteps = policy_is_spiral(c.policy) ? (h->frame_steps / 2) * 2 : h->frame_steps;
    int nchunks = h->n_frames / min_chunk;
    if (nchunks > B200V_MAXCHUNKS) nchunks = B200V_MAXCHUNKS;
    const int chunk = (((h->n_frames + nchunks - 1) / nchunks) + 127) / 128 * 128;
    for (int i = 0; i < b200v_batch::NPIPE; i++) CUDA_TRY(cudaStreamWaitEvent(h->pipe[i], h->ev_pipe, 0));
    int k = 0;
    rc = B200V_OK;
    cudaError_t ce = cudaSuccess;
    for (int f0 = 0; f0 < h->n_frames && rc == B200V_OK && ce == cudaSuccess; f0 += chunk, k++) {
        const int cnt = (h->n_frames - f0 < chunk) ? h->n_frames - f0 : chunk;
        cudaStream_t st = h->pipe[k % b200v_batch::NPIPE];
        ce = cudaMemcpyAsync(d_in + (size_t)f0 * irow, (const char*)syms + (size_t)f0 * irow,
                             (size_t)cnt * irow, cudaMemcpyHostToDevice, st);
        if (ce != cudaSuccess) break;
        rc = launch_update_range(h, d_in, irow, nsteps, 0, f0, cnt, st, punct ? h->d_pidx : nullptr);
        if (rc) break;
        rc = launch_chainback_range(h, h->d_out_stage, orow, nbits, 0, f0, cnt, st, nullptr, true);
        if (rc) break;
        ce = cudaMemcpyAsync(data + (size_t)f0 * orow, h->d_out_stage + (size_t)f0 * orow, (size_t)cnt * orow,
                             cudaMemcpyDeviceToHost, st);
    }
    // success or not, nothing of this call may still be running on the pipeline streams when it returns: the caller
    // owns `syms` and `data` again
    for (int i = 0; i < b200v_batch::NPIPE; i++) {
        const cudaError_t se = cudaStreamSynchronize(h->pipe[i]);
        if (ce == cudaSuccess) ce = se;
    }
    if (rc) return rc;
    if (ce != cudaSuccess)
        return fail(ce == cudaErrorMemoryAllocation ? B200V_ERR_NOMEM : B200V_ERR_CUDA, "pipelined decode failed: %s",
                    cudaGetErrorString(ce));
    h->cursor = nsteps;
    return B200V_OK;
}

extern "C" int b200v_batch_decode(b200v_batch* h, const void* syms, uint8_t* data, unsigned nbits) {
    return decode_host(h, syms, data, nbits, false);
}
extern "C" int b200v_batch_decode_punctured(b200v_batch* h, const void* punctured, uint8_t* data, unsigned nbits) {
    return decode_host(h, punctured, data, nbits, true);
}

// Device-pointer batch entry: init + update(frame_steps) + chainback(nbits) with symbols and output already in
// HBM; asynchronous, on the instance's stream. Measured on the B200: running the HBM-bound chainback of one chunk
// of frames beside the issue-bound update of the next (priority streams) is SLOWER than back-to-back whole-batch
// launches (23.9 vs 20.5 ms for 1 Mi K=7 frames), and so is a single warp-specialised launch in which every block
// updates 128 frames on four warps while a fifth warp traces back the frames of a block that finished one residency
// earlier (ticket + release/acquire flags; bit-exact, but 40.4 vs 19.5 ms: one traceback warp per block is latency
// bound at ~0.9 ms per 32 frames and holds the block's slot 4x longer than the update needs). So the three kernels
// simply follow each other.
extern "C" int b200v_batch_decode_dev(b200v_batch* h, const void* d_syms, size_t sym_stride_bytes, uint8_t* d_data,
                                      size_t out_stride_bytes, unsigned nbits) {
    if (!h) return fail(B200V_ERR_ARG, "NULL instance");
    if (!d_syms || (!d_data && nbits)) return fail(B200V_ERR_ARG, "NULL buffer");
    int rc = b200v_batch_init(h, 0);
    if (rc) return rc;
    rc = b200v_batch_update_dev(h, d_syms, sym_stride_bytes, h->frame_steps);
    if (rc) return rc;
    return b200v_batch_chainback_dev(h, d_data, out_stride_bytes, nbits, 0);
}

extern "C" int b200v_batch_sync(b200v_batch* h) {
    if (!h) return fail(B200V_ERR_ARG, "NULL instance");
    CUDA_TRY(cudaSetDevice(h->device));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    return B200V_OK;
}

// ---------------------------------------------------------------------------------------------------
// sliding-window decode of continuous streams
// ---------------------------------------------------------------------------------------------------
// A b200v_batch whose decision rows are used as a ring: logical row r (trellis step since reset) lives at row r % W,
// W = depth + chunk. depth is a multiple of chunk, so an update launch of `chunk` steps never wraps, and every ring
// position is always written by the same step-in-launch (the row phases of the register kernels stay valid).
struct b200v_stream {
    b200v_batch* b = nullptr;
    int depth = 0, chunk = 0, W = 0;
    long long steps = 0, emitted = 0;
};

extern "C" b200v_stream* b200v_stream_create(int codec, int K, int R, const int* poly, int n_streams, int depth, int chunk,
                                             int device, const b200v_wy_config* wy_config) {
    CodecParams c;
    if (fill_params(c, codec, K, R, poly, wy_config) != B200V_OK) return nullptr;
    if (c.K > 16) { fail(B200V_ERR_ARG, "stream decoding exists for K <= 16 (K=%d)", c.K); return nullptr; }
    if (chunk <= 0 || chunk % 8 || depth <= 0 || depth % chunk || depth < c.K - 1) {
        fail(B200V_ERR_ARG, "stream geometry: chunk=%d must be a positive multiple of 8, depth=%d a multiple of chunk and "
             ">= K-1 = %d", chunk, depth, c.K - 1);
        return nullptr;
    }
    b200v_stream* s = new (std::nothrow) b200v_stream();
    if (!s) { fail(B200V_ERR_NOMEM, "host allocation failed"); return nullptr; }
    s->depth = depth; s->chunk = chunk; s->W = depth + chunk;
    // b200v_batch_create allocates frame_steps + K-1 rows
    s->b = b200v_batch_create(codec, K, R, poly, s->W - (c.K - 1), n_streams, device, wy_config);
    if (!s->b) { delete s; return nullptr; }
    if (s->b->c.policy == P_WY_U16) {      // exact for any int16 input: no per-update range scan, no family change mid-stream
        s->b->variant = exact_variant(s->b); s->b->layout = DL_BALLOT_FRAME_MAJOR; s->b->user_variant = true;
    }
    return s;
}

extern "C" void b200v_stream_destroy(b200v_stream* s) {
    if (!s) return;
    b200v_batch_destroy(s->b);
    delete s;
}

extern "C" int b200v_stream_reset(b200v_stream* s, int starting_state) {
    if (!s) return fail(B200V_ERR_ARG, "NULL stream");
    s->steps = 0; s->emitted = 0;
    return b200v_batch_init(s->b, starting_state);
}

static int stream_traceback(b200v_stream* s, int endstate, unsigned long long first_bit, unsigned nbits, int state_bits,
                            uint8_t* d_out, size_t out_stride) {
    b200v_batch* h = s->b;
    if (nbits == 0) return B200V_OK;
    if (!d_out) return fail(B200V_ERR_ARG, "NULL output buffer");
    if (out_stride < (nbits + 7) / 8 && h->n_frames > 1) return fail(B200V_ERR_ARG, "output stride %zu smaller than %u bits", out_stride, nbits);
    StreamTracebackArgs t{};
    ChainbackArgs& a = t.a;
    a.c = h->c; a.layout = h->layout; a.dec = h->d_dec; a.dec_rows = h->dec_rows; a.row_words = h->row_words;
    a.row_phase = h->d_row_phase; a.n_frames = h->n_frames; a.frame0 = 0; a.count = h->n_frames;
    a.metrics = h->d_metrics; a.out = d_out; a.out_stride = out_stride; a.nbits = nbits;
    a.endstate = endstate < 0 ? 0u : (unsigned)endstate; a.endstates = nullptr; a.results = nullptr;
    if (endstate < 0) {
        if (!h->d_best) CUDA_TRY(cudaMalloc(&h->d_best, (size_t)h->n_frames * sizeof(unsigned)));
        CUDA_TRY(launch_best_state(h->c, h->d_metrics, h->d_best, h->n_frames, h->stream));
        h->launches++;
        a.endstates = h->d_best;
    }
    t.ring_rows = (unsigned)s->W; t.top_row = (unsigned long long)(s->steps - 1); t.first_bit = first_bit; t.state_bits = state_bits;
    CUDA_TRY(launch_stream_traceback(t, h->stream));
    h->launches++;
    return B200V_OK;
}

extern "C" int b200v_stream_push_dev(b200v_stream* s, const void* d_syms, size_t sym_stride_bytes, uint8_t* d_out,
                                     size_t out_stride_bytes, unsigned* bits_out) {
    if (!s) return fail(B200V_ERR_ARG, "NULL stream");
    if (!d_syms) return fail(B200V_ERR_ARG, "NULL symbol buffer");
    b200v_batch* h = s->b;
    if (bits_out) *bits_out = 0;
    if (h->n_frames > 1 && sym_stride_bytes < (size_t)s->chunk * h->c.R * h->c.sym_bytes)
        return fail(B200V_ERR_ARG, "stream stride %zu smaller than one chunk's symbols", sym_stride_bytes);
    CUDA_TRY(cudaSetDevice(h->device));
    const int pos = (int)(s->steps % s->W);
    int rc = launch_update_range(h, d_syms, sym_stride_bytes, s->chunk, pos, 0, h->n_frames, h->stream);
    if (rc) return rc;
    s->steps += s->chunk;
    h->cursor = pos + s->chunk;
    if (s->steps < s->W) return B200V_OK;
    const unsigned long long first = (unsigned long long)(s->steps - s->W);
    rc = stream_traceback(s, -1, first, (unsigned)s->chunk, 0, d_out, out_stride_bytes);
    if (rc) return rc;
    s->emitted = s->steps - s->depth;
    if (bits_out) *bits_out = (unsigned)s->chunk;
    return B200V_OK;
}

extern "C" int b200v_stream_flush_dev(b200v_stream* s, int endstate, uint8_t* d_out, size_t out_stride_bytes, unsigned* bits_out) {
    if (!s) return fail(B200V_ERR_ARG, "NULL stream");
    b200v_batch* h = s->b;
    if (bits_out) *bits_out = 0;
    if (endstate >= h->c.S) return fail(B200V_ERR_ARG, "end state %d out of range", endstate);
    const long long end = endstate < 0 ? s->steps : s->steps - (h->c.K - 1);
    if (end <= s->emitted) return B200V_OK;
    CUDA_TRY(cudaSetDevice(h->device));
    const unsigned nbits = (unsigned)(end - s->emitted);
    int rc = stream_traceback(s, endstate, (unsigned long long)s->emitted, nbits, endstate < 0 ? 1 : 0, d_out, out_stride_bytes);
    if (rc) return rc;
    s->emitted = end;
    if (bits_out) *bits_out = nbits;
    return B200V_OK;
}

// host-buffer variants: copy in, push / flush on the device, copy the emitted bytes out; synchronous like the reference's calls
extern "C" int b200v_stream_push(b200v_stream* s, const void* syms, size_t sym_stride_bytes, uint8_t* out,
                                 size_t out_stride_bytes, unsigned* bits_out) {
    if (!s) return fail(B200V_ERR_ARG, "NULL stream");
    if (!syms) return fail(B200V_ERR_ARG, "NULL symbol buffer");
    b200v_batch* h = s->b;
    CUDA_TRY(cudaSetDevice(h->device));
    const size_t row = (size_t)s->chunk * h->c.R * h->c.sym_bytes, orow = (size_t)s->chunk / 8;
    if (h->n_frames == 1) { sym_stride_bytes = row; out_stride_bytes = orow; }
    int rc = ensure_stage(&h->d_syms_stage, &h->syms_stage_bytes, row * h->n_frames);
    if (rc) return rc;
    rc = ensure_stage((void**)&h->d_out_stage, &h->out_stage_bytes, (size_t)(s->W / 8 + 1) * h->n_frames);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpy2DAsync(h->d_syms_stage, row, syms, sym_stride_bytes, row, (size_t)h->n_frames, cudaMemcpyHostToDevice, h->stream));
    unsigned n = 0;
    rc = b200v_stream_push_dev(s, h->d_syms_stage, row, h->d_out_stage, orow, &n);
    if (rc) return rc;
    if (n) {
        if (!out) return fail(B200V_ERR_ARG, "NULL output buffer");
        CUDA_TRY(cudaMemcpy2DAsync(out, out_stride_bytes, h->d_out_stage, orow, orow, (size_t)h->n_frames, cudaMemcpyDeviceToHost, h->stream));
    }
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    if (bits_out) *bits_out = n;
    return B200V_OK;
}

extern "C" int b200v_stream_flush(b200v_stream* s, int endstate, uint8_t* out, size_t out_stride_bytes, unsigned* bits_out) {
    if (!s) return fail(B200V_ERR_ARG, "NULL stream");
    b200v_batch* h = s->b;
    CUDA_TRY(cudaSetDevice(h->device));
    const size_t orow = (size_t)s->W / 8 + 1;
    int rc = ensure_stage((void**)&h->d_out_stage, &h->out_stage_bytes, orow * h->n_frames);
    if (rc) return rc;
    unsigned n = 0;
    rc = b200v_stream_flush_dev(s, endstate, h->d_out_stage, orow, &n);
    if (rc) return rc;
    if (n) {
        if (!out) return fail(B200V_ERR_ARG, "NULL output buffer");
        const size_t nb = (n + 7) / 8;
        if (h->n_frames == 1) out_stride_bytes = nb;
        CUDA_TRY(cudaMemcpy2DAsync(out, out_stride_bytes, h->d_out_stage, orow, nb, (size_t)h->n_frames, cudaMemcpyDeviceToHost, h->stream));
    }
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    if (bits_out) *bits_out = n;
    return B200V_OK;
}

extern "C" long long b200v_stream_steps(const b200v_stream* s) { return s ? s->steps : 0; }
extern "C" long long b200v_stream_emitted(const b200v_stream* s) { return s ? s->emitted : 0; }
extern "C" int b200v_stream_sync(b200v_stream* s) { return s ? b200v_batch_sync(s->b) : fail(B200V_ERR_ARG, "NULL stream"); }
extern "C" b200v_batch* b200v_stream_batch(b200v_stream* s) { return s ? s->b : nullptr; }

// ---------------------------------------------------------------------------------------------------
// introspection
// ---------------------------------------------------------------------------------------------------
extern "C" int b200v_batch_get_decisions(b200v_batch* h, int frame, int step0, int nsteps, uint8_t* out) {
    if (!h || !out) return fail(B200V_ERR_ARG, "NULL argument");
    if (frame < 0 || frame >= h->n_frames || step0 < 0 || nsteps < 0 || step0 + nsteps > h->dec_rows)
        return fail(B200V_ERR_ARG, "frame/step range out of bounds");
    CUDA_TRY(cudaSetDevice(h->device));
    const int S = h->c.S;
    const size_t nb = S >= 8 ? (size_t)S / 8 : 1;     // K = 3: four states still take one byte per step
    memset(out, 0, nb * nsteps);
    if (h->layout == DL_CANONICAL) {
        const uint32_t* src = (const uint32_t*)h->d_dec + ((size_t)frame * h->dec_rows + step0) * h->row_words;
        if (nsteps > 0) {
            CUDA_TRY(cudaMemcpyAsync(out, src, nb * nsteps, cudaMemcpyDeviceToHost, h->stream));
            CUDA_TRY(cudaStreamSynchronize(h->stream));
        }
        return B200V_OK;
    }
    if (h->layout == DL_BALLOT_FRAME_MAJOR) {
        std::vector<uint32_t> rows((size_t)nsteps * h->row_words);
        const uint32_t* src = (const uint32_t*)h->d_dec + ((size_t)frame * h->dec_rows + step0) * h->row_words;
        CUDA_TRY(cudaMemcpyAsync(rows.data(), src, rows.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
        CUDA_TRY(cudaStreamSynchronize(h->stream));
        for (int t = 0; t < nsteps; t++) {
            const uint32_t* r = rows.data() + (size_t)t * h->row_words;
            uint8_t* o = out + (size_t)t * nb;
            for (int n = 0; n < S; n++) {
                unsigned w, b;
                ballot_position((unsigned)n, w, b);
                o[n >> 3] |= (uint8_t)(((r[w] >> b) & 1u) << (n & 7));
            }
        }
        return B200V_OK;
    }
    if (h->layout == DL_SMALLK_INTERLEAVED) {
        std::vector<uint32_t> rows((size_t)nsteps);
        const uint32_t* src = (const uint32_t*)h->d_dec + (size_t)step0 * h->n_frames + frame;
        if (nsteps > 0) {
            CUDA_TRY(cudaMemcpy2DAsync(rows.data(), sizeof(uint32_t), src, (size_t)h->n_frames * sizeof(uint32_t),
                                       sizeof(uint32_t), (size_t)nsteps, cudaMemcpyDeviceToHost, h->stream));
            CUDA_TRY(cudaStreamSynchronize(h->stream));
        }
        for (int t = 0; t < nsteps; t++) memcpy(out + (size_t)t * nb, &rows[t], nb);    // little-endian bit n = state n
        return B200V_OK;
    }
    if (h->layout == DL_K7_INTERLEAVED) {
        std::vector<unsigned long long> rows((size_t)nsteps);
        std::vector<uint8_t> phase((size_t)nsteps);
        const unsigned long long* src = (const unsigned long long*)h->d_dec + (size_t)step0 * h->n_frames + frame;
        if (nsteps > 0) {
            CUDA_TRY(cudaMemcpy2DAsync(rows.data(), 8, src, (size_t)h->n_frames * 8, 8, (size_t)nsteps,
                                       cudaMemcpyDeviceToHost, h->stream));
            CUDA_TRY(cudaMemcpyAsync(phase.data(), h->d_row_phase + step0, (size_t)nsteps, cudaMemcpyDeviceToHost,
                                     h->stream));
            CUDA_TRY(cudaStreamSynchronize(h->stream));
        }
        for (int t = 0; t < nsteps; t++) {
            uint8_t* o = out + (size_t)t * nb;
            for (int n = 0; n < S; n++)
                o[n >> 3] |= (uint8_t)(((rows[t] >> k7::decision_bit_of(phase[t], n, k7::decisions_bytewise(h->c.policy))) & 1ull) << (n & 7));
        }
        return B200V_OK;
    }
    if (h->layout == DL_GRP_INTERLEAVED) {
        const int tpf = grp_threads_per_frame(h->c), nxb = h->c.K - 1 - grp::NL;
        std::vector<unsigned long long> rows((size_t)nsteps * tpf);
        std::vector<uint8_t> phase((size_t)nsteps);
        const unsigned long long* src = (const unsigned long long*)h->d_dec + ((size_t)step0 * h->n_frames + frame) * tpf;
        if (nsteps > 0) {
            CUDA_TRY(cudaMemcpy2DAsync(rows.data(), (size_t)tpf * 8, src, (size_t)h->n_frames * tpf * 8, (size_t)tpf * 8,
                                       (size_t)nsteps, cudaMemcpyDeviceToHost, h->stream));
            CUDA_TRY(cudaMemcpyAsync(phase.data(), h->d_row_phase + step0, (size_t)nsteps, cudaMemcpyDeviceToHost,
                                     h->stream));
            CUDA_TRY(cudaStreamSynchronize(h->stream));
        }
        for (int t = 0; t < nsteps; t++) {
            uint8_t* o = out + (size_t)t * nb;
            for (int n = 0; n < S; n++) {
                unsigned tx, b;
                grp::decision_pos(nxb, phase[t], (unsigned)n, tx, b);
                o[n >> 3] |= (uint8_t)(((rows[(size_t)t * tpf + tx] >> b) & 1ull) << (n & 7));
            }
        }
        return B200V_OK;
    }
    return fail(B200V_ERR_STATE, "unknown decision layout %d", h->layout);
}

extern "C" int b200v_batch_get_metrics(b200v_batch* h, int frame, void* out) {
    if (!h || !out) return fail(B200V_ERR_ARG, "NULL argument");
    if (frame < 0 || frame >= h->n_frames) return fail(B200V_ERR_ARG, "frame out of bounds");
    CUDA_TRY(cudaSetDevice(h->device));
    const int S = h->c.S;
    std::vector<uint16_t> m((size_t)S);
    CUDA_TRY(cudaMemcpyAsync(m.data(), (h->cur_metrics ? h->d_metrics2 : h->d_metrics) + (size_t)frame * S, (size_t)S * 2,
                             cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    if (h->c.metric_bytes == 2) memcpy(out, m.data(), (size_t)S * 2);
    else for (int i = 0; i < S; i++) ((uint8_t*)out)[i] = (uint8_t)m[i];
    return B200V_OK;
}

extern "C" int b200v_batch_steps_done(const b200v_batch* h) { return h ? h->cursor : 0; }

extern "C" int b200v_batch_accumulated_error(b200v_batch* h, uint64_t* out) {
    if (!h || !out) return fail(B200V_ERR_ARG, "NULL argument");
    CUDA_TRY(cudaSetDevice(h->device));
    CUDA_TRY(cudaMemcpyAsync(out, h->d_err, (size_t)h->n_frames * sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    return B200V_OK;
}

extern "C" int b200v_batch_last_kernel_ms(b200v_batch* h, float* update_ms, float* chainback_ms) {
    if (!h) return fail(B200V_ERR_ARG, "NULL instance");
    CUDA_TRY(cudaSetDevice(h->device));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    if (update_ms) {
        *update_ms = 0.f;
        if (h->timed_update) CUDA_TRY(cudaEventElapsedTime(update_ms, h->ev_u0, h->ev_u1));
    }
    if (chainback_ms) {
        *chainback_ms = 0.f;
        if (h->timed_chainback) CUDA_TRY(cudaEventElapsedTime(chainback_ms, h->ev_c0, h->ev_c1));
    }
    return B200V_OK;
}

extern "C" long long b200v_batch_launch_count(const b200v_batch* h) { return h ? h->launches : 0; }

// ---------------------------------------------------------------------------------------------------
// synthetic input
// ---------------------------------------------------------------------------------------------------
extern "C" int b200v_batch_generate_dev(b200v_batch* h, void* d_syms, size_t sym_frame_stride_bytes,
                                        uint8_t* d_data, size_t data_frame_stride_bytes, unsigned nbits,
                                        uint64_t seed, uint64_t frame0, int amp_q8, int sigma_q8, int mean_q8,
                                        int lo, int hi) {
    if (!h || !d_syms) return fail(B200V_ERR_ARG, "NULL argument");
    if (nbits % 8) return fail(B200V_ERR_ARG, "nbits=%u must be a multiple of 8", nbits);
    CUDA_TRY(cudaSetDevice(h->device));
    GenerateArgs a{};
    a.c = h->c; a.syms = d_syms; a.sym_stride = sym_frame_stride_bytes; a.data = d_data;
    a.data_stride = data_frame_stride_bytes; a.nbits = nbits; a.seed = seed; a.frame0 = frame0;
    a.amp_q8 = amp_q8; a.sigma_q8 = sigma_q8; a.mean_q8 = mean_q8; a.lo = lo; a.hi = hi; a.n_frames = h->n_frames;
    CUDA_TRY(launch_generate(a, h->stream));
    h->launches++;
    return B200V_OK;
}

// ---------------------------------------------------------------------------------------------------
// depuncturing front-end
// ---------------------------------------------------------------------------------------------------
extern "C" int b200v_batch_depuncture_dev(b200v_batch* h, const void* d_punctured, size_t in_stride_bytes,
                                          size_t in_offset_syms, const uint8_t* puncture_code, int code_len,
                                          void* d_syms, size_t out_stride_bytes, size_t out_offset_syms,
                                          int n_out_syms, int unpunctured_value, int* consumed) {
    if (!h || !d_punctured || !d_syms || !puncture_code) return fail(B200V_ERR_ARG, "NULL argument");
    if (code_len <= 0 || code_len > MAX_PUNCTURE_LEN)
        return fail(B200V_ERR_ARG, "puncture code length %d out of range [1,%d]", code_len, MAX_PUNCTURE_LEN);
    if (n_out_syms < 0) return fail(B200V_ERR_ARG, "negative symbol count");
    DepunctureArgs a{};
    a.in = d_punctured; a.in_stride = in_stride_bytes; a.in_offset = in_offset_syms;
    a.out = d_syms; a.out_stride = out_stride_bytes; a.out_offset = out_offset_syms;
    a.n_out = n_out_syms; a.code_len = code_len; a.fill = unpunctured_value;
    a.prefix[0] = 0;
    for (int j = 0; j < code_len; j++) a.prefix[j + 1] = (uint16_t)(a.prefix[j] + (puncture_code[j] ? 1 : 0));
    if (consumed) *consumed = (n_out_syms / code_len) * a.prefix[code_len] + a.prefix[n_out_syms % code_len];
    CUDA_TRY(cudaSetDevice(h->device));
    CUDA_TRY(launch_depuncture(a, h->c.sym_bytes, h->n_frames, h->stream));
    h->launches++;
    return B200V_OK;
}

// The puncturing pattern of a whole frame as a gather table: entry k = position of full-rate symbol k in the frame's
// transmitted stream, or -1. Segment by segment exactly as decode_punctured_symbols walks them
// (puncture_code_helpers.h:17-55: the code index restarts at 0 in every segment; run_punctured_decoder.cpp:260-282).
extern "C" int b200v_batch_set_puncturing(b200v_batch* h, const b200v_puncture_segment* segments, int n_segments,
                                          int unpunctured_value, int* n_transmitted) {
    if (!h) return fail(B200V_ERR_ARG, "NULL instance");
    if (n_transmitted) *n_transmitted = 0;
    CUDA_TRY(cudaSetDevice(h->device));
    if (n_segments <= 0) {
        CUDA_TRY(cudaStreamSynchronize(h->stream));
        cudaFree(h->d_pidx); h->d_pidx = nullptr; h->punct_steps = h->punct_tx = 0;
        return B200V_OK;
    }
    if (!segments) return fail(B200V_ERR_ARG, "NULL segment list");
    std::vector<int> idx;
    int tx = 0;
    for (int g = 0; g < n_segments; g++) {
        const b200v_puncture_segment& sg = segments[g];
        if (!sg.code || sg.code_len <= 0 || sg.n_out_syms < 0)
            return fail(B200V_ERR_ARG, "puncture segment %d: code=%p code_len=%d n_out_syms=%d", g, (const void*)sg.code,
                        sg.code_len, sg.n_out_syms);
        for (int k = 0; k < sg.n_out_syms; k++) idx.push_back(sg.code[k % sg.code_len] ? tx++ : -1);
    }
    if (idx.empty() || idx.size() % (size_t)h->c.R)
        return fail(B200V_ERR_ARG, "the segments cover %zu full-rate symbols: not a positive multiple of R = %d", idx.size(), h->c.R);
    const int steps = (int)(idx.size() / (size_t)h->c.R);
    if (steps > h->dec_rows) return fail(B200V_ERR_ARG, "the segments cover %d steps, the instance holds %d", steps, h->dec_rows);
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    cudaFree(h->d_pidx); h->d_pidx = nullptr;
    CUDA_TRY(cudaMalloc(&h->d_pidx, idx.size() * sizeof(int)));
    CUDA_TRY(cudaMemcpy(h->d_pidx, idx.data(), idx.size() * sizeof(int), cudaMemcpyHostToDevice));
    h->punct_steps = steps; h->punct_tx = tx; h->punct_fill = unpunctured_value;
    if (n_transmitted) *n_transmitted = tx;
    return B200V_OK;
}

extern "C" int b200v_batch_update_punctured_dev(b200v_batch* h, const void* d_punctured, size_t frame_stride_bytes) {
    if (!h) return fail(B200V_ERR_ARG, "NULL instance");
    if (!h->d_pidx) return fail(B200V_ERR_STATE, "no puncturing pattern installed (b200v_batch_set_puncturing)");
    if (!d_punctured) return fail(B200V_ERR_ARG, "NULL symbol buffer");
    const CodecParams& c = h->c;
    int nsteps = h->punct_steps;
    if (policy_is_spiral(c.policy)) { h->cursor = 0; nsteps = (nsteps / 2) * 2; }     // spiral27.cpp:383-386
    if (h->cursor + nsteps > h->dec_rows)
        return fail(B200V_ERR_STATE, "update of %d steps at step %d exceeds the %d decision rows allocated", nsteps,
                    h->cursor, h->dec_rows);
    if (h->n_frames > 1 && frame_stride_bytes < (size_t)h->punct_tx * c.sym_bytes)
        return fail(B200V_ERR_ARG, "frame stride %zu smaller than one frame's %d transmitted symbols", frame_stride_bytes, h->punct_tx);
    CUDA_TRY(cudaSetDevice(h->device));
    if (c.policy == P_WY_U16 && h->variant == 2 && h->layout != DL_SMALLK_INTERLEAVED && h->layout != DL_CANONICAL) {
        int ok = wy16_symbols_in_range(h, d_punctured, frame_stride_bytes, h->punct_tx);
        if (ok < 0) return ok;
        const int f = h->punct_fill < 0 ? -h->punct_fill : h->punct_fill;
        if (f > wy16_symbol_limit(c)) ok = 0;
        if (!ok) {
            if (h->cursor != 0)
                return fail(B200V_ERR_ARG, "int16 symbols beyond +-%d in a continued update: select the generic family "
                            "(variant 1) after init", wy16_symbol_limit(c));
            h->variant = exact_variant(h); h->layout = DL_BALLOT_FRAME_MAJOR;
        }
    }
    const bool fused = h->variant == 2 && (h->layout == DL_GRP_INTERLEAVED || h->layout == DL_K7_INTERLEAVED);
    if (!fused) {
        int rc = ensure_stage(&h->d_syms_stage, &h->syms_stage_bytes, (size_t)nsteps * c.R * c.sym_bytes * h->n_frames);
        if (rc) return rc;
    }
    CUDA_TRY(cudaEventRecord(h->ev_u0, h->stream));
    int rc = launch_update_range(h, d_punctured, frame_stride_bytes, nsteps, h->cursor, 0, h->n_frames, h->stream, h->d_pidx);
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(h->ev_u1, h->stream));
    h->timed_update = true;
    h->cursor += nsteps;
    return B200V_OK;
}

extern "C" int b200v_batch_decode_punctured_dev(b200v_batch* h, const void* d_punctured, size_t in_stride_bytes,
                                                uint8_t* d_data, size_t out_stride_bytes, unsigned nbits) {
    if (!h) return fail(B200V_ERR_ARG, "NULL instance");
    if (!d_punctured || (!d_data && nbits)) return fail(B200V_ERR_ARG, "NULL buffer");
    int rc = b200v_batch_init(h, 0);
    if (rc) return rc;
    rc = b200v_batch_update_punctured_dev(h, d_punctured, in_stride_bytes);
    if (rc) return rc;
    return b200v_batch_chainback_dev(h, d_data, out_stride_bytes, nbits, 0);
}

// ---------------------------------------------------------------------------------------------------
// memory helpers
// ---------------------------------------------------------------------------------------------------
extern "C" void* b200v_device_alloc(int device, size_t bytes) {
    void* p = nullptr;
    if (cudaSetDevice(device) != cudaSuccess || cudaMalloc(&p, bytes) != cudaSuccess) {
        fail(B200V_ERR_NOMEM, "cudaMalloc(%zu) on device %d failed: %s", bytes, device,
             cudaGetErrorString(cudaGetLastError()));
        return nullptr;
    }
    return p;
}
extern "C" void b200v_device_free(int device, void* p) {
    if (p && cudaSetDevice(device) == cudaSuccess) cudaFree(p);
}
extern "C" void* b200v_host_alloc_pinned(size_t bytes) {
    void* p = nullptr;
    if (cudaMallocHost(&p, bytes) != cudaSuccess) {
        fail(B200V_ERR_NOMEM, "cudaMallocHost(%zu) failed: %s", bytes, cudaGetErrorString(cudaGetLastError()));
        return nullptr;
    }
    return p;
}
extern "C" void b200v_host_free_pinned(void* p) { if (p) cudaFreeHost(p); }
extern "C" int b200v_host_is_pinned(const void* p) {
    cudaPointerAttributes at;
    if (!p || cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return 0; }
    return at.type == cudaMemoryTypeHost ? 1 : 0;
}
extern "C" int b200v_memcpy_h2d(int device, void* dst, const void* src, size_t bytes) {
    CUDA_TRY(cudaSetDevice(device));
    CUDA_TRY(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice));
    return B200V_OK;
}
extern "C" int b200v_memcpy_d2h(int device, void* dst, const void* src, size_t bytes) {
    CUDA_TRY(cudaSetDevice(device));
    CUDA_TRY(cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost));
    return B200V_OK;
}
