// noalloc.cu -- b200v_decode_noalloc_dev: init + update + chainback of a batch of K=7 frames without a single device
// allocation, for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
// that forbid it). It has no counterpart in the reference.
//
// The K=7 register kernel (acs_k7.cu) needs scratch memory between the update and the traceback: one u64 of
// decisions per (trellis step, frame), the path metrics, and the phase byte of every decision row. b200v_batch_create
// cudaMallocs them per instance; here they live in ONE file-scope __device__ array of B200V_NOALLOC_SCRATCH_BYTES
// that every GPU loading the library holds until the process exits.
//
// g_scratch below is, in HOST code, only a stand-in for the symbol: its device address comes from
// cudaGetSymbolAddress, asked once per device with that device current (scratch_base), and that address is what the
// kernels receive. Device code (noalloc_clear_kernel) may name the array directly.
//
// All calls share the array and it keeps the last call's contents: the entry point clears what must start at zero
// (the row phases) and rewrites the metrics on every call; decision rows are fully rewritten by the
// update before the traceback reads them. Two calls that overlap on one GPU (different streams, or two host threads)
// RACE on the scratch: serialise them.
#include "../../include/b200_viterbi.h"
#include "api_internal.h"
#include "kernels.h"

#include <cstdio>
#include <mutex>

using namespace b200v;

static __device__ __align__(256) unsigned char g_scratch[B200V_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
unsigned char* g_base[MAX_DEVICES];      // device address of g_scratch on each device, nullptr until asked

// Device address of the scratch array on `device`, which must be the current device.
cudaError_t scratch_base(int device, unsigned char** base) {
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_base[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_scratch);
        if (e != cudaSuccess) return e;
        g_base[device] = static_cast<unsigned char*>(p);
    }
    *base = g_base[device];
    return cudaSuccess;
}

// what must start at zero: the row phases at the head of the array
__global__ void noalloc_clear_kernel(size_t bytes) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < bytes / 16; i += (size_t)gridDim.x * blockDim.x)
        reinterpret_cast<uint4*>(g_scratch)[i] = make_uint4(0u, 0u, 0u, 0u);
}

size_t round256(size_t x) { return (x + 255) & ~(size_t)255; }

int failf(int code, const char* fmt, long long a = 0, long long b = 0, long long c = 0) {
    char msg[384];
    snprintf(msg, sizeof(msg), fmt, a, b, c);
    return api_fail(code, msg);
}

}  // namespace

#define NA_TRY(expr)                                                                                            \
    do {                                                                                                        \
        const cudaError_t e_ = (expr);                                                                          \
        if (e_ != cudaSuccess) {                                                                                \
            char msg_[384];                                                                                     \
            snprintf(msg_, sizeof(msg_), "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
            return api_fail(B200V_ERR_CUDA, msg_);                                                              \
        }                                                                                                       \
    } while (0)

extern "C" int b200v_decode_noalloc_dev(int codec, int K, int R, const int* poly, const b200v_wy_config* wy_config,
                                        int frame_steps, int n_frames, const void* d_syms, size_t sym_stride_bytes,
                                        uint8_t* d_data, size_t out_stride_bytes, unsigned nbits, int device,
                                        void* cuda_stream) {
    CodecParams c;
    const int rc = api_fill_params(c, codec, K, R, poly, wy_config);
    if (rc) return rc;
    if (!k7_supported(c) || c.policy == P_WY_U16)
        return failf(B200V_ERR_ARG, "b200v_decode_noalloc_dev serves the K=7 codes of the register kernel with 8-bit symbols "
                     "(ka9q27, spiral27, spiral47, wy-u8 K=7); got codec %lld K=%lld R=%lld", codec, c.K, c.R);
    if (frame_steps <= 0 || n_frames <= 0) return failf(B200V_ERR_ARG, "frame_steps=%lld n_frames=%lld must be positive", frame_steps, n_frames);
    if (!d_syms || (!d_data && nbits)) return api_fail(B200V_ERR_ARG, "NULL buffer");
    if (device < 0 || device >= MAX_DEVICES) return failf(B200V_ERR_ARG, "device %lld out of range", device);
    const int nsteps = policy_is_spiral(c.policy) ? (frame_steps / 2) * 2 : frame_steps;   // spiral27.cpp:385
    if (nbits + (unsigned)(c.K - 1) > (unsigned)nsteps)
        return failf(B200V_ERR_STATE, "chainback of %lld bits needs %lld decision rows, the update writes %lld", nbits,
                     nbits + (unsigned)(c.K - 1), nsteps);
    if (n_frames > 1 && sym_stride_bytes < (size_t)frame_steps * c.R)
        return failf(B200V_ERR_ARG, "frame stride %lld smaller than one frame's symbols", (long long)sym_stride_bytes);

    // carve the array: [row phases] (cleared) [metrics][decision rows] (rewritten)
    const size_t phase_bytes = round256((size_t)nsteps);
    const size_t metric_bytes = round256((size_t)n_frames * c.S * sizeof(uint16_t));
    const size_t dec_bytes = (size_t)nsteps * n_frames * sizeof(unsigned long long);
    const size_t need = phase_bytes + metric_bytes + dec_bytes;
    if (need > (size_t)B200V_NOALLOC_SCRATCH_BYTES)
        return failf(B200V_ERR_ARG, "%lld frames of %lld steps need %lld bytes of scratch; the allocation-free entry point "
                     "holds B200V_NOALLOC_SCRATCH_BYTES = 32 MiB (about frames * (steps * 8 + 128) bytes): use b200v_batch_*",
                     n_frames, nsteps, (long long)need);

    NA_TRY(cudaSetDevice(device));
    unsigned char* base = nullptr;
    NA_TRY(scratch_base(device, &base));
    cudaStream_t st = (cudaStream_t)cuda_stream;
    uint8_t* row_phase = base;
    uint16_t* metrics = reinterpret_cast<uint16_t*>(base + phase_bytes);
    unsigned long long* dec = reinterpret_cast<unsigned long long*>(base + phase_bytes + metric_bytes);

    const size_t clear = phase_bytes;
    noalloc_clear_kernel<<<(unsigned)((clear / 16 + 255) / 256), 256, 0, st>>>(clear);
    NA_TRY(cudaGetLastError());
    NA_TRY(launch_init_metrics(metrics, n_frames, c, 0, st));
    NA_TRY(launch_acs_k7(c, d_syms, sym_stride_bytes, nsteps, 0, dec, row_phase, metrics, n_frames, 0, n_frames, st));
    ChainbackArgs a{};
    a.c = c; a.layout = DL_K7_INTERLEAVED; a.dec = dec; a.dec_rows = nsteps; a.row_words = 2;
    a.row_phase = row_phase; a.n_frames = n_frames; a.frame0 = 0; a.count = n_frames;
    a.metrics = metrics; a.out = d_data; a.out_stride = out_stride_bytes;
    a.nbits = nbits; a.endstate = 0; a.endstates = nullptr; a.results = nullptr;
    a.par_mode = 0; a.par_warmup = 0;
    NA_TRY(launch_chainback(a, st));
    return B200V_OK;
}
