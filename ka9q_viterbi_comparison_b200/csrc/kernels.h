// kernels.h -- host-callable launchers of the CUDA kernels (one .cu per family).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "codec.h"

namespace b200v {

// ---- generic shared-memory ACS kernel (acs_generic.cu) -------------------------------------------
struct GenericArgs {
    CodecParams c;
    const void* syms;           // device; frame f at syms + f*sym_stride bytes
    size_t sym_stride;          // bytes
    int nsteps;                 // steps to run in this launch
    int step0;                  // decision row of the first step (the cursor)
    uint32_t* dec;              // device; dec[(f*dec_rows + row)*row_words + w]
    int dec_rows;               // rows allocated per frame
    int row_words;              // max(2, S/32)
    uint16_t* metrics;          // device; [n_frames][S] raw 16-bit (u8 policies zero-extended)
    const uint8_t* cw;          // device; [S/2] code word per butterfly
    int n_frames;
    int frame0, count;          // this launch works on frames [frame0, frame0 + count)
    unsigned long long* err;    // device [n_frames] or nullptr: += the minima this launch's renormalisations subtract
};
struct GenericPlan { int threads_per_frame, frames_per_block, smem_bytes; };
GenericPlan generic_plan(const CodecParams& c);
cudaError_t launch_acs_generic(const GenericArgs& a, cudaStream_t st);

// ---- K = 3 ... 6 register kernel (acs_smallk.cu): thread per frame, u32 decision word per (row, frame) ----
bool smallk_supported(const CodecParams& c);
cudaError_t launch_acs_smallk(const GenericArgs& a, cudaStream_t st);

// ---- K=7 warp-per-frame kernel (acs_k7_warp.cu): the alternative design, kept for the measured comparison ----
bool k7_warp_supported(const CodecParams& c);
cudaError_t launch_acs_k7_warp(const GenericArgs& a, cudaStream_t st);

// metrics[f][s] = init_other, metrics[f][start] = init_start (all families share this buffer format)
cudaError_t launch_init_metrics(uint16_t* metrics, int n_frames, const CodecParams& c, int starting_state,
                                cudaStream_t st);

// ---- register-resident K=7 kernels (acs_k7.cu): decisions as one u64 per (row, frame), dec[row*n_frames+f] ----
bool k7_supported(const CodecParams& c);
cudaError_t launch_acs_k7(const CodecParams& c, const void* syms, size_t sym_stride, int nsteps, int step0,
                          unsigned long long* dec, uint8_t* row_phase, uint16_t* metrics, int n_frames,
                          int frame0, int count, cudaStream_t st, const int* pidx = nullptr, int fill = 0,
                          unsigned long long* err = nullptr);
// err (every family): device [n_frames] or nullptr; each frame adds the minima its renormalisations subtracted in this
// launch -- the reference's update<sum_error_t>() return value (viterbi_decoder_sse_u8.h:57-69).
// pidx != nullptr: depuncturing fused into the symbol fetch -- `syms` holds only each frame's transmitted symbols and
// pidx[k] is the position of full-rate symbol k of this launch in that stream, or < 0 for a symbol that was not transmitted
// (the decoder then sees `fill`). The K=7 register kernels and the group kernels take it.

// ---- group kernels for K=9 / K=15 (acs_grp.cu): decisions as one u64 per (row, frame, thread of the frame) ----
bool grp_supported(const CodecParams& c);
int grp_threads_per_frame(const CodecParams& c);
cudaError_t launch_acs_grp(const CodecParams& c, const void* syms, size_t sym_stride, int nsteps, int step0,
                           unsigned long long* dec, uint8_t* row_phase, uint16_t* metrics, int n_frames,
                           int frame0, int count, cudaStream_t st, const int* pidx = nullptr, int fill = 0,
                           unsigned long long* err = nullptr);

// ---- K=24 (acs_k24.cu): one launch per trellis step, ping-pong metric buffers, canonical decision rows ----
// replay cache of the captured step loop of one update call (acs_k24.cu)
struct K24GraphCache {
    void* exec = nullptr;            // cudaGraphExec_t
    const void* syms = nullptr; size_t sym_stride = 0;
    int nsteps = 0, step0 = 0, cur = 0, n_frames = 0, cur_after = 0;
    long long kernels = 0;
    cudaStream_t stream = nullptr;
};
cudaError_t launch_acs_k24(const CodecParams& c, const void* syms, size_t sym_stride, int nsteps, int step0,
                           uint32_t* dec, int dec_rows, uint16_t* metrics0, uint16_t* metrics1, int* cur,
                           int* gmin, int* adj, int n_frames, long long* launches, K24GraphCache* cache,
                           cudaStream_t st, unsigned long long* err = nullptr);
void k24_graph_cache_release(K24GraphCache* cache);
bool k24_supported(const CodecParams& c);       // K = 17 ... 24, r = 1/2: ka9q viterbi224 and the williamyang policies

// ---- chainback (chainback.cu) ---------------------------------------------------------------------
struct ChainbackArgs {
    CodecParams c;
    int layout;                 // DecisionLayout
    const void* dec;
    int dec_rows, row_words;    // DL_BALLOT_FRAME_MAJOR geometry
    const uint8_t* row_phase;   // DL_K7_INTERLEAVED / DL_GRP_INTERLEAVED: phase of the step that wrote each row
    int n_frames;
    int frame0, count;          // this launch works on frames [frame0, frame0 + count)
    const uint16_t* metrics;    // for the KA9Q615 return value
    uint8_t* out;               // device; frame f at out + f*out_stride
    size_t out_stride;
    unsigned nbits, endstate;
    const unsigned* endstates;  // per-frame end states (device, [n_frames]) or nullptr: use `endstate` for all
    int* results;               // device [n_frames] or nullptr
    int par_mode, par_warmup;   // b200v_batch_set_traceback: 0 auto / 1 per-frame threads / 2 block-parallel; warm-up rows or 0
};
cudaError_t launch_chainback(const ChainbackArgs& a, cudaStream_t st);
// ---- sliding-window traceback over a ring of decision rows (chainback.cu) ----
// Logical row r (the trellis step counted since init) lives at ring position r % ring_rows. Every stream is walked from
// its end state (a.endstates[f], or a.endstate) at logical row top_row down to the row of the first emitted bit;
// the bits [first_bit, first_bit + a.nbits) are written MSB-first to a.out + f * a.out_stride (bit first_bit = bit 7 of
// byte 0). state_bits: the K-1 newest input bits [top_row + 2 - K, top_row] are read from the end state itself (best-state
// flush of an unterminated stream); otherwise the newest bit a walk can produce is top_row - (K-1).
struct StreamTracebackArgs {
    ChainbackArgs a;            // decision layout / buffers (dec_rows = ring_rows), output, end states; a.nbits = bits emitted
    unsigned ring_rows;
    unsigned long long top_row; // newest row written (logical)
    unsigned long long first_bit;
    int state_bits;
};
cudaError_t launch_stream_traceback(const StreamTracebackArgs& s, cudaStream_t st);

// best[f] = index of the smallest path metric of frame f (ties: lowest state); ka9q modular metrics are compared
// as (int8) distances from state 0, s16 metrics as signed values
cudaError_t launch_best_state(const CodecParams& c, const uint16_t* metrics, unsigned* best, int n_frames, cudaStream_t st);

// ---- depuncturing front-end (depuncture.cu) -------------------------------------------------------------
constexpr int MAX_PUNCTURE_LEN = 256;
struct DepunctureArgs {
    const void* in; size_t in_stride; size_t in_offset;     // transmitted symbols; stride in bytes, offset in symbols
    void* out; size_t out_stride; size_t out_offset;        // full-rate symbols
    int n_out;                                              // output symbols to produce per frame
    int code_len;
    int fill;                                               // value of a symbol that was not transmitted
    uint16_t prefix[MAX_PUNCTURE_LEN + 1];                  // prefix[j] = kept positions among code[0..j)
};
cudaError_t launch_depuncture(const DepunctureArgs& a, int sym_bytes, int n_frames, cudaStream_t st);
// expansion of frames [frame0, frame0 + count) through a gather table: out[f][k] = idx[k] >= 0 ? in[f][idx[k]] : fill
cudaError_t launch_depuncture_gather(const void* in, size_t in_stride, void* out, size_t out_stride, const int* idx,
                                     int n_out, int fill, int sym_bytes, int frame0, int count, cudaStream_t st);

// ---- synthetic input generator (generate.cu) ------------------------------------------------------
struct GenerateArgs {
    CodecParams c;
    void* syms; size_t sym_stride;
    uint8_t* data; size_t data_stride;     // may be nullptr
    unsigned nbits;                         // data bits per frame (multiple of 8)
    uint64_t seed, frame0;
    int amp_q8, sigma_q8, mean_q8, lo, hi;
    int n_frames;
};
cudaError_t launch_generate(const GenerateArgs& a, cudaStream_t st);

}  // namespace b200v
