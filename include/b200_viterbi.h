/* b200_viterbi.h -- C-ABI of libb200viterbi.so: batched Viterbi update + chainback on one B200 (sm_100a).
 *
 * This is the drop-in boundary for the reference's decode hot path. The reference calls its decoders
 * through five free functions per codec, wrapped by the adapter templates in
 *   /root/reference/src/ka9q_interface.h:12-56   (create/init/update_blk/chainback/delete _viterbiXX_sse2)
 *   /root/reference/src/spiral_interface.h:12-56 (create/init/update/chainback/delete _spiralXX)
 * and the williamyang decoders through
 *   /root/reference/src/main.cpp:163-187         (ViterbiDecoder_Core::reset / decoder_t::update / chainback).
 * Every entry point below cites the reference function whose role it takes. The same five calls exist
 * here on a *batch* of independent frames (b200v_batch_*); the per-codec single-frame names of the
 * reference are provided on top of it by include/ref_compat/ (C++ linkage, as in the reference).
 *
 * Plain pointers and sizes only. Functions return 0 on success and a negative b200v_status on failure;
 * b200v_last_error() gives the message of the calling thread's last failure. There is no CPU fallback:
 * without a CUDA device every compute entry point fails with B200V_ERR_CUDA.
 */
#ifndef B200_VITERBI_H
#define B200_VITERBI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200V_VERSION 1
#define B200V_MAX_R 8

typedef enum b200v_status {
    B200V_OK = 0,
    B200V_ERR_ARG = -1,      /* bad argument (unknown codec, K/R mismatch, sizes out of range) */
    B200V_ERR_CUDA = -2,     /* CUDA runtime failure (no device, launch error, ...) */
    B200V_ERR_NOMEM = -3,    /* device or host allocation failed */
    B200V_ERR_STATE = -4     /* call sequence error (e.g. update past the allocated length) */
} b200v_status;

/* Arithmetic policy + code: which reference decoder's metric width, saturation, renormalisation and
 * tie-break are reproduced bit for bit. */
typedef enum b200v_codec {
    B200V_KA9Q27 = 0,     /* ka9q_libfec_port/viterbi27_sse2.cpp   K=7  r=1/2 u8 modular            */
    B200V_KA9Q29 = 1,     /* ka9q_libfec_port/viterbi29_sse2.cpp   K=9  r=1/2 u8 modular            */
    B200V_KA9Q615 = 2,    /* ka9q_libfec_port/viterbi615_sse2.cpp  K=15 r=1/6 s16 saturating        */
    B200V_KA9Q224 = 3,    /* ka9q_libfec_port/viterbi224_sse2.cpp  K=24 r=1/2 s16 saturating        */
    B200V_SPIRAL27 = 4,   /* spiral/spiral27.cpp                   K=7  r=1/2 u8 saturating         */
    B200V_SPIRAL29 = 5,   /* spiral/spiral29.cpp                   K=9  r=1/2                       */
    B200V_SPIRAL47 = 6,   /* spiral/spiral47.cpp                   K=7  r=1/4                       */
    B200V_SPIRAL49 = 7,   /* spiral/spiral49.cpp                   K=9  r=1/4                       */
    B200V_SPIRAL615 = 8,  /* spiral/spiral615.cpp                  K=15 r=1/6                       */
    B200V_WY_U8 = 9,      /* viterbi/x86/viterbi_decoder_{sse,avx}_u8.h   any K,R; int8 symbols      */
    B200V_WY_U16 = 10     /* viterbi/x86/viterbi_decoder_{sse,avx}_u16.h  any K,R; int16 symbols     */
} b200v_codec;

/* Numeric configuration of the williamyang decoders: ViterbiDecoder_Config<error_t>
 * (viterbi/viterbi_decoder_config.h:12-17) plus the branch-table symbol values
 * (viterbi/viterbi_branch_table.h:34-36). NULL in b200v_batch_create selects main.cpp's defaults
 * (src/viterbi_configs.h:22-50: soft8 for WY_U8, soft16 for WY_U16). Ignored by ka9q/spiral codecs. */
typedef struct b200v_wy_config {
    int soft_decision_high, soft_decision_low;
    unsigned soft_decision_max_error;
    unsigned initial_start_error;
    unsigned initial_non_start_error;
    unsigned renormalisation_threshold;
} b200v_wy_config;

typedef struct b200v_batch b200v_batch;   /* opaque, like the reference's struct v27 / spiral27 */

/* ---- library ---- */
int b200v_version(void);
const char* b200v_last_error(void);
int b200v_device_count(void);
/* K and R of a fixed codec (0,0 for the WY_* families). */
int b200v_codec_shape(int codec, int* K, int* R);
/* Bytes per input symbol (1, or 2 for WY_U16) and per path metric (1 or 2). */
int b200v_codec_symbol_bytes(int codec);
int b200v_codec_metric_bytes(int codec);

/* ---- instance lifetime ----
 * Takes the role of create_viterbiXX_sse2(poly,len) (viterbi27_sse2.cpp:57-75), create_spiralXX
 * (spiral27.cpp:62-78) and of ViterbiBranchTable + ViterbiDecoder_Core construction (main.cpp:163-165),
 * for n_frames independent frames at once.
 *   K,R          must match the codec (or be 0) for ka9q/spiral; required for WY_*
 *   poly[R]      generator polynomials, LSB = newest input bit; spiral codecs accept a negative value as
 *                "inverted output" (spiral27.cpp:69). Unlike the reference (process-global table filled by
 *                the first create), the table is per instance.
 *   frame_steps  trellis steps per frame incl. the K-1 tail steps (= `len` as main.cpp:247 passes it)
 *   device       CUDA device ordinal
 * Returns NULL on failure (see b200v_last_error). The instance is initialised as by init(.., 0). */
b200v_batch* b200v_batch_create(int codec, int K, int R, const int* poly, int frame_steps, int n_frames,
                                int device, const b200v_wy_config* wy_config);
/* delete_viterbiXX_sse2 (viterbi27_sse2.cpp:108-115). NULL is accepted. */
void b200v_batch_destroy(b200v_batch* h);

/* Run the instance's work on a caller-owned CUDA stream (cudaStream_t passed as void*); NULL restores the
 * instance's own stream. */
int b200v_batch_set_stream(b200v_batch* h, void* cuda_stream);
/* Choose the kernel family: 0 = automatic (fastest available for the codec/polynomials),
 * 1 = generic shared-memory kernel (any K<=16, any polynomial), 2 = register-resident specialised kernel,
 * 3 = warp-per-frame shuffle/ballot kernel (any K=7 or K=9 code and policy; the automatic family for small batches
 *     (K=7: up to 4096 frames, K=9: up to 2048) and for polynomial sets without a family-2 instance, and the
 *     measured alternative to family 2).
 * Returns B200V_ERR_ARG if the requested family does not exist for this instance. */
int b200v_batch_set_variant(b200v_batch* h, int variant);
int b200v_batch_get_variant(const b200v_batch* h);
/* WY_U16 only. The specialised u16 kernels size their 16-bit lane window for |symbol| up to about three times the
 * soft-decision range; the reference accepts any int16. Without a promise every update of such an instance first reduces
 * max |symbol| on the device and reads it back -- b200v_batch_update_dev then SYNCHRONISES the stream, and the host-buffer
 * batch decode runs unpipelined. A caller that knows its symbols (a quantiser's clamp) states the bound once:
 * updates become asynchronous again and b200v_batch_decode uses the chunk pipeline. A bound above the kernels' range
 * selects the exact (generic) family instead. 0 withdraws the promise. Symbols beyond a promised bound give undefined
 * metrics (never out-of-bounds accesses). */
int b200v_batch_set_symbol_bound(b200v_batch* h, int max_abs_symbol);
/* Traceback kernel selection. mode 0 = automatic (block-parallel speculative traceback for small and medium
 * batches, one thread per frame for large ones), 1 = always one thread per frame, 2 = block-parallel whenever
 * the frame shape allows it (nbits a multiple of 8). warmup_rows = decision rows a block walks before its own
 * bits to merge with the survivor path (0 = default, max(64, 12*(K-1))); every block boundary is verified and a
 * mismatch is re-traced sequentially, so the output is bit-exact for ANY warm-up (a short one only costs time). */
int b200v_batch_set_traceback(b200v_batch* h, int mode, int warmup_rows);

/* init_viterbiXX_sse2(p, starting_state) (viterbi27_sse2.cpp:42-54), init_spiralXX (spiral27.cpp:53-59),
 * ViterbiDecoder_Core::reset (viterbi_decoder_core.h:202-211): all frames. Asynchronous on the stream. */
int b200v_batch_init(b200v_batch* h, int starting_state);

/* update_viterbiXX_blk_sse2(p, syms, nbits) (viterbi27_sse2.cpp:119-175 ...), update_spiralXX
 * (spiral27.cpp:383-386), decoder_t::update (sse_u8.h:45-70). `nsteps` trellis steps for every frame.
 * ka9q/WY codecs continue from the previous call's position (streaming); spiral codecs restart at step 0
 * and drop an odd trailing step, exactly as the reference does.
 *   _dev : syms is a device pointer; frame f's symbols start at syms + f*frame_stride_bytes and hold
 *          nsteps*R symbols (u8, int8 or int16 per codec). Asynchronous on the stream -- except WY_U16 instances
 *          without b200v_batch_set_symbol_bound, whose symbol range check reads one int back per update.
 *   host : syms is a host pointer with the same layout; copies in, runs, and returns after completion
 *          (the reference's callers time each call on the host, main.cpp:264-278). */
int b200v_batch_update_dev(b200v_batch* h, const void* d_syms, size_t frame_stride_bytes, int nsteps);
int b200v_batch_update(b200v_batch* h, const void* syms, size_t frame_stride_bytes, int nsteps);

/* chainback_viterbiXX_sse2(p, data, nbits, endstate) (viterbi27_sse2.cpp:78-105, viterbi615_sse2.cpp:65-91,
 * viterbi224_sse2.cpp:79-121 -- reproduced literally), chainback_spiralXX (spiral27.cpp:81-120),
 * ViterbiDecoder_Core::chainback (viterbi_decoder_core.h:214-236). Frame f's (nbits+7)/8 output bytes go
 * to data + f*frame_stride_bytes. Decoded bit n lands in byte n/8, bit 7-(n%8). */
int b200v_batch_chainback_dev(b200v_batch* h, uint8_t* d_data, size_t frame_stride_bytes, unsigned nbits,
                              unsigned endstate);
int b200v_batch_chainback(b200v_batch* h, uint8_t* data, size_t frame_stride_bytes, unsigned nbits,
                          unsigned endstate);
/* Traceback of an unterminated stream (the reference's resumable update + get_error(),
 * viterbi_decoder_core.h:193-197, give the pieces): every frame is traced back from its own best state -- the
 * smallest path metric after the steps consumed so far, ties to the lowest state -- through all rows written
 * since init: nbits must equal steps_done - (K-1). d_best_states (device, n_frames unsigned, may be NULL) receives
 * the states used. Decoded bit n is the input bit of trellis step n (the last K-1 input bits are the best state
 * itself); a sliding-window caller keeps all but the last `depth` bits. Asynchronous on the stream. */
int b200v_batch_chainback_best_dev(b200v_batch* h, uint8_t* d_data, size_t frame_stride_bytes, unsigned nbits,
                                   unsigned* d_best_states);
/* Per-frame return value of the reference's chainback: the end-state path metric for KA9Q615
 * (viterbi615_sse2.cpp:76,90), 0 otherwise. Valid after a chainback; copies n_frames ints to the host. */
int b200v_batch_chainback_results(b200v_batch* h, int* results);

/* Batch entry point (new): init + update(frame_steps) + chainback(nbits) for all frames from
 * host buffers, host<->device copies included and overlapped with compute in chunks of frames.
 * syms: [n_frames][frame_steps*R] symbols, data: [n_frames][(nbits+7)/8] bytes, both densely packed. */
int b200v_batch_decode(b200v_batch* h, const void* syms, uint8_t* data, unsigned nbits);

/* The same with symbols and output already in HBM (device pointers, frame strides in bytes). Asynchronous; ordered
 * on the instance's stream. */
int b200v_batch_decode_dev(b200v_batch* h, const void* d_syms, size_t sym_stride_bytes, uint8_t* d_data,
                           size_t out_stride_bytes, unsigned nbits);

/* Allocation-free decode. NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * init(0) + update(frame_steps) + chainback(nbits, endstate 0) for n_frames frames, as b200v_batch_decode_dev computes
 * them, but without an instance: the decision rows, path metrics and row phases that b200v_batch_create cudaMallocs live
 * in one static __device__ array of B200V_NOALLOC_SCRATCH_BYTES inside the library (held by every GPU that loads it,
 * until exit). Calls no cudaMalloc / cudaFree / allocating routine. Serves the K=7 codecs of the register kernel with
 * 8-bit symbols (KA9Q27, SPIRAL27, SPIRAL47, WY_U8 with a K=7 polynomial set it is instantiated for); the batch must fit:
 *   round256(steps) + round256(128*n_frames) + 8*steps*n_frames <= B200V_NOALLOC_SCRATCH_BYTES
 * (e.g. 2026 frames x 2054 steps, or 510 x 8198); larger batches are rejected with B200V_ERR_ARG and a message.
 * d_syms / d_data are device pointers laid out as for b200v_batch_decode_dev; cuda_stream is a cudaStream_t (NULL = the
 * default stream). Asynchronous on that stream. Every call on a GPU shares the one array: calls that overlap on the GPU
 * (two streams, two host threads) race -- serialise them. */
#define B200V_NOALLOC_SCRATCH_BYTES (32u << 20)
int b200v_decode_noalloc_dev(int codec, int K, int R, const int* poly, const b200v_wy_config* wy_config,
                             int frame_steps, int n_frames, const void* d_syms, size_t sym_stride_bytes,
                             uint8_t* d_data, size_t out_stride_bytes, unsigned nbits, int device, void* cuda_stream);

/* Wait for everything queued on the instance's stream. */
int b200v_batch_sync(b200v_batch* h);

/* ---- sliding-window decode of continuous streams ----
 * The reference's decoders resume across update calls (viterbi27_sse2.cpp:121,174; sse_u8.h:59,67) and trace back over a
 * configurable length (viterbi_decoder_core.h:180-189,214-236), but their decision storage is finite: an unbounded stream
 * needs a window. b200v_stream keeps, for n_streams independent streams, a RING of depth + chunk decision rows:
 *   push:  `chunk` more trellis steps per stream (metrics continue, as a resumed update), then every stream is traced back
 *          from its own best state (smallest path metric, ties to the lowest state) through the ring, and the `chunk` oldest
 *          input bits are emitted -- each with at least `depth` newer trellis steps behind it, the usual meaning of a
 *          traceback depth. Push p (1-based) emits bits [p*chunk - depth - chunk, p*chunk - depth) once that range is
 *          non-negative; earlier pushes emit nothing (*bits_out = 0).
 *   flush: end of the stream: the remaining bits. endstate >= 0: the stream was tail-terminated into that state: emits bits
 *          [emitted, steps - (K-1)) traced from it. endstate < 0: unterminated: traced from the best state, and the K-1
 *          newest bits are the state itself: emits [emitted, steps).
 * depth must be a multiple of chunk, chunk a multiple of 8 (and even: spiral codecs step in pairs), depth >= K-1.
 * Bits are packed MSB-first, (bits+7)/8 bytes per stream at d_out + stream * out_stride_bytes; the first emitted bit of a
 * call is bit 7 of byte 0. All codecs with K <= 16. Asynchronous on the instance's stream. */
typedef struct b200v_stream b200v_stream;
b200v_stream* b200v_stream_create(int codec, int K, int R, const int* poly, int n_streams, int depth, int chunk,
                                  int device, const b200v_wy_config* wy_config);
void b200v_stream_destroy(b200v_stream* s);
/* init(starting_state) of every stream; forgets the window. */
int b200v_stream_reset(b200v_stream* s, int starting_state);
int b200v_stream_push_dev(b200v_stream* s, const void* d_syms, size_t sym_stride_bytes, uint8_t* d_out,
                          size_t out_stride_bytes, unsigned* bits_out);
int b200v_stream_flush_dev(b200v_stream* s, int endstate, uint8_t* d_out, size_t out_stride_bytes, unsigned* bits_out);
/* The same with host buffers (copies in and out, returns after completion, like the reference's synchronous calls). */
int b200v_stream_push(b200v_stream* s, const void* syms, size_t sym_stride_bytes, uint8_t* out, size_t out_stride_bytes,
                      unsigned* bits_out);
int b200v_stream_flush(b200v_stream* s, int endstate, uint8_t* out, size_t out_stride_bytes, unsigned* bits_out);
/* trellis steps consumed / bits emitted per stream since the last reset */
long long b200v_stream_steps(const b200v_stream* s);
long long b200v_stream_emitted(const b200v_stream* s);
int b200v_stream_sync(b200v_stream* s);
/* the batch instance behind the stream (introspection: b200v_batch_get_metrics, b200v_batch_launch_count ...) */
b200v_batch* b200v_stream_batch(b200v_stream* s);

/* ---- introspection (parity tests, SURVEY 8b "b200v_batch_get_decisions") ---- */
/* Canonical decision vectors of one frame: max(S/8, 1) bytes per step, bit n of the little-endian vector is the
 * decision of new state n (the layout every reference decoder uses). Host output buffer. */
int b200v_batch_get_decisions(b200v_batch* h, int frame, int step0, int nsteps, uint8_t* out);
/* Current path metrics of one frame: S elements of b200v_codec_metric_bytes() each. Host buffer. */
int b200v_batch_get_metrics(b200v_batch* h, int frame, void* out);
/* Per frame: the sum of the minima that renormalisations subtracted since the last init -- what the williamyang decoders'
 * update<sum_error_t>() returns and their callers add to get_error() (viterbi_decoder_sse_u8.h:57-69;
 * run_snr_ber.cpp:369-371, run_punctured_decoder.cpp:176, run_simple.cpp:78-79). Defined the same way for every policy
 * that renormalises (ka9q s16: min - SHRT_MIN per renormalisation); 0 for the modular ka9q policy. Copies n_frames
 * uint64 to the host; synchronises the stream. */
int b200v_batch_accumulated_error(b200v_batch* h, uint64_t* out);
/* Trellis steps consumed since the last init (the reference's dp / m_current_decoded_bit cursor). */
int b200v_batch_steps_done(const b200v_batch* h);

/* Device time of the most recent update / chainback launches on this instance, measured with CUDA events
 * on the launching stream (ms). Synchronises the stream. Either pointer may be NULL. */
int b200v_batch_last_kernel_ms(b200v_batch* h, float* update_ms, float* chainback_ms);
/* Number of kernels this instance has launched since creation. */
long long b200v_batch_launch_count(const b200v_batch* h);

/* ---- synthetic input on the device (encoder + integer AWGN channel + quantiser) ----
 * Fills frame f (global index frame0+f) of d_syms with the encoded, noisy, quantised symbols of
 * pseudo-random data derived from `seed`, and optionally the data bytes themselves into d_data
 * ((nbits)/8 bytes per frame, may be NULL). Restates src/util.h:15-62 +
 * convolutional_encoder_shift_register.h:42-62 with an integer-only channel so that host and device
 * generators agree bit for bit. amp_q8 / sigma_q8 / mean_q8 are Q8 fixed-point channel constants,
 * lo/hi the clamp range (0/255, +-3, +-127). Asynchronous on the instance's stream. */
int b200v_batch_generate_dev(b200v_batch* h, void* d_syms, size_t sym_frame_stride_bytes,
                             uint8_t* d_data, size_t data_frame_stride_bytes,
                             unsigned nbits, uint64_t seed, uint64_t frame0,
                             int amp_q8, int sigma_q8, int mean_q8, int lo, int hi);

/* ---- depuncturing front-end ----
 * The symbol loop of decode_punctured_symbols (williamyang_viterbi_lib/examples/helpers/puncture_code_helpers.h:17-55)
 * for all frames: output symbol k (k < n_out_syms) = the next transmitted symbol if puncture_code[k % code_len] is
 * non-zero, else unpunctured_value; the code index starts at 0 on every call, so a frame made of segments with
 * different codes (run_punctured_decoder.cpp:260-282) is one call per segment, advancing in_offset_syms by *consumed
 * and out_offset_syms by n_out_syms. Frame f reads d_punctured + f*in_stride_bytes and writes d_syms +
 * f*out_stride_bytes (symbols of the instance's codec: u8, int8 or int16). puncture_code is a host array of
 * code_len <= 256 flags. *consumed (may be NULL) = transmitted symbols used per frame. Asynchronous on the stream. */
int b200v_batch_depuncture_dev(b200v_batch* h, const void* d_punctured, size_t in_stride_bytes, size_t in_offset_syms,
                               const uint8_t* puncture_code, int code_len, void* d_syms, size_t out_stride_bytes,
                               size_t out_offset_syms, int n_out_syms, int unpunctured_value, int* consumed);

/* Depuncturing FUSED into the decode: only the transmitted symbols exist in memory / cross PCIe.
 * b200v_batch_set_puncturing installs the puncturing pattern of one update call -- normally a whole frame -- as the list
 * of segments decode_punctured_symbols is called with (puncture_code_helpers.h:17-55; run_punctured_decoder.cpp:260-282:
 * code, code length, full-rate symbols requested; the code index restarts at 0 in every segment). The segments must cover
 * a multiple of R symbols. *n_transmitted (may be NULL) = transmitted symbols per frame. n_segments = 0 removes the pattern.
 * b200v_batch_update_punctured_dev then consumes (sum of n_out_syms) / R trellis steps per frame from the TRANSMITTED
 * symbols at d_punctured + f * frame_stride_bytes: the K=7 register kernels and the K=9 / K=15 group kernels gather
 * through the pattern inside their symbol fetch (a symbol that was not transmitted becomes unpunctured_value in
 * registers); the other kernel families expand into a staging buffer first. Results are bit-identical to
 * b200v_batch_depuncture_dev + b200v_batch_update_dev.
 * b200v_batch_decode_punctured[_dev]: init + update_punctured + chainback(nbits, 0), the pattern must cover frame_steps;
 * the host entry copies n_transmitted symbols per frame, densely packed, in pipelined chunks of frames. */
typedef struct b200v_puncture_segment {
    const uint8_t* code;    /* code_len flags: non-zero = the symbol at this position was transmitted */
    int code_len;
    int n_out_syms;         /* full-rate symbols this segment produces */
} b200v_puncture_segment;
int b200v_batch_set_puncturing(b200v_batch* h, const b200v_puncture_segment* segments, int n_segments,
                               int unpunctured_value, int* n_transmitted);
int b200v_batch_update_punctured_dev(b200v_batch* h, const void* d_punctured, size_t frame_stride_bytes);
int b200v_batch_decode_punctured_dev(b200v_batch* h, const void* d_punctured, size_t in_stride_bytes, uint8_t* d_data,
                                     size_t out_stride_bytes, unsigned nbits);
int b200v_batch_decode_punctured(b200v_batch* h, const void* punctured, uint8_t* data, unsigned nbits);

/* ---- multi-GPU inside one process (frames are independent; no collective) ----
 * Shard g of G owns the contiguous frames [g*B/G, (g+1)*B/G). */
int b200v_shard_range(int n_frames, int n_shards, int shard, int* frame0, int* count);
typedef struct b200v_sharded b200v_sharded;
/* One b200v_batch per entry of devices[] (a device may appear more than once), sized for its shard. */
b200v_sharded* b200v_sharded_create(int codec, int K, int R, const int* poly, int frame_steps, int n_frames,
                                    const int* devices, int n_devices, const b200v_wy_config* wy_config);
/* The same with shards proportional to weights[] -- e.g. the host-to-device rate each GPU's link sustains when all copy at
 * once (tools/h2d_probe.py: on an 8-GPU box four links gave 20 GB/s and four 35 GB/s; the end-to-end rate of equal shards is
 * set by the slowest). b200v_shard_range_weighted gives the ranges: contiguous, every frame once. */
int b200v_shard_range_weighted(int n_frames, const double* weights, int n_shards, int shard, int* frame0, int* count);
b200v_sharded* b200v_sharded_create_weighted(int codec, int K, int R, const int* poly, int frame_steps, int n_frames,
                                             const int* devices, const double* weights, int n_devices,
                                             const b200v_wy_config* wy_config);
void b200v_sharded_destroy(b200v_sharded* s);
int b200v_sharded_num_shards(const b200v_sharded* s);
int b200v_sharded_shard(const b200v_sharded* s, int shard, int* frame0, int* count);
/* b200v_batch_decode of every shard, one host thread per shard, each writing its slice of `data`. */
int b200v_sharded_decode(b200v_sharded* s, const void* syms, uint8_t* data, unsigned nbits);
const char* b200v_sharded_last_error(const b200v_sharded* s);

/* Bind the calling host thread (and what it allocates from now on, pinned buffers included) to the CPUs and the memory
 * node of `device`'s PCIe root complex (sysfs topology; sched_setaffinity + set_mempolicy(MPOL_PREFERRED)). The
 * host-buffer path is PCIe-bound end to end, so where the staging memory lives decides the multi-GPU rate: call it
 * once per rank / shard thread BEFORE b200v_host_alloc_pinned. Returns the node (>= 0), or -1 if the topology is not
 * exposed (nothing changed). b200v_sharded_decode's shard threads call it themselves. */
int b200v_bind_host_to_device(int device);

/* ---- device / pinned-host memory helpers (so a C host needs nothing but this library) ---- */
void* b200v_device_alloc(int device, size_t bytes);
void  b200v_device_free(int device, void* p);
void* b200v_host_alloc_pinned(size_t bytes);
void  b200v_host_free_pinned(void* p);
/* 1 if p is page-locked host memory known to CUDA (b200v_host_alloc_pinned, cudaMallocHost, cudaHostRegister), else 0.
 * b200v_batch_decode / b200v_sharded_decode accept pageable buffers, but the driver then stages every chunk copy
 * synchronously: copies stop overlapping the kernels and the end-to-end rate drops to the staging rate. */
int   b200v_host_is_pinned(const void* p);
int   b200v_memcpy_h2d(int device, void* dst, const void* src, size_t bytes);
int   b200v_memcpy_d2h(int device, void* dst, const void* src, size_t bytes);

#ifdef __cplusplus
}
#endif
#endif /* B200_VITERBI_H */
