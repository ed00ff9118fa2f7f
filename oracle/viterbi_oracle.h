/* TEST INFRASTRUCTURE ONLY -- never linked into, imported by or called from the product path.
 *
 * Plain-C scalar restatement of the reference's Viterbi update + chainback arithmetic for every policy
 * on the hot path. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may use it, and only as the checker.
 *
 * PARITY PIN: this restatement is checked bit-for-bit (per-step decision vectors, final metrics, decoded
 * bytes) against the UNMODIFIED reference compiled from /root/reference into oracle/_ref/libkvref.so
 * (tests/test_oracle_vs_ref.py) and against golden vectors generated from that build
 * (tests/golden/, tests/golden/make_golden.py). The reference itself holds no golden vectors.
 */
#pragma once
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum kvo_policy {
    KVO_KA9Q_U8_MOD = 0,   /* ka9q_libfec_port/viterbi27_sse2.cpp:119-175, viterbi29_sse2.cpp:108-164 */
    KVO_KA9Q_S16_615 = 1,  /* ka9q_libfec_port/viterbi615_sse2.cpp:104-191 */
    KVO_KA9Q_S16_224 = 2,  /* ka9q_libfec_port/viterbi224_sse2.cpp:135-258 */
    KVO_SPIRAL_R2 = 3,     /* spiral/spiral27.cpp:130-381, spiral29.cpp */
    KVO_SPIRAL_R4 = 4,     /* spiral/spiral47.cpp:131-, spiral49.cpp */
    KVO_SPIRAL_615 = 5,    /* spiral/spiral615.cpp:130-411 */
    KVO_WY_U8 = 6,         /* viterbi/x86/viterbi_decoder_sse_u8.h:45-154 (== avx_u8.h) */
    KVO_WY_U16 = 7         /* viterbi/x86/viterbi_decoder_sse_u16.h:45-177 (== avx_u16.h) */
} kvo_policy;

#define KVO_MAX_R 8

typedef struct kvo_codec {
    int policy;            /* kvo_policy */
    int K, R;
    int poly[KVO_MAX_R];   /* LSB = newest input bit; spiral honours a negative poly as inversion */
    /* numeric parameters (filled by kvo_codec_defaults) */
    int renorm_threshold;  /* policy-specific compare constant */
    int init_start, init_other;
    int max_error;         /* C (wy: soft_decision_max_error) */
    int soft_hi, soft_lo;  /* wy branch-table symbol values */
} kvo_codec;

/* Fill policy constants for (policy,K,R); poly is left untouched. Returns 0, or -1 if unsupported. */
int kvo_codec_defaults(kvo_codec* c, int policy, int K, int R);
/* Codec by registry name ("ka9q27", "spiral49", "wy_sse_u8_k7r2", "wy_avx_u16_k15r6", ...) with the
 * polynomials src/main.cpp:363-419 uses. Returns 0 / -1. */
int kvo_codec_by_name(kvo_codec* c, const char* name);

static inline int kvo_sym_bytes(const kvo_codec* c) { return c->policy == KVO_WY_U16 ? 2 : 1; }
static inline int kvo_metric_bytes(const kvo_codec* c) {
    return (c->policy == KVO_KA9Q_S16_615 || c->policy == KVO_KA9Q_S16_224 || c->policy == KVO_WY_U16) ? 2 : 1;
}

typedef struct kvo_decoder kvo_decoder;

/* Same five calls as the reference's per-codec API (src/ka9q_interface.h:12-56). len = trellis steps. */
kvo_decoder* kvo_create(const kvo_codec* c, int len);
int  kvo_init(kvo_decoder* d, int starting_state);
void kvo_update(kvo_decoder* d, const void* syms, int nsteps);
int  kvo_chainback(kvo_decoder* d, uint8_t* data, unsigned nbits, unsigned endstate);
void kvo_destroy(kvo_decoder* d);
/* canonical decision vectors, S/8 bytes per step */
void kvo_get_decisions(kvo_decoder* d, int step0, int nsteps, uint8_t* out);
/* current path metrics as S elements of kvo_metric_bytes() each (u8 / i16 / u16 raw bits) */
void kvo_get_metrics(kvo_decoder* d, void* out);
int  kvo_steps_done(const kvo_decoder* d);
/* Sum of the minima subtracted by the renormalisations since kvo_init: what the williamyang decoders' update<>()
 * returns, summed over the calls (viterbi_decoder_sse_u8.h:57-69; consumed as accumulated_error + get_error() by
 * run_snr_ber.cpp:369-371). Tracked for every renormalising policy. */
uint64_t kvo_renorm_sum(const kvo_decoder* d);

/* ---- test-input generation (restates src/util.h:15-62 + convolutional_encoder_shift_register.h:42-62) ---- */

/* Encode nbytes of data (MSB first) plus K-1 zero tail bits. out_bits[(nbytes*8+K-1)*R] receives 0/1. */
void kvo_encode_bits(int K, int R, const int* poly, const uint8_t* data, size_t nbytes, uint8_t* out_bits);

/* Deterministic counter-based 64-bit hash (splitmix64 finaliser of seed ^ golden*counter). */
uint64_t kvo_hash64(uint64_t seed, uint64_t counter);
/* Uniform data bytes for frame `frame`: byte i = low 8 bits of kvo_hash64(seed ^ 0xDA7A, frame*2^32+i). */
void kvo_random_bytes(uint64_t seed, uint64_t frame, uint8_t* data, size_t nbytes);

/* Integer-only noisy channel. For coded bit b (0/1) at global symbol counter n:
 *   g  = (sum of 12 uniform u16 drawn from 3 hashes) - 6*65535            ~ N(0, 65536^2)
 *   x  = (b ? +amp_q8 : -amp_q8) + ((sigma_q8 * g) >> 16)                  Q8 fixed point
 *   s  = clamp(floor((x + mean_q8 + 128) / 256), lo, hi)      mean_q8 = 256*(hi+lo)/2, e.g. 32640 for 0/255
 * amp_q8/sigma_q8 are chosen by the caller (tests: amp = mag/sqrt(1+sigma^2), run_snr_ber.cpp:317-365).
 * sigma_q8 = 0 gives the noiseless hard symbols main.cpp uses. out is int32 per symbol. */
void kvo_channel(const uint8_t* bits, size_t nsyms, uint64_t seed, uint64_t counter0,
                 int amp_q8, int sigma_q8, int mean_q8, int lo, int hi, int32_t* out);

#ifdef __cplusplus
}
#endif
