/* TEST INFRASTRUCTURE ONLY -- never linked into or called by the product library.
 *
 * C-ABI over the *unmodified* reference decoders, compiled from the sources where they lie under
 * /root/reference (see oracle/Makefile; outputs only into oracle/_ref/, which is git-ignored).
 * It exists so tests can (a) pin oracle/viterbi_oracle.c against the real reference and
 * (b) time the reference's CPU decoders as bench.py's `cpu_baseline` / `--impl reference` arm.
 *
 * One shim TU per reference .cpp: the shim #includes the reference translation unit so that it can see
 * the opaque instance struct (decision buffer, metric buffers) -- no reference source is copied into
 * this repository.
 */
#pragma once
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Uniform table of entry points for one reference decoder. */
typedef struct kvref_vtable {
    const char* name;      /* "ka9q27", "spiral615", "wy_sse_u8_k7r2", ... */
    int K, R;
    int sym_bytes;         /* bytes per input symbol: 1 (u8 / int8) or 2 (int16) */
    int metric_bytes;      /* bytes per path metric: 1 or 2 */
    void* (*create)(const int* poly, int len);                 /* len = trellis steps incl. tail */
    int   (*init)(void* h, int starting_state);
    void  (*update)(void* h, const void* syms, int nsteps);    /* nsteps trellis steps */
    int   (*chainback)(void* h, uint8_t* data, unsigned nbits, unsigned endstate);
    void  (*destroy)(void* h);
    /* canonical decision vectors: S/8 bytes per step, bit n of the little-endian vector = decision of
     * new state n (SURVEY 8 notation). Copies steps [step0, step0+nsteps). */
    void  (*get_decisions)(void* h, int step0, int nsteps, uint8_t* out);
    /* current ("old") path metrics, S elements of metric_bytes each */
    void  (*get_metrics)(void* h, void* out);
} kvref_vtable;

/* Lookup by name; NULL if unknown. */
const kvref_vtable* kvref_find(const char* name);
int kvref_count(void);
const kvref_vtable* kvref_at(int i);

/* williamyang decoders take a numeric config; defaults follow src/viterbi_configs.h (soft8 / soft16).
 * Set before create; applies to wy_* decoders created afterwards (process-global, test use only).
 * kind: 0 = defaults from main.cpp (soft8 for u8, soft16 for u16), 1 = hard8 (u8 only). */
void kvref_wy_set_config_kind(int kind);
/* williamyang instances only: the sum of decoder_t::update<uint64_t>()'s return values since the last init. */
uint64_t kvref_wy_error_sum(void* h);

#ifdef __cplusplus
}
#endif

#ifdef __cplusplus
extern "C" {
#endif
/* bench.py's CPU arm: decode a pool of frames in a C loop for `seconds` (see registry.cpp). */
double kvref_time_decode(const kvref_vtable* vt, void* h, const void* syms_pool, int n_pool,
                         size_t frame_stride_bytes, int T, unsigned nbits, double seconds,
                         long long* frames_done, uint8_t* out);
#ifdef __cplusplus
}
#endif
