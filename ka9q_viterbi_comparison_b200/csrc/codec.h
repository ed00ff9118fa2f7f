// codec.h -- host/device shared description of one decoder configuration.
// Policy numbers follow the table in SURVEY.md 8(a); each constant cites where the reference fixes it.
#pragma once
#include <stdint.h>

namespace b200v {

enum Policy : int {
    P_KA9Q_U8_MOD = 0,   // ka9q_libfec_port/viterbi27_sse2.cpp:119-175, viterbi29_sse2.cpp:108-164
    P_KA9Q_S16_615 = 1,  // ka9q_libfec_port/viterbi615_sse2.cpp:104-191
    P_KA9Q_S16_224 = 2,  // ka9q_libfec_port/viterbi224_sse2.cpp:135-258
    P_SPIRAL_R2 = 3,     // spiral/spiral27.cpp:130-381, spiral29.cpp
    P_SPIRAL_R4 = 4,     // spiral/spiral47.cpp, spiral49.cpp
    P_SPIRAL_615 = 5,    // spiral/spiral615.cpp:130-411
    P_WY_U8 = 6,         // viterbi/x86/viterbi_decoder_sse_u8.h:45-154
    P_WY_U16 = 7,        // viterbi/x86/viterbi_decoder_sse_u16.h:45-177
    P_COUNT = 8
};

constexpr int MAX_R = 8;

struct CodecParams {
    int policy;
    int K, R;
    int S;                  // 2^(K-1) states
    int poly[MAX_R];
    int renorm_threshold;
    int init_start, init_other;
    int max_error;          // C: BM + complement
    int soft_hi, soft_lo;   // wy branch table values
    int sym_bytes;          // 1 or 2
    int metric_bytes;       // 1 or 2
};

// How one trellis step's S decision bits are laid out in memory by a kernel family.
enum DecisionLayout : int {
    // generic kernel: row of max(2,S/32) u32 words; state n -> word ((n>>6)<<1)|(n&1), bit (n>>1)&31
    // (two __ballot_sync results per 32 butterflies), frame-major: dec[frame][step][word]
    DL_BALLOT_FRAME_MAJOR = 0,
    // register-resident K=7 kernel: one u64 per (step, frame), frame-interleaved: dec[step][frame];
    // bit positions given by a per-phase permutation table (see acs_k7.cuh)
    DL_K7_INTERLEAVED = 1,
    // group family (K=9, K=15; acs_grp.cuh): one u64 per (step, frame, thread-of-frame), frame-interleaved:
    // dec[(step*n_frames + frame)*TPF + tx]; bit positions depend on the phase (step mod 5) of the row
    DL_GRP_INTERLEAVED = 2,
    // K=24 family (acs_k24.cu): the canonical vector itself, frame-major: bit n of dec[frame][step][S/32 words]
    DL_CANONICAL = 3,
    // small-K register kernel (acs_smallk.cu, K <= 6): the canonical vector in one u32 per (step, frame),
    // frame-interleaved: bit n of dec[step][frame]
    DL_SMALLK_INTERLEAVED = 4,
};

static inline bool policy_is_spiral(int p) { return p == P_SPIRAL_R2 || p == P_SPIRAL_R4 || p == P_SPIRAL_615; }
static inline bool policy_is_wy(int p) { return p == P_WY_U8 || p == P_WY_U16; }

}  // namespace b200v
