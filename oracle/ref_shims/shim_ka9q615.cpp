/* TEST INFRASTRUCTURE ONLY. Wraps the unmodified reference TU ka9q_libfec_port/viterbi615_sse2.cpp (compiled from
 * /root/reference via -I, never copied) so tests can read its decision words and metrics. */
#include "kvref.h"
#include <stdlib.h>
#include <stdio.h>
#include <string.h>
#include <memory.h>
#include <limits.h>
#include <assert.h>
#include <stddef.h>
#include <stdint.h>
#include <immintrin.h>
/* LP64 fix (SURVEY 4 defect 1): the reference's decision_t uses 'unsigned long w[512]' assuming 32-bit long;
 * on Linux that makes chainback read the wrong words. The system headers are already included above, so
 * this define only affects the reference TU below. */
#define long int
#include "ka9q_libfec_port/viterbi615_sse2.cpp"
#undef long
#include "shim_c_style.h"
KVREF_DEFINE_C_STYLE(ka9q615, v615, 15, 6, 2,
    create_viterbi615_sse2, init_viterbi615_sse2, update_viterbi615_blk_sse2, chainback_viterbi615_sse2, delete_viterbi615_sse2,
    (void*)vp->old_metrics)
