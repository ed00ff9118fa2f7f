/* TEST INFRASTRUCTURE ONLY. Wraps the unmodified reference TU ka9q_libfec_port/viterbi29_sse2.cpp (compiled from
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

#include "ka9q_libfec_port/viterbi29_sse2.cpp"
#include "shim_c_style.h"
KVREF_DEFINE_C_STYLE(ka9q29, v29, 9, 2, 1,
    create_viterbi29_sse2, init_viterbi29_sse2, update_viterbi29_blk_sse2, chainback_viterbi29_sse2, delete_viterbi29_sse2,
    (void*)vp->old_metrics)
