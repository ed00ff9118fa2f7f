/* TEST INFRASTRUCTURE ONLY. Wraps the unmodified reference TU spiral/spiral29.cpp (compiled from
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
#include <pmmintrin.h>
/* 'renormalize' is a non-static inline with a different body in every spiral TU (ODR hazard when all are
 * linked into one library): give each TU its own name. */
#define renormalize renormalize_spiral29
#include "spiral/spiral29.cpp"
#include "shim_c_style.h"
KVREF_DEFINE_C_STYLE(spiral29, spiral29, 9, 2, 1,
    create_spiral29, init_spiral29, update_spiral29, chainback_spiral29, delete_spiral29,
    (void*)vp->old_metrics)
