/* TEST INFRASTRUCTURE ONLY. Wraps the unmodified reference TU spiral/spiral47.cpp (compiled from
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
#define renormalize renormalize_spiral47
#include "spiral/spiral47.cpp"
#include "shim_c_style.h"
KVREF_DEFINE_C_STYLE(spiral47, spiral47, 7, 4, 1,
    create_spiral47, init_spiral47, update_spiral47, chainback_spiral47, delete_spiral47,
    (void*)vp->old_metrics)
