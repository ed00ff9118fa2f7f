/* TEST INFRASTRUCTURE ONLY. Boilerplate to wrap one ka9q / spiral reference TU (already #included above
 * this header by the shim .cpp, so its opaque struct and decision_t are visible) into a kvref_vtable. */
#pragma once
#include "kvref.h"
#include <string.h>

/* STRUCT      : instance struct tag (v27, spiral27, ...)
 * PFX_*       : the five reference functions
 * NAME        : registry name
 * KK, RR      : code parameters
 * MBYTES      : bytes per metric
 * METRIC_PTR  : expression (given `vp`) yielding pointer to the current metrics
 */
#define KVREF_DEFINE_C_STYLE(NAME, STRUCT, KK, RR, MBYTES, CREATE, INIT, UPDATE, CHAINBACK, DELETE, METRIC_PTR) \
    static void* NAME##_create(const int* poly, int len) { return (void*)CREATE(poly, len); }                   \
    static int NAME##_init(void* h, int s) { return INIT((STRUCT*)h, s); }                                       \
    static void NAME##_update(void* h, const void* syms, int n) {                                                \
        UPDATE((STRUCT*)h, (unsigned char*)syms, n);                                                             \
    }                                                                                                            \
    static int NAME##_chainback(void* h, uint8_t* data, unsigned nbits, unsigned endstate) {                     \
        return CHAINBACK((STRUCT*)h, data, nbits, endstate);                                                     \
    }                                                                                                            \
    static void NAME##_destroy(void* h) { DELETE((STRUCT*)h); }                                                  \
    static void NAME##_get_decisions(void* h, int step0, int nsteps, uint8_t* out) {                             \
        STRUCT* vp = (STRUCT*)h;                                                                                 \
        const size_t nb = ((size_t)1 << ((KK)-1)) / 8;                                                           \
        const decision_t* d = (const decision_t*)vp->decisions;                                                  \
        for (int t = 0; t < nsteps; t++) memcpy(out + (size_t)t * nb, &d[step0 + t], nb);                        \
    }                                                                                                            \
    static void NAME##_get_metrics(void* h, void* out) {                                                         \
        STRUCT* vp = (STRUCT*)h;                                                                                 \
        memcpy(out, (METRIC_PTR), ((size_t)1 << ((KK)-1)) * (MBYTES));                                           \
    }                                                                                                            \
    extern "C" const kvref_vtable kvref_vt_##NAME = {                                                            \
        #NAME, KK, RR, 1, MBYTES, NAME##_create, NAME##_init, NAME##_update, NAME##_chainback,                   \
        NAME##_destroy, NAME##_get_decisions, NAME##_get_metrics };
