/* TEST INFRASTRUCTURE ONLY. Name -> vtable registry for the compiled reference decoders. */
#include "kvref.h"
#include <string.h>

#define DECL(N) extern "C" const kvref_vtable kvref_vt_##N;
#define WY4(KR) DECL(wy_sse_u8_##KR) DECL(wy_avx_u8_##KR) DECL(wy_sse_u16_##KR) DECL(wy_avx_u16_##KR)
DECL(ka9q27) DECL(ka9q29) DECL(ka9q615) DECL(ka9q224)
DECL(spiral27) DECL(spiral29) DECL(spiral47) DECL(spiral49) DECL(spiral615)
WY4(k7r2) WY4(k7r3) WY4(k7r4) WY4(k9r2) WY4(k9r4) WY4(k15r6) WY4(k24r2)
DECL(wy_sse_u16_k5r2)
#undef DECL
#define DECL(N) &kvref_vt_##N,
static const kvref_vtable* const g_all[] = {
    DECL(ka9q27) DECL(ka9q29) DECL(ka9q615) DECL(ka9q224)
    DECL(spiral27) DECL(spiral29) DECL(spiral47) DECL(spiral49) DECL(spiral615)
    WY4(k7r2) WY4(k7r3) WY4(k7r4) WY4(k9r2) WY4(k9r4) WY4(k15r6) WY4(k24r2)
    DECL(wy_sse_u16_k5r2)
};
extern "C" int kvref_count(void) { return (int)(sizeof(g_all) / sizeof(g_all[0])); }
extern "C" const kvref_vtable* kvref_at(int i) { return (i >= 0 && i < kvref_count()) ? g_all[i] : 0; }
extern "C" const kvref_vtable* kvref_find(const char* name) {
    for (int i = 0; i < kvref_count(); i++) if (!strcmp(g_all[i]->name, name)) return g_all[i];
    return 0;
}

/* Timed decode loop for bench.py's CPU arm: init + update(T steps) + chainback(nbits) over a pool of frames,
 * repeated until `seconds` have elapsed. Runs entirely in C so that one Python thread per core measures the
 * reference decoder and not the interpreter. Returns the elapsed seconds; *frames_done = frames decoded. */
#include <chrono>
extern "C" double kvref_time_decode(const kvref_vtable* vt, void* h, const void* syms_pool, int n_pool,
                                    size_t frame_stride_bytes, int T, unsigned nbits, double seconds,
                                    long long* frames_done, uint8_t* out) {
    using clk = std::chrono::steady_clock;
    const clk::time_point t0 = clk::now();
    long long n = 0;
    double dt = 0.0;
    do {
        for (int f = 0; f < n_pool; f++) {
            vt->init(h, 0);
            vt->update(h, (const char*)syms_pool + (size_t)f * frame_stride_bytes, T);
            vt->chainback(h, out, nbits, 0);
            n++;
        }
        dt = std::chrono::duration<double>(clk::now() - t0).count();
    } while (dt < seconds);
    if (frames_done) *frames_done = n;
    return dt;
}
