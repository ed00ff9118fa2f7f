/* TEST INFRASTRUCTURE ONLY. Instantiates the reference's header-only williamyang decoders
 * (williamyang_viterbi_lib/include/viterbi/x86/*.h, used by src/main.cpp:149-237) for the (K,R)
 * pairs main.cpp benchmarks plus LTE K7R3, and exposes them through kvref_vtable.
 * Numeric configs follow src/viterbi_configs.h:22-65 (restated here because that header drags in
 * main.cpp-only types). Nothing from the reference is copied: headers are found via -I. */
#include "kvref.h"
#include <limits>
#include <memory>
#include <string.h>
#include "viterbi/viterbi_decoder_core.h"
#include "viterbi/x86/viterbi_decoder_sse_u8.h"
#include "viterbi/x86/viterbi_decoder_avx_u8.h"
#include "viterbi/x86/viterbi_decoder_sse_u16.h"
#include "viterbi/x86/viterbi_decoder_avx_u16.h"

static int g_cfg_kind = 0;
extern "C" void kvref_wy_set_config_kind(int kind) { g_cfg_kind = kind; }

template <typename soft_t, typename error_t> struct WyCfg {
    soft_t hi, lo;
    ViterbiDecoder_Config<error_t> cfg;
};

/* src/viterbi_configs.h:37-50 (soft8), :52-65 (hard8) */
static WyCfg<int8_t, uint8_t> make_cfg_u8(size_t R) {
    const int8_t hi = (g_cfg_kind == 1) ? 1 : 3, lo = (int8_t)-hi;
    const uint8_t max_error = uint8_t(uint8_t(hi - lo) * uint8_t(R));
    const uint8_t margin = uint8_t(max_error * uint8_t(g_cfg_kind == 1 ? 3u : 2u));
    ViterbiDecoder_Config<uint8_t> c;
    c.soft_decision_max_error = max_error;
    c.initial_start_error = 0;
    c.initial_non_start_error = margin;
    c.renormalisation_threshold = uint8_t(255 - margin);
    return { hi, lo, c };
}
/* src/viterbi_configs.h:22-35 (soft16) */
static WyCfg<int16_t, uint16_t> make_cfg_u16(size_t R) {
    const int16_t hi = 127, lo = -127;
    const uint16_t max_error = uint16_t(uint16_t(hi - lo) * uint16_t(R));
    const uint16_t margin = uint16_t(max_error * uint16_t(5u));
    ViterbiDecoder_Config<uint16_t> c;
    c.soft_decision_max_error = max_error;
    c.initial_start_error = 0;
    c.initial_non_start_error = margin;
    c.renormalisation_threshold = uint16_t(65535 - margin);
    return { hi, lo, c };
}

/* sum of update<>()'s return values since the last init: the renormalisation minima the reference's callers add to
 * get_error() (run_snr_ber.cpp:369-371, run_simple.cpp:78-79) */
struct WyBase { uint64_t err_sum = 0; };
extern "C" uint64_t kvref_wy_error_sum(void* h) { return ((WyBase*)h)->err_sum; }

template <size_t K, size_t R, typename soft_t, typename error_t, class decoder_t>
struct WyInst : WyBase {
    using Core = ViterbiDecoder_Core<K, R, error_t, soft_t>;
    using Table = ViterbiBranchTable<K, R, soft_t>;
    std::unique_ptr<Table> table;
    std::unique_ptr<Core> core;

    static void* create(const int* poly, int len) {
        auto* p = new WyInst();
        WyCfg<soft_t, error_t> c;
        if constexpr (sizeof(error_t) == 1) c = make_cfg_u8(R); else c = make_cfg_u16(R);
        p->table = std::make_unique<Table>(poly, c.hi, c.lo);
        p->core = std::make_unique<Core>(*p->table, c.cfg);
        /* len = trellis steps incl. tail, as main.cpp passes to the third-party decoders */
        p->core->set_traceback_length(size_t(len) - (K - 1));
        return p;
    }
    static int init(void* h, int s) { ((WyInst*)h)->core->reset(size_t(s)); ((WyInst*)h)->err_sum = 0; return 0; }
    static void update(void* h, const void* syms, int nsteps) {
        auto* p = (WyInst*)h;
        p->err_sum += decoder_t::template update<uint64_t>(*p->core, (const soft_t*)syms, size_t(nsteps) * R);
    }
    static int chainback(void* h, uint8_t* data, unsigned nbits, unsigned endstate) {
        ((WyInst*)h)->core->chainback(data, nbits, endstate);
        return 0;
    }
    static void destroy(void* h) { delete (WyInst*)h; }
    static void get_decisions(void* h, int step0, int nsteps, uint8_t* out) {
        auto* p = (WyInst*)h;
        const size_t nb = (size_t(1) << (K - 1)) / 8;
        for (int t = 0; t < nsteps; t++) memcpy(out + size_t(t) * nb, p->core->m_decisions[step0 + t], nb);
    }
    static void get_metrics(void* h, void* out) {
        auto* p = (WyInst*)h;
        memcpy(out, p->core->m_metrics.get_old(), (size_t(1) << (K - 1)) * sizeof(error_t));
    }
};

#define WY_VT(NAME, KK, RR, SOFT, ERR, DEC)                                                         \
    extern "C" const kvref_vtable kvref_vt_##NAME = {                                                \
        #NAME, KK, RR, (int)sizeof(SOFT), (int)sizeof(ERR),                                          \
        WyInst<KK, RR, SOFT, ERR, DEC<KK, RR>>::create, WyInst<KK, RR, SOFT, ERR, DEC<KK, RR>>::init, \
        WyInst<KK, RR, SOFT, ERR, DEC<KK, RR>>::update, WyInst<KK, RR, SOFT, ERR, DEC<KK, RR>>::chainback, \
        WyInst<KK, RR, SOFT, ERR, DEC<KK, RR>>::destroy, WyInst<KK, RR, SOFT, ERR, DEC<KK, RR>>::get_decisions, \
        WyInst<KK, RR, SOFT, ERR, DEC<KK, RR>>::get_metrics };

#define WY_ALL(KK, RR)                                                                    \
    WY_VT(wy_sse_u8_k##KK##r##RR, KK, RR, int8_t, uint8_t, ViterbiDecoder_SSE_u8)          \
    WY_VT(wy_avx_u8_k##KK##r##RR, KK, RR, int8_t, uint8_t, ViterbiDecoder_AVX_u8)          \
    WY_VT(wy_sse_u16_k##KK##r##RR, KK, RR, int16_t, uint16_t, ViterbiDecoder_SSE_u16)      \
    WY_VT(wy_avx_u16_k##KK##r##RR, KK, RR, int16_t, uint16_t, ViterbiDecoder_AVX_u16)

WY_ALL(7, 2)
WY_ALL(7, 3)
WY_ALL(7, 4)
WY_ALL(9, 2)
WY_ALL(9, 4)
WY_ALL(15, 6)
WY_ALL(24, 2)
/* common_codes.h:24 "Basic K=5 R=1/2": the smallest K a SIMD decoder accepts (SSE u16; the others need K >= 6 or 7) */
WY_VT(wy_sse_u16_k5r2, 5, 2, int16_t, uint16_t, ViterbiDecoder_SSE_u16)
