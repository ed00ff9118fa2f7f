/* TEST INFRASTRUCTURE ONLY. Name -> vtable registry for the compiled reference decoders. */
#include "kvref.h"
#include <string.h>

#define DECL(N) extern "C" const kvref_vtable kvref_vt_##N;
#define WY4(KR) DECL(wy_sse_u8_##KR) DECL(wy_avx_u8_##KR) DECL(wy_sse_u16_##KR) DECL(wy_avx_u16_##KR)
DECL(ka9q27) DECL(ka9q29) DECL(ka9q615) DECL(ka9q224)
DECL(spiral27) DECL(spiral29) DECL(spiral47) DECL(spiral49) DECL(spiral615)
WY4(k7r2) WY4(k7r3) WY4(k7r4) WY4(k9r2) WY4(k9r4) WY4(k15r6) WY4(k24r2)
#undef DECL
#define DECL(N) &kvref_vt_##N,
static const kvref_vtable* const g_all[] = {
    DECL(ka9q27) DECL(ka9q29) DECL(ka9q615) DECL(ka9q224)
    DECL(spiral27) DECL(spiral29) DECL(spiral47) DECL(spiral49) DECL(spiral615)
    WY4(k7r2) WY4(k7r3) WY4(k7r4) WY4(k9r2) WY4(k9r4) WY4(k15r6) WY4(k24r2)
};
extern "C" int kvref_count(void) { return (int)(sizeof(g_all) / sizeof(g_all[0])); }
extern "C" const kvref_vtable* kvref_at(int i) { return (i >= 0 && i < kvref_count()) ? g_all[i] : 0; }
extern "C" const kvref_vtable* kvref_find(const char* name) {
    for (int i = 0; i < kvref_count(); i++) if (!strcmp(g_all[i]->name, name)) return g_all[i];
    return 0;
}
