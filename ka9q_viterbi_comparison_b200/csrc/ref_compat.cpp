// ref_compat.cpp -- the reference's per-codec decoder functions (include/ref_compat/b200_ref_compat.h) as thin
// forwards to the batch C-ABI with one frame per instance. Built into libb200viterbi_compat.so (C++ linkage,
// like the reference's own objects) so that src/main.cpp links against it unmodified.
#include "../../include/ref_compat/b200_ref_compat.h"
#include "../../include/b200_viterbi.h"

#include <cstdio>
#include <cstdlib>

static int g_device = -1;

void b200v_compat_set_device(int device) { g_device = device; }

static int compat_device() {
    if (g_device < 0) {
        const char* e = std::getenv("B200V_DEVICE");
        g_device = e ? std::atoi(e) : 0;
    }
    return g_device;
}

struct CompatInstance {
    b200v_batch* h;
    int R;
    bool failed;        // an update failed: the reference's update is void, so the failure surfaces at the next chainback
};

static CompatInstance* compat_create(int codec, int K, int R, const int* poly, int len) {
    b200v_batch* h = b200v_batch_create(codec, K, R, poly, len, 1, compat_device(), nullptr);
    if (!h) return nullptr;
    return new CompatInstance{ h, R, false };
}
static int compat_init(CompatInstance* p, int starting_state) {
    if (!p) return -1;
    p->failed = false;
    if (b200v_batch_init(p->h, starting_state) != B200V_OK) return -1;
    return b200v_batch_sync(p->h) == B200V_OK ? 0 : -1;       // callers time each call on the host
}
static void compat_update(CompatInstance* p, unsigned char* syms, int nbits) {
    if (!p || nbits <= 0) return;
    if (b200v_batch_update(p->h, syms, (size_t)nbits * (size_t)p->R, nbits) != B200V_OK) {
        // the reference's signature has no error channel: say why, and make the following chainback fail instead of
        // tracing back stale decisions
        fprintf(stderr, "b200v compat: update of %d steps failed: %s\n", nbits, b200v_last_error());
        p->failed = true;
    }
}
static int compat_chainback(CompatInstance* p, unsigned char* data, unsigned nbits, unsigned endstate) {
    if (!p || p->failed) return -1;
    if (b200v_batch_chainback(p->h, data, (nbits + 7) / 8, nbits, endstate) != B200V_OK) return -1;
    int rc = 0;
    if (b200v_batch_chainback_results(p->h, &rc) != B200V_OK) return -1;
    return rc;
}
static void compat_delete(CompatInstance* p) {
    if (!p) return;
    b200v_batch_destroy(p->h);
    delete p;
}

#define B200V_COMPAT_DEFINE(TAG, CODEC, K, R, CREATE, INIT, UPDATE, CHAINBACK, DELETE)                          \
    struct TAG : CompatInstance {};                                                                             \
    TAG* CREATE(const int* poly, int len) { return static_cast<TAG*>(compat_create(CODEC, K, R, poly, len)); } \
    int INIT(TAG* p, int starting_state) { return compat_init(p, starting_state); }                            \
    void UPDATE(TAG* p, unsigned char* syms, int nbits) { compat_update(p, syms, nbits); }                     \
    int CHAINBACK(TAG* p, unsigned char* data, unsigned int nbits, unsigned int endstate) {                    \
        return compat_chainback(p, data, nbits, endstate);                                                     \
    }                                                                                                           \
    void DELETE(TAG* p) { compat_delete(p); }

#define B200V_COMPAT_DEFINE_KA9Q(N, CODEC, K, R)                                                                \
    B200V_COMPAT_DEFINE(v##N, CODEC, K, R, create_viterbi##N##_sse2, init_viterbi##N##_sse2,                    \
                        update_viterbi##N##_blk_sse2, chainback_viterbi##N##_sse2, delete_viterbi##N##_sse2)
#define B200V_COMPAT_DEFINE_SPIRAL(N, CODEC, K, R)                                                              \
    B200V_COMPAT_DEFINE(spiral##N, CODEC, K, R, create_spiral##N, init_spiral##N, update_spiral##N,             \
                        chainback_spiral##N, delete_spiral##N)

B200V_COMPAT_DEFINE_KA9Q(27, B200V_KA9Q27, 7, 2)       // ka9q_libfec_port/viterbi27_sse2.cpp:42-175
B200V_COMPAT_DEFINE_KA9Q(29, B200V_KA9Q29, 9, 2)       // viterbi29_sse2.cpp
B200V_COMPAT_DEFINE_KA9Q(615, B200V_KA9Q615, 15, 6)    // viterbi615_sse2.cpp
B200V_COMPAT_DEFINE_KA9Q(224, B200V_KA9Q224, 24, 2)    // viterbi224_sse2.cpp
B200V_COMPAT_DEFINE_SPIRAL(27, B200V_SPIRAL27, 7, 2)   // spiral/spiral27.cpp:53-386
B200V_COMPAT_DEFINE_SPIRAL(29, B200V_SPIRAL29, 9, 2)
B200V_COMPAT_DEFINE_SPIRAL(47, B200V_SPIRAL47, 7, 4)
B200V_COMPAT_DEFINE_SPIRAL(49, B200V_SPIRAL49, 9, 4)
B200V_COMPAT_DEFINE_SPIRAL(615, B200V_SPIRAL615, 15, 6)
