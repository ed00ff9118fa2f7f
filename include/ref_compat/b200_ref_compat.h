/* b200_ref_compat.h -- the reference's per-codec decoder API (C++ linkage, opaque instance structs), served by
 * libb200viterbi_compat.so on top of the batch C-ABI of b200_viterbi.h with a one-frame batch per instance.
 *
 * Each block of five functions replaces the same-named functions of
 *   ka9q_libfec_port/viterbi{27,29,615,224}_sse2.h:3-8   (create_/init_/chainback_/delete_/update_..._blk_)
 *   spiral/spiral{27,29,47,49,615}.h:3-9                 (create_/init_/update_/chainback_/delete_)
 * so that src/ka9q_interface.h:58-61 and src/spiral_interface.h:58-62 -- and with them src/main.cpp -- compile
 * and link unmodified when include/ref_compat/ is on the include path instead of the reference's decoder
 * directories. Signatures, argument meaning and return values follow the reference: create returns NULL on
 * failure (b200v_last_error() says why), init/chainback return 0 (chainback_viterbi615_sse2 returns the
 * end-state path metric, viterbi615_sse2.cpp:90), -1 on a CUDA error; update is void and synchronous.
 */
#ifndef B200_REF_COMPAT_H
#define B200_REF_COMPAT_H

#define B200V_COMPAT_DECLARE(TAG, CREATE, INIT, UPDATE, CHAINBACK, DELETE)                                  \
    struct TAG;                                                                                             \
    TAG* CREATE(const int* poly, int len);                                                                  \
    int INIT(TAG* p, int starting_state);                                                                   \
    void UPDATE(TAG* p, unsigned char* syms, int nbits);                                                    \
    int CHAINBACK(TAG* p, unsigned char* data, unsigned int nbits, unsigned int endstate);                  \
    void DELETE(TAG* p);

#define B200V_COMPAT_KA9Q(N) \
    B200V_COMPAT_DECLARE(v##N, create_viterbi##N##_sse2, init_viterbi##N##_sse2, update_viterbi##N##_blk_sse2, \
                         chainback_viterbi##N##_sse2, delete_viterbi##N##_sse2)
#define B200V_COMPAT_SPIRAL(N) \
    B200V_COMPAT_DECLARE(spiral##N, create_spiral##N, init_spiral##N, update_spiral##N, chainback_spiral##N, \
                         delete_spiral##N)

B200V_COMPAT_KA9Q(27)
B200V_COMPAT_KA9Q(29)
B200V_COMPAT_KA9Q(615)
B200V_COMPAT_KA9Q(224)
B200V_COMPAT_SPIRAL(27)
B200V_COMPAT_SPIRAL(29)
B200V_COMPAT_SPIRAL(47)
B200V_COMPAT_SPIRAL(49)
B200V_COMPAT_SPIRAL(615)

/* CUDA device the compat instances are created on (default 0, or the B200V_DEVICE environment variable). */
void b200v_compat_set_device(int device);

#endif
