/* viterbi/x86/viterbi_decoder_sse_u16.h as served by include/ref_compat/ -- ViterbiDecoder_SSE_u16<K,R>::update with the
 * signature of the reference (williamyang_viterbi_lib/include/viterbi/x86/viterbi_decoder_sse_u16.h:45-70), forwarding to
 * the CUDA path through ViterbiDecoder_Core (the SSE and AVX decoders of the reference give identical results; so do
 * these). */
#pragma once
#include "../viterbi_decoder_core.h"

template <size_t constraint_length, size_t code_rate>
class ViterbiDecoder_SSE_u16 {
public:
    using Base = ViterbiDecoder_Core<constraint_length, code_rate, uint16_t, int16_t>;
    static constexpr bool is_valid = constraint_length >= 3;
    template <typename sum_error_t>
    static sum_error_t update(Base& base, const int16_t* symbols, const size_t N) {
        return base.template update_symbols<sum_error_t>(symbols, N);
    }
};
