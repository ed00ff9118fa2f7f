/* TEST INFRASTRUCTURE ONLY. Runs the reference's own puncturing helpers
 * (williamyang_viterbi_lib/examples/helpers/puncture_code_helpers.h, found via -I) so that the depuncturing
 * restatement in oracle/puncture_oracle.py and the CUDA front-end can be pinned against them:
 *   kvref_depuncture_*  : decode_punctured_symbols (:17-55) with a recording stand-in for the decoder -> the
 *                         full-rate symbol stream the reference feeds to decoder_t::update
 *   kvref_puncture_data : encode_punctured_data (:57-101) on a real ConvolutionalEncoder_ShiftRegister */
#include "kvref.h"
#include <vector>
#include "helpers/puncture_code_helpers.h"
#include "viterbi/convolutional_encoder_shift_register.h"

template <typename soft_t>
struct Recorder {
    static std::vector<soft_t>& sink() { static thread_local std::vector<soft_t> v; return v; }
    template <typename sum_t, class core_t>
    static sum_t update(core_t&, const soft_t* symbols, const size_t n) {
        sink().insert(sink().end(), symbols, symbols + n);
        return sum_t(0);
    }
};

template <size_t R, typename error_t, typename soft_t>
static long depuncture_impl(const soft_t* punctured, size_t total, const bool* code, size_t code_len, size_t requested,
                            soft_t fill, soft_t* out) {
    constexpr size_t K = 7;
    ViterbiBranchTable<K, R, soft_t> table((const int[]){ 109, 79, 83, 109, 1, 1, 1, 1 }, soft_t(1), soft_t(-1));
    ViterbiDecoder_Config<error_t> cfg;
    cfg.soft_decision_max_error = 1; cfg.initial_start_error = 0; cfg.initial_non_start_error = 1;
    cfg.renormalisation_threshold = 100;
    ViterbiDecoder_Core<K, R, error_t, soft_t> core(table, cfg);
    Recorder<soft_t>::sink().clear();
    const PuncturedDecodeResult res = decode_punctured_symbols<Recorder<soft_t>, K, R, error_t, soft_t, soft_t>(
        core, fill, punctured, total, code, code_len, requested);
    const std::vector<soft_t>& v = Recorder<soft_t>::sink();
    for (size_t i = 0; i < v.size(); i++) out[i] = v[i];
    return (long)res.index_punctured_symbol;
}

/* R must be 2 or 4 (the loop advances R symbols per decoder call; requested should be a multiple of R) */
extern "C" long kvref_depuncture_s8(int R, const int8_t* punctured, size_t total, const uint8_t* code, size_t code_len,
                                    size_t requested, int fill, int8_t* out) {
    std::vector<char> c(code, code + code_len);
    const bool* b = reinterpret_cast<const bool*>(c.data());
    if (R == 2) return depuncture_impl<2, uint8_t, int8_t>(punctured, total, b, code_len, requested, (int8_t)fill, out);
    if (R == 4) return depuncture_impl<4, uint8_t, int8_t>(punctured, total, b, code_len, requested, (int8_t)fill, out);
    return -1;
}
extern "C" long kvref_depuncture_s16(int R, const int16_t* punctured, size_t total, const uint8_t* code, size_t code_len,
                                     size_t requested, int fill, int16_t* out) {
    std::vector<char> c(code, code + code_len);
    const bool* b = reinterpret_cast<const bool*>(c.data());
    if (R == 2) return depuncture_impl<2, uint16_t, int16_t>(punctured, total, b, code_len, requested, (int16_t)fill, out);
    if (R == 4) return depuncture_impl<4, uint16_t, int16_t>(punctured, total, b, code_len, requested, (int16_t)fill, out);
    return -1;
}

/* encode_punctured_data + encode_punctured_tail for one code: returns the number of transmitted symbols */
extern "C" long kvref_puncture_encode_s8(int K, int R, const int* poly, const uint8_t* data, size_t nbytes,
                                         const uint8_t* code, size_t code_len, int with_tail, const uint8_t* tail_code,
                                         size_t tail_code_len, int hi, int lo, int8_t* out, size_t max_out) {
    ConvolutionalEncoder_ShiftRegister<uint32_t> enc((size_t)K, (size_t)R, poly);
    std::vector<char> c(code, code + code_len), t(tail_code, tail_code + tail_code_len);
    size_t n = encode_punctured_data<int8_t>(&enc, data, nbytes, out, max_out, reinterpret_cast<const bool*>(c.data()),
                                             code_len, (int8_t)hi, (int8_t)lo);
    if (with_tail)
        n += encode_punctured_tail<int8_t>(&enc, out + n, max_out - n, reinterpret_cast<const bool*>(t.data()),
                                           tail_code_len, (int8_t)hi, (int8_t)lo);
    return (long)n;
}
