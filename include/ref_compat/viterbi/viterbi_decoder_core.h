/* viterbi/viterbi_decoder_core.h as served by include/ref_compat/ -- ViterbiDecoder_Core with the interface the
 * reference's callers use (williamyang_viterbi_lib/include/viterbi/viterbi_decoder_core.h:154-242: constructor,
 * set_traceback_length, get_traceback_length, reset, chainback, get_error; src/main.cpp:163-187,
 * examples/run_simple.cpp), implemented as a one-frame b200v_batch of the WY_U8 / WY_U16 codec: metrics and decision rows
 * live on the GPU, every call ends in the CUDA kernels of libb200viterbi.so. error_t selects the policy (uint8_t ->
 * viterbi_decoder_sse_u8.h arithmetic, uint16_t -> ..._u16.h). The CUDA device is $B200V_DEVICE (default 0). Failures
 * (no device, out of memory) print b200v_last_error() and abort: the reference's interface has no error channel. */
#pragma once
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#include "./viterbi_branch_table.h"
#include "viterbi/viterbi_decoder_config.h"      /* the reference's own plain struct (not shadowed here) */
#include "../../b200_viterbi.h"

template <size_t constraint_length, size_t code_rate, typename error_t, typename soft_t>
class ViterbiDecoder_Core {
public:
    static constexpr size_t K = constraint_length;
    static constexpr size_t R = code_rate;
    static constexpr size_t TOTAL_STATE_BITS = K - 1;
    static constexpr size_t NUMSTATES = size_t(1) << TOTAL_STATE_BITS;
    using BranchTable = ViterbiBranchTable<K, R, soft_t>;
    using Config = ViterbiDecoder_Config<error_t>;
    static_assert(sizeof(error_t) == sizeof(soft_t) && (sizeof(error_t) == 1 || sizeof(error_t) == 2),
                  "the CUDA path serves the u8 / int8 and u16 / int16 decoders");

    ViterbiDecoder_Core(const BranchTable& branch_table, const Config& config) : m_branch_table(branch_table), m_config(config) {
        set_traceback_length(0);
    }
    ~ViterbiDecoder_Core() { b200v_batch_destroy(m_batch); }
    ViterbiDecoder_Core(const ViterbiDecoder_Core&) = delete;
    ViterbiDecoder_Core& operator=(const ViterbiDecoder_Core&) = delete;

    /* decoded bits to keep decisions for, not counting the K-1 tail steps (viterbi_decoder_core.h:180-187) */
    void set_traceback_length(const size_t traceback_length) {
        b200v_batch_destroy(m_batch);
        m_traceback_length = traceback_length;
        b200v_wy_config cfg;
        cfg.soft_decision_high = (int)m_branch_table.soft_decision_high;
        cfg.soft_decision_low = (int)m_branch_table.soft_decision_low;
        cfg.soft_decision_max_error = (unsigned)m_config.soft_decision_max_error;
        cfg.initial_start_error = (unsigned)m_config.initial_start_error;
        cfg.initial_non_start_error = (unsigned)m_config.initial_non_start_error;
        cfg.renormalisation_threshold = (unsigned)m_config.renormalisation_threshold;
        const char* dev = getenv("B200V_DEVICE");
        m_batch = b200v_batch_create(sizeof(error_t) == 1 ? B200V_WY_U8 : B200V_WY_U16, (int)K, (int)R, m_branch_table.poly,
                                     (int)(traceback_length + TOTAL_STATE_BITS), 1, dev ? atoi(dev) : 0, &cfg);
        if (!m_batch) die("b200v_batch_create");
        m_error_seen = 0;
    }
    size_t get_traceback_length() const { return m_traceback_length; }

    error_t get_error(const size_t end_state = 0u) {
        std::vector<error_t> m(NUMSTATES);
        if (b200v_batch_get_metrics(m_batch, 0, m.data()) != B200V_OK) die("b200v_batch_get_metrics");
        return m[end_state % NUMSTATES];
    }
    void reset(const size_t starting_state = 0u) {
        if (b200v_batch_init(m_batch, (int)starting_state) != B200V_OK) die("b200v_batch_init");
        m_error_seen = 0;
    }
    void chainback(uint8_t* bytes_out, const size_t total_bits, const size_t end_state = 0u) {
        if (b200v_batch_chainback(m_batch, bytes_out, (total_bits + 7) / 8, (unsigned)total_bits, (unsigned)end_state) != B200V_OK)
            die("b200v_batch_chainback");
    }

    /* decoder_t::update of the x86/ headers lands here: N symbols = N / R trellis steps, continuing from the previous
     * call; returns what the reference's update returns, the sum of the minima its renormalisations subtracted
     * (viterbi_decoder_sse_u8.h:57-69) */
    template <typename sum_error_t>
    sum_error_t update_symbols(const soft_t* symbols, const size_t N) {
        if (b200v_batch_update(m_batch, symbols, N * sizeof(soft_t), (int)(N / R)) != B200V_OK) die("b200v_batch_update");
        uint64_t total = 0;
        if (b200v_batch_accumulated_error(m_batch, &total) != B200V_OK) die("b200v_batch_accumulated_error");
        const uint64_t delta = total - m_error_seen;
        m_error_seen = total;
        return (sum_error_t)delta;
    }
    b200v_batch* batch() { return m_batch; }

private:
    [[noreturn]] static void die(const char* what) {
        fprintf(stderr, "%s failed: %s\n", what, b200v_last_error());
        abort();
    }
    const BranchTable& m_branch_table;
    const Config m_config;
    b200v_batch* m_batch = nullptr;
    size_t m_traceback_length = 0;
    uint64_t m_error_seen = 0;
};
