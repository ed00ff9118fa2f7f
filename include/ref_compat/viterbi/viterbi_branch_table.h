/* viterbi/viterbi_branch_table.h as served by include/ref_compat/ -- the class the reference's callers construct before
 * a decoder (williamyang_viterbi_lib/include/viterbi/viterbi_branch_table.h:19-55; src/main.cpp:163), on top of the
 * C-ABI of b200_viterbi.h: it only records the generator polynomials and the two soft-decision levels; the per-state
 * code-word table itself is built on the device side by b200v_batch_create. Put include/ref_compat/ BEFORE the
 * reference's include directory and src/main.cpp's williamyang rows run on the CUDA path unmodified. */
#pragma once
#include <stddef.h>
#include <stdint.h>

template <size_t constraint_length, size_t code_rate, typename soft_t>
class ViterbiBranchTable {
public:
    static constexpr size_t K = constraint_length;
    static constexpr size_t R = code_rate;
    template <typename code_t>
    ViterbiBranchTable(const code_t* G, const soft_t high, const soft_t low) : soft_decision_high(high), soft_decision_low(low) {
        static_assert(K > 1u && R > 1u, "constraint length and code rate");
        for (size_t i = 0; i < R; i++) poly[i] = (int)G[i];
    }
    int poly[R];
    const soft_t soft_decision_high;
    const soft_t soft_decision_low;
};
