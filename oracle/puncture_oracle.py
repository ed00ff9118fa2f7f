"""TEST INFRASTRUCTURE ONLY -- numpy/pure-Python restatement of the reference's puncturing helpers
(williamyang_viterbi_lib/examples/helpers/puncture_code_helpers.h). Pinned against the compiled helpers through
oracle/_ref (tests/test_puncture.py). Never imported by the product."""
import numpy as np

# DAB fast information channel codes (run_punctured_decoder.cpp:39-76, ETSI EN 300 401 table 13 rows 16 and 15)
PI_16 = np.array([1, 1, 1, 0] * 8, dtype=np.uint8)
PI_15 = np.array([1, 1, 1, 0] * 7 + [1, 1, 0, 0], dtype=np.uint8)
PI_X = np.array([1, 1, 0, 0] * 6, dtype=np.uint8)


def depuncture(punctured, code, requested, fill, R):
    """decode_punctured_symbols (:17-55), symbol stream only: literal loop. Returns (symbols, consumed).
    The reference produces symbols R at a time, so it overshoots `requested` up to the next multiple of R."""
    punctured = np.asarray(punctured)
    out = []
    ip = ic = 0
    while len(out) < requested:
        for _ in range(R):
            if code[ic]:
                if ip >= len(punctured):
                    return np.array(out, dtype=punctured.dtype), ip
                out.append(punctured[ip])
                ip += 1
            else:
                out.append(fill)
            ic = (ic + 1) % len(code)
    return np.array(out, dtype=punctured.dtype), ip


def puncture(full_rate, code):
    """encode_punctured_data / _tail (:57-147): keep full-rate symbol i where code[i % L] is set (index from 0)."""
    full_rate = np.asarray(full_rate)
    keep = np.resize(np.asarray(code, dtype=bool), full_rate.size)
    return full_rate[keep]


def dab_fic_segments(R=4):
    """(code, full-rate symbols) per segment of a DAB FIC frame: 21 x 32 bits PI_16, 3 x 32 bits PI_15, 6 tail bits
    PI_X (run_punctured_decoder.cpp:194-285)."""
    return [(PI_16, 32 * R * 21), (PI_15, 32 * R * 3), (PI_X, 24)]
