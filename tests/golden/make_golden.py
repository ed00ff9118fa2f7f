"""Generate tests/golden/reference_vectors.npz from the UNMODIFIED reference (oracle/_ref/libkvref.so,
compiled from /root/reference by oracle/Makefile). Run in the container that has /root/reference:

    python tests/golden/make_golden.py

For every reference decoder and three channel conditions it stores the exact input symbols, the decoded
bytes, the chainback return value, a SHA-256 of the final path metrics (verbatim for K<=9) and a SHA-256 of the per-step decision vectors
(canonical S/8 bytes per step). The reference ships no golden vectors of its own, so these
are the known-answer tests for both oracle/viterbi_oracle.c and the CUDA path.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import _oracle as o  # noqa: E402

OUT = os.path.join(HERE, "reference_vectors.npz")


def frame_bytes(K):
    return {24: 1, 15: 24}.get(K, 96)


K24_RENORM_BYTES = 30     # K=24 case 3: 240 + 23 = 263 steps of erased symbols, enough for every K=24 policy to renormalise


def cases(c):
    """(case index, data, symbols): three channel conditions; K=24 decoders get a fourth, renormalising frame -- erased
    symbols (the middle of the soft range) cost every path the same, so new[0] climbs to the threshold of
    viterbi224_sse2.cpp:226-246 / viterbi_decoder_sse_u8.h:63-65,127-154 within 263 steps."""
    nbytes = frame_bytes(c.K)
    for ci, sigma in enumerate((0.0, 0.7, 1.0)):
        data, syms = o.make_frame(c, nbytes, sigma, seed=0x601D + ci, frame=ci)
        yield ci, data, syms
    if c.K == 24:
        lo, hi = o.soft_range(c)
        T = K24_RENORM_BYTES * 8 + c.K - 1
        yield 3, np.zeros(K24_RENORM_BYTES, dtype=np.uint8), np.full(T * c.R, (lo + hi + 1) // 2, dtype=o.sym_dtype(c))


def generate(names=None):
    """name -> array dict of the known answers, computed with the compiled reference"""
    store = {}
    names = names or o.ref_names()
    for name in names:
        c = o.codec_by_name(name)
        for ci, data, syms in cases(c):
            nbytes = data.size
            T = nbytes * 8 + c.K - 1
            if name.startswith("spiral") and T % 2:
                raise AssertionError("spiral needs an even step count")
            r = o.Ref(name, T)
            r.init(0)
            r.update(syms, T)
            out, rc = r.chainback(nbytes * 8, 0, return_code=True)
            dec = r.decisions(0, T)
            if ci == 3:     # the frame exists to renormalise: un-renormalised metrics would sit far above these bounds
                top = int(r.metrics().astype(np.int64 if c.policy != 2 else np.int16).max())
                assert top < {2: 0, 6: 231, 7: 62995}[c.policy], (name, top)
            key = f"{name}/{ci}"
            store[key + "/data"] = data
            store[key + "/syms"] = syms
            store[key + "/out"] = out
            store[key + "/rc"] = np.int64(rc)
            met = r.metrics()
            store[key + "/metrics_sha256"] = np.frombuffer(hashlib.sha256(met.tobytes()).digest(), dtype=np.uint8)
            if c.K <= 9:
                store[key + "/metrics"] = met
            store[key + "/dec_sha256"] = np.frombuffer(hashlib.sha256(dec.tobytes()).digest(), dtype=np.uint8)
            # the first and last 4 decision rows verbatim, for debuggability
            if c.K <= 15:
                store[key + "/dec_head"] = dec[:4].copy()
                store[key + "/dec_tail"] = dec[-4:].copy()
            if name.startswith("wy_"):      # decoder_t::update<uint64_t>()'s return value (sum of renormalisation minima)
                store[key + "/error_sum"] = np.uint64(r.renorm_sum())
            r.close()
    store["names"] = np.array(names)
    return store


def main():
    o.build_oracle()
    assert o.ref_available(), "oracle/_ref/libkvref.so missing: needs /root/reference"
    store = generate()
    np.savez_compressed(OUT, **store)
    print(f"wrote {OUT}: {len(store['names'])} decoders, {len(store)} arrays, {os.path.getsize(OUT)} bytes")


if __name__ == "__main__":
    main()
