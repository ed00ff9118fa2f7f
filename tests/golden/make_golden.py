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


def main():
    o.build_oracle()
    assert o.ref_available(), "oracle/_ref/libkvref.so missing: needs /root/reference"
    store = {}
    names = o.ref_names()
    for name in names:
        c = o.codec_by_name(name)
        nbytes = frame_bytes(c.K)
        T = nbytes * 8 + c.K - 1
        if name.startswith("spiral") and T % 2:
            raise AssertionError("spiral needs an even step count")
        for ci, sigma in enumerate((0.0, 0.7, 1.0)):
            data, syms = o.make_frame(c, nbytes, sigma, seed=0x601D + ci, frame=ci)
            r = o.Ref(name, T)
            r.init(0)
            r.update(syms, T)
            out, rc = r.chainback(nbytes * 8, 0, return_code=True)
            dec = r.decisions(0, T)
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
            r.close()
    store["names"] = np.array(names)
    np.savez_compressed(OUT, **store)
    print(f"wrote {OUT}: {len(names)} decoders x 3 cases, {os.path.getsize(OUT)} bytes")


if __name__ == "__main__":
    main()
