#!/usr/bin/env python
"""cpu_table.py -- the CPU table of SURVEY 8d: every CPU decoder the reference has for each benchmarked code (ka9q port,
spiral, williamyang sse_u8 / avx_u8 / sse_u16 / avx_u16), one instance per thread, threads pinned, built with the
reference preset's flags (oracle/Makefile) -- per core (one thread) and all cores. Runs the UNMODIFIED reference objects
of oracle/_ref (test infrastructure; this tool is the other measurement that may execute them, like bench.py's
cpu_baseline leg). One JSON line per decoder.

    python tools/cpu_table.py > gpurun_out/cpu_table.jsonl
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=1.5)
    ap.add_argument("--frame-bits", type=int, default=2048)
    a = ap.parse_args()
    o = bench.checker_module()
    if not o.ref_available():
        sys.exit("oracle/_ref is not built: the CPU table times the compiled reference")
    ncpu = len(bench.usable_cpus())
    for K, R in ((7, 2), (7, 4), (9, 2), (9, 4), (15, 6), (24, 2)):
        fb = 64 if K == 24 else a.frame_bits
        sigma = bench.default_sigma(K)
        best = None
        rows = []
        for name in bench.cpu_family(K, R):
            try:
                one = bench.cpu_reference_rate(name, fb, sigma, a.seconds, n_threads=1)[0]
                allc, n, _, sample = bench.cpu_reference_rate(name, fb, sigma, a.seconds, n_threads=ncpu)
            except Exception as ex:
                print(json.dumps({"K": K, "R": R, "decoder": name, "error": str(ex)[:200]}), flush=True)
                continue
            rows.append({"K": K, "R": R, "frame_bits": fb, "sigma": sigma, "decoder": name, "per_core_mbit_s": round(one, 6),
                         "all_cores_mbit_s": round(allc, 6), "cores": n, "scaling": round(allc / one / n, 3), "host": bench.cpu_model(),
                         "build": bench._REF_BUILD})
            if best is None or allc > best[1]:
                best = (name, allc)
        for r in rows:
            r["fastest_of_this_code"] = best[0] if best else None
            print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
