#!/usr/bin/env python
"""Small-batch latency of the reference-style call sequence (init / update / chainback with HOST buffers, each call
synchronous, as src/main.cpp:264-278 times them) for every kernel family an instance offers. GPU box only.

    python tools/latency.py [--codecs ka9q27,spiral27,...] [--frames 1,32,1024] > gpurun_out/latency.jsonl

One JSON line per (codec, frame bytes, n_frames, variant): mean microseconds of each call over --reps repetitions
after a warm-up, and whether the decoded bytes equal the transmitted ones (noiseless symbols like main.cpp's).
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

# main.cpp's default frame sizes in bytes per codec shape (src/main.cpp:363-419)
MAIN_BYTES = {7: 1024, 9: 512, 15: 256, 24: 8}
CODECS = ["ka9q27", "spiral27", "spiral47", "wy_u8:7:2", "wy_u16:7:2", "ka9q29", "spiral29", "spiral49", "wy_u8:9:2",
          "ka9q615", "spiral615", "wy_u8:15:6", "ka9q224"]


def main():
    import _oracle as o
    import ka9q_viterbi_comparison_b200 as kv
    p = argparse.ArgumentParser()
    p.add_argument("--codecs", default=",".join(CODECS))
    p.add_argument("--frames", default="1,64,1024")
    p.add_argument("--variants", default="0,1,2,3")
    p.add_argument("--reps", type=int, default=20)
    args = p.parse_args()
    for spec in args.codecs.split(","):
        parts = spec.split(":")
        name = parts[0]
        K, R = (int(parts[1]), int(parts[2])) if len(parts) == 3 else (0, 0)
        oname = name if not K else f"wy_avx_{name[3:]}_k{K}r{R}"
        c = o.codec_by_name(oname)
        nbytes = MAIN_BYTES[c.K]
        T = nbytes * 8 + c.K - 1
        for n_frames in [int(x) for x in args.frames.split(",")]:
            if c.K >= 15 and n_frames > 64:
                continue
            frames = [o.make_frame(c, nbytes, 0.0, seed=0x1A7, frame=f) for f in range(min(n_frames, 8))]
            data = np.stack([frames[f % len(frames)][0] for f in range(n_frames)])
            syms = np.stack([frames[f % len(frames)][1] for f in range(n_frames)])
            for variant in [int(x) for x in args.variants.split(",")]:
                try:
                    dec = kv.BatchDecoder(name, T, n_frames, K=K, R=R)
                    if variant:
                        dec.set_variant(variant)
                except Exception:
                    continue
                used = dec.variant
                if variant == 0:
                    auto = used
                t_i, t_u, t_c = [], [], []
                out = None
                for rep in range(args.reps + 3):
                    t0 = time.perf_counter()
                    dec.reset()
                    t1 = time.perf_counter()
                    dec.update(syms)
                    t2 = time.perf_counter()
                    out = dec.chainback(nbytes * 8)
                    t3 = time.perf_counter()
                    if rep >= 3:
                        t_i.append(t1 - t0), t_u.append(t2 - t1), t_c.append(t3 - t2)
                ok = bool(np.array_equal(out, data))
                dec.close()
                print(json.dumps({"codec": spec, "K": c.K, "R": c.R, "frame_bytes": nbytes, "n_frames": n_frames,
                                  "variant": "auto" if variant == 0 else variant, "family": used,
                                  "init_us": round(1e6 * float(np.mean(t_i)), 1),
                                  "update_us": round(1e6 * float(np.mean(t_u)), 1),
                                  "chainback_us": round(1e6 * float(np.mean(t_c)), 1), "decoded_ok": ok}), flush=True)


if __name__ == "__main__":
    main()
