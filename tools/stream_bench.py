#!/usr/bin/env python
"""stream_bench.py -- throughput of the sliding-window stream decoder (b200v_stream_*): n_streams continuous
streams, pushes of `chunk` steps, traceback depth `depth`; symbols resident in HBM (generated on the device), CUDA events
around the pushes. Reports decoded Mbit/s and the share of the window traceback. One JSON line per configuration.

    python tools/stream_bench.py --codec ka9q27 --streams 262144 --chunk 512 --depth 512
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--codec", default="ka9q27")
    ap.add_argument("--K", type=int, default=7)
    ap.add_argument("--R", type=int, default=2)
    ap.add_argument("--streams", type=int, default=262144)
    ap.add_argument("--chunk", type=int, default=512)
    ap.add_argument("--depth", type=int, default=512)
    ap.add_argument("--pushes", type=int, default=16)
    a = ap.parse_args()
    import torch
    import ka9q_viterbi_comparison_b200 as kv
    import _oracle as o
    codec = kv.CODEC_BY_NAME[a.codec]
    with kv.StreamDecoder(codec, a.streams, a.depth, a.chunk, K=a.K, R=a.R) as sd:
        K, R = sd.K, sd.R
        name = a.codec if a.codec not in ("wy_u8", "wy_u16") else f"wy_sse_u{8 if a.codec == 'wy_u8' else 16}_k{K}r{R}"
        c = o.codec_by_name(name)
        sb = np.dtype(o.sym_dtype(c)).itemsize
        T = a.chunk * a.pushes
        nbits = T - 8                                  # the generator wants a byte multiple of data bits + a tail
        amp_q8, sigma_q8, mean_q8, lo, hi = o.channel_params(c, 0.5)
        srow = T * R * sb
        with kv.BatchDecoder(codec, T, a.streams, K=K, R=R) as gen:
            d_syms = torch.zeros(a.streams * srow, dtype=torch.uint8, device="cuda")
            d_data = torch.zeros(a.streams * (nbits // 8), dtype=torch.uint8, device="cuda")
            gen.generate_dev(d_syms, srow, d_data, nbits // 8, nbits, 9, 0, amp_q8, sigma_q8, mean_q8, lo, hi)
            gen.sync()
        d_out = torch.zeros(a.streams * (a.chunk // 8), dtype=torch.uint8, device="cuda")
        st = torch.cuda.Stream()
        lib = kv.load_library()
        lib.b200v_batch_set_stream(lib.b200v_stream_batch(sd.h), st.cuda_stream)
        outs = []

        def run(collect):
            sd.reset(0)
            for p in range(a.pushes):
                n = sd.push_dev(d_syms.data_ptr() + p * a.chunk * R * sb, srow, d_out, a.chunk // 8)
                if collect and n:
                    st.synchronize()
                    outs.append(d_out.view(a.streams, -1)[:64].cpu().numpy().copy())

        run(True)                                      # warm-up + output check
        got = np.concatenate(outs, axis=1)
        want = d_data.view(a.streams, -1)[:64, :got.shape[1]].cpu().numpy()
        ber = float(np.unpackbits(np.bitwise_xor(got, want)).sum()) / got.size / 8
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        with torch.cuda.stream(st):
            e0.record()
            run(False)
            e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        emitted = a.streams * (T - a.depth)
        print(json.dumps({"tool": "stream_bench", "codec": name, "streams": a.streams, "chunk": a.chunk, "depth": a.depth,
                          "pushes": a.pushes, "steps_per_stream": T, "ms": ms, "decoded_mbit_s": emitted / ms / 1e3,
                          "steps_mstep_s": a.streams * T / ms / 1e3, "window_rows_walked_per_emitted_bit": (a.depth + a.chunk) / a.chunk,
                          "sample_ber": ber}))


if __name__ == "__main__":
    main()
