#!/usr/bin/env python
"""main.cpp's benchmark, batched: every codec of src/main.cpp:363-419 at main.cpp's own frame sizes and
noiseless symbols, decoded as a BATCH of frames on the GPU, written in the JSON schema of the reference's driver
(src/main.cpp:80-118; scripts/sample_loader.py:3-20), so that scripts/tabulate_data.py and scripts/plot_data.py
tabulate / plot it unchanged: one "b200_<impl>" column per reference implementation next to (optionally) the
reference's own CPU columns measured on this host.

    python tools/batch_benchmark.py [--frames 4096] [--samples 8] [--cpu] [-o gpurun_out/benchmark_b200.json]

Sizes are scaled by the batch so that the reference's rate formulas hold:
    update symbol rate  = total_output_symbols / update_ns      (scripts/tabulate_data.py:33)
    chainback bit rate  = total_input_bytes * 8 / chainback_ns  (scripts/tabulate_data.py:54)
with total_input_bytes = n_frames * frame bytes and total_output_symbols = n_frames * T * R. Times are CUDA-event
device times of the init / update / chainback launches with symbols resident in HBM. Extra keys (ignored by the
reference's loader): n_frames, frame_input_bytes, kernel_family, device.
GPU box only for the b200 columns; --cpu adds the compiled reference (oracle/_ref, test infrastructure) per codec, one
frame on one core exactly like main.cpp.
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

# (K, R, frame bytes, [(column name of the reference implementation, product codec, oracle registry name)])
TESTS = [
    (7, 2, 1024, [("ka9q", "ka9q27", "ka9q27"), ("spiral", "spiral27", "spiral27"),
                  ("sse_u8", "wy_u8", "wy_sse_u8_k7r2"), ("sse_u16", "wy_u16", "wy_sse_u16_k7r2")]),
    (7, 4, 1024, [("spiral", "spiral47", "spiral47"), ("sse_u8", "wy_u8", "wy_sse_u8_k7r4"),
                  ("sse_u16", "wy_u16", "wy_sse_u16_k7r4")]),
    (9, 2, 512, [("ka9q", "ka9q29", "ka9q29"), ("spiral", "spiral29", "spiral29"),
                 ("sse_u8", "wy_u8", "wy_sse_u8_k9r2"), ("sse_u16", "wy_u16", "wy_sse_u16_k9r2")]),
    (9, 4, 512, [("spiral", "spiral49", "spiral49"), ("sse_u8", "wy_u8", "wy_sse_u8_k9r4"),
                 ("sse_u16", "wy_u16", "wy_sse_u16_k9r4")]),
    (15, 6, 256, [("ka9q", "ka9q615", "ka9q615"), ("spiral", "spiral615", "spiral615"),
                  ("sse_u8", "wy_u8", "wy_sse_u8_k15r6"), ("sse_u16", "wy_u16", "wy_sse_u16_k15r6")]),
    (24, 2, 8, [("ka9q", "ka9q224", "ka9q224"), ("sse_u8", "wy_u8", "wy_sse_u8_k24r2"),
                ("sse_u16", "wy_u16", "wy_sse_u16_k24r2")]),
]
REQUIRED_KEYS = ["name", "K", "R", "poly", "total_input_bytes", "total_transmit_bits", "total_output_symbols",
                 "sampling_time", "minimum_samples", "total_samples", "init_ns", "update_ns", "chainback_ns",
                 "total_bits", "total_bit_errors", "bit_error_rate"]


def entry(name, c, nbytes, n_frames, init_ns, update_ns, chainback_ns, bit_errors, sampling_time, min_samples, extra):
    """One object of the reference's JSON array (src/main.cpp:80-118), sizes scaled by the batch."""
    T = nbytes * 8 + c.K - 1
    total_bits = n_frames * nbytes * 8
    d = {
        "name": name, "K": int(c.K), "R": int(c.R), "poly": [int(c.poly[i]) for i in range(c.R)],
        "total_input_bytes": n_frames * nbytes, "total_transmit_bits": n_frames * T,
        "total_output_symbols": n_frames * T * c.R, "sampling_time": float(sampling_time),
        "minimum_samples": int(min_samples), "total_samples": len(update_ns),
        "init_ns": [int(x) for x in init_ns], "update_ns": [int(x) for x in update_ns],
        "chainback_ns": [int(x) for x in chainback_ns],
        "total_bits": total_bits, "total_bit_errors": int(bit_errors),
        "bit_error_rate": float(bit_errors) / float(total_bits),
    }
    d.update(extra)
    return d


def default_frames(K, frames):
    # keep the decision buffer of a batch within a few GB: K=15 2 KiB / step, K=24 1 MiB / step
    return frames if K <= 9 else min(frames, 1024) if K <= 16 else min(frames, 8)


def run_b200(args, out):
    import torch
    import _oracle as o
    import ka9q_viterbi_comparison_b200 as kv
    dev = torch.device("cuda", 0)
    for K, R, nbytes, impls in TESTS:
        for col, codec, oname in impls:
            c = o.codec_by_name(oname)
            n_frames = default_frames(K, args.frames)
            T = nbytes * 8 + K - 1
            uniq = [o.make_frame(c, nbytes, 0.0, seed=0xBE7C, frame=f) for f in range(min(n_frames, 4))]
            data = np.stack([uniq[f % len(uniq)][0] for f in range(n_frames)])
            syms = np.stack([uniq[f % len(uniq)][1] for f in range(n_frames)])
            srow = syms.strides[0]
            d_syms = torch.from_numpy(syms.view(np.uint8).reshape(-1)).to(dev)
            d_out = torch.zeros(n_frames * nbytes, dtype=torch.uint8, device=dev)
            stream = torch.cuda.Stream(device=dev)
            with torch.cuda.stream(stream), kv.BatchDecoder(codec, T, n_frames, K=K, R=R) as dec:
                dec.set_stream(stream.cuda_stream)
                init_ns, upd_ns, cb_ns = [], [], []
                for rep in range(args.samples + 2):
                    e = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
                    e[0].record()
                    dec.reset()
                    e[1].record()
                    dec.update_dev(d_syms, srow, T)
                    dec.chainback_dev(d_out, nbytes, nbytes * 8)
                    torch.cuda.synchronize()
                    u, cb = dec.last_kernel_ms()
                    if rep >= 2:
                        init_ns.append(e[0].elapsed_time(e[1]) * 1e6), upd_ns.append(u * 1e6), cb_ns.append(cb * 1e6)
                got = d_out.cpu().numpy().reshape(n_frames, nbytes)
                errs = int(np.unpackbits(np.bitwise_xor(got, data)).sum())
                out.append(entry("b200_" + col, c, nbytes, n_frames, init_ns, upd_ns, cb_ns, errs, 0.0, args.samples,
                                 {"n_frames": n_frames, "frame_input_bytes": nbytes, "kernel_family": dec.variant,
                                  "device": torch.cuda.get_device_name(0)}))
            print(f"b200_{col} K={K} R={R}: {n_frames} frames, update {np.mean(upd_ns) / 1e6:.3f} ms, "
                  f"chainback {np.mean(cb_ns) / 1e6:.3f} ms, bit errors {errs}", file=sys.stderr)


def run_cpu(args, out):
    """The reference's own decoders, one frame on one core, timed call by call like src/main.cpp:264-278."""
    import _oracle as o
    if not o.ref_available():
        print("oracle/_ref not built: no CPU columns", file=sys.stderr)
        return
    for K, R, nbytes, impls in TESTS:
        for col, _, oname in impls:
            c = o.codec_by_name(oname)
            T = nbytes * 8 + K - 1
            data, syms = o.make_frame(c, nbytes, 0.0, seed=0xBE7C, frame=0)
            d = o.Ref(oname, T)
            init_ns, upd_ns, cb_ns = [], [], []
            t_end = time.perf_counter() + args.cpu_seconds
            got = None
            while time.perf_counter() < t_end or len(upd_ns) < args.samples:
                t0 = time.perf_counter_ns()
                d.init(0)
                t1 = time.perf_counter_ns()
                d.update(syms, T)
                t2 = time.perf_counter_ns()
                got = d.chainback(nbytes * 8)
                t3 = time.perf_counter_ns()
                init_ns.append(t1 - t0), upd_ns.append(t2 - t1), cb_ns.append(t3 - t2)
            d.close()
            errs = int(np.unpackbits(np.bitwise_xor(np.asarray(got).reshape(-1)[:nbytes], data)).sum())
            out.append(entry(col, c, nbytes, 1, init_ns, upd_ns, cb_ns, errs, args.cpu_seconds, args.samples,
                             {"n_frames": 1, "frame_input_bytes": nbytes, "timer": "python perf_counter_ns around ctypes calls"}))


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--frames", type=int, default=4096, help="frames per batch (K=15: at most 1024, K=24: at most 8)")
    p.add_argument("--samples", type=int, default=8)
    p.add_argument("--cpu", action="store_true", help="add the reference's CPU columns (oracle/_ref)")
    p.add_argument("--cpu-seconds", type=float, default=0.2)
    p.add_argument("--no-gpu", action="store_true")
    p.add_argument("-o", "--output", default="-")
    args = p.parse_args()
    out = []
    if args.cpu:
        run_cpu(args, out)
    if not args.no_gpu:
        run_b200(args, out)
    text = "[\n" + ",\n".join(json.dumps(x) for x in out) + "\n]\n"
    if args.output == "-":
        sys.stdout.write(text)
    else:
        with open(args.output, "w") as fp:
            fp.write(text)


if __name__ == "__main__":
    main()
