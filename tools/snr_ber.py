#!/usr/bin/env python
"""BER against Eb/N0 on the GPU: the device-side counterpart of the reference's
williamyang_viterbi_lib/examples/run_snr_ber.cpp. Data, encoder, AWGN channel and quantiser run on the device
(b200v_batch_generate_dev), the frames are decoded with b200v_batch_decode_dev, and the bit errors are counted on
the device; nothing but the error count crosses PCIe. The channel follows run_snr_ber.cpp:326-364:
    snr_dB = EbN0_dB + 10 log10(R);  sigma^2 = 10^(-snr_dB/10);  soft = clamp(round(x * mag / sqrt(1 + sigma^2) + mean))
and the output is the same JSON list (name, decode_type, simd_type, K, R, G, EbNo_dB, ber; run_snr_ber.cpp:420-447)
so that the reference's scripts/plot_snr_ber.py reads it, with the bit and error counts added.

    python tools/snr_ber.py > gpurun_out/snr_ber.json          (GPU box only)
"""
import argparse
import json
import math
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

# williamyang_viterbi_lib/examples/helpers/common_codes.h:23-30
COMMON_CODES = [
    ("Basic K=3 R=1/2", 3, 2, [0b111, 0b101]), ("Basic K=5 R=1/2", 5, 2, [0b10111, 0b11001]),
    ("Voyager", 7, 2, [109, 79]), ("LTE", 7, 3, [91, 117, 121]), ("DAB Radio", 7, 4, [109, 79, 83, 109]),
    ("CDMA IS-95A", 9, 2, [491, 369]), ("CDMA 2000", 9, 4, [501, 441, 331, 315]),
    ("Cassini", 15, 6, [17817, 20133, 23879, 30451, 32439, 26975]),
]


def channel_q8(lo, hi, ebno_db, R):
    snr_db = ebno_db + 10.0 * math.log10(R)
    var = 10.0 ** (-snr_db / 10.0)
    mag, mean = (hi - lo) / 2.0, (hi + lo) / 2.0
    amp = mag / math.sqrt(1.0 + var)
    return int(round(amp * 256)), int(round(amp * math.sqrt(var) * 256)), int(round(mean * 256))


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--frame-bits", type=int, default=1024)
    p.add_argument("--ebno-start", type=float, default=None, help="default: the reference's rule, run_snr_ber.cpp:229-236")
    p.add_argument("--ebno-step", type=float, default=None)
    p.add_argument("--max-points", type=int, default=60)
    p.add_argument("--min-errors", type=int, default=100, help="stop a curve once a point collects fewer errors")
    p.add_argument("--target-bits", type=float, default=2.0e9, help="decoded bits per point (K=15: / 64)")
    p.add_argument("--widths", default="8,16")
    p.add_argument("--codes", default="", help="comma-separated substrings of code names (default: all)")
    args = p.parse_args()

    import torch
    import ka9q_viterbi_comparison_b200 as kv

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    popcnt = torch.tensor([bin(i).count("1") for i in range(256)], dtype=torch.int32, device=dev)
    results = []
    for name, K, R, poly in COMMON_CODES:
        if args.codes and not any(s.lower() in name.lower() for s in args.codes.split(",")):
            continue
        for width in [int(w) for w in args.widths.split(",")]:
            codec = kv.WY_U8 if width == 8 else kv.WY_U16
            lo, hi = (-3, 3) if width == 8 else (-127, 127)
            nbits = args.frame_bits
            T = nbits + K - 1
            sym_bytes = width // 8
            srow, orow = T * R * sym_bytes, nbits // 8
            budget = args.target_bits / (64 if K >= 15 else 1)
            frames = max(256, min(1 << 18, int(budget // nbits)))
            reps = max(1, int(round(budget / (frames * nbits))))
            dec = kv.BatchDecoder(codec, T, frames, K=K, R=R, poly=poly, device=0)
            stream = torch.cuda.Stream(device=dev)
            torch.cuda.set_stream(stream)
            dec.set_stream(stream.cuda_stream)
            d_syms = torch.empty(frames * srow, dtype=torch.uint8, device=dev)
            d_data = torch.empty(frames * orow, dtype=torch.uint8, device=dev)
            d_out = torch.empty(frames * orow, dtype=torch.uint8, device=dev)
            ebnos, bers, bits_l, errs_l = [], [], [], []
            # run_snr_ber.cpp:229-236: start = -ceil((K R)^0.8) (K >= 9: -17), step 0.5 dB when K R > 20 else 1 dB
            ebno0 = args.ebno_start if args.ebno_start is not None else (-17.0 if K >= 9 else -math.ceil((K * R) ** 0.8))
            ebno_step = args.ebno_step if args.ebno_step is not None else (0.5 if K * R > 20 else 1.0)
            t0 = time.time()
            for pt in range(args.max_points):
                ebno = ebno0 + pt * ebno_step
                amp_q8, sigma_q8, mean_q8 = channel_q8(lo, hi, ebno, R)
                errors = torch.zeros((), dtype=torch.int64, device=dev)
                for rep in range(reps):
                    dec.generate_dev(d_syms, srow, d_data, orow, nbits, 0x5EED + 1000 * pt + rep, 0,
                                     amp_q8, sigma_q8, mean_q8, lo, hi)
                    dec.decode_dev(d_syms, srow, d_out, orow, nbits)
                    errors += popcnt[torch.bitwise_xor(d_out, d_data).long()].sum()
                n_err = int(errors.item())
                n_bits = reps * frames * nbits
                ebnos.append(round(ebno, 1))
                bers.append(n_err / n_bits)
                bits_l.append(n_bits)
                errs_l.append(n_err)
                if n_err < args.min_errors and bers[-1] < 1e-3:
                    break
            torch.cuda.synchronize()
            results.append({"name": name, "decode_type": f"soft{width}", "simd_type": "b200",
                            "K": K, "R": R, "G": poly, "EbNo_dB": ebnos, "ber": bers,
                            "bits": bits_l, "bit_errors": errs_l, "kernel_family": dec.variant,
                            "frames_per_launch": frames, "seconds": round(time.time() - t0, 2)})
            print(f"{name} soft{width}: {len(ebnos)} points in {results[-1]['seconds']} s", file=sys.stderr, flush=True)
            dec.close()
            del d_syms, d_data, d_out
    print(json.dumps(results, indent=1))


if __name__ == "__main__":
    main()
