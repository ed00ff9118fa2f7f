#!/usr/bin/env python
"""punctured_bench.py -- end-to-end rate of a punctured stream: the same frames decoded
(a) from full-rate symbols with the erasures already filled in on the host (what b200v_batch_decode needs), and
(b) from the transmitted symbols only, through b200v_batch_decode_punctured (depuncturing fused into the symbol fetch).
Both through host (pinned) buffers, copies inside the timed region. Also the device-resident rates of the fused update
against expand-then-update. One JSON line.

    python tools/punctured_bench.py --codec ka9q27 --frames 1048576 --code 1,1,0,1,1,0      # rate 3/4 from the r=1/2 code
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--codec", default="ka9q27")
    ap.add_argument("--K", type=int, default=7)
    ap.add_argument("--R", type=int, default=2)
    ap.add_argument("--frames", type=int, default=1 << 20)
    ap.add_argument("--frame-bits", type=int, default=2048)
    ap.add_argument("--code", default="1,1,0,1,1,0")
    ap.add_argument("--steps", type=int, default=3)
    args = ap.parse_args()
    import torch
    import ka9q_viterbi_comparison_b200 as kv
    import _oracle as o
    lib = kv.load_library()
    codec = kv.CODEC_BY_NAME[args.codec]
    code = np.array([int(x) for x in args.code.split(",")], dtype=np.uint8)
    with kv.BatchDecoder(codec, args.frame_bits + args.K - 1, args.frames, K=args.K, R=args.R) as dec:
        K, R, T, frames, nbits = dec.K, dec.R, dec.frame_steps, args.frames, args.frame_bits
        name = args.codec if args.codec not in ("wy_u8", "wy_u16") else f"wy_sse_u{8 if args.codec == 'wy_u8' else 16}_k{K}r{R}"
        c = o.codec_by_name(name)
        sb = np.dtype(o.sym_dtype(c)).itemsize
        amp_q8, sigma_q8, mean_q8, lo, hi = o.channel_params(c, 0.5)
        fill = (lo + hi + 1) // 2
        srow, orow = T * R * sb, nbits // 8
        d_syms = torch.zeros(frames * srow, dtype=torch.uint8, device="cuda")
        d_data = torch.zeros(frames * orow, dtype=torch.uint8, device="cuda")
        dec.generate_dev(d_syms, srow, d_data, orow, nbits, 3, 0, amp_q8, sigma_q8, mean_q8, lo, hi)
        dec.sync()
        n_tx = dec.set_puncturing([(code, T * R)], fill)
        keep = torch.from_numpy(np.resize(code.astype(bool), T * R)).cuda()
        full = d_syms.view(frames, T * R, sb)
        d_tx = full[:, keep, :].contiguous().view(-1)                     # transmitted symbols only
        erased = full.clone()
        fb = torch.tensor(list(int(fill & (0xffff if sb == 2 else 0xff)).to_bytes(sb, "little")), dtype=torch.uint8, device="cuda")
        erased[:, ~keep, :] = fb
        d_erased = erased.view(-1)
        irow = n_tx * sb
        h_full = lib.b200v_host_alloc_pinned(frames * srow)
        h_tx = lib.b200v_host_alloc_pinned(frames * irow)
        h_out = lib.b200v_host_alloc_pinned(frames * orow)
        h_out2 = lib.b200v_host_alloc_pinned(frames * orow)
        lib.b200v_memcpy_d2h(0, h_full, d_erased.data_ptr(), frames * srow)
        lib.b200v_memcpy_d2h(0, h_tx, d_tx.data_ptr(), frames * irow)

        def timed(fn):
            assert fn() == 0, lib.b200v_last_error()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(args.steps):
                assert fn() == 0, lib.b200v_last_error()
            torch.cuda.synchronize()
            return (time.perf_counter() - t0) / args.steps

        t_full = timed(lambda: lib.b200v_batch_decode(dec.h, h_full, h_out, nbits))
        t_punct = timed(lambda: lib.b200v_batch_decode_punctured(dec.h, h_tx, h_out2, nbits))
        a = np.ctypeslib.as_array(C.cast(h_out, C.POINTER(C.c_uint8)), shape=(frames * orow,))
        b = np.ctypeslib.as_array(C.cast(h_out2, C.POINTER(C.c_uint8)), shape=(frames * orow,))
        same = bool(np.array_equal(a, b))
        ber = float(np.unpackbits(np.bitwise_xor(b[:4096 * orow], d_data[:4096 * orow].cpu().numpy())).sum()) / (4096 * nbits)
        # device-resident: fused update against expansion + update
        d_out = torch.zeros(frames * orow, dtype=torch.uint8, device="cuda")
        dec.decode_punctured_dev(d_tx, irow, d_out, orow, nbits); dec.sync()
        fused_ms = dec.last_kernel_ms()[0]
        d_exp = torch.zeros(frames * srow, dtype=torch.uint8, device="cuda")
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        st = torch.cuda.Stream()           # a non-default stream: a NULL handle means "the instance's own stream" to the library
        torch.cuda.synchronize()
        dec.set_stream(st.cuda_stream)
        with torch.cuda.stream(st):
            dec.depuncture_dev(d_tx, irow, 0, code, d_exp, srow, 0, T * R, fill)       # warm-up
            ev[0].record()
            dec.depuncture_dev(d_tx, irow, 0, code, d_exp, srow, 0, T * R, fill)
            ev[1].record()
        torch.cuda.synchronize()
        expand_ms = ev[0].elapsed_time(ev[1])
        dec.decode_dev(d_exp, srow, d_out, orow, nbits); dec.sync()
        plain_ms = dec.last_kernel_ms()[0]
        print(json.dumps({
            "tool": "punctured_bench", "codec": name, "frames": frames, "frame_bits": nbits, "puncture_code": args.code,
            "rate": f"{int(code.sum())}/{len(code)} of the symbols transmitted", "variant": dec.variant,
            "e2e_full_rate_host_gbit_s": frames * nbits / t_full / 1e9, "e2e_punctured_host_gbit_s": frames * nbits / t_punct / 1e9,
            "e2e_speedup": t_full / t_punct, "byte_ratio": srow / irow, "h2d_bytes_full": frames * srow,
            "h2d_bytes_punctured": frames * irow, "outputs_identical": same, "sample_ber": ber,
            "device_update_ms": {"fused": fused_ms, "expand_kernel": expand_ms, "update_after_expand": plain_ms}}))
        for p in (h_full, h_tx, h_out, h_out2):
            lib.b200v_host_free_pinned(p)


if __name__ == "__main__":
    main()
