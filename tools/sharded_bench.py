#!/usr/bin/env python
"""sharded_bench.py -- end-to-end rate of the IN-PROCESS multi-GPU entry (b200v_sharded_decode, one host thread and one
instance per GPU, pinned host buffers) with equal shards and with shards weighted by the host-to-device rate each GPU's
link sustains when all of them copy at once (measured first, in this process). One JSON line.

    python tools/sharded_bench.py --frames-per-gpu 262144          # on a multi-GPU box
"""
import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--codec", default="ka9q27")
    ap.add_argument("--frames-per-gpu", type=int, default=262144)
    ap.add_argument("--frame-bits", type=int, default=2048)
    ap.add_argument("--reps", type=int, default=2)
    a = ap.parse_args()
    import ka9q_viterbi_comparison_b200 as kv
    import _oracle as o
    lib = kv.load_library()
    ndev = lib.b200v_device_count()
    codec = kv.CODEC_BY_NAME[a.codec]
    c = o.codec_by_name(a.codec)
    K, R = c.K, c.R
    nbits, T = a.frame_bits, a.frame_bits + K - 1
    srow, orow = T * R, nbits // 8
    n_frames = a.frames_per_gpu * ndev
    block = min(65536, a.frames_per_gpu)
    h_syms = lib.b200v_host_alloc_pinned(n_frames * srow)
    h_out = lib.b200v_host_alloc_pinned(n_frames * orow)
    assert h_syms and h_out, lib.b200v_last_error()
    # one block of frames from the device generator, repeated over the whole batch
    import torch
    amp_q8, sigma_q8, mean_q8, lo, hi = o.channel_params(c, 0.7)
    with kv.BatchDecoder(codec, T, block) as gen:
        d = torch.zeros(block * srow, dtype=torch.uint8, device="cuda:0")
        d_data = torch.zeros(block * orow, dtype=torch.uint8, device="cuda:0")
        gen.generate_dev(d, srow, d_data, orow, nbits, 11, 0, amp_q8, sigma_q8, mean_q8, lo, hi)
        gen.sync()
        lib.b200v_memcpy_d2h(0, h_syms, d.data_ptr(), block * srow)
        data = d_data.cpu().numpy()
        del d, d_data
    for i in range(1, n_frames // block):
        C.memmove(h_syms + i * block * srow, h_syms, block * srow)

    # link rates with every GPU copying at once
    per = a.frames_per_gpu * srow
    d_bufs = [lib.b200v_device_alloc(g, per) for g in range(ndev)]
    rates = [0.0] * ndev

    def copy(g):
        lib.b200v_memcpy_h2d(g, d_bufs[g], h_syms + g * per, per)
        t0 = time.perf_counter()
        for _ in range(2):
            lib.b200v_memcpy_h2d(g, d_bufs[g], h_syms + g * per, per)
        rates[g] = 2 * per / (time.perf_counter() - t0) / 1e9

    th = [threading.Thread(target=copy, args=(g,)) for g in range(ndev)]
    [t.start() for t in th]
    [t.join() for t in th]
    for g in range(ndev):
        lib.b200v_device_free(g, d_bufs[g])

    devs = (C.c_int * ndev)(*range(ndev))
    poly = kv.STANDARD_POLYS[(K, R)]
    parr = (C.c_int * len(poly))(*poly)
    res = {}
    for label, w in (("equal", None), ("weighted", rates)):
        if w is None:
            h = lib.b200v_sharded_create(codec, K, R, parr, T, n_frames, devs, ndev, None)
        else:
            h = lib.b200v_sharded_create_weighted(codec, K, R, parr, T, n_frames, devs, (C.c_double * ndev)(*w), ndev, None)
        assert h, lib.b200v_last_error()
        assert lib.b200v_sharded_decode(h, h_syms, h_out, nbits) == 0, lib.b200v_sharded_last_error(h)     # warm-up
        t0 = time.perf_counter()
        for _ in range(a.reps):
            assert lib.b200v_sharded_decode(h, h_syms, h_out, nbits) == 0, lib.b200v_sharded_last_error(h)
        dt = (time.perf_counter() - t0) / a.reps
        f0, cnt = C.c_int(), C.c_int()
        shards = []
        for g in range(ndev):
            lib.b200v_sharded_shard(h, g, C.byref(f0), C.byref(cnt))
            shards.append(cnt.value)
        out = np.ctypeslib.as_array(C.cast(h_out, C.POINTER(C.c_uint8)), shape=(n_frames * orow,))
        ok = all(np.array_equal(out[i * block * orow:(i * block + 256) * orow], out[:256 * orow]) for i in range(1, n_frames // block))
        ber = float(np.unpackbits(np.bitwise_xor(out[:4096 * orow], data[:4096 * orow])).sum()) / (4096 * nbits)
        res[label] = {"e2e_gbit_s": n_frames * nbits / dt / 1e9, "h2d_gbs": n_frames * srow / dt / 1e9, "seconds": dt,
                      "frames_per_shard": shards, "blocks_identical": bool(ok), "sample_ber": ber}
        lib.b200v_sharded_destroy(h)
    print(json.dumps({"tool": "sharded_bench", "codec": a.codec, "n_gpus": ndev, "frames": n_frames, "frame_bits": nbits,
                      "link_gbs_all_copying": [round(r, 2) for r in rates], "sum_gbs": round(sum(rates), 1), **res}))
    lib.b200v_host_free_pinned(h_syms)
    lib.b200v_host_free_pinned(h_out)


if __name__ == "__main__":
    main()
