#!/usr/bin/env python
"""h2d_probe.py -- the fabric ceiling of the end-to-end path: N processes, one per GPU, each
copying a pinned host buffer to its device with cudaMemcpy, all at once. Reports per-rank and aggregate GB/s with and
without binding each rank to the CPUs / memory node of its GPU (b200v_bind_host_to_device).

    python tools/h2d_probe.py                                   # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
        tools/h2d_probe.py --gb 4

Prints one JSON line per mode on rank 0. GPU box only (nothing here computes: the copies are the measurement).
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def topology(lib, n):
    import ctypes as C
    out = {}
    try:
        out["nodes_online"] = open("/sys/devices/system/node/online").read().strip()
    except OSError:
        out["nodes_online"] = None
    out["affinity"] = len(os.sched_getaffinity(0))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gb", type=float, default=4.0)
    ap.add_argument("--seconds", type=float, default=3.0)
    ap.add_argument("--modes", default="unbound,bound")
    args = ap.parse_args()
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("gloo")
    import torch
    import ka9q_viterbi_comparison_b200 as kv
    lib = kv.load_library()
    torch.cuda.set_device(local)
    nbytes = int(args.gb * (1 << 30))
    d_buf = lib.b200v_device_alloc(local, nbytes)
    assert d_buf, lib.b200v_last_error()
    full = os.sched_getaffinity(0)
    for mode in args.modes.split(","):
        node = None
        if mode == "bound":
            node = lib.b200v_bind_host_to_device(local)
        else:
            os.sched_setaffinity(0, full)
        h_buf = lib.b200v_host_alloc_pinned(nbytes)
        assert h_buf, lib.b200v_last_error()
        import ctypes as C
        C.memset(h_buf, rank + 1, nbytes)                 # touch every page from this thread
        lib.b200v_memcpy_h2d(local, d_buf, h_buf, nbytes)
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        t0 = time.perf_counter()
        n = 0
        while time.perf_counter() - t0 < args.seconds:
            lib.b200v_memcpy_h2d(local, d_buf, h_buf, nbytes)
            n += 1
        dt = time.perf_counter() - t0
        gbs = n * nbytes / dt / 1e9
        vals = [gbs]
        nodes = [node]
        if dist is not None:
            t = torch.tensor([gbs, float(-1 if node is None else node)], dtype=torch.float64)
            gathered = [torch.zeros(2, dtype=torch.float64) for _ in range(world)]
            dist.all_gather(gathered, t)
            vals = [float(g[0]) for g in gathered]
            nodes = [int(g[1]) for g in gathered]
            dist.barrier()
        if rank == 0:
            print(json.dumps({"probe": "h2d", "n_gpus": world, "mode": mode, "gb_per_copy": args.gb,
                              "per_rank_gbs": [round(v, 2) for v in vals], "aggregate_gbs": round(sum(vals), 2),
                              "numa_node_per_rank": nodes, "host": topology(lib, world)}), flush=True)
        lib.b200v_host_free_pinned(h_buf)
    lib.b200v_device_free(local, d_buf)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
