#!/usr/bin/env python
"""bench.py -- decoded Mbit/s (update + chainback) of the batched Viterbi decoder on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--codec ka9q27] ...

One "step" = init + update + chainback of one batch of frames (BASELINE.json configs[1]: 1 Mi frames x
2048 bits of K=7 r=1/2 per GPU; frames are independent, so ranks shard them with no collective: weak
scaling). `value` is timed with inputs resident in HBM; `e2e` goes through b200v_batch_decode with host
(pinned) buffers, copies included. Prints ONE JSON line on rank 0.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "decoded Mbit/s (update+chainback)"
UNIT = "Mbit/s"
ISSUE_PEAK = 148 * 4 * 1.965e9      # warp-instructions/s: one issue slot per SM sub-partition and clock
ALU_PEAK = 148 * 128 * 1.965e9      # dual-pipe integer issue peak, thread-instructions/s (measured 96-121 of 128 lanes)


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--codec", default="ka9q27")
    p.add_argument("--K", type=int, default=7, help="constraint length (wy_u8 / wy_u16 only)")
    p.add_argument("--R", type=int, default=2, help="code rate 1/R (wy_u8 / wy_u16 only)")
    p.add_argument("--frames", type=int, default=0, help="frames per GPU (default: 1 Mi for K<=9, 8192 for K=15)")
    p.add_argument("--frame-bits", type=int, default=2048)
    p.add_argument("--sigma", type=float, default=0.7, help="channel noise (0 = main.cpp's noiseless symbols)")
    p.add_argument("--variant", type=int, default=0, help="kernel family: 0 auto, 1 generic, 2 specialised")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--unfused", action="store_true", help="time init/update/chainback as three calls instead of b200v_batch_decode_dev")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--cpu-seconds", type=float, default=6.0)
    return p.parse_args()


# oracle registry name (CPU reference arm) for each product codec
REF_NAME = {"ka9q27": "ka9q27", "ka9q29": "ka9q29", "ka9q615": "ka9q615", "spiral27": "spiral27",
            "spiral29": "spiral29", "spiral47": "spiral47", "spiral49": "spiral49", "spiral615": "spiral615",
            "ka9q224": "ka9q224"}


def codec_shape(codec, K=7, R=2):
    """(K, R) of a fixed codec; the williamyang families take them from --K / --R."""
    import ka9q_viterbi_comparison_b200 as kv
    k, r = C.c_int(), C.c_int()
    kv.load_library().b200v_codec_shape(kv.CODEC_BY_NAME[codec], C.byref(k), C.byref(r))
    return (k.value, r.value) if k.value else (K, R)


def ref_name(codec, K, R):
    """oracle registry name of the reference decoder that matches a product codec"""
    if codec == "wy_u8":
        return f"wy_avx_u8_k{K}r{R}"
    if codec == "wy_u16":
        return f"wy_avx_u16_k{K}r{R}"
    return REF_NAME[codec]


def default_frames(K):
    return 1 << 20 if K <= 9 else 8192 if K <= 16 else 16


def channel_params(lo, hi, sigma):
    """Q8 fixed-point channel constants (run_snr_ber.cpp:317-365 shape): amp = mag/sqrt(1+sigma^2)."""
    mag = (hi - lo) / 2.0
    amp = mag / (1.0 + sigma * sigma) ** 0.5 if sigma > 0 else mag
    return int(round(amp * 256)), int(round(amp * sigma * 256)), int(round(256 * (hi + lo) / 2.0)), lo, hi


# ---------------------------------------------------------------------------------------------------
# clocks sampling during the timed region
# ---------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.samples, self.stop_flag, self.thread = index, [], False, None

    def _run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                self.samples.append([x.strip() for x in out.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def start(self):
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()

    def stop(self):
        self.stop_flag = True
        if self.thread:
            self.thread.join(timeout=6)
        sm, mx, reasons = [], 0.0, set()
        for s in self.samples:
            if len(s) < 6:
                continue
            try:
                sm.append(float(s[0]))
                mx = max(mx, float(s[1]))
            except ValueError:
                continue
            for nm, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), s[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------
# CPU reference arm: the reference's own decoder on the host cores (oracle/_ref, else the oracle port)
# ---------------------------------------------------------------------------------------------------
def cpu_model():
    """model name and logical / physical core counts of the host the CPU arm runs on"""
    model, cores = "unknown CPU", set()
    try:
        phys = core = None
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name") and model == "unknown CPU":
                model = line.split(":", 1)[1].strip()
            elif line.startswith("physical id"):
                phys = line.split(":", 1)[1].strip()
            elif line.startswith("core id"):
                core = line.split(":", 1)[1].strip()
                cores.add((phys, core))
    except OSError:
        pass
    return f"{model}, {os.cpu_count()} threads / {len(cores) or '?'} cores"


def cpu_reference_rate(codec, frame_bits, sigma, seconds, n_threads=None, K=7, R=2):
    """Decode frames of `frame_bits` with one decoder instance per host thread for ~`seconds`.
    Returns (Mbit/s aggregate, threads, kind, sample description)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle as o
    name = ref_name(codec, K, R)
    c = o.codec_by_name(name)
    nbytes = frame_bits // 8
    T = frame_bits + c.K - 1
    kind = "reference" if o.ref_available() else "port"
    n_threads = n_threads or (os.cpu_count() or 1)
    pool = [o.make_frame(c, nbytes, sigma, seed=0xB200, frame=f)[1] for f in range(16)]
    # create serially: the reference's create_* fills process-global tables unguarded
    decs = [(o.Ref(name, T) if kind == "reference" else o.Oracle(c, T)) for _ in range(n_threads)]
    counts = [0] * n_threads
    deadline = [0.0]
    c_loop = kind == "reference" and hasattr(o.ref_lib(), "kvref_time_decode")
    pool_arr = np.stack(pool)

    def work(i):
        d = decs[i]
        if c_loop:      # whole loop in C (GIL released): measures the decoder, not the interpreter
            counts[i], _ = d.time_decode(pool_arr, T, frame_bits, seconds)
            return
        n = 0
        while time.perf_counter() < deadline[0]:
            for s in pool:
                d.init(0)
                d.update(s, T)
                d.chainback(frame_bits)
                n += 1
        counts[i] = n

    t0 = time.perf_counter()
    deadline[0] = t0 + seconds
    th = [threading.Thread(target=work, args=(i,)) for i in range(n_threads)]
    [t.start() for t in th]
    [t.join() for t in th]
    dt = time.perf_counter() - t0
    for d in decs:
        d.close()
    frames = sum(counts)
    return frames * frame_bits / dt / 1e6, n_threads, kind, \
        f"{frames} frames x {frame_bits} bits of {name}, sigma={sigma}, {n_threads} threads, {dt:.1f} s, host {cpu_model()}"


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    K, R = codec_shape(args.codec, args.K, args.R)
    per_step = max(1.0, min(args.cpu_seconds, 240.0 / max(1, args.steps + args.warmup)))
    vals = []
    info = None
    for i in range(args.warmup + args.steps):
        v, cores, kind, sample = cpu_reference_rate(args.codec, args.frame_bits, args.sigma, per_step, K=K, R=R)
        if i >= args.warmup:
            vals.append(v)
            info = (cores, kind, sample)
    value = float(np.mean(vals))
    frames = args.frames or default_frames(K)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": f"{args.codec} K={K} r=1/{R}: {frames} frames x {args.frame_bits} bits per GPU",
                   "sigma": args.sigma},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": info[0], "kind": info[1], "sample": info[2]},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ---------------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------------
def run_b200_arm(args):
    import torch
    import ka9q_viterbi_comparison_b200 as kv

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)

    codec = kv.CODEC_BY_NAME[args.codec]
    K, R = codec_shape(args.codec, args.K, args.R)
    S = 1 << (K - 1)
    nbits = args.frame_bits
    T = nbits + K - 1
    frames = args.frames or default_frames(K)
    sym_bytes = kv.load_library().b200v_codec_symbol_bytes(codec)
    lo, hi = (-3, 3) if args.codec == "wy_u8" else {1: (0, 255)}.get(sym_bytes, (-127, 127))
    srow, orow = T * R * sym_bytes, nbits // 8

    dec = kv.BatchDecoder(codec, T, frames, K=K, R=R, device=local)
    if args.variant:
        dec.set_variant(args.variant)
    # a non-default torch stream: the library treats a NULL stream handle as "use the instance's own stream",
    # and torch.cuda.Event only sees work on torch's current stream
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    dec.set_stream(stream.cuda_stream)

    d_syms = torch.empty(frames * srow, dtype=torch.uint8, device=dev)
    d_data = torch.empty(frames * orow, dtype=torch.uint8, device=dev)
    d_out = torch.empty(frames * orow, dtype=torch.uint8, device=dev)
    amp_q8, sigma_q8, mean_q8, lo, hi = channel_params(lo, hi, args.sigma)
    dec.generate_dev(d_syms, srow, d_data, orow, nbits, 0xB200, rank * frames, amp_q8, sigma_q8, mean_q8, lo, hi)
    torch.cuda.synchronize()

    def step_calls():
        dec.reset()
        dec.update_dev(d_syms, srow, T)
        dec.chainback_dev(d_out, orow, nbits)

    def step():
        if args.unfused:
            step_calls()
        else:
            dec.decode_dev(d_syms, srow, d_out, orow, nbits)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    launches0 = dec.launch_count
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    upd_ms, cb_ms = [], []
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    total_ms = e0.elapsed_time(e1)
    launches = dec.launch_count - launches0
    # per-kernel device time (CUDA events on the launching stream, recorded inside the library)
    for _ in range(3):
        step_calls()
        u, c = dec.last_kernel_ms()
        upd_ms.append(u)
        cb_ms.append(c)
    clocks = sampler.stop() if rank == 0 else None

    # correctness guard inside the bench: decoded output vs generated data
    err_bits = int(np.unpackbits(torch.bitwise_xor(d_out[: 4096 * orow], d_data[: 4096 * orow]).cpu().numpy()).sum())
    ber = err_bits / float(4096 * nbits)

    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    value = world * frames * nbits / (ms_per_step * 1e-3) / 1e6

    # ---- end-to-end through the C-ABI batch entry point with host buffers ----
    e2e = None
    if not args.no_e2e:
        lib = kv.load_library()
        h_syms = lib.b200v_host_alloc_pinned(frames * srow)
        h_out = lib.b200v_host_alloc_pinned(frames * orow)
        if not h_syms or not h_out:
            raise RuntimeError("pinned allocation failed")
        lib.b200v_memcpy_d2h(local, h_syms, d_syms.data_ptr(), frames * srow)
        e2e_steps = max(1, min(args.steps, 3))
        for _ in range(1):
            rc = lib.b200v_batch_decode(dec.h, h_syms, h_out, nbits)
            assert rc == 0, lib.b200v_last_error()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            rc = lib.b200v_batch_decode(dec.h, h_syms, h_out, nbits)
            assert rc == 0, lib.b200v_last_error()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
        e2e = {"value": world * frames * nbits * e2e_steps / dt / 1e6, "unit": UNIT,
               "h2d_bytes_per_step": frames * srow, "d2h_bytes_per_step": frames * orow, "steps": e2e_steps}
        lib.b200v_host_free_pinned(h_syms)
        lib.b200v_host_free_pinned(h_out)

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (update): algorithmic bytes = symbols in + decision bits out ----
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_gbs = float(peaks.get("hbm_gbs", 6650.0))
    upd = float(np.median(upd_ms))
    cb = float(np.median(cb_ms))
    bytes_per_step_frame = R * sym_bytes + S / 8.0
    achieved = frames * T * bytes_per_step_frame / (upd * 1e-3) / 1e9
    lane_ops = frames * T * 4.0 * S                      # 4*S integer lane-ops per step
    traffic, traffic_src, warp_inst = None, None, None
    try:     # DRAM bytes per launch as ncu measured them (profiles/ncu_traffic.json: per frame-step, scaled to this launch)
        tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(
            f"{args.codec}_k{K}r{R}" if args.codec.startswith("wy_") else args.codec)
        if tr:
            traffic, traffic_src = tr["bytes_per_frame_step"] * frames * T, tr["capture"]
            if tr.get("warp_inst_per_frame_step"):
                warp_inst = tr["warp_inst_per_frame_step"] * frames * T
    except Exception:
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak_gbs, "unit": "GB/s", "frac": achieved / peak_gbs,
                "traffic": traffic, "traffic_source": traffic_src, "algorithmic_bytes": frames * T * bytes_per_step_frame, "kernel": "update", "kernel_ms": upd, "chainback_ms": cb,
                "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback",
                # the roofline that actually binds the update kernels: integer issue, two pipes of 64 lanes
                # per clock and SM at the measured SM clock limit (tools/ubench_int.cu, profiles/r01_ubench_int_pipes.jsonl)
                "alu": {"lane_ops_per_s": lane_ops / (upd * 1e-3), "peak_lane_ops_per_s": ALU_PEAK,
                        "frac": lane_ops / (upd * 1e-3) / ALU_PEAK,
                        # executed warp-instructions (ncu count of the same kernel, scaled to this launch) per issue slot
                        "issue_frac": (warp_inst / (upd * 1e-3) / ISSUE_PEAK) if warp_inst else None,
                        "note": "frac: 4*S scalar lane-ops per trellis step (SURVEY 8d) vs 148 SMs x 128 thread-instructions/clk "
                                "x 1.965 GHz -- packed u16x2 instructions carry two lanes, so frac can exceed 1; issue_frac: "
                                "warp-instructions ncu counted for this kernel / (148 x 4 issue slots x 1.965 GHz x kernel time)"}}

    cpu = None
    if not args.no_cpu_baseline:
        v, cores, kind, sample = cpu_reference_rate(args.codec, nbits, args.sigma, args.cpu_seconds, K=K, R=R)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8" if sym_bytes == 1 else "u16", "data": "synthetic",
        "config": {"workload": f"{args.codec} K={K} r=1/{R}: {frames} frames x {nbits} bits per GPU",
                   "sigma": args.sigma, "kernel_variant": dec.variant, "entry": "calls" if args.unfused else "decode_dev", "l2": "inputs larger than L2"
                   if frames * srow > 126e6 else "inputs smaller than L2", "sample_ber": ber},
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu,
    }
    emit(line)
    if dist is not None:
        dist.destroy_process_group()


_JSON_OUT = None


def emit(line):
    """the ONE JSON line, on the process's original stdout"""
    os.write(_JSON_OUT if _JSON_OUT is not None else 1, (json.dumps(line) + "\n").encode())


def main():
    global _JSON_OUT
    args = parse()
    # libraries print to stdout behind our back (NCCL: "NCCL version ..." at communicator creation): everything but
    # the JSON line goes to stderr
    sys.stdout.flush()
    _JSON_OUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()
