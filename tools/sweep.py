#!/usr/bin/env python
"""sweep.py -- the configs[4] sweep of BASELINE.json: every codec x frame length x batch size on one
GPU (run it under torchrun for more: every rank decodes its own batch, values are whole-job), each point with the
device-timed rate, the end-to-end rate through the host-buffer entry, both roofline fractions and a parity sample.
One JSON line per point. In-process: reuses bench.py's measure().

    python tools/sweep.py > gpurun_out/sweep.jsonl
    python tools/sweep.py --codecs ka9q27,ka9q615 --frame-bits 2048 --batches 1,1024
"""
import argparse
import json
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

CODECS = [("ka9q27", 7, 2), ("spiral27", 7, 2), ("spiral47", 7, 4), ("wy_u8", 7, 2), ("wy_u16", 7, 2),
          ("ka9q29", 9, 2), ("spiral29", 9, 2), ("spiral49", 9, 4), ("wy_u8", 9, 2),
          ("ka9q615", 15, 6), ("spiral615", 15, 6), ("wy_u8", 15, 6), ("wy_u16", 15, 6), ("ka9q224", 24, 2)]
HBM_BUDGET = 100e9          # decisions + symbols per point: the "1 Mi (capped by HBM)" of SURVEY 8d


def cap_frames(K, R, sb, frame_bits, batch):
    T = frame_bits + K - 1
    S = 1 << (K - 1)
    per_frame = T * (max(S // 8, 8) + R * sb) + (S * 4 if K > 16 else S * 2)
    return max(1, min(batch, int(HBM_BUDGET // per_frame)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--codecs", default="")
    ap.add_argument("--frame-bits", default="256,2048,16384")
    ap.add_argument("--batches", default="1,1024,65536,1048576")
    ap.add_argument("--steps", type=int, default=3)
    a = ap.parse_args()
    want = set(a.codecs.split(",")) if a.codecs else None
    ranks = bench.Ranks()
    import torch
    torch.cuda.set_device(ranks.local)
    args = types.SimpleNamespace(variant=0, unfused=False, no_e2e=False, no_cpu_baseline=True, parity_frames=8, K=7, R=2)
    for codec, K, R in CODECS:
        if want and codec not in want:
            continue
        sb = 2 if codec == "wy_u16" else 1
        for fb in [int(x) for x in a.frame_bits.split(",")]:
            if K == 24 and fb > 256:
                continue                                   # 1 MiB of decisions per step and frame
            if K == 24:
                fb = 64
            seen = set()
            for batch in [int(x) for x in a.batches.split(",")]:
                frames = cap_frames(K, R, sb, fb, batch)
                if K == 24:
                    frames = min(frames, 16)
                if frames in seen:
                    continue
                seen.add(frames)
                w = {"codec": codec, "K": K, "R": R, "frames": frames, "frame_bits": fb, "sigma": bench.default_sigma(K)}
                try:
                    d = bench.measure(args, ranks, w, a.steps, 3, 0.0, False)
                except Exception as ex:      # a point that does not fit is reported, the sweep goes on
                    d = {"error": f"{type(ex).__name__}: {ex}"[:300]}
                    torch.cuda.empty_cache()
                if ranks.rank != 0:
                    continue
                if "error" in d:
                    print(json.dumps({"codec": codec, "K": K, "R": R, "frame_bits": fb, "frames": frames, **d}), flush=True)
                    continue
                r, e = d["roofline"], d.get("e2e") or {}
                print(json.dumps({
                    "codec": codec, "K": K, "R": R, "frame_bits": fb, "frames_per_gpu": frames, "batch_requested": batch,
                    "n_gpus": ranks.world, "variant": d["detail"]["kernel_variant"],
                    "value_mbit_s": round(d["value"], 1), "e2e_mbit_s": round(e.get("value", 0.0), 1),
                    "ms_per_step": round(d["ms_per_step"], 4), "update_ms": round(r["kernel_ms"], 4),
                    "chainback_ms": round(r["chainback"]["kernel_ms"], 4),
                    "int_issue_frac": round(r["frac"], 4), "hbm_frac": round(r["hbm"]["frac"], 4),
                    "parity_frames": d["parity_sample"]["frames"], "parity_mismatches": d["parity_sample"]["mismatches"]
                    + d["parity_sample"]["decision_mismatches"], "sample_ber": d["detail"]["sample_ber"],
                    "rejected": d.get("rejected")}), flush=True)
    ranks.close()


if __name__ == "__main__":
    main()
