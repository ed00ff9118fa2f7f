#!/usr/bin/env python
"""Run bench.py over every codec (one JSON line each) -> a table for profiles/. GPU box only.
    python tools/sweep.py > gpurun_out/sweep.jsonl"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RUNS = [
    ("ka9q27", []), ("spiral27", []), ("spiral47", []), ("wy_u8", ["--K", "7", "--R", "2"]),
    ("wy_u16", ["--K", "7", "--R", "2", "--frames", "262144"]),
    ("ka9q29", ["--frames", "262144"]), ("spiral29", ["--frames", "262144"]), ("spiral49", ["--frames", "262144"]),
    ("wy_u8", ["--K", "9", "--R", "2", "--frames", "262144"]),
    ("ka9q615", ["--sigma", "1.0"]), ("spiral615", ["--sigma", "1.0"]), ("wy_u8", ["--K", "15", "--R", "6", "--sigma", "1.0"]),
    ("wy_u16", ["--K", "15", "--R", "6", "--sigma", "1.0", "--frames", "1024"]),
    ("ka9q224", ["--frame-bits", "64", "--sigma", "1.0"]),
]
extra = sys.argv[1:]
for codec, opts in RUNS:
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--codec", codec, "--steps", "3", "--warmup", "3",
           "--cpu-seconds", "2"] + opts + extra
    res = subprocess.run(cmd, capture_output=True, text=True)
    line = res.stdout.strip().splitlines()[-1] if res.stdout.strip() else ""
    try:
        d = json.loads(line)
        r, c, e = d["roofline"], d.get("cpu_baseline") or {}, d.get("e2e") or {}
        print(json.dumps({"codec": codec, "opts": opts, "workload": d["config"]["workload"], "variant": d["config"]["kernel_variant"],
                          "value_mbit_s": round(d["value"], 1), "e2e_mbit_s": round(e.get("value", 0), 1),
                          "update_ms": round(r["kernel_ms"], 3), "chainback_ms": round(r["chainback_ms"], 3),
                          "hbm_frac": round(r["frac"], 4), "lane_ops_per_s": r["alu"]["lane_ops_per_s"],
                          "cpu_mbit_s": round(c.get("value", 0), 1), "cpu_cores": c.get("cores"),
                          "ber": d["config"]["sample_ber"]}), flush=True)
    except Exception as ex:
        print(json.dumps({"codec": codec, "opts": opts, "error": str(ex), "stderr": res.stderr[-400:]}), flush=True)
