"""tools/batch_benchmark.py writes the JSON schema of the reference's benchmark driver (src/main.cpp:80-118), the
one scripts/sample_loader.py:3-20 reads: every key present, array lengths consistent, and the reference's rate
formulas (scripts/tabulate_data.py:33,54) meaningful on batch-scaled sizes."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import _oracle as o

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOOL = os.path.join(ROOT, "tools", "batch_benchmark.py")
sys.path.insert(0, os.path.join(ROOT, "tools"))


def check_schema(objs):
    import batch_benchmark as bb
    assert objs
    for v in objs:
        for k in bb.REQUIRED_KEYS:          # sample_loader.Sample.__init__ indexes exactly these
            assert k in v, (v.get("name"), k)
        assert len(v["poly"]) == v["R"]
        assert v["total_samples"] == len(v["init_ns"]) == len(v["update_ns"]) == len(v["chainback_ns"]) > 0
        T = v["frame_input_bytes"] * 8 + v["K"] - 1
        assert v["total_input_bytes"] == v["n_frames"] * v["frame_input_bytes"]
        assert v["total_transmit_bits"] == v["n_frames"] * T
        assert v["total_output_symbols"] == v["n_frames"] * T * v["R"]
        assert v["total_bits"] == v["total_input_bytes"] * 8
        assert abs(v["bit_error_rate"] - v["total_bit_errors"] / v["total_bits"]) < 1e-12
        rate = v["total_output_symbols"] / (np.array(v["update_ns"]) * 1e-9)      # tabulate_data.py:33
        assert np.all(np.isfinite(rate)) and np.all(rate > 0)


@pytest.mark.skipif(not o.ref_available(), reason="oracle/_ref not built")
def test_cpu_columns_follow_the_reference_schema(tmp_path):
    out = tmp_path / "bench.json"
    subprocess.run([sys.executable, TOOL, "--cpu", "--no-gpu", "--cpu-seconds", "0.02", "--samples", "2", "-o", str(out)],
                   check=True, timeout=600)
    objs = json.loads(out.read_text())
    check_schema(objs)
    names = {(v["name"], v["K"], v["R"]) for v in objs}
    assert ("ka9q", 7, 2) in names and ("spiral", 9, 4) in names and ("sse_u16", 15, 6) in names
    # main.cpp's symbols are noiseless: every decoder but the reference's own broken K=24 chainback returns the input
    for v in objs:
        if not (v["name"] == "ka9q" and v["K"] == 24):
            assert v["total_bit_errors"] == 0, (v["name"], v["K"], v["R"])


@pytest.mark.gpu
def test_b200_columns_follow_the_reference_schema(tmp_path):
    out = tmp_path / "bench.json"
    subprocess.run([sys.executable, TOOL, "--frames", "64", "--samples", "2", "-o", str(out)], check=True, timeout=900)
    objs = json.loads(out.read_text())
    check_schema(objs)
    assert len(objs) == 21 and all(v["name"].startswith("b200_") for v in objs)
    for v in objs:
        if not (v["name"] == "b200_ka9q" and v["K"] == 24):     # literal reproduction of chainback_viterbi224_sse2
            assert v["total_bit_errors"] == 0, (v["name"], v["K"], v["R"])
