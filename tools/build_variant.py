#!/usr/bin/env python
"""build_variant.py -- A/B builds of one kernel file: recompiles ONE source of the library with extra -D flags and links it
with the objects of the regular build into scratch/lib_<tag>.so. Run a measurement against it with B200V_LIB=<path>.

    python tools/build_variant.py acs_k7.cu v7 -DK7M_MINBLOCKS=7
    python tools/build_variant.py acs_k7.cu,chainback.cu,api.cu sat1 -DK7_SAT_FORM=1     # a macro several sources see
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ka9q_viterbi_comparison_b200 import build as b  # noqa: E402


def main():
    srcs, tag, flags = sys.argv[1].split(","), sys.argv[2], sys.argv[3:]
    os.makedirs(os.path.join(ROOT, "scratch"), exist_ok=True)
    cflags = [f for f in b.NVCC_FLAGS if f != "-shared"]
    objs = []
    for src in srcs:
        obj = os.path.join(ROOT, "scratch", f"{src}.{tag}.o")
        r = subprocess.run([b.nvcc_path()] + cflags + flags + ["-Xptxas", "-v", "-I", b.INCLUDE_DIR, "-c", "-o", obj,
                                                              os.path.join(b.CSRC, src)], capture_output=True, text=True)
        if r.returncode:
            sys.exit(r.stdout + r.stderr)
        open(os.path.join(ROOT, "scratch", f"{src}.{tag}.ptxas.txt"), "w").write(r.stderr)
        objs.append(obj)
    others = [os.path.join(b.OBJ_DIR, os.path.basename(s) + ".o") for s in b.sources() if os.path.basename(s) not in srcs]
    out = os.path.join(ROOT, "scratch", f"lib_{tag}.so")
    r = subprocess.run([b.nvcc_path(), "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", out] + objs + others,
                       capture_output=True, text=True)
    if r.returncode:
        sys.exit(r.stdout + r.stderr)
    print(out)


if __name__ == "__main__":
    main()
