"""Build libb200viterbi.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
import os
import shutil
import subprocess

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "libb200viterbi.so")
# the reference's per-codec C++ API on top of the C-ABI (include/ref_compat/): a separate library so that its
# C++-linkage symbols never meet the reference's own objects in one link
COMPAT_SRC = os.path.join(CSRC, "ref_compat.cpp")
COMPAT_LIB_PATH = os.path.join(PKG_DIR, "libb200viterbi_compat.so")
INCLUDE_DIR = os.path.join(os.path.dirname(PKG_DIR), "include")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared",
    "-diag-suppress", "128",
]


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC)
                  if f.endswith((".cu", ".cpp")) and os.path.join(CSRC, f) != COMPAT_SRC)


def _deps():
    out = []
    for root in (CSRC, INCLUDE_DIR):
        for dp, _, files in os.walk(root):
            out += [os.path.join(dp, f) for f in files if f.endswith((".cu", ".cuh", ".h", ".cpp", ".hpp"))]
    return out


def needs_build():
    if not os.path.exists(LIB_PATH) or not os.path.exists(COMPAT_LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(p) > t for p in _deps())


def nvcc_path():
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    return None


def build_library(force=False, verbose=False):
    """Compile every CUDA source of the package into one shared library. Returns its path."""
    if not force and not needs_build():
        return LIB_PATH
    nvcc = nvcc_path()
    if nvcc is None:
        raise RuntimeError("nvcc not found: cannot build libb200viterbi.so")
    cmd = [nvcc] + NVCC_FLAGS + ["-I", INCLUDE_DIR, "-o", LIB_PATH] + sources()
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    cxx = shutil.which("g++") or "g++"
    res = subprocess.run([cxx, "-std=c++17", "-O2", "-fPIC", "-shared", "-I", INCLUDE_DIR, "-o", COMPAT_LIB_PATH,
                          COMPAT_SRC, "-L", PKG_DIR, "-lb200viterbi", "-Wl,-rpath,$ORIGIN"],
                         capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("building libb200viterbi_compat.so failed:\n" + res.stdout + res.stderr)
    return LIB_PATH


if __name__ == "__main__":
    import sys
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
