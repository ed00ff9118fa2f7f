"""Build libb200viterbi.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
import os
import shutil
import subprocess

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
# B200V_LIB: load another build of the library instead (A/B measurements of kernel variants; see tools/build_variant.py)
LIB_PATH = os.environ.get("B200V_LIB") or os.path.join(PKG_DIR, "libb200viterbi.so")
# the reference's per-codec C++ API on top of the C-ABI (include/ref_compat/): a separate library so that its
# C++-linkage symbols never meet the reference's own objects in one link
COMPAT_SRC = os.path.join(CSRC, "ref_compat.cpp")
COMPAT_LIB_PATH = os.path.join(PKG_DIR, "libb200viterbi_compat.so")
INCLUDE_DIR = os.path.join(os.path.dirname(PKG_DIR), "include")
OBJ_DIR = os.path.join(PKG_DIR, "build")

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
    # one object per source, compiled in parallel and reused while it is newer than every header and its source
    from concurrent.futures import ThreadPoolExecutor
    os.makedirs(OBJ_DIR, exist_ok=True)
    # acs_grp.cu counts as a header: acs_grp_punct.cu includes it
    hdr_time = max([os.path.getmtime(p) for p in _deps() if p.endswith((".cuh", ".h", ".hpp", "acs_grp.cu"))] +
                   [os.path.getmtime(os.path.abspath(__file__))])
    compile_flags = [f for f in NVCC_FLAGS if f != "-shared"]

    def compile_one(src):
        obj = os.path.join(OBJ_DIR, os.path.basename(src) + ".o")
        if not force and os.path.exists(obj) and os.path.getmtime(obj) > max(hdr_time, os.path.getmtime(src)):
            return obj, None
        cmd = [nvcc] + compile_flags + (["-Xptxas", "-v"] if verbose else []) + ["-I", INCLUDE_DIR, "-c", "-o", obj, src]
        return obj, subprocess.run(cmd, capture_output=True, text=True)

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as pool:
        results = list(pool.map(compile_one, sources()))
    for obj, res in results:
        if res is not None and res.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
        if res is not None and verbose:
            print(res.stderr)
    tmp = LIB_PATH + f".tmp{os.getpid()}"          # link under a temporary name: a concurrent dlopen never sees half a file
    res = subprocess.run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", tmp] +
                         [obj for obj, _ in results], capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc link failed:\n" + res.stdout + res.stderr)
    os.replace(tmp, LIB_PATH)
    cxx = shutil.which("g++") or "g++"
    tmp = COMPAT_LIB_PATH + f".tmp{os.getpid()}"
    res = subprocess.run([cxx, "-std=c++17", "-O2", "-fPIC", "-shared", "-I", INCLUDE_DIR, "-o", tmp,
                          COMPAT_SRC, "-L", PKG_DIR, "-lb200viterbi", "-Wl,-rpath,$ORIGIN"],
                         capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("building libb200viterbi_compat.so failed:\n" + res.stdout + res.stderr)
    os.replace(tmp, COMPAT_LIB_PATH)
    return LIB_PATH


def build_library_locked(force=False, verbose=False):
    """build_library under an inter-process lock: the ranks of a multi-GPU run start together, and only one of them may
    run nvcc into the shared object directory; the others wait and then find the library up to date."""
    import fcntl
    os.makedirs(OBJ_DIR, exist_ok=True)
    with open(os.path.join(OBJ_DIR, ".lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            return build_library(force=force, verbose=verbose)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


if __name__ == "__main__":
    import sys
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
