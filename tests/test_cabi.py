"""CPU-side checks of the drop-in boundary: the library builds/loads, exports every symbol that
include/b200_viterbi.h declares, validates arguments, and refuses to compute without a GPU."""
import ctypes as C
import os
import re

import pytest

import ka9q_viterbi_comparison_b200 as kv

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200_viterbi.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200v_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = kv.load_library()
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/b200_viterbi.h but not exported"
    assert sorted(kv.EXPORTED_SYMBOLS) == names, "python binding list out of sync with the header"


def test_version_and_codec_table():
    lib = kv.load_library()
    assert lib.b200v_version() == 1
    k, r = C.c_int(), C.c_int()
    expect = {kv.KA9Q27: (7, 2), kv.KA9Q29: (9, 2), kv.KA9Q615: (15, 6), kv.KA9Q224: (24, 2),
              kv.SPIRAL27: (7, 2), kv.SPIRAL29: (9, 2), kv.SPIRAL47: (7, 4), kv.SPIRAL49: (9, 4),
              kv.SPIRAL615: (15, 6), kv.WY_U8: (0, 0), kv.WY_U16: (0, 0)}
    for codec, kr in expect.items():
        assert lib.b200v_codec_shape(codec, C.byref(k), C.byref(r)) == 0
        assert (k.value, r.value) == kr
    assert lib.b200v_codec_shape(99, C.byref(k), C.byref(r)) < 0
    assert lib.b200v_codec_symbol_bytes(kv.WY_U16) == 2 and lib.b200v_codec_symbol_bytes(kv.KA9Q27) == 1
    assert lib.b200v_codec_metric_bytes(kv.KA9Q615) == 2 and lib.b200v_codec_metric_bytes(kv.SPIRAL615) == 1


def test_argument_validation_needs_no_gpu():
    lib = kv.load_library()
    poly = (C.c_int * 2)(0x6d, 0x4f)
    assert not lib.b200v_batch_create(99, 0, 0, poly, 10, 1, 0, None)
    assert b"unknown codec" in lib.b200v_last_error()
    assert not lib.b200v_batch_create(kv.KA9Q27, 9, 2, poly, 10, 1, 0, None)      # K mismatch
    assert not lib.b200v_batch_create(kv.KA9Q27, 7, 2, poly, 0, 1, 0, None)       # no steps
    assert not lib.b200v_batch_create(kv.WY_U8, 7, 2, None, 10, 1, 0, None)       # NULL poly
    assert not lib.b200v_batch_create(kv.WY_U8, 2, 2, poly, 10, 1, 0, None)       # K too small
    assert lib.b200v_batch_init(None, 0) < 0
    lib.b200v_batch_destroy(None)                                                 # accepted, like delete_*(NULL)
    # stream geometry: chunk a multiple of 8, depth a multiple of the chunk and >= K-1, K <= 16
    for depth, chunk, msg in ((48, 20, b"chunk"), (40, 16, b"depth"), (0, 16, b"depth")):
        assert not lib.b200v_stream_create(kv.KA9Q27, 7, 2, poly, 4, depth, chunk, 0, None)
        assert msg in lib.b200v_last_error()
    poly24 = (C.c_int * 2)(0o62650457, 0o62650455)
    assert not lib.b200v_stream_create(kv.KA9Q224, 24, 2, poly24, 1, 64, 32, 0, None)
    assert b"K <= 16" in lib.b200v_last_error()
    assert lib.b200v_stream_push_dev(None, None, 0, None, 0, None) < 0
    lib.b200v_stream_destroy(None)
    # the allocation-free entry rejects what it does not serve before it touches the device
    poly9 = (C.c_int * 2)(0x1af, 0x11d)
    assert lib.b200v_decode_noalloc_dev(kv.KA9Q29, 9, 2, poly9, None, 72, 4, 1, 144, 1, 8, 64, 0, None) == -1
    assert b"K=7" in lib.b200v_last_error()
    assert lib.b200v_decode_noalloc_dev(kv.KA9Q27, 7, 2, poly, None, 2054, 4096, 1, 4108, 1, 256, 2048, 0, None) == -1
    assert b"B200V_NOALLOC_SCRATCH_BYTES" in lib.b200v_last_error()
    assert lib.b200v_batch_set_puncturing(None, None, 0, 0, None) < 0
    assert lib.b200v_batch_accumulated_error(None, None) < 0


def test_no_cpu_fallback():
    """Without a CUDA device the product must fail loudly, never compute on the host."""
    lib = kv.load_library()
    if lib.b200v_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(kv.B200ViterbiError):
        kv.BatchDecoder("ka9q27", 100, 2)


def test_product_does_not_reference_oracle():
    """The oracle is test infrastructure: nothing under the package may import, link or execute it."""
    pkg = os.path.join(ROOT, "ka9q_viterbi_comparison_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dp, f)).read()
                assert "oracle" not in text.lower() or f == "__init__.py" and False, f"{f} mentions the oracle"


def test_k7_register_kernels_do_not_spill():
    """The K=7 register kernels hold 32 metric registers per thread and are issue-bound: a spill or a register count
    that costs a resident block shows up as 5-10 %, so the build is checked. ka9q27 (mask decisions):
    at most 80 registers = 6 blocks of 128 threads per SM."""
    import re
    import shutil
    import subprocess
    obj = os.path.join(ROOT, "ka9q_viterbi_comparison_b200", "build", "acs_k7.cu.o")
    tool = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(tool):
        pytest.skip("cuobjdump not available")
    assert os.path.exists(obj), "acs_k7.cu.o missing: the library was not built by ka9q_viterbi_comparison_b200.build"
    out = subprocess.run([tool, "-res-usage", obj], capture_output=True, text=True).stdout
    m = re.search(r"acs_k7_kernelILi0ENS0_4CodeILi2ELi109ELi79ELi0ELi0EEELb0EEE\S*:\s*\n\s*REG:(\d+) STACK:(\d+)", out)
    assert m, "ka9q27 kernel not found in acs_k7.cu.o"
    assert int(m.group(1)) <= 80, m.groups()
