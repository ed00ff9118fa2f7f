"""ctypes bindings for the test-side checkers (TEST INFRASTRUCTURE ONLY):

* ``Oracle``  -> oracle/liboracle.so   (C restatement, oracle/viterbi_oracle.c)
* ``Ref``     -> oracle/_ref/libkvref.so (the unmodified reference compiled from /root/reference)

plus the shared synthetic-input generator (encoder + integer AWGN channel) so that the oracle, the compiled
reference and the CUDA path all consume byte-identical symbol buffers.
"""
import ctypes as C
import math
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "liboracle.so")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "libkvref.so")
REF_NATIVE_SO = os.path.join(ORACLE_DIR, "_ref", "libkvref_native.so")          # the reference preset's -march=native build
REF_NATIVE_FLAGS = os.path.join(ORACLE_DIR, "_ref", "native_cpu_flags.txt")     # ISA flags of the host that built it


def native_ref_usable():
    """libkvref_native.so may run here iff this host has every CPU flag of the host that compiled it."""
    import re
    isa = re.compile(r"^(sse|ssse|avx|amx|bmi|fma|f16c|adx|aes|vaes|sha|gfni|movbe|movdir|popcnt|abm|pclmul|vpclmul|rdrand|rdseed|"
                     r"clflushopt|clwb|cldemote|xsave|fsgsbase|rdpid|pku|ospke|waitpkg|serialize|lahf|cx16|erms|fsrm|invpcid|rtm|hle)")
    try:
        need = set(f for f in open(REF_NATIVE_FLAGS).read().split() if isa.match(f))
        have = set()
        for line in open("/proc/cpuinfo"):
            if line.startswith("flags"):
                have = set(line.split(":", 1)[1].split())
                break
        return os.path.exists(REF_NATIVE_SO) and bool(need) and need <= have
    except OSError:
        return False


def use_native_ref():
    """Point Ref at the -march=native build (bench.py's CPU arm; call before the first ref_lib()). Returns True if done."""
    global REF_SO
    if _ref is None and native_ref_usable():
        REF_SO = REF_NATIVE_SO
        return True
    return False

MAX_R = 8


class KvoCodec(C.Structure):
    _fields_ = [("policy", C.c_int), ("K", C.c_int), ("R", C.c_int), ("poly", C.c_int * MAX_R),
                ("renorm_threshold", C.c_int), ("init_start", C.c_int), ("init_other", C.c_int),
                ("max_error", C.c_int), ("soft_hi", C.c_int), ("soft_lo", C.c_int)]


def build_oracle():
    """(Re)build the checker libraries. liboracle.so always; _ref only where /root/reference exists."""
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "oracle", "ref"], check=True,
                   stdout=subprocess.DEVNULL)


_lib = None


def oracle_lib():
    global _lib
    if _lib is None:
        if not os.path.exists(ORACLE_SO):
            build_oracle()
        lib = C.CDLL(ORACLE_SO)
        lib.kvo_codec_by_name.argtypes = [C.POINTER(KvoCodec), C.c_char_p]
        lib.kvo_codec_defaults.argtypes = [C.POINTER(KvoCodec), C.c_int, C.c_int, C.c_int]
        lib.kvo_create.restype = C.c_void_p
        lib.kvo_create.argtypes = [C.POINTER(KvoCodec), C.c_int]
        lib.kvo_init.argtypes = [C.c_void_p, C.c_int]
        lib.kvo_update.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        lib.kvo_update.restype = None
        lib.kvo_chainback.argtypes = [C.c_void_p, C.c_void_p, C.c_uint, C.c_uint]
        lib.kvo_destroy.argtypes = [C.c_void_p]
        lib.kvo_destroy.restype = None
        lib.kvo_get_decisions.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        lib.kvo_get_decisions.restype = None
        lib.kvo_get_metrics.argtypes = [C.c_void_p, C.c_void_p]
        lib.kvo_get_metrics.restype = None
        lib.kvo_renorm_sum.argtypes = [C.c_void_p]
        lib.kvo_renorm_sum.restype = C.c_uint64
        lib.kvo_encode_bits.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_int), C.c_void_p, C.c_size_t, C.c_void_p]
        lib.kvo_encode_bits.restype = None
        lib.kvo_random_bytes.argtypes = [C.c_uint64, C.c_uint64, C.c_void_p, C.c_size_t]
        lib.kvo_random_bytes.restype = None
        lib.kvo_channel.argtypes = [C.c_void_p, C.c_size_t, C.c_uint64, C.c_uint64,
                                    C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        lib.kvo_channel.restype = None
        lib.kvo_hash64.argtypes = [C.c_uint64, C.c_uint64]
        lib.kvo_hash64.restype = C.c_uint64
        _lib = lib
    return _lib


def codec_by_name(name, poly=None):
    c = KvoCodec()
    if oracle_lib().kvo_codec_by_name(C.byref(c), name.encode()) != 0:
        raise ValueError(f"unknown codec {name}")
    if poly is not None:
        assert len(poly) == c.R
        for i, p in enumerate(poly):
            c.poly[i] = int(p)
    return c


def codec_wy(width, K, R, poly):
    """williamyang policy (width 8 or 16) for any (K, R, polynomials), e.g. the common_codes.h entries."""
    c = KvoCodec()
    if oracle_lib().kvo_codec_defaults(C.byref(c), 6 if width == 8 else 7, K, R) != 0:
        raise ValueError((width, K, R))
    assert len(poly) == R
    for i in range(MAX_R):
        c.poly[i] = int(poly[i]) if i < R else 0
    return c


# williamyang_viterbi_lib/examples/helpers/common_codes.h:23-30
COMMON_CODES = [
    ("Basic K=3 R=1/2", 3, 2, [0b111, 0b101]), ("Basic K=5 R=1/2", 5, 2, [0b10111, 0b11001]),
    ("Voyager", 7, 2, [109, 79]), ("LTE", 7, 3, [91, 117, 121]), ("DAB Radio", 7, 4, [109, 79, 83, 109]),
    ("CDMA IS-95A", 9, 2, [491, 369]), ("CDMA 2000", 9, 4, [501, 441, 331, 315]),
    ("Cassini", 15, 6, [17817, 20133, 23879, 30451, 32439, 26975]),
]


def sym_dtype(codec):
    return {6: np.int8, 7: np.int16}.get(codec.policy, np.uint8)


def metric_dtype(codec):
    return {1: np.int16, 2: np.int16, 7: np.uint16}.get(codec.policy, np.uint8)


class _Decoder:
    """Common five-call surface: init / update / chainback / decisions / metrics."""

    def __init__(self, codec, len_steps):
        self.codec = codec
        self.K, self.R = codec.K, codec.R
        self.S = 1 << (codec.K - 1)
        self.len = len_steps

    def decode(self, syms, nbits):
        """reset + update(all) + chainback -> (decoded bytes, decisions[T,S/8])."""
        T = syms.size // self.R
        self.init(0)
        self.update(syms, T)
        out = self.chainback(nbits)
        return out, self.decisions(0, T)


class Oracle(_Decoder):
    def __init__(self, codec, len_steps):
        super().__init__(codec, len_steps)
        self.lib = oracle_lib()
        self.h = self.lib.kvo_create(C.byref(codec), len_steps)
        assert self.h

    def init(self, start=0):
        return self.lib.kvo_init(self.h, start)

    def update(self, syms, nsteps):
        syms = np.ascontiguousarray(syms, dtype=sym_dtype(self.codec))
        self.lib.kvo_update(self.h, syms.ctypes.data, nsteps)

    def chainback(self, nbits, endstate=0, return_code=False):
        out = np.zeros((nbits + 7) // 8, dtype=np.uint8)
        rc = self.lib.kvo_chainback(self.h, out.ctypes.data, nbits, endstate)
        return (out, rc) if return_code else out

    def decisions(self, step0, nsteps):
        out = np.zeros((nsteps, max(self.S // 8, 1)), dtype=np.uint8)
        self.lib.kvo_get_decisions(self.h, step0, nsteps, out.ctypes.data)
        return out

    def metrics(self):
        out = np.zeros(self.S, dtype=metric_dtype(self.codec))
        self.lib.kvo_get_metrics(self.h, out.ctypes.data)
        return out

    def renorm_sum(self):
        """sum of the renormalisation minima since init (the williamyang update()'s return value, accumulated)"""
        return int(self.lib.kvo_renorm_sum(self.h))

    def close(self):
        if self.h:
            self.lib.kvo_destroy(self.h)
            self.h = None

    __del__ = close


class _KvrefVtable(C.Structure):
    _fields_ = [("name", C.c_char_p), ("K", C.c_int), ("R", C.c_int), ("sym_bytes", C.c_int),
                ("metric_bytes", C.c_int),
                ("create", C.CFUNCTYPE(C.c_void_p, C.POINTER(C.c_int), C.c_int)),
                ("init", C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int)),
                ("update", C.CFUNCTYPE(None, C.c_void_p, C.c_void_p, C.c_int)),
                ("chainback", C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_uint, C.c_uint)),
                ("destroy", C.CFUNCTYPE(None, C.c_void_p)),
                ("get_decisions", C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_int, C.c_void_p)),
                ("get_metrics", C.CFUNCTYPE(None, C.c_void_p, C.c_void_p))]


_ref = None


def ref_available():
    return os.path.exists(REF_SO)


def ref_lib():
    global _ref
    if _ref is None:
        lib = C.CDLL(REF_SO)
        lib.kvref_find.restype = C.POINTER(_KvrefVtable)
        lib.kvref_find.argtypes = [C.c_char_p]
        lib.kvref_count.restype = C.c_int
        lib.kvref_at.restype = C.POINTER(_KvrefVtable)
        lib.kvref_at.argtypes = [C.c_int]
        if hasattr(lib, "kvref_wy_error_sum"):
            lib.kvref_wy_error_sum.restype = C.c_uint64
            lib.kvref_wy_error_sum.argtypes = [C.c_void_p]
        if hasattr(lib, "kvref_time_decode"):
            lib.kvref_time_decode.restype = C.c_double
            lib.kvref_time_decode.argtypes = [C.POINTER(_KvrefVtable), C.c_void_p, C.c_void_p, C.c_int, C.c_size_t,
                                              C.c_int, C.c_uint, C.c_double, C.POINTER(C.c_longlong), C.c_void_p]
        _ref = lib
    return _ref


def ref_names():
    lib = ref_lib()
    return [lib.kvref_at(i).contents.name.decode() for i in range(lib.kvref_count())]


class Ref(_Decoder):
    """The compiled, unmodified reference decoder `name` (standard polys only: the reference keeps its
    branch table in a process-global that the first create() fills)."""

    def __init__(self, name, len_steps, codec=None):
        """`codec`: a williamyang codec with its own polynomials (their branch table is per instance)."""
        assert codec is None or name.startswith("wy_")
        codec = codec or codec_by_name(name)
        super().__init__(codec, len_steps)
        vt = ref_lib().kvref_find(name.encode())
        if not vt:
            raise ValueError(name)
        self.vt = vt.contents
        poly = (C.c_int * self.R)(*[codec.poly[i] for i in range(self.R)])
        self.h = self.vt.create(poly, len_steps)
        assert self.h

    def init(self, start=0):
        return self.vt.init(self.h, start)

    def update(self, syms, nsteps):
        syms = np.ascontiguousarray(syms, dtype=sym_dtype(self.codec))
        self.vt.update(self.h, syms.ctypes.data, nsteps)

    def chainback(self, nbits, endstate=0, return_code=False):
        out = np.zeros((nbits + 7) // 8, dtype=np.uint8)
        rc = self.vt.chainback(self.h, out.ctypes.data, nbits, endstate)
        return (out, rc) if return_code else out

    def decisions(self, step0, nsteps):
        out = np.zeros((nsteps, max(self.S // 8, 1)), dtype=np.uint8)
        self.vt.get_decisions(self.h, step0, nsteps, out.ctypes.data)
        return out

    def metrics(self):
        out = np.zeros(self.S, dtype=metric_dtype(self.codec))
        self.vt.get_metrics(self.h, out.ctypes.data)
        return out

    def renorm_sum(self):
        """williamyang decoders: sum of update<uint64_t>()'s return values since init"""
        assert self.vt.name.decode().startswith("wy_")
        return int(ref_lib().kvref_wy_error_sum(self.h))

    def time_decode(self, pool, T, nbits, seconds):
        """Decode the frames of `pool` ([n][T*R] symbols) in a C loop for ~`seconds` (GIL released).
        Returns (frames decoded, elapsed seconds)."""
        pool = np.ascontiguousarray(pool, dtype=sym_dtype(self.codec))
        out = np.zeros((nbits + 7) // 8 + 8, dtype=np.uint8)
        n = C.c_longlong(0)
        dt = ref_lib().kvref_time_decode(self.vt, self.h, pool.ctypes.data, pool.shape[0], pool.strides[0], T, nbits,
                                         seconds, C.byref(n), out.ctypes.data)
        return n.value, dt

    def close(self):
        if getattr(self, "h", None):
            self.vt.destroy(self.h)
            self.h = None

    __del__ = close


# ------------------------------------------------------------------------------------------------
# synthetic inputs
# ------------------------------------------------------------------------------------------------

def soft_range(codec):
    """(lo, hi) symbol alphabet: ka9q/spiral 0..255, wy soft8 +-3, wy soft16 +-127."""
    if codec.policy == 6:
        return -3, 3
    if codec.policy == 7:
        return -127, 127
    return 0, 255


def random_bytes(seed, frame, nbytes):
    out = np.zeros(nbytes, dtype=np.uint8)
    oracle_lib().kvo_random_bytes(seed, frame, out.ctypes.data, nbytes)
    return out


def encode_bits(codec, data):
    data = np.ascontiguousarray(data, dtype=np.uint8)
    n = (data.size * 8 + codec.K - 1) * codec.R
    out = np.zeros(n, dtype=np.uint8)
    oracle_lib().kvo_encode_bits(codec.K, codec.R, codec.poly, data.ctypes.data, data.size, out.ctypes.data)
    return out


def channel_params(codec, sigma):
    """Fixed-point channel constants following run_snr_ber.cpp:317-365:
    soft = clamp(round(x*mag/sqrt(1+sigma^2) + mean)), x = +-1 + N(0, sigma^2)."""
    lo, hi = soft_range(codec)
    mag = (hi - lo) / 2.0
    mean_q8 = int(round(256 * (hi + lo) / 2.0))
    amp = mag / math.sqrt(1.0 + sigma * sigma) if sigma > 0 else mag
    return int(round(amp * 256)), int(round(amp * sigma * 256)), mean_q8, lo, hi


def make_symbols(codec, data, sigma=0.0, seed=0xB200, counter0=0):
    """Encode `data` and pass it through the integer channel. sigma=0 -> main.cpp's noiseless symbols."""
    bits = encode_bits(codec, data)
    amp_q8, sigma_q8, mean_q8, lo, hi = channel_params(codec, sigma)
    out = np.zeros(bits.size, dtype=np.int32)
    oracle_lib().kvo_channel(bits.ctypes.data, bits.size, seed, counter0, amp_q8, sigma_q8, mean_q8, lo, hi,
                             out.ctypes.data)
    return out.astype(sym_dtype(codec))


def make_frame(codec, nbytes, sigma=0.0, seed=0xB200, frame=0):
    data = random_bytes(seed, frame, nbytes)
    T = nbytes * 8 + codec.K - 1
    syms = make_symbols(codec, data, sigma, seed, counter0=frame * T * codec.R)
    return data, syms
