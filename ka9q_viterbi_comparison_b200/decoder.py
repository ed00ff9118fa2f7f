"""Host-side mirror of the reference's decoder interface over the C-ABI of libb200viterbi.so.

The reference wraps each decoder in an adapter with ``reset() / update(sym, total_syms) /
chainback(data, total_bits)`` (/root/reference/src/ka9q_interface.h:22-56, src/spiral_interface.h:22-56);
:class:`BatchDecoder` offers the same three calls for a batch of independent frames, plus ``decode`` (the
batch entry point). Everything here is ctypes plumbing: there is no Python/NumPy compute path, and loading
fails loudly when the CUDA library is missing.
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

# b200v_codec (include/b200_viterbi.h)
KA9Q27, KA9Q29, KA9Q615, KA9Q224, SPIRAL27, SPIRAL29, SPIRAL47, SPIRAL49, SPIRAL615, WY_U8, WY_U16 = range(11)

CODEC_BY_NAME = {
    "ka9q27": KA9Q27, "ka9q29": KA9Q29, "ka9q615": KA9Q615, "ka9q224": KA9Q224,
    "spiral27": SPIRAL27, "spiral29": SPIRAL29, "spiral47": SPIRAL47, "spiral49": SPIRAL49,
    "spiral615": SPIRAL615, "wy_u8": WY_U8, "wy_u16": WY_U16,
}

# Polynomials the reference benchmarks (src/main.cpp:363-419) and LTE K7R3 (helpers/common_codes.h:23)
STANDARD_POLYS = {
    (7, 2): [0x6d, 0x4f],
    (7, 3): [91, 117, 121],
    (7, 4): [121, 117, 91, 111],
    (9, 2): [0x1af, 0x11d],
    (9, 4): [501, 441, 331, 315],
    (15, 6): [0o42631, 0o47245, 0o56507, 0o73363, 0o77267, 0o64537],
    (24, 2): [0o62650457, 0o62650455],
}


class B200ViterbiError(RuntimeError):
    pass


class WyConfig(C.Structure):
    """ViterbiDecoder_Config + branch-table symbol values (b200v_wy_config)."""
    _fields_ = [("soft_decision_high", C.c_int), ("soft_decision_low", C.c_int),
                ("soft_decision_max_error", C.c_uint), ("initial_start_error", C.c_uint),
                ("initial_non_start_error", C.c_uint), ("renormalisation_threshold", C.c_uint)]


class PunctureSegment(C.Structure):
    """b200v_puncture_segment: one call of the reference's decode_punctured_symbols (code, length, symbols requested)."""
    _fields_ = [("code", C.POINTER(C.c_uint8)), ("code_len", C.c_int), ("n_out_syms", C.c_int)]


_lib = None


def library_path():
    return _build.LIB_PATH


def load_library():
    """dlopen libb200viterbi.so (building it first if the sources are newer). No fallback."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB_PATH
    if not os.path.exists(path) or (_build.nvcc_path() and _build.needs_build()):
        _build.build_library_locked()
    if not os.path.exists(path):
        raise B200ViterbiError(f"{path} is missing and cannot be built: the CUDA extension is required")
    lib = C.CDLL(path)
    vp, sz, i32, u32, u64 = C.c_void_p, C.c_size_t, C.c_int, C.c_uint, C.c_uint64
    sig = {
        "b200v_version": (i32, []),
        "b200v_last_error": (C.c_char_p, []),
        "b200v_device_count": (i32, []),
        "b200v_codec_shape": (i32, [i32, C.POINTER(i32), C.POINTER(i32)]),
        "b200v_codec_symbol_bytes": (i32, [i32]),
        "b200v_codec_metric_bytes": (i32, [i32]),
        "b200v_batch_create": (vp, [i32, i32, i32, C.POINTER(i32), i32, i32, i32, C.POINTER(WyConfig)]),
        "b200v_batch_destroy": (None, [vp]),
        "b200v_batch_set_stream": (i32, [vp, vp]),
        "b200v_batch_set_variant": (i32, [vp, i32]),
        "b200v_batch_set_traceback": (i32, [vp, i32, i32]),
        "b200v_batch_set_symbol_bound": (i32, [vp, i32]),
        "b200v_batch_get_variant": (i32, [vp]),
        "b200v_batch_init": (i32, [vp, i32]),
        "b200v_batch_update_dev": (i32, [vp, vp, sz, i32]),
        "b200v_batch_update": (i32, [vp, vp, sz, i32]),
        "b200v_batch_chainback_dev": (i32, [vp, vp, sz, u32, u32]),
        "b200v_batch_chainback": (i32, [vp, vp, sz, u32, u32]),
        "b200v_batch_chainback_results": (i32, [vp, vp]),
        "b200v_batch_chainback_best_dev": (i32, [vp, vp, sz, u32, vp]),
        "b200v_batch_decode": (i32, [vp, vp, vp, u32]),
        "b200v_batch_decode_dev": (i32, [vp, vp, sz, vp, sz, u32]),
        "b200v_decode_noalloc_dev": (i32, [i32, i32, i32, C.POINTER(i32), C.POINTER(WyConfig), i32, i32, vp, sz, vp, sz, u32,
                                           i32, vp]),
        "b200v_batch_sync": (i32, [vp]),
        "b200v_stream_create": (vp, [i32, i32, i32, C.POINTER(i32), i32, i32, i32, i32, C.POINTER(WyConfig)]),
        "b200v_stream_destroy": (None, [vp]),
        "b200v_stream_reset": (i32, [vp, i32]),
        "b200v_stream_push_dev": (i32, [vp, vp, sz, vp, sz, C.POINTER(u32)]),
        "b200v_stream_flush_dev": (i32, [vp, i32, vp, sz, C.POINTER(u32)]),
        "b200v_stream_push": (i32, [vp, vp, sz, vp, sz, C.POINTER(u32)]),
        "b200v_stream_flush": (i32, [vp, i32, vp, sz, C.POINTER(u32)]),
        "b200v_stream_steps": (C.c_longlong, [vp]),
        "b200v_stream_emitted": (C.c_longlong, [vp]),
        "b200v_stream_sync": (i32, [vp]),
        "b200v_stream_batch": (vp, [vp]),
        "b200v_batch_get_decisions": (i32, [vp, i32, i32, i32, vp]),
        "b200v_batch_get_metrics": (i32, [vp, i32, vp]),
        "b200v_batch_steps_done": (i32, [vp]),
        "b200v_batch_accumulated_error": (i32, [vp, vp]),
        "b200v_batch_last_kernel_ms": (i32, [vp, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
        "b200v_batch_launch_count": (C.c_longlong, [vp]),
        "b200v_batch_generate_dev": (i32, [vp, vp, sz, vp, sz, u32, u64, u64, i32, i32, i32, i32, i32]),
        "b200v_batch_depuncture_dev": (i32, [vp, vp, sz, sz, vp, i32, vp, sz, sz, i32, i32, C.POINTER(i32)]),
        "b200v_batch_set_puncturing": (i32, [vp, C.POINTER(PunctureSegment), i32, i32, C.POINTER(i32)]),
        "b200v_batch_update_punctured_dev": (i32, [vp, vp, sz]),
        "b200v_batch_decode_punctured_dev": (i32, [vp, vp, sz, vp, sz, u32]),
        "b200v_batch_decode_punctured": (i32, [vp, vp, vp, u32]),
        "b200v_shard_range": (i32, [i32, i32, i32, C.POINTER(i32), C.POINTER(i32)]),
        "b200v_sharded_create": (vp, [i32, i32, i32, C.POINTER(i32), i32, i32, C.POINTER(i32), i32, C.POINTER(WyConfig)]),
        "b200v_shard_range_weighted": (i32, [i32, C.POINTER(C.c_double), i32, i32, C.POINTER(i32), C.POINTER(i32)]),
        "b200v_sharded_create_weighted": (vp, [i32, i32, i32, C.POINTER(i32), i32, i32, C.POINTER(i32), C.POINTER(C.c_double), i32,
                                               C.POINTER(WyConfig)]),
        "b200v_sharded_destroy": (None, [vp]),
        "b200v_sharded_num_shards": (i32, [vp]),
        "b200v_sharded_shard": (i32, [vp, i32, C.POINTER(i32), C.POINTER(i32)]),
        "b200v_sharded_decode": (i32, [vp, vp, vp, u32]),
        "b200v_sharded_last_error": (C.c_char_p, [vp]),
        "b200v_bind_host_to_device": (i32, [i32]),
        "b200v_device_alloc": (vp, [i32, sz]),
        "b200v_device_free": (None, [i32, vp]),
        "b200v_host_alloc_pinned": (vp, [sz]),
        "b200v_host_free_pinned": (None, [vp]),
        "b200v_host_is_pinned": (i32, [vp]),
        "b200v_memcpy_h2d": (i32, [i32, vp, vp, sz]),
        "b200v_memcpy_d2h": (i32, [i32, vp, vp, sz]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


EXPORTED_SYMBOLS = [
    "b200v_version", "b200v_last_error", "b200v_device_count", "b200v_codec_shape",
    "b200v_codec_symbol_bytes", "b200v_codec_metric_bytes", "b200v_batch_create", "b200v_batch_destroy",
    "b200v_batch_set_stream", "b200v_batch_set_variant", "b200v_batch_get_variant", "b200v_batch_set_traceback", "b200v_batch_set_symbol_bound", "b200v_batch_init",
    "b200v_batch_update_dev", "b200v_batch_update", "b200v_batch_chainback_dev", "b200v_batch_chainback",
    "b200v_batch_chainback_results", "b200v_batch_chainback_best_dev", "b200v_batch_decode", "b200v_batch_decode_dev", "b200v_decode_noalloc_dev", "b200v_batch_sync",
    "b200v_stream_create", "b200v_stream_destroy", "b200v_stream_reset", "b200v_stream_push_dev", "b200v_stream_flush_dev", "b200v_stream_push", "b200v_stream_flush",
    "b200v_stream_steps", "b200v_stream_emitted", "b200v_stream_sync", "b200v_stream_batch", "b200v_batch_get_decisions",
    "b200v_batch_get_metrics", "b200v_batch_steps_done", "b200v_batch_accumulated_error", "b200v_batch_last_kernel_ms",
    "b200v_batch_launch_count", "b200v_batch_generate_dev", "b200v_batch_depuncture_dev", "b200v_batch_set_puncturing", "b200v_batch_update_punctured_dev",
    "b200v_batch_decode_punctured_dev", "b200v_batch_decode_punctured", "b200v_shard_range", "b200v_shard_range_weighted", "b200v_sharded_create", "b200v_sharded_create_weighted",
    "b200v_sharded_destroy", "b200v_sharded_num_shards", "b200v_sharded_shard", "b200v_sharded_decode",
    "b200v_sharded_last_error", "b200v_bind_host_to_device", "b200v_device_alloc", "b200v_device_free",
    "b200v_host_alloc_pinned", "b200v_host_free_pinned", "b200v_host_is_pinned", "b200v_memcpy_h2d", "b200v_memcpy_d2h",
]


def _check(rc):
    if rc != 0:
        raise B200ViterbiError(f"b200v error {rc}: {load_library().b200v_last_error().decode()}")


def symbol_dtype(codec):
    return {WY_U8: np.int8, WY_U16: np.int16}.get(codec, np.uint8)


def metric_dtype(codec):
    return {KA9Q615: np.int16, KA9Q224: np.int16, WY_U16: np.uint16}.get(codec, np.uint8)


def _ptr(x):
    """Device pointer of a torch tensor / raw int, or host pointer of a numpy array."""
    if x is None:
        return None
    if isinstance(x, int):
        return x
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    return x.data_ptr()   # torch.Tensor


class BatchDecoder:
    """n_frames independent frames of one codec on one GPU.

    ``frame_steps`` is the number of trellis steps per frame including the K-1 tail steps -- the ``len`` the
    reference's adapters pass to ``create_*`` (src/ka9q_interface.h:28, src/main.cpp:247).
    """

    def __init__(self, codec, frame_steps, n_frames=1, K=0, R=0, poly=None, device=0, wy_config=None):
        self.lib = load_library()
        if isinstance(codec, str):
            codec = CODEC_BY_NAME[codec]
        self.codec = codec
        k, r = C.c_int(0), C.c_int(0)
        _check(self.lib.b200v_codec_shape(codec, C.byref(k), C.byref(r)))
        self.K = k.value or K
        self.R = r.value or R
        if poly is None:
            poly = STANDARD_POLYS.get((self.K, self.R))
            if poly is None:
                raise ValueError(f"no standard polynomials for K={self.K} R={self.R}")
        self.poly = list(poly)
        self.S = 1 << (self.K - 1)
        self.frame_steps, self.n_frames, self.device = frame_steps, n_frames, device
        self.sym_dtype = symbol_dtype(codec)
        arr = (C.c_int * len(self.poly))(*self.poly)
        cfg = C.byref(wy_config) if wy_config is not None else None
        self.h = self.lib.b200v_batch_create(codec, self.K, self.R, arr, frame_steps, n_frames, device, cfg)
        if not self.h:
            raise B200ViterbiError("b200v_batch_create failed: " + self.lib.b200v_last_error().decode())

    # -- the reference adapter's three calls, batched ------------------------------------------------
    def reset(self, starting_state=0):
        _check(self.lib.b200v_batch_init(self.h, starting_state))

    def update(self, sym, total_syms=None):
        """sym: host array [n_frames, total_syms] (or flat for one frame). total_syms must be a multiple of R."""
        sym = np.ascontiguousarray(sym, dtype=self.sym_dtype).reshape(self.n_frames, -1)
        if total_syms is None:
            total_syms = sym.shape[1]
        if total_syms % self.R:
            raise ValueError("total_syms must be a multiple of the code rate")   # ka9q_interface.h:49
        _check(self.lib.b200v_batch_update(self.h, sym.ctypes.data, sym.strides[0], total_syms // self.R))

    def chainback(self, total_bits, endstate=0):
        out = np.zeros((self.n_frames, (total_bits + 7) // 8), dtype=np.uint8)
        _check(self.lib.b200v_batch_chainback(self.h, out.ctypes.data, out.strides[0], total_bits, endstate))
        return out

    # -- batch entry point ----------------------------------------------------------------------------
    def decode(self, sym, total_bits):
        sym = np.ascontiguousarray(sym, dtype=self.sym_dtype).reshape(self.n_frames, -1)
        assert sym.shape[1] == self.frame_steps * self.R
        out = np.zeros((self.n_frames, (total_bits + 7) // 8), dtype=np.uint8)
        _check(self.lib.b200v_batch_decode(self.h, sym.ctypes.data, out.ctypes.data, total_bits))
        return out

    # -- device-pointer variants (inputs already in HBM) ------------------------------------------------
    def update_dev(self, d_syms, frame_stride_bytes, nsteps):
        _check(self.lib.b200v_batch_update_dev(self.h, _ptr(d_syms), frame_stride_bytes, nsteps))

    def chainback_dev(self, d_out, frame_stride_bytes, total_bits, endstate=0):
        _check(self.lib.b200v_batch_chainback_dev(self.h, _ptr(d_out), frame_stride_bytes, total_bits, endstate))

    def chainback_best_dev(self, d_out, out_stride, nbits, d_best=None):
        """Traceback from every frame's best state (unterminated streams)."""
        _check(self.lib.b200v_batch_chainback_best_dev(self.h, _ptr(d_out), out_stride, nbits, _ptr(d_best)))

    def decode_dev(self, d_syms, sym_stride, d_out, out_stride, nbits):
        """init + update + chainback with device buffers (b200v_batch_decode_dev); asynchronous."""
        _check(self.lib.b200v_batch_decode_dev(self.h, _ptr(d_syms), sym_stride, _ptr(d_out), out_stride, nbits))

    def depuncture_dev(self, d_punctured, in_stride, in_offset, code, d_syms, out_stride, out_offset, n_out, fill=0):
        """One segment of decode_punctured_symbols' symbol expansion for all frames; returns symbols consumed."""
        code = np.ascontiguousarray(code, dtype=np.uint8)
        used = C.c_int(0)
        _check(self.lib.b200v_batch_depuncture_dev(self.h, _ptr(d_punctured), in_stride, in_offset, code.ctypes.data,
                                                   len(code), _ptr(d_syms), out_stride, out_offset, n_out, fill,
                                                   C.byref(used)))
        return used.value

    def set_puncturing(self, segments, fill=0):
        """segments: [(code, n_out_syms), ...] as the reference passes them to decode_punctured_symbols, one per segment
        of a frame; None / [] removes the pattern. Returns the transmitted symbols per frame."""
        segments = list(segments or [])
        keep = [np.ascontiguousarray(code, dtype=np.uint8) for code, _ in segments]
        arr = (PunctureSegment * max(len(segments), 1))()
        for i, (code, n_out) in enumerate(segments):
            arr[i].code = keep[i].ctypes.data_as(C.POINTER(C.c_uint8))
            arr[i].code_len = len(keep[i])
            arr[i].n_out_syms = int(n_out)
        n_tx = C.c_int(0)
        _check(self.lib.b200v_batch_set_puncturing(self.h, arr, len(segments), fill, C.byref(n_tx)))
        self.n_transmitted = n_tx.value
        return n_tx.value

    def update_punctured_dev(self, d_punctured, frame_stride_bytes):
        _check(self.lib.b200v_batch_update_punctured_dev(self.h, _ptr(d_punctured), frame_stride_bytes))

    def decode_punctured_dev(self, d_punctured, in_stride, d_out, out_stride, nbits):
        _check(self.lib.b200v_batch_decode_punctured_dev(self.h, _ptr(d_punctured), in_stride, _ptr(d_out), out_stride, nbits))

    def decode_punctured(self, punctured, total_bits):
        """Host buffers: punctured [n_frames, n_transmitted] symbols -> decoded bytes; only those symbols cross PCIe."""
        punctured = np.ascontiguousarray(punctured, dtype=self.sym_dtype).reshape(self.n_frames, -1)
        assert punctured.shape[1] == self.n_transmitted
        out = np.zeros((self.n_frames, (total_bits + 7) // 8), dtype=np.uint8)
        _check(self.lib.b200v_batch_decode_punctured(self.h, punctured.ctypes.data, out.ctypes.data, total_bits))
        return out

    def generate_dev(self, d_syms, sym_stride, d_data, data_stride, nbits, seed, frame0, amp_q8, sigma_q8,
                     mean_q8, lo, hi):
        _check(self.lib.b200v_batch_generate_dev(self.h, _ptr(d_syms), sym_stride, _ptr(d_data), data_stride,
                                                 nbits, seed, frame0, amp_q8, sigma_q8, mean_q8, lo, hi))

    def set_stream(self, cuda_stream):
        _check(self.lib.b200v_batch_set_stream(self.h, cuda_stream))

    def set_variant(self, variant):
        _check(self.lib.b200v_batch_set_variant(self.h, variant))

    def set_symbol_bound(self, max_abs_symbol):
        """WY_U16: promise max |symbol| so that updates skip the device-side range scan (and stay asynchronous)."""
        _check(self.lib.b200v_batch_set_symbol_bound(self.h, max_abs_symbol))

    def set_traceback(self, mode=0, warmup_rows=0):
        """0 = automatic, 1 = one thread per frame, 2 = block-parallel speculative traceback (verified: bit-exact)."""
        _check(self.lib.b200v_batch_set_traceback(self.h, mode, warmup_rows))

    @property
    def variant(self):
        return self.lib.b200v_batch_get_variant(self.h)

    def sync(self):
        _check(self.lib.b200v_batch_sync(self.h))

    # -- introspection -----------------------------------------------------------------------------------
    def chainback_results(self):
        out = np.zeros(self.n_frames, dtype=np.int32)
        _check(self.lib.b200v_batch_chainback_results(self.h, out.ctypes.data))
        return out

    def decisions(self, frame, step0, nsteps):
        out = np.zeros((nsteps, max(self.S // 8, 1)), dtype=np.uint8)
        _check(self.lib.b200v_batch_get_decisions(self.h, frame, step0, nsteps, out.ctypes.data))
        return out

    def metrics(self, frame):
        out = np.zeros(self.S, dtype=metric_dtype(self.codec))
        _check(self.lib.b200v_batch_get_metrics(self.h, frame, out.ctypes.data))
        return out

    def accumulated_error(self):
        """Per frame: sum of the minima subtracted by renormalisations since init (the reference's update() return value)."""
        out = np.zeros(self.n_frames, dtype=np.uint64)
        _check(self.lib.b200v_batch_accumulated_error(self.h, out.ctypes.data))
        return out

    @property
    def steps_done(self):
        return self.lib.b200v_batch_steps_done(self.h)

    def last_kernel_ms(self):
        u, c = C.c_float(0), C.c_float(0)
        _check(self.lib.b200v_batch_last_kernel_ms(self.h, C.byref(u), C.byref(c)))
        return u.value, c.value

    @property
    def launch_count(self):
        return self.lib.b200v_batch_launch_count(self.h)

    def close(self):
        if getattr(self, "h", None):
            self.lib.b200v_batch_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


class StreamDecoder:
    """Sliding-window decoder of ``n_streams`` continuous streams (b200v_stream_*): every ``push_dev`` consumes ``chunk``
    trellis steps per stream and, once ``depth + chunk`` steps are in the ring, emits the ``chunk`` oldest bits traced
    back from every stream's best state. ``flush_dev`` emits what is left at the end of the stream."""

    def __init__(self, codec, n_streams, depth, chunk, K=0, R=0, poly=None, device=0, wy_config=None):
        self.lib = load_library()
        if isinstance(codec, str):
            codec = CODEC_BY_NAME[codec]
        k, r = C.c_int(0), C.c_int(0)
        _check(self.lib.b200v_codec_shape(codec, C.byref(k), C.byref(r)))
        self.K, self.R = k.value or K, r.value or R
        poly = list(poly if poly is not None else STANDARD_POLYS[(self.K, self.R)])
        self.codec, self.n_streams, self.depth, self.chunk = codec, n_streams, depth, chunk
        self.sym_dtype = symbol_dtype(codec)
        cfg = C.byref(wy_config) if wy_config is not None else None
        self.h = self.lib.b200v_stream_create(codec, self.K, self.R, (C.c_int * len(poly))(*poly), n_streams, depth, chunk,
                                              device, cfg)
        if not self.h:
            raise B200ViterbiError("b200v_stream_create failed: " + self.lib.b200v_last_error().decode())

    def reset(self, starting_state=0):
        _check(self.lib.b200v_stream_reset(self.h, starting_state))

    def push_dev(self, d_syms, sym_stride, d_out, out_stride):
        """Returns the number of bits emitted per stream by this push (0 or ``chunk``)."""
        n = C.c_uint(0)
        _check(self.lib.b200v_stream_push_dev(self.h, _ptr(d_syms), sym_stride, _ptr(d_out), out_stride, C.byref(n)))
        return n.value

    def push(self, syms):
        """Host symbols [n_streams, chunk*R] -> emitted bytes [n_streams, chunk/8] (or None while the window fills)."""
        syms = np.ascontiguousarray(syms, dtype=self.sym_dtype).reshape(self.n_streams, -1)
        assert syms.shape[1] == self.chunk * self.R
        out = np.zeros((self.n_streams, self.chunk // 8), dtype=np.uint8)
        n = C.c_uint(0)
        _check(self.lib.b200v_stream_push(self.h, syms.ctypes.data, syms.strides[0], out.ctypes.data, out.strides[0], C.byref(n)))
        return out if n.value else None

    def flush(self, endstate=-1):
        """Remaining bits at the end of the stream: (bytes [n_streams, ceil(bits/8)], bits)."""
        out = np.zeros((self.n_streams, (self.depth + self.chunk) // 8 + 1), dtype=np.uint8)
        n = C.c_uint(0)
        _check(self.lib.b200v_stream_flush(self.h, endstate, out.ctypes.data, out.strides[0], C.byref(n)))
        return out[:, :(n.value + 7) // 8], n.value

    def flush_dev(self, d_out, out_stride, endstate=-1):
        n = C.c_uint(0)
        _check(self.lib.b200v_stream_flush_dev(self.h, endstate, _ptr(d_out), out_stride, C.byref(n)))
        return n.value

    @property
    def steps(self):
        return self.lib.b200v_stream_steps(self.h)

    @property
    def emitted(self):
        return self.lib.b200v_stream_emitted(self.h)

    def sync(self):
        _check(self.lib.b200v_stream_sync(self.h))

    def metrics(self, stream):
        out = np.zeros(1 << (self.K - 1), dtype=metric_dtype(self.codec))
        _check(self.lib.b200v_batch_get_metrics(self.lib.b200v_stream_batch(self.h), stream, out.ctypes.data))
        return out

    def close(self):
        if getattr(self, "h", None):
            self.lib.b200v_stream_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


NOALLOC_SCRATCH_BYTES = 32 << 20     # B200V_NOALLOC_SCRATCH_BYTES


def decode_noalloc_dev(codec, frame_steps, n_frames, d_syms, sym_stride, d_out, out_stride, nbits, K=0, R=0, poly=None,
                       device=0, cuda_stream=None, wy_config=None):
    """b200v_decode_noalloc_dev: init + update + chainback of a K=7 batch with device buffers and no device allocation
    (scratch in a static __device__ array of the library). Asynchronous on ``cuda_stream`` (None = default stream)."""
    lib = load_library()
    if isinstance(codec, str):
        codec = CODEC_BY_NAME[codec]
    k, r = C.c_int(0), C.c_int(0)
    _check(lib.b200v_codec_shape(codec, C.byref(k), C.byref(r)))
    K, R = k.value or K, r.value or R
    poly = list(poly if poly is not None else STANDARD_POLYS[(K, R)])
    cfg = C.byref(wy_config) if wy_config is not None else None
    _check(lib.b200v_decode_noalloc_dev(codec, K, R, (C.c_int * len(poly))(*poly), cfg, frame_steps, n_frames,
                                        _ptr(d_syms), sym_stride, _ptr(d_out), out_stride, nbits, device, cuda_stream))


def shard_range(n_frames, n_shards, shard):
    """Contiguous frame range [frame0, frame0 + count) of shard `shard` of `n_shards`."""
    f0, cnt = C.c_int(), C.c_int()
    rc = load_library().b200v_shard_range(n_frames, n_shards, shard, C.byref(f0), C.byref(cnt))
    if rc != 0:
        raise ValueError(f"bad shard arguments ({n_frames}, {n_shards}, {shard})")
    return f0.value, cnt.value


def shard_range_weighted(n_frames, weights, shard):
    """Contiguous frame range of shard `shard` when shards are proportional to `weights`."""
    w = (C.c_double * len(weights))(*[float(x) for x in weights])
    f0, cnt = C.c_int(), C.c_int()
    rc = load_library().b200v_shard_range_weighted(n_frames, w, len(weights), shard, C.byref(f0), C.byref(cnt))
    if rc != 0:
        raise ValueError(f"bad weighted shard arguments ({n_frames}, {list(weights)}, {shard})")
    return f0.value, cnt.value


class ShardedDecoder:
    """One batch instance per device inside this process (b200v_sharded_*): frames are cut into contiguous
    ranges, each decoded on its device by its own host thread; no collective."""

    def __init__(self, codec, frame_steps, n_frames, devices, K=0, R=0, poly=None, weights=None):
        self.lib = load_library()
        if isinstance(codec, str):
            codec = CODEC_BY_NAME[codec]
        k, r = C.c_int(), C.c_int()
        self.lib.b200v_codec_shape(codec, C.byref(k), C.byref(r))
        self.K, self.R = k.value or K, r.value or R
        poly = list(poly if poly is not None else STANDARD_POLYS[(self.K, self.R)])
        self.codec, self.n_frames, self.frame_steps = codec, n_frames, frame_steps
        self.sym_dtype = symbol_dtype(codec)
        devs = (C.c_int * len(devices))(*devices)
        if weights is not None:
            w = (C.c_double * len(devices))(*[float(x) for x in weights])
            self.h = self.lib.b200v_sharded_create_weighted(codec, self.K, self.R, (C.c_int * len(poly))(*poly), frame_steps,
                                                            n_frames, devs, w, len(devices), None)
        else:
            self.h = self.lib.b200v_sharded_create(codec, self.K, self.R, (C.c_int * len(poly))(*poly), frame_steps,
                                                   n_frames, devs, len(devices), None)
        if not self.h:
            raise B200ViterbiError(f"b200v_sharded_create failed: {self.lib.b200v_last_error().decode()}")

    def shards(self):
        out = []
        f0, cnt = C.c_int(), C.c_int()
        for g in range(self.lib.b200v_sharded_num_shards(self.h)):
            self.lib.b200v_sharded_shard(self.h, g, C.byref(f0), C.byref(cnt))
            out.append((f0.value, cnt.value))
        return out

    def decode(self, syms, nbits):
        syms = np.ascontiguousarray(syms, dtype=self.sym_dtype).reshape(self.n_frames, -1)
        assert syms.shape[1] == self.frame_steps * self.R
        out = np.zeros((self.n_frames, (nbits + 7) // 8), dtype=np.uint8)
        rc = self.lib.b200v_sharded_decode(self.h, syms.ctypes.data, out.ctypes.data, nbits)
        if rc != 0:
            raise B200ViterbiError(f"b200v error {rc}: {self.lib.b200v_sharded_last_error(self.h).decode()}")
        return out

    def close(self):
        if getattr(self, "h", None):
            self.lib.b200v_sharded_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
