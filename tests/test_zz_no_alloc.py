"""b200v_decode_noalloc_dev (no counterpart in the reference): the
allocation-free K=7 decode, whose decision rows / metrics / row phases live in a static __device__ array of the
library, against the existing entry point b200v_batch_decode_dev and the oracle -- bit-exact (byte work), on two calls
in a row that share the scratch array. Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import _oracle as o
import ka9q_viterbi_comparison_b200 as kv

pytestmark = pytest.mark.gpu

CASES = {"ka9q27": (kv.KA9Q27, 7, 2), "spiral27": (kv.SPIRAL27, 7, 2), "spiral47": (kv.SPIRAL47, 7, 4),
         "wy_sse_u8_k7r2": (kv.WY_U8, 7, 2)}


def device_symbols(c, codec, K, R, T, n_frames, nbits, seed):
    import torch
    sb = np.dtype(o.sym_dtype(c)).itemsize
    amp_q8, sigma_q8, mean_q8, lo, hi = o.channel_params(c, 0.7)
    d_syms = torch.zeros(n_frames * T * R * sb, dtype=torch.uint8, device="cuda")
    with kv.BatchDecoder(codec, T, n_frames, K=K, R=R) as gen:
        gen.generate_dev(d_syms, T * R * sb, None, 0, nbits, seed, 0, amp_q8, sigma_q8, mean_q8, lo, hi)
        gen.sync()
    return d_syms


@pytest.mark.parametrize("name", list(CASES))
def test_noalloc_decode_matches_the_batch_entry_point_twice_in_a_row(name):
    import torch
    codec, K, R = CASES[name]
    c = o.codec_by_name(name)
    # two different shapes back to back: the second call finds the first call's rows, metrics and phases in the array
    for call, (n_frames, nbytes, seed) in enumerate([(777, 64, 11), (300, 41, 12)]):
        nbits = nbytes * 8
        T = nbits + K - 1
        d_syms = device_symbols(c, codec, K, R, T, n_frames, nbits, seed)
        d_ref = torch.zeros(n_frames * nbytes, dtype=torch.uint8, device="cuda")
        with kv.BatchDecoder(codec, T, n_frames, K=K, R=R) as dec:
            dec.set_variant(2)                 # the register kernel, as the allocation-free entry uses
            dec.decode_dev(d_syms, T * R, d_ref, nbytes, nbits)
            dec.sync()
        d_out = torch.full((n_frames * nbytes,), 0xA5, dtype=torch.uint8, device="cuda")
        before = torch.cuda.memory_allocated()
        kv.decode_noalloc_dev(codec, T, n_frames, d_syms, T * R, d_out, nbytes, nbits, K=K, R=R)
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        ref = d_ref.cpu().numpy().reshape(n_frames, nbytes)
        out = d_out.cpu().numpy().reshape(n_frames, nbytes)
        assert np.array_equal(out, ref), (name, call)
        syms = d_syms.cpu().numpy().view(o.sym_dtype(c)).reshape(n_frames, T * R)
        for f in (0, n_frames // 2, n_frames - 1):
            q = o.Oracle(c, T)
            qo, _ = q.decode(syms[f], nbits)
            assert np.array_equal(out[f], qo), (name, call, f)
            q.close()


def test_noalloc_decode_on_a_caller_stream_and_repeated():
    """Same input twice on a non-default stream: identical output both times (nothing of call 1 leaks into call 2)."""
    import torch
    codec, K, R = CASES["ka9q27"]
    c = o.codec_by_name("ka9q27")
    n_frames, nbytes = 1024, 256
    nbits, T = nbytes * 8, nbytes * 8 + K - 1
    d_syms = device_symbols(c, codec, K, R, T, n_frames, nbits, 5)
    d_ref = torch.zeros(n_frames * nbytes, dtype=torch.uint8, device="cuda")
    with kv.BatchDecoder(codec, T, n_frames) as dec:
        dec.decode_dev(d_syms, T * R, d_ref, nbytes, nbits)
        dec.sync()
    st = torch.cuda.Stream()
    outs = []
    for _ in range(2):
        d_out = torch.zeros(n_frames * nbytes, dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        kv.decode_noalloc_dev(codec, T, n_frames, d_syms, T * R, d_out, nbytes, nbits, cuda_stream=st.cuda_stream)
        st.synchronize()
        outs.append(d_out.cpu().numpy())
    assert np.array_equal(outs[0], d_ref.cpu().numpy())
    assert np.array_equal(outs[1], d_ref.cpu().numpy())


def test_noalloc_rejects_what_does_not_fit_or_is_not_served():
    import torch
    lib = kv.load_library()
    d = torch.zeros(64, dtype=torch.uint8, device="cuda")
    poly = (C.c_int * 2)(0x6d, 0x4f)
    # 4096 frames x 2054 steps x 8 B = 67 MB of decision rows > 32 MiB
    rc = lib.b200v_decode_noalloc_dev(kv.KA9Q27, 7, 2, poly, None, 2054, 4096, d.data_ptr(), 4108, d.data_ptr(), 256, 2048,
                                      0, None)
    assert rc == -1 and b"B200V_NOALLOC_SCRATCH_BYTES" in lib.b200v_last_error()
    poly9 = (C.c_int * 2)(0x1af, 0x11d)
    rc = lib.b200v_decode_noalloc_dev(kv.KA9Q29, 9, 2, poly9, None, 72, 4, d.data_ptr(), 144, d.data_ptr(), 8, 64, 0, None)
    assert rc == -1 and b"K=7" in lib.b200v_last_error()
    # the largest batch that fits still decodes (frames x (steps*8 + 128) + small change <= 32 MiB)
    assert kv.NOALLOC_SCRATCH_BYTES == 32 << 20
