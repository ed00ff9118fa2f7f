"""Depuncturing front-end. CPU: the restatement in oracle/puncture_oracle.py equals the reference's
own helpers (compiled into oracle/_ref). GPU: b200v_batch_depuncture_dev equals the restatement; a DAB fast
information channel frame (K=7 r=1/4, codes PI_16 / PI_15 / PI_X) punctured on the host decodes to the data, and
matches the oracle decoder fed with the restated depunctured stream."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

import _oracle as o

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import puncture_oracle as po  # noqa: E402


def _ref():
    if not o.ref_available():
        pytest.skip("oracle/_ref not built")
    lib = o.ref_lib()
    if not hasattr(lib, "kvref_depuncture_s8"):
        pytest.skip("oracle/_ref predates the puncture shim")
    lib.kvref_depuncture_s8.restype = C.c_long
    lib.kvref_depuncture_s16.restype = C.c_long
    lib.kvref_puncture_encode_s8.restype = C.c_long
    return lib


@pytest.mark.parametrize("R", [2, 4])
def test_depuncture_restatement_matches_reference_helper(R):
    lib = _ref()
    rng = np.random.default_rng(7)
    for code in (po.PI_16, po.PI_15, po.PI_X, np.array([1, 0, 1, 1, 0, 1, 1], dtype=np.uint8)):
        for requested in (R, 24, 32 * R * 3, 100 * R):
            n_in = int(code.sum()) * (requested // len(code) + 2)
            for dt, fn in ((np.int8, lib.kvref_depuncture_s8), (np.int16, lib.kvref_depuncture_s16)):
                tx = rng.integers(-100, 100, n_in).astype(dt)
                out = np.zeros(requested + R, dtype=dt)
                used = fn(C.c_int(R), C.c_void_p(tx.ctypes.data), C.c_size_t(n_in), C.c_void_p(code.ctypes.data),
                          C.c_size_t(len(code)), C.c_size_t(requested), C.c_int(5), C.c_void_p(out.ctypes.data))
                want, want_used = po.depuncture(tx, code, requested, 5, R)
                assert used == want_used
                assert np.array_equal(out[:len(want)], want)


def test_puncture_restatement_matches_reference_encoder():
    lib = _ref()
    c = o.codec_by_name("wy_sse_u8_k7r4")
    poly = (C.c_int * 4)(*[c.poly[i] for i in range(4)])
    data = o.random_bytes(3, 0, 12)
    out = np.zeros(12 * 8 * 4 + 64, dtype=np.int8)
    n = lib.kvref_puncture_encode_s8(C.c_int(7), C.c_int(4), poly, C.c_void_p(data.ctypes.data), C.c_size_t(12),
                                     C.c_void_p(po.PI_15.ctypes.data), C.c_size_t(32), C.c_int(1),
                                     C.c_void_p(po.PI_X.ctypes.data), C.c_size_t(24), C.c_int(3), C.c_int(-3),
                                     C.c_void_p(out.ctypes.data), C.c_size_t(out.size))
    _, full = o.make_frame(c, 12, 0.0, seed=3, frame=0)          # noiseless full-rate symbols incl. the tail
    want = np.concatenate([po.puncture(full[:12 * 8 * 4], po.PI_15), po.puncture(full[12 * 8 * 4:], po.PI_X)])
    assert n == want.size
    assert np.array_equal(out[:n], want)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["wy_sse_u8_k7r4", "wy_sse_u16_k7r4", "spiral47"])
def test_gpu_depuncture_and_dab_fic_round_trip(name):
    import torch
    import ka9q_viterbi_comparison_b200 as kv
    from test_gpu_parity import CASES
    codec, K, R = CASES[name]
    c = o.codec_by_name(name)
    nbytes, n_frames = 96, 5                                   # 768 bits = 21 + 3 blocks of 32 bits
    nbits = nbytes * 8
    T = nbits + K - 1
    lo, hi = o.soft_range(c)
    fill = (lo + hi + 1) // 2                                  # erasure: 0 for the wy alphabets, 128 for 0..255
    dt = o.sym_dtype(c)
    segs = po.dab_fic_segments(R)
    datas, txs, fulls = [], [], []
    for f in range(n_frames):
        d, full = o.make_frame(c, nbytes, 0.4 if f else 0.0, seed=21, frame=f)
        pos, parts = 0, []
        for code, n_out in segs:
            parts.append(po.puncture(full[pos:pos + n_out], code))
            pos += n_out
        assert pos == T * R
        datas.append(d); txs.append(np.concatenate(parts)); fulls.append(full)
    n_tx = txs[0].size
    tx = np.stack(txs).astype(dt)
    sb = np.dtype(dt).itemsize
    d_tx = torch.from_numpy(tx.view(np.uint8).reshape(n_frames, -1).copy()).cuda()
    d_syms = torch.zeros(n_frames * T * R * sb, dtype=torch.uint8, device="cuda")
    d_out = torch.zeros(n_frames * nbytes, dtype=torch.uint8, device="cuda")
    with kv.BatchDecoder(codec, T, n_frames, K=K, R=R) as dec:
        ip = op = 0
        for code, n_out in segs:
            ip += dec.depuncture_dev(d_tx, n_tx * sb, ip, code, d_syms, T * R * sb, op, n_out, fill)
            op += n_out
        assert ip == n_tx and op == T * R
        dec.sync()
        got = d_syms.cpu().numpy().view(dt).reshape(n_frames, T * R)
        dec.decode_dev(d_syms, T * R * sb, d_out, nbytes, nbits)
        dec.sync()
        out = d_out.cpu().numpy().reshape(n_frames, nbytes)
    for f in range(n_frames):
        ip, parts = 0, []
        for code, n_out in segs:
            s, used = po.depuncture(tx[f, ip:], code, n_out, fill, R)
            parts.append(s); ip += used
        want = np.concatenate(parts)
        assert np.array_equal(got[f], want), f
        q = o.Oracle(c, T)
        qo, _ = q.decode(want, nbits)
        assert np.array_equal(out[f], qo), f
        q.close()
    assert np.array_equal(out[0], datas[0])                    # the noiseless frame decodes to the data


@pytest.mark.gpu
def test_gpu_depuncture_more_than_65535_frames():
    """The headline batch has 1 Mi frames: the frame index must not live on a 65535-capped grid dimension."""
    import torch
    import ka9q_viterbi_comparison_b200 as kv
    n_frames, n_out, R = 70001, 96, 4
    code = po.PI_15
    kept = int(sum(code[k % len(code)] for k in range(n_out)))
    rng = np.random.default_rng(5)
    tx = rng.integers(-3, 4, (n_frames, kept)).astype(np.int8)
    d_tx = torch.from_numpy(tx.view(np.uint8).copy()).cuda()
    d_syms = torch.zeros(n_frames * n_out, dtype=torch.uint8, device="cuda")
    with kv.BatchDecoder(kv.WY_U8, n_out // R, n_frames, K=7, R=R, poly=[109, 79, 83, 109]) as dec:
        used = dec.depuncture_dev(d_tx, kept, 0, code, d_syms, n_out, 0, n_out, 0)
        dec.sync()
    assert used == kept
    got = d_syms.cpu().numpy().view(np.int8).reshape(n_frames, n_out)
    for f in (0, 65535, 65536, n_frames - 1):
        want, _ = po.depuncture(tx[f], code, n_out, 0, R)
        assert np.array_equal(got[f], want), f


def _punctured_batch(c, R, K, nbytes, n_frames, segs, sigma, seed):
    """Full-rate frames from the generator, punctured on the host segment by segment. Returns (data, full, tx)."""
    datas, fulls, txs = [], [], []
    for f in range(n_frames):
        d, full = o.make_frame(c, nbytes, sigma if f else 0.0, seed=seed, frame=f)
        pos, parts = 0, []
        for code, n_out in segs:
            parts.append(po.puncture(full[pos:pos + n_out], code))
            pos += n_out
        assert pos == full.size
        datas.append(d); fulls.append(full); txs.append(np.concatenate(parts))
    return np.stack(datas), np.stack(fulls), np.stack(txs).astype(o.sym_dtype(c))


@pytest.mark.gpu
@pytest.mark.parametrize("name,variant", [("wy_sse_u8_k7r4", 2), ("wy_sse_u8_k7r4", 3), ("wy_sse_u16_k7r4", 2), ("spiral47", 2),
                                          ("ka9q27", 2), ("ka9q27", 1), ("ka9q29", 2), ("spiral49", 2), ("wy_sse_u8_k9r2", 2),
                                          ("ka9q615", 2), ("wy_sse_u16_k15r6", 2)])
def test_fused_depuncturing_equals_expand_then_update(name, variant):
    """b200v_batch_update_punctured_dev (gather inside the symbol fetch of the register / group kernels, expansion for the
    other families) against b200v_batch_depuncture_dev + b200v_batch_update_dev and against the oracle fed with the
    restated depunctured stream: decision vectors, metrics and decoded bytes, bit for bit."""
    import torch
    import ka9q_viterbi_comparison_b200 as kv
    from test_gpu_parity import CASES
    codec, K, R = CASES[name]
    c = o.codec_by_name(name)
    n_frames = 3 if K == 15 else 37
    nbytes = 24 if K == 15 else 96
    nbits = nbytes * 8
    T = nbits + K - 1
    lo, hi = o.soft_range(c)
    fill = (lo + hi + 1) // 2
    dt = o.sym_dtype(c)
    sb = np.dtype(dt).itemsize
    # three segments with different codes, the last one ragged against its code length
    a, b = (T * R * 5 // 8) // R * R, (T * R // 4) // R * R
    segs = [(po.PI_16, a), (po.PI_15, b), (np.array([1, 1, 0, 1, 0, 1, 1], dtype=np.uint8), T * R - a - b)]
    data, full, tx = _punctured_batch(c, R, K, nbytes, n_frames, segs, 0.35, 77)
    n_tx = tx.shape[1]
    d_tx = torch.from_numpy(tx.view(np.uint8).reshape(n_frames, -1).copy()).cuda()
    d_out = torch.zeros(n_frames * nbytes, dtype=torch.uint8, device="cuda")
    with kv.BatchDecoder(codec, T, n_frames, K=K, R=R) as dec:
        dec.reset()
        dec.set_variant(variant)
        assert dec.set_puncturing(segs, fill) == n_tx
        dec.decode_punctured_dev(d_tx, n_tx * sb, d_out, nbytes, nbits)
        dec.sync()
        assert dec.variant == variant
        out = d_out.cpu().numpy().reshape(n_frames, nbytes)
        decs = [dec.decisions(f, 0, T) for f in range(n_frames)]
        mets = [dec.metrics(f) for f in range(n_frames)]
        # the unfused route on the same instance
        d_syms = torch.zeros(n_frames * T * R * sb, dtype=torch.uint8, device="cuda")
        ip = op = 0
        for code, n_out in segs:
            ip += dec.depuncture_dev(d_tx, n_tx * sb, ip, code, d_syms, T * R * sb, op, n_out, fill)
            op += n_out
        d_out2 = torch.zeros_like(d_out)
        dec.decode_dev(d_syms, T * R * sb, d_out2, nbytes, nbits)
        dec.sync()
        assert np.array_equal(d_out2.cpu().numpy().reshape(n_frames, nbytes), out)
        for f in (0, n_frames - 1):
            assert np.array_equal(dec.decisions(f, 0, T), decs[f])
        # host entry: only the transmitted symbols are handed over
        assert np.array_equal(dec.decode_punctured(tx, nbits), out)
        # a second pattern replaces the first; none = error
        dec.set_puncturing(None)
        with pytest.raises(kv.B200ViterbiError):
            dec.update_punctured_dev(d_tx, n_tx * sb)
    for f in range(n_frames):
        ip, parts = 0, []
        for code, n_out in segs:
            s, used = po.depuncture(tx[f, ip:], code, n_out, fill, R)
            parts.append(s[:n_out]); ip += used
        want = np.concatenate(parts)
        q = o.Oracle(c, T)
        qo, qd = q.decode(want, nbits)
        assert np.array_equal(decs[f], qd), f
        assert np.array_equal(mets[f], q.metrics()), f
        assert np.array_equal(out[f], qo), f
        q.close()
    assert np.array_equal(out[0], data[0])                     # the noiseless frame decodes to the data


@pytest.mark.gpu
def test_pipelined_punctured_host_decode():
    """b200v_batch_decode_punctured on a batch large enough for the chunk pipeline (copy / compute overlap): equals the
    full-rate b200v_batch_decode of the expanded symbols; rate 3/4 of a rate-1/4 mother code as DAB's PI_16."""
    import torch
    import ka9q_viterbi_comparison_b200 as kv
    codec, K, R = kv.WY_U8, 7, 4
    poly = [109, 79, 83, 109]
    n_frames, nbytes = 9000, 32
    nbits = nbytes * 8
    T = nbits + K - 1
    segs = [(po.PI_16, T * R)]
    c = o.codec_wy(8, K, R, poly)
    amp_q8, sigma_q8, mean_q8, lo, hi = o.channel_params(c, 0.4)
    with kv.BatchDecoder(codec, T, n_frames, K=K, R=R, poly=poly) as dec:
        d_syms = torch.zeros(n_frames * T * R, dtype=torch.uint8, device="cuda")
        dec.generate_dev(d_syms, T * R, None, 0, nbits, 5, 0, amp_q8, sigma_q8, mean_q8, lo, hi)
        dec.sync()
        full = d_syms.cpu().numpy().view(np.int8).reshape(n_frames, T * R)
        keep = np.resize(po.PI_16.astype(bool), T * R)
        tx = np.ascontiguousarray(full[:, keep])
        erased = full.copy()
        erased[:, ~keep] = 0
        assert dec.set_puncturing(segs, 0) == tx.shape[1]
        assert dec.variant == 2
        got = dec.decode_punctured(tx, nbits)
        want = dec.decode(erased, nbits)
        assert np.array_equal(got, want)
        for f in (0, 4999, n_frames - 1):
            q = o.Oracle(c, T)
            qo, _ = q.decode(erased[f], nbits)
            assert np.array_equal(got[f], qo), f
            q.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["ka9q27", "wy_sse_u8_k9r2", "ka9q615"])
def test_punctured_update_continues_across_calls(name):
    """A pattern that covers half a frame, applied twice (ka9q / williamyang updates continue from the previous call's
    position): equals one full-rate update of the expanded symbols."""
    import torch
    import ka9q_viterbi_comparison_b200 as kv
    from test_gpu_parity import CASES
    codec, K, R = CASES[name]
    c = o.codec_by_name(name)
    n_frames = 3 if K == 15 else 11
    half = 60 if K == 15 else 132                       # steps per call
    T = 2 * half
    nbits = T - (K - 1)
    lo, hi = o.soft_range(c)
    fill = (lo + hi + 1) // 2
    dt = o.sym_dtype(c)
    sb = np.dtype(dt).itemsize
    code = np.array([1, 0, 1, 1, 1, 0, 1], dtype=np.uint8)
    rng = np.random.default_rng(4)
    full = rng.integers(lo, hi + 1, (n_frames, T * R)).astype(dt)
    keep = np.resize(code.astype(bool), half * R)        # the code index restarts in every call
    tx = [np.ascontiguousarray(full[:, h * half * R:(h + 1) * half * R][:, keep]) for h in (0, 1)]
    erased = full.copy()
    for h in (0, 1):
        part = erased[:, h * half * R:(h + 1) * half * R]
        part[:, ~keep] = fill
    with kv.BatchDecoder(codec, T, n_frames, K=K, R=R) as dec:
        dec.reset()
        dec.set_variant(2)
        assert dec.set_puncturing([(code, half * R)], fill) == tx[0].shape[1]
        for h in (0, 1):
            d_tx = torch.from_numpy(tx[h].view(np.uint8).reshape(n_frames, -1).copy()).cuda()
            dec.update_punctured_dev(d_tx, tx[h].shape[1] * sb)
            dec.sync()
        assert dec.steps_done == T
        got = dec.chainback(nbits)
        mets = [dec.metrics(f) for f in range(n_frames)]
        dec.reset()
        dec.update(erased)
        assert np.array_equal(dec.chainback(nbits), got)
    for f in (0, n_frames - 1):
        q = o.Oracle(c, T)
        qo, _ = q.decode(erased[f], nbits)
        assert np.array_equal(got[f], qo), f
        assert np.array_equal(mets[f], q.metrics()), f
        q.close()
