"""Sliding-window decode of continuous streams (b200v_stream_*) against the same window run on the oracle:
the oracle is updated chunk by chunk (the reference's resumable update: viterbi27_sse2.cpp:121,174; sse_u8.h:59,67), its
metrics give every stream's best state after each push, its canonical decision vectors are walked back `depth + chunk`
rows, and every emitted byte must match. At low noise the emitted bits are the transmitted ones."""
import numpy as np
import pytest

import _oracle as o
import ka9q_viterbi_comparison_b200 as kv

pytestmark = pytest.mark.gpu

CASES = {
    "ka9q27": (kv.KA9Q27, 7, 2), "spiral27": (kv.SPIRAL27, 7, 2), "wy_sse_u8_k7r4": (kv.WY_U8, 7, 4),
    "wy_sse_u16_k7r2": (kv.WY_U16, 7, 2), "ka9q29": (kv.KA9Q29, 9, 2), "spiral49": (kv.SPIRAL49, 9, 4),
    "ka9q615": (kv.KA9Q615, 15, 6), "spiral615": (kv.SPIRAL615, 15, 6),
}


def best_state(name, m):
    m = m.astype(np.int64)
    if name in ("ka9q27", "ka9q29"):                       # modular metrics: int8 distance from state 0
        m = ((m - m[0] + 128) % 256) - 128
    return int(np.argmin(m))


def walk(dec_rows, K, state, top_row, first_bit, nbits, state_bits):
    """dec_rows[r] = canonical decision vector of logical row r. Returns nbits bits [first_bit, first_bit + nbits)."""
    bits = {}
    b = top_row + 1
    if state_bits:
        for j in range(K - 1):
            b -= 1
            bits[b] = (state >> j) & 1
    else:
        b = top_row + 1 - (K - 1)
    r = top_row
    while b > first_bit:
        b -= 1
        k = (int(dec_rows[r][state >> 3]) >> (state & 7)) & 1
        state = (state >> 1) | (k << (K - 2))
        bits[b] = k
        r -= 1
    out = np.zeros((nbits + 7) // 8, dtype=np.uint8)
    for i in range(nbits):
        out[i >> 3] |= bits[first_bit + i] << (7 - (i & 7))
    return out


@pytest.mark.parametrize("terminated", [True, False], ids=["tail", "unterminated"])
@pytest.mark.parametrize("name", list(CASES))
def test_stream_window_matches_the_oracle_window(name, terminated):
    import torch
    codec, K, R = CASES[name]
    c = o.codec_by_name(name)
    chunk = 16
    depth = 48 if K <= 9 else 96                 # >= 5 (K-1), a multiple of the chunk
    W = depth + chunk
    n_streams = 3 if K == 15 else 5
    npush = (10 * W) // chunk + 3                # the stream is more than ten ring lengths long
    T = npush * chunk
    nbytes = (T - (K - 1)) // 8 + 1              # enough data bits to cover every step; cut to T steps below
    sigma = 0.35
    frames = [o.make_frame(c, nbytes, sigma, seed=0x57E, frame=f) for f in range(n_streams)]
    data = np.stack([d for d, _ in frames])
    syms = np.stack([s[:T * R] for _, s in frames])
    sb = np.dtype(o.sym_dtype(c)).itemsize
    # oracle: chunked update, metrics after every push, all decision rows
    oracles = []
    for f in range(n_streams):
        q = o.Oracle(c, T)
        q.init(0)
        oracles.append(q)
    emitted = [[] for _ in range(n_streams)]
    hist = [[] for _ in range(n_streams)]        # canonical decision rows of every step so far, per stream
    spiral = name.startswith("spiral")           # spiral's update rewrites its rows from index 0 on every call
    with kv.StreamDecoder(codec, n_streams, depth, chunk, K=K, R=R) as sd:
        sd.reset(0)
        d_syms = torch.from_numpy(syms.view(np.uint8).reshape(n_streams, -1).copy()).cuda()
        d_out = torch.zeros(n_streams, W, dtype=torch.uint8, device="cuda")
        row = T * R * sb
        for p in range(npush):
            chunk_ptr = d_syms.data_ptr() + p * chunk * R * sb
            n = sd.push_dev(chunk_ptr, row, d_out, W)
            sd.sync()
            steps = (p + 1) * chunk
            assert n == (chunk if steps >= W else 0)
            got = d_out.cpu().numpy()
            for f in range(n_streams):
                q = oracles[f]
                q.update(syms[f, p * chunk * R:(p + 1) * chunk * R], chunk)
                assert np.array_equal(sd.metrics(f), q.metrics()), (p, f)
                hist[f].extend(q.decisions(0 if spiral else p * chunk, chunk))
                if n:
                    rows = hist[f]
                    st = best_state(name, q.metrics())
                    w = walk(rows, K, st, steps - 1, steps - W, chunk, False)
                    assert np.array_equal(got[f, :chunk // 8], w), (p, f)
                    emitted[f].append(got[f, :chunk // 8].copy())
        assert sd.steps == T and sd.emitted == T - depth
        # end of the stream
        n = sd.flush_dev(d_out, W, endstate=0 if terminated else -1)
        sd.sync()
        got = d_out.cpu().numpy()
        assert n == (depth - (K - 1) if terminated else depth)
        for f in range(n_streams):
            q = oracles[f]
            rows = hist[f]
            st = 0 if terminated else best_state(name, q.metrics())
            w = walk(rows, K, st, T - 1, T - depth, n, not terminated)
            assert np.array_equal(got[f, :(n + 7) // 8], w), f
            emitted[f].append(got[f, :(n + 7) // 8].copy())
            q.close()
        assert sd.flush_dev(d_out, W, endstate=0 if terminated else -1) == 0      # nothing left
    # low noise: the window decoder recovers the transmitted data (all but the freshest, still uncertain bits)
    for f in range(n_streams):
        stream = np.concatenate(emitted[f])
        nfull = (T - depth) // 8
        assert np.array_equal(stream[:nfull], data[f, :nfull]), f


def test_stream_many_streams_register_kernel():
    """4608 streams: the K=7 register kernel (decision words interleaved across streams) writes the ring; the emitted
    bits equal a whole-frame decode of the same symbols wherever the window is deep enough -- at low noise everywhere."""
    import torch
    codec, K, R = kv.KA9Q27, 7, 2
    c = o.codec_by_name("ka9q27")
    n_streams, chunk, depth = 4608, 64, 64
    W = depth + chunk
    nbytes = 126
    nbits = nbytes * 8
    T = nbits + K - 1 + 2                          # 1016 steps: not a multiple of the chunk -> feed 15 chunks (960 steps)
    npush = 15
    amp_q8, sigma_q8, mean_q8, lo, hi = o.channel_params(c, 0.3)
    with kv.BatchDecoder(codec, T, n_streams) as gen, kv.StreamDecoder(codec, n_streams, depth, chunk) as sd:
        d_syms = torch.zeros(n_streams * T * R, dtype=torch.uint8, device="cuda")
        d_data = torch.zeros(n_streams * nbytes, dtype=torch.uint8, device="cuda")
        gen.generate_dev(d_syms, T * R, d_data, nbytes, nbits, 99, 0, amp_q8, sigma_q8, mean_q8, lo, hi)
        gen.sync()
        assert gen.variant == 2
        data = d_data.cpu().numpy().reshape(n_streams, nbytes)
        sd.reset(0)
        d_out = torch.zeros(n_streams, chunk // 8, dtype=torch.uint8, device="cuda")
        got = []
        for p in range(npush):
            n = sd.push_dev(d_syms.data_ptr() + p * chunk * R, T * R, d_out, chunk // 8)
            sd.sync()
            if n:
                got.append(d_out.cpu().numpy().copy())
        stream = np.concatenate(got, axis=1)
        assert stream.shape[1] == (npush * chunk - depth) // 8
        assert np.array_equal(stream, data[:, :stream.shape[1]])


def test_stream_argument_errors():
    with pytest.raises(kv.B200ViterbiError):
        kv.StreamDecoder(kv.KA9Q27, 4, 48, 20)          # chunk not a multiple of 8
    with pytest.raises(kv.B200ViterbiError):
        kv.StreamDecoder(kv.KA9Q27, 4, 40, 16)          # depth not a multiple of the chunk
    with pytest.raises(kv.B200ViterbiError):
        kv.StreamDecoder(kv.KA9Q615, 4, 8, 8)           # depth below K-1
    with pytest.raises(kv.B200ViterbiError):
        kv.StreamDecoder(kv.KA9Q224, 1, 64, 32)         # K=24 has no stream decoder


def test_stream_host_buffer_entries_match_the_device_entries():
    """b200v_stream_push / _flush (host buffers, synchronous) against push_dev / flush_dev on the same symbols."""
    import torch
    codec, K, R = kv.SPIRAL27, 7, 2
    c = o.codec_by_name("spiral27")
    n_streams, chunk, depth, npush = 7, 32, 64, 9
    T = chunk * npush
    syms = np.stack([o.make_frame(c, T // 8 + 1, 0.5, seed=3, frame=f)[1][:T * R] for f in range(n_streams)])
    with kv.StreamDecoder(codec, n_streams, depth, chunk) as a, kv.StreamDecoder(codec, n_streams, depth, chunk) as b:
        a.reset(0); b.reset(0)
        d_syms = torch.from_numpy(syms.copy()).cuda()
        d_out = torch.zeros(n_streams, (depth + chunk) // 8 + 1, dtype=torch.uint8, device="cuda")
        for p in range(npush):
            got = a.push(syms[:, p * chunk * R:(p + 1) * chunk * R])
            n = b.push_dev(d_syms.data_ptr() + p * chunk * R, T * R, d_out, d_out.shape[1])
            b.sync()
            assert (got is None) == (n == 0)
            if n:
                assert np.array_equal(got, d_out.cpu().numpy()[:, :chunk // 8]), p
        tail, nbits = a.flush()
        n = b.flush_dev(d_out, d_out.shape[1])
        b.sync()
        assert nbits == n == depth
        assert np.array_equal(tail, d_out.cpu().numpy()[:, :(n + 7) // 8])


@pytest.mark.parametrize("name", ["ka9q27", "spiral49", "ka9q615"])
def test_short_stream_flushed_before_the_window_fills_equals_a_frame_decode(name):
    """A tail-terminated stream shorter than the ring: no push emits anything, flush(endstate 0) emits every data bit and
    must equal the ordinary init + update + chainback of the same symbols (and the data at low noise); reset and reuse."""
    import torch
    codec, K, R = CASES[name]
    c = o.codec_by_name(name)
    chunk, depth, n_streams = 16, 96, 4
    npush = 5                                          # 80 steps < depth + chunk = 112
    T = npush * chunk
    nbits = T - (K - 1)
    nbytes = (nbits + 7) // 8
    for seed in (1, 2):                                # the second pass reuses the stream object after reset
        frames = [o.make_frame(c, nbits // 8, 0.3, seed=seed, frame=f) for f in range(n_streams)]
        # make_frame gives (nbits//8)*8 + K-1 steps: pad with "strong zero" symbols up to T steps (more tail)
        lo, hi = o.soft_range(c)
        zero_sym = lo
        syms = np.full((n_streams, T * R), zero_sym, dtype=o.sym_dtype(c))
        for f, (_, s) in enumerate(frames):
            syms[f, :s.size] = s
        if seed == 1:
            sd = kv.StreamDecoder(codec, n_streams, depth, chunk, K=K, R=R)
        sd.reset(0)
        for p in range(npush):
            assert sd.push(syms[:, p * chunk * R:(p + 1) * chunk * R]) is None
        tail, n = sd.flush(endstate=0)
        assert n == nbits and sd.emitted == nbits
        with kv.BatchDecoder(codec, T, n_streams, K=K, R=R) as dec:
            dec.reset()
            dec.update(syms)
            want = dec.chainback(nbits)
        assert np.array_equal(tail[:, :nbytes], want[:, :nbytes])
        for f, (d, _) in enumerate(frames):
            assert np.array_equal(tail[f, :d.size], d), f
    sd.close()
