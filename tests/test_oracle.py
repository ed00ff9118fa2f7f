"""CPU tests of the checker itself: oracle/viterbi_oracle.c against (a) the compiled reference
(oracle/_ref, only where it was built) and (b) the committed golden vectors generated from it."""
import hashlib
import os

import numpy as np
import pytest

import _oracle as o

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "reference_vectors.npz")
ALL_NAMES = [
    "ka9q27", "ka9q29", "ka9q615", "ka9q224", "spiral27", "spiral29", "spiral47", "spiral49", "spiral615",
] + [f"wy_{simd}_{w}_{kr}" for kr in ("k7r2", "k7r3", "k7r4", "k9r2", "k9r4", "k15r6", "k24r2")
     for simd in ("sse", "avx") for w in ("u8", "u16")] + ["wy_sse_u16_k5r2"]


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLDEN)


def extreme_patterns(c, T):
    """Symbol sequences that saturate, tie on every butterfly and renormalise often, plus hard-decision adversarial ones
    (runs and alternations of the two extreme values at several periods)."""
    lo, hi = o.soft_range(c)
    n = T * c.R
    k = np.arange(n)
    pats = [np.full(n, lo), np.full(n, hi), np.where(k % 2, lo, hi), np.full(n, (lo + hi) // 2),
            np.where((k // c.R) % 3 == 0, hi, lo),
            np.where((k // c.R) % 2, hi, lo), np.where((k // (c.R * (c.K - 1))) % 2, hi, lo),
            np.where((k * 7 // 5) % 3 == 1, hi, lo), np.where((k // c.R) % (c.K + 1) < 2, lo, (lo + hi + 1) // 2)]
    return [p.astype(o.sym_dtype(c)) for p in pats]


def sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), dtype=np.uint8)


@pytest.mark.parametrize("name", ALL_NAMES)
def test_oracle_matches_golden(golden, name):
    """Known-answer test: decisions, metrics, decoded bytes and return code of the reference."""
    c = o.codec_by_name(name)
    for ci in range(4):
        key = f"{name}/{ci}"
        if key + "/syms" not in golden:
            assert ci == 3 and c.K != 24      # the renormalising fourth case exists for K=24 only
            continue
        syms = golden[key + "/syms"]
        T = syms.size // c.R
        nbits = T - (c.K - 1)
        q = o.Oracle(c, T)
        q.init(0)
        q.update(syms, T)
        out, rc = q.chainback(nbits, 0, return_code=True)
        dec = q.decisions(0, T)
        assert np.array_equal(out, golden[key + "/out"]), f"{key}: decoded bytes differ"
        assert rc == int(golden[key + "/rc"]), f"{key}: chainback return value differs"
        assert np.array_equal(sha(dec), golden[key + "/dec_sha256"]), f"{key}: decision vectors differ"
        assert np.array_equal(sha(q.metrics()), golden[key + "/metrics_sha256"]), f"{key}: metrics differ"
        if key + "/dec_head" in golden:
            assert np.array_equal(dec[:4], golden[key + "/dec_head"])
            assert np.array_equal(dec[-4:], golden[key + "/dec_tail"])
        if key + "/error_sum" in golden:      # decoder_t::update<uint64_t>()'s return value
            assert q.renorm_sum() == int(golden[key + "/error_sum"]), f"{key}: sum of renormalisation minima differs"
            if ci == 3:
                assert q.renorm_sum() > 0, f"{key}: the K=24 renormalisation case did not renormalise"
        q.close()


@pytest.mark.skipif(not o.ref_available(), reason="oracle/_ref/libkvref.so not built (needs /root/reference)")
def test_golden_recipe_reproduces_the_committed_fixtures(golden):
    """tests/golden/make_golden.py must run at HEAD and give exactly the committed .npz (every array)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(os.path.dirname(GOLDEN), "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    store = mg.generate()
    assert sorted(store) == sorted(golden.files)
    assert sorted(store["names"].tolist()) == sorted(ALL_NAMES)
    for k in golden.files:
        assert np.array_equal(np.asarray(store[k]), golden[k]), k


K24_PIN = [("ka9q224", "erased"), ("ka9q224", "random"), ("wy_sse_u8_k24r2", "random"), ("wy_avx_u8_k24r2", "erased"),
           ("wy_sse_u16_k24r2", "erased"), ("wy_avx_u16_k24r2", "random")]


@pytest.mark.skipif(not o.ref_available(), reason="oracle/_ref/libkvref.so not built (needs /root/reference)")
@pytest.mark.parametrize("name,kind", K24_PIN)
def test_oracle_k24_renormalisation_matches_compiled_reference(name, kind):
    """viterbi224_sse2.cpp:226-246 and the williamyang renormalise (sse_u8.h:127-154, sse_u16.h:152-177) at K=24 need
    hundreds of steps before new[0] reaches the threshold: 263-step frames of erased symbols (every path costs the
    same) and of uniformly random symbols, split over two update calls, against the compiled reference -- decision
    vectors of every step, final metrics, decoded bytes and (williamyang) the sum of the subtracted minima."""
    c = o.codec_by_name(name)
    lo, hi = o.soft_range(c)
    T = 263
    rng = np.random.default_rng(224)
    syms = (np.full(T * c.R, (lo + hi + 1) // 2) if kind == "erased" else rng.integers(lo, hi + 1, T * c.R)).astype(o.sym_dtype(c))
    r, q = o.Ref(name, T), o.Oracle(c, T)
    for d in (r, q):
        d.init(0)
        d.update(syms[:211 * c.R], 211)
        d.update(syms[211 * c.R:], T - 211)
    assert np.array_equal(r.decisions(0, T), q.decisions(0, T)), "decision vectors differ"
    assert np.array_equal(r.metrics(), q.metrics()), "metrics differ"
    nbits = T - (c.K - 1)
    assert np.array_equal(r.chainback(nbits), q.chainback(nbits))
    if name.startswith("wy_"):
        assert r.renorm_sum() == q.renorm_sum()
    if kind == "erased":        # must have renormalised: otherwise new[0] would have passed the threshold long ago
        assert q.renorm_sum() > 0
        thr = c.renorm_threshold
        assert int(q.metrics().astype(np.int16 if c.policy == 2 else np.int64)[0]) < thr
    r.close(); q.close()


def test_golden_noiseless_round_trip(golden):
    """The only thing the reference's own tests pin (run_tests.cpp:184-186): clean channel -> 0 errors.
    ka9q224 is the documented exception (SURVEY.md 4 defect 2)."""
    for name in ALL_NAMES:
        if name == "ka9q224":
            continue
        key = f"{name}/0"
        assert np.array_equal(golden[key + "/out"], golden[key + "/data"]), name


@pytest.mark.skipif(not o.ref_available(), reason="oracle/_ref/libkvref.so not built (needs /root/reference)")
@pytest.mark.parametrize("name", [n for n in ALL_NAMES if "k24" not in n and n != "ka9q224"])
@pytest.mark.parametrize("sigma", [0.0, 0.5, 0.9, 1.3])
def test_oracle_matches_compiled_reference(name, sigma):
    c = o.codec_by_name(name)
    nbytes = 16 if c.K == 15 else 64
    T = nbytes * 8 + c.K - 1
    for frame in range(2):
        data, syms = o.make_frame(c, nbytes, sigma, seed=0xABCD, frame=frame)
        r, q = o.Ref(name, T), o.Oracle(c, T)
        ro, rd = r.decode(syms, nbytes * 8)
        qo, qd = q.decode(syms, nbytes * 8)
        assert np.array_equal(rd, qd), "decision vectors differ"
        assert np.array_equal(r.metrics(), q.metrics()), "metrics differ"
        assert np.array_equal(ro, qo), "decoded bytes differ"
        if name.startswith("wy_"):
            assert r.renorm_sum() == q.renorm_sum(), "sum of renormalisation minima (update()'s return value) differs"
        r.close(); q.close()


@pytest.mark.skipif(not o.ref_available(), reason="oracle/_ref/libkvref.so not built")
@pytest.mark.parametrize("name", ["ka9q27", "ka9q29", "wy_sse_u8_k7r2", "wy_avx_u16_k9r4", "ka9q615"])
def test_oracle_streaming_updates_match_reference(name):
    """ka9q / williamyang update() continues across calls (viterbi27_sse2.cpp:121,174; sse_u8.h:59,67)."""
    c = o.codec_by_name(name)
    nbytes = 8 if c.K == 15 else 40
    T = nbytes * 8 + c.K - 1
    data, syms = o.make_frame(c, nbytes, 0.8, seed=7, frame=1)
    r, q = o.Ref(name, T), o.Oracle(c, T)
    for d in (r, q):
        d.init(0)
        pos = 0
        for chunk in (1, 2, 5, 17, T):
            n = min(chunk, T - pos)
            d.update(syms[pos * c.R:(pos + n) * c.R], n)
            pos += n
        assert pos == T
    assert np.array_equal(r.decisions(0, T), q.decisions(0, T))
    assert np.array_equal(r.chainback(nbytes * 8), q.chainback(nbytes * 8))
    r.close(); q.close()


@pytest.mark.skipif(not o.ref_available(), reason="oracle/_ref/libkvref.so not built")
def test_oracle_extreme_inputs_match_reference():
    """All-zero / all-max / alternating symbols: saturation, ties on every butterfly, renormalisation."""
    for name in ["ka9q27", "ka9q29", "spiral27", "spiral47", "spiral29", "spiral49", "wy_sse_u8_k7r2", "wy_sse_u8_k9r2",
                 "wy_sse_u16_k7r4", "spiral615", "ka9q615", "wy_avx_u8_k15r6"]:
        c = o.codec_by_name(name)
        T = 200 if c.K == 15 else 600
        for p in extreme_patterns(c, T):
            syms = p.astype(o.sym_dtype(c))
            r, q = o.Ref(name, T), o.Oracle(c, T)
            nbits = T - (c.K - 1)
            ro, rd = r.decode(syms, nbits)
            qo, qd = q.decode(syms, nbits)
            assert np.array_equal(rd, qd), name
            assert np.array_equal(r.metrics(), q.metrics()), name
            assert np.array_equal(ro, qo), name
            r.close(); q.close()


def test_modular_compare_bound_k7_k9():
    """ka9q27 / ka9q29 decide with (int8)(m0 - m1) > 0 on wrapping u8 metrics (viterbi27_sse2.cpp:155-156). The CUDA
    kernels keep unwrapped metrics and compare plainly: the same decision iff |m0 - m1| <= 127 at every butterfly.
    Proof (csrc/acs_k7.cu, csrc/acs_grp.cu headers): the two candidates of a butterfly are the best paths into the new
    state through old states j and j + S/2; from a common state K-1 steps back those paths take the same inputs but
    the first, so their codewords differ at step m exactly by the taps of register position m-1: both taps -> the
    branch metrics differ by <= 15, one tap -> <= 8, none -> 0. Summed over the K positions of main.cpp's polynomials:
    76 (K=7), 92 (K=9) once K-1 steps have run; before that every metric lies in [0, 63 + 7t], so the difference is at
    most 78 + 7t <= 113 (K=7), 127 (K=9). This test tracks the unwrapped recurrence on random, hard-decision,
    erasure-mixed and the extreme patterns and checks both bounds."""
    for name, steady, transient in (("ka9q27", 76, 113), ("ka9q29", 92, 127)):
        c = o.codec_by_name(name)
        S, H = 1 << (c.K - 1), 1 << (c.K - 2)
        # the analytic steady-state bound from the tap patterns of the two polynomials
        taps = [((c.poly[0] >> p) & 1) + ((c.poly[1] >> p) & 1) for p in range(c.K)]
        assert sum({2: 15, 1: 8, 0: 0}[x] for x in taps) == steady
        cw = np.array([sum((bin((2 * j) & c.poly[i]).count("1") & 1) << i for i in range(2)) for j in range(H)])
        T = 400
        rng = np.random.default_rng(29)
        seqs = list(extreme_patterns(c, T)) + [rng.integers(0, 256, T * 2), rng.choice([0, 255], T * 2),
                                              rng.choice([0, 128, 255], T * 2), rng.choice([0, 15, 240, 255], T * 2)]
        worst_t, worst_s = 0, 0
        for syms in seqs:
            syms = np.asarray(syms, dtype=np.int64)
            for start in (0, 5):
                m = np.full(S, 63, dtype=np.int64)
                m[start] = 0
                for t in range(T):
                    s0, s1 = syms[2 * t], syms[2 * t + 1]
                    x0 = np.where(cw & 1, s0 ^ 255, s0)
                    x1 = np.where(cw & 2, s1 ^ 255, s1)
                    bm = ((x0 + x1 + 1) >> 1) >> 4
                    lo_, hi_ = m[:H], m[H:]
                    a0, a1, b0, b1 = lo_ + bm, hi_ + 15 - bm, lo_ + 15 - bm, hi_ + bm
                    d = max(int(np.abs(a0 - a1).max()), int(np.abs(b0 - b1).max()))
                    if t < c.K - 1:
                        worst_t = max(worst_t, d)
                        assert d <= 78 + 7 * t
                    else:
                        worst_s = max(worst_s, d)
                    n = np.empty(S, dtype=np.int64)
                    n[0::2] = np.minimum(a0, a1)
                    n[1::2] = np.minimum(b0, b1)
                    m = n
                    assert int(m.max() - m.min()) < 128        # what the kernels' unwrap around state 0 needs
        assert worst_t <= transient and worst_s <= steady, (name, worst_t, worst_s)


def test_encoder_matches_definition():
    """Symbol i of a step = parity(G[i] & reg), reg = (reg << 1) | bit, MSB-first, K-1 zero tail bits
    (convolutional_encoder_shift_register.h:52-60)."""
    c = o.codec_by_name("ka9q27")
    data = np.array([0x80, 0x01], dtype=np.uint8)
    bits = o.encode_bits(c, data)
    assert bits.size == (16 + 6) * 2
    reg, exp = 0, []
    stream = [int(b) for b in np.unpackbits(data)] + [0] * 6
    for b in stream:
        reg = (reg << 1) | b
        for g in (0x6d, 0x4f):
            exp.append(bin(reg & g).count("1") & 1)
    assert list(bits) == exp


def test_channel_noiseless_levels():
    for name, lo, hi in (("ka9q27", 0, 255), ("wy_sse_u8_k7r2", -3, 3), ("wy_sse_u16_k7r2", -127, 127)):
        c = o.codec_by_name(name)
        data, syms = o.make_frame(c, 8, 0.0)
        assert set(np.unique(syms).tolist()) == {lo, hi}
        bits = o.encode_bits(c, data)
        assert np.array_equal(syms == hi, bits == 1)


def _hard8(c):
    """src/viterbi_configs.h:52-65 (hard8): symbols +-1, max error 2R, margin 3x."""
    c.soft_hi, c.soft_lo = 1, -1
    c.max_error = 2 * c.R
    c.init_start, c.init_other = 0, 3 * c.max_error
    c.renorm_threshold = 255 - 3 * c.max_error
    return c


@pytest.mark.parametrize("name", ["wy_sse_u8_k7r2", "wy_avx_u8_k9r4", "wy_sse_u8_k15r6"])
def test_oracle_matches_reference_with_the_hard8_config(name):
    """The williamyang decoders take a numeric config; pin the restatement on main.cpp's other config too."""
    if not o.ref_available():
        pytest.skip("oracle/_ref not built")
    c = _hard8(o.codec_by_name(name))
    nbytes = 6 if c.K == 15 else 40
    T = nbytes * 8 + c.K - 1
    rng = np.random.default_rng(8)
    o.ref_lib().kvref_wy_set_config_kind(1)
    try:
        r = o.Ref(name, T)
    finally:
        o.ref_lib().kvref_wy_set_config_kind(0)
    q = o.Oracle(c, T)
    for trial in range(3):
        syms = rng.integers(-1, 2, T * c.R).astype(np.int8) if trial else np.where(rng.random(T * c.R) < 0.5, 1, -1).astype(np.int8)
        ro, rd = r.decode(syms, nbytes * 8)
        qo, qd = q.decode(syms, nbytes * 8)
        assert np.array_equal(rd, qd) and np.array_equal(ro, qo)
        assert np.array_equal(r.metrics(), q.metrics())
    r.close(); q.close()


# common_codes.h entries that main.cpp does not benchmark: other polynomials (DAB, IS-95A) and the smallest K a
# SIMD decoder accepts (K=5, SSE u16 only). The williamyang decoders build their branch table per instance, so the compiled
# reference can be asked for them directly.
COMMON_EXTRA = [("k5r2", 5, 2, [0b10111, 0b11001], ("wy_sse_u16",)),
                ("k7r4", 7, 4, [109, 79, 83, 109], ("wy_sse_u8", "wy_avx_u8", "wy_sse_u16", "wy_avx_u16")),
                ("k9r2", 9, 2, [491, 369], ("wy_sse_u8", "wy_avx_u8", "wy_sse_u16", "wy_avx_u16"))]


@pytest.mark.skipif(not o.ref_available(), reason="oracle/_ref/libkvref.so not built")
@pytest.mark.parametrize("kr,K,R,poly,decoders", COMMON_EXTRA, ids=[c[0] for c in COMMON_EXTRA])
def test_oracle_matches_reference_on_common_codes(kr, K, R, poly, decoders):
    nbytes = 64
    T = nbytes * 8 + K - 1
    for prefix in decoders:
        c = o.codec_wy(8 if prefix.endswith("u8") else 16, K, R, poly)
        for sigma in (0.0, 0.7, 1.3):
            data, syms = o.make_frame(c, nbytes, sigma, seed=0xC0DE, frame=3)
            r, q = o.Ref(f"{prefix}_{kr}", T, codec=c), o.Oracle(c, T)
            ro, rd = r.decode(syms, nbytes * 8)
            qo, qd = q.decode(syms, nbytes * 8)
            assert np.array_equal(rd, qd), (prefix, sigma, "decision vectors differ")
            assert np.array_equal(r.metrics(), q.metrics()), (prefix, sigma, "metrics differ")
            assert np.array_equal(ro, qo), (prefix, sigma, "decoded bytes differ")
            if sigma == 0.0:
                assert np.array_equal(ro, data)
            r.close(); q.close()


@pytest.mark.parametrize("name", ["ka9q27", "spiral49", "wy_sse_u16_k9r2", "ka9q615"])
def test_window_walk_of_the_stream_tests_equals_the_oracle_chainback(name):
    """tests/test_stream.py checks the sliding-window decoder against `walk` (a traceback through the oracle's canonical
    decision vectors from a given state). Pin that helper: over a whole terminated frame, from end state 0, it must
    reproduce the oracle's own chainback (which is pinned against the compiled reference above) -- with and without the
    K-1 newest bits read from the state -- and a window in the middle must equal the same bits of the full traceback."""
    import importlib
    ts = importlib.import_module("test_stream")
    c = o.codec_by_name(name)
    K = c.K
    nbytes = 12 if K == 15 else 40
    nbits = nbytes * 8
    T = nbits + K - 1
    data, syms = o.make_frame(c, nbytes, 0.4, seed=9, frame=1)
    q = o.Oracle(c, T)
    want, rows = q.decode(syms, nbits)
    q.close()
    got = ts.walk(rows, K, 0, T - 1, 0, nbits, False)
    assert np.array_equal(got, want)
    # the newest K-1 bits of a walk that starts at state s are s itself: here the tail (zeros)
    full = ts.walk(rows, K, 0, T - 1, 0, T, True)
    assert np.array_equal(np.unpackbits(full)[:nbits], np.unpackbits(want)[:nbits])
    assert not np.unpackbits(full)[nbits:T].any()
    # a window: bits [16, 48) traced from the true state after row 79 (the K-1 newest input bits up to step 79)
    bits = np.unpackbits(want)
    top = 79
    state = 0
    for j in range(K - 1):
        state |= int(bits[top - j]) << j
    win = ts.walk(rows, K, state, top, 16, 32, False)
    assert np.array_equal(np.unpackbits(win)[:32], bits[16:48])
