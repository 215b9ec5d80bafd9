"""Pins oracle/cascaded_oracle.c against the reference's own known answers.

Mirrors tests/test_cascaded_batch.cpp of the reference:
  test_predefined_cases (:212-481)  -> test_kat_*
  fallback path (:491-715)          -> test_fallback_random
  out-of-bounds matrix (:717-916)   -> test_oob_matrix
  invalid size query (:950-990)     -> test_size_query_short
"""
import numpy as np
import pytest

import oracle_lib as o

DTYPES = {0: np.int8, 1: np.uint8, 2: np.int16, 3: np.uint16, 4: np.int32, 5: np.uint32, 6: np.int64, 7: np.uint64}


def up(x, a):
    return (x + a - 1) // a * a


def predefined(vals, counts, dt):
    return np.repeat(np.array(vals, dtype=np.int64), counts).astype(dt)


KAT = [
    # input values, run lengths, runs0, runs1, final array, delta seed (test_cascaded_batch.cpp:217-223,336-378)
    ([3, 9, 4, 0, 1], [1, 20, 13, 25, 6], [1, 20, 13, 25, 6], [1, 1, 1, 1], [6, -5, -4, 1], 3),
    ([1, 2, 3, 4, 5, 6], [10, 6, 15, 1, 13, 9], [10, 6, 15, 1, 13, 9], [5], [1], 1),
]


@pytest.mark.parametrize("type_id", sorted(DTYPES))
@pytest.mark.parametrize("case", range(len(KAT)))
def test_kat_layout_no_bitpacking(type_id, case):
    dt = DTYPES[type_id]
    w = np.dtype(dt).itemsize
    vals, cnts, runs0, runs1, final, seed = KAT[case]
    x = predefined(vals, cnts, dt)
    comp = o.casc_compress(x, type_id, 2, 1, 0)
    assert len(comp) % 4 == 0 and len(comp) % w == 0
    words = comp.view(np.uint32)
    assert words[0] == 2 + (1 << 8) + (0 << 16) + (type_id << 24)
    assert words[1] == x.nbytes
    chunk = up(8, w)
    cw = comp[chunk:].view(np.uint32)
    assert cw[1] == len(runs0) * 2 and cw[2] == len(runs1) * 2 and cw[3] == len(runs1) * w
    pos = chunk + up(16, w)
    assert comp[pos : pos + w].view(dt)[0] == np.array(seed).astype(dt)
    pos = up(pos + w, 4)
    assert list(comp[pos : pos + 2 * len(runs0)].view(np.uint16)) == runs0
    pos = up(pos + 2 * len(runs0), 4)
    assert list(comp[pos : pos + 2 * len(runs1)].view(np.uint16)) == runs1
    pos = up(up(pos + 2 * len(runs1), 4), w)
    assert list(comp[pos : pos + w * len(final)].view(dt)) == list(np.array(final).astype(dt))
    assert cw[0] == len(comp) - chunk  # single chunk: chunk_bytes spans to the end
    st, back = o.casc_decompress(comp, x.nbytes, w)
    assert st == o.SUCCESS and back.tobytes() == x.tobytes()


@pytest.mark.parametrize("type_id", sorted(DTYPES))
@pytest.mark.parametrize("use_bp", [0, 1])
def test_kat_roundtrip(type_id, use_bp):
    dt = DTYPES[type_id]
    w = np.dtype(dt).itemsize
    for vals, cnts, *_ in KAT:
        x = predefined(vals, cnts, dt)
        comp = o.casc_compress(x, type_id, 2, 1, use_bp)
        assert len(comp) % 4 == 0 and len(comp) % w == 0
        assert o.casc_get_size(comp) == x.nbytes
        st, back = o.casc_decompress(comp, x.nbytes, w)
        assert st == o.SUCCESS and back.tobytes() == x.tobytes()


@pytest.mark.parametrize("type_id", sorted(DTYPES))
def test_fallback_random(type_id):
    dt = DTYPES[type_id]
    w = np.dtype(dt).itemsize
    rng = np.random.default_rng(7)
    x = rng.integers(0, 256, size=10000 * w, dtype=np.uint8).view(dt)
    comp = o.casc_compress(x, type_id, 2, 1, 1)
    assert comp[:4].view(np.uint32)[0] == type_id << 24  # header bytes 0..2 zero => raw copy
    assert len(comp) == up(8, w) + up(x.nbytes, 4)
    st, back = o.casc_decompress(comp, x.nbytes, w)
    assert st == o.SUCCESS and back.tobytes() == x.tobytes()


@pytest.mark.parametrize("type_id", [1, 3, 4, 7])
def test_oob_matrix(type_id):
    dt = DTYPES[type_id]
    w = np.dtype(dt).itemsize
    x = np.repeat(np.arange(300), 11).astype(dt)
    comp = o.casc_compress(x, type_id, 2, 1, 1)
    cb, ub = len(comp), x.nbytes
    for comp_bytes, cap, expect in [
        (cb // 2, ub, o.CANNOT_DECOMPRESS),
        (cb, ub // 2, o.CANNOT_DECOMPRESS),
        (cb, ub - 1, o.CANNOT_DECOMPRESS),
        (cb, ub, o.SUCCESS),
    ]:
        st, back = o.casc_decompress(comp[:comp_bytes], cap, w)
        assert st == expect
        assert len(back) == (ub if expect == o.SUCCESS else 0)


def test_size_query_short():
    assert o.casc_get_size(np.zeros(4, np.uint8)) == 0


@pytest.mark.parametrize("opts", [(0, 0, 1), (1, 0, 0), (0, 1, 0), (1, 1, 1), (2, 2, 1), (3, 1, 0), (1, 3, 1), (7, 0, 1)])
@pytest.mark.parametrize("type_id", [0, 3, 4, 6])
def test_option_matrix_roundtrip(opts, type_id):
    dt = DTYPES[type_id]
    w = np.dtype(dt).itemsize
    rng = np.random.default_rng(hash(opts) & 0xFFFF)
    steps = np.where(rng.random(9000) < 0.7, 0, rng.geometric(0.05, 9000))
    x = np.cumsum(steps).astype(dt)
    for n in (0, 1, 2, 5, 1000, 4096 // w, 4096 // w + 1, 9000):
        comp = o.casc_compress(x[:n], type_id, *opts)
        if n == 0:
            assert len(comp) == 0
            continue
        st, back = o.casc_decompress(comp, n * w, w)
        assert st == o.SUCCESS and back.tobytes() == x[:n].tobytes(), (n, opts)


def test_all_equal_chunk_has_empty_final_array():
    # one run, one delta -> final array of 0 elements (SURVEY A.3 caveat 1): FOR / width written as 0
    x = np.zeros(1024, np.int32)
    comp = o.casc_compress(x, 4, 1, 1, 1)
    cw = comp[8:].view(np.uint32)
    assert cw[2] == 8  # final array = bit-pack header only
    final = comp[8 + cw[0] - 8 :]
    assert final.view(np.uint32).tolist() == [0, 0]
    st, back = o.casc_decompress(comp, x.nbytes, 4)
    assert st == o.SUCCESS and not back.any()


def test_truncation_never_succeeds_wrongly():
    x = np.repeat(np.arange(5000, dtype=np.int32), 3)
    comp = o.casc_compress(x, 4, 1, 1, 1)
    for cut in range(0, len(comp), 97):
        st, back = o.casc_decompress(comp[:cut], x.nbytes, 4)
        assert st == o.CANNOT_DECOMPRESS and len(back) == 0


# ---- the GPU encoder's look-ahead for incompressible partitions (csrc/cascaded.cu: distinct_chunk_bytes,
# partition_cannot_fit) restated in numpy and held against the oracle: for a chunk of all-distinct elements the
# formula must give exactly the bytes the format spends on that chunk, and the verdict "cannot fit" must agree
# with the oracle falling back to a raw partition.
def distinct_chunk_bytes_model(x, R, use_bp):
    """x: one full chunk of 1024 uint32.  Bytes of the chunk (metadata + arrays), or 0 when the chunk is not
    all distinct (R = 2: or two neighbouring deltas are equal)."""
    n = 1024
    assert len(x) == n
    d = (x[1:].astype(np.int64) - x[:-1].astype(np.int64)).astype(np.uint32).view(np.int32).astype(np.int64)
    if (d == 0).any():
        return 0
    if R == 2 and (d[1:] == d[:-1]).any():
        return 0
    if not use_bp:
        return 12 + 4 * R + 2 * n + (up(2 * (n - 1), 4) if R == 2 else 0) + 4 * (n - 1)
    span = int(d.max() - d.min()) & 0xFFFFFFFF
    vbw = span.bit_length()
    return 12 + 4 * R + 8 * R + 8 + 4 * (((n - 1) * vbw + 31) >> 5)


def partition_cannot_fit_model(x, R, use_bp):
    """x: a partition of uint32 (>= 2 full chunks) whose first chunk is all distinct; the verdict of the look-ahead
    that runs behind the (encoded) first chunk."""
    n = 1024
    limit = 8 + up(4 * len(x), 4)
    meta = 12 + 4 * R
    floor_bytes = meta + (8 * (R + 1) if use_bp else 4)
    first = distinct_chunk_bytes_model(x[:n], R, use_bp)
    assert first, "the model is called for partitions whose first chunk the encoder found all distinct and exact"
    total = 8 + first + (meta if len(x) % n else 0)
    for j in range(1, len(x) // n):
        exact = distinct_chunk_bytes_model(x[j * n:(j + 1) * n], R, use_bp)
        if j == 1 and exact == 0:
            return False
        total += exact if exact else floor_bytes
        if total > limit:
            return True
    return total > limit


@pytest.mark.parametrize("opts", [(1, 1, 1), (2, 1, 1), (1, 1, 0), (2, 1, 0)])
def test_lookahead_chunk_size_formula_equals_the_oracle(opts):
    R, D, bp = opts
    rng = np.random.default_rng(100 + R + 2 * bp)
    for trial in range(60):
        bits = int(rng.integers(11, 33))
        if bits <= 16:
            x = rng.permutation(1 << bits)[:1024].astype(np.uint32)
        else:  # distinct with overwhelming probability; the model returns 0 (skipped) otherwise
            x = rng.integers(0, 1 << bits, 1024, dtype=np.uint64).astype(np.uint32)
        want = distinct_chunk_bytes_model(x, R, bp)
        if want == 0:
            continue
        # a partition of this chunk plus a constant chunk (which always fits in a few bytes): when the oracle
        # keeps the partition compressed, the first chunk's metadata word 0 is its size
        part = np.concatenate([x, np.full(1024, 7, np.uint32)])
        comp = o.casc_compress(part.view(np.uint8), 5, R, D, bp)
        w = comp.view(np.uint32)
        if (w[0] & 0xFFFFFF) == 0:  # raw: then the chunk alone must already overflow the partition's budget
            assert 8 + want + (12 + 4 * R) > 8 + 8192 or not bp
            continue
        assert w[2] == want, f"trial {trial}: chunk bytes {w[2]} vs formula {want}"


@pytest.mark.parametrize("opts", [(1, 1, 1), (2, 1, 1), (1, 1, 0), (2, 1, 0)])
def test_lookahead_verdict_agrees_with_the_oracle_fallback(opts):
    """'Cannot fit' is only ever said of partitions the oracle stores raw (the look-ahead is a lower bound), and it is
    said of every partition of full chunks of random words."""
    R, D, bp = opts
    rng = np.random.default_rng(200 + R + 2 * bp)
    said = 0
    for trial in range(40):
        chunks = int(rng.integers(2, 17))
        tail = int(rng.integers(0, 2)) * int(rng.integers(1, 1024))
        kind = trial % 4
        if kind == 0:    # random words
            x = rng.integers(0, 1 << 32, chunks * 1024 + tail, dtype=np.uint64).astype(np.uint32)
        elif kind == 1:  # distinct but narrow: fits
            x = rng.permutation(1 << 15)[: chunks * 1024 + tail].astype(np.uint32)
        elif kind == 2:  # random first chunks, compressible rest
            x = np.concatenate([rng.integers(0, 1 << 32, 2048, dtype=np.uint64).astype(np.uint32),
                                np.zeros((chunks - 2) * 1024 + tail, np.uint32)])
        else:            # wide deltas in the first chunk only
            x = np.concatenate([rng.integers(0, 1 << 32, 1024, dtype=np.uint64).astype(np.uint32),
                                rng.permutation(1 << 12)[:1024].astype(np.uint32),
                                np.arange((chunks - 2) * 1024 + tail, dtype=np.uint32)])
        if len(x) < 2048 or distinct_chunk_bytes_model(x[:1024], R, bp) == 0:
            continue
        comp = o.casc_compress(x.view(np.uint8), 5, R, D, bp)
        raw = (comp.view(np.uint32)[0] & 0xFFFFFF) == 0
        # the encoder only looks ahead when the first chunk itself fitted
        first_fits = 8 + distinct_chunk_bytes_model(x[:1024], R, bp) <= 8 + 4 * len(x)
        verdict = first_fits and partition_cannot_fit_model(x, R, bp)
        if verdict:
            said += 1
            assert raw, f"trial {trial}: look-ahead says 'cannot fit' of a partition the oracle compresses"
        if kind == 0:
            assert raw
            if tail == 0:  # (a ragged tail is only counted with its metadata: the bound may fall short)
                assert verdict or not first_fits
    assert said > 0


# ---- the GPU decoder's formulation for the default options (csrc/cascaded.cu: decompress_chunk_r2d1 +
# decompress_tail_d1) restated in numpy on the oracle's streams: both run-length layers expand by "scatter onto run
# starts, then one running sum" instead of marking and gathering.
def unpack_bp(words, elem_bits):
    """one bit-packed array of the format: frame of reference, (width << 16 | count), packed words"""
    fr, meta = int(words[0]), int(words[1])
    bw, n = meta >> 16, meta & 0xFFFF
    bits = np.unpackbits(words[2:].view(np.uint8), bitorder="little")
    vals = np.zeros(n, np.uint64)
    for b in range(bw):
        vals |= bits[b:n * bw:bw].astype(np.uint64)[:n] << np.uint64(b) if bw else 0
    return ((vals + np.uint64(fr)) & np.uint64((1 << elem_bits) - 1)).astype(np.uint64)


def decode_chunk_r2d1_model(w):
    """w: the uint32 words of one (2,1,1) chunk (metadata first) -> its 1024 uint32 outputs"""
    cb1, cb2, vb, seed = int(w[1]), int(w[2]), int(w[3]), int(w[4])
    a1, a2, av = 5, 5 + cb1 // 4, 5 + cb1 // 4 + cb2 // 4
    len1 = unpack_bp(w[a1:a1 + cb1 // 4], 16).astype(np.int64)
    len2 = unpack_bp(w[a2:a2 + cb2 // 4], 16).astype(np.int64)
    val2 = unpack_bp(w[av:av + vb // 4], 32).astype(np.int64)
    n2 = len(val2)
    len2 = len2[:n2]
    assert (len2 >= 1).all()
    # layer 2: run k stores val2[k] - val2[k-1] at its start; the running sum is the delta array
    nv = int(len2.sum())
    e = np.zeros(nv, np.int64)
    starts2 = np.cumsum(len2) - len2
    e[starts2] = np.diff(val2, prepend=0)
    deltas = np.cumsum(e) & 0xFFFFFFFF
    # layer 1 + delta: run j adds its delta (the seed for run 0) at its start; the running sum is the output
    m = nv + 1
    starts1 = (np.cumsum(len1[:m]) - len1[:m])
    assert int(len1[:m].sum()) == 1024
    o_ = np.zeros(1024, np.int64)
    np.add.at(o_, starts1, np.concatenate([[seed], deltas]))
    return (np.cumsum(o_) & 0xFFFFFFFF).astype(np.uint32)


@pytest.mark.parametrize("seed", range(6))
def test_default_options_decoder_formulation(seed):
    import datagen

    rng = np.random.default_rng(seed)
    cols = [datagen.sorted_ints(4096, seed).astype(np.uint32),
            np.repeat(rng.integers(0, 1 << 20, 300), rng.integers(1, 30, 300))[:4096].astype(np.uint32),
            np.cumsum(np.repeat(rng.integers(1, 5, 700), rng.integers(1, 12, 700))[:4096]).astype(np.uint32),
            np.zeros(4096, np.uint32),
            (np.arange(4096) // 3 * 7).astype(np.uint32)]
    for x in cols:
        assert len(x) == 4096
        comp = o.casc_compress(x.view(np.uint8), 5, 2, 1, 1)
        w = comp.view(np.uint32)
        assert (w[0] & 0xFFFFFF) == (2 | 1 << 8 | 1 << 16), "stays compressed"
        pos, out = 2, []
        while pos < len(w):
            nbytes = int(w[pos])
            out.append(decode_chunk_r2d1_model(w[pos:pos + nbytes // 4]))
            pos += nbytes // 4
        assert np.array_equal(np.concatenate(out), x)
