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
