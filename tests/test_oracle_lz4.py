"""Pins oracle/lz4_oracle.c at the LZ4 block format: liblz4 must decode the
oracle encoder's output and the oracle decoder must decode liblz4's output
(the reference holds no byte-level LZ4 vectors)."""
import ctypes as C

import numpy as np
import pytest

import datagen
import oracle_lib as o

_lz4 = C.CDLL("liblz4.so.1")
_lz4.LZ4_compressBound.restype = C.c_int
_lz4.LZ4_compress_default.restype = C.c_int
_lz4.LZ4_compress_default.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
_lz4.LZ4_decompress_safe.restype = C.c_int
_lz4.LZ4_decompress_safe.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]


def stock_compress(raw):
    raw = np.ascontiguousarray(raw)
    out = np.zeros(_lz4.LZ4_compressBound(len(raw)) + 8, np.uint8)
    n = _lz4.LZ4_compress_default(raw.ctypes.data, out.ctypes.data, len(raw), len(out))
    assert n > 0 or len(raw) == 0
    return out[:n].copy()


def stock_decompress(comp, size):
    comp = np.ascontiguousarray(comp)
    out = np.zeros(max(size, 1), np.uint8)
    n = _lz4.LZ4_decompress_safe(comp.ctypes.data, out.ctypes.data, len(comp), size)
    return n, out[: max(n, 0)]


@pytest.mark.parametrize("size", [1, 2, 4, 5, 12, 13, 14, 31, 32, 33, 45, 64, 255, 256, 1000, 4096, 65536, 70003])
@pytest.mark.parametrize("tsize", [1, 2, 4])
def test_cross_decode_with_liblz4(size, tsize):
    for chunk in datagen.mixed_chunks(size, 6, 900 + size):
        comp = o.lz4_compress(chunk, tsize, max_chunk=size)
        assert len(comp) <= o.lz4_max_compressed(size)
        n, out = stock_decompress(comp, size)
        assert n == size and out.tobytes() == chunk.tobytes()
        st, out = o.lz4_decompress(comp, size)
        assert st == o.SUCCESS and out.tobytes() == chunk.tobytes()
        st, n = o.lz4_decompress(comp, 0, write_output=False)
        assert st == o.SUCCESS and n == size
        stock = stock_compress(chunk)
        st, out = o.lz4_decompress(stock, size)
        assert st == o.SUCCESS and out.tobytes() == chunk.tobytes()


def test_empty_chunk():
    assert len(o.lz4_compress(np.zeros(0, np.uint8))) == 0
    st, out = o.lz4_decompress(np.zeros(0, np.uint8), 0)
    assert st == o.SUCCESS and len(out) == 0


def test_last_token_low_nibble_matches_reference():
    # the reference emits (0 - 4) & 0xF in the literal-only last token (LZ4Kernels.hiph:324-331)
    comp = o.lz4_compress(np.arange(10, dtype=np.uint8))
    assert comp[0] == (10 << 4) | 0xC


def test_hash_table_size_rule():
    assert o.lz4_hash_table_size(100) == 128
    assert o.lz4_hash_table_size(4096) == 4096
    assert o.lz4_hash_table_size(65536) == 16384
    assert o.lz4_hash_table_size(1 << 24) == 16384
    raw = datagen.log_lines(4096, 5)
    a = o.lz4_compress(raw, 1, max_chunk=4096)
    b = o.lz4_compress(raw, 1, max_chunk=65536)
    for c in (a, b):
        st, out = o.lz4_decompress(c, 4096)
        assert st == o.SUCCESS and out.tobytes() == raw.tobytes()


def test_decoder_error_conditions():
    raw = datagen.log_lines(3000, 2)
    comp = o.lz4_compress(raw)
    assert o.lz4_decompress(comp, 2999)[0] == o.CANNOT_DECOMPRESS
    st, out = o.lz4_decompress(comp, 2999)
    assert len(out) == 0
    # offset pointing before the start of the output
    bad = bytes([0x10, ord("a"), 5, 0]) + bytes([0x50]) + b"bcdef"
    assert o.lz4_decompress(bad, 100)[0] == o.CANNOT_DECOMPRESS
    # offset 0
    bad = bytes([0x40]) + b"abcd" + bytes([0, 0]) + bytes([0x50]) + b"bcdef"
    assert o.lz4_decompress(bad, 100)[0] == o.CANNOT_DECOMPRESS
    # raw ints as "compressed" data (reference corrupt-input test): never succeeds into size-1
    rng = np.random.default_rng(0)
    for _ in range(50):
        junk = (rng.integers(0, 4, 300) + 300).astype(np.int32).view(np.uint8)
        st, n = o.lz4_decompress(junk, 0, write_output=False)
        cap = len(junk) if (st != o.SUCCESS or n == 0 or n > len(junk)) else n - 1
        assert o.lz4_decompress(junk, cap)[0] == o.CANNOT_DECOMPRESS


def test_ratio_close_to_liblz4():
    raw = datagen.log_lines(1 << 20, 12)
    ours = sum(len(o.lz4_compress(raw[i : i + 65536])) for i in range(0, len(raw), 65536))
    stock = sum(len(stock_compress(raw[i : i + 65536])) for i in range(0, len(raw), 65536))
    assert ours < 1.3 * stock


@pytest.mark.parametrize("size", [1, 4, 5, 11, 12, 13, 14, 31, 32, 33, 64, 255, 1000, 1023, 1024, 1025, 2048, 4096, 5000,
                                  16384, 40000, 65535, 65536])
def test_team_model_streams_are_lz4_blocks(size):
    """oracle/lz_team_model.c (the sequential statement of the CUDA team encoder) emits plain LZ4 blocks:
    liblz4 and the reference-policy oracle decoder decode them, and they fit GetMaxOutputChunkSize."""
    for chunk in datagen.mixed_chunks(size, 6, 300 + size):
        comp = o.lz4_team_compress(chunk, size)
        assert 0 < len(comp) <= o.lz4_max_compressed(size)
        n, out = stock_decompress(comp, size)
        assert n == size and out.tobytes() == chunk.tobytes()
        st, out = o.lz4_decompress(comp, size)
        assert st == o.SUCCESS and out.tobytes() == chunk.tobytes()


def test_team_model_ratio_not_below_reference_policy():
    for gen, seed in ((datagen.log_lines, 5), (datagen.zipf_text, 6)):
        raw = gen(1 << 20, seed)
        chunks = [raw[i : i + 65536] for i in range(0, len(raw), 65536)]
        team = sum(len(o.lz4_team_compress(c, 65536)) for c in chunks)
        ref = sum(len(o.lz4_compress(c, 1, 65536)) for c in chunks)
        assert team <= ref * 1.01
    z = np.zeros(65536, np.uint8)
    assert len(o.lz4_team_compress(z, 65536)) == len(o.lz4_compress(z, 1, 65536))  # runs merge across segments


def test_maximum_chunk_size_of_the_format():
    """16 MiB, the largest chunk the C ABI accepts (reference LZ4Batch.cpp:154-170): the reference-policy oracle's stream
    for one such chunk decodes with liblz4 and with the oracle's own decoder, and stays inside GetMaxOutputChunkSize."""
    n = 1 << 24
    raw = np.resize(datagen.log_lines(1 << 20, 31), n).copy()
    raw[5 << 20:6 << 20] = np.random.default_rng(3).integers(0, 256, 1 << 20, dtype=np.uint8)  # an incompressible stretch
    comp = o.lz4_compress(raw, 1, n)
    assert len(comp) <= o.lz4_max_compressed(n)
    m, out = stock_decompress(comp, n)
    assert m == n and out.tobytes() == raw.tobytes()
    st, back = o.lz4_decompress(comp, n)
    assert st == 0 and back.tobytes() == raw.tobytes()
