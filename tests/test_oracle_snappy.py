"""Pins oracle/snappy_oracle.c: reference golden streams, reference encoder
known answers (src/test/SnappyLargeTokens_test.cpp), and stock snappy (pyarrow)."""
import os

import numpy as np
import pyarrow as pa
import pytest

import datagen
import oracle_lib as o

GOLD = os.path.join(os.path.dirname(__file__), "golden")
CODEC = pa.Codec("snappy")


def varint(n):
    out = bytearray()
    while n > 0x7F:
        out.append((n & 0x7F) | 0x80)
        n >>= 7
    out.append(n)
    return bytes(out)


def lit_header(n):
    n -= 1
    if n < 60:
        return bytes([n << 2])
    body = bytearray()
    while n > 0xFF:
        body.append(n & 0xFF)
        n >>= 8
    body.append(n)
    return bytes([(59 + len(body)) << 2]) + bytes(body)


def copy_elem(offset, length):
    if 4 <= length <= 11 and offset <= 2047:
        return bytes([((offset >> 8) << 5) | ((length - 4) << 2) | 1, offset & 0xFF])
    if offset < (1 << 16):
        return bytes([((length - 1) << 2) | 2, offset & 0xFF, offset >> 8])
    return bytes([((length - 1) << 2) | 3]) + int(offset).to_bytes(4, "little")


@pytest.mark.parametrize("i", [1, 2])
def test_reference_golden_streams(i):
    comp = open(os.path.join(GOLD, f"snappy_app_{i}.comp"), "rb").read()
    raw = open(os.path.join(GOLD, f"snappy_app_{i}.raw"), "rb").read()
    assert o.snappy_get_size(comp) == len(raw)
    st, out = o.snappy_decompress(comp, len(raw))
    assert st == o.SUCCESS and out.tobytes() == raw
    assert CODEC.decompress(comp, decompressed_size=len(raw)).to_pybytes() == raw


def test_encoder_kat_256_literals():
    # test_mock_literal_compressor (:381-405)
    raw = bytes(range(256))
    expect = varint(256) + lit_header(256) + raw
    assert o.snappy_compress(raw).tobytes() == expect


def test_encoder_kat_literals_then_copy():
    # test_mock_match_compressor (:410-450)
    raw = bytes(range(256)) + bytes(range(64))
    expect = varint(320) + lit_header(256) + bytes(range(256)) + copy_elem(256, 64)
    assert o.snappy_compress(raw).tobytes() == expect


def test_decode_large_literal():
    raw = bytes(i % 256 for i in range(512))
    comp = varint(512) + lit_header(512) + raw
    st, out = o.snappy_decompress(comp, 512)
    assert st == o.SUCCESS and out.tobytes() == raw


@pytest.mark.parametrize("n", [(1 << 15) + 100, (1 << 16) + 100])
def test_decode_long_offset_copies(n):
    rng = np.random.default_rng(42)
    head = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
    raw = head + head[:35]
    comp = varint(len(raw)) + lit_header(n) + head + copy_elem(n, 35)
    st, out = o.snappy_decompress(comp, len(raw))
    assert st == o.SUCCESS and out.tobytes() == raw
    assert CODEC.decompress(comp, decompressed_size=len(raw)).to_pybytes() == raw


@pytest.mark.parametrize("size", [0, 1, 3, 4, 5, 31, 32, 33, 255, 256, 257, 1000, 4096, 65536, 70001])
def test_cross_decode_with_stock_snappy(size):
    for k, chunk in enumerate(datagen.mixed_chunks(size, 6, 100 + size)):
        raw = chunk.tobytes()
        comp = o.snappy_compress(chunk)
        assert len(comp) <= o.snappy_max_compressed(size)
        assert CODEC.decompress(comp.tobytes(), decompressed_size=size).to_pybytes() == raw
        st, out = o.snappy_decompress(comp, size)
        assert st == o.SUCCESS and out.tobytes() == raw
        stock = CODEC.compress(raw).to_pybytes()
        st, out = o.snappy_decompress(stock, size)
        assert st == o.SUCCESS and out.tobytes() == raw
        assert o.snappy_get_size(stock) == size


def test_errors():
    raw = datagen.zipf_text(5000, 3)
    comp = o.snappy_compress(raw)
    assert o.snappy_decompress(b"", 10)[0] == o.CANNOT_DECOMPRESS
    assert o.snappy_decompress(comp, 4999)[0] == o.CANNOT_DECOMPRESS  # declared > capacity
    for cut in range(1, len(comp), 53):
        st, out = o.snappy_decompress(comp[:cut], 5000)
        assert st == o.CANNOT_DECOMPRESS
        assert raw[: len(out)].tobytes() == out.tobytes()  # partial progress is a true prefix
    bad = varint(8) + bytes([3 << 2]) + b"abcd" + copy_elem(5, 4)  # offset > produced
    assert o.snappy_decompress(bad, 8)[0] == o.CANNOT_DECOMPRESS
    bad = varint(8) + bytes([3 << 2]) + b"abcd" + bytes([1 | (0 << 2), 0])  # offset 0
    assert o.snappy_decompress(bad, 8)[0] == o.CANNOT_DECOMPRESS
    assert o.snappy_get_size(b"\xff\xff\xff\xff\x08") == 0  # >= 2^31


def test_ratio_close_to_stock_snappy():
    raw = datagen.log_lines(1 << 20, 11)
    ours = sum(len(o.snappy_compress(raw[i : i + 65536])) for i in range(0, len(raw), 65536))
    stock = sum(len(CODEC.compress(raw[i : i + 65536].tobytes())) for i in range(0, len(raw), 65536))
    assert ours < 1.35 * stock


@pytest.mark.parametrize("size", [0, 1, 4, 5, 11, 12, 13, 60, 61, 64, 255, 256, 257, 1024, 1025, 5000, 16384, 65535, 65536])
def test_team_model_streams_are_snappy_streams(size):
    """oracle/lz_team_model.c (the sequential statement of the CUDA team encoder) emits plain Snappy streams:
    stock snappy and the reference-grammar oracle decoder decode them, and they fit GetMaxOutputChunkSize."""
    import pyarrow as pa

    stock = pa.Codec("snappy")
    for chunk in datagen.mixed_chunks(max(size, 1), 6, 400 + size):
        chunk = chunk[:size]
        comp = o.snappy_team_compress(chunk, max(size, 1))
        assert 0 < len(comp) <= o.snappy_max_compressed(size)
        assert o.snappy_get_size(comp) == size
        if size:
            assert stock.decompress(comp.tobytes(), decompressed_size=size).to_pybytes() == chunk.tobytes()
            st, out = o.snappy_decompress(comp, size)
            assert st == o.SUCCESS and out.tobytes() == chunk.tobytes()


def test_team_model_ratio_not_below_reference_policy():
    for gen, seed in ((datagen.log_lines, 5), (datagen.zipf_text, 6)):
        raw = gen(1 << 20, seed)
        chunks = [raw[i : i + 65536] for i in range(0, len(raw), 65536)]
        team = sum(len(o.snappy_team_compress(c, 65536)) for c in chunks)
        ref = sum(len(o.snappy_compress(c)) for c in chunks)
        assert team <= ref * 1.01
