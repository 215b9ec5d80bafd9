"""-m gpu: the LZ4 / Snappy decoders on inputs chosen for their copy paths: ragged sizes, periodic data with every
period 1..40 (overlapping matches shorter / longer than a warp step), long matches and long literal runs (the
16-byte vector paths), liblz4 / stock snappy streams, and corrupt streams (statuses must equal the oracle's)."""
import numpy as np
import pytest

import datagen
import oracle_lib as o

pytestmark = pytest.mark.gpu


def _g():
    import gpu_common

    return gpu_common


def periodic_chunks(n):
    rng = np.random.default_rng(99)
    out = []
    for period in list(range(1, 41)) + [63, 64, 65, 127, 128, 129, 255, 256, 257, 511, 512, 513, 1000, 4096]:
        pat = rng.integers(0, 256, period, dtype=np.uint8)
        reps = n // period + 1
        c = np.tile(pat, reps)[:n].copy()
        # a few random bytes so that the stream has literal runs between the long matches
        pos = rng.integers(0, n, 6)
        c[pos] = rng.integers(0, 256, 6, dtype=np.uint8)
        out.append(c)
    out.append(rng.integers(0, 256, n, dtype=np.uint8))  # one long literal run
    out.append(np.zeros(n, np.uint8))
    return out


@pytest.fixture
def width():
    return 0  # one decoder per codec; the name is kept in the assertion messages


@pytest.mark.parametrize("codec", ["lz4", "snappy"])
def test_roundtrip_every_group_width(codec, width):
    g = _g()
    for max_chunk, chunks in ((9000, periodic_chunks(9000)), (65536, datagen.mixed_chunks(65536, 13, 5)),
                              (2048, [c[: (37 * i) % 2049] for i, c in enumerate(datagen.mixed_chunks(2048, 67, 8))])):
        comp, dec, status, sizes, actual = g.roundtrip(codec, chunks, max_chunk=max_chunk)
        for i, c in enumerate(chunks):
            assert status[i] == 0 and actual[i] == len(c) and sizes[i] == len(c), (width, i)
            assert dec[i].tobytes() == c.tobytes(), f"G={width} chunk {i} (len {len(c)})"


def test_lz4_stock_and_corrupt_streams_every_group_width(width):
    g = _g()
    from test_oracle_lz4 import stock_compress

    chunks = periodic_chunks(5000) + datagen.mixed_chunks(4096, 12, 3)
    comp = [stock_compress(c) for c in chunks]
    out, actual, status = g.decode_raw("lz4", comp, [len(c) for c in chunks])
    assert (status.cpu().numpy() == 0).all()
    for d, c in zip(out.to_host(actual), chunks):
        assert d.tobytes() == c.tobytes()
    # corrupt: flip bytes / truncate; statuses and sizes must equal the oracle's verdicts
    rng = np.random.default_rng(11)
    bad, caps = [], []
    for c, s in zip(chunks, comp):
        s = np.array(s, copy=True)
        if len(s) > 8:
            k = rng.integers(0, 3)
            if k == 0:
                s[rng.integers(0, len(s), 3)] ^= rng.integers(1, 256, 3, dtype=np.uint8)
            elif k == 1:
                s = s[: rng.integers(1, len(s))]
            else:
                s[rng.integers(0, min(len(s), 16))] = 0xFF
        bad.append(s)
        caps.append(len(c))
    out, actual, status = g.decode_raw("lz4", bad, caps)
    actual, status = actual.cpu().numpy(), status.cpu().numpy()
    host = out.to_host(actual)
    for i, (s, cap) in enumerate(zip(bad, caps)):
        st, exp = o.lz4_decompress(s, cap)
        assert status[i] == st, (width, i)
        assert actual[i] == len(exp), (width, i)
        if st == 0:
            assert host[i].tobytes() == exp.tobytes()


def test_snappy_stock_and_corrupt_streams_every_group_width(width):
    g = _g()
    import pyarrow as pa

    codec = pa.Codec("snappy")
    chunks = periodic_chunks(5000) + datagen.mixed_chunks(4096, 12, 3)
    comp = [np.frombuffer(codec.compress(c.tobytes()).to_pybytes(), np.uint8) for c in chunks]
    out, actual, status = g.decode_raw("snappy", comp, [len(c) for c in chunks])
    assert (status.cpu().numpy() == 0).all()
    for d, c in zip(out.to_host(actual), chunks):
        assert d.tobytes() == c.tobytes()
    rng = np.random.default_rng(12)
    bad, caps = [], []
    for c, s in zip(chunks, comp):
        s = np.array(s, copy=True)
        if len(s) > 8:
            if rng.integers(0, 2):
                s[rng.integers(2, len(s), 3)] ^= rng.integers(1, 256, 3, dtype=np.uint8)
            else:
                s = s[: rng.integers(3, len(s))]
        bad.append(s)
        caps.append(len(c))
    out, actual, status = g.decode_raw("snappy", bad, caps)
    actual, status = actual.cpu().numpy(), status.cpu().numpy()
    host = out.to_host(actual)
    for i, (s, cap) in enumerate(zip(bad, caps)):
        st, exp = o.snappy_decompress(s, cap)
        assert status[i] == st, (width, i)
        assert actual[i] == len(exp), (width, i)
        assert host[i].tobytes() == exp.tobytes()


@pytest.mark.parametrize("codec", ["lz4", "snappy"])
def test_streams_flush_against_the_buffer_ends(codec):
    """Streams that start at byte 1 of an allocation and end on its last byte, at every 16-byte phase: the decoders stage
    whole aligned blocks and must not read the bytes around the stream (run this file under
    `PYTORCH_NO_CUDA_MEMORY_CACHING=1 compute-sanitizer --tool memcheck`, summary in profiles/)."""
    import torch

    g = _g()
    rng = np.random.default_rng(5)
    for phase in range(0, 16, 3):
        raw = datagen.log_lines(3000 + 7 * phase, 40 + phase)
        comp = (o.lz4_compress(raw) if codec == "lz4" else o.snappy_compress(raw))
        buf = torch.empty(phase + 1 + len(comp), dtype=torch.uint8, device="cuda")  # its own allocation, exact size
        buf[: phase + 1] = 0xEE
        buf[phase + 1:] = torch.from_numpy(np.ascontiguousarray(comp)).cuda()
        batch = g.B.DeviceBatch(buf, np.array([phase + 1], dtype=np.int64), np.array([len(comp)], dtype=np.int64))
        out = g.B.DeviceBatch.empty(1, len(raw) + 16)
        out.sizes = torch.tensor([len(raw)], dtype=torch.int64, device="cuda")
        actual, status = g.B.decompress(codec, batch, out)
        torch.cuda.synchronize()
        assert int(status[0]) == 0 and int(actual[0]) == len(raw)
        assert out.to_host(actual)[0].tobytes() == raw.tobytes()
