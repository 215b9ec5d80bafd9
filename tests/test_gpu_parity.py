"""-m gpu parity tests: the CUDA path through the C ABI vs the CPU oracle (bit-exact)."""
import numpy as np
import pytest
import torch

import datagen
import oracle_lib as o

pytestmark = pytest.mark.gpu

DTYPES = {0: np.int8, 1: np.uint8, 2: np.int16, 3: np.uint16, 4: np.int32, 5: np.uint32, 6: np.int64, 7: np.uint64}


def _g():
    import gpu_common

    return gpu_common


def ragged_chunks(max_size, n, seed):
    rng = np.random.default_rng(seed)
    base = datagen.mixed_chunks(max_size, n, seed)
    sizes = [0, 1, 2, 3, 4, 5, 11, 12, 13, 14, 31, 32, 33, 63, 64, 65, 255, 256, 257, 1021] + list(
        rng.integers(1, max_size + 1, n))
    return [c[: int(s)] for c, s in zip(base, sizes[:n])]


# ------------------------------------------------------------------ LZ4
@pytest.mark.parametrize("tsize,type_id", [(1, 0), (1, 0xFF), (2, 3), (4, 4)])
@pytest.mark.parametrize("max_chunk", [4096, 16384, 65536])
def test_lz4_team_encoder_bit_exact_with_model(tsize, type_id, max_chunk):
    """Chunks <= 64 KiB: the stream is byte-identical to the sequential statement of our own policy
    (oracle/lz_team_model.c), decodes with liblz4 and with the reference-policy oracle decoder, and is never
    larger than GetMaxOutputChunkSize."""
    from test_oracle_lz4 import stock_decompress

    g = _g()
    chunks = [c[: len(c) // tsize * tsize] for c in ragged_chunks(max_chunk, 48, 1000 + max_chunk + tsize)]
    chunks += datagen.mixed_chunks(max_chunk, 12, 77)
    comp, dec, status, sizes, actual = g.roundtrip("lz4", chunks, opts=g.L.LZ4Opts(type_id), max_chunk=max_chunk)
    for i, c in enumerate(chunks):
        exp = o.lz4_team_compress(c, max_chunk)
        assert comp[i].tobytes() == exp.tobytes(), f"chunk {i} (len {len(c)}) compressed stream differs"
        assert len(comp[i]) <= o.lz4_max_compressed(max_chunk)
        assert status[i] == 0 and sizes[i] == len(c) and actual[i] == len(c)
        assert dec[i].tobytes() == c.tobytes()
        if len(c):
            n, out = stock_decompress(comp[i], len(c))
            assert n == len(c) and out.tobytes() == c.tobytes()


@pytest.mark.parametrize("kind", ["logs", "text", "ints", "zeros", "random", "line37"])
def test_lz4_team_encoder_ratio_vs_reference_policy(kind):
    """North star: compression ratio within 1 % of the reference encoder on every input class.  The reference
    policy is oracle/lz4_oracle.c (byte-identical to the reference GPU encoder, tests/test_gpu_reference.py)."""
    g = _g()
    n, cb = 2 << 20, 65536
    rng = np.random.default_rng(11)
    raw = {"logs": lambda: datagen.log_lines(n, 21), "text": lambda: datagen.zipf_text(n, 22),
           "ints": lambda: datagen.sorted_ints(n // 4, 23).view(np.uint8), "zeros": lambda: np.zeros(n, np.uint8),
           "random": lambda: rng.integers(0, 256, n, dtype=np.uint8),
           "line37": lambda: np.frombuffer((b"the quick brown fox jumps over a dog\n" * (n // 37 + 1))[:n], np.uint8)}[kind]()
    chunks = [raw[i : i + cb] for i in range(0, n, cb)]
    comp, dec, status, sizes, actual = g.roundtrip("lz4", chunks, max_chunk=cb)
    ours = sum(len(c) for c in comp)
    ref = sum(len(o.lz4_compress(c, 1, cb)) for c in chunks)
    assert all(d.tobytes() == c.tobytes() for d, c in zip(dec, chunks)) and (status == 0).all()
    assert n / ours >= 0.99 * (n / ref), f"{kind}: ratio {n / ours:.4f} vs reference policy {n / ref:.4f}"


def test_lz4_decodes_liblz4_streams():
    g = _g()
    from test_oracle_lz4 import stock_compress

    chunks = ragged_chunks(65536, 40, 5)
    comp = [stock_compress(c) for c in chunks]
    out, actual, status = g.decode_raw("lz4", comp, [len(c) for c in chunks])
    assert (status.cpu().numpy() == 0).all()
    for d, c in zip(out.to_host(actual), chunks):
        assert d.tobytes() == c.tobytes()


def test_lz4_large_chunk_and_null_outputs():
    g = _g()
    chunks = [datagen.log_lines(1 << 20, 3), datagen.zipf_text(300000, 4), np.zeros(1 << 20, np.uint8)]
    comp, dec, status, sizes, actual = g.roundtrip("lz4", chunks, max_chunk=1 << 20)
    for i, c in enumerate(chunks):
        # above 64 KiB the encoder keeps the reference's policy: same bytes as the reference-policy oracle
        assert comp[i].tobytes() == o.lz4_compress(c, 1, 1 << 20).tobytes()
        assert dec[i].tobytes() == c.tobytes() and status[i] == 0
    # NULL actual_bytes / statuses are allowed (reference SUPPORT_NULLPTR_APIS)
    out, actual, status = g.decode_raw("lz4", comp, [len(c) for c in chunks], want_actual=False, want_status=False)
    assert actual is None and status is None
    for d, c in zip(out.to_host([len(c) for c in chunks]), chunks):
        assert d.tobytes() == c.tobytes()


def test_lz4_corrupt_input_matches_oracle_and_never_faults():
    g = _g()
    rng = np.random.default_rng(0)
    junk = [(rng.integers(0, 4, int(n)) + 300).astype(np.int32).view(np.uint8) for n in rng.integers(25, 250, 64)]
    junk += [rng.integers(0, 256, int(n), dtype=np.uint8) for n in rng.integers(1, 2000, 64)]
    compb = g.B.DeviceBatch.from_host(junk)
    sizes = g.B.decompress_sizes("lz4", compb).cpu().numpy()
    caps = []
    for s, j in zip(sizes, junk):
        st, n = o.lz4_decompress(j, 0, write_output=False)
        assert s == (n if st == 0 else 0)
        caps.append(len(j) if (s == 0 or s > len(j)) else int(s) - 1)
    out, actual, status = g.decode_raw("lz4", junk, caps)
    st = status.cpu().numpy()
    for i, (j, cap) in enumerate(zip(junk, caps)):
        est, eout = o.lz4_decompress(j, cap)
        assert st[i] == est
        assert actual.cpu().numpy()[i] == len(eout)
    assert (st == 12).all()  # the reference test requires every chunk to be flagged
    # good stream into a too-small buffer
    good = o.lz4_compress(datagen.log_lines(5000, 1))
    _, actual, status = g.decode_raw("lz4", [good, good], [4999, 5000])
    assert status.cpu().tolist() == [12, 0] and actual.cpu().tolist() == [0, 5000]


# ------------------------------------------------------------------ Snappy
@pytest.mark.parametrize("max_chunk", [4096, 16384, 65536, 1 << 20])
def test_snappy_bit_exact_with_oracle(max_chunk):
    """Chunks <= 64 KiB: byte-identical to the sequential statement of our own policy (oracle/lz_team_model.c);
    above: to the reference-policy oracle.  Every stream decodes with stock snappy and fits GetMaxOutputChunkSize."""
    import pyarrow as pa

    stock = pa.Codec("snappy")
    g = _g()
    chunks = ragged_chunks(max_chunk, 48, 2000 + max_chunk) + datagen.mixed_chunks(max_chunk, 12, 78)
    comp, dec, status, sizes, actual = g.roundtrip("snappy", chunks, max_chunk=max_chunk)
    for i, c in enumerate(chunks):
        exp = o.snappy_product_compress(c, max_chunk)
        assert comp[i].tobytes() == exp.tobytes(), f"chunk {i} (len {len(c)}) compressed stream differs"
        assert len(comp[i]) <= o.snappy_max_compressed(max_chunk)
        assert sizes[i] == len(c)
        if len(c):
            assert status[i] == 0 and actual[i] == len(c) and dec[i].tobytes() == c.tobytes()
            assert stock.decompress(comp[i].tobytes(), decompressed_size=len(c)).to_pybytes() == c.tobytes()


@pytest.mark.parametrize("kind", ["logs", "text", "ints", "zeros", "random", "line37"])
def test_snappy_team_encoder_ratio_vs_reference_policy(kind):
    """North star: compression ratio within 1 % of the reference encoder on every input class (reference policy =
    oracle/snappy_oracle.c, byte-identical to the reference GPU encoder)."""
    g = _g()
    n, cb = 2 << 20, 65536
    rng = np.random.default_rng(11)
    raw = {"logs": lambda: datagen.log_lines(n, 21), "text": lambda: datagen.zipf_text(n, 22),
           "ints": lambda: datagen.sorted_ints(n // 4, 23).view(np.uint8), "zeros": lambda: np.zeros(n, np.uint8),
           "random": lambda: rng.integers(0, 256, n, dtype=np.uint8),
           "line37": lambda: np.frombuffer((b"the quick brown fox jumps over a dog\n" * (n // 37 + 1))[:n], np.uint8)}[kind]()
    chunks = [raw[i : i + cb] for i in range(0, n, cb)]
    comp, dec, status, sizes, actual = g.roundtrip("snappy", chunks, max_chunk=cb)
    ours = sum(len(c) for c in comp)
    ref = sum(len(o.snappy_compress(c)) for c in chunks)
    assert all(d.tobytes() == c.tobytes() for d, c in zip(dec, chunks)) and (status == 0).all()
    assert n / ours >= 0.99 * (n / ref), f"{kind}: ratio {n / ours:.4f} vs reference policy {n / ref:.4f}"


def test_snappy_reference_golden_and_unusual_streams():
    g = _g()
    import os

    from test_oracle_snappy import GOLD, copy_elem, lit_header, varint

    comp, raw = [], []
    for i in (1, 2):
        comp.append(np.frombuffer(open(os.path.join(GOLD, f"snappy_app_{i}.comp"), "rb").read(), np.uint8))
        raw.append(open(os.path.join(GOLD, f"snappy_app_{i}.raw"), "rb").read())
    big = bytes(i % 256 for i in range(512))
    comp.append(np.frombuffer(varint(512) + lit_header(512) + big, np.uint8))
    raw.append(big)
    for n in ((1 << 15) + 100, (1 << 16) + 100):
        head = np.random.default_rng(42).integers(0, 256, n, dtype=np.uint8).tobytes()
        comp.append(np.frombuffer(varint(n + 35) + lit_header(n) + head + copy_elem(n, 35), np.uint8))
        raw.append(head + head[:35])
    out, actual, status = g.decode_raw("snappy", comp, [len(r) for r in raw])
    assert (status.cpu().numpy() == 0).all()
    for d, r in zip(out.to_host(actual), raw):
        assert d.tobytes() == r


def test_snappy_decodes_stock_streams_and_errors():
    g = _g()
    import pyarrow as pa

    codec = pa.Codec("snappy")
    chunks = [c for c in ragged_chunks(65536, 40, 6) if len(c)]
    comp = [np.frombuffer(codec.compress(c.tobytes()).to_pybytes(), np.uint8) for c in chunks]
    out, actual, status = g.decode_raw("snappy", comp, [len(c) for c in chunks])
    assert (status.cpu().numpy() == 0).all()
    for d, c in zip(out.to_host(actual), chunks):
        assert d.tobytes() == c.tobytes()
    # truncated / undersized: status and partial progress equal the oracle's
    good = o.snappy_compress(datagen.zipf_text(5000, 3))
    cases = [(good[:cut], 5000) for cut in range(1, len(good), 211)] + [(good, 4999), (good, 5000)]
    out, actual, status = g.decode_raw("snappy", [c for c, _ in cases], [cap for _, cap in cases])
    for i, (c, cap) in enumerate(cases):
        est, eout = o.snappy_decompress(c, cap)
        assert status.cpu().numpy()[i] == est and actual.cpu().numpy()[i] == len(eout)


# ------------------------------------------------------------------ Cascaded
@pytest.mark.parametrize("type_id", sorted(DTYPES))
@pytest.mark.parametrize("opts", [(2, 1, 1), (1, 1, 1), (2, 1, 0), (0, 0, 1), (1, 0, 0), (0, 1, 0), (0, 0, 0), (2, 2, 1),
                                  (3, 1, 1), (1, 2, 1), (7, 0, 1)])
def test_cascaded_bit_exact_with_oracle(type_id, opts):
    g = _g()
    dt = DTYPES[type_id]
    w = np.dtype(dt).itemsize
    rng = np.random.default_rng(type_id * 100 + sum(opts))
    parts = []
    for n in (1, 2, 3, 31, 32, 33, 100, 4096 // w - 1, 4096 // w, 4096 // w + 1, 3 * 4096 // w + 17, 20000):
        parts.append(datagen.sorted_ints(n, int(rng.integers(1 << 30))).astype(dt))
    parts.append(np.zeros(5000, dt))                                     # one run per chunk
    parts.append(rng.integers(0, 256, 6000 * w, dtype=np.uint8).view(dt))  # incompressible -> raw fallback
    parts.append(np.repeat(np.array([3, 9, 4, 0, 1]), [1, 20, 13, 25, 6]).astype(dt))  # reference KAT input
    parts.append((np.arange(3000) * 7 - 5000).astype(dt))                # negative values, constant delta
    chunks = [p.view(np.uint8) for p in parts]
    co = g.L.CascadedOpts(4096, type_id, *opts)
    comp, dec, status, sizes, actual = g.roundtrip("cascaded", chunks, opts=co)
    for i, c in enumerate(chunks):
        exp = o.casc_compress(c, type_id, *opts)
        assert comp[i].tobytes() == exp.tobytes(), f"partition {i} ({len(c)} B) compressed bytes differ from oracle"
        assert sizes[i] == len(c) and status[i] == 0 and actual[i] == len(c)
        assert dec[i].tobytes() == c.tobytes()


@pytest.mark.parametrize("opts", [(1, 1, 1), (2, 1, 1), (1, 1, 0), (2, 1, 0)])
def test_cascaded_incompressible_and_default_option_fast_paths(opts):
    """The encoder's look-ahead on partitions whose first chunk is all distinct (sizes added up from the adjacent
    differences before anything is encoded) and the fused num_RLEs = 2 decoder: same bytes as the oracle on the
    cases that steer them."""
    g = _g()
    rng = np.random.default_rng(sum(opts) + 77)
    n = 1024
    parts = [
        rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32),           # one chunk, does not fit
        rng.integers(0, 1 << 32, 2 * n, dtype=np.uint64).astype(np.uint32),       # first chunk of two does not fit
        rng.integers(0, 1 << 32, n + 5, dtype=np.uint64).astype(np.uint32),       # full chunk + ragged tail
        rng.permutation(1 << 20)[:n].astype(np.uint32),                           # all distinct, 21-bit deltas: fits
        rng.permutation(1 << 12)[: 3 * n].astype(np.uint32),                      # all distinct, small range
        (np.arange(n) * 3 + 11).astype(np.uint32),                                # all distinct, every delta equal
        np.cumsum(np.tile([1, 1, 2, 5, 5, 5, 9, 1], n // 8)).astype(np.uint32),   # all distinct, deltas repeat in places
        np.cumsum(rng.integers(1, 3, 4 * n)).astype(np.uint32),                   # all distinct, deltas 1 or 2
        np.concatenate([rng.integers(0, 1 << 32, n - 2, dtype=np.uint64).astype(np.uint32), [7, 7]]).astype(np.uint32),
        np.concatenate([[5, 5], rng.integers(0, 1 << 32, n - 2, dtype=np.uint64).astype(np.uint32)]).astype(np.uint32),
        datagen.sorted_ints(8 * n, 5).astype(np.uint32),                          # the usual compressible column
        np.repeat(rng.integers(0, 1000, 400), rng.integers(1, 9, 400))[: 2 * n].astype(np.uint32),  # runs in both layers
    ]
    # the last delta pair of a chunk (elements N-3 .. N-1) equal / the pair that straddles two lanes equal
    x = rng.permutation(1 << 24)[:n].astype(np.int64)
    x[-1] = 2 * x[-2] - x[-3]
    parts.append(x.astype(np.uint32))
    y = rng.permutation(1 << 24)[:n].astype(np.int64)
    y[5] = 2 * y[4] - y[3]
    parts.append(y.astype(np.uint32))
    chunks = [p.view(np.uint8) for p in parts]
    co = g.L.CascadedOpts(4096, g.L.TYPE_UINT, *opts)
    comp, dec, status, sizes, actual = g.roundtrip("cascaded", chunks, opts=co)
    for i, c in enumerate(chunks):
        exp = o.casc_compress(c, g.L.TYPE_UINT, *opts)
        assert comp[i].tobytes() == exp.tobytes(), f"partition {i} ({len(c)} B) compressed bytes differ from oracle"
        assert sizes[i] == len(c) and status[i] == 0 and actual[i] == len(c)
        assert dec[i].tobytes() == c.tobytes()
    # destinations that are only 4-byte aligned
    src = g.B.DeviceBatch.from_host(chunks[:3])
    stride = (g.B.max_compressed_chunk_bytes("cascaded", 8192 + 64, co) + 15) // 16 * 16
    out = g.B.DeviceBatch.empty(src.n, stride)
    shifted = g.B.DeviceBatch(out.buf, out.offsets + 4, np.full(src.n, stride - 4, np.int64))
    comp2 = g.B.compress("cascaded", src, opts=co, max_chunk=8192 + 64, out=shifted)
    torch.cuda.synchronize()
    for got, c in zip(comp2.to_host(), chunks[:3]):
        assert got.tobytes() == o.casc_compress(c, g.L.TYPE_UINT, *opts).tobytes()


def test_cascaded_default_options_decoder_rejects_what_the_generic_layers_reject():
    """Streams the fused num_RLEs = 2 decoder hands back to the generic layers: an empty run in the second layer,
    truncated arrays; statuses and sizes equal the oracle's."""
    g = _g()
    x = datagen.sorted_ints(1024, 9).astype(np.uint32)
    good = o.casc_compress(x.view(np.uint8), g.L.TYPE_UINT, 2, 1, 1)
    cases = [good.copy()]
    for cut in range(8, len(good), 37):
        cases.append(good[:cut].copy())
    n_exact = len(cases)  # whole + truncated: the verdict equals the oracle's
    for word in range(2, min(len(good) // 4, 40)):
        for flip in (0x1, 0x10000, 0xFFFF):
            bad = good.copy()
            bad.view(np.uint32)[word] ^= flip
            cases.append(bad)
    out, actual, status = g.decode_raw("cascaded", cases, [4096] * len(cases))
    st, ac = status.cpu().numpy(), actual.cpu().numpy()
    host = out.to_host(np.maximum(ac, 0))
    for i, c in enumerate(cases):
        est, eout = o.casc_decompress(c, 4096, 4)
        if i < n_exact:
            assert (st[i] == 0) == (est == 0), f"case {i}"
        # damaged header words: never a fault, and a stream both accept decodes to the same bytes
        assert st[i] in (0, 12)
        if est == 0 and st[i] == 0:
            assert ac[i] == len(eout) and host[i].tobytes() == eout.tobytes(), f"case {i}"


def test_cascaded_reference_kat_words():
    """tests/test_cascaded_batch.cpp:336-378 through the C ABI (int32, RLE=2, delta=1, no bit-packing)."""
    g = _g()
    x = np.repeat(np.array([3, 9, 4, 0, 1], np.int32), [1, 20, 13, 25, 6])
    comp, *_ = g.roundtrip("cascaded", [x.view(np.uint8)], opts=g.L.CascadedOpts(4096, 4, 2, 1, 0))
    w = comp[0].view(np.uint32)
    assert w[0] == 2 + (1 << 8) + (4 << 24) and w[1] == x.nbytes
    assert list(w[3:6]) == [10, 8, 16] and w[6].view(np.int32) == 3
    assert list(comp[0][28:38].view(np.uint16)) == [1, 20, 13, 25, 6]
    assert list(comp[0][40:48].view(np.uint16)) == [1, 1, 1, 1]
    assert list(comp[0][48:64].view(np.int32)) == [6, -5, -4, 1]


def test_cascaded_oob_matrix_and_size_query():
    g = _g()
    x = np.repeat(np.arange(3000, dtype=np.int32), 5).view(np.uint8)
    comp = o.casc_compress(x, 4, 2, 1, 1)
    cb, ub = len(comp), len(x)
    cases = [(comp[: cb // 2 // 4 * 4], ub, 12), (comp, ub // 2, 12), (comp, ub - 1, 12), (comp, ub, 0)]
    out, actual, status = g.decode_raw("cascaded", [c for c, _, _ in cases], [cap for _, cap, _ in cases])
    assert status.cpu().tolist() == [e for _, _, e in cases]
    assert actual.cpu().tolist() == [0, 0, 0, ub]
    # truncations: same verdict as the oracle, never a fault
    cuts = [comp[:k] for k in range(8, cb, 52)]
    out, actual, status = g.decode_raw("cascaded", cuts, [ub] * len(cuts))
    for i, c in enumerate(cuts):
        est, _ = o.casc_decompress(c, ub, 4)
        assert status.cpu().numpy()[i] == est
    short = g.B.DeviceBatch.from_host([comp[:4], comp])
    assert g.B.decompress_sizes("cascaded", short).cpu().tolist() == [0, ub]


# ------------------------------------------------------------------ properties at scale
@pytest.mark.parametrize("codec", ["lz4", "snappy"])
def test_roundtrip_property_many_chunks(codec):
    g = _g()
    raw = np.concatenate([datagen.log_lines(8 << 20, 21), datagen.zipf_text(8 << 20, 22)])
    chunks = [raw[i : i + 65536] for i in range(0, len(raw), 65536)]
    comp, dec, status, sizes, actual = g.roundtrip(codec, chunks, max_chunk=65536)
    assert (status == 0).all() and (actual == 65536).all() and (sizes == 65536).all()
    assert np.concatenate(dec).tobytes() == raw.tobytes()
    ratio = len(raw) / sum(len(c) for c in comp)
    assert ratio > 1.5


# ------------------------------------------------------------------ boundary
@pytest.mark.parametrize("codec", ["lz4", "snappy", "cascaded"])
def test_host_pointer_arrays_are_rejected_not_dereferenced(codec):
    """The pointer / size arrays must be device-accessible: plain host memory comes back as hipcompErrorInvalidValue
    (reference HipUtils::device_pointer, src/HipUtils.hip:91-108), it does not fault in the kernel."""
    import ctypes as C

    g = _g()
    lib = g.L.lib()
    name = {"lz4": "LZ4", "snappy": "Snappy", "cascaded": "Cascaded"}[codec]
    opts = g.L.CascadedOpts(4096, g.L.TYPE_INT, 1, 1, 1) if codec == "cascaded" else g.B.default_opts(codec)
    host = (C.c_size_t * 8)()
    dev = torch.zeros(64, dtype=torch.int64, device="cuda")
    temp = torch.zeros(1024, dtype=torch.uint8, device="cuda")
    f = getattr(lib, f"hipcompBatched{name}CompressAsync")
    st = f(C.addressof(host), dev.data_ptr(), 4096, 4, temp.data_ptr(), temp.numel(), dev.data_ptr(), dev.data_ptr(), opts, None)
    assert st == g.L.ERROR_INVALID_VALUE
    f = getattr(lib, f"hipcompBatched{name}DecompressAsync")
    st = f(dev.data_ptr(), C.addressof(host), dev.data_ptr(), dev.data_ptr(), 4, temp.data_ptr(), temp.numel(), dev.data_ptr(),
           None, None)
    assert st == g.L.ERROR_INVALID_VALUE
    torch.cuda.synchronize()


# ------------------------------------------------------------------ team encoder at scale
@pytest.mark.parametrize("codec", ["lz4", "snappy"])
def test_team_encoder_matches_model_at_scale_and_is_deterministic(codec):
    """1024 chunks of 64 KiB (every SM busy, many chunks per CTA): every stream equals the sequential model's, and
    a second run gives the same bytes - the encoder's cross-warp protocol (mbarrier chains, shared-memory atomic max,
    in-order shared-memory accesses inside the serial section) leaves no room for timing to show."""
    g = _g()
    cb, n = 65536, 1024
    base = np.concatenate([datagen.log_lines(8 << 20, 91), datagen.zipf_text(8 << 20, 92),
                           datagen.sorted_ints((4 << 20) // 4, 93).view(np.uint8),
                           np.frombuffer((b"the quick brown fox jumps over a dog\n" * ((2 << 20) // 37 + 1))[: 2 << 20], np.uint8),
                           np.zeros(1 << 20, np.uint8), np.random.default_rng(94).integers(0, 256, 1 << 20, dtype=np.uint8)])
    raw = np.resize(base, cb * n)
    chunks = [raw[i * cb:(i + 1) * cb] for i in range(n)]
    src = g.B.DeviceBatch.from_host(chunks)
    first = g.B.compress(codec, src, max_chunk=cb).to_host()
    second = g.B.compress(codec, src, max_chunk=cb).to_host()
    model = (lambda c: o.lz4_team_compress(c, cb)) if codec == "lz4" else (lambda c: o.snappy_team_compress(c, cb))
    seen = {}
    for i, c in enumerate(chunks):
        key = c.tobytes()
        if key not in seen:  # np.resize repeats the base: model each distinct chunk once
            seen[key] = model(c).tobytes()
        assert first[i].tobytes() == seen[key], f"chunk {i}: stream differs from the model"
        assert second[i].tobytes() == first[i].tobytes(), f"chunk {i}: two runs differ"
