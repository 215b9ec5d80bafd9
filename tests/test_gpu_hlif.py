"""HLIF managers (LZ4Manager / SnappyManager / CascadedManager / create_manager) on the GPU.

Modelled on the reference's HLIF tests (tests/test_lz4.cpp, tests/test_cascaded.cpp use
hipcomp.hpp managers: configure -> compress -> size -> configure_decompression -> decompress).
Parity: the container is parsed on the host (SURVEY A.4 layout) and every chunk is (1) decoded by
the CPU oracle and (2) compared byte for byte with the oracle ENCODER's stream for the same chunk.
"""
import importlib
import os
import struct

import numpy as np
import pytest
import torch

import datagen
import oracle_lib as o

pytestmark = pytest.mark.gpu

L = importlib.import_module("hipcomp-core_b200._lib")
M = importlib.import_module("hipcomp-core_b200.managers")

FMT = {"lz4": 0, "snappy": 1, "cascaded": 4}
SPEC_BYTES = {"lz4": 4, "snappy": 1, "cascaded": 24}


def make_manager(kind, chunk, opts=None):
    if kind == "lz4":
        return M.LZ4Manager(chunk, L.TYPE_CHAR)
    if kind == "snappy":
        return M.SnappyManager(chunk)
    return M.CascadedManager(opts or L.CascadedOpts(chunk, L.TYPE_INT, 1, 1, 1))


def make_data(kind, nbytes, seed):
    if kind == "cascaded":
        n = (nbytes + 3) // 4
        return datagen.sorted_ints(n, seed).view(np.uint8)[:nbytes].copy()
    if nbytes == 0:
        return np.zeros(0, np.uint8)
    base = datagen.log_lines(min(nbytes, 1 << 20), seed)
    return np.resize(base, nbytes).copy()


def parse_container(blob, kind):
    """-> (header dict, [chunk bytes]) following SURVEY A.4 / reference hlif_shared_types.hpp:68-84."""
    magic, major, minor, fmt = struct.unpack_from("<IBBB", blob, 0)
    comp_size, decomp_size, nchunks = struct.unpack_from("<QQQ", blob, 8)
    starts = blob[32]
    chunk_size, = struct.unpack_from("<Q", blob, 48)
    data_off, = struct.unpack_from("<I", blob, 56)
    hdr = dict(magic=magic, major=major, minor=minor, format=fmt, comp_data_size=comp_size, decomp_data_size=decomp_size,
               num_chunks=nchunks, include_chunk_starts=starts, uncomp_chunk_size=chunk_size, comp_data_offset=data_off)
    arr = (64 + SPEC_BYTES[kind] + 7) // 8 * 8
    offs = np.frombuffer(blob, np.uint64, nchunks, arr)
    sizes = np.frombuffer(blob, np.uint64, nchunks, arr + 8 * nchunks)
    assert data_off == arr + 24 * nchunks
    chunks = [bytes(blob[data_off + int(a): data_off + int(a) + int(s)]) for a, s in zip(offs, sizes)]
    hdr["offsets"], hdr["sizes"] = offs, sizes
    return hdr, chunks


def oracle_encode(kind, chunk, opts=None, max_chunk=None):
    if kind == "lz4":
        return o.lz4_product_compress(chunk, 1, max_chunk)
    if kind == "snappy":
        return o.snappy_product_compress(chunk, max_chunk)
    return o.casc_compress(chunk, opts.type, opts.num_RLEs, opts.num_deltas, opts.use_bp)


def oracle_decode(kind, comp, cap):
    comp = np.frombuffer(comp, np.uint8)
    if kind == "lz4":
        return o.lz4_decompress(comp, cap)
    if kind == "snappy":
        return o.snappy_decompress(comp, cap)
    return o.casc_decompress(comp, cap, 4)


def run_roundtrip(kind, chunk, nbytes, seed=7, check_oracle=True, opts=None):
    data = make_data(kind, nbytes, seed)
    mgr = make_manager(kind, chunk, opts)
    opts = opts or L.CascadedOpts(chunk, L.TYPE_INT, 1, 1, 1)
    src = torch.from_numpy(data).cuda() if nbytes else torch.empty(0, dtype=torch.uint8, device="cuda")
    cc = mgr.configure_compression(nbytes)
    assert cc.num_chunks == (nbytes + chunk - 1) // chunk
    comp = torch.full((cc.max_compressed_buffer_size + 64,), 0xCD, dtype=torch.uint8, device="cuda")
    mgr.compress(src, comp, cc)
    torch.cuda.synchronize()
    assert cc.get_status() == L.SUCCESS
    total = mgr.get_compressed_output_size(comp)
    assert total <= cc.max_compressed_buffer_size
    blob = comp[:total].cpu().numpy().tobytes()
    hdr, chunks = parse_container(blob, kind)
    assert (hdr["magic"], hdr["major"], hdr["minor"], hdr["format"]) == (0, 2, 2, FMT[kind])
    assert hdr["decomp_data_size"] == nbytes and hdr["num_chunks"] == cc.num_chunks
    assert hdr["uncomp_chunk_size"] == chunk and hdr["include_chunk_starts"] == 1
    assert hdr["comp_data_size"] + hdr["comp_data_offset"] == total
    assert bytes(comp[total:total + 64].cpu().numpy()) == b"\xCD" * 64, "wrote past the reported size"
    if check_oracle:
        for i, c in enumerate(chunks):
            raw = data[i * chunk:(i + 1) * chunk]
            assert c == oracle_encode(kind, raw, opts, chunk).tobytes(), f"chunk {i}: stream differs from the oracle encoder"
            st, dec = oracle_decode(kind, c, len(raw))
            assert st == 0 and dec.tobytes() == raw.tobytes(), f"chunk {i}: oracle decode differs"
    # decode through a manager built from the container itself
    mgr2 = M.create_manager(comp)
    dc = mgr2.configure_decompression(comp)
    assert dc.decomp_data_size == nbytes and dc.num_chunks == cc.num_chunks
    out = torch.full((nbytes + 64,), 0xEE, dtype=torch.uint8, device="cuda")
    mgr2.decompress(out, comp, dc)
    torch.cuda.synchronize()
    assert dc.get_status() == L.SUCCESS
    assert torch.equal(out[:nbytes], src)
    assert bytes(out[nbytes:].cpu().numpy()) == b"\xEE" * 64
    # and through the same manager with a config derived from the compression config
    dc2 = mgr.configure_decompression(cc)
    out.fill_(0)
    mgr.decompress(out, comp, dc2)
    torch.cuda.synchronize()
    assert dc2.get_status() == L.SUCCESS and torch.equal(out[:nbytes], src)
    return hdr


@pytest.mark.parametrize("kind", ["lz4", "snappy", "cascaded"])
@pytest.mark.parametrize("nbytes", [0, 4, 4092, 4096, 4100, 65536 * 3 + 1236])
def test_roundtrip_sizes(kind, nbytes):
    chunk = 4096 if kind == "cascaded" else 16384
    run_roundtrip(kind, chunk, nbytes)


@pytest.mark.parametrize("kind", ["lz4", "snappy", "cascaded"])
def test_multi_wave(kind):
    """Three waves (the managers cut a buffer into waves of 128 MiB, Cascaded 512 MiB): offsets continue across waves."""
    chunk = 4096 if kind == "cascaded" else 65536
    nbytes = (1100 << 20) + 4096 * 3 + 8 if kind == "cascaded" else (300 << 20) + 300_000
    hdr = run_roundtrip(kind, chunk, nbytes, check_oracle=False)
    offs, sizes = hdr["offsets"].astype(np.int64), hdr["sizes"].astype(np.int64)
    assert offs[0] == 0 and (np.diff(offs) >= sizes[:-1]).all()
    if kind != "cascaded":
        assert (np.diff(offs) == sizes[:-1]).all(), "LZ4/Snappy data region is gapless"
    assert hdr["comp_data_size"] >= offs[-1] + sizes[-1]


def test_lz4_64k_chunks_large():
    run_roundtrip("lz4", 65536, 24 * (1 << 20) + 17, check_oracle=False)


def test_cascaded_default_options():
    opts = L.CascadedOpts(4096, L.TYPE_INT, 2, 1, 1)
    run_roundtrip("cascaded", 4096, 4096 * 37 + 64, opts=opts)


def test_user_scratch_buffer():
    mgr = M.LZ4Manager(16384, L.TYPE_CHAR)
    need = mgr.get_required_scratch_buffer_size()
    assert need > 0
    scratch = torch.empty(need, dtype=torch.uint8, device="cuda")
    mgr.set_scratch_buffer(scratch)
    data = make_data("lz4", 1 << 20, 3)
    src = torch.from_numpy(data).cuda()
    cc = mgr.configure_compression(len(data))
    comp = torch.empty(cc.max_compressed_buffer_size, dtype=torch.uint8, device="cuda")
    mgr.compress(src, comp, cc)
    dc = mgr.configure_decompression(comp)
    out = torch.empty_like(src)
    mgr.decompress(out, comp, dc)
    torch.cuda.synchronize()
    assert torch.equal(out, src)


@pytest.mark.parametrize("kind", ["lz4", "snappy"])
def test_corrupt_chunk_sets_status(kind):
    """A damaged chunk must surface in the config's status word and never fault
    (reference hlif_shared.hiph:332-337)."""
    chunk, nbytes = 16384, 16384 * 8
    data = make_data(kind, nbytes, 11)
    mgr = make_manager(kind, chunk)
    src = torch.from_numpy(data).cuda()
    cc = mgr.configure_compression(nbytes)
    comp = torch.zeros(cc.max_compressed_buffer_size, dtype=torch.uint8, device="cuda")
    mgr.compress(src, comp, cc)
    torch.cuda.synchronize()
    blob = comp.cpu().numpy()
    hdr, _ = parse_container(blob.tobytes(), kind)
    start = hdr["comp_data_offset"] + int(hdr["offsets"][3])
    if kind == "lz4":
        blob[start] = 0xF0  # 15+ literals ...
        blob[start + 1:start + 40] = 0xFF  # ... with a length that runs past the chunk
    else:
        blob[start:start + 5] = [0xFF, 0xFF, 0xFF, 0xFF, 0x7F]  # absurd uncompressed length
    bad = torch.from_numpy(blob).cuda()
    dc = mgr.configure_decompression(bad)
    out = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
    mgr.decompress(out, bad, dc)
    torch.cuda.synchronize()
    assert dc.get_status() == L.ERROR_CANNOT_DECOMPRESS
    good = np.ones(8, bool)
    good[3] = False
    got = out.cpu().numpy().reshape(8, chunk)
    assert (got[good] == data.reshape(8, chunk)[good]).all(), "undamaged chunks still decode"


def test_create_manager_rejects_unknown_format():
    buf = torch.zeros(256, dtype=torch.uint8, device="cuda")
    buf[6] = 3  # GDeflate
    with pytest.raises(M.ManagerError):
        M.create_manager(buf)


@pytest.mark.gpu
@pytest.mark.parametrize("kind,chunk,nbytes", [("cascaded", 4096, 10000 * 4096 + 1236), ("lz4", 4096, 9000 * 4096 + 77)])
def test_large_wave_two_level_placement(kind, chunk, nbytes):
    """Waves of more than 8192 chunks place their chunks with the two-level scan (hlif.cu place_sum / place_kernel /
    place_scatter): offsets must be the exclusive sum of the padded sizes, gapless, and the container must round-trip."""
    raw = make_data(kind, nbytes, 99)
    mgr = make_manager(kind, chunk)
    d_raw = torch.from_numpy(raw).cuda()
    cc = mgr.configure_compression(nbytes)
    assert cc.num_chunks > 8192
    comp = torch.zeros(cc.max_compressed_buffer_size, dtype=torch.uint8, device="cuda")
    mgr.compress(d_raw, comp, cc)
    total = mgr.get_compressed_output_size(comp)
    blob = comp[:total].cpu().numpy().tobytes()
    hdr, _ = parse_container(blob, kind)
    align = 8 if kind == "cascaded" else 1
    padded = (hdr["sizes"].astype(np.int64) + align - 1) // align * align
    expect = np.cumsum(padded) - padded
    assert np.array_equal(hdr["offsets"].astype(np.int64), expect)
    assert hdr["comp_data_size"] == int(padded.sum()) and total == hdr["comp_data_offset"] + int(padded.sum())
    dc = mgr.configure_decompression(comp)
    back = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    mgr.decompress(back, comp, dc)
    torch.cuda.synchronize()
    assert dc.get_status() == 0 and torch.equal(back, d_raw)
