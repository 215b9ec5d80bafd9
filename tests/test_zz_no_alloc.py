"""b200compBatchedLZ4CompressNoAllocAsync (include/hipcomp/b200_ext.h): the LZ4 compress entry point that takes
no workspace and allocates nothing - its hash tables for chunks above 64 KiB sit in a static device array inside
the library.  Checked against hipcompBatchedLZ4CompressAsync (byte-identical streams: same kernel, same policy)
on calls in a row, since every call shares the array and finds the previous call's tables in it.

Named test_zz_* so that pytest collects it last; it must pass alone in a fresh process and after the other tests.
"""
import ctypes as C

import numpy as np
import pytest
import torch

import datagen

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 16 << 20  # B200COMP_LZ4_NOALLOC_SCRATCH_BYTES
MAX_BATCH = 512           # B200COMP_LZ4_NOALLOC_MAX_BATCH (chunks above 64 KiB)


def _mods():
    import gpu_common as g

    return g.B, g.L


def _noalloc_compress(src, max_chunk, opts=None, stream=None):
    """Same outputs as B.compress('lz4', ...), through the entry point without a workspace."""
    B, L = _mods()
    opts = opts or B.default_opts("lz4")
    stride = (B.max_compressed_chunk_bytes("lz4", max_chunk, opts) + 15) // 16 * 16
    out = B.DeviceBatch.empty(src.n, stride, src.buf.device)
    out.buf.fill_(0xCD)
    comp_sizes = torch.zeros(src.n, dtype=torch.int64, device=src.buf.device)
    st = L.lib().b200compBatchedLZ4CompressNoAllocAsync(src.ptrs.data_ptr(), src.sizes.data_ptr(), max_chunk, src.n,
                                                         out.ptrs.data_ptr(), comp_sizes.data_ptr(), opts,
                                                         B._stream_ptr(stream))
    out.sizes = comp_sizes
    return st, out


def _ragged(sizes, seed):
    mixed = datagen.mixed_chunks(max(sizes), len(sizes), seed)
    return [c[:n].copy() for c, n in zip(mixed, sizes)]


def _assert_same_streams(a, b):
    ha, hb = a.to_host(), b.to_host()
    assert a.sizes.cpu().tolist() == b.sizes.cpu().tolist()
    for i, (x, y) in enumerate(zip(ha, hb)):
        assert x.tobytes() == y.tobytes(), f"chunk {i}: stream differs from hipcompBatchedLZ4CompressAsync"


def test_noalloc_matches_workspace_entry_point_two_calls_in_a_row():
    B, L = _mods()
    max_chunk = 1 << 18
    batches = [
        _ragged([1 << 18, 200000, 65537, 131072, 1 << 18, 99999, 1, 180000, 1 << 18, 70001, 150000], 11),
        # different data, different batch size: every warp finds the first call's tables in the array
        _ragged([131072, 1 << 18, 250000, 66000, 1 << 18, 123457, 1 << 18], 29),
    ]
    batches.append(batches[0])  # and the first batch again, behind the second
    for chunks in batches:
        src = B.DeviceBatch.from_host(chunks)
        want = B.compress("lz4", src, max_chunk=max_chunk)
        st, got = _noalloc_compress(src, max_chunk)
        assert st == L.SUCCESS
        torch.cuda.synchronize()
        _assert_same_streams(got, want)
        # and the streams decode
        out = B.DeviceBatch.empty(src.n, max_chunk + 16)
        out.sizes = src.sizes.clone()
        actual, status = B.decompress("lz4", got, out)
        torch.cuda.synchronize()
        assert (status.cpu().numpy() == L.SUCCESS).all()
        for d, c in zip(out.to_host(actual), chunks):
            assert d.tobytes() == c.tobytes()


def test_noalloc_back_to_back_on_one_stream_without_sync():
    """Two calls queued behind each other on one stream (no host sync between them) do not disturb each other."""
    B, L = _mods()
    max_chunk = 1 << 17
    a = B.DeviceBatch.from_host(_ragged([1 << 17] * 40, 5))
    b = B.DeviceBatch.from_host(_ragged([100000] * 33, 6))
    sa, ga = _noalloc_compress(a, max_chunk)
    sb, gb = _noalloc_compress(b, max_chunk)
    assert sa == L.SUCCESS and sb == L.SUCCESS
    torch.cuda.synchronize()
    _assert_same_streams(ga, B.compress("lz4", a, max_chunk=max_chunk))
    _assert_same_streams(gb, B.compress("lz4", b, max_chunk=max_chunk))


def test_noalloc_full_batch_typed_and_no_device_allocation():
    """The largest batch the scratch holds (512 tables), 4-byte match granularity; device memory in use is the
    same before and after the call (the first call above / below already loaded the module)."""
    B, L = _mods()
    max_chunk = 65536 + 4096
    rng = np.random.default_rng(77)
    base = datagen.sorted_ints(MAX_BATCH * (max_chunk // 4), 3).view(np.uint8)
    chunks = [base[i * max_chunk:(i + 1) * max_chunk] for i in range(MAX_BATCH)]
    chunks[7] = rng.integers(0, 256, max_chunk, dtype=np.uint8)
    src = B.DeviceBatch.from_host(chunks)
    opts = L.LZ4Opts(L.TYPE_INT)
    want = B.compress("lz4", src, opts=opts, max_chunk=max_chunk)
    st, got = _noalloc_compress(src, max_chunk, opts)  # warm: module and symbol address are set up by now
    assert st == L.SUCCESS
    torch.cuda.synchronize()
    _assert_same_streams(got, want)
    got.buf.fill_(0)
    got.sizes.zero_()
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    st = L.lib().b200compBatchedLZ4CompressNoAllocAsync(src.ptrs.data_ptr(), src.sizes.data_ptr(), max_chunk, src.n,
                                                         got.ptrs.data_ptr(), got.sizes.data_ptr(), opts,
                                                         B._stream_ptr(None))
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert st == L.SUCCESS
    assert free_after >= free_before, "the call allocated device memory"
    _assert_same_streams(got, want)


def test_noalloc_small_chunks_take_any_batch_size():
    """Up to 64 KiB the encoder keeps its tables in shared memory: no scratch, no batch limit."""
    B, L = _mods()
    chunks = datagen.mixed_chunks(16384, 1500, 123)
    src = B.DeviceBatch.from_host(chunks)
    st, got = _noalloc_compress(src, 16384)
    assert st == L.SUCCESS
    torch.cuda.synchronize()
    _assert_same_streams(got, B.compress("lz4", src, max_chunk=16384))


def test_noalloc_rejects_requests_beyond_the_scratch():
    B, L = _mods()
    lib = L.lib()
    need = B.compress_temp_bytes("lz4", MAX_BATCH, 1 << 17)
    assert need == SCRATCH_BYTES
    assert B.compress_temp_bytes("lz4", MAX_BATCH + 1, 1 << 17) > SCRATCH_BYTES
    dummy = torch.zeros(MAX_BATCH + 1, dtype=torch.int64, device="cuda")
    args = (dummy.data_ptr(), dummy.data_ptr())
    opts = B.default_opts("lz4")
    st = lib.b200compBatchedLZ4CompressNoAllocAsync(*args, 1 << 17, MAX_BATCH + 1, dummy.data_ptr(), dummy.data_ptr(),
                                                    opts, C.c_void_p(0))
    assert st == L.ERROR_INVALID_VALUE
    # chunk size beyond the format's limit: the error of the existing entry point
    st = lib.b200compBatchedLZ4CompressNoAllocAsync(*args, (1 << 24) + 1, 1, dummy.data_ptr(), dummy.data_ptr(),
                                                    opts, C.c_void_p(0))
    assert st == L.ERROR_INVALID_VALUE
    # an empty batch succeeds without touching anything
    st = lib.b200compBatchedLZ4CompressNoAllocAsync(None, None, 1 << 17, 0, None, None, opts, C.c_void_p(0))
    assert st == L.SUCCESS
