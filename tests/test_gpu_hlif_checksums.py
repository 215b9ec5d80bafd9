"""-m gpu: the container's per-chunk checksum arrays (extension b200compManagerSetChecksumMode).
The function is zlib's CRC-32, so the expected values come from Python's zlib."""
import importlib
import struct
import zlib

import numpy as np
import pytest
import torch

from test_gpu_hlif import FMT, SPEC_BYTES, make_data, make_manager, parse_container

pytestmark = pytest.mark.gpu

L = importlib.import_module("hipcomp-core_b200._lib")
M = importlib.import_module("hipcomp-core_b200.managers")


def checksum_arrays(blob, kind):
    hdr, chunks = parse_container(blob, kind)
    n = hdr["num_chunks"]
    arr = (64 + SPEC_BYTES[kind] + 7) // 8 * 8 + 16 * n
    comp = np.frombuffer(blob, np.uint32, n, arr)
    decomp = np.frombuffer(blob, np.uint32, n, arr + 4 * n)
    flags = (blob[44], blob[45])  # include_per_chunk_{comp,decomp}_buffer_checksums
    return hdr, chunks, comp, decomp, flags


def compress_with(mgr, data):
    src = torch.from_numpy(data).cuda() if len(data) else torch.empty(0, dtype=torch.uint8, device="cuda")
    cc = mgr.configure_compression(len(data))
    comp = torch.zeros(cc.max_compressed_buffer_size + 64, dtype=torch.uint8, device="cuda")
    mgr.compress(src, comp, cc)
    torch.cuda.synchronize()
    assert cc.get_status() == L.SUCCESS
    return src, comp, mgr.get_compressed_output_size(comp)


def decode_status(mgr, comp, nbytes, want=None):
    dc = mgr.configure_decompression(comp)
    out = torch.zeros(nbytes + 16, dtype=torch.uint8, device="cuda")
    mgr.decompress(out, comp, dc)
    torch.cuda.synchronize()
    if want is not None:
        assert torch.equal(out[:nbytes], want)
    return dc.get_status()


@pytest.mark.parametrize("kind,chunk,nbytes", [("lz4", 65536, 65536 * 5 + 1237), ("lz4", 4096, 4096 * 40 + 1),
                                               ("snappy", 16384, 16384 * 9 + 77), ("cascaded", 4096, 4096 * 33 + 8),
                                               ("lz4", 65536, 100), ("snappy", 65536, 0)])
def test_checksums_are_zlib_crc32_of_every_chunk(kind, chunk, nbytes):
    data = make_data(kind, nbytes, 5)
    mgr = make_manager(kind, chunk)
    mgr.set_checksum_mode(1)
    src, comp, total = compress_with(mgr, data)
    blob = comp[:total].cpu().numpy().tobytes()
    hdr, chunks, ccrc, dcrc, flags = checksum_arrays(blob, kind)
    assert flags == (1, 1)
    for i, c in enumerate(chunks):
        raw = data[i * chunk:(i + 1) * chunk].tobytes()
        assert ccrc[i] == zlib.crc32(c), f"chunk {i}: compressed-bytes checksum"
        assert dcrc[i] == zlib.crc32(raw), f"chunk {i}: uncompressed-bytes checksum"
    # verified on the way back, by the same manager and by one built from the container
    assert decode_status(mgr, comp, nbytes, src) == L.SUCCESS
    m2 = M.create_manager(comp)
    m2.set_checksum_mode(1)
    assert decode_status(m2, comp, nbytes, src) == L.SUCCESS
    # a manager that does not verify decodes it as well
    assert decode_status(M.create_manager(comp), comp, nbytes, src) == L.SUCCESS


def test_default_mode_leaves_arrays_zero_and_is_accepted_by_a_verifying_manager():
    kind, chunk, nbytes = "lz4", 16384, 16384 * 6
    data = make_data(kind, nbytes, 6)
    mgr = make_manager(kind, chunk)
    src, comp, total = compress_with(mgr, data)
    hdr, chunks, ccrc, dcrc, flags = checksum_arrays(comp[:total].cpu().numpy().tobytes(), kind)
    assert flags == (0, 0) and not ccrc.any() and not dcrc.any()
    verifying = make_manager(kind, chunk)
    verifying.set_checksum_mode(1)
    assert decode_status(verifying, comp, nbytes, src) == L.SUCCESS


@pytest.mark.parametrize("which", ["comp_checksum", "decomp_checksum", "literal_byte"])
def test_mismatch_is_reported_only_when_verifying(which):
    kind, chunk, nbytes = "lz4", 16384, 16384 * 8
    data = make_data(kind, nbytes, 8)
    mgr = make_manager(kind, chunk)
    mgr.set_checksum_mode(1)
    src, comp, total = compress_with(mgr, data)
    blob = comp.cpu().numpy().copy()
    hdr, chunks, ccrc, dcrc, flags = checksum_arrays(blob[:total].tobytes(), kind)
    n = hdr["num_chunks"]
    arr = (64 + SPEC_BYTES[kind] + 7) // 8 * 8 + 16 * n
    if which == "comp_checksum":
        blob[arr + 4 * 2] ^= 1
    elif which == "decomp_checksum":
        blob[arr + 4 * n + 4 * 5] ^= 0x80
    else:
        # a literal byte of chunk 3 (first sequence: token, [length bytes], literals): the stream still decodes,
        # to other bytes - only the checksums can tell
        start = hdr["comp_data_offset"] + int(hdr["offsets"][3])
        tok = blob[start]
        assert tok >> 4 >= 1, "the chunk starts with literals"
        pos = start + 1
        if tok >> 4 == 15:
            while blob[pos] == 255:
                pos += 1
            pos += 1
        blob[pos] ^= 0x20
    bad = torch.from_numpy(blob).cuda()
    assert decode_status(mgr, bad, nbytes) == L.ERROR_CANNOT_DECOMPRESS
    plain = make_manager(kind, chunk)
    st = decode_status(plain, bad, nbytes)
    assert st == L.SUCCESS, "without verification the damaged container decodes (to wrong bytes for literal_byte)"


def test_mode_argument_is_checked():
    mgr = make_manager("lz4", 4096)
    with pytest.raises(M.ManagerError):
        mgr.set_checksum_mode(7)
    mgr.set_checksum_mode(0)
