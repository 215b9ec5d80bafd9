"""-m gpu: HLIF containers cross-decoded with the reference's own managers, the Workspace unit
test (row a20) and untrusted container headers.  The C++ driver tests/cpp/hlif_xcodec.cpp is linked once against
this library and once against oracle/_ref/libhipcomp_ref.so by oracle/build_ref_tests.sh."""
import os
import struct
import subprocess

import numpy as np
import pytest

import datagen

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "tests")
OURS, REF = os.path.join(BIN, "hlif_xcodec_ours"), os.path.join(BIN, "hlif_xcodec_ref")


def run(exe, *args, ok=True):
    out = subprocess.run([exe, *map(str, args)], capture_output=True, text=True, timeout=600)
    if ok:
        assert out.returncode == 0, (out.stdout + out.stderr)[-2000:]
    return out


def payload(codec, n):
    if codec == "cascaded":
        return datagen.sorted_ints(n // 4, 41).view(np.uint8).tobytes()
    return np.concatenate(datagen.mixed_chunks(65536, (n + 65535) // 65536, 42))[:n].tobytes()


def test_workspace_reserve_align_exhaust_and_private_pool():
    if not os.path.exists(OURS):
        pytest.skip("driver not built")
    out = run(OURS, "workspace")
    assert "workspace: ok" in out.stdout


@pytest.mark.parametrize("codec,chunk,n", [("lz4", 65536, 5 * 65536 + 777), ("lz4", 4096, 100000), ("cascaded", 4096, 1 << 20),
                                           ("cascaded", 4096, 4096 * 3 + 8)])
def test_containers_cross_decode_with_the_reference_managers(tmp_path, codec, chunk, n):
    if not (os.path.exists(OURS) and os.path.exists(REF)):
        pytest.skip("drivers not built")
    raw = payload(codec, n)
    (tmp_path / "raw").write_bytes(raw)
    run(OURS, "compress", codec, chunk, tmp_path / "raw", tmp_path / "ours.hlif", 0)
    run(REF, "compress", codec, chunk, tmp_path / "raw", tmp_path / "ref.hlif", 0)
    run(REF, "decompress", codec, chunk, tmp_path / "ours.hlif", tmp_path / "ours_by_ref", 0)
    run(OURS, "decompress", codec, chunk, tmp_path / "ref.hlif", tmp_path / "ref_by_ours", 0)
    run(OURS, "decompress", codec, chunk, tmp_path / "ours.hlif", tmp_path / "ours_by_ours", 0)
    assert (tmp_path / "ours_by_ref").read_bytes() == raw, "the reference's manager does not decode our container"
    assert (tmp_path / "ref_by_ours").read_bytes() == raw, "our manager does not decode the reference's container"
    assert (tmp_path / "ours_by_ours").read_bytes() == raw
    ours, ref = (tmp_path / "ours.hlif").stat().st_size, (tmp_path / "ref.hlif").stat().st_size
    assert ours <= 1.01 * ref + 64 * ((n + chunk - 1) // chunk), f"container {ours} B vs reference {ref} B"


@pytest.mark.parametrize("field", ["num_chunks", "uncomp_chunk_size", "offset", "size", "decomp_size"])
def test_untrusted_container_header_is_rejected_without_faulting(tmp_path, field):
    """CommonHeader (64 B): comp_data_size @8, decomp_data_size @16, num_chunks @24, uncomp_chunk_size @48; then the
    format header, offsets[n], sizes[n].  Inconsistent values give a CannotDecompress status, never a wild write."""
    if not os.path.exists(OURS):
        pytest.skip("driver not built")
    chunk, n = 65536, 4 * 65536
    (tmp_path / "raw").write_bytes(payload("lz4", n))
    run(OURS, "compress", "lz4", chunk, tmp_path / "raw", tmp_path / "c", 0)
    blob = bytearray((tmp_path / "c").read_bytes())
    nchunks = struct.unpack_from("<Q", blob, 24)[0]
    assert nchunks == 4
    arr = (64 + 4 + 7) // 8 * 8  # LZ4 format header is one 4-byte word
    if field == "num_chunks":
        struct.pack_into("<Q", blob, 24, 1 << 20)
    elif field == "uncomp_chunk_size":
        struct.pack_into("<Q", blob, 48, 4096)
    elif field == "decomp_size":
        struct.pack_into("<Q", blob, 16, 1 << 34)
    elif field == "offset":
        struct.pack_into("<Q", blob, arr + 8, 1 << 40)
    else:
        struct.pack_into("<Q", blob, arr + 8 * nchunks + 16, 1 << 40)
    (tmp_path / "bad").write_bytes(bytes(blob))
    out = run(OURS, "decompress", "lz4", chunk, tmp_path / "bad", tmp_path / "o", 0, ok=False)
    assert out.returncode == 3, f"expected a CannotDecompress status, got rc {out.returncode}: {out.stderr[-500:]}"
