"""-m gpu: the host-buffer staging pipeline (b200compStager*): streams in the host blob are the
oracle's streams, both directions and the fused round trip return the input bit-exactly."""
import importlib

import numpy as np
import pytest

import datagen
import oracle_lib as o

pytestmark = pytest.mark.gpu


def _mods():
    return importlib.import_module("hipcomp-core_b200.staging"), importlib.import_module("hipcomp-core_b200._lib")


@pytest.mark.parametrize("codec,chunk,nbytes,slices", [("lz4", 65536, 3 * 65536 + 1234, 4), ("lz4", 4096, 1 << 20, 8),
                                                       ("snappy", 65536, (1 << 20) + 17, 3), ("cascaded", 65536, 1 << 20, 5),
                                                       ("lz4", 65536, 100, 8)])
def test_stager_directions_and_roundtrip(codec, chunk, nbytes, slices):
    S, L = _mods()
    opts = L.CascadedOpts(4096, L.TYPE_INT, 1, 1, 1) if codec == "cascaded" else None
    if codec == "cascaded":
        data = datagen.sorted_ints(nbytes // 4, 5).view(np.uint8)
    else:
        data = np.concatenate(datagen.mixed_chunks(65536, (nbytes + 65535) // 65536, 9))[:nbytes]
    st = S.Stager(codec, chunk, 2 * nbytes + 4096, opts=opts, slices=slices)
    n = st.nchunks(nbytes)
    raw, out = S.PinnedArray(nbytes), S.PinnedArray(nbytes)
    raw.array[:] = data
    bound = {"lz4": o.lz4_max_compressed, "snappy": o.snappy_max_compressed, "cascaded": lambda c: c + 64}[codec](chunk)
    blob, sizes, stat = S.PinnedArray(n * (bound + 16)), S.PinnedArray(n, np.uint64), S.PinnedArray(n, np.int32)
    total = st.compress(raw, blob, sizes)
    assert total == int(sizes.array.sum())
    offs = np.concatenate([[0], np.cumsum(sizes.array)]).astype(np.int64)
    for i in range(n):
        c = data[i * chunk:(i + 1) * chunk]
        got = blob.array[offs[i]:offs[i + 1]]
        if codec == "lz4":
            exp = o.lz4_product_compress(c, 1, chunk)
        elif codec == "snappy":
            exp = o.snappy_product_compress(c, chunk)
        else:
            exp = o.casc_compress(c, L.TYPE_INT, 1, 1, 1)
        assert got.tobytes() == exp.tobytes(), f"chunk {i}: blob bytes differ from the oracle stream"
    stat.array[:] = -1
    st.decompress(blob, sizes, out, stat)
    assert (stat.array == 0).all() and out.array.tobytes() == data.tobytes()
    out.array[:] = 0
    blob.array[:total] = 0
    stat.array[:] = -1
    assert st.roundtrip(raw, blob, sizes, out, stat) == total
    assert (stat.array == 0).all() and out.array.tobytes() == data.tobytes()
    st.close()


def test_stager_rejects_oversize_and_bad_arguments():
    S, L = _mods()
    st = S.Stager("lz4", 65536, 1 << 20)
    raw = S.PinnedArray((1 << 20) + 1)
    blob, sizes = S.PinnedArray(1 << 21), S.PinnedArray(32, np.uint64)
    with pytest.raises(RuntimeError):
        st.compress(raw, blob, sizes)
    with pytest.raises(RuntimeError):
        S.Stager("lz4", 0, 1 << 20)
    st.close()
