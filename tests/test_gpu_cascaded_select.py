"""-m gpu: the Cascaded per-partition option selector (b200compBatchedCascadedSelectOptsAsync +
b200compBatchedCascadedCompressWithOptsAsync).  The choice is restated here in numpy; every partition's stream
is the oracle's stream for the chosen options, and the ordinary decoder returns the input."""
import ctypes as C

import numpy as np
import pytest
import torch

import datagen
import oracle_lib as o

pytestmark = pytest.mark.gpu

DT = {0: np.int8, 1: np.uint8, 2: np.int16, 3: np.uint16, 4: np.int32, 5: np.uint32, 6: np.int64, 7: np.uint64}


def _g():
    import gpu_common

    return gpu_common


def range_bits(span):
    return int(span).bit_length()


def model_select(part, type_id):
    """numpy restatement of cascaded::select_kernel (csrc/cascaded.cu)."""
    w = np.dtype(DT[type_id]).itemsize
    udt = {1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64}[w]
    sdt = {1: np.int8, 2: np.int16, 4: np.int32, 8: np.int64}[w]
    v = part.view(udt)[:1024]
    m = len(v)
    if m < 2:
        return (0, 0, 0)
    d = (v[1:] - v[:-1]).view(sdt).astype(object)  # wrap-around difference, read as signed
    starts = np.concatenate([[True], v[1:] != v[:-1]])
    idx = np.flatnonzero(starts)
    runs = np.diff(np.concatenate([idx, [m]]))
    nr, maxrun = len(idx), int(runs.max())
    vb = range_bits(int(v.max()) - int(v.min()))
    db = range_bits(int(max(d)) - int(min(d)))
    nz = [x for x, s in zip(d, starts[1:]) if s]
    rb = range_bits(int(max(nz)) - int(min(nz))) if nz else 0
    lb = range_bits(maxrun - 1)
    w8 = 8 * w
    best, choice = m * w8, (0, 0, 0)
    for size, c in ((m * vb + 64, (0, 0, 1)), ((m - 1) * db + w8 + 64, (0, 1, 1)), (nr * (vb + lb) + 128, (1, 0, 1)),
                    ((nr - 1) * rb + nr * lb + w8 + 128, (1, 1, 1))):
        if size < best:
            best, choice = size, c
    return choice


def partitions(kind, type_id, nparts, part_bytes, seed):
    rng = np.random.default_rng(seed)
    dt = DT[type_id]
    n = part_bytes // np.dtype(dt).itemsize
    out = []
    for i in range(nparts):
        if kind == "sorted":
            a = datagen.sorted_ints(n, seed + i).astype(dt)
        elif kind == "runs":
            a = np.repeat(rng.integers(0, 50, n // 64 + 1), 64)[:n].astype(dt)
        elif kind == "random":
            a = rng.integers(0, 256, part_bytes, dtype=np.uint8).view(dt)
        elif kind == "small":
            a = rng.integers(0, 7, n).astype(dt)
        elif kind == "ramp":
            a = (np.arange(n) * 3 + 1000).astype(dt)
        else:  # mixed: every partition its own kind
            return sum((partitions(k, type_id, 1, part_bytes, seed + 10 * j) for j, k in
                        enumerate(["sorted", "runs", "random", "small", "ramp"] * (nparts // 5 + 1))), [])[:nparts]
        out.append(np.ascontiguousarray(a).view(np.uint8))
    return out


@pytest.mark.parametrize("type_id", [1, 3, 4, 7])
@pytest.mark.parametrize("kind", ["sorted", "runs", "random", "small", "ramp", "mixed"])
def test_selector_matches_model_and_streams_match_oracle(kind, type_id):
    g = _g()
    lib = g.L.lib()
    parts = partitions(kind, type_id, 10, 8192, 17)
    parts[-1] = parts[-1][: 8192 - 8 * 37]  # a ragged last partition
    src = g.B.DeviceBatch.from_host(parts)
    opts = torch.zeros(4 * src.n, dtype=torch.uint8, device="cuda")
    assert lib.b200compBatchedCascadedSelectOptsAsync(src.ptrs.data_ptr(), src.sizes.data_ptr(), src.n, type_id,
                                                      opts.data_ptr(), None) == 0
    chosen = opts.cpu().numpy().reshape(-1, 4)
    for i, p in enumerate(parts):
        assert tuple(int(x) for x in chosen[i][:3]) == model_select(p, type_id), f"partition {i} ({kind})"
    copts = g.L.CascadedOpts(4096, type_id, 1, 1, 1)
    stride = (g.B.max_compressed_chunk_bytes("cascaded", 8192, copts) + 15) // 16 * 16
    comp = g.B.DeviceBatch.empty(src.n, stride)
    comp_sizes = torch.zeros(src.n, dtype=torch.int64, device="cuda")
    assert lib.b200compBatchedCascadedCompressWithOptsAsync(src.ptrs.data_ptr(), src.sizes.data_ptr(), src.n,
                                                            comp.ptrs.data_ptr(), comp_sizes.data_ptr(), type_id,
                                                            opts.data_ptr(), None) == 0
    comp.sizes = comp_sizes
    got = comp.to_host()
    total_sel = 0
    for i, p in enumerate(parts):
        r, d, bp = (int(x) for x in chosen[i][:3])
        assert got[i].tobytes() == o.casc_compress(p, type_id, r, d, bp).tobytes(), f"partition {i}: stream differs from the oracle"
        total_sel += len(got[i])
    out = g.B.DeviceBatch.empty(src.n, 8192 + 16)
    out.sizes = torch.from_numpy(np.array([len(p) for p in parts], dtype=np.int64)).cuda()
    actual, status = g.B.decompress("cascaded", comp, out)
    torch.cuda.synchronize()
    assert (status.cpu().numpy() == 0).all()
    for d_, p in zip(out.to_host(actual), parts):
        assert d_.tobytes() == p.tobytes()
    # the selection is never much worse than the best fixed option set of the candidates
    fixed = min(sum(len(o.casc_compress(p, type_id, *c)) for p in parts) for c in [(0, 0, 1), (0, 1, 1), (1, 0, 1), (1, 1, 1)])
    assert total_sel <= 1.05 * fixed + 64 * len(parts)
