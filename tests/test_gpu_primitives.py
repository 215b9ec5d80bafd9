"""Device-wide RLE / delta / bit-packing primitives against CPU loops, following the reference's own
unit tests (src/test/RunLengthEncodeGPU_test.cpp:259-560, DeltaGPU_test.cpp:117-190,
BitPackGPU_test.cpp:251-560: random / run-heavy inputs, every value and count type, var-bit ranges,
results compared with a host loop and with unpackBytes)."""
import importlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

L = importlib.import_module("hipcomp-core_b200._lib")
P = importlib.import_module("hipcomp-core_b200.primitives")

DTYPES = ["int8", "uint8", "int16", "uint16", "int32", "uint32", "int64", "uint64"]


def dev(a):
    # torch has no uint16/32/64 arithmetic but can carry the bytes
    return torch.from_numpy(a.view(np.uint8).copy()).cuda()


def host(t, dtype, n=None):
    a = t.cpu().numpy().view(dtype)
    return a if n is None else a[:n]


def rle_oracle(a):
    if len(a) == 0:
        return a[:0], np.zeros(0, np.int64)
    starts = np.flatnonzero(np.concatenate([[True], a[1:] != a[:-1]]))
    counts = np.diff(np.concatenate([starts, [len(a)]]))
    return a[starts], counts


def runs_input(n, dtype, seed, max_run):
    rng = np.random.default_rng(seed)
    info = np.iinfo(dtype)
    out = np.empty(n, dtype)
    i = 0
    while i < n:
        r = int(rng.integers(1, max_run + 1))
        out[i:i + r] = rng.integers(max(info.min, -1000), min(info.max, 1000), endpoint=True)
        i += r
    return out


@pytest.mark.parametrize("vt", DTYPES)
@pytest.mark.parametrize("ct", ["uint16", "int32", "uint64"])
@pytest.mark.parametrize("downstream", [False, True])
def test_rle_types(vt, ct, downstream):
    n = 10_000
    a = runs_input(n, vt, 5, 40)
    ev, ec = rle_oracle(a)
    ws = torch.empty(P.RunLengthEncodeGPU.requiredWorkspaceSize(n, P.NP_TYPE[vt], P.NP_TYPE[ct]), dtype=torch.uint8, device="cuda")
    ov = torch.zeros(n * a.itemsize, dtype=torch.uint8, device="cuda")
    oc = torch.zeros(n * np.dtype(ct).itemsize, dtype=torch.uint8, device="cuda")
    nout = torch.zeros(1, dtype=torch.int64, device="cuda")
    d_in = dev(a)
    if downstream:
        nd = torch.tensor([n], dtype=torch.int64, device="cuda")
        keep = P.RunLengthEncodeGPU.compressDownstream(ws, P.NP_TYPE[vt], ov, P.NP_TYPE[ct], oc, nout, d_in, nd, n + 500)
    else:
        P.RunLengthEncodeGPU.compress(ws, P.NP_TYPE[vt], ov, P.NP_TYPE[ct], oc, nout, d_in, n)
    torch.cuda.synchronize()
    k = int(nout.item())
    assert k == len(ev)
    assert (host(ov, vt, k) == ev).all()
    assert (host(oc, ct, k).astype(np.int64) == ec).all()


@pytest.mark.parametrize("n,max_run", [(1, 1), (2, 1), (4095, 3), (4096, 1), (4097, 5000), (1 << 20, 1), (3_000_001, 100_000),
                                       (10_000_000, 30)])
def test_rle_sizes(n, max_run):
    """compress_10Million_Test and the tile-boundary cases (runs longer than a tile, all-distinct)."""
    a = runs_input(n, "int32", n, max_run) if max_run > 1 else np.arange(n, dtype=np.int32)
    ev, ec = rle_oracle(a)
    ws = torch.empty(P.RunLengthEncodeGPU.requiredWorkspaceSize(n, L.TYPE_INT, L.TYPE_UINT), dtype=torch.uint8, device="cuda")
    ov = torch.zeros(n, dtype=torch.int32, device="cuda")
    oc = torch.zeros(n, dtype=torch.int32, device="cuda")
    nout = torch.zeros(1, dtype=torch.int64, device="cuda")
    P.RunLengthEncodeGPU.compress(ws, L.TYPE_INT, ov, L.TYPE_UINT, oc, nout, dev(a), n)
    torch.cuda.synchronize()
    k = int(nout.item())
    assert k == len(ev)
    assert (ov[:k].cpu().numpy() == ev).all() and (oc[:k].cpu().numpy() == ec).all()
    assert int(oc[:k].sum().item()) == n


def test_rle_workspace_too_small():
    ws = torch.empty(8, dtype=torch.uint8, device="cuda")
    t = torch.zeros(100_000, dtype=torch.int32, device="cuda")
    nout = torch.zeros(1, dtype=torch.int64, device="cuda")
    with pytest.raises(P.PrimitiveError):
        P.RunLengthEncodeGPU.compress(ws, L.TYPE_INT, t, L.TYPE_INT, t, nout, t, 100_000)


@pytest.mark.parametrize("vt", DTYPES)
@pytest.mark.parametrize("n", [1, 7, 10_000, 1_000_003])
def test_delta(vt, n):
    """compress_10Thousand_Test: out[0] = in[0], out[i] = in[i] - in[i-1] (wrapping)."""
    rng = np.random.default_rng(n)
    info = np.iinfo(vt)
    a = rng.integers(info.min, info.max, n, dtype=vt, endpoint=True)
    expect = a.copy()
    expect[1:] = a[1:] - a[:-1]
    out = torch.zeros(n * a.itemsize + 16, dtype=torch.uint8, device="cuda")
    nd = torch.tensor([n], dtype=torch.int64, device="cuda")
    keep = P.DeltaGPU.compress(P.NP_TYPE[vt], out, dev(a), nd, n + 100)
    torch.cuda.synchronize()
    assert (host(out, vt)[:n] == expect).all()
    # an unaligned view takes the scalar path
    if n > 16:
        src = dev(a)
        off = a.itemsize
        out2 = torch.zeros_like(out)
        nd2 = torch.tensor([n - 1], dtype=torch.int64, device="cuda")
        keep2 = P.DeltaGPU.compress(P.NP_TYPE[vt], out2[off:], src[off:], nd2, n)
        torch.cuda.synchronize()
        e2 = a[1:].copy()
        e2[1:] = a[2:] - a[1:-1]
        assert (host(out2[off:off + (n - 1) * a.itemsize], vt) == e2).all()


def unpack_all(words_bytes, nbits, minv, n, dtype):
    """numpy restatement of unpackBytes (reference src/unpack.h:68-112) over a whole stream."""
    if nbits == 0:
        return np.full(n, minv, dtype)
    bits = np.unpackbits(words_bytes, bitorder="little")
    need = n * nbits
    vals = bits[:need].reshape(n, nbits).astype(np.uint64)
    weights = (np.uint64(1) << np.arange(nbits, dtype=np.uint64))
    u = (vals * weights).sum(axis=1, dtype=np.uint64)
    udt = np.dtype(dtype).str.replace("i", "u")
    return (u.astype(udt) + np.array(minv, dtype).astype(udt)).astype(udt).view(dtype)


def pack_oracle(a):
    """CPU loop of the reference's test: min, numBits = bits(max - min), LSB-first packing."""
    dtype = a.dtype
    udt = np.dtype(dtype.str.replace("i", "u"))
    mn, mx = a.min(), a.max()
    rng_ = int(mx) - int(mn)
    nbits = rng_.bit_length()
    diff = (a.astype(udt) - np.array(mn, dtype).astype(udt)).astype(np.uint64)
    if nbits == 0:
        return mn, 0, np.zeros(0, np.uint8)
    bits = ((diff[:, None] >> np.arange(nbits, dtype=np.uint64)) & np.uint64(1)).astype(np.uint8).reshape(-1)
    word_bits = 64 if a.itemsize == 8 else 32
    total = (len(bits) + word_bits - 1) // word_bits * word_bits
    bits = np.concatenate([bits, np.zeros(total - len(bits), np.uint8)])
    return mn, nbits, np.packbits(bits, bitorder="little")


def run_bitpack(a, slack=0):
    n = len(a)
    vt = a.dtype.name
    ws = torch.empty(P.BitPackGPU.requiredWorkspaceSize(n + slack, P.NP_TYPE[vt]), dtype=torch.uint8, device="cuda")
    out = torch.full((n * a.itemsize + 64,), 0xAA, dtype=torch.uint8, device="cuda")
    mn = torch.zeros(8, dtype=torch.uint8, device="cuda")
    nb = torch.full((8,), 0x77, dtype=torch.uint8, device="cuda")
    nd = torch.tensor([n], dtype=torch.int64, device="cuda")
    keep = P.BitPackGPU.compress(ws, P.NP_TYPE[vt], out, dev(a), nd, n + slack, mn, nb)
    torch.cuda.synchronize()
    return host(mn, vt)[0], int(nb[0].item()), out.cpu().numpy()


@pytest.mark.parametrize("vt", DTYPES)
@pytest.mark.parametrize("bits", [0, 1, 3, 7, 8, 13, 16, 24, 31, 32, 33, 47, 63, 64])
def test_bitpack_var_bits(vt, bits):
    """compress{Int16,Uint32,Int64}VarBitTest / compressType*RangeTest: every width a type allows."""
    width = np.dtype(vt).itemsize * 8
    if bits > width:
        pytest.skip("wider than the type")
    n = 10_000 + bits
    rng = np.random.default_rng(bits * 8 + width)
    info = np.iinfo(vt)
    span = (1 << bits) - 1
    lo = info.min if bits == width else info.min + int(rng.integers(0, 1 << 62)) % (info.max - span - info.min + 1)
    r = rng.integers(0, span, n, dtype=np.uint64, endpoint=True)
    r[0], r[-1] = 0, span  # the range is exactly `bits` wide
    a = (r.astype(np.dtype(vt).str.replace("i", "u")) + np.array(lo, vt).astype(np.dtype(vt).str.replace("i", "u"))).view(vt) \
        if np.dtype(vt).kind == "i" else (r + np.uint64(lo)).astype(vt)
    emn, enb, ebytes = pack_oracle(a)
    assert enb == bits
    mn, nb, out = run_bitpack(a, slack=77)
    assert mn == emn and nb == bits
    assert (out[:len(ebytes)] == ebytes).all(), "packed words differ from the CPU loop"
    assert (out[len(ebytes):] == 0xAA).all(), "wrote past the last word"
    assert (unpack_all(out[:len(ebytes)], nb, mn, n, vt) == a).all()


@pytest.mark.parametrize("n", [1, 31, 32, 33, 4095, 4096, 4097, 1_000_000, 5_000_011])
def test_bitpack_sizes(n):
    """compressInt32VarSizeTest."""
    rng = np.random.default_rng(n)
    a = rng.integers(-5000, 70000, n, dtype=np.int32)
    emn, enb, ebytes = pack_oracle(a)
    mn, nb, out = run_bitpack(a)
    assert mn == emn and nb == enb
    assert (out[:len(ebytes)] == ebytes).all()
    assert (out[len(ebytes):] == 0xAA).all()


def test_bitpack_workspace_reuse():
    """The fold's ticket is reset by the kernel: the same workspace serves back-to-back calls."""
    a = np.arange(100_000, dtype=np.int32) * 3 - 1000
    n = len(a)
    ws = torch.empty(P.BitPackGPU.requiredWorkspaceSize(n, L.TYPE_INT), dtype=torch.uint8, device="cuda")
    for rep in range(3):
        out = torch.zeros(n * 4, dtype=torch.uint8, device="cuda")
        mn = torch.zeros(2, dtype=torch.int32, device="cuda")
        nb = torch.zeros(8, dtype=torch.uint8, device="cuda")
        nd = torch.tensor([n - rep], dtype=torch.int64, device="cuda")
        keep = P.BitPackGPU.compress(ws, L.TYPE_INT, out, dev(a), nd, n, mn, nb)
        torch.cuda.synchronize()
        emn, enb, ebytes = pack_oracle(a[:n - rep])
        assert int(mn[0].item()) == emn and int(nb[0].item()) == enb
        assert (out.cpu().numpy()[:len(ebytes)] == ebytes).all()
