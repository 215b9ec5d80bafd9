"""Host-side harness over the C ABI: device buffers (torch), pointer/size arrays,
and one call per codec direction.  Mirrors how the reference's C tests drive the
low-level interface (tests/test_batch_c_api.h:225-503)."""
import ctypes as C

import numpy as np
import torch

from . import _lib as L

CODECS = ("lz4", "snappy", "cascaded")
_NAME = {"lz4": "LZ4", "snappy": "Snappy", "cascaded": "Cascaded"}


class CodecError(RuntimeError):
    def __init__(self, fn, status):
        super().__init__(f"{fn} returned hipcompStatus_t {status}")
        self.status = status


def _fn(codec, what):
    return getattr(L.lib(), f"hipcompBatched{_NAME[codec]}{what}")


def default_opts(codec):
    if codec == "lz4":
        return L.LZ4Opts(L.TYPE_CHAR)
    if codec == "snappy":
        return L.SnappyOpts(0)
    return L.CascadedOpts(4096, L.TYPE_INT, 2, 1, 1)


def _check(status, fn):
    if status != L.SUCCESS:
        raise CodecError(fn, status)


def _stream_ptr(stream):
    if stream is None:
        stream = torch.cuda.current_stream()
    return C.c_void_p(stream.cuda_stream)


def max_compressed_chunk_bytes(codec, max_chunk, opts=None):
    out = C.c_size_t(0)
    _check(_fn(codec, "CompressGetMaxOutputChunkSize")(max_chunk, opts or default_opts(codec), C.byref(out)),
           "CompressGetMaxOutputChunkSize")
    return out.value


def compress_temp_bytes(codec, batch, max_chunk, opts=None):
    out = C.c_size_t(0)
    _check(_fn(codec, "CompressGetTempSize")(batch, max_chunk, opts or default_opts(codec), C.byref(out)),
           "CompressGetTempSize")
    return out.value


def decompress_temp_bytes(codec, batch, max_chunk):
    out = C.c_size_t(0)
    _check(_fn(codec, "DecompressGetTempSize")(batch, max_chunk, C.byref(out)), "DecompressGetTempSize")
    return out.value


class DeviceBatch:
    """A batch of chunks living in one device buffer, with device pointer/size arrays."""

    def __init__(self, buf, offsets, sizes, device=None):
        self.buf = buf  # torch.uint8, device
        dev = buf.device
        self.offsets = np.asarray(offsets, dtype=np.int64)
        self.n = len(self.offsets)
        self.ptrs = torch.from_numpy(self.offsets + buf.data_ptr()).to(dev)
        self.sizes = sizes if isinstance(sizes, torch.Tensor) else torch.from_numpy(
            np.asarray(sizes, dtype=np.int64)).to(dev)

    @staticmethod
    def from_host(chunks, device="cuda", align=16):
        sizes = np.array([len(c) for c in chunks], dtype=np.int64)
        strides = (sizes + align - 1) // align * align
        offsets = np.concatenate([[0], np.cumsum(strides)[:-1]]) if len(chunks) else np.zeros(0, np.int64)
        host = np.zeros(int(strides.sum()) + align, dtype=np.uint8)
        for c, o in zip(chunks, offsets):
            host[o : o + len(c)] = np.frombuffer(bytes(c), np.uint8) if not isinstance(c, np.ndarray) else c.view(np.uint8).reshape(-1)
        return DeviceBatch(torch.from_numpy(host).to(device), offsets, sizes)

    @staticmethod
    def empty(n, stride, device="cuda"):
        buf = torch.empty(n * stride + 16, dtype=torch.uint8, device=device)
        return DeviceBatch(buf, np.arange(n, dtype=np.int64) * stride, np.full(n, stride, np.int64))

    def to_host(self, sizes=None):
        sizes = (self.sizes if sizes is None else sizes)
        sizes = sizes.cpu().numpy() if isinstance(sizes, torch.Tensor) else np.asarray(sizes)
        host = self.buf.cpu().numpy()
        return [host[o : o + s].copy() for o, s in zip(self.offsets, sizes)]


def compress(codec, src, opts=None, max_chunk=None, out=None, temp=None, stream=None):
    """src: DeviceBatch of uncompressed chunks.  Returns DeviceBatch whose .sizes are the compressed sizes (device)."""
    opts = opts or default_opts(codec)
    if max_chunk is None:
        max_chunk = int(src.sizes.max().item()) if src.n else 0
    if out is None:
        stride = (max_compressed_chunk_bytes(codec, max_chunk, opts) + 15) // 16 * 16
        out = DeviceBatch.empty(src.n, stride, src.buf.device)
    if temp is None:
        tb = compress_temp_bytes(codec, src.n, max_chunk, opts)
        temp = torch.empty(max(tb, 1), dtype=torch.uint8, device=src.buf.device)
    comp_sizes = torch.zeros(src.n, dtype=torch.int64, device=src.buf.device)
    st = _fn(codec, "CompressAsync")(src.ptrs.data_ptr(), src.sizes.data_ptr(), max_chunk, src.n, temp.data_ptr(),
                                     temp.numel(), out.ptrs.data_ptr(), comp_sizes.data_ptr(), opts,
                                     _stream_ptr(stream))
    _check(st, f"hipcompBatched{_NAME[codec]}CompressAsync")
    out.sizes = comp_sizes
    out._keepalive = (temp, src)
    return out


def decompress_sizes(codec, comp, stream=None):
    sizes = torch.full((comp.n,), -1, dtype=torch.int64, device=comp.buf.device)
    st = _fn(codec, "GetDecompressSizeAsync")(comp.ptrs.data_ptr(), comp.sizes.data_ptr(), sizes.data_ptr(), comp.n,
                                              _stream_ptr(stream))
    _check(st, f"hipcompBatched{_NAME[codec]}GetDecompressSizeAsync")
    return sizes


def decompress(codec, comp, out, want_actual=True, want_status=True, stream=None):
    """comp: DeviceBatch (compressed), out: DeviceBatch whose .sizes are the capacities."""
    dev = comp.buf.device
    actual = torch.full((comp.n,), -1, dtype=torch.int64, device=dev) if want_actual else None
    status = torch.full((comp.n,), -1, dtype=torch.int32, device=dev) if want_status else None
    tb = decompress_temp_bytes(codec, comp.n, int(out.sizes.max().item()) if comp.n else 0)
    temp = torch.empty(max(tb, 1), dtype=torch.uint8, device=dev)
    st = _fn(codec, "DecompressAsync")(comp.ptrs.data_ptr(), comp.sizes.data_ptr(), out.sizes.data_ptr(),
                                       actual.data_ptr() if want_actual else None, comp.n, temp.data_ptr(), tb,
                                       out.ptrs.data_ptr(), status.data_ptr() if want_status else None,
                                       _stream_ptr(stream))
    _check(st, f"hipcompBatched{_NAME[codec]}DecompressAsync")
    return actual, status
