"""ctypes view of the host-buffer staging pipeline (include/hipcomp/b200_ext.h: b200compStager*): pinned host
buffers as numpy arrays, one call per direction.  Harness only — the pipeline itself is csrc/stager.cu."""
import ctypes as C

import numpy as np

from . import _lib as L

CODEC_ID = {"lz4": 0, "snappy": 1, "cascaded": 2}


def _check(st, what):
    if st != 0:
        raise RuntimeError(f"{what} -> status {st}")


class PinnedArray:
    """A pinned + mapped host allocation (b200compHostAlloc) seen as a numpy array."""

    def __init__(self, count, dtype=np.uint8):
        self.dtype = np.dtype(dtype)
        self.nbytes = int(count) * self.dtype.itemsize
        self._p = C.c_void_p()
        _check(L.lib().b200compHostAlloc(max(self.nbytes, 1), C.byref(self._p)), "b200compHostAlloc")
        buf = (C.c_uint8 * max(self.nbytes, 1)).from_address(self._p.value)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(count))

    @property
    def ptr(self):
        return self._p.value

    def free(self):
        if self._p:
            self.array = None
            L.lib().b200compHostFree(self._p)
            self._p = C.c_void_p()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Stager:
    def __init__(self, codec, chunk_bytes, max_bytes, opts=None, slices=8):
        self.codec, self.chunk, self.max_bytes = codec, chunk_bytes, max_bytes
        self._h = C.c_void_p()
        self._opts = opts
        _check(L.lib().b200compStagerCreate(CODEC_ID[codec], chunk_bytes, max_bytes, C.byref(opts) if opts is not None else None,
                                            slices, C.byref(self._h)), "b200compStagerCreate")

    def nchunks(self, nbytes):
        return (nbytes + self.chunk - 1) // self.chunk

    def compress(self, raw, blob, sizes):
        """raw: PinnedArray uint8; blob: PinnedArray uint8 (>= worst case); sizes: PinnedArray uint64.  Returns total."""
        total = C.c_size_t(0)
        _check(L.lib().b200compStagerCompress(self._h, raw.ptr, raw.nbytes, blob.ptr, sizes.ptr, C.byref(total)),
               "b200compStagerCompress")
        return total.value

    def decompress(self, blob, sizes, out, statuses=None):
        _check(L.lib().b200compStagerDecompress(self._h, blob.ptr, sizes.ptr, self.nchunks(out.nbytes), out.ptr, out.nbytes,
                                                statuses.ptr if statuses is not None else None), "b200compStagerDecompress")

    def roundtrip(self, raw, blob, sizes, out, statuses=None):
        total = C.c_size_t(0)
        _check(L.lib().b200compStagerRoundTrip(self._h, raw.ptr, raw.nbytes, blob.ptr, sizes.ptr, out.ptr,
                                               statuses.ptr if statuses is not None else None, C.byref(total)),
               "b200compStagerRoundTrip")
        return total.value

    def close(self):
        if self._h:
            L.lib().b200compStagerDestroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
