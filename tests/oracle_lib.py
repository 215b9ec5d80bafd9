"""ctypes view of oracle/liboracle.so (CPU restatements of the reference codecs).

TEST INFRASTRUCTURE: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py only.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_ORACLE_DIR = os.path.join(_ROOT, "oracle")
_SO = os.path.join(_ORACLE_DIR, "liboracle.so")

TYPE_WIDTH = {0: 1, 1: 1, 2: 2, 3: 2, 4: 4, 5: 4, 6: 8, 7: 8, 0xFF: 1}
SUCCESS = 0
CANNOT_DECOMPRESS = 12


def build():
    srcs = [os.path.join(_ORACLE_DIR, f)
            for f in ("cascaded_oracle.c", "lz4_oracle.c", "snappy_oracle.c", "lz_team_model.c", "cpu_bench.c", "Makefile")]
    if os.path.exists(_SO) and all(os.path.getmtime(_SO) >= os.path.getmtime(s) for s in srcs):
        return _SO
    subprocess.check_call(["make", "-C", _ORACLE_DIR, "liboracle.so"], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        sz, u8p, vp = C.c_size_t, C.POINTER(C.c_uint8), C.c_void_p
        L = _lib
        L.casc_oracle_compress.restype = sz
        L.casc_oracle_compress.argtypes = [vp, sz, vp, C.c_int, C.c_int, C.c_int, C.c_int]
        L.casc_oracle_decompress.restype = C.c_int
        L.casc_oracle_decompress.argtypes = [vp, sz, vp, sz, C.c_int, C.POINTER(sz)]
        L.casc_oracle_get_size.restype = sz
        L.casc_oracle_get_size.argtypes = [vp, sz]
        L.lz4_oracle_hash_table_size.restype = C.c_uint32
        L.lz4_oracle_hash_table_size.argtypes = [sz]
        L.lz4_oracle_max_compressed.restype = sz
        L.lz4_oracle_max_compressed.argtypes = [sz]
        L.lz4_oracle_compress.restype = sz
        L.lz4_oracle_compress.argtypes = [vp, sz, vp, C.c_int, C.c_uint32]
        L.lz4_oracle_decompress.restype = C.c_int
        L.lz4_oracle_decompress.argtypes = [vp, sz, vp, sz, C.c_int, C.POINTER(sz)]
        L.lz4_team_model_compress.restype = sz
        L.lz4_team_model_compress.argtypes = [vp, sz, vp, C.c_int]
        L.snappy_team_model_compress.restype = sz
        L.snappy_team_model_compress.argtypes = [vp, sz, vp, C.c_int]
        L.snappy_oracle_max_compressed.restype = sz
        L.snappy_oracle_max_compressed.argtypes = [sz]
        L.snappy_oracle_compress.restype = sz
        L.snappy_oracle_compress.argtypes = [vp, sz, vp]
        L.snappy_oracle_get_size.restype = sz
        L.snappy_oracle_get_size.argtypes = [vp, sz]
        L.snappy_oracle_decompress.restype = C.c_int
        L.snappy_oracle_decompress.argtypes = [vp, sz, vp, sz, C.POINTER(sz)]
        for name in ("casc", "lz4", "snappy"):
            getattr(L, f"{name}_oracle_compress_batch").restype = None
            getattr(L, f"{name}_oracle_decompress_batch").restype = None
    return _lib


def _buf(b):
    a = np.frombuffer(bytes(b), dtype=np.uint8) if not isinstance(b, np.ndarray) else b
    a = np.ascontiguousarray(a.view(np.uint8).reshape(-1))
    return a


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


# ---------------------------------------------------------------- cascaded
def casc_compress(data, type_id, num_rles, num_deltas, use_bp):
    a = _buf(data)
    out = np.zeros((len(a) + 3) // 4 * 4 + 8 + 8, dtype=np.uint8)
    n = lib().casc_oracle_compress(_p(a), len(a), _p(out), type_id, num_rles, num_deltas, use_bp)
    return out[:n].copy()


def casc_decompress(comp, out_cap, width):
    a = _buf(comp)
    out = np.zeros(max(out_cap, 1), dtype=np.uint8)
    actual = C.c_size_t(0)
    st = lib().casc_oracle_decompress(_p(a), len(a), _p(out), out_cap, width, C.byref(actual))
    return st, out[: actual.value].copy()


def casc_get_size(comp):
    a = _buf(comp)
    return lib().casc_oracle_get_size(_p(a), len(a))


# ---------------------------------------------------------------- lz4
def lz4_hash_table_size(max_chunk):
    return lib().lz4_oracle_hash_table_size(max_chunk)


def lz4_max_compressed(n):
    return lib().lz4_oracle_max_compressed(n)


def lz4_compress(data, tsize=1, max_chunk=None):
    a = _buf(data)
    ht = lz4_hash_table_size(max_chunk if max_chunk is not None else max(len(a), 1))
    out = np.zeros(lz4_max_compressed(len(a)) + 16, dtype=np.uint8)
    n = lib().lz4_oracle_compress(_p(a), len(a), _p(out), tsize, ht)
    return out[:n].copy()


def lz4_team_hash_log(max_chunk):
    """Hash-table bits the team encoder uses for a declared maximum chunk size (csrc/lz4.cu launch_team)."""
    return 12 if max_chunk <= 4096 else 13 if max_chunk <= 16384 else 14


def lz4_team_compress(data, max_chunk=None):
    """b200comp's own encoder policy for chunks <= 64 KiB (oracle/lz_team_model.c)."""
    a = _buf(data)
    assert len(a) <= 65536
    out = np.zeros(lz4_max_compressed(len(a)) + 16, dtype=np.uint8)
    n = lib().lz4_team_model_compress(_p(a), len(a), _p(out), lz4_team_hash_log(max_chunk or max(len(a), 1)))
    return out[:n].copy()


def lz4_product_compress(data, tsize=1, max_chunk=None):
    """The stream libhipcomp.so is expected to produce: team policy up to 64 KiB declared chunk size,
    the reference's policy above."""
    mc = max_chunk if max_chunk is not None else max(len(_buf(data)), 1)
    return lz4_team_compress(data, mc) if mc <= 65536 else lz4_compress(data, tsize, mc)


def lz4_decompress(comp, cap, write_output=True):
    a = _buf(comp)
    out = np.zeros(max(cap, 1) if write_output else 1, dtype=np.uint8)
    actual = C.c_size_t(0)
    st = lib().lz4_oracle_decompress(_p(a), len(a), _p(out), cap, 1 if write_output else 0, C.byref(actual))
    return st, (out[: actual.value].copy() if write_output else actual.value)


# ---------------------------------------------------------------- snappy
def snappy_max_compressed(n):
    return lib().snappy_oracle_max_compressed(n)


def snappy_compress(data):
    a = _buf(data)
    out = np.zeros(snappy_max_compressed(len(a)) + 16, dtype=np.uint8)
    n = lib().snappy_oracle_compress(_p(a), len(a), _p(out))
    return out[:n].copy()


def snappy_team_compress(data, max_chunk=None):
    """b200comp's own Snappy encoder policy for chunks <= 64 KiB (oracle/lz_team_model.c)."""
    a = _buf(data)
    assert len(a) <= 65536
    out = np.zeros(snappy_max_compressed(len(a)) + 16, dtype=np.uint8)
    n = lib().snappy_team_model_compress(_p(a), len(a), _p(out), lz4_team_hash_log(max_chunk or max(len(a), 1)))
    return out[:n].copy()


def snappy_product_compress(data, max_chunk=None):
    """The stream libhipcomp.so is expected to produce: team policy up to 64 KiB declared chunk size, the
    reference's policy above (or when the caller declares no maximum)."""
    mc = max_chunk if max_chunk is not None else max(len(_buf(data)), 1)
    return snappy_team_compress(data, mc) if 0 < mc <= 65536 else snappy_compress(data)


def snappy_get_size(comp):
    a = _buf(comp)
    return lib().snappy_oracle_get_size(_p(a), len(a))


def snappy_decompress(comp, cap):
    a = _buf(comp)
    out = np.zeros(max(cap, snappy_get_size(comp) if cap == 0 else cap, 1), dtype=np.uint8)
    actual = C.c_size_t(0)
    st = lib().snappy_oracle_decompress(_p(a), len(a), _p(out), cap, C.byref(actual))
    return st, out[: actual.value].copy()
