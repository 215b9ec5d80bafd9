"""Loads libhipcomp.so (the sm_100a C-ABI library) and declares its 18 entry points.

There is no CPU fallback: if the shared library is missing this raises.
Build it with `python -c "import __graft_entry__ as g; g.build()"` or
`make -C hipcomp-core_b200/csrc`.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libhipcomp.so")

SUCCESS = 0
ERROR_INVALID_VALUE = 10
ERROR_NOT_SUPPORTED = 11
ERROR_CANNOT_DECOMPRESS = 12
ERROR_CUDA = 1000
ERROR_INTERNAL = 10000

TYPE_CHAR, TYPE_UCHAR, TYPE_SHORT, TYPE_USHORT, TYPE_INT, TYPE_UINT, TYPE_LONGLONG, TYPE_ULONGLONG = range(8)
TYPE_BITS = 0xFF


class LZ4Opts(C.Structure):
    _fields_ = [("data_type", C.c_int)]


class SnappyOpts(C.Structure):
    _fields_ = [("reserved", C.c_int)]


class CascadedOpts(C.Structure):
    _fields_ = [("chunk_size", C.c_size_t), ("type", C.c_int), ("num_RLEs", C.c_int), ("num_deltas", C.c_int),
                ("use_bp", C.c_int)]


_sz, _vp, _szp = C.c_size_t, C.c_void_p, C.POINTER(C.c_size_t)

# name -> argtypes (restype is always hipcompStatus_t == int); order as in include/hipcomp/*.h
PROTOTYPES = {}
for _c, _opts in (("LZ4", LZ4Opts), ("Snappy", SnappyOpts), ("Cascaded", CascadedOpts)):
    PROTOTYPES[f"hipcompBatched{_c}CompressGetTempSize"] = [_sz, _sz, _opts, _szp]
    PROTOTYPES[f"hipcompBatched{_c}CompressGetMaxOutputChunkSize"] = [_sz, _opts, _szp]
    PROTOTYPES[f"hipcompBatched{_c}CompressAsync"] = [_vp, _vp, _sz, _sz, _vp, _sz, _vp, _vp, _opts, _vp]
    PROTOTYPES[f"hipcompBatched{_c}DecompressGetTempSize"] = [_sz, _sz, _szp]
    PROTOTYPES[f"hipcompBatched{_c}DecompressAsync"] = [_vp, _vp, _vp, _vp, _sz, _vp, _sz, _vp, _vp, _vp]
    PROTOTYPES[f"hipcompBatched{_c}GetDecompressSizeAsync"] = [_vp, _vp, _vp, _sz, _vp]

# extensions (include/hipcomp/b200_ext.h): gather + the C view of the HLIF managers
_vpp = C.POINTER(C.c_void_p)
EXT_PROTOTYPES = {
    "b200compGatherChunksAsync": [_vp, _vp, _vp, _vp, _sz, _vp],
    "b200compBatchedCascadedSelectOptsAsync": [_vp, _vp, _sz, C.c_int, _vp, _vp],
    "b200compBatchedCascadedCompressWithOptsAsync": [_vp, _vp, _sz, _vp, _vp, C.c_int, _vp, _vp],
    "b200compBatchedLZ4CompressNoAllocAsync": [_vp, _vp, _sz, _sz, _vp, _vp, LZ4Opts, _vp],
    "b200compHostAlloc": [_sz, _vpp],
    "b200compStagerCreate": [C.c_int, _sz, _sz, _vp, C.c_int, _vpp],
    "b200compStagerCompress": [_vp, _vp, _sz, _vp, _vp, _szp],
    "b200compStagerDecompress": [_vp, _vp, _vp, _sz, _vp, _sz, _vp],
    "b200compStagerRoundTrip": [_vp, _vp, _sz, _vp, _vp, _vp, _vp, _szp],
    "b200compManagerCreateLZ4": [_sz, C.c_int, _vp, C.c_int, _vpp],
    "b200compManagerCreateSnappy": [_sz, _vp, C.c_int, _vpp],
    "b200compManagerCreateCascaded": [C.POINTER(CascadedOpts), _vp, C.c_int, _vpp],
    "b200compManagerCreateFromBuffer": [_vp, _vp, C.c_int, _vpp],
    "b200compManagerSetChecksumMode": [_vp, C.c_int],
    "b200compManagerConfigureCompression": [_vp, _sz, _vpp, _szp, _szp],
    "b200compManagerCompress": [_vp, _vp, _vp, _vp],
    "b200compManagerGetCompressedOutputSize": [_vp, _vp, _szp],
    "b200compManagerConfigureDecompression": [_vp, _vp, _vpp, _szp, _szp],
    "b200compManagerConfigureDecompressionFromComp": [_vp, _vp, _vpp, _szp, _szp],
    "b200compManagerDecompress": [_vp, _vp, _vp, _vp],
    "b200compManagerGetRequiredScratchBufferSize": [_vp, _szp],
    "b200compManagerSetScratchBuffer": [_vp, _vp],
    "b200compRleRequiredWorkspaceSize": [_sz, C.c_int, C.c_int, _szp],
    "b200compRleCompress": [_vp, _sz, C.c_int, _vp, C.c_int, _vp, _vp, _vp, _sz, _vp],
    "b200compRleCompressDownstream": [_vp, _sz, C.c_int, _vp, C.c_int, _vp, _vp, _vp, _vp, _sz, _vp],
    "b200compDeltaCompress": [C.c_int, _vp, _vp, _vp, _sz, _vp],
    "b200compBitPackRequiredWorkspaceSize": [_sz, C.c_int, _szp],
    "b200compBitPackCompress": [_vp, _sz, C.c_int, _vp, _vp, _vp, _sz, _vp, _vp, _vp],
    "b200compCompConfigStatus": [_vp],
    "b200compDecompConfigStatus": [_vp],
}
EXT_VOID = {
    "b200compHostFree": [_vp],
    "b200compStagerDestroy": [_vp],
    "b200compManagerDestroy": [_vp],
    "b200compCompConfigDestroy": [_vp],
    "b200compDecompConfigDestroy": [_vp],
}

_lib = None


def load_from(path):
    """Bind the same 18 prototypes on another build of the same ABI (tests and bench.py use this
    to drive the reference's own GPU library, oracle/_ref/libhipcomp_ref.so, as an oracle)."""
    other = C.CDLL(path)
    for name, args in PROTOTYPES.items():
        fn = getattr(other, name)
        fn.restype = C.c_int
        fn.argtypes = args
    return other


def use(handle):
    """Route this module's lib() to `handle` (None restores the product library)."""
    global _lib
    _lib = handle


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: the CUDA library has not been built "
                "(run __graft_entry__.build()). There is no CPU fallback.")
        _lib = C.CDLL(LIB_PATH)
        for name, args in list(PROTOTYPES.items()) + list(EXT_PROTOTYPES.items()):
            fn = getattr(_lib, name)
            fn.restype = C.c_int
            fn.argtypes = args
        for name, args in EXT_VOID.items():
            fn = getattr(_lib, name)
            fn.restype = None
            fn.argtypes = args
    return _lib
