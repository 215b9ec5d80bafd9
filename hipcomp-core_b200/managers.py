"""Python mirror of the high-level interface (include/hipcomp/hipcompManager.hpp): LZ4Manager,
SnappyManager, CascadedManager and create_manager(), over the C view in b200_ext.h.

Same call sequence as the reference's HLIF tests (tests/test_lz4.cpp, test_cascaded.cpp via
hipcomp.hpp): configure_compression -> compress -> get_compressed_output_size ->
configure_decompression -> decompress, with a status word per config.  Buffers are torch uint8
CUDA tensors; nothing here computes — every call lands in libhipcomp.so.
"""
import ctypes as C
import weakref

import torch

from . import _lib as L


class ManagerError(RuntimeError):
    def __init__(self, fn, status):
        super().__init__(f"{fn} returned hipcompStatus_t {status}")
        self.status = status


def _ck(status, fn):
    if status != L.SUCCESS:
        raise ManagerError(fn, status)


def _sp(stream):
    return C.c_void_p((stream or torch.cuda.current_stream()).cuda_stream)


class CompressionConfig:
    def __init__(self, handle, uncompressed_buffer_size, max_compressed_buffer_size, num_chunks):
        self._h = handle
        self.uncompressed_buffer_size = uncompressed_buffer_size
        self.max_compressed_buffer_size = max_compressed_buffer_size
        self.num_chunks = num_chunks

    def get_status(self):
        return L.lib().b200compCompConfigStatus(self._h)

    def _destroy(self):
        if getattr(self, "_h", None):
            L.lib().b200compCompConfigDestroy(self._h)
            self._h = None

    __del__ = _destroy


class DecompressionConfig:
    def __init__(self, handle, decomp_data_size, num_chunks):
        self._h = handle
        self.decomp_data_size = decomp_data_size
        self.num_chunks = num_chunks

    def get_status(self):
        return L.lib().b200compDecompConfigStatus(self._h)

    def _destroy(self):
        if getattr(self, "_h", None):
            L.lib().b200compDecompConfigDestroy(self._h)
            self._h = None

    __del__ = _destroy


class Manager:
    """hipcompManagerBase (hipcompManager.hpp:141-236)."""

    def __init__(self, handle, stream=None):
        self._h = handle
        self._stream = stream
        # a config hands its pinned status word back to the manager's pool when it dies (as in the reference,
        # PinnedPtrs.hpp:70-211), so the manager must outlive its configs: close() retires them first
        self._configs = weakref.WeakSet()

    def _track(self, cfg):
        self._configs.add(cfg)
        return cfg

    def configure_compression(self, decomp_buffer_size):
        h, mx, n = C.c_void_p(), C.c_size_t(), C.c_size_t()
        _ck(L.lib().b200compManagerConfigureCompression(self._h, decomp_buffer_size, C.byref(h), C.byref(mx), C.byref(n)),
            "configure_compression")
        return self._track(CompressionConfig(h, decomp_buffer_size, mx.value, n.value))

    def compress(self, decomp_buffer, comp_buffer, comp_config):
        _ck(L.lib().b200compManagerCompress(self._h, decomp_buffer.data_ptr(), comp_buffer.data_ptr(), comp_config._h), "compress")

    def configure_decompression(self, arg):
        h, sz, n = C.c_void_p(), C.c_size_t(), C.c_size_t()
        if isinstance(arg, CompressionConfig):
            _ck(L.lib().b200compManagerConfigureDecompressionFromComp(self._h, arg._h, C.byref(h), C.byref(sz), C.byref(n)),
                "configure_decompression")
        else:
            _ck(L.lib().b200compManagerConfigureDecompression(self._h, arg.data_ptr(), C.byref(h), C.byref(sz), C.byref(n)),
                "configure_decompression")
        return self._track(DecompressionConfig(h, sz.value, n.value))

    def decompress(self, decomp_buffer, comp_buffer, decomp_config):
        _ck(L.lib().b200compManagerDecompress(self._h, decomp_buffer.data_ptr(), comp_buffer.data_ptr(), decomp_config._h),
            "decompress")

    def get_compressed_output_size(self, comp_buffer):
        out = C.c_size_t()
        _ck(L.lib().b200compManagerGetCompressedOutputSize(self._h, comp_buffer.data_ptr(), C.byref(out)),
            "get_compressed_output_size")
        return out.value

    def get_required_scratch_buffer_size(self):
        out = C.c_size_t()
        _ck(L.lib().b200compManagerGetRequiredScratchBufferSize(self._h, C.byref(out)), "get_required_scratch_buffer_size")
        return out.value

    def set_scratch_buffer(self, scratch):
        self._scratch = scratch  # keep it alive
        _ck(L.lib().b200compManagerSetScratchBuffer(self._h, scratch.data_ptr()), "set_scratch_buffer")

    def set_checksum_mode(self, mode):
        """0: none (the reference's containers); 1: CRC-32 per compressed / uncompressed chunk, filled by compress()
        and verified by decompress() (extension, include/hipcomp/b200_ext.h)."""
        _ck(L.lib().b200compManagerSetChecksumMode(self._h, int(mode)), "set_checksum_mode")

    def close(self):
        if getattr(self, "_h", None):
            for cfg in list(getattr(self, "_configs", ())):
                cfg._destroy()
            L.lib().b200compManagerDestroy(self._h)
            self._h = None

    def __del__(self):
        self.close()


class LZ4Manager(Manager):
    def __init__(self, uncomp_chunk_size, data_type=L.TYPE_CHAR, stream=None, device_id=0):
        h = C.c_void_p()
        _ck(L.lib().b200compManagerCreateLZ4(uncomp_chunk_size, data_type, _sp(stream), device_id, C.byref(h)), "LZ4Manager")
        super().__init__(h, stream)


class SnappyManager(Manager):
    def __init__(self, uncomp_chunk_size, stream=None, device_id=0):
        h = C.c_void_p()
        _ck(L.lib().b200compManagerCreateSnappy(uncomp_chunk_size, _sp(stream), device_id, C.byref(h)), "SnappyManager")
        super().__init__(h, stream)


class CascadedManager(Manager):
    def __init__(self, options=None, stream=None, device_id=0):
        opts = options or L.CascadedOpts(4096, L.TYPE_INT, 2, 1, 1)
        h = C.c_void_p()
        _ck(L.lib().b200compManagerCreateCascaded(C.byref(opts), _sp(stream), device_id, C.byref(h)), "CascadedManager")
        super().__init__(h, stream)


def create_manager(comp_buffer, stream=None, device_id=0):
    """hipcomp::create_manager (hipcompManagerFactory.hpp:70): manager for an existing container."""
    h = C.c_void_p()
    _ck(L.lib().b200compManagerCreateFromBuffer(comp_buffer.data_ptr(), _sp(stream), device_id, C.byref(h)), "create_manager")
    return Manager(h, stream)
