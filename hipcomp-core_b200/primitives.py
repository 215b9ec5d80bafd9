"""Python mirror of include/hipcomp/primitives.hpp (RunLengthEncodeGPU, DeltaGPU, BitPackGPU):
same static-method names and argument meaning, over the C view in b200_ext.h.  Tensors are
torch CUDA tensors; "device pointer to pointer" arguments are built here from int64 tensors."""
import ctypes as C

import torch

from . import _lib as L

NP_TYPE = {"int8": L.TYPE_CHAR, "uint8": L.TYPE_UCHAR, "int16": L.TYPE_SHORT, "uint16": L.TYPE_USHORT,
           "int32": L.TYPE_INT, "uint32": L.TYPE_UINT, "int64": L.TYPE_LONGLONG, "uint64": L.TYPE_ULONGLONG}


class PrimitiveError(RuntimeError):
    pass


def _ck(st, fn):
    if st != L.SUCCESS:
        raise PrimitiveError(f"{fn} returned hipcompStatus_t {st}")


def _sp(stream=None):
    return C.c_void_p((stream or torch.cuda.current_stream()).cuda_stream)


def _ptr_cell(t):
    """device int64 cell holding t.data_ptr() (a `void**` argument)"""
    return torch.tensor([t.data_ptr()], dtype=torch.int64, device=t.device)


class RunLengthEncodeGPU:
    @staticmethod
    def requiredWorkspaceSize(num, value_type, count_type):
        out = C.c_size_t()
        _ck(L.lib().b200compRleRequiredWorkspaceSize(num, value_type, count_type, C.byref(out)), "requiredWorkspaceSize")
        return out.value

    @staticmethod
    def compress(workspace, value_type, out_values, count_type, out_counts, num_out_device, inp, num, stream=None):
        _ck(L.lib().b200compRleCompress(workspace.data_ptr(), workspace.numel(), value_type, out_values.data_ptr(), count_type,
                                        out_counts.data_ptr(), num_out_device.data_ptr(), inp.data_ptr(), num, _sp(stream)),
            "RunLengthEncodeGPU::compress")

    @staticmethod
    def compressDownstream(workspace, value_type, out_values, count_type, out_counts, num_out_device, inp, num_device,
                           max_num, stream=None):
        vp, cp = _ptr_cell(out_values), _ptr_cell(out_counts)
        _ck(L.lib().b200compRleCompressDownstream(workspace.data_ptr(), workspace.numel(), value_type, vp.data_ptr(),
                                                  count_type, cp.data_ptr(), num_out_device.data_ptr(), inp.data_ptr(),
                                                  num_device.data_ptr(), max_num, _sp(stream)),
            "RunLengthEncodeGPU::compressDownstream")
        return vp, cp  # keep alive until the stream has run


class DeltaGPU:
    @staticmethod
    def requiredWorkspaceSize(num, value_type):
        return 0

    @staticmethod
    def compress(value_type, out_values, inp, num_device, max_num, stream=None):
        vp = _ptr_cell(out_values)
        _ck(L.lib().b200compDeltaCompress(value_type, vp.data_ptr(), inp.data_ptr(), num_device.data_ptr(), max_num, _sp(stream)),
            "DeltaGPU::compress")
        return vp


class BitPackGPU:
    @staticmethod
    def requiredWorkspaceSize(num, in_type):
        out = C.c_size_t()
        _ck(L.lib().b200compBitPackRequiredWorkspaceSize(num, in_type, C.byref(out)), "requiredWorkspaceSize")
        return out.value

    @staticmethod
    def compress(workspace, in_type, out, inp, num_device, max_num, min_value_device, num_bits_device, stream=None):
        op, mp, bp = _ptr_cell(out), _ptr_cell(min_value_device), _ptr_cell(num_bits_device)
        _ck(L.lib().b200compBitPackCompress(workspace.data_ptr(), workspace.numel(), in_type, op.data_ptr(), inp.data_ptr(),
                                            num_device.data_ptr(), max_num, mp.data_ptr(), bp.data_ptr(), _sp(stream)),
            "BitPackGPU::compress")
        return op, mp, bp
