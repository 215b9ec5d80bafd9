"""Helpers for the -m gpu parity tests (call the C ABI through ctypes)."""
import importlib

import numpy as np
import torch

B = importlib.import_module("hipcomp-core_b200.batched")
L = importlib.import_module("hipcomp-core_b200._lib")


def roundtrip(codec, chunks, opts=None, max_chunk=None, cap_extra=0):
    """Compress + size query + decompress; returns (compressed chunks, decoded chunks, statuses, size-query)."""
    src = B.DeviceBatch.from_host(chunks)
    if max_chunk is None:
        max_chunk = max((len(c) for c in chunks), default=0)
    comp = B.compress(codec, src, opts=opts, max_chunk=max_chunk)
    sizes = B.decompress_sizes(codec, comp)
    stride = (max_chunk + cap_extra + 15) // 16 * 16 + 16
    out = B.DeviceBatch.empty(src.n, stride)
    out.sizes = torch.from_numpy(np.array([len(c) + cap_extra for c in chunks], dtype=np.int64)).cuda()
    actual, status = B.decompress(codec, comp, out)
    torch.cuda.synchronize()
    return comp.to_host(), out.to_host(actual), status.cpu().numpy(), sizes.cpu().numpy(), actual.cpu().numpy()


def decode_raw(codec, comp_chunks, caps, want_actual=True, want_status=True):
    comp = B.DeviceBatch.from_host(comp_chunks)
    stride = (max(list(caps) + [1]) + 15) // 16 * 16 + 16
    out = B.DeviceBatch.empty(comp.n, stride)
    out.buf.fill_(0xAB)
    out.sizes = torch.from_numpy(np.asarray(caps, dtype=np.int64)).cuda()
    actual, status = B.decompress(codec, comp, out, want_actual=want_actual, want_status=want_status)
    torch.cuda.synchronize()
    return out, actual, status
