"""b200comp: B200-native batched LZ4 / Snappy / Cascaded codecs behind the hipCOMP C ABI.

The product is hipcomp-core_b200/lib/libhipcomp.so (C ABI, see include/).  This
package is the Python-side harness used by the tests and bench.py: ctypes
bindings, torch device buffers, the multi-GPU sharder.
"""
from . import _lib  # noqa: F401
from ._lib import lib  # noqa: F401
