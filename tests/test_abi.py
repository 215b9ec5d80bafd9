"""CPU-only checks of the drop-in boundary: libhipcomp.so loads, exports every symbol the
headers under include/ declare, and its host-only entry points (sizes, argument errors) behave
like the reference's (src/lowlevel/LZ4Batch.cpp:71-187, SnappyBatch.cpp:72-206,
CascadedBatch.hip:306-364).  No kernel is launched here."""
import ctypes as C
import importlib
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
L = importlib.import_module("hipcomp-core_b200._lib")


def declared_symbols():
    names = set()
    inc = os.path.join(ROOT, "include")
    for dirpath, _, files in os.walk(inc):
        for f in files:
            if f.endswith(".h"):
                text = open(os.path.join(dirpath, f)).read()
                names.update(re.findall(r"\b((?:hipcomp|b200comp)[A-Za-z0-9_]+)\s*\(", text))
    return sorted(names)


def test_headers_declare_the_18_reference_symbols():
    names = [n for n in declared_symbols() if n.startswith("hipcompBatched")]
    assert len(names) == 18
    assert set(names) == set(L.PROTOTYPES)


def test_library_exports_every_declared_symbol():
    out = subprocess.check_output(["nm", "-D", "--defined-only", L.LIB_PATH], text=True)
    exported = {line.split()[-1] for line in out.splitlines() if line.strip()}
    missing = [n for n in declared_symbols() if n not in exported]
    assert not missing, f"declared in include/ but not exported: {missing}"
    lib = L.lib()
    for n in declared_symbols():
        assert getattr(lib, n) is not None


def test_soname():
    out = subprocess.check_output(["readelf", "-d", L.LIB_PATH], text=True)
    assert "libhipcomp.so" in out


@pytest.mark.parametrize("n", [0, 1, 255, 4096, 65536, 1 << 24])
def test_lz4_sizes(n):
    lib = L.lib()
    out = C.c_size_t(0)
    assert lib.hipcompBatchedLZ4CompressGetMaxOutputChunkSize(n, L.LZ4Opts(L.TYPE_CHAR), C.byref(out)) == 0
    assert out.value == (n + 1 + (n + 254) // 255 + 7) // 8 * 8  # LZ4Kernels.hiph:198-202
    assert lib.hipcompBatchedLZ4CompressGetTempSize(7, n, L.LZ4Opts(L.TYPE_CHAR), C.byref(out)) == 0
    ht = 1
    while ht < n:
        ht *= 2
    ref = 7 * min(ht, 16384) * 2  # LZ4CompressionKernels.hip:142-156,287-296
    # never more than the reference asks for (callers' allocations stay valid); chunks <= 64 KiB need no workspace
    assert out.value == (ref if n > 65536 else 8)
    assert lib.hipcompBatchedLZ4DecompressGetTempSize(5, n, C.byref(out)) == 0
    assert out.value == 120  # roundUp8(24 * chunks)


def test_lz4_argument_errors():
    lib = L.lib()
    out = C.c_size_t(0)
    too_big = (1 << 24) + 1
    assert lib.hipcompBatchedLZ4CompressGetTempSize(1, too_big, L.LZ4Opts(0), C.byref(out)) == L.ERROR_INVALID_VALUE
    assert lib.hipcompBatchedLZ4CompressGetMaxOutputChunkSize(too_big, L.LZ4Opts(0), C.byref(out)) == L.ERROR_INVALID_VALUE
    assert lib.hipcompBatchedLZ4CompressGetTempSize(1, 10, L.LZ4Opts(0), None) == L.ERROR_INVALID_VALUE
    assert lib.hipcompBatchedLZ4DecompressGetTempSize(1, 10, None) == L.ERROR_INVALID_VALUE
    # insufficient temp space is diagnosed before any launch (LZ4Batch.cpp:189-224)
    assert lib.hipcompBatchedLZ4CompressAsync(None, None, 65536, 4, None, 0, None, None, L.LZ4Opts(0), None) \
        == L.ERROR_INVALID_VALUE


@pytest.mark.parametrize("n", [0, 1, 5, 65536, 1 << 20])
def test_snappy_sizes(n):
    lib = L.lib()
    out = C.c_size_t(99)
    assert lib.hipcompBatchedSnappyCompressGetMaxOutputChunkSize(n, L.SnappyOpts(0), C.byref(out)) == 0
    assert out.value == 32 + n + n // 6  # SnappyBatch.cpp:72-76
    assert lib.hipcompBatchedSnappyCompressGetTempSize(3, n, L.SnappyOpts(0), C.byref(out)) == 0 and out.value == 0
    assert lib.hipcompBatchedSnappyDecompressGetTempSize(3, n, C.byref(out)) == 0 and out.value == 0
    assert lib.hipcompBatchedSnappyCompressGetTempSize(3, n, L.SnappyOpts(0), None) == L.ERROR_INVALID_VALUE


@pytest.mark.parametrize("n", [0, 1, 4095, 4096, 65537])
def test_cascaded_sizes(n):
    lib = L.lib()
    out = C.c_size_t(99)
    o = L.CascadedOpts(4096, L.TYPE_INT, 2, 1, 1)
    assert lib.hipcompBatchedCascadedCompressGetMaxOutputChunkSize(n, o, C.byref(out)) == 0
    assert out.value == (n + 3) // 4 * 4 + 8  # CascadedBatch.hip:318-327
    assert lib.hipcompBatchedCascadedCompressGetTempSize(3, n, o, C.byref(out)) == 0 and out.value == 0
    assert lib.hipcompBatchedCascadedDecompressGetTempSize(3, n, C.byref(out)) == 0 and out.value == 0


def test_empty_batches_succeed_without_a_device():
    lib = L.lib()
    for c, o in (("LZ4", L.LZ4Opts(0)), ("Snappy", L.SnappyOpts(0)), ("Cascaded", L.CascadedOpts(4096, 4, 1, 1, 1))):
        dummy = C.c_void_p(16)
        assert getattr(lib, f"hipcompBatched{c}CompressAsync")(dummy, dummy, 0, 0, None, 0, dummy, dummy, o, None) == 0
        assert getattr(lib, f"hipcompBatched{c}DecompressAsync")(dummy, dummy, dummy, dummy, 0, None, 0, dummy, dummy,
                                                                 None) == 0


def test_library_exports_the_hlif_classes():
    """C++ side of the boundary: manager ctors/dtors, config special members, factory."""
    out = subprocess.check_output(["nm", "-DC", "--defined-only", L.LIB_PATH], text=True)
    for sym in ("hipcomp::LZ4Manager::LZ4Manager(", "hipcomp::LZ4Manager::~LZ4Manager(",
                "hipcomp::SnappyManager::SnappyManager(", "hipcomp::CascadedManager::CascadedManager(",
                "hipcomp::CascadedManager::~CascadedManager(", "hipcomp::create_manager(",
                "hipcomp::CompressionConfig::CompressionConfig(", "hipcomp::CompressionConfig::get_status(",
                "hipcomp::DecompressionConfig::DecompressionConfig(", "hipcomp::DecompressionConfig::get_status("):
        assert sym in out, f"{sym} not exported"


def test_public_headers_compile_as_plain_cxx_and_c(tmp_path):
    """A caller's translation unit: C++14 against the .hpp headers, C99 against the .h headers."""
    inc = ["-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "include", "hipcomp"), "-I", "/usr/local/cuda/include"]
    cxx = tmp_path / "user.cpp"
    cxx.write_text('#include "hipcomp/hipcompManagerFactory.hpp"\n#include "hipcomp/lz4.h"\n#include "hipcomp/snappy.h"\n'
                   '#include "hipcomp/cascaded.h"\n#include "hipcomp/b200_ext.h"\n'
                   "size_t f(hipcomp::hipcompManagerBase& m, const uint8_t* in, uint8_t* out, size_t n) {\n"
                   "  auto cfg = m.configure_compression(n); m.compress(in, out, cfg);\n"
                   "  return m.get_compressed_output_size(out); }\n"
                   "hipcomp::LZ4Manager* g() { return new hipcomp::LZ4Manager(65536, HIPCOMP_TYPE_CHAR); }\n")
    subprocess.check_call(["g++", "-std=c++14", "-fsyntax-only", *inc, str(cxx)])
    c = tmp_path / "user.c"
    c.write_text('#include "hipcomp/lz4.h"\n#include "hipcomp/snappy.h"\n#include "hipcomp/cascaded.h"\n'
                 '#include "hipcomp/b200_ext.h"\nint main(void) { size_t t; return (int)hipcompBatchedLZ4DecompressGetTempSize(1, 1, &t); }\n')
    subprocess.check_call(["gcc", "-std=c99", "-fsyntax-only", *inc, str(c)])
