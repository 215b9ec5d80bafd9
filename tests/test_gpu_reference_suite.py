"""-m gpu: the reference's own test programs (tests/test_*_c_api.c, test_cascaded_batch.cpp, test_lz4.cpp,
test_random_lz4.cpp, test_cascaded.cpp), compiled where they lie against this repo's include/ + libhipcomp.so by
oracle/build_ref_tests.sh, run unchanged: the drop-in boundary as the reference's maintainers test it."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "tests")
PROGRAMS = ["test_lz4batch_c_api", "test_snappy_batch_c_api", "test_cascadedbatch_c_api", "test_cascaded_batch", "test_lz4",
            "test_random_lz4", "test_cascaded"]


@pytest.mark.parametrize("name", PROGRAMS)
def test_reference_program_passes_against_this_library(name):
    exe = os.path.join(BIN, name)
    if not os.path.exists(exe):
        pytest.skip(f"{exe} not built (oracle/build_ref_tests.sh needs /root/reference)")
    out = subprocess.run([exe], capture_output=True, text=True, timeout=900, cwd=BIN)
    tail = (out.stdout + out.stderr)[-3000:]
    assert out.returncode == 0, f"{name} failed:\n{tail}"
    assert "FAILED" not in out.stdout, tail


def test_programs_link_this_library_not_the_reference():
    exe = os.path.join(BIN, PROGRAMS[0])
    if not os.path.exists(exe):
        pytest.skip("not built")
    ldd = subprocess.run(["ldd", exe], capture_output=True, text=True).stdout
    line = [l for l in ldd.splitlines() if "libhipcomp.so" in l]
    assert line and os.path.join("hipcomp-core_b200", "lib") in os.path.realpath(line[0].split("=>")[1].split("(")[0].strip())
