#!/bin/sh
# TEST INFRASTRUCTURE.  Compiles the reference's OWN test programs, where they lie under /root/reference/tests,
# against THIS repo's include/ tree and libhipcomp.so: the proof that a C / C++ caller of the reference relinks
# unchanged (VERDICT r1 "missing" #6).  Nothing is copied; binaries go to oracle/_ref/tests/
# (git-ignored, travel to the GPU box, where tests/test_gpu_reference_suite.py runs them).
#   C ABI      : test_lz4batch_c_api.c, test_snappy_batch_c_api.c, test_cascadedbatch_c_api.c (test_batch_c_api.h)
#   C ABI, C++ : test_cascaded_batch.cpp (known-answer words, OOB matrix, size queries)
#   HLIF       : test_lz4.cpp, test_random_lz4.cpp, test_cascaded.cpp (hipcomp::LZ4Manager / CascadedManager)
# Not built: test_snappy_app.cpp and src/test/*_test.cpp include the reference's private headers
# (lowlevel/SnappyBatchKernels.h, TempSpaceBroker.h, ...), which are not part of the drop-in boundary; their
# vectors are restated in tests/golden/ and tests/test_gpu_primitives.py.
set -e
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
ROOT=$(dirname "$HERE")
OUT=$HERE/_ref/tests
[ -d "$REF/tests" ] || { echo "no reference tree at $REF: keeping any prebuilt $OUT"; exit 0; }
mkdir -p "$OUT"
CUDA=${CUDA_HOME:-/usr/local/cuda}
INC="-I$ROOT/include -I$ROOT/include/hipcomp -I$CUDA/include -D__HIP_PLATFORM_NVIDIA__"
LNK="-L$ROOT/hipcomp-core_b200/lib -lhipcomp -L$CUDA/lib64 -lcudart -Wl,-rpath,\$ORIGIN/../../../hipcomp-core_b200/lib -Wl,-rpath,$CUDA/lib64"
for t in test_lz4batch_c_api test_snappy_batch_c_api test_cascadedbatch_c_api; do
  [ "$OUT/$t" -nt "$REF/tests/$t.c" ] && [ "$OUT/$t" -nt "$ROOT/include/hipcomp.h" ] && continue
  gcc -std=gnu11 -O1 -w $INC "$REF/tests/$t.c" -o "$OUT/$t" $LNK -lm
done
for t in test_cascaded_batch test_lz4 test_random_lz4 test_cascaded; do
  [ "$OUT/$t" -nt "$REF/tests/$t.cpp" ] && [ "$OUT/$t" -nt "$ROOT/include/hipcomp.hpp" ] && continue
  g++ -std=c++17 -O1 -w $INC -I"$REF/tests" "$REF/tests/$t.cpp" -o "$OUT/$t" $LNK
done
# HLIF cross-codec driver (tests/cpp/hlif_xcodec.cpp, this repo's own source): once against this library (with the
# Workspace unit test compiled in), once against the reference's build - both from THIS repo's headers
g++ -std=c++17 -O1 -w $INC -DXCODEC_WORKSPACE_TEST -I"$ROOT/hipcomp-core_b200/csrc" "$ROOT/tests/cpp/hlif_xcodec.cpp" -o "$OUT/hlif_xcodec_ours" $LNK
if [ -f "$HERE/_ref/libhipcomp_ref.so" ]; then
  g++ -std=c++17 -O1 -w $INC "$ROOT/tests/cpp/hlif_xcodec.cpp" -o "$OUT/hlif_xcodec_ref" -L"$HERE/_ref" -l:libhipcomp_ref.so \
    -L$CUDA/lib64 -lcudart -Wl,-rpath,\$ORIGIN/.. -Wl,-rpath,$CUDA/lib64
fi
ls -la "$OUT"
