#!/bin/sh
# TEST INFRASTRUCTURE.  Compiles the reference's own sources, where they lie under
# /root/reference, into oracle/_ref/libhipcomp_ref.so (CUDA backend, sm_100 only).
# Nothing is copied; the reference's build system is not run.  Two work-arounds are
# needed for the reference's CUDA backend to build at all:
#   --pre-include src/device_types.h   (its "device_types.h" is shadowed by CUDA's)
#   --allow-multiple-definition        (a non-inline __device__ helper in a header)
# The .so is git-ignored but travels to the GPU box, where tests/test_gpu_reference.py
# and bench_ref.py use it as the GPU oracle and the kernel-to-beat.
set -e
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/_ref
[ -d "$REF/src" ] || { echo "no reference tree at $REF: keeping any prebuilt $OUT"; exit 0; }
mkdir -p "$OUT/obj"
if [ -f "$OUT/libhipcomp_ref.so" ] && [ "$OUT/libhipcomp_ref.so" -nt "$HERE/build_ref.sh" ]; then
  echo "oracle/_ref/libhipcomp_ref.so is up to date"; exit 0
fi
FLAGS="-std=c++14 -O3 --expt-extended-lambda -Wno-deprecated-gpu-targets -w \
 -D__HIP_PLATFORM_NVCC__ -D__HIP_PLATFORM_NVIDIA__ \
 -I$REF/src -I$REF/include -I$REF/include/hipcomp --pre-include=$REF/src/device_types.h \
 -gencode arch=compute_100,code=sm_100 -Xcompiler -fPIC"
SRCS="$(ls $REF/src/*.hip $REF/src/lowlevel/*.hip $REF/src/highlevel/*.hip $REF/src/*.cpp $REF/src/lowlevel/*.cpp $REF/src/highlevel/*.cpp)"
for s in $SRCS; do
  o="$OUT/obj/$(echo "$s" | sed "s#$REF/src/##; s#/#_#g").o"
  echo "nvcc -x cu $FLAGS -c $s -o $o"
done | xargs -P 8 -I{} sh -c '{}'
nvcc -shared -gencode arch=compute_100,code=sm_100 -o "$OUT/libhipcomp_ref.so" "$OUT"/obj/*.o -Xlinker --allow-multiple-definition
rm -rf "$OUT/obj"
ls -la "$OUT/libhipcomp_ref.so"
