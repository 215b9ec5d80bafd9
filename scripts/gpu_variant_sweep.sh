# usage: gpu_variant_sweep.sh <codec> file1.so file2.so ...   (swaps the library in the box's scratch copy)
c=$1; shift
cp hipcomp-core_b200/lib/libhipcomp.so /tmp/orig.so
for f in "$@"; do
cp $f hipcomp-core_b200/lib/libhipcomp.so
timeout 300 python bench.py --codec $c --steps 3 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$f', 'comp', round(d['compress_GBps'],1), 'decomp', round(d['decompress_GBps'],1), 'ratio', round(d['compression_ratio'],3))"
done
cp /tmp/orig.so hipcomp-core_b200/lib/libhipcomp.so
