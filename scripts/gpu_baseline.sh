# tests + default bench of every codec (one gpurun call)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/gpu_tests.log
for c in lz4 snappy cascaded; do
timeout 600 python bench.py --codec $c --steps 5 --warmup 3 > gpurun_out/bench_${c}.json 2> gpurun_out/bench_${c}.err; echo "bench $c rc=$?"
cat gpurun_out/bench_${c}.json
done
