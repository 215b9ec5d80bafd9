# flush test + timing check (compute-sanitizer is closed on this pool)
set -x
timeout 600 python -m pytest tests/test_gpu_decoders.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2i_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r2i_tests.log
for c in lz4 snappy; do
timeout 300 python bench.py --codec $c --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2i_bench_$c.json 2> gpurun_out/r2i_bench_$c.err; echo "bench rc=$?"
done
