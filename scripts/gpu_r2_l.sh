set -x
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2l_tests.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2l_tests.log
for c in lz4 snappy; do
timeout 300 python bench.py --codec $c --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2l_bench_$c.json 2> gpurun_out/r2l_bench_$c.err; echo "bench rc=$?"; tail -3 gpurun_out/r2l_bench_$c.err
done
timeout 300 python bench.py --codec lz4 --chunk 4096 --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2l_bench_lz4_4k.json 2>/dev/null
