set -x
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_decoders.py tests/test_gpu_reference.py tests/test_gpu_hlif.py -m gpu -x -q > gpurun_out/r2o_tests.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2o_tests.log
for c in lz4 snappy; do
timeout 300 python bench.py --codec $c --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2o_bench_$c.json 2> gpurun_out/r2o_bench_$c.err; echo "bench rc=$?"; tail -3 gpurun_out/r2o_bench_$c.err
done
timeout 600 python bench.py --workload c5 --c5-gib 1 --steps 2 --no-cpu > gpurun_out/r2o_c5_small.jsonl 2> gpurun_out/r2o_c5.err; echo "c5 rc=$?"
