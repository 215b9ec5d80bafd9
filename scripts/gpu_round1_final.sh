mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/gpu_tests.log
for c in lz4 snappy cascaded; do
timeout 600 python bench.py --codec $c --steps 5 --warmup 3 > gpurun_out/bench_${c}.json 2> gpurun_out/bench_${c}.err; echo "bench $c rc=$?"
cat gpurun_out/bench_${c}.json
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01_launches_lz4.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_launch.log 2>&1; echo "ncu rc=$?"
