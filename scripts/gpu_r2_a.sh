# round-2 opening measurement: tests, the three bench lines, ncu of the LZ4 team encoder + window decoder
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2a_gpu_tests.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/r2a_gpu_tests.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r2a_bench_lz4.json 2> gpurun_out/r2a_bench_lz4.err; echo "bench rc=$?"
for c in snappy cascaded; do
python bench.py --codec $c --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2a_bench_${c}.json 2> gpurun_out/r2a_bench_${c}.err; echo "bench $c rc=$?"
done
python bench.py --codec lz4 --size-mib 256 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2a_plain_lz4.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'compress_team_kernel|decompress_window_kernel' -s 6 -c 2 -o gpurun_out/r2a_prof_lz4 python bench.py --codec lz4 --size-mib 256 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2a_ncu_lz4.log 2>&1
ls -la gpurun_out
