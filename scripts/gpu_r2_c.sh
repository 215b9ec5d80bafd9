# full GPU suite + the three bench lines + ncu of the Snappy team encoder
set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2c_tests.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/r2c_tests.log
for c in lz4 snappy cascaded; do
timeout 300 python bench.py --codec $c --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2c_bench_$c.json 2> gpurun_out/r2c_bench_$c.err; echo "bench $c rc=$?"
tail -3 gpurun_out/r2c_bench_$c.err
done
timeout 300 python bench.py --codec lz4 --chunk 4096 --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2c_bench_lz4_4k.json 2> gpurun_out/r2c_bench_lz4_4k.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'compress_team_kernel|decompress_window' -s 6 -c 2 -o gpurun_out/r2c_prof_snappy python bench.py --codec snappy --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2c_ncu_snappy.log 2>&1
