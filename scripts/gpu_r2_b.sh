# LZ4 team encoder v2: parity + bench + ncu
set -x
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "lz4" > gpurun_out/r2b_tests.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/r2b_tests.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2b_bench_lz4.json 2> gpurun_out/r2b_bench_lz4.err; echo "bench rc=$?"
cat gpurun_out/r2b_bench_lz4.json | cut -c1-400
tail -5 gpurun_out/r2b_bench_lz4.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'compress_team_kernel' -s 3 -c 1 -o gpurun_out/r2b_prof_lz4 python bench.py --codec lz4 --size-mib 256 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2b_ncu_lz4.log 2>&1
