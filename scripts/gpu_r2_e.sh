# decoder iteration: parity + LZ4 bench + ncu of the decoder
set -x
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_decoders.py tests/test_gpu_reference.py -m gpu -x -q > gpurun_out/r2e_tests.log 2>&1; echo "pytest rc=$?"
tail -12 gpurun_out/r2e_tests.log
for c in lz4 snappy; do
timeout 300 python bench.py --codec $c --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2e_bench_$c.json 2> gpurun_out/r2e_bench_$c.err; echo "bench rc=$?"; tail -3 gpurun_out/r2e_bench_$c.err
done
timeout 300 python bench.py --codec lz4 --chunk 4096 --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2e_bench_lz4_4k.json 2>/dev/null
timeout 600 ncu --set full --clock-control none --import-source on -k regex:decompress_window -s 3 -c 1 -o gpurun_out/r2e_prof_snappydec python bench.py --codec snappy --size-mib 256 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2e_ncu.log 2>&1
