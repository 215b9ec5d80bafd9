set -x
python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_4.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/gpu_tests_4.log
for c in lz4 snappy cascaded; do
python bench.py --codec $c --steps 5 --warmup 3 > gpurun_out/bench_${c}_r1c.json 2> gpurun_out/bench_${c}_r1c.err; echo "bench $c rc=$?"
done
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference_lz4_r1c.json 2>&1
python bench.py --codec lz4 --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain_lz4.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_lz4_r1.csv python bench.py --codec lz4 --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_lz4.log 2>&1
python bench.py --codec lz4 --size-mib 256 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain_lz4b.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'compress_kernel|decompress_kernel' -s 6 -c 2 -o gpurun_out/prof_lz4_r1 python bench.py --codec lz4 --size-mib 256 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_lz4b.log 2>&1
python bench.py --codec cascaded --size-mib 512 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain_casc.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'compress_kernel|decompress_kernel' -s 15 -c 5 -o gpurun_out/prof_casc_r1 python bench.py --codec cascaded --size-mib 512 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_casc.log 2>&1
python bench.py --codec snappy --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain_snappy.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'compress_kernel|decompress_kernel' -s 6 -c 2 -o gpurun_out/prof_snappy_r1 python bench.py --codec snappy --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_snappy.log 2>&1
ls -la gpurun_out
