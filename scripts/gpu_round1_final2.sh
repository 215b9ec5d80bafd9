# round-1 closing run: GPU tests, the three bench lines, the CPU reference arm, C5 input-class sweep,
# launch list and full ncu captures of the LZ4 / Snappy kernels (one gpurun call)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/gpu_tests.log
for c in lz4 snappy cascaded; do
timeout 600 python bench.py --codec $c --steps 5 --warmup 3 > gpurun_out/bench_${c}.json 2> gpurun_out/bench_${c}.err; echo "bench $c rc=$?"
done
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference_lz4.json 2> gpurun_out/bench_reference_lz4.err; echo "ref rc=$?"
: > gpurun_out/class_sweep.jsonl
for c in lz4 snappy cascaded; do for d in zeros random text logs ints; do
timeout 300 python bench.py --codec $c --data $d --size-mib 512 --steps 3 --warmup 3 --no-e2e --no-cpu 2>>gpurun_out/class_sweep.err | tail -1 >> gpurun_out/class_sweep.jsonl; echo "sweep $c $d rc=$?"
done; done
timeout 300 python bench.py --codec lz4 --chunk 4096 --size-mib 1024 --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/bench_lz4_4k.json 2>gpurun_out/bench_lz4_4k.err; echo "4k rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01_launches_lz4.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_launch.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'compress_kernel|decompress' -s 6 -c 2 -f -o gpurun_out/prof_lz4_final python bench.py --codec lz4 --size-mib 256 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_lz4_full.log 2>&1; echo "ncu lz4 rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'compress_kernel|decompress' -s 6 -c 2 -f -o gpurun_out/prof_snappy_final python bench.py --codec snappy --size-mib 256 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_snappy_full.log 2>&1; echo "ncu snappy rc=$?"
ls -la gpurun_out | head -40
