set -x
python -m pytest tests -m gpu -x -q -k "cascaded or casc" > gpurun_out/gpu_tests_casc.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/gpu_tests_casc.log
python bench.py --codec cascaded --steps 5 --warmup 3 --no-e2e --no-cpu 2>gpurun_out/bench_casc.err | tee gpurun_out/bench_casc_new.json | python -c "import json,sys; j=json.loads(sys.stdin.read()); print(j['value'], j['compress_GBps'], j['decompress_GBps'], j['roofline_compress']['frac'], j['roofline_decompress']['frac'])"
tail -5 gpurun_out/bench_casc.err
