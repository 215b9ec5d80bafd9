# cascaded parity tests + kernel-only bench (one gpurun call)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_hlif.py tests/test_gpu_reference.py -m gpu -x -q -k "cascaded or Cascaded" > gpurun_out/casc_tests.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/casc_tests.log
for ch in 65536 4096; do
timeout 300 python bench.py --codec cascaded --chunk $ch --steps 5 --warmup 3 --no-e2e --no-cpu 2>gpurun_out/casc_bench.err | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('chunk $ch', 'comp', round(d['compress_GBps'],1), 'dec', round(d['decompress_GBps'],1), 'ratio', round(d['compression_ratio'],3), 'frac', round(d['roofline_compress']['frac'],3), round(d['roofline_decompress']['frac'],3))"
done
