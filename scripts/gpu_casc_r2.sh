mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_hlif.py tests/test_gpu_reference.py -m gpu -x -q -k "cascaded or Cascaded" > gpurun_out/casc_tests.log 2>&1; echo "pytest rc=$?"
tail -2 gpurun_out/casc_tests.log
run() { timeout 300 python bench.py --codec cascaded --size-mib 1024 --steps 3 --warmup 3 --no-e2e --no-cpu "$@" 2>>gpurun_out/casc_cls.err | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$*', 'comp', round(d['compress_GBps'],1), 'dec', round(d['decompress_GBps'],1), 'ratio', round(d['compression_ratio'],3))"; }
run --data ints
run --data ints --casc-opts 2,1,1
run --data ints --casc-opts 2,1,0
run --data zeros --casc-opts 2,1,1
