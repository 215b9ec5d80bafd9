mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_hlif.py tests/test_gpu_reference.py tests/test_gpu_primitives.py -m gpu -x -q > gpurun_out/casc_tests.log 2>&1; echo "pytest rc=$?"
tail -2 gpurun_out/casc_tests.log
run() { timeout 300 python bench.py --codec cascaded --steps 3 --warmup 3 --no-e2e --no-cpu "$@" 2>>gpurun_out/casc_cls.err | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$*', 'comp', round(d['compress_GBps'],1), 'dec', round(d['decompress_GBps'],1), 'ratio', round(d['compression_ratio'],3))"; }
run
run --size-mib 1024 --casc-opts 2,1,1
run --size-mib 1024 --data zeros --casc-opts 2,1,1
run --size-mib 1024 --casc-opts 3,1,1
run --size-mib 1024 --data random --casc-opts 2,1,1
timeout 300 python scripts/bench_hlif.py 1024 2>&1 | tail -1
