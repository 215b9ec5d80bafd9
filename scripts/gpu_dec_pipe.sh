timeout 900 python -m pytest tests/test_gpu_decoder_groups.py tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/gpu_tests_pipe.log 2>&1; tail -5 gpurun_out/gpu_tests_pipe.log
for c in lz4 snappy; do
  timeout 300 python bench.py --codec $c --steps 5 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$c pipe', d['compress_GBps'], d['decompress_GBps'])"
done 2>&1 | tee gpurun_out/dec_pipe.log
