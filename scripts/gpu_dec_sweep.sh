# decoder group-width sweep (LZ4 C2 / Snappy C1), after the group-width tests
timeout 900 python -m pytest tests/test_gpu_decoder_groups.py tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/gpu_tests_groups.log 2>&1; tail -3 gpurun_out/gpu_tests_groups.log
for G in 4 8 16 32; do
  B200COMP_LZ4_DEC_G=$G timeout 300 python bench.py --codec lz4 --steps 5 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('lz4 G=$G', d['compress_GBps'], d['decompress_GBps'])"
  B200COMP_SNAPPY_DEC_G=$G timeout 300 python bench.py --codec snappy --steps 5 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('snappy G=$G', d['compress_GBps'], d['decompress_GBps'])"
done 2>&1 | tee gpurun_out/dec_sweep.log
