timeout 900 python -m pytest tests/test_gpu_decoder_groups.py tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/gpu_tests_variants.log 2>&1; tail -5 gpurun_out/gpu_tests_variants.log
for v in round pipe; do
for c in lz4 snappy; do
  B200COMP_LZ4_DEC_G=$v B200COMP_SNAPPY_DEC_G=$v timeout 300 python bench.py --codec $c --steps 5 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$c $v', d['compress_GBps'], d['decompress_GBps'])"
done; done 2>&1 | tee gpurun_out/dec_variants.log
