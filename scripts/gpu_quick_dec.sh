# decoder parity tests + lz4/snappy bench (no e2e, no cpu)
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_decoder_groups.py -x -q -m gpu 2>&1 | tail -3
for c in lz4 snappy; do
timeout 300 python bench.py --codec $c --steps 5 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$c', 'comp', round(d['compress_GBps'],1), 'decomp', round(d['decompress_GBps'],1), 'ratio', round(d['compression_ratio'],3))"
done
