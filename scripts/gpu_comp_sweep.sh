for n in 2 3 4 6 8 10 16; do
B200COMP_LZ4_COMP_CTAS=$n timeout 300 python bench.py --codec lz4 --steps 3 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ctas/sm $n', 'comp', round(d['compress_GBps'],1), 'decomp', round(d['decompress_GBps'],1), 'ratio', round(d['compression_ratio'],3))"
done
