# closing run for the final library: full GPU tests, the three bench lines, HLIF bench, cascaded option sets
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "pytest rc=$?"
tail -2 gpurun_out/gpu_tests.log
for c in lz4 snappy cascaded; do
timeout 600 python bench.py --codec $c --steps 5 --warmup 3 > gpurun_out/bench_${c}.json 2> gpurun_out/bench_${c}.err; echo "bench $c rc=$?"
python -c "import json; d=json.loads(open('gpurun_out/bench_${c}.json').read().strip().splitlines()[-1]); print('$c', round(d['value'],1), 'comp', round(d['compress_GBps'],1), 'dec', round(d['decompress_GBps'],1), 'e2e', round(d['e2e']['value'],2), 'frac', round(d['roofline']['frac'],4), 'traffic', d['roofline'].get('traffic'))"
done
timeout 300 python bench.py --codec cascaded --size-mib 1024 --casc-opts 2,1,1 --steps 3 --no-e2e --no-cpu 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('casc 2,1,1 comp', round(d['compress_GBps'],1), 'dec', round(d['decompress_GBps'],1))"
timeout 300 python scripts/bench_hlif.py 1024 > gpurun_out/hlif.log 2>&1; tail -4 gpurun_out/hlif.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
