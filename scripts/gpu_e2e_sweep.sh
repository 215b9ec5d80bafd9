# e2e pipeline depth sweep (LZ4 C2)
for S in 1 4 8 16; do
  timeout 300 python bench.py --codec lz4 --steps 3 --warmup 3 --no-cpu --e2e-slices $S 2>gpurun_out/e2e_$S.err | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('slices $S', d['value'], d['e2e'])"
done 2>&1 | tee gpurun_out/e2e_sweep.log
tail -3 gpurun_out/e2e_1.err
