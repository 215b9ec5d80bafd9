# usage: gpu_lz4_iter.sh <tag>   — LZ4 parity tests, quick C2 bench line, ncu capture of the team encoder
tag=$1
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "lz4" 2>&1 | tail -5
timeout 200 python bench.py --no-e2e --no-cpu --steps 5 --warmup 3 2>gpurun_out/b_$tag.err | tee gpurun_out/b_$tag.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('value','compress_GBps','decompress_GBps','compression_ratio')})"
sh scripts/gpu_ncu_kernel.sh lz4 compress_team $tag 256 3
