# size-walk kernel + manager waves: tests, C4 line (with the size query timed), HLIF lines
set -x
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_decoders.py tests/test_gpu_hlif.py tests/test_gpu_hlif_interop.py tests/test_gpu_reference.py -m gpu -x -q > gpurun_out/r2m_tests.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2m_tests.log
timeout 900 python bench.py --workload c4 > gpurun_out/r2z_c4_1gpu.json 2> gpurun_out/r2z_c4.err; echo "c4 rc=$?"; tail -3 gpurun_out/r2z_c4.err
timeout 600 python scripts/bench_hlif.py 1024 > gpurun_out/r2z_hlif.jsonl 2> gpurun_out/r2z_hlif.err; echo "hlif rc=$?"
