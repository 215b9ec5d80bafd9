# last check of the committed library: full GPU suite, smoke(), the default bench line, the reference arm
set -x
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r2c_tests.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2c_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2c_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2c_smoke.log
( time timeout 900 python bench.py > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err ) 2>&1 | grep real; echo "bench rc=$?"
( time timeout 900 python bench.py --impl reference > gpurun_out/r2c_reference.json 2> gpurun_out/r2c_reference.err ) 2>&1 | grep real
