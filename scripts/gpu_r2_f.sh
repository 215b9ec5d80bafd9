# full GPU suite, C5 sweep at N=1, cascaded element types / option sets
set -x
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2f_tests.log 2>&1; echo "pytest rc=$?"
tail -8 gpurun_out/r2f_tests.log
timeout 1200 python bench.py --workload c5 --steps 2 > gpurun_out/r2f_c5_1gpu.jsonl 2> gpurun_out/r2f_c5.err; echo "c5 rc=$?"; tail -5 gpurun_out/r2f_c5.err
: > gpurun_out/r2f_casc_variants.jsonl
for t in char short longlong; do
timeout 300 python bench.py --codec cascaded --casc-type $t --size-mib 1024 --steps 5 --warmup 3 --no-e2e --no-cpu >> gpurun_out/r2f_casc_variants.jsonl 2>> gpurun_out/r2f_casc.err; echo "casc $t rc=$?"
done
for o in 2,1,1 3,1,1 1,0,1 0,1,1 0,0,1 1,1,0; do
timeout 300 python bench.py --codec cascaded --casc-opts $o --size-mib 1024 --steps 5 --warmup 3 --no-e2e --no-cpu >> gpurun_out/r2f_casc_variants.jsonl 2>> gpurun_out/r2f_casc.err; echo "casc $o rc=$?"
done
tail -5 gpurun_out/r2f_casc.err
