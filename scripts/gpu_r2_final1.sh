# final single-GPU records of the round: tests, smoke, default bench, reference arm, C4, C5, refgpu lines, launch list
set -x
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r2z_tests.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2z_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2z_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2z_smoke.log
timeout 900 python bench.py > gpurun_out/r2z_bench.json 2> gpurun_out/r2z_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r2z_bench.err
timeout 900 python bench.py --impl reference > gpurun_out/r2z_reference.json 2> gpurun_out/r2z_reference.err; echo "ref rc=$?"
timeout 900 python bench.py --workload c4 > gpurun_out/r2z_c4_1gpu.json 2> gpurun_out/r2z_c4.err; echo "c4 rc=$?"
timeout 1200 python bench.py --workload c5 --steps 2 > gpurun_out/r2z_c5_1gpu.jsonl 2> gpurun_out/r2z_c5.err; echo "c5 rc=$?"
timeout 600 python scripts/bench_hlif.py 1024 > gpurun_out/r2z_hlif.jsonl 2> gpurun_out/r2z_hlif.err; echo "hlif rc=$?"
: > gpurun_out/r2z_refgpu.jsonl
for c in lz4 snappy cascaded; do timeout 600 python bench.py --impl refgpu --codec $c --steps 3 --warmup 3 >> gpurun_out/r2z_refgpu.jsonl 2>> gpurun_out/r2z_refgpu.err; done
: > gpurun_out/r2z_casc_variants.jsonl
for t in char short longlong; do timeout 300 python bench.py --codec cascaded --casc-type $t --size-mib 1024 --steps 5 --warmup 3 --no-e2e --no-cpu >> gpurun_out/r2z_casc_variants.jsonl 2>> gpurun_out/r2z_casc.err; done
for o in 2,1,1 3,1,1 1,0,1 0,1,1 0,0,1 1,1,0; do timeout 300 python bench.py --codec cascaded --casc-opts $o --size-mib 1024 --steps 5 --warmup 3 --no-e2e --no-cpu >> gpurun_out/r2z_casc_variants.jsonl 2>> gpurun_out/r2z_casc.err; done
timeout 300 python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > /dev/null 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:compress -c 400 --csv --log-file gpurun_out/r2z_launches.csv python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/r2z_ncu_launches.log 2>&1; echo "launch list rc=$?"
