# Cascaded: tests, default bench line, option sets, C5 sweep at 1 GiB (incompressible class)
set -x
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_reference.py tests/test_gpu_hlif.py tests/test_gpu_cascaded_select.py tests/test_gpu_reference_suite.py tests/test_gpu_staging.py -m gpu -x -q -k "cascaded or Cascaded or casc" > gpurun_out/r2q_tests.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2q_tests.log
: > gpurun_out/r2q_casc.jsonl
timeout 300 python bench.py --codec cascaded --steps 5 --warmup 3 --no-e2e --no-cpu >> gpurun_out/r2q_casc.jsonl 2>> gpurun_out/r2q_casc.err
for o in 2,1,1 3,1,1 1,1,0; do timeout 300 python bench.py --codec cascaded --casc-opts $o --size-mib 1024 --steps 5 --warmup 3 --no-e2e --no-cpu >> gpurun_out/r2q_casc.jsonl 2>> gpurun_out/r2q_casc.err; done
timeout 600 python bench.py --workload c5 --c5-gib 1 --steps 2 --no-cpu > gpurun_out/r2q_c5_small.jsonl 2> gpurun_out/r2q_c5.err; echo "c5 rc=$?"
python - <<'PY'
import json
for f in ("gpurun_out/r2q_casc.jsonl", "gpurun_out/r2q_c5_small.jsonl"):
    for l in open(f):
        if l.startswith("{"):
            r = json.loads(l)
            if r.get("codec", "cascaded") != "cascaded": continue
            print(r.get("input_class") or r["config"]["workload"][-24:], round(r["compress_GBps"]), round(r["decompress_GBps"]), round(r["compression_ratio"], 2))
PY
