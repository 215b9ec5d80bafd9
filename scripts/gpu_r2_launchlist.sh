set -x
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'compress' -c 400 --csv --log-file gpurun_out/r2z_launches.csv python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/r2z_ncu_launches.log 2>&1; echo "launch list rc=$?"
