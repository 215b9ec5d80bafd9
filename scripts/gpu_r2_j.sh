set -x
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'compress_team' -s 3 -c 1 -o gpurun_out/r2j_prof_lz4comp python bench.py --codec lz4 --size-mib 256 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2j_ncu.log 2>&1
