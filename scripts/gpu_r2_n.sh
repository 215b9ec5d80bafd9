set -x
timeout 600 ncu --set full --clock-control none --import-source on -k regex:decompress_window -s 3 -c 1 -o gpurun_out/r2n_prof_snappydec python bench.py --codec snappy --size-mib 256 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2n_ncu.log 2>&1
