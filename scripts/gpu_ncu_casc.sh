python bench.py --codec cascaded --size-mib 512 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain_casc.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'compress_kernel|decompress_kernel' -s 15 -c 2 -f -o gpurun_out/prof_casc_r1b python bench.py --codec cascaded --size-mib 512 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_casc.log 2>&1
tail -3 gpurun_out/ncu_casc.log
