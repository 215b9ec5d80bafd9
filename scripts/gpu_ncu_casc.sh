# full capture of one timed step of the cascaded bench (1 compress + 5 decompress launches)
python bench.py --codec cascaded --size-mib 1024 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain_casc.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:'compress_kernel|decompress_kernel' -s 18 -c 6 -f -o gpurun_out/prof_casc python bench.py --codec cascaded --size-mib 1024 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_casc.log 2>&1
tail -2 gpurun_out/ncu_casc.log
