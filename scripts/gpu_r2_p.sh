# full capture of the Cascaded kernels with the library's default option set (num_RLEs 2, num_deltas 1, bit-packed)
python bench.py --codec cascaded --casc-opts 2,1,1 --size-mib 1024 --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2p_casc211.json 2> gpurun_out/r2p_casc211.err || exit 1
tail -c 600 gpurun_out/r2p_casc211.json
ncu --set full --clock-control none --import-source on -k regex:'compress_kernel|decompress_kernel' -s 18 -c 6 -f -o gpurun_out/prof_casc211 python bench.py --codec cascaded --casc-opts 2,1,1 --size-mib 1024 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_casc211.log 2>&1
tail -2 gpurun_out/ncu_casc211.log
