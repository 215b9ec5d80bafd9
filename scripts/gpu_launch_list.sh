# launch list of the default bench command (all launches of the process; data generation by torch comes first)
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/plain_default.log 2>&1 || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r01z_launches_lz4.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_launch.log 2>&1; echo "ncu list rc=$?"
grep -c "compress_kernel\|decompress_window" gpurun_out/r01z_launches_lz4.csv
