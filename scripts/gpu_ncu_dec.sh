# instruction counts + durations of the LZ4 / Snappy decoder variants (256 MiB), and a full capture of the round decoder
for v in round pipe 8; do
for c in lz4 snappy; do
B200COMP_LZ4_DEC_G=$v B200COMP_SNAPPY_DEC_G=$v ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,sm__inst_issued.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:'decompress' -s 3 -c 1 --csv python bench.py --codec $c --size-mib 256 --steps 1 --warmup 3 --no-e2e --no-cpu 2>/dev/null | grep -E '"decompress|"void' | awk -F'","' -v t="$c $v" '{print t, $(NF-2), $(NF)}'
done; done > gpurun_out/dec_inst_counts.txt 2>&1
cat gpurun_out/dec_inst_counts.txt
ncu --set full --clock-control none --import-source on -k regex:'decompress_round_kernel' -s 3 -c 1 -f -o gpurun_out/prof_lz4_round2 python bench.py --codec lz4 --size-mib 256 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_lz4c.log 2>&1
tail -2 gpurun_out/ncu_lz4c.log
