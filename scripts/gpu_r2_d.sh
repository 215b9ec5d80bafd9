# new bench: default line (3 codecs), reference arm, C4 (small then full), refgpu lines
set -x
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; echo "bench rc=$?"; tail -5 gpurun_out/r2d_bench.err
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2d_reference.json 2> gpurun_out/r2d_reference.err; echo "ref rc=$?"; tail -5 gpurun_out/r2d_reference.err
timeout 600 python bench.py --workload c4 --size-mib 2048 --steps 5 --warmup 3 > gpurun_out/r2d_c4_small.json 2> gpurun_out/r2d_c4_small.err; echo "c4 small rc=$?"; tail -5 gpurun_out/r2d_c4_small.err
timeout 900 python bench.py --workload c4 --steps 5 --warmup 3 > gpurun_out/r2d_c4.json 2> gpurun_out/r2d_c4.err; echo "c4 rc=$?"; tail -5 gpurun_out/r2d_c4.err
for c in lz4 snappy cascaded; do
timeout 600 python bench.py --impl refgpu --codec $c --steps 3 --warmup 3 > gpurun_out/r2d_refgpu_$c.json 2> gpurun_out/r2d_refgpu_$c.err; echo "refgpu $c rc=$?"; tail -3 gpurun_out/r2d_refgpu_$c.err
done
