# usage: gpu_ncu_kernel.sh <codec> <kernel-regex> <out-name> [size-mib] [skip]
# full ncu capture (with source) of one launch of a kernel inside bench.py
c=$1; k=$2; o=$3; sz=${4:-256}; skip=${5:-3}
python bench.py --codec $c --size-mib $sz --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain_$o.log 2>&1 || { tail -5 gpurun_out/plain_$o.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"$k" -s $skip -c 1 -f -o gpurun_out/$o python bench.py --codec $c --size-mib $sz --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_$o.log 2>&1
tail -2 gpurun_out/ncu_$o.log
