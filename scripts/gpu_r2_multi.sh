# N-GPU records: default bench, C4 (strong scaling through the sharder), C5 sweep, reference arm
N=${1:-2}
set -x
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2m_bench_${N}gpu.json 2> gpurun_out/r2m_bench_${N}gpu.err; echo "bench rc=$?"; tail -3 gpurun_out/r2m_bench_${N}gpu.err | cut -c1-300
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --workload c4 --gpus $N --steps 5 --warmup 3 > gpurun_out/r2m_c4_${N}gpu.json 2> gpurun_out/r2m_c4_${N}gpu.err; echo "c4 rc=$?"; tail -3 gpurun_out/r2m_c4_${N}gpu.err | cut -c1-300
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --workload c5 --gpus $N --steps 2 > gpurun_out/r2m_c5_${N}gpu.jsonl 2> gpurun_out/r2m_c5_${N}gpu.err; echo "c5 rc=$?"; tail -3 gpurun_out/r2m_c5_${N}gpu.err | cut -c1-300
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus $N --steps 2 --warmup 1 --codec lz4 > gpurun_out/r2m_reference_${N}gpu.json 2> gpurun_out/r2m_reference_${N}gpu.err; echo "ref rc=$?"
