#!/bin/bash
# round 2, GPU call F (N GPUs): streaming multi-GPU route -- parity, bench, exchange route beside it
N=${1:-8}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 tests/mgpu_slab_check.py > gpurun_out/r2f_${N}gpu_parity.log 2>&1
echo "parity rc=$?" >> gpurun_out/r2f_${N}gpu_parity.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r2f_bench_${N}gpu.json 2> gpurun_out/r2f_bench_${N}gpu.err
PF_STREAM=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus $N --steps 20 --warmup 3 --no-e2e > gpurun_out/r2f_bench_${N}gpu_exchange.json 2> gpurun_out/r2f_bench_${N}gpu_exchange.err
tail -n 3 gpurun_out/r2f_${N}gpu_parity.log
