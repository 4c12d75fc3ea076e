#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
timeout 300 oracle/_ref/dropin_test > gpurun_out/r2g_dropin.log 2>&1; echo "rc=$?" >> gpurun_out/r2g_dropin.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29711 tests/mgpu_slab_check.py > gpurun_out/r2g_${N}gpu_parity.log 2>&1
echo "parity rc=$?" >> gpurun_out/r2g_${N}gpu_parity.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29712 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r2g_bench_${N}gpu.json 2> gpurun_out/r2g_bench_${N}gpu.err
