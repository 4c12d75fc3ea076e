#!/bin/bash
# round 2, GPU call D (2 GPUs): streaming multi-GPU route -- parity, bench; kp_gen3 with shared twiddles on one GPU
mkdir -p gpurun_out
timeout 120 python tools/passbench.py --workload K3 --masks default --reps 5 > gpurun_out/r2d_passbench_1gpu.log 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/mgpu_slab_check.py > gpurun_out/r2d_2gpu_parity.log 2>&1
echo "parity rc=$?" >> gpurun_out/r2d_2gpu_parity.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r2d_bench_2gpu.json 2> gpurun_out/r2d_bench_2gpu.err
PF_STREAM=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r2d_bench_2gpu_exchange.json 2> gpurun_out/r2d_bench_2gpu_exchange.err
tail -n 3 gpurun_out/r2d_2gpu_parity.log
