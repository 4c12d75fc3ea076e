#!/bin/bash
# round 2, GPU call M (N GPUs): streaming-route tests on one GPU, slab parity
# over N ranks, bench at N GPUs (streaming route) and the exchange route beside it
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_stream.py -m gpu -q > gpurun_out/r2m_stream_tests.log 2>&1
echo "stream tests rc=$?" >> gpurun_out/r2m_stream_tests.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29711 tests/mgpu_slab_check.py > gpurun_out/r2m_${N}gpu_parity.log 2>&1
echo "parity rc=$?" >> gpurun_out/r2m_${N}gpu_parity.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29712 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r2m_bench_${N}gpu.json 2> gpurun_out/r2m_bench_${N}gpu.err
echo "bench rc=$?" >> gpurun_out/r2m_bench_${N}gpu.err
tail -n 2 gpurun_out/r2m_stream_tests.log gpurun_out/r2m_${N}gpu_parity.log gpurun_out/r2m_bench_${N}gpu.err
