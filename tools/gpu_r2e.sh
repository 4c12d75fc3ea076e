#!/bin/bash
# round 2, GPU call E: whole GPU suite on the current state + bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --maxfail=40 > gpurun_out/r2e_gpu_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2e_gpu_tests.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench.err
echo "bench rc=$?" >> gpurun_out/r2e_bench.err
