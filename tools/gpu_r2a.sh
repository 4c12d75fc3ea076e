#!/bin/bash
# round 2, GPU call A: whole GPU suite, bench (both arms), ncu source-level
# capture of pass A.  Run from the repo root on the GPU box.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/r2a_smi.txt 2>&1
free -g > gpurun_out/r2a_host.txt; nproc >> gpurun_out/r2a_host.txt
python -m pytest tests -m gpu -q --maxfail=30 -x --deselect tests/test_gpu_fullsize.py > gpurun_out/r2a_gpu_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2a_gpu_tests.log
python bench.py --steps 20 --warmup 3 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err
echo "bench rc=$?" >> gpurun_out/r2a_bench.err
python tools/passbench.py --workload K3 --masks default --reps 2 > gpurun_out/r2a_passbench.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:kp_gen3 -s 2 -c 1 -o gpurun_out/r2a_gen3 python tools/passbench.py --workload K3 --masks default --reps 2 > gpurun_out/r2a_ncu.log 2>&1
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2a_bench_ref.json 2> gpurun_out/r2a_bench_ref.err
python -m pytest tests/test_gpu_fullsize.py -m gpu -q > gpurun_out/r2a_fullsize.log 2>&1
ls -la gpurun_out
