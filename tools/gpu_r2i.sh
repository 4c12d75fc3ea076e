#!/bin/bash
# round 2, GPU call I: whole GPU suite + both bench arms + launch list of the bench command
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q --maxfail=40 > gpurun_out/r2i_gpu_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2i_gpu_tests.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r2i_bench.json 2> gpurun_out/r2i_bench.err
rc=$?; echo "bench rc=$rc" >> gpurun_out/r2i_bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2i_bench_reference.json 2> gpurun_out/r2i_bench_reference.err
echo "ref rc=$?" >> gpurun_out/r2i_bench_reference.err
if [ $rc -eq 0 ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file gpurun_out/r2i_launches.csv python bench.py --steps 3 --warmup 3 --no-others --no-fftlib --no-e2e --no-cpu-baseline \
    > gpurun_out/r2i_ncu.log 2>&1
fi
tail -n 3 gpurun_out/r2i_gpu_tests.log
