#!/bin/bash
# round 2, GPU call B: streaming route -- smoke, parity, bench, ncu
mkdir -p gpurun_out
timeout 120 python tools/passbench.py --workload K5 --masks default --reps 3 > gpurun_out/r2b_smoke_k5.log 2>&1
echo "smoke rc=$?" >> gpurun_out/r2b_smoke_k5.log
timeout 600 python -m pytest tests/test_gpu_stream.py -m gpu -q -x > gpurun_out/r2b_stream_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2b_stream_tests.log
for ring in 4 6 8 12; do
  PF_RING=$ring timeout 120 python tools/passbench.py --workload K3 --masks default --reps 5 >> gpurun_out/r2b_passbench.log 2>&1
done
PF_STREAM=0 timeout 120 python tools/passbench.py --workload K3 --masks default --reps 5 >> gpurun_out/r2b_passbench.log 2>&1
for wl in K2 K5; do timeout 120 python tools/passbench.py --workload $wl --masks default --reps 5 >> gpurun_out/r2b_passbench.log 2>&1; PF_STREAM=0 timeout 120 python tools/passbench.py --workload $wl --masks default --reps 5 >> gpurun_out/r2b_passbench.log 2>&1; done
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err
timeout 120 python tools/passbench.py --workload K3 --masks default --reps 2 > gpurun_out/r2b_pb_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'kp_ab|kp_c2rt' -s 4 -c 2 -o gpurun_out/r2b_stream python tools/passbench.py --workload K3 --masks default --reps 2 > gpurun_out/r2b_ncu.log 2>&1
timeout 1200 python -m pytest tests -m gpu -q --maxfail=40 > gpurun_out/r2b_gpu_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2b_gpu_tests.log
ls -la gpurun_out
