#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_stream.py -m gpu -q -x > gpurun_out/r2c_stream_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2c_stream_tests.log
for wl in K3 K5; do
  timeout 120 python tools/passbench.py --workload $wl --masks default --reps 5 >> gpurun_out/r2c_passbench.log 2>&1
  PF_ZT=1 timeout 120 python tools/passbench.py --workload $wl --masks default --reps 5 >> gpurun_out/r2c_passbench.log 2>&1
  PF_LAG=2 PF_RING=5 timeout 120 python tools/passbench.py --workload $wl --masks default --reps 5 >> gpurun_out/r2c_passbench.log 2>&1
  PF_STREAM=0 timeout 120 python tools/passbench.py --workload $wl --masks default --reps 5 >> gpurun_out/r2c_passbench.log 2>&1
done
timeout 120 python tools/passbench.py --workload K3 --masks default --reps 2 > gpurun_out/r2c_pb_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'kp_ab|kp_c2rt' -s 4 -c 2 -o gpurun_out/r2c_stream python tools/passbench.py --workload K3 --masks default --reps 2 > gpurun_out/r2c_ncu.log 2>&1
timeout 900 python -m pytest tests/test_gpu_noise.py -m gpu -q --maxfail=40 > gpurun_out/r2c_noise_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2c_noise_tests.log
