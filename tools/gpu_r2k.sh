#!/bin/bash
# round 2, GPU call K: ncu --set full capture (with source) of the three K3
# passes in their current state
mkdir -p gpurun_out
python tools/passbench.py --workload K3 --masks default --reps 2 > gpurun_out/r2k_passbench.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:kp_(gen3|c2t|c2r)' -s 3 -c 3 -o gpurun_out/r2k_k3_passes python tools/passbench.py --workload K3 --masks default --reps 2 > gpurun_out/r2k_ncu.log 2>&1
tail -n 5 gpurun_out/r2k_passbench.log; tail -n 3 gpurun_out/r2k_ncu.log; ls -la gpurun_out
