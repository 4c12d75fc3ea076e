#!/bin/bash
# round 2, GPU call R: ncu --set full capture (with source) of the three K3
# passes in their current state
mkdir -p gpurun_out
python tools/passbench.py --workload K3 --masks default --reps 2 > gpurun_out/r2r_passbench.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:kp_gen3' -s 1 -c 1 -o gpurun_out/r2r_k3_passes python tools/passbench.py --workload K3 --masks default --reps 2 > gpurun_out/r2r_ncu.log 2>&1
tail -n 5 gpurun_out/r2r_passbench.log; tail -n 3 gpurun_out/r2r_ncu.log; ls -la gpurun_out
