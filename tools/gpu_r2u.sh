#!/bin/bash
# round 2, GPU call U: ncu --set full summaries of K2's passes (N = 512 kernels)
mkdir -p gpurun_out
python tools/passbench.py --workload K2 --masks default --reps 2 > gpurun_out/r2u_passbench.log 2>&1 &&
ncu --set full --clock-control none -k 'regex:k_gen_tile|kp_c2t|kp_c2r' -s 3 -c 3 -o gpurun_out/r2u_k2 python tools/passbench.py --workload K2 --masks default --reps 2 > gpurun_out/r2u_ncu.log 2>&1
tail -n 1 gpurun_out/r2u_passbench.log; tail -n 2 gpurun_out/r2u_ncu.log
