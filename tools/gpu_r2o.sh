#!/bin/bash
# round 2, GPU call O: ncu --set full of K4's kernels (long-line route)
mkdir -p gpurun_out
python tools/passbench.py --workload K4 --masks default --reps 2 > gpurun_out/r2o_passbench.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:k_long' -s 4 -c 3 -o gpurun_out/r2o_k4 python tools/passbench.py --workload K4 --masks default --reps 2 > gpurun_out/r2o_ncu.log 2>&1
tail -n 2 gpurun_out/r2o_passbench.log; tail -n 2 gpurun_out/r2o_ncu.log
