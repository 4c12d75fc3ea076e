#!/bin/bash
# round 2, GPU call P: ncu --set full of the fused first-two-axes kernel (streaming route on one GPU)
mkdir -p gpurun_out
PF_STREAM=1 python tools/passbench.py --workload K3 --masks default --reps 2 > gpurun_out/r2p_passbench.log 2>&1 &&
PF_STREAM=1 ncu --set full --clock-control none --import-source on -k 'regex:kp_ab' -s 1 -c 1 -o gpurun_out/r2p_kp_ab python tools/passbench.py --workload K3 --masks default --reps 2 > gpurun_out/r2p_ncu.log 2>&1
tail -n 1 gpurun_out/r2p_passbench.log; tail -n 2 gpurun_out/r2p_ncu.log
