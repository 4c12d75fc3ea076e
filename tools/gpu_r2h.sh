#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r2h_variants.log
: > $L
TWS=$PWD/dune-parafields_b200/csrc/libparafields_cuda_tws.so
for wl in K3 K2 K5; do
  echo "== $wl default" >> $L; timeout 120 python tools/passbench.py --workload $wl --masks default --reps 5 >> $L 2>&1
  echo "== $wl tws" >> $L; PF_LIB=$TWS timeout 120 python tools/passbench.py --workload $wl --masks default --reps 5 >> $L 2>&1
done
for wl in K2 K5; do
  echo "== $wl gen3 force" >> $L; PF_GEN3=force timeout 120 python tools/passbench.py --workload $wl --masks default --reps 5 >> $L 2>&1
  echo "== $wl gen3 force + tws" >> $L; PF_LIB=$TWS PF_GEN3=force timeout 120 python tools/passbench.py --workload $wl --masks default --reps 5 >> $L 2>&1
done
echo "== opbench default" >> $L; timeout 200 python tools/opbench.py >> $L 2>&1
echo "== opbench tws" >> $L; PF_LIB=$TWS timeout 200 python tools/opbench.py >> $L 2>&1
