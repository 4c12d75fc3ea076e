#!/bin/bash
# development aid: build a variant of the N=1024 kernels (PF_VAR_N="512 1024" for more) with extra -D flags and
# link it with the other objects into csrc/variants/libpf_<name>.so
# (select with PF_LIB=...).  usage: tools/variant.sh name "-DPF_X=1 ..."
set -e
cd "$(dirname "$0")/../dune-parafields_b200/csrc"
name=$1; shift
mkdir -p variants build/var_$name
ns=${PF_VAR_N:-1024}
skip=""
for n in $ns; do
  nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Wno-deprecated-gpu-targets -DPF_N=$n "$@" -c pf_inst.cu -o build/var_$name/pf_inst_$n.o &
  skip="$skip -e pf_inst_$n.o"
done
wait
objs=$(ls build/pf_inst_*.o | grep -v $skip)
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o variants/libpf_$name.so build/var_$name/pf_inst_*.o $objs build/pf_api.o
echo built variants/libpf_$name.so
