#!/bin/bash
# development aid: build a variant of the N=1024 kernels with extra -D flags and
# link it with the other objects into csrc/variants/libpf_<name>.so
# (select with PF_LIB=...).  usage: tools/variant.sh name "-DPF_X=1 ..."
set -e
cd "$(dirname "$0")/../dune-parafields_b200/csrc"
name=$1; shift
mkdir -p variants build/var_$name
nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Wno-deprecated-gpu-targets -DPF_N=1024 "$@" -c pf_inst.cu -o build/var_$name/pf_inst_1024.o
objs=$(ls build/pf_inst_*.o | grep -v pf_inst_1024.o)
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o variants/libpf_$name.so build/var_$name/pf_inst_1024.o $objs build/pf_api.o
echo built variants/libpf_$name.so
