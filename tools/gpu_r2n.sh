#!/bin/bash
mkdir -p gpurun_out
for n in 1 2 4 8; do
  timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29800+n)) tools/d2h_ceiling.py 2>/dev/null | grep gbs >> gpurun_out/r2n_d2h_ceiling.jsonl
done
nvidia-smi topo -m > gpurun_out/r2n_topo.txt 2>&1
cat gpurun_out/r2n_d2h_ceiling.jsonl
