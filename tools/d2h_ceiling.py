"""Platform ceiling for the end-to-end figure: concurrent device -> pinned host
copies from every rank (plain cudaMemcpyAsync through torch, none of this
repository's code), to set beside bench.py's e2e d2h_gbs_per_gpu in slab mode.
Launch: python -m torch.distributed.run --nproc-per-node N tools/d2h_ceiling.py"""
import json
import os
import time

import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", 0))
world = int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
if world > 1:
    dist.init_process_group("gloo")
nbytes = 128 << 20  # one rank's share of a K3 realization at 8 GPUs
src = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
dst = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
for _ in range(3):
    dst.copy_(src, non_blocking=True)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
reps = 40
t0 = time.perf_counter()
for _ in range(reps):
    dst.copy_(src, non_blocking=True)
torch.cuda.synchronize()
dt = time.perf_counter() - t0
gbs = torch.tensor([nbytes * reps / dt / 1e9], dtype=torch.float64)
if world > 1:
    all_ = [torch.zeros(1, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(all_, gbs)
else:
    all_ = [gbs]
if rank == 0:
    v = [float(x) for x in all_]
    print(json.dumps({"what": "concurrent cudaMemcpyAsync device -> pinned host, 128 MiB x %d per rank" % reps,
                      "ranks": world, "gbs_per_gpu_min": min(v), "gbs_per_gpu_max": max(v), "gbs_total": sum(v)}))
