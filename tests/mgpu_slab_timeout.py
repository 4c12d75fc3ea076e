#!/usr/bin/env python
"""A rank that falls behind its peers must not hang them: the wait of a slab
plan is bounded (PF_SLAB_TIMEOUT_S) and a timeout is FATAL for the plan --
the call fails, and so does every later call on that plan.
Two ranks: both generate once (fine), then rank 1 sleeps past the limit while
rank 0 issues its next call.  Launched by tests/test_multigpu.py under
torch.distributed.run with PF_SLAB_TIMEOUT_S = 2."""
import importlib
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist

    rank = int(os.environ["RANK"])
    world = int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("gloo")
    pfb = importlib.import_module("dune-parafields_b200")
    limit = float(os.environ["PF_SLAB_TIMEOUT_S"])
    ok = True
    with pfb.Plan([64, 64, 64], covariance="exponential", corr_length=0.1, device=local, n_ranks=world, rank=rank) as slab:
        slab.connect()
        slab.generate(seed=3, realization=0)  # both ranks in step: works
        slab.synchronize()
        dist.barrier()
        if rank == 1:
            time.sleep(3 * limit)  # falls behind by more than the limit
        t0 = time.perf_counter()
        failed, msg = False, ""
        try:
            slab.generate(seed=3, realization=1)
            slab.synchronize()
        except pfb.PfError as e:
            failed, msg = True, str(e)
        dt = time.perf_counter() - t0
        print("rank %d: second call %s after %.2f s: %s" % (rank, "FAILED" if failed else "returned", dt, msg[:120]), flush=True)
        if rank == 0:
            # the rank that waited: bounded wait, loud failure, plan stays failed
            ok = failed and "PF_SLAB_TIMEOUT_S" in msg and dt < limit + 5.0
            again = False
            try:
                slab.generate(seed=3, realization=2)
            except pfb.PfError:
                again = True
            ok = ok and again
            print("rank 0: a later call on the same plan %s" % ("fails too" if again else "WORKED (timeout not fatal)"), flush=True)
        # nobody frees its exchange buffers while the late rank still stores to them
        torch.cuda.synchronize()
        dist.barrier()
    flag = torch.tensor([1 if ok else 0])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("SLAB TIMEOUT %s" % ("OK" if int(flag) else "FAILED"), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(flag) else 1)


if __name__ == "__main__":
    main()
