#!/usr/bin/env python
"""Multi-GPU parity of the slab mode: the P-GPU result must equal the 1-GPU
result on the same global noise (Philox counters are global frequency indices).
Launched by tests/test_multigpu.py as

  python -m torch.distributed.run --nproc-per-node P --master-addr 127.0.0.1 \
      --master-port <port> tests/mgpu_slab_check.py
"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CASES = [
    dict(cells=[64, 64, 64], covariance="exponential", corr_length=0.1),
    dict(cells=[128, 128, 128], covariance="spherical", corr_length=0.1),
    dict(cells=[32, 16, 8], covariance="gaussian", corr_length=[0.05, 0.03, 0.02], anisotropy="axiparallel", approximate=True),
    dict(cells=[256, 256], covariance="exponential", corr_length=0.05),
    dict(cells=[64, 32], covariance="cubic", corr_length=[0.08, 0.03, 0.03, 0.05], anisotropy="geometric", approximate=True),
    # 512^3 torus: the streaming multi-GPU route (every rank takes its planes of
    # k0 through both first axes, lines stored to the layer owners); the
    # operators keep the exchange route
    dict(cells=[256, 256, 256], covariance="exponential", corr_length=0.05),
    dict(cells=[256, 256, 256], covariance="exponential", corr_length=[0.1, 0.02, 0.0, 0.02, 0.2, 0.01, 0.0, 0.01, 0.15], anisotropy="geometric", approximate=True, transform="logNormal"),
]


def main():
    import torch
    import torch.distributed as dist

    rank = int(os.environ["RANK"])
    world = int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pfb = importlib.import_module("dune-parafields_b200")
    worst = 0.0
    cg_worst = 0.0  # scalar products are summed per rank first: rounding-level differences
    for case in CASES:
        if case["cells"][-1] % world:
            continue
        with pfb.Plan(device=local, **case) as single, pfb.Plan(device=local, n_ranks=world, rank=rank, **case) as slab:
            slab.connect()
            layers = case["cells"][-1] // world
            for r in range(5):  # > 2 calls: exercises the buffer-reuse flags
                full = single.generate(seed=11, realization=r)
                mine = slab.generate(seed=11, realization=r)
                slab.synchronize()
                ref = full[rank * layers : (rank + 1) * layers]
                err = float(np.max(np.abs(mine - ref)) / np.max(np.abs(full)))
                worst = max(worst, err)
            # pipelined batch on a device buffer
            nb = 4
            dev = torch.empty((nb, slab.local_domain_size), dtype=torch.float64, device="cuda")
            slab.generate_batch(11, 100, nb, out=dev)
            slab.synchronize()
            for r in range(nb):
                full = single.generate(seed=11, realization=100 + r)
                ref = full[rank * layers : (rank + 1) * layers].ravel()
                err = float(np.max(np.abs(dev[r].cpu().numpy() - ref)) / np.max(np.abs(full)))
                worst = max(worst, err)
            # injected (reference-order) noise takes the same route
            rng = np.random.default_rng(5)
            noise = rng.standard_normal(2 * single.noise_entries)
            full = single.generate_from_noise(noise)
            mine = slab.generate_from_noise(noise)
            err = float(np.max(np.abs(mine - full[rank * layers : (rank + 1) * layers])) / np.max(np.abs(full)))
            worst = max(worst, err)
            # slab -> block redistribution with overlap (dataToEval +
            # exchangeOverlap, stochastic.hh:830-925,1016-1127) and back
            cells = case["cells"]
            dim = len(cells)
            full = single.generate(seed=11, realization=3)
            mine = slab.generate(seed=11, realization=3)
            tried = []
            for ppd in (pfb.default_load_balance(cells, world), [1] * (dim - 1) + [world], ([world] + [1] * (dim - 1)) if cells[0] % world == 0 else None):
                if ppd is None or ppd in tried:
                    continue
                tried.append(ppd)
                le, lo = pfb.eval_block(cells, ppd, rank)
                ev, ov = slab.slab_to_blocks(ppd, mine)
                sl = tuple(slice(o, o + e) for o, e in zip(reversed(lo), reversed(le)))
                if not np.array_equal(ev, full[sl].ravel()):
                    worst = max(worst, 1.0)
                # faces: lower / upper periodic neighbour layer of every dimension
                pos = 0
                for i in range(dim):
                    ax = dim - 1 - i  # numpy axis of dimension i
                    for coord in ((lo[i] - 1) % cells[i], (lo[i] + le[i]) % cells[i]):
                        idx = list(sl)
                        idx[ax] = slice(coord, coord + 1)
                        face = np.squeeze(full[tuple(idx)], axis=ax)
                        # face index: e_next + e_nextnext * le[next] (3D), e_next (2D)
                        if dim == 3:
                            nx, nn = (i + 1) % 3, (i + 2) % 3
                            # `face` axes are the remaining dims in descending order
                            rem = [d for d in (2, 1, 0) if d != i]
                            want = np.transpose(face, (rem.index(nn), rem.index(nx))).ravel()
                        else:
                            want = face.ravel()
                        got = ov[pos : pos + want.size]
                        pos += want.size
                        if not np.array_equal(got, want):
                            worst = max(worst, 2.0)
                assert pos == ov.size
                back = slab.blocks_to_slab(ppd, ev)
                if not np.array_equal(back, mine):
                    worst = max(worst, 3.0)
            # covariance operators on the distributed field (two exchanges per
            # matvec) and the PCG inverse with cross-rank scalar products
            x_full = np.random.default_rng(9).standard_normal(single.shape)
            x_mine = np.ascontiguousarray(x_full[rank * layers : (rank + 1) * layers])
            for op in ("matrix", "root", "matrix"):
                y_full = single.apply(op, x_full)
                y_mine = slab.apply(op, x_mine)
                err = float(np.max(np.abs(y_mine - y_full[rank * layers : (rank + 1) * layers])) / np.max(np.abs(y_full)))
                worst = max(worst, err)
            mine = slab.generate(seed=11, realization=50)  # interleaves with the operator epochs
            full = single.generate(seed=11, realization=50)
            worst = max(worst, float(np.max(np.abs(mine - full[rank * layers : (rank + 1) * layers])) / np.max(np.abs(full))))
            dist.barrier()
        if not case.get("approximate"):
            kw = dict(case, cg_iterations=3)
            with pfb.Plan(device=local, **kw) as single, pfb.Plan(device=local, n_ranks=world, rank=rank, **kw) as slab:
                slab.connect()
                layers = case["cells"][-1] // world
                b_full = np.random.default_rng(4).standard_normal(single.shape)
                x_full = single.solve(b_full)
                x_mine = slab.solve(np.ascontiguousarray(b_full[rank * layers : (rank + 1) * layers]))
                err = float(np.max(np.abs(x_mine - x_full[rank * layers : (rank + 1) * layers])) / np.max(np.abs(x_full)))
                cg_worst = max(cg_worst, err)
                assert slab.cg_iterations_used == single.cg_iterations_used
                dist.barrier()
        if rank == 0:
            print("case %s ok so far, worst rel err %.3e" % ("x".join(map(str, case["cells"])), worst), flush=True)
    t = torch.tensor([worst, cg_worst], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print("slab PCG worst rel err %.3e" % t[1].item(), flush=True)
    assert t[1].item() < 1e-9
    t = t[:1]
    if rank == 0:
        print("SLAB PARITY worst rel err %.3e over %d ranks: %s" % (t.item(), world, "OK" if t.item() < 1e-12 else "FAIL"), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if t.item() < 1e-12 else 1)


if __name__ == "__main__":
    main()
