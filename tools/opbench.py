#!/usr/bin/env python
"""Device-resident timing of the covariance operators on workload K5 of
SURVEY.md section 8 (3D 256^3 cells, exponential, 512^3 torus): multiply by the
covariance matrix, approximate square root, inverse on the extended domain, and
the PCG inverse (RandomField::timesInverseMatrix).  One JSON line per operator
with the fraction of the HBM roofline (algorithmic bytes of SURVEY 8(d):
3 638 040 584 per matvec at K5).  Development / measurement tool.

  python tools/opbench.py [--workload K5] [--reps 20] [--cg 10]
"""
import argparse
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def matvec_bytes(cells, factor=2):
    c = list(cells)
    n = [factor * x for x in c]
    m0 = n[0] // 2 + 1
    if len(c) == 3:
        lam = 8 * (n[0] // 2 + 1) * (n[1] // 2 + 1) * (n[2] // 2 + 1)
        f = 8 * c[0] * c[1] * c[2]
        u = 16 * m0 * c[1] * c[2]
        v = 16 * m0 * n[1] * c[2]
        return 2 * f + 4 * u + 4 * v + lam
    lam = 8 * (n[0] // 2 + 1) * (n[1] // 2 + 1)
    f = 8 * c[0] * c[1]
    u = 16 * m0 * c[1]
    return 2 * f + 4 * u + lam


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="K5")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--cg", type=int, default=10)
    args = ap.parse_args()
    import torch

    import bench

    pfb = importlib.import_module("dune-parafields_b200")
    wl = {k: v for k, v in bench.WORKLOADS[args.workload].items() if k not in ("label", "transform")}
    peaks, _ = bench.measured_peaks()
    peak = float(peaks["hbm_gbs"])
    nbytes = matvec_bytes(wl["cells"])
    with pfb.Plan(device=0, cg_iterations=args.cg, **wl) as plan:
        plan.compute_eigenvalues()
        x = torch.empty(plan.domain_size, dtype=torch.float64, device="cuda")
        y = torch.empty_like(x)
        plan.generate(1, 0, out=x)
        plan.synchronize()
        for op in ("matrix", "root", "inverse_ext"):
            for _ in range(3):
                plan.apply(op, x, out=y)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(args.reps):
                plan.apply(op, x, out=y)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.reps
            print(json.dumps(dict(workload=args.workload, op=op, ms=round(ms, 4), per_s=round(1e3 / ms, 1), algorithmic_bytes=nbytes, achieved_gbs=round(nbytes / ms / 1e6, 1), frac=round(nbytes / ms / 1e6 / peak, 3))), flush=True)
        # PCG with a pinned iteration count (2 matvecs + sweeps per iteration)
        plan.solve(x, out=y)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        plan.solve(x, out=y)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        its = max(1, plan.cg_iterations_used)
        print(json.dumps(dict(workload=args.workload, op="solve (PCG)", iterations=its, ms=round(ms, 3), ms_per_iteration=round(ms / its, 3))), flush=True)


if __name__ == "__main__":
    main()
