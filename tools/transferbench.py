#!/usr/bin/env python
"""Device-resident timing of the field transfer kernels (SURVEY 8(f) rows 1
and 3) at the cell counts of K5 / K3: refine 256^3 -> 512^3, coarsen
512^3 -> 256^3, localize 512^3 in place, point evaluation at every cell
centre of 512^3 from a location array, and the fused bulkEvaluate sweep
(trend + value transform).  One JSON line per kernel
with its algorithmic bytes and the fraction of the measured HBM copy
bandwidth.  Each call synchronises the plan's stream on return, so host
timers around a call see the kernel plus one synchronisation.
Development / measurement tool.

  python tools/transferbench.py [--n 256] [--reps 20]
"""
import argparse
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=256)
    ap.add_argument("--reps", type=int, default=20)
    args = ap.parse_args()
    import torch

    import bench

    pfb = importlib.import_module("dune-parafields_b200")
    peaks, _ = bench.measured_peaks()
    peak = float(peaks["hbm_gbs"])
    n, k = args.n, 2 * args.n
    dev = torch.device("cuda")
    coarse = torch.randn((n, n, n), dtype=torch.float64, device=dev)
    fine = torch.empty((k, k, k), dtype=torch.float64, device=dev)
    c = (torch.arange(k, dtype=torch.float64, device=dev) + 0.5) / k
    loc = torch.stack(torch.meshgrid(c, c, c, indexing="ij"), dim=-1).reshape(-1, 3).flip(-1).contiguous()
    out = torch.empty(k**3, dtype=torch.float64, device=dev)
    trend = [("mean", [0.5]), ("slope", [1.0, -1.0, 0.5]), ("disk", [0.5, 0.5, 0.5, 0.2, 1.0]), ("block", [0.3, 0.3, 0.3, 0.2, 0.2, 0.2, 2.0])]
    out = out.reshape(k, k, k)
    with pfb.Plan([8, 8], covariance="exponential", corr_length=0.3, transform="logNormal", device=0) as p:
        jobs = [
            ("refine", lambda: p.field_refine([n] * 3, coarse, fine), 8 * n**3 + 8 * k**3),
            ("coarsen", lambda: p.field_coarsen([k] * 3, fine, coarse), 8 * n**3 + 8 * k**3),
            ("localize", lambda: p.field_localize([k] * 3, [1.0] * 3, [0.5] * 3, 10.0, fine), 16 * k**3),
            ("evaluate", lambda: p.field_evaluate([k] * 3, [1.0] * 3, fine, loc, apply_transform=False, out=out), (24 + 8 + 8) * k**3),
            ("bulk_evaluate (logNormal)", lambda: p.field_bulk_evaluate([k] * 3, [1.0] * 3, fine, out=out), 16 * k**3),
            ("bulk_evaluate (logNormal, mean+slope+disk+block)", lambda: p.field_bulk_evaluate([k] * 3, [1.0] * 3, fine, trend, out=out), 16 * k**3),
        ]
        for name, fn, nbytes in jobs:
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(args.reps):
                fn()
            torch.cuda.synchronize()
            ms = (time.perf_counter() - t0) * 1e3 / args.reps
            print(json.dumps(dict(kernel=name, cells=[n if name in ("refine",) else k] * 3, ms=round(ms, 4), algorithmic_bytes=nbytes, achieved_gbs=round(nbytes / ms / 1e6, 1), frac=round(nbytes / ms / 1e6 / peak, 3))), flush=True)


if __name__ == "__main__":
    main()
