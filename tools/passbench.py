#!/usr/bin/env python
"""Per-pass device times of the generate pipeline for one workload under
several kernel selections (PF_PIPE masks).  Development tool; prints one JSON
line per configuration.

  python tools/passbench.py [--workload K3] [--masks 0,0x1d,...] [--reps 5]
"""
import argparse
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="K3")
    ap.add_argument("--masks", default="0,0x1d")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--rng", default="fast")
    args = ap.parse_args()
    import torch

    import bench

    pfb = importlib.import_module("dune-parafields_b200")
    wl = {k: v for k, v in bench.WORKLOADS[args.workload].items() if k != "label"}
    names = {0: "A", 1: "A1d", 2: "B", 4: "C", 12: "AB"}
    for mask in args.masks.split(","):
        if mask == "default":
            os.environ.pop("PF_PIPE", None)
        else:
            os.environ["PF_PIPE"] = mask
        with pfb.Plan(device=0, rng_mode=args.rng, **wl) as plan:
            plan.compute_eigenvalues()
            out = torch.empty(plan.domain_size, dtype=torch.float64, device="cuda")
            for r in range(2):
                plan.generate(1, r, out=out)
            plan.synchronize()
            plan.set_profiling(True)
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
            for r in range(args.reps):
                plan.generate(1, 10 + r, out=out)
            ev1.record()
            torch.cuda.synchronize()
            t = plan.pass_times()
            res = {names.get(k, str(k)): round(v[0] / max(v[1], 1), 4) for k, v in t.items()}
            res["step_ms"] = round(ev0.elapsed_time(ev1) / args.reps, 4)
            res["mask"] = mask
            res["workload"] = args.workload
            res["checksum"] = float(out.double().abs().sum().item())
            print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
