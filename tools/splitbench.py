#!/usr/bin/env python
"""One GPU: sweep of the SM split of pf_generate_batch's pipeline (pass A of
realization r+1 beside the tail of realization r; PF_SPLIT = SMs for pass A,
0 = sequential).  Checks that the pipelined batch is bit-identical to
realization-by-realization calls.  Development tool; one JSON line per split.

  python tools/splitbench.py [--workload K3] [--splits 0,96,104,108,112,116] [--reps 16]
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
    ap.add_argument("--splits", default="0,96,104,108,112,116,120")
    ap.add_argument("--reps", type=int, default=16)
    ap.add_argument("--rng", default="fast")
    args = ap.parse_args()
    import torch

    import bench

    pfb = importlib.import_module("dune-parafields_b200")
    wl = {k: v for k, v in bench.WORKLOADS[args.workload].items() if k != "label"}
    apply_tr = wl.get("transform", "none") != "none"
    names = {0: "A", 2: "B", 4: "C"}
    ref = None
    for split in args.splits.split(","):
        os.environ["PF_SPLIT"] = split
        with pfb.Plan(device=0, rng_mode=args.rng, **wl) as plan:
            stream = torch.cuda.Stream()
            plan.set_stream(stream)
            plan.compute_eigenvalues()
            D = plan.domain_size
            K = args.reps
            fields = torch.empty((K, D), dtype=torch.float64, device="cuda")
            plan.generate_batch(1, 100, 3, out=fields[:3], apply_transform=apply_tr)
            plan.synchronize()
            torch.cuda.synchronize()
            plan.set_profiling(True)
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record(stream)
            plan.generate_batch(1, 0, K, out=fields, apply_transform=apply_tr)
            ev1.record(stream)
            torch.cuda.synchronize()
            t = plan.pass_times()
            plan.set_profiling(False)
            res = {names.get(k, str(k)): round(v[0] / max(v[1], 1), 4) for k, v in t.items()}
            res["step_ms"] = round(ev0.elapsed_time(ev1) / K, 4)
            res["per_s"] = round(K / (ev0.elapsed_time(ev1) * 1e-3), 2)
            res["split"] = int(split)
            res["workload"] = args.workload
            if ref is None:
                # reference: one call per realization (never pipelined)
                ref = torch.empty((3, D), dtype=torch.float64, device="cuda")
                for r, idx in enumerate((0, 1, K - 1)):
                    plan.generate(1, idx, out=ref[r], apply_transform=apply_tr)
                plan.synchronize()
                torch.cuda.synchronize()
            res["identical_to_single_calls"] = bool(
                torch.equal(fields[0], ref[0]) and torch.equal(fields[1], ref[1]) and torch.equal(fields[K - 1], ref[2])
            )
            print(json.dumps(res), flush=True)
            del fields


if __name__ == "__main__":
    main()
