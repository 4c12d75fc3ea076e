#!/usr/bin/env python
"""Generates tests/golden/transfer_*.npz from the UNMODIFIED reference
(oracle/_ref/libpf_ref.so): RandomField::refineMatrix + refine,
coarsenMatrix + coarsen (randomfield.hh:825-945 -> stochastic.hh:597-755) and
RandomField::evaluate (randomfield.hh:585-595 -> stochastic.hh:480-574) at
random locations, on the upper faces, edges and the far corner.

Run in the build container (the reference is not present on the GPU box):

    make -C oracle && python tests/golden/make_golden_transfer.py

RandomField::localize is absent: the reference's own code for it does not
compile (stochastic.hh:806 calls the non-static Traits::indicesToCoords
without an object), so there is nothing to record.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import ref_lib as rl  # noqa: E402

CASES = {
    "2d": dict(cells=[16, 8], covariance="exponential", corr_length=0.2),
    "2d_ragged_lognormal": dict(cells=[12, 6], covariance="exponential", corr_length=0.1, extensions=[2.0, 0.5], transform="logNormal"),
    "3d": dict(cells=[8, 4, 6], covariance="spherical", corr_length=0.3),
    "3d_sign": dict(cells=[4, 10, 2], covariance="exponential", corr_length=0.2, approximate=True, extensions=[1.0, 3.0, 0.25], transform="sign"),
}
SEED = 20240611

# trend part (trend.hh): ini sections of the reference with zero variance, so
# that meanVector + shiftVector = mean; the same components as
# (kind, coefficients) for the C ABI
TREND_CASES = {
    "2d": dict(
        case=dict(cells=[16, 12], covariance="exponential", corr_length=0.2, extensions=[1.0, 2.0]),
        ini="[mean]\nmean = 0.7\nvariance = 0\n[slope]\nmean = 0.3 -0.2\nvariance = 0 0\n[disk0]\nmean = 0.4 1.1 0.2 1.5\nvariance = 0 0 0 0\n[block0]\nmean = 0.7 0.3 0.3 0.2 -2.0\nvariance = 0 0 0 0 0\n[block1]\nmean = 0.1 1.7 0.15 0.35 4.0\nvariance = 0 0 0 0 0\n",
        trend=[("mean", [0.7]), ("slope", [0.3, -0.2]), ("disk", [0.4, 1.1, 0.2, 1.5]), ("block", [0.7, 0.3, 0.3, 0.2, -2.0]), ("block", [0.1, 1.7, 0.15, 0.35, 4.0])],
    ),
    "3d_lognormal": dict(
        case=dict(cells=[8, 6, 10], covariance="exponential", corr_length=0.2, transform="logNormal"),
        ini="[mean]\nmean = -0.5\nvariance = 0\n[slope]\nmean = 0.3 -0.2 0.9\nvariance = 0 0 0\n[disk0]\nmean = 0.4 0.6 0.5 0.3 0.8\nvariance = 0 0 0 0 0\n[disk1]\nmean = 0.9 0.1 0.2 0.15 0.4\nvariance = 0 0 0 0 0\n[block0]\nmean = 0.5 0.5 0.5 0.4 0.6 0.2 1.25\nvariance = 0 0 0 0 0 0 0\n",
        trend=[("mean", [-0.5]), ("slope", [0.3, -0.2, 0.9]), ("disk", [0.4, 0.6, 0.5, 0.3, 0.8]), ("disk", [0.9, 0.1, 0.2, 0.15, 0.4]), ("block", [0.5, 0.5, 0.5, 0.4, 0.6, 0.2, 1.25])],
    ),
    "1d": dict(
        case=dict(cells=[64], covariance="exponential", corr_length=0.1, extensions=[3.0]),
        ini="[mean]\nmean = 2.0\nvariance = 0\n[slope]\nmean = -0.75\nvariance = 0\n[disk0]\nmean = 1.0 0.5 2.0\nvariance = 0 0 0\n[block0]\nmean = 2.5 0.5 3.0\nvariance = 0 0 0\n",
        trend=[("mean", [2.0]), ("slope", [-0.75]), ("disk", [1.0, 0.5, 2.0]), ("block", [2.5, 0.5, 3.0])],
    ),
}


def trend_fixtures():
    import json

    for name, t in TREND_CASES.items():
        rng = np.random.default_rng(100 + len(name))
        case = t["case"]
        cells = case["cells"]
        ext = case.get("extensions", [1.0] * len(cells))
        r = rl.Reference(cells, extra=t["ini"], **case_kw(case))
        field = r.generate(SEED)
        loc = locations(cells, ext, rng)
        np.savez(
            os.path.join(HERE, "trend_%s.npz" % name),
            ini=r.text,
            cells=np.array(cells),
            extensions=np.array(ext),
            transform=case.get("transform", "none"),
            trend=json.dumps(t["trend"]),
            field=field,
            bulk=r.bulk_evaluate(),
            locations=loc,
            values=r.evaluate(loc),
        )
        print("trend", name, field.shape)


def locations(cells, ext, rng):
    d = len(cells)
    loc = rng.random((64, d)) * np.asarray(ext)
    # upper face of one dimension, an edge, the far corner, the origin,
    # a cell boundary and a cell centre
    for i in range(d):
        loc[i, i] = ext[i]
    loc[d, :] = ext
    if d > 1:
        loc[d + 1, :2] = ext[:2]
    loc[d + 2, :] = 0.0
    loc[d + 3, :] = [ext[i] / cells[i] * min(3, cells[i] - 1) for i in range(d)]
    loc[d + 4, :] = [ext[i] / cells[i] * min(2.5, cells[i] - 0.5) for i in range(d)]
    return loc


def main():
    if not rl.available():
        raise SystemExit("oracle/_ref/libpf_ref.so missing: run `make -C oracle` where /root/reference exists")
    for name, case in CASES.items():
        rng = np.random.default_rng(len(name))
        cells = case["cells"]
        ext = case.get("extensions", [1.0] * len(cells))
        r = rl.Reference(cells, **case_kw(case))
        field = r.generate(SEED)
        loc = locations(cells, ext, rng)
        values = r.evaluate(loc)
        refined = r.refine()
        loc_fine = locations([2 * c for c in cells], ext, rng)
        values_fine = r.evaluate(loc_fine)
        back = r.coarsen()
        # coarsening of a field that is not piecewise constant
        rough = rng.standard_normal(field.shape)
        r.set(rough)
        coarse = r.coarsen()
        np.savez(
            os.path.join(HERE, "transfer_%s.npz" % name),
            ini=r.text,
            cells=np.array(cells),
            extensions=np.array(ext),
            transform=case.get("transform", "none"),
            field=field,
            locations=loc,
            values=values,
            refined=refined,
            locations_fine=loc_fine,
            values_fine=values_fine,
            refined_coarsened=back,
            rough=rough,
            rough_coarsened=coarse,
        )
        print(name, field.shape, "->", refined.shape, "->", back.shape, "| rough ->", coarse.shape)
    trend_fixtures()


def case_kw(case):
    kw = dict(case)
    kw.pop("cells")
    return kw


if __name__ == "__main__":
    main()
