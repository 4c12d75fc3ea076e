#!/usr/bin/env python
"""Small invocations of every kernel family for compute-sanitizer
(`compute-sanitizer --tool memcheck python tools/sanitize_small.py`;
`--new` runs only the merge periodisation and the field transfer kernels)."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["PF_GEN3"] = "force"
pfb = importlib.import_module("dune-parafields_b200")

CASES = [
    dict(cells=[16, 4, 512], covariance="spherical", corr_length=0.1),  # kp_gen3 N=1024, kp_c2t, kp_c2r
    dict(cells=[8, 2, 256], covariance="exponential", corr_length=0.1),  # kp_gen3 N=512
    dict(cells=[64, 512], covariance="exponential", corr_length=0.05),  # 2D, N=1024
    dict(cells=[1024, 8], covariance="exponential", corr_length=0.05),  # long c2r axis
    dict(cells=[8, 1024], covariance="cubic", corr_length=[0.08, 0.03, 0.03, 0.05], anisotropy="geometric", approximate=True),  # long pass A axis
    dict(cells=[32, 16, 8], covariance="exponential", corr_length=0.1),  # short lines, operators
    dict(cells=[32], covariance="exponential", corr_length=0.05),
]
for case in [] if "--new" in sys.argv else CASES:
    with pfb.Plan(device=0, cg_iterations=2, **case) as p:
        p.compute_eigenvalues()
        f = p.generate(seed=1, realization=0)
        assert np.isfinite(f).all()
        noise = np.random.default_rng(0).standard_normal(2 * p.noise_entries)
        p.generate_from_noise(noise)
        if max(case["cells"]) * 2 <= 1024 or len(case["cells"]) == 3:
            p.apply("matrix", f)
            p.solve(f)
    print("ok", case["cells"], flush=True)

# merge periodisation (cov_torus_merge in the first eigenvalue pass)
for case in [
    dict(cells=[32, 16], covariance="exponential", corr_length=0.3, factor=4, sigmoid_combine="radial"),
    dict(cells=[8, 8, 8], covariance="matern32", corr_length=0.2, sigmoid="smooth", approximate=True),
    dict(cells=[16, 8], covariance="exponential", corr_length=[0.2, 0.05, 0.05, 0.1], anisotropy="geometric", approximate=True, factor=4),
]:
    with pfb.Plan(device=0, periodization="merge", **case) as p:
        p.compute_eigenvalues()
        assert np.isfinite(p.generate(seed=1, realization=0)).all()
    print("ok merge", case["cells"], flush=True)

# field transfer / evaluation kernels on ragged shapes (host arrays are staged)
rng = np.random.default_rng(1)
trend = [("mean", [0.5]), ("slope", [1.0, -1.0, 0.5]), ("disk", [0.5, 0.5, 0.5, 0.2, 1.0]), ("block", [0.3, 0.3, 0.3, 0.2, 0.2, 0.2, 2.0])]
with pfb.Plan([8, 8], device=0, transform="logNormal") as p:
    for cells in ([33, 7], [1, 5], [5, 3, 7], [34, 6, 10], [300, 2, 3]):
        d = len(cells)
        f = rng.standard_normal(tuple(reversed(cells)))
        fine = p.field_refine(cells, f)
        p.field_coarsen([2 * c for c in cells], fine)
        p.field_localize(cells, [1.0] * d, [0.5] * d, 0.3, f.copy())
        loc = rng.random((1001, d))
        loc[0, :] = 1.0
        t = [(k, c[: (1 if k == "mean" else d if k == "slope" else d + 2 if k == "disk" else 2 * d + 1)]) for k, c in trend]
        p.field_evaluate(cells, [1.0] * d, f, loc, trend=t)
        p.field_evaluate(cells, [1.0] * d, f, loc, apply_transform=False)
        p.field_bulk_evaluate(cells, [1.0] * d, f, t)
        p.field_bulk_evaluate(cells, [1.0] * d, f)
        print("ok transfer", cells, flush=True)
    x = p.field_localize([17], [2.0], [0.5], 0.3, rng.standard_normal(17))
    p.field_evaluate([17], [2.0], x, rng.random((50, 1)) * 2.0, trend=[("slope", [1.0]), ("disk", [1.0, 0.2, 1.0]), ("block", [0.5, 0.5, 3.0])])
    p.field_bulk_evaluate([17], [2.0], x, [("mean", [1.0])])
    print("ok transfer 1d", flush=True)
