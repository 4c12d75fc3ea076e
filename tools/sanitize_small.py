#!/usr/bin/env python
"""Small invocations of every kernel family for compute-sanitizer
(`compute-sanitizer --tool memcheck python tools/sanitize_small.py`)."""
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
for case in CASES:
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
