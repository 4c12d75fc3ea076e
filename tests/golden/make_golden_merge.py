#!/usr/bin/env python
"""Generates tests/golden/merge_*.npz from the UNMODIFIED reference
(oracle/_ref/libpf_ref.so) with embedding.periodization = merge
(Matrix::computeMatrixEntriesWithMerge, matrix.hh:662-787): eigenvalues as
stored by the matrix backend and the field generated from the reference's own
RNG stream.  `case` holds the keyword arguments shared by the oracle and the
CUDA plan, `ini` the extra ini lines given to the reference.

Run in the build container (the reference is not present on the GPU box):

    make -C oracle && python tests/golden/make_golden_merge.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import ref_lib as rl  # noqa: E402

GEO2 = [0.2, 0.05, 0.05, 0.1]
GEO3 = [0.2, 0.05, 0.0, 0.05, 0.1, 0.0, 0.0, 0.0, 0.3]
CASES = {
    "2d_default": (dict(cells=[16, 16], covariance="exponential", corr_length=0.2, approximate=True), ""),
    "2d_smooth_matern32": (dict(cells=[16, 8], covariance="matern32", corr_length=0.2, approximate=True, sigmoid="smooth"), "sigmoid = smooth\n"),
    "3d_default": (dict(cells=[8, 8, 8], covariance="exponential", corr_length=0.2, approximate=True), ""),
    "2d_radial_f4": (dict(cells=[16, 16], covariance="exponential", corr_length=0.2, approximate=True, factor=4, sigmoid_combine="radial"), "[stochastic]\nsigmoidCombine = radial\n"),
    "2d_tensor_cauchy_f4": (dict(cells=[16, 16], covariance="cauchy", corr_length=0.2, approximate=True, factor=4, sigmoid_combine="tensor"), "[stochastic]\nsigmoidCombine = tensor\n"),
    "2d_geometric_f4_params": (
        dict(cells=[16, 8], covariance="exponential", corr_length=GEO2, anisotropy="geometric", approximate=True, factor=4, merge_recursions=5, sigmoid_start=0.8, sigmoid_end=2.5),
        "mergeRecursions = 5\nsigmoidStart = 0.8\nsigmoidEnd = 2.5\n",
    ),
    "3d_geometric": (dict(cells=[8, 4, 4], covariance="exponential", corr_length=GEO3, anisotropy="geometric", approximate=True), ""),
    "3d_radialplus_f4_matern52": (dict(cells=[8, 8, 8], covariance="matern52", corr_length=0.2, approximate=True, factor=4), ""),
    "1d_default": (dict(cells=[32], covariance="exponential", corr_length=0.2, approximate=True), ""),
}
REF_KEYS = ("covariance", "corr_length", "approximate", "factor", "anisotropy")
SEED = 20240611


def main():
    if not rl.available():
        raise SystemExit("oracle/_ref/libpf_ref.so missing: run `make -C oracle` where /root/reference exists")
    for name, (case, ini) in CASES.items():
        kw = {k: v for k, v in case.items() if k in REF_KEYS}
        r = rl.Reference(case["cells"], extra="[embedding]\nperiodization = merge\n" + ini, **kw)
        lam = r.lambda_stored()
        field = r.generate(SEED)
        np.savez(os.path.join(HERE, "merge_%s.npz" % name), ini=r.text, case=json.dumps(case), seed=SEED, lambda_stored=lam, field=field)
        print(name, field.shape, "lambda min %.3e" % lam.min())


if __name__ == "__main__":
    main()
