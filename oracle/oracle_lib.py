"""ctypes binding of oracle/libpf_oracle.so -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs may import this module (see pf_oracle.cpp header).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libpf_oracle.so")

COV = {
    "exponential": 0,
    "gaussian": 1,
    "spherical": 2,
    "separableExponential": 3,
    "matern32": 4,
    "matern52": 5,
    "dampedOscillation": 6,
    "gammaExponential": 7,
    "cauchy": 8,
    "generalizedCauchy": 9,
    "cubic": 10,
    "whiteNoise": 11,
    "matern": 12,  # general nu = cov_param[0] (stochastic.maternNu)
}
ANISO = {"none": 0, "axiparallel": 1, "geometric": 2}
TRANSFORM = {"none": 0, "logNormal": 1, "foldedNormal": 2, "sign": 3, "boxCox": 4}


TREND = {"mean": 0, "slope": 1, "disk": 2, "block": 3}


class OrcTrend(C.Structure):
    """one TrendComponent: kind + (meanVector + shiftVector), trend.hh:601-634"""

    _fields_ = [("kind", C.c_int), ("coeff", C.c_double * 7)]


def trend_array(trend, cls=None):
    """[(kind name, [coefficients...]), ...] -> ctypes array"""
    cls = cls or OrcTrend
    arr = (cls * max(len(trend), 1))()
    for a, (kind, coeff) in zip(arr, trend):
        a.kind = TREND[kind]
        assert len(coeff) <= 7
        for i, c in enumerate(coeff):
            a.coeff[i] = float(c)
    return arr


PERIODIZATION = {"classical": 0, "merge": 1}
SIGMOID = {"smoothstep": 0, "smooth": 1}
SIGMOID_COMBINE = {"radialPlus": 0, "radial": 1, "tensor": 2}


def merge_recursions_for(covariance, dim, recursions=99, p0=0.0):
    """embedding.mergeRecursions with the magic value 99, matrix.hh:693-719"""
    import math

    if recursions != 99:
        return recursions
    if covariance == "matern":
        return int(math.floor(p0 * 2.0 + dim / 2.0))
    if covariance in ("gammaExponential", "generalizedCauchy"):
        return int(math.floor(p0 + dim / 2.0))
    table = {"exponential": 2, "matern32": 4, "matern52": 6, "cauchy": 3}
    if covariance not in table:
        raise RuntimeError("magic recursions value 99 not implemented for choice of covariance")
    return table[covariance]


class OrcCfg(C.Structure):
    _fields_ = [
        ("dim", C.c_int),
        ("cells", C.c_uint * 3),
        ("ext", C.c_double * 3),
        ("factor", C.c_uint),
        ("cov", C.c_int),
        ("variance", C.c_double),
        ("p0", C.c_double),
        ("p1", C.c_double),
        ("aniso", C.c_int),
        ("corr", C.c_double * 9),
        ("threshold", C.c_double),
        ("approximate", C.c_int),
        ("matrix_backend", C.c_int),
        ("field_backend", C.c_int),
        ("transposed", C.c_int),
        ("cg_iterations", C.c_uint),
        ("threads", C.c_int),
        ("periodization", C.c_int),
        ("sigmoid", C.c_int),
        ("sigmoid_combine", C.c_int),
        ("sigmoid_start", C.c_double),
        ("sigmoid_end", C.c_double),
        ("recursions", C.c_uint),
    ]


def build(force=False):
    """Compile the oracle with the committed Makefile (gcc only)."""
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(
        os.path.join(_HERE, "pf_oracle.cpp")
    ):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        L.orc_last_error.restype = C.c_char_p
        L.orc_lambda_size.restype = C.c_size_t
        L.orc_noise_entries.restype = C.c_size_t
        L.orc_lambda_size.argtypes = [C.POINTER(OrcCfg)]
        L.orc_noise_entries.argtypes = [C.POINTER(OrcCfg)]
        dp = C.POINTER(C.c_double)
        L.orc_eigenvalues.argtypes = [C.POINTER(OrcCfg), dp, C.POINTER(C.c_int), dp, C.c_int]
        L.orc_lambda_natural.argtypes = [C.POINTER(OrcCfg), dp, dp]
        L.orc_generate_from_noise.argtypes = [C.POINTER(OrcCfg), dp, dp, dp, dp]
        L.orc_reference_noise.argtypes = [C.c_uint, C.c_size_t, dp]
        L.orc_reference_noise_ranks.argtypes = [C.c_uint, C.c_size_t, C.c_int, dp]
        L.orc_max_threads.restype = C.c_int
        u32 = C.POINTER(C.c_uint32)
        L.orc_philox4x32.argtypes = [C.c_int, u32, u32, u32]
        L.orc_philox_noise.argtypes = [C.POINTER(OrcCfg), C.c_uint64, C.c_uint64, C.c_int, dp]
        L.orc_philox_uncorrelated.argtypes = [C.c_uint64, C.c_size_t, C.c_uint64, dp]
        L.orc_generate_uncorrelated.argtypes = [C.c_uint, C.c_size_t, dp]
        L.orc_multiply.argtypes = [C.POINTER(OrcCfg), dp, C.c_int, dp, dp]
        L.orc_multiply_inverse.argtypes = [C.POINTER(OrcCfg), dp, dp, dp, C.POINTER(C.c_uint)]
        L.orc_transform.argtypes = [C.c_int, C.c_double, C.c_size_t, dp]
        L.orc_torus_covariance.argtypes = [C.POINTER(OrcCfg), dp]
        up = C.POINTER(C.c_uint)
        L.orc_field_refine.argtypes = [C.c_int, up, dp, dp]
        L.orc_field_coarsen.argtypes = [C.c_int, up, dp, dp]
        L.orc_field_localize.argtypes = [C.c_int, up, up, up, dp, dp, C.c_double, dp]
        L.orc_field_evaluate.argtypes = [C.c_int, up, dp, dp, C.c_size_t, dp, C.c_int, C.c_double, dp]
        tp = C.POINTER(OrcTrend)
        L.orc_field_evaluate_trend.argtypes = [C.c_int, up, dp, dp, C.c_size_t, dp, C.c_int, tp, C.c_int, C.c_double, dp]
        L.orc_field_bulk_evaluate.argtypes = [C.c_int, up, dp, dp, C.c_int, tp, C.c_int, C.c_double, dp]
        _lib = L
    return _lib


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _check(rc):
    if rc != 0:
        raise RuntimeError(lib().orc_last_error().decode())


class Oracle:
    """One reference configuration (single rank).

    Defaults follow the reference's backend selection
    (matrix.hh:1668-1767, randomfield.hh:151-162): 1D -> DFT matrix + DFT
    field backend; nD iso/axiparallel -> DCT matrix + R2C field; nD geometric
    -> R2C matrix + R2C field; transposed layout only in 3D
    (fieldtraits.hh:226-264).
    """

    def __init__(
        self,
        cells,
        extensions=None,
        covariance="exponential",
        corr_length=0.05,
        variance=1.0,
        anisotropy="none",
        factor=2,
        threshold=1e-14,
        approximate=False,
        matrix_backend=None,
        field_backend=None,
        transposed=None,
        cg_iterations=100,
        threads=1,
        p0=0.0,
        p1=0.0,
        periodic=False,
        periodization="classical",
        sigmoid="smoothstep",
        sigmoid_combine="radialPlus",
        sigmoid_start=1.0,
        sigmoid_end=None,
        merge_recursions=99,
    ):
        if periodic:
            factor = 1  # fieldtraits.hh:196-202
        cells = list(cells)
        dim = len(cells)
        self.dim = dim
        self.cells = cells
        self.factor = factor
        self.n = [factor * c for c in cells]
        cfg = OrcCfg()
        cfg.dim = dim
        ext = list(extensions) if extensions is not None else [1.0] * dim
        for i in range(dim):
            cfg.cells[i] = cells[i]
            cfg.ext[i] = ext[i]
        cfg.factor = factor
        cfg.cov = COV[covariance]
        cfg.variance = variance
        cfg.p0, cfg.p1 = p0, p1
        cfg.aniso = ANISO[anisotropy]
        cl = np.atleast_1d(np.asarray(corr_length, dtype=np.float64)).ravel()
        need = {0: 1, 1: dim, 2: dim * dim}[cfg.aniso]
        if cl.size != need:
            raise ValueError("corrLength arity %d, expected %d" % (cl.size, need))
        for i in range(cl.size):
            cfg.corr[i] = cl[i]
        cfg.threshold = threshold
        cfg.approximate = int(approximate)
        if matrix_backend is None:
            matrix_backend = "DFT" if dim == 1 else ("R2C" if anisotropy == "geometric" else "DCT")
        if field_backend is None:
            field_backend = "DFT" if dim == 1 else "R2C"
        cfg.matrix_backend = {"DCT": 0, "R2C": 1, "DFT": 2}[matrix_backend]
        cfg.field_backend = {"R2C": 0, "DFT": 1}[field_backend]
        cfg.transposed = int(dim == 3) if transposed is None else int(transposed)
        cfg.cg_iterations = cg_iterations
        cfg.threads = threads
        cfg.periodization = PERIODIZATION[periodization]
        cfg.sigmoid = SIGMOID[sigmoid]
        cfg.sigmoid_combine = SIGMOID_COMBINE[sigmoid_combine]
        cfg.sigmoid_start = sigmoid_start
        cfg.sigmoid_end = factor - 1.0 if sigmoid_end is None else sigmoid_end
        cfg.recursions = merge_recursions_for(covariance, dim, merge_recursions, p0) if periodization == "merge" else 0
        self.cfg = cfg
        self.matrix_backend = matrix_backend
        self.field_backend = field_backend
        self._lambda = None
        self.stats = None

    # -- sizes -----------------------------------------------------------
    @property
    def domain_size(self):
        return int(np.prod(self.cells))

    @property
    def noise_entries(self):
        return int(lib().orc_noise_entries(C.byref(self.cfg)))

    @property
    def lambda_size(self):
        return int(lib().orc_lambda_size(C.byref(self.cfg)))

    # -- eigenvalues ------------------------------------------------------
    def eigenvalues(self, census_quirk=False):
        lam = np.empty(self.lambda_size, dtype=np.float64)
        counts = (C.c_int * 3)()
        smallest = C.c_double()
        rc = lib().orc_eigenvalues(
            C.byref(self.cfg), _dp(lam), counts, C.byref(smallest), int(census_quirk)
        )
        self.stats = dict(small=counts[0], zero=counts[1], negative=counts[2], smallest=smallest.value)
        _check(rc)
        self._lambda = lam
        return lam

    def lambda_stored(self):
        if self._lambda is None:
            self.eigenvalues()
        return self._lambda

    def lambda_natural(self):
        """lambda on the field backend's index set in natural (k0,k1,k2) order."""
        lam = self.lambda_stored()
        n = self.n + [1] * (3 - self.dim)
        b0 = n[0] if self.field_backend == "DFT" else n[0] // 2 + 1
        out = np.empty(b0 * n[1] * n[2], dtype=np.float64)
        _check(lib().orc_lambda_natural(C.byref(self.cfg), _dp(lam), _dp(out)))
        return out.reshape(n[2], n[1], b0)

    def torus_covariance(self):
        n = self.n + [1] * (3 - self.dim)
        out = np.empty(n[0] * n[1] * n[2], dtype=np.float64)
        _check(lib().orc_torus_covariance(C.byref(self.cfg), _dp(out)))
        return out.reshape(n[2], n[1], n[0])

    # -- generate ----------------------------------------------------------
    def reference_noise(self, seed):
        noise = np.empty(2 * self.noise_entries, dtype=np.float64)
        _check(lib().orc_reference_noise(C.c_uint(seed), self.noise_entries, _dp(noise)))
        return noise

    def reference_noise_ranks(self, seed, ranks):
        """noise stream of a `ranks`-rank MPI run of the reference"""
        noise = np.empty(2 * self.noise_entries, dtype=np.float64)
        _check(lib().orc_reference_noise_ranks(C.c_uint(seed), self.noise_entries, ranks, _dp(noise)))
        return noise

    def philox_noise(self, seed, realization=0, mode="fp64"):
        """the CUDA path's device noise (Philox4x32 counter stream + Box-Muller)
        for (seed, realization), restated on the CPU, in reference stream order"""
        noise = np.empty(2 * self.noise_entries, dtype=np.float64)
        _check(lib().orc_philox_noise(C.byref(self.cfg), seed, realization, {"fast": 0, "fp64": 1}[mode], _dp(noise)))
        return noise

    def generate_from_noise(self, noise, want_spare=False):
        lam = self.lambda_stored()
        noise = np.ascontiguousarray(noise, dtype=np.float64)
        assert noise.size == 2 * self.noise_entries
        out = np.empty(self.domain_size, dtype=np.float64)
        spare = np.empty(self.domain_size, dtype=np.float64) if want_spare else None
        _check(
            lib().orc_generate_from_noise(
                C.byref(self.cfg), _dp(lam), _dp(noise), _dp(out), _dp(spare) if want_spare else None
            )
        )
        shape = tuple(reversed(self.cells))
        if want_spare:
            return out.reshape(shape), spare.reshape(shape)
        return out.reshape(shape)

    def generate(self, seed, want_spare=False):
        return self.generate_from_noise(self.reference_noise(seed), want_spare)

    # -- operators ----------------------------------------------------------
    def multiply(self, x, op="matrix"):
        lam = self.lambda_stored()
        x = np.ascontiguousarray(x, dtype=np.float64).ravel()
        out = np.empty_like(x)
        opi = {"matrix": 0, "root": 1, "inverse_ext": 2}[op]
        _check(lib().orc_multiply(C.byref(self.cfg), _dp(lam), opi, _dp(x), _dp(out)))
        return out.reshape(tuple(reversed(self.cells)))

    def multiply_inverse(self, x):
        lam = self.lambda_stored()
        x = np.ascontiguousarray(x, dtype=np.float64).ravel()
        out = np.empty_like(x)
        it = C.c_uint()
        _check(lib().orc_multiply_inverse(C.byref(self.cfg), _dp(lam), _dp(x), _dp(out), C.byref(it)))
        self.cg_iterations_used = it.value
        return out.reshape(tuple(reversed(self.cells)))


def generate_uncorrelated(seed, n):
    out = np.empty(n, dtype=np.float64)
    _check(lib().orc_generate_uncorrelated(C.c_uint(seed), n, _dp(out)))
    return out


def philox4x32(rounds, ctr, key):
    """one Philox4x32-R block of the oracle's restatement (Salmon et al., SC'11)"""
    c = (C.c_uint32 * 4)(*ctr)
    k = (C.c_uint32 * 2)(*key)
    o = (C.c_uint32 * 4)()
    _check(lib().orc_philox4x32(rounds, c, k, o))
    return tuple(int(x) for x in o)


def philox_uncorrelated(seed, n, q0=0):
    """pf_generate_uncorrelated's specification restated (device Philox, fp64)"""
    out = np.empty(n, dtype=np.float64)
    _check(lib().orc_philox_uncorrelated(seed, n, q0, _dp(out)))
    return out


def transform(kind, values, boxcox_lambda=0.0):
    v = np.array(values, dtype=np.float64).ravel().copy()
    _check(lib().orc_transform(TRANSFORM[kind], boxcox_lambda, v.size, _dp(v)))
    return v.reshape(np.shape(values))


def _up(seq):
    return (C.c_uint * 3)(*(list(seq) + [1] * (3 - len(seq))))


def field_refine(cells, field):
    """StochasticPart::refine, stochastic.hh:597-670; cells = old cells (x fastest)."""
    f = np.ascontiguousarray(field, dtype=np.float64).ravel()
    out = np.empty(f.size * 2 ** len(cells))
    _check(lib().orc_field_refine(len(cells), _up(cells), _dp(f), _dp(out)))
    return out.reshape(tuple(2 * c for c in reversed(cells)))


def field_coarsen(cells, field):
    """StochasticPart::coarsen, stochastic.hh:679-755; cells = old (fine) cells."""
    f = np.ascontiguousarray(field, dtype=np.float64).ravel()
    out = np.empty(max(f.size // 2 ** len(cells), 1))
    _check(lib().orc_field_coarsen(len(cells), _up(cells), _dp(f), _dp(out)))
    return out.reshape(tuple(c // 2 for c in reversed(cells)))


def field_localize(cells, extensions, center, radius, field, offset=None, global_cells=None):
    """StochasticPart::localize, stochastic.hh:798-817."""
    d = len(cells)
    f = np.array(field, dtype=np.float64).ravel().copy()
    ext = np.asarray(extensions, dtype=np.float64)
    cen = np.asarray(center, dtype=np.float64)
    _check(lib().orc_field_localize(d, _up(cells), _up(offset or [0] * d), _up(global_cells or cells), _dp(ext), _dp(cen), radius, _dp(f)))
    return f.reshape(tuple(reversed(cells)))


def field_evaluate(cells, extensions, field, locations, transform="none", boxcox_lambda=0.0):
    """StochasticPart::evaluate (stochastic.hh:480-574) + value transform, one rank."""
    d = len(cells)
    f = np.ascontiguousarray(field, dtype=np.float64).ravel()
    loc = np.ascontiguousarray(locations, dtype=np.float64).reshape(-1, d)
    ext = np.asarray(extensions, dtype=np.float64)
    out = np.empty(loc.shape[0])
    _check(lib().orc_field_evaluate(d, _up(cells), _dp(ext), _dp(f), loc.shape[0], _dp(loc), TRANSFORM[transform], boxcox_lambda, _dp(out)))
    return out


def field_evaluate_trend(cells, extensions, field, locations, trend, transform="none", boxcox_lambda=0.0):
    """RandomField::evaluate = T(stochastic + trend), randomfield.hh:585-595."""
    d = len(cells)
    f = np.ascontiguousarray(field, dtype=np.float64).ravel()
    loc = np.ascontiguousarray(locations, dtype=np.float64).reshape(-1, d)
    ext = np.asarray(extensions, dtype=np.float64)
    out = np.empty(loc.shape[0])
    _check(lib().orc_field_evaluate_trend(d, _up(cells), _dp(ext), _dp(f), loc.shape[0], _dp(loc), len(trend), trend_array(trend), TRANSFORM[transform], boxcox_lambda, _dp(out)))
    return out


def field_bulk_evaluate(cells, extensions, field, trend=(), transform="none", boxcox_lambda=0.0):
    """RandomField::bulkEvaluate, randomfield.hh:597-631."""
    d = len(cells)
    f = np.ascontiguousarray(field, dtype=np.float64).ravel()
    ext = np.asarray(extensions, dtype=np.float64)
    out = np.empty(f.size)
    _check(lib().orc_field_bulk_evaluate(d, _up(cells), _dp(ext), _dp(f), len(trend), trend_array(list(trend)), TRANSFORM[transform], boxcox_lambda, _dp(out)))
    return out.reshape(tuple(reversed(cells)))
