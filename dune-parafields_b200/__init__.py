"""dune-parafields_b200 -- B200-native circulant-embedding sampler.

Python plumbing over the C ABI of ``csrc/libparafields_cuda.so``
(``include/parafields_cuda.h``).  The product is the CUDA library and the
C++17 headers under ``include/``; this module exists for the tests, the
benchmark and ``torch.distributed`` launches.  It never falls back to a CPU
implementation: if the shared library is missing or no CUDA device is
present, calls raise.

The package directory name contains a hyphen (it is the reference's project
name), so import it with::

    import importlib
    pfb = importlib.import_module("dune-parafields_b200")
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.environ.get("PF_LIB") or os.path.join(CSRC, "libparafields_cuda.so")
INCLUDE_DIR = os.path.join(os.path.dirname(_HERE), "include")

COVARIANCE = {
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
    "custom": 13,  # custom-iso / custom-aniso: values through set_covariance_table
}
ANISOTROPY = {"none": 0, "axiparallel": 1, "geometric": 2}
TRANSFORM = {"none": 0, "logNormal": 1, "foldedNormal": 2, "sign": 3, "boxCox": 4}
OP = {"matrix": 0, "root": 1, "inverse_ext": 2}
LAMBDA_NATURAL, LAMBDA_REFERENCE = 0, 1

PF_OK, PF_ERR_INVALID, PF_ERR_NEGATIVE_EIGENVALUES, PF_ERR_CUDA, PF_ERR_NOMEM = range(5)

# every entry point declared in include/parafields_cuda.h
ABI_SYMBOLS = [
    "pf_version",
    "pf_last_error",
    "pf_device_count",
    "pf_device_info",
    "pf_desc_default",
    "pf_plan_create",
    "pf_plan_destroy",
    "pf_plan_set_stream",
    "pf_plan_synchronize",
    "pf_plan_launch_count",
    "pf_plan_set_profiling",
    "pf_plan_get_pass_times",
    "pf_domain_size",
    "pf_extended_size",
    "pf_noise_entries",
    "pf_local_domain_size",
    "pf_slab_handle",
    "pf_slab_connect",
    "pf_set_covariance_table",
    "pf_compute_eigenvalues",
    "pf_eigenvalue_count",
    "pf_get_eigenvalues",
    "pf_generate",
    "pf_generate_batch",
    "pf_generate_from_noise",
    "pf_generate_uncorrelated",
    "pf_set_variance_as_field",
    "pf_apply",
    "pf_solve",
    "pf_transform",
    "pf_vec_axpy",
    "pf_vec_scale",
    "pf_vec_dot",
    "pf_vec_norm1",
    "pf_vec_norminf",
    "pf_vec_equal",
    "pf_vec_fill",
    "pf_vec_dot_static",
    "pf_field_refine",
    "pf_field_coarsen",
    "pf_field_localize",
    "pf_field_evaluate",
    "pf_field_evaluate_trend",
    "pf_field_bulk_evaluate",
    "pf_field_overlap_size",
    "pf_field_slab_to_blocks",
    "pf_field_blocks_to_slab",
    "pf_device_alloc",
    "pf_device_free",
    "pf_memcpy",
]


TREND_KINDS = {"mean": 0, "slope": 1, "disk": 2, "block": 3}


class PfTrendComponent(C.Structure):
    """pf_trend_component: kind + (meanVector + shiftVector), trend.hh:601-634"""

    _fields_ = [("kind", C.c_int), ("coeff", C.c_double * 7)]


def _trend_array(trend):
    """[(kind name, [coefficients...]), ...] -> (count, ctypes array or None)"""
    trend = list(trend or ())
    if not trend:
        return 0, None
    arr = (PfTrendComponent * len(trend))()
    for a, (kind, coeff) in zip(arr, trend):
        a.kind = TREND_KINDS[kind]
        for i, c in enumerate(coeff):
            a.coeff[i] = float(c)
    return len(trend), arr


class PfDesc(C.Structure):
    _fields_ = [
        ("dim", C.c_int),
        ("cells", C.c_uint * 3),
        ("extensions", C.c_double * 3),
        ("embedding_factor", C.c_uint),
        ("covariance", C.c_int),
        ("variance", C.c_double),
        ("cov_param", C.c_double * 2),
        ("anisotropy", C.c_int),
        ("corr_length", C.c_double * 9),
        ("threshold", C.c_double),
        ("approximate", C.c_int),
        ("periodic", C.c_int),
        ("transform", C.c_int),
        ("boxcox_lambda", C.c_double),
        ("cg_iterations", C.c_uint),
        ("parity_mode", C.c_int),
        ("rng_mode", C.c_int),
        ("census_quirk", C.c_int),
        ("max_factor", C.c_uint),
        ("device", C.c_int),
        ("n_ranks", C.c_int),
        ("rank", C.c_int),
        ("periodization", C.c_int),
        ("sigmoid", C.c_int),
        ("sigmoid_combine", C.c_int),
        ("sigmoid_start", C.c_double),
        ("sigmoid_end", C.c_double),
        ("merge_recursions", C.c_uint),
    ]


class PfEigStats(C.Structure):
    _fields_ = [
        ("small_count", C.c_ulonglong),
        ("zero_count", C.c_ulonglong),
        ("negative_count", C.c_ulonglong),
        ("smallest", C.c_double),
        ("factor_used", C.c_uint),
    ]


class PfError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


def build(verbose=False):
    """Compile csrc/ into libparafields_cuda.so (nvcc, sm_100a, -lineinfo)."""
    jobs = str(max(1, min(10, os.cpu_count() or 1)))
    cmd = ["make", "-C", CSRC, "-j", jobs]
    out = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or out.returncode != 0:
        print(out.stdout)
    if out.returncode != 0:
        raise RuntimeError("building libparafields_cuda.so failed")
    return LIB_PATH


_lib = None


def lib():
    """Load the CUDA library; fail loudly if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "%s not found: build it with `make -C %s` (nvcc, sm_100a). "
            "There is no CPU fallback." % (LIB_PATH, CSRC)
        )
    L = C.CDLL(LIB_PATH)
    vp, dp, sz = C.c_void_p, C.c_void_p, C.c_size_t
    L.pf_version.restype = C.c_char_p
    L.pf_last_error.restype = C.c_char_p
    L.pf_device_info.argtypes = [C.c_int, C.c_char_p, C.POINTER(C.c_int), C.POINTER(sz)]
    L.pf_desc_default.argtypes = [C.POINTER(PfDesc)]
    L.pf_desc_default.restype = None
    L.pf_plan_create.argtypes = [C.POINTER(PfDesc), C.POINTER(vp)]
    L.pf_plan_destroy.argtypes = [vp]
    L.pf_plan_set_stream.argtypes = [vp, vp]
    L.pf_plan_synchronize.argtypes = [vp]
    L.pf_plan_launch_count.argtypes = [vp]
    L.pf_plan_launch_count.restype = C.c_ulonglong
    L.pf_plan_set_profiling.argtypes = [vp, C.c_int]
    L.pf_plan_get_pass_times.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_ulonglong)]
    L.pf_slab_handle.argtypes = [vp, vp]
    L.pf_slab_connect.argtypes = [vp, vp, C.c_int]
    for f in (L.pf_domain_size, L.pf_extended_size, L.pf_noise_entries, L.pf_local_domain_size):
        f.argtypes = [vp]
        f.restype = sz
    L.pf_set_covariance_table.argtypes = [vp, dp, sz]
    L.pf_compute_eigenvalues.argtypes = [vp, C.POINTER(PfEigStats)]
    L.pf_eigenvalue_count.argtypes = [vp, C.c_int]
    L.pf_eigenvalue_count.restype = sz
    L.pf_get_eigenvalues.argtypes = [vp, dp, sz, C.c_int]
    L.pf_generate.argtypes = [vp, C.c_uint64, C.c_uint64, dp, C.c_int]
    L.pf_generate_batch.argtypes = [vp, C.c_uint64, C.c_uint64, sz, dp, C.c_int]
    L.pf_generate_from_noise.argtypes = [vp, dp, dp, dp]
    L.pf_generate_uncorrelated.argtypes = [vp, C.c_uint64, dp]
    L.pf_set_variance_as_field.argtypes = [vp, dp]
    L.pf_apply.argtypes = [vp, C.c_int, dp, dp]
    L.pf_solve.argtypes = [vp, dp, dp, C.POINTER(C.c_uint)]
    L.pf_transform.argtypes = [vp, dp, dp, sz]
    L.pf_vec_axpy.argtypes = [vp, sz, C.c_double, dp, dp]
    L.pf_vec_scale.argtypes = [vp, sz, C.c_double, dp]
    L.pf_vec_dot.argtypes = [vp, sz, dp, dp, C.POINTER(C.c_double)]
    L.pf_vec_norm1.argtypes = [vp, sz, dp, C.POINTER(C.c_double)]
    L.pf_vec_norminf.argtypes = [vp, sz, dp, C.POINTER(C.c_double)]
    L.pf_vec_equal.argtypes = [vp, sz, dp, dp, C.POINTER(C.c_int)]
    L.pf_vec_fill.argtypes = [vp, sz, C.c_double, dp]
    L.pf_vec_dot_static.argtypes = [vp, sz, dp, dp, dp]
    up = C.POINTER(C.c_uint)
    L.pf_field_refine.argtypes = [vp, C.c_int, up, dp, dp]
    L.pf_field_coarsen.argtypes = [vp, C.c_int, up, dp, dp]
    L.pf_field_localize.argtypes = [vp, C.c_int, up, up, up, dp, dp, C.c_double, dp]
    L.pf_field_evaluate.argtypes = [vp, C.c_int, up, dp, dp, sz, dp, C.c_int, dp]
    tp = C.POINTER(PfTrendComponent)
    L.pf_field_evaluate_trend.argtypes = [vp, C.c_int, up, dp, dp, sz, dp, C.c_int, tp, C.c_int, dp]
    L.pf_field_bulk_evaluate.argtypes = [vp, C.c_int, up, dp, dp, C.c_int, tp, C.c_int, dp]
    ip = C.POINTER(C.c_int)
    L.pf_field_overlap_size.argtypes = [vp, ip]
    L.pf_field_overlap_size.restype = sz
    L.pf_field_slab_to_blocks.argtypes = [vp, ip, dp, dp, dp]
    L.pf_field_blocks_to_slab.argtypes = [vp, ip, dp, dp]
    L.pf_device_alloc.argtypes = [vp, sz, C.POINTER(vp)]
    L.pf_device_free.argtypes = [vp, vp]
    L.pf_memcpy.argtypes = [vp, vp, vp, sz]
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise PfError(rc, lib().pf_last_error().decode())


def _ptr(a):
    """Address of a numpy array, a torch tensor (host or CUDA) or a raw int."""
    if a is None:
        return None
    if isinstance(a, int):
        return a
    if isinstance(a, np.ndarray):
        assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
        return a.ctypes.data
    if hasattr(a, "data_ptr"):  # torch tensor
        assert a.is_contiguous()
        return a.data_ptr()
    raise TypeError("unsupported buffer type %r" % type(a))


def make_desc(
    cells,
    extensions=None,
    covariance="exponential",
    corr_length=0.05,
    variance=1.0,
    anisotropy="none",
    factor=2,
    threshold=1e-14,
    approximate=False,
    periodic=False,
    transform="none",
    boxcox_lambda=0.0,
    cg_iterations=100,
    parity_mode="reference",
    rng_mode="fast",
    census_quirk=False,
    max_factor=0,
    device=-1,
    cov_param=(0.0, 0.0),
    n_ranks=1,
    rank=0,
    periodization="classical",
    sigmoid="smoothstep",
    sigmoid_combine="radialPlus",
    sigmoid_start=1.0,
    sigmoid_end=None,
    merge_recursions=99,
):
    d = PfDesc()
    lib().pf_desc_default(C.byref(d))
    cells = list(cells)
    dim = len(cells)
    d.dim = dim
    ext = list(extensions) if extensions is not None else [1.0] * dim
    for i in range(dim):
        d.cells[i] = int(cells[i])
        d.extensions[i] = float(ext[i])
    d.embedding_factor = factor
    d.covariance = COVARIANCE[covariance]
    d.variance = variance
    d.cov_param[0], d.cov_param[1] = cov_param
    d.anisotropy = ANISOTROPY[anisotropy]
    cl = np.atleast_1d(np.asarray(corr_length, dtype=np.float64)).ravel()
    need = {0: 1, 1: dim, 2: dim * dim}[d.anisotropy]
    if cl.size != need:
        raise ValueError("stochastic.corrLength needs %d values for anisotropy %s" % (need, anisotropy))
    for i in range(cl.size):
        d.corr_length[i] = cl[i]
    d.threshold = threshold
    d.approximate = int(approximate)
    d.periodic = int(periodic)
    d.transform = TRANSFORM[transform]
    d.boxcox_lambda = boxcox_lambda
    d.cg_iterations = cg_iterations
    d.parity_mode = {"reference": 0, "corrected": 1}[parity_mode]
    d.rng_mode = {"fast": 0, "fp64": 1}[rng_mode]
    d.census_quirk = int(census_quirk)
    d.max_factor = max_factor
    d.device = device
    d.n_ranks = int(n_ranks)
    d.rank = int(rank)
    # embedding.periodization / sigmoid, stochastic.sigmoidCombine (matrix.hh:530-590,662-719)
    d.periodization = {"classical": 0, "merge": 1}[periodization]
    d.sigmoid = {"smoothstep": 0, "smooth": 1}[sigmoid]
    d.sigmoid_combine = {"radialPlus": 0, "radial": 1, "tensor": 2}[sigmoid_combine]
    d.sigmoid_start = sigmoid_start
    d.sigmoid_end = -1.0 if sigmoid_end is None else sigmoid_end
    d.merge_recursions = merge_recursions
    return d


def slab_layout(cells, factor, n_ranks, rank):
    """Host-side statement of the slab decomposition the plan uses (pf_api.cu
    plan_resize; replaces fftw_mpi_local_size of fieldtraits.hh:271-282):
    which lines of the first inverse pass this rank computes and which layers
    of the last dimension it owns afterwards."""
    cells = list(cells)
    dim = len(cells)
    if dim < 2:
        raise ValueError("slab mode needs dim >= 2")
    if cells[-1] % n_ranks:
        raise ValueError("number of cells in last dimension has to be multiple of numProc")
    n = [factor * c for c in cells]
    m0 = n[0] // 2 + 1
    A = n[1] if dim == 3 else m0  # pass A lines: k1 (3D) or k0 (2D)
    N = n[2] if dim == 3 else n[1]  # pass A line length
    lpc = max(8, min(128, 4096 // N))  # lines per tile of the pass kernels
    chunk = -(-A // n_ranks)
    chunk = -(-chunk // lpc) * lpc
    b0, b1 = min(A, chunk * rank), min(A, chunk * (rank + 1))
    layers = cells[-1] // n_ranks
    return dict(a_begin=b0, a_count=b1 - b0, layers=layers, layer_begin=rank * layers, lines=A, line_length=N)


def default_load_balance(cells, n_ranks):
    """DefaultLoadBalance::loadbalance (fieldtraits.hh:497-561): processors per
    dimension minimising the largest local extent, exact divisibility only;
    the recursion tries the last dimension first and keeps the FIRST optimum."""
    dim = len(cells)
    best = [None, 1e100]
    trial = [0] * dim

    def rec(i, P):
        if i > 0:
            for k in range(1, P + 1):
                if P % k == 0 and cells[i] % k == 0:
                    trial[i] = k
                    rec(i - 1, P // k)
        elif cells[0] % P == 0:
            trial[0] = P
            m = max(cells[k] / trial[k] for k in range(dim))
            if m < best[1]:
                best[1] = m
                best[0] = list(trial)

    rec(dim - 1, n_ranks)
    return best[0]


def eval_block(cells, proc_per_dim, rank):
    """localEvalCells / localEvalOffset of StochasticPart::update (stochastic.hh:237-257)"""
    dim = len(cells)
    le = [c // p for c, p in zip(cells, proc_per_dim)]
    p = list(proc_per_dim) + [1] * (3 - dim)
    if dim == 3:
        lo = [(rank % (p[0] * p[1])) % p[0] * le[0], (rank % (p[0] * p[1])) // p[0] * le[1], rank // (p[0] * p[1]) * le[2]]
    elif dim == 2:
        lo = [rank % p[0] * le[0], rank // p[0] * le[1]]
    else:
        lo = [rank * le[0]]
    return le, lo


def stream_slab_layout(cells, factor, n_ranks, rank):
    """Decomposition of the streaming multi-GPU route (pf_api.cu plan_resize,
    pf_stream.cuh): rank r takes the planes [plane_begin, plane_begin + planes)
    of k0 through BOTH first inverse axes and stores the twice-pruned lines
    (k0, x2, x1 < c1) into the buffer of the rank owning layer x2; every rank
    then finishes the last axis on its layers.  None if the shape is not
    eligible (cubic torus of 512 / 1024 per axis, factor 2, whole B-items per
    owner)."""
    cells = list(cells)
    if len(cells) != 3 or factor != 2 or cells[2] % n_ranks:
        return None
    n = [factor * c for c in cells]
    if n[0] not in (512, 1024) or n[1] != n[0] or n[2] != n[0]:
        return None
    lpg = 4 if n[0] == 1024 else 8
    layers = cells[2] // n_ranks
    m0 = n[0] // 2 + 1
    if layers % lpg or m0 < n_ranks:
        return None
    base, extra = divmod(m0, n_ranks)
    planes = base + (1 if rank < extra else 0)
    begin = base * rank + min(rank, extra)
    # bytes this rank sends to its peers per realization
    sent = 16 * planes * (cells[2] - layers) * cells[1]
    return dict(plane_begin=begin, planes=planes, layers=layers, layer_begin=rank * layers, sent_bytes=sent)


class Plan:
    """Owner of one pf_plan (the CUDA counterpart of parafields::Matrix)."""

    def __init__(self, cells, **kw):
        self.cells = list(cells)
        self.dim = len(self.cells)
        self.desc = make_desc(cells, **kw)
        self._h = C.c_void_p()
        _check(lib().pf_plan_create(C.byref(self.desc), C.byref(self._h)))

    def close(self):
        if self._h:
            lib().pf_plan_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- plumbing ---------------------------------------------------------
    def set_stream(self, stream):
        """stream: torch.cuda.Stream, raw cudaStream_t int, or None."""
        h = 0 if stream is None else getattr(stream, "cuda_stream", stream)
        _check(lib().pf_plan_set_stream(self._h, C.c_void_p(h)))

    def synchronize(self):
        _check(lib().pf_plan_synchronize(self._h))

    def set_profiling(self, on=True):
        _check(lib().pf_plan_set_profiling(self._h, int(on)))

    def pass_times(self):
        """{pass kind: (total ms, launches)} measured with CUDA events."""
        ms = (C.c_double * 16)()
        cnt = (C.c_ulonglong * 16)()
        _check(lib().pf_plan_get_pass_times(self._h, ms, cnt))
        return {i: (ms[i], int(cnt[i])) for i in range(16) if cnt[i]}

    @property
    def launch_count(self):
        return int(lib().pf_plan_launch_count(self._h))

    @property
    def domain_size(self):
        return int(lib().pf_domain_size(self._h))

    @property
    def extended_size(self):
        return int(lib().pf_extended_size(self._h))

    @property
    def noise_entries(self):
        return int(lib().pf_noise_entries(self._h))

    @property
    def local_domain_size(self):
        return int(lib().pf_local_domain_size(self._h))

    @property
    def shape(self):
        """shape of what this plan's generate() returns: the whole field, or
        this rank's slab (last dimension split) in slab mode"""
        sh = list(reversed(self.cells))
        if self.desc.n_ranks > 1:
            sh[0] //= self.desc.n_ranks
        return tuple(sh)

    # -- slab mode (one process per GPU) ---------------------------------------
    def slab_handle(self):
        buf = C.create_string_buffer(64)
        _check(lib().pf_slab_handle(self._h, buf))
        return buf.raw

    def slab_connect(self, handles):
        """handles: list of the 64-byte handles of all ranks, in rank order"""
        blob = b"".join(handles)
        _check(lib().pf_slab_connect(self._h, blob, len(handles)))

    def connect(self, group=None):
        """exchange the IPC handles through torch.distributed and map the peers"""
        import torch.distributed as dist

        handles = [None] * dist.get_world_size(group)
        dist.all_gather_object(handles, self.slab_handle(), group=group)
        self.slab_connect(handles)
        dist.barrier(group)

    # -- slab <-> block redistribution with overlap (multi-GPU plans) -----------
    def _ppd(self, proc_per_dim):
        return (C.c_int * 3)(*(list(proc_per_dim) + [1] * (3 - self.dim)))

    def overlap_size(self, proc_per_dim):
        return int(lib().pf_field_overlap_size(self._h, self._ppd(proc_per_dim)))

    def slab_to_blocks(self, proc_per_dim, slab, eval_out=None, overlap_out=None):
        """StochasticPart::dataToEval + exchangeOverlap (stochastic.hh:830-925,
        1016-1127): returns (evalVector, overlap) of this rank"""
        if isinstance(slab, np.ndarray):
            slab = np.ascontiguousarray(slab, dtype=np.float64)
            if eval_out is None:
                eval_out = np.empty(self.local_domain_size)
            if overlap_out is None:
                overlap_out = np.empty(self.overlap_size(proc_per_dim))
        _check(lib().pf_field_slab_to_blocks(self._h, self._ppd(proc_per_dim), _ptr(slab), _ptr(eval_out), _ptr(overlap_out)))
        return eval_out, overlap_out

    def blocks_to_slab(self, proc_per_dim, eval_in, slab_out=None):
        """StochasticPart::evalToData (stochastic.hh:935-1005)"""
        if isinstance(eval_in, np.ndarray):
            eval_in = np.ascontiguousarray(eval_in, dtype=np.float64)
            if slab_out is None:
                slab_out = np.empty(self.shape)
        _check(lib().pf_field_blocks_to_slab(self._h, self._ppd(proc_per_dim), _ptr(eval_in), _ptr(slab_out)))
        return slab_out

    # -- eigenvalues ---------------------------------------------------------
    def set_covariance_table(self, values):
        """custom covariance: its values on the extended domain, natural order"""
        if isinstance(values, np.ndarray):
            values = np.ascontiguousarray(values, dtype=np.float64)
            n = values.size
        else:
            n = values.numel()
        _check(lib().pf_set_covariance_table(self._h, _ptr(values), n))

    def compute_eigenvalues(self):
        st = PfEigStats()
        rc = lib().pf_compute_eigenvalues(self._h, C.byref(st))
        self.stats = dict(
            small=int(st.small_count),
            zero=int(st.zero_count),
            negative=int(st.negative_count),
            smallest=float(st.smallest),
            factor_used=int(st.factor_used),
        )
        _check(rc)
        return self.stats

    def eigenvalues(self, layout=LAMBDA_NATURAL):
        n = int(lib().pf_eigenvalue_count(self._h, layout))
        out = np.empty(n, dtype=np.float64)
        _check(lib().pf_get_eigenvalues(self._h, _ptr(out), n, layout))
        return out

    # -- sampling ------------------------------------------------------------
    def generate(self, seed, realization=0, out=None, apply_transform=False):
        if out is None:
            out = np.empty(self.shape, dtype=np.float64)
        _check(lib().pf_generate(self._h, seed, realization, _ptr(out), int(apply_transform)))
        return out

    def generate_batch(self, seed, first, count, out=None, apply_transform=False):
        if out is None:
            out = np.empty((count,) + self.shape, dtype=np.float64)
        _check(lib().pf_generate_batch(self._h, seed, first, count, _ptr(out), int(apply_transform)))
        return out

    def generate_from_noise(self, noise, out=None, spare=None):
        if isinstance(noise, np.ndarray):
            noise = np.ascontiguousarray(noise, dtype=np.float64)
            assert noise.size == 2 * self.noise_entries
        if out is None:
            out = np.empty(self.shape, dtype=np.float64)
        _check(lib().pf_generate_from_noise(self._h, _ptr(noise), _ptr(out), _ptr(spare)))
        return out

    def generate_uncorrelated(self, seed, out=None):
        if out is None:
            out = np.empty(self.shape, dtype=np.float64)
        _check(lib().pf_generate_uncorrelated(self._h, seed, _ptr(out)))
        return out

    def variance_field(self, out=None):
        if out is None:
            out = np.empty(self.shape, dtype=np.float64)
        _check(lib().pf_set_variance_as_field(self._h, _ptr(out)))
        return out

    # -- operators -------------------------------------------------------------
    def apply(self, op, x, out=None):
        if isinstance(x, np.ndarray):
            x = np.ascontiguousarray(x, dtype=np.float64)
        if out is None:
            out = np.empty(self.shape, dtype=np.float64)
        _check(lib().pf_apply(self._h, OP[op], _ptr(x), _ptr(out)))
        return out

    def solve(self, b, out=None):
        if isinstance(b, np.ndarray):
            b = np.ascontiguousarray(b, dtype=np.float64)
        if out is None:
            out = np.empty(self.shape, dtype=np.float64)
        it = C.c_uint()
        _check(lib().pf_solve(self._h, _ptr(b), _ptr(out), C.byref(it)))
        self.cg_iterations_used = it.value
        return out

    def transform(self, x, out=None):
        if isinstance(x, np.ndarray):
            x = np.ascontiguousarray(x, dtype=np.float64)
            n = x.size
            if out is None:
                out = np.empty_like(x)
        else:
            n = x.numel()
        _check(lib().pf_transform(self._h, _ptr(x), _ptr(out), n))
        return out

    # -- resolution transfer, localisation, point evaluation ---------------------
    # cells are given x-fastest (like grid.cells); numpy arrays are C-ordered,
    # i.e. shaped reversed(cells)
    @staticmethod
    def _cells(cells):
        return (C.c_uint * 3)(*(list(cells) + [1] * (3 - len(cells))))

    def field_refine(self, cells, x, out=None):
        """StochasticPart::refine (stochastic.hh:597-670); cells = old cells."""
        if isinstance(x, np.ndarray):
            x = np.ascontiguousarray(x, dtype=np.float64)
            if out is None:
                out = np.empty(tuple(2 * c for c in reversed(cells)), dtype=np.float64)
        _check(lib().pf_field_refine(self._h, len(cells), self._cells(cells), _ptr(x), _ptr(out)))
        return out

    def field_coarsen(self, cells, x, out=None):
        """StochasticPart::coarsen (stochastic.hh:679-755); cells = old (fine) cells."""
        if isinstance(x, np.ndarray):
            x = np.ascontiguousarray(x, dtype=np.float64)
            if out is None:
                out = np.empty(tuple(c // 2 for c in reversed(cells)), dtype=np.float64)
        _check(lib().pf_field_coarsen(self._h, len(cells), self._cells(cells), _ptr(x), _ptr(out)))
        return out

    def field_localize(self, cells, extensions, center, radius, x, offset=None, global_cells=None):
        """StochasticPart::localize (stochastic.hh:798-817), in place."""
        d = len(cells)
        ext = (C.c_double * 3)(*(list(extensions) + [1.0] * (3 - d)))
        cen = (C.c_double * 3)(*(list(center) + [0.0] * (3 - d)))
        off = self._cells(offset) if offset is not None else None
        gc = self._cells(global_cells) if global_cells is not None else None
        _check(lib().pf_field_localize(self._h, d, self._cells(cells), off, gc, ext, cen, radius, _ptr(x)))
        return x

    def field_evaluate(self, cells, extensions, field, locations, apply_transform=True, out=None, trend=None):
        """RandomField::evaluate = T(stochastic + trend) for n x dim locations
        (stochastic.hh:480-574, trend.hh:601-634,1491-1505, randomfield.hh:585-595)."""
        d = len(cells)
        ext = (C.c_double * 3)(*(list(extensions) + [1.0] * (3 - d)))
        if isinstance(locations, np.ndarray):
            locations = np.ascontiguousarray(locations, dtype=np.float64)
            npts = locations.size // d
            if out is None:
                out = np.empty(npts, dtype=np.float64)
        else:
            npts = locations.numel() // d
        if isinstance(field, np.ndarray):
            field = np.ascontiguousarray(field, dtype=np.float64)
        nt, tr = _trend_array(trend)
        _check(lib().pf_field_evaluate_trend(self._h, d, self._cells(cells), ext, _ptr(field), npts, _ptr(locations), nt, tr, int(bool(apply_transform)), _ptr(out)))
        return out

    def field_bulk_evaluate(self, cells, extensions, field, trend=None, apply_transform=True, out=None):
        """RandomField::bulkEvaluate (randomfield.hh:597-631) in one fused sweep."""
        d = len(cells)
        ext = (C.c_double * 3)(*(list(extensions) + [1.0] * (3 - d)))
        if isinstance(field, np.ndarray):
            field = np.ascontiguousarray(field, dtype=np.float64)
            if out is None:
                out = np.empty(tuple(reversed(cells)), dtype=np.float64)
        nt, tr = _trend_array(trend)
        _check(lib().pf_field_bulk_evaluate(self._h, d, self._cells(cells), ext, _ptr(field), nt, tr, int(bool(apply_transform)), _ptr(out)))
        return out

    # -- vector space -----------------------------------------------------------
    def axpy(self, alpha, x, y):
        n = x.size if isinstance(x, np.ndarray) else x.numel()
        _check(lib().pf_vec_axpy(self._h, n, alpha, _ptr(x), _ptr(y)))
        return y

    def scale(self, alpha, x):
        n = x.size if isinstance(x, np.ndarray) else x.numel()
        _check(lib().pf_vec_scale(self._h, n, alpha, _ptr(x)))
        return x

    def _red(self, fn, x, y=None):
        n = x.size if isinstance(x, np.ndarray) else x.numel()
        r = C.c_double()
        if y is None:
            _check(fn(self._h, n, _ptr(x), C.byref(r)))
        else:
            _check(fn(self._h, n, _ptr(x), _ptr(y), C.byref(r)))
        return r.value

    def dot(self, x, y):
        return self._red(lib().pf_vec_dot, x, y)

    def norm1(self, x):
        return self._red(lib().pf_vec_norm1, x)

    def norminf(self, x):
        return self._red(lib().pf_vec_norminf, x)

    def dot_static(self, x, y, result=None):
        """pf_vec_dot_static: no allocation inside the library; x, y device
        tensors, result an optional 1-element device tensor (asynchronous)."""
        n = x.numel()
        if result is not None:
            _check(lib().pf_vec_dot_static(self._h, n, _ptr(x), _ptr(y), _ptr(result)))
            return result
        r = C.c_double()
        _check(lib().pf_vec_dot_static(self._h, n, _ptr(x), _ptr(y), C.byref(r)))
        return r.value

    def equal(self, x, y):
        n = x.size if isinstance(x, np.ndarray) else x.numel()
        same = C.c_int()
        _check(lib().pf_vec_equal(self._h, n, _ptr(x), _ptr(y), C.byref(same)))
        return bool(same.value)

    def fill(self, value, x):
        n = x.size if isinstance(x, np.ndarray) else x.numel()
        _check(lib().pf_vec_fill(self._h, n, value, _ptr(x)))
        return x


def device_count():
    return int(lib().pf_device_count())


def device_info(device=0):
    name = C.create_string_buffer(256)
    sm = C.c_int()
    mem = C.c_size_t()
    _check(lib().pf_device_info(device, name, C.byref(sm), C.byref(mem)))
    return dict(name=name.value.decode(), sm_count=sm.value, mem_bytes=mem.value)
