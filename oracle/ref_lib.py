"""ctypes wrapper over oracle/_ref/libpf_ref.so -- the UNMODIFIED reference
headers compiled against single-rank stand-ins for FFTW3-MPI / MPI /
dune-common (oracle/ref_driver.cpp, oracle/shim/).  TEST INFRASTRUCTURE: used
only by tests/ and tests/golden/make_golden.py to pin the oracle.  The library
exists only where /root/reference was present at build time (this container);
callers must skip when `available()` is false."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_ref", "libpf_ref.so")
_lib = None


def available():
    return os.path.exists(LIB)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB)
        L.ref_last_error.restype = C.c_char_p
        L.ref_create.restype = C.c_void_p
        L.ref_create.argtypes = [C.c_char_p]
        L.ref_destroy.argtypes = [C.c_void_p]
        for name in ("ref_generate", "ref_generate_uncorrelated"):
            getattr(L, name).argtypes = [C.c_void_p, C.c_uint]
        L.ref_size.restype = C.c_size_t
        L.ref_size.argtypes = [C.c_void_p]
        for name in ("ref_get", "ref_set", "ref_lambda", "ref_stored_lambda", "ref_bulk_evaluate"):
            getattr(L, name).argtypes = [C.c_void_p, C.c_void_p]
        L.ref_times.argtypes = [C.c_void_p, C.c_int]
        for name in ("ref_lambda_size", "ref_stored_lambda_size"):
            getattr(L, name).restype = C.c_longlong
            getattr(L, name).argtypes = [C.c_void_p]
        L.ref_dot_self.restype = C.c_double
        L.ref_dot_self.argtypes = [C.c_void_p]
        L.ref_rng.argtypes = [C.c_uint, C.c_size_t, C.c_void_p]
        L.ref_refine.argtypes = [C.c_void_p]
        L.ref_coarsen.argtypes = [C.c_void_p]
        L.ref_evaluate.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
        L.ref_write.argtypes = [C.c_void_p, C.c_char_p]
        L.ref_read.restype = C.c_void_p
        L.ref_read.argtypes = [C.c_char_p, C.c_int]
        _lib = L
    return _lib


def ini_text(cells, covariance="exponential", corr_length=0.05, extensions=None, variance=1.0, anisotropy="none", factor=None, approximate=False, periodic=False, transform=None, boxcox_lambda=None, cg_iterations=None, cov_params=None, cache_inv=False, cache_inv_root=False, threshold=None, extra=None):
    """ini file contents in the reference's own keys."""
    dim = len(cells)
    ext = extensions or [1.0] * dim
    cl = corr_length if isinstance(corr_length, (list, tuple)) else [corr_length]
    lines = ["[grid]", "extensions = " + " ".join(repr(float(e)) for e in ext), "cells = " + " ".join(str(int(c)) for c in cells)]
    lines += ["[stochastic]", "variance = %r" % float(variance), "covariance = " + covariance, "corrLength = " + " ".join(repr(float(x)) for x in cl)]
    if anisotropy != "none":
        lines.append("anisotropy = " + anisotropy)
    for k, v in (cov_params or {}).items():
        lines.append("%s = %r" % (k, float(v)))
    lines.append("[embedding]")
    if factor is not None:
        lines.append("factor = %d" % factor)
    if approximate:
        lines.append("approximate = 1")
    if threshold is not None:
        lines.append("threshold = %r" % threshold)
    lines.append("[randomField]")
    if periodic:
        lines.append("periodic = 1")
    if transform:
        lines.append("transform = " + transform)
    if cg_iterations is not None:
        lines.append("cgIterations = %d" % cg_iterations)
    if cache_inv:
        lines.append("cacheInvMatvec = 1")
    if cache_inv_root:
        lines.append("cacheInvRootMatvec = 1")
    if boxcox_lambda is not None:
        lines += ["[transform]", "boxCoxLambda = %r" % float(boxcox_lambda)]
    if extra:
        lines.append(extra.strip())
    return "\n".join(lines) + "\n"


class RefError(RuntimeError):
    pass


class Reference:
    """One parafields::RandomField<GridTraits<double,double,dim>> of the reference."""

    OPS = {"matrix": 0, "root": 1, "inverse": 2, "inv_root": 3}

    def __init__(self, cells, **kw):
        self.cells = list(cells)
        self.text = ini_text(cells, **kw)
        self.h = lib().ref_create(self.text.encode())
        if not self.h:
            raise RefError(lib().ref_last_error().decode())

    def close(self):
        if self.h:
            lib().ref_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise RefError(lib().ref_last_error().decode())

    @property
    def shape(self):
        return tuple(reversed(self.cells))

    def generate(self, seed):
        self._ck(lib().ref_generate(self.h, seed))
        return self.get()

    def generate_uncorrelated(self, seed):
        self._ck(lib().ref_generate_uncorrelated(self.h, seed))
        return self.get()

    def get(self):
        out = np.empty(lib().ref_size(self.h))
        self._ck(lib().ref_get(self.h, out.ctypes.data))
        return out.reshape(self.shape)

    def set(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64).ravel()
        assert x.size == int(np.prod(self.cells))
        self._ck(lib().ref_set(self.h, x.ctypes.data))

    def times(self, op):
        self._ck(lib().ref_times(self.h, self.OPS[op]))
        return self.get()

    def lambda_stream(self):
        """eigenvalues in the order generateField consumes them."""
        n = lib().ref_lambda_size(self.h)
        if n < 0:
            raise RefError(lib().ref_last_error().decode())
        out = np.empty(n)
        self._ck(lib().ref_lambda(self.h, out.ctypes.data))
        return out

    def lambda_stored(self):
        n = lib().ref_stored_lambda_size(self.h)
        if n < 0:
            raise RefError(lib().ref_last_error().decode())
        out = np.empty(n)
        self._ck(lib().ref_stored_lambda(self.h, out.ctypes.data))
        return out

    def bulk_evaluate(self):
        out = np.empty(int(np.prod(self.cells)))
        self._ck(lib().ref_bulk_evaluate(self.h, out.ctypes.data))
        return out.reshape(self.shape)

    def dot_self(self):
        return lib().ref_dot_self(self.h)

    def refine(self):
        """refineMatrix + refine (randomfield.hh:825-880); cells double."""
        self._ck(lib().ref_refine(self.h))
        self.cells = [2 * c for c in self.cells]
        return self.get()

    def coarsen(self):
        """coarsenMatrix + coarsen (randomfield.hh:890-945); cells halve."""
        self._ck(lib().ref_coarsen(self.h))
        self.cells = [c // 2 for c in self.cells]
        return self.get()

    def write(self, base):
        """RandomField::writeToFile (randomfield.hh:656-663): <base>.stoch.h5
        (+ .xdmf), <base>.trend, <base>.field"""
        self._ck(lib().ref_write(self.h, str(base).encode()))

    @classmethod
    def read(cls, base, cells):
        """the file-reading constructor RandomField(fileName) (randomfield.hh:190-225)"""
        self = cls.__new__(cls)
        self.cells = list(cells)
        self.text = None
        self.h = lib().ref_read(str(base).encode(), len(cells))
        if not self.h:
            raise RefError(lib().ref_last_error().decode())
        return self

    def evaluate(self, locations):
        """RandomField::evaluate (randomfield.hh:585-595) at n x dim locations."""
        loc = np.ascontiguousarray(locations, dtype=np.float64).reshape(-1, len(self.cells))
        out = np.empty(loc.shape[0])
        self._ck(lib().ref_evaluate(self.h, loc.shape[0], loc.ctypes.data, out.ctypes.data))
        return out


def rng(seed, n):
    out = np.empty(n)
    lib().ref_rng(seed, n, out.ctypes.data)
    return out
