"""pf_vec_dot_static: the scalar product with its scratch in a file-scope
device array (no cudaMalloc inside the call).  No counterpart in the
reference.  Checked through the
C ABI against pf_vec_dot -- same grid, same summation order, so the results
must agree bit for bit -- on two calls in a row (the second call finds the
first call's partial sums in the shared array), with a host and with a device
result pointer, and the size limit is checked to be refused.

Collected last on purpose (zz).
"""
import ctypes as C

import numpy as np
import pytest

from util import pfb

pytestmark = pytest.mark.gpu


def test_dot_static_matches_dot_twice_in_a_row():
    import torch

    m = pfb()
    with m.Plan([16, 16], corr_length=0.1) as p:
        g = torch.Generator(device="cuda").manual_seed(7)
        # second size: fewer blocks than the first, so stale partial sums of
        # call one sit behind the entries call two uses
        for n in (1 << 20, 12345, 1):
            x = torch.randn(n, dtype=torch.float64, device="cuda", generator=g)
            y = torch.randn(n, dtype=torch.float64, device="cuda", generator=g)
            want = p.dot(x, y)
            first = p.dot_static(x, y)
            second = p.dot_static(x, y)
            assert first == want and second == want  # bit-identical
            host = float(np.dot(x.cpu().numpy(), y.cpu().numpy()))
            scale = float(torch.sum(torch.abs(x * y)))
            assert abs(first - host) <= 1e-13 * scale  # tolerance of test_gpu_noise.py
            # device result pointer: asynchronous on the plan's stream
            r = torch.zeros(1, dtype=torch.float64, device="cuda")
            p.dot_static(x, y, result=r)
            p.dot_static(x, y, result=r)
            p.synchronize()
            assert float(r[0]) == want


def test_dot_static_limits():
    import torch

    m = pfb()
    L = m.lib()
    with m.Plan([16, 16], corr_length=0.1) as p:
        x = torch.ones(8, dtype=torch.float64, device="cuda")
        r = C.c_double()
        # beyond the stated maximum: refused before anything is read
        rc = L.pf_vec_dot_static(p._h, (1 << 28) + 1, x.data_ptr(), x.data_ptr(), C.byref(r))
        assert rc == m.PF_ERR_INVALID
        assert b"maximum" in L.pf_last_error()
        # host operands would need a staging buffer
        h = np.ones(8)
        rc = L.pf_vec_dot_static(p._h, 8, h.ctypes.data, h.ctypes.data, C.byref(r))
        assert rc == m.PF_ERR_INVALID
        assert b"device arrays" in L.pf_last_error()
        # and the array still works afterwards
        assert p.dot_static(x, x) == 8.0
