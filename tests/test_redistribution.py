"""Host-side restatement of the reference's slab -> block redistribution
(StochasticPart::dataToEval, stochastic.hh:830-925: slice re-sort + exchange
inside groups of procPerDim[0] * procPerDim[1] ranks) in numpy, used to PIN what
the device implementation (pf_field_slab_to_blocks: a direct gather from the
peers' slabs) must produce: every rank's evalVector is the block of the global
field at localEvalOffset in natural order, and the overlap faces are the
periodic neighbours' boundary layers.  Also DefaultLoadBalance
(fieldtraits.hh:497-561).  No GPU needed."""
import numpy as np
import pytest

from util import pfb


def reference_data_to_eval(global_field, cells, ppd, P):
    """dataToEval of all P ranks, transcribed index formula by index formula"""
    dim = len(cells)
    cl = cells[-1] // P
    out = []
    resorted_all = []
    for rank in range(P):
        data = global_field[rank * cl : (rank + 1) * cl].ravel()  # this rank's slab
        local_cells = list(cells[:-1]) + [cl]
        D = data.size
        num_slices = ppd[0] * D // local_cells[0]
        slice_size = D // num_slices
        res = np.zeros(D)
        if dim == 3:
            px, py = ppd[0], ppd[1]
            ny, nz = local_cells[1], local_cells[2]
            dy = ny // py
            for i in range(num_slices):
                i_new = (i % px) * (dy * nz) + ((i // (dy * px) * dy) % ny) * (nz * px) + (i // (ny * px)) * dy + (i // px) % dy
                res[i_new * slice_size : (i_new + 1) * slice_size] = data[i * slice_size : (i + 1) * slice_size]
        else:
            for i in range(num_slices):
                i_new = i // ppd[0] + (i % ppd[0]) * local_cells[-1]
                res[i_new * slice_size : (i_new + 1) * slice_size] = data[i * slice_size : (i + 1) * slice_size]
        resorted_all.append(res)
    num_comms = ppd[0] * ppd[1] if dim == 3 else ppd[0]
    for rank in range(P):
        D = resorted_all[rank].size
        chunk = D // num_comms
        ev = np.zeros(D)
        group = rank // num_comms * num_comms
        me = rank - group
        for i in range(num_comms):  # receive chunk `me` of rank group + i
            ev[i * chunk : (i + 1) * chunk] = resorted_all[group + i][me * chunk : (me + 1) * chunk]
        out.append(ev)
    return out


@pytest.mark.parametrize("cells,P,ppd", [([8, 12, 16], 4, [2, 2, 1]), ([8, 12, 16], 8, [2, 2, 2]), ([6, 8, 8], 8, [1, 2, 4]), ([12, 16], 4, [2, 2]), ([12, 16], 4, [1, 4]), ([8, 6, 12], 6, [2, 3, 1])])
def test_reference_resort_yields_natural_blocks(cells, P, ppd):
    m = pfb()
    rng = np.random.default_rng(0)
    field = rng.standard_normal(tuple(reversed(cells)))
    evs = reference_data_to_eval(field, cells, ppd, P)
    for rank in range(P):
        le, lo = m.eval_block(cells, ppd, rank)
        sl = tuple(slice(o, o + e) for o, e in zip(reversed(lo), reversed(le)))
        assert np.array_equal(evs[rank], field[sl].ravel()), (rank, le, lo)


def test_default_load_balance():
    m = pfb()
    # exact divisibility only, largest local extent minimised (fieldtraits.hh:529-556)
    assert m.default_load_balance([512, 512, 512], 8) == [2, 2, 2]
    assert m.default_load_balance([512, 512], 4) == [2, 2]
    assert m.default_load_balance([64], 4) == [4]
    assert m.default_load_balance([30, 20, 8], 8) == [2, 4, 1]  # max(15, 5, 8) = 15, first optimum found
    ppd = m.default_load_balance([9, 10, 8], 4)
    assert np.prod(ppd) == 4 and all(c % p == 0 for c, p in zip([9, 10, 8], ppd))
    assert m.default_load_balance([7, 7], 4) is None  # no exact subdivision
