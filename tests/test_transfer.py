"""Resolution transfer (refine / coarsen), localize and point evaluation
(stochastic.hh:480-817, randomfield.hh:585-595,
825-945).

CPU part: the oracle restatement reproduces the fixtures recorded from the
UNMODIFIED reference (tests/golden/transfer_*.npz, made by
tests/golden/make_golden_transfer.py) bit for bit, and the live reference
where oracle/_ref exists.  GPU part (-m gpu): the CUDA kernels through the C
ABI reproduce the same fixtures and the oracle on seeded inputs: bit-exact for
refine / coarsen / evaluate without a transcendental transform, 1e-14 where a
device exp() is involved."""
import glob
import os

import numpy as np
import pytest

from oracle import oracle_lib as ol
from oracle import ref_lib as rl
from util import pfb, rel

HERE = os.path.dirname(os.path.abspath(__file__))
FIXTURES = sorted(glob.glob(os.path.join(HERE, "golden", "transfer_*.npz")))
EXP_TOL = 1e-14


def fid(f):
    return os.path.basename(f)[len("transfer_") : -4]


def load(f):
    g = np.load(f)
    cells = [int(c) for c in g["cells"]]
    ext = [float(e) for e in g["extensions"]]
    return g, cells, ext, str(g["transform"])


def same(a, b, transform):
    """bit-exact unless the value went through exp()."""
    if transform == "logNormal":
        return rel(a, b) < EXP_TOL
    return np.array_equal(np.asarray(a).ravel(), np.asarray(b).ravel())


# ---------------------------------------------------------------------------
# CPU: oracle against the reference
# ---------------------------------------------------------------------------
def test_fixtures_present():
    assert len(FIXTURES) == 4


@pytest.mark.parametrize("fixture", FIXTURES, ids=fid)
def test_oracle_matches_reference_fixtures(fixture):
    g, cells, ext, transform = load(fixture)
    fine = [2 * c for c in cells]
    refined = ol.field_refine(cells, g["field"])
    assert np.array_equal(refined, g["refined"])
    assert np.array_equal(ol.field_coarsen(fine, refined), g["refined_coarsened"])
    assert np.array_equal(ol.field_coarsen(cells, g["rough"]), g["rough_coarsened"])
    v = ol.field_evaluate(cells, ext, g["field"], g["locations"], transform)
    assert same(v, g["values"], transform)
    v = ol.field_evaluate(fine, ext, g["refined"], g["locations_fine"], transform)
    assert same(v, g["values_fine"], transform)


@pytest.mark.skipif(not rl.available(), reason="oracle/_ref not built (no /root/reference)")
@pytest.mark.parametrize("cells", [[6, 10], [32, 2], [2, 6, 4], [10, 4, 2]], ids=lambda c: "x".join(map(str, c)))
def test_oracle_matches_live_reference(cells):
    rng = np.random.default_rng(sum(cells))
    r = rl.Reference(cells, covariance="exponential", corr_length=0.3, approximate=True)
    f = rng.standard_normal(tuple(reversed(cells)))
    r.set(f)
    loc = rng.random((200, len(cells)))
    loc[0, :] = 1.0
    loc[1, 0] = 1.0
    assert np.array_equal(r.evaluate(loc), ol.field_evaluate(cells, [1.0] * len(cells), f, loc))
    assert np.array_equal(r.refine(), ol.field_refine(cells, f))
    r.coarsen()
    assert np.array_equal(r.coarsen(), ol.field_coarsen(cells, r_get_before_coarsen(f, cells)))


def r_get_before_coarsen(f, cells):
    # refine followed by coarsen of the reference: what the second coarsen saw
    return ol.field_coarsen([2 * c for c in cells], ol.field_refine(cells, f))


def test_oracle_error_texts_follow_the_reference():
    with pytest.raises(RuntimeError, match="not implemented"):
        ol.field_refine([8], np.zeros(8))
    with pytest.raises(RuntimeError, match="not implemented"):
        ol.field_coarsen([8], np.zeros(8))
    with pytest.raises(RuntimeError, match="cannot coarsen odd number of cells"):
        ol.field_coarsen([6, 3], np.zeros(18))
    with pytest.raises(RuntimeError, match="outside the domain"):
        ol.field_evaluate([4, 4], [1.0, 1.0], np.zeros(16), [[0.5, 1.5]])


@pytest.mark.parametrize("cells,ext", [([16], [2.0]), ([8, 6], [1.0, 0.5]), ([4, 6, 2], [1.0, 2.0, 3.0])], ids=["1d", "2d", "3d"])
def test_oracle_localize_closed_form(cells, ext):
    """stochastic.hh:798-817 restated with numpy (the reference's own code for
    it does not compile, stochastic.hh:806)."""
    d = len(cells)
    rng = np.random.default_rng(d)
    f = rng.standard_normal(tuple(reversed(cells)))
    center = [0.3 * e for e in ext]
    radius = 0.4
    out = ol.field_localize(cells, ext, center, radius, f)
    grids = np.meshgrid(*[(np.arange(cells[i]) + 0.5) * ext[i] / cells[i] for i in reversed(range(d))], indexing="ij")
    dist = sum((g - center[i]) ** 2 for g, i in zip(grids, reversed(range(d))))
    want = f * (2.0 * 3.14159) ** (-(d / 2.0)) * np.exp(-0.5 * dist / radius**2)
    assert rel(out, want) < 1e-15
    # a slab of the same field (offset in the last dimension) agrees with the whole
    if d > 1:
        half = cells[:-1] + [cells[-1] // 2]
        off = [0] * (d - 1) + [cells[-1] // 2]
        part = ol.field_localize(half, ext, center, radius, f[cells[-1] // 2 :], offset=off, global_cells=cells)
        assert np.array_equal(part, out[cells[-1] // 2 :])


# ---------------------------------------------------------------------------
# GPU: CUDA kernels through the C ABI
# ---------------------------------------------------------------------------
@pytest.fixture(scope="module")
def m():
    return pfb()


def plan_for(m, transform="none"):
    kw = {} if transform == "none" else {"transform": transform}
    return m.Plan([8, 8], covariance="exponential", corr_length=0.3, **kw)


@pytest.mark.gpu
@pytest.mark.parametrize("fixture", FIXTURES, ids=fid)
def test_cuda_matches_reference_fixtures(m, fixture):
    g, cells, ext, transform = load(fixture)
    fine = [2 * c for c in cells]
    with plan_for(m, transform) as p:
        refined = p.field_refine(cells, g["field"])
        assert np.array_equal(refined, g["refined"])
        assert np.array_equal(p.field_coarsen(fine, refined), g["refined_coarsened"])
        assert np.array_equal(p.field_coarsen(cells, g["rough"]), g["rough_coarsened"])
        v = p.field_evaluate(cells, ext, g["field"], g["locations"])
        assert same(v, g["values"], transform)
        v = p.field_evaluate(fine, ext, g["refined"], g["locations_fine"])
        assert same(v, g["values_fine"], transform)
        # without the transform: the stochastic part itself
        raw = p.field_evaluate(cells, ext, g["field"], g["locations"], apply_transform=False)
        assert np.array_equal(raw, ol.field_evaluate(cells, ext, g["field"], g["locations"]))


SHAPES = [[2, 2], [6, 10], [33, 7], [1, 5], [130, 66], [2, 2, 2], [5, 3, 7], [34, 6, 10], [1, 1, 3], [64, 32, 16]]


@pytest.mark.gpu
@pytest.mark.parametrize("cells", SHAPES, ids=lambda c: "x".join(map(str, c)))
def test_cuda_refine_coarsen_match_oracle(m, cells):
    import torch

    rng = np.random.default_rng(sum(cells))
    f = rng.standard_normal(tuple(reversed(cells)))
    fine = [2 * c for c in cells]
    with plan_for(m) as p:
        want = ol.field_refine(cells, f)
        assert np.array_equal(p.field_refine(cells, f), want)
        rough = rng.standard_normal(want.shape)
        want_c = ol.field_coarsen(fine, rough)
        assert np.array_equal(p.field_coarsen(fine, rough), want_c)
        # device-resident arrays take the same kernels without staging
        dev = torch.device("cuda")
        d_in = torch.from_numpy(f).to(dev)
        d_out = torch.empty(want.shape, dtype=torch.float64, device=dev)
        p.field_refine(cells, d_in, d_out)
        assert np.array_equal(d_out.cpu().numpy(), want)
        d_r = torch.from_numpy(rough).to(dev)
        d_c = torch.empty(want_c.shape, dtype=torch.float64, device=dev)
        p.field_coarsen(fine, d_r, d_c)
        assert np.array_equal(d_c.cpu().numpy(), want_c)


@pytest.mark.gpu
@pytest.mark.parametrize("cells,ext", [([16], [2.0]), ([33, 6], [1.0, 0.5]), ([12, 6, 20], [1.0, 2.0, 3.0])], ids=["1d", "2d", "3d"])
def test_cuda_localize_and_evaluate_match_oracle(m, cells, ext):
    d = len(cells)
    rng = np.random.default_rng(7 + d)
    f = rng.standard_normal(tuple(reversed(cells)))
    center = [0.3 * e for e in ext]
    with plan_for(m, "boxCox" if d == 2 else "foldedNormal") as p:
        got = p.field_localize(cells, ext, center, 0.4, f.copy())
        assert rel(got, ol.field_localize(cells, ext, center, 0.4, f)) < EXP_TOL
        if d > 1:
            half = cells[:-1] + [cells[-1] // 2]
            off = [0] * (d - 1) + [cells[-1] // 2]
            part = p.field_localize(half, ext, center, 0.4, f[cells[-1] // 2 :].copy(), offset=off, global_cells=cells)
            assert np.array_equal(part, got[cells[-1] // 2 :])
        loc = rng.random((5000, d)) * np.asarray(ext)
        loc[0, :] = ext
        for i in range(d):
            loc[1 + i, i] = ext[i]
        loc[10, :] = 0.0
        raw = p.field_evaluate(cells, ext, f, loc, apply_transform=False)
        assert np.array_equal(raw, ol.field_evaluate(cells, ext, f, loc))
        if d != 2:
            tv = p.field_evaluate(cells, ext, f, loc)
            assert np.array_equal(tv, ol.field_evaluate(cells, ext, f, loc, "foldedNormal"))


@pytest.mark.gpu
def test_cuda_evaluate_on_cell_boundaries(m):
    """coordsToIndices truncates x * (cells + 1e-6) / ext (fieldtraits.hh:445):
    locations on cell boundaries and their floating-point neighbours must land
    in the reference's cell (the kernel's reciprocal fast path falls back to
    the division there)."""
    cells, ext = [48, 37], [0.7, 3.0]
    rng = np.random.default_rng(3)
    f = rng.standard_normal(tuple(reversed(cells)))
    pts = []
    for i in range(2):
        edges = np.arange(cells[i]) * ext[i] / cells[i]
        scaled = np.arange(1, cells[i]) * ext[i] / (cells[i] + 1e-6)
        for e in (edges, scaled):
            for v in (e, np.nextafter(e, np.inf), np.nextafter(e, -np.inf)[1:]):
                loc = rng.random((v.size, 2)) * np.asarray(ext)
                loc[:, i] = v
                pts.append(loc)
    loc = np.concatenate(pts)
    with plan_for(m) as p:
        got = p.field_evaluate(cells, ext, f, loc, apply_transform=False)
    assert np.array_equal(got, ol.field_evaluate(cells, ext, f, loc))


@pytest.mark.gpu
def test_cuda_error_behaviour(m):
    with plan_for(m) as p:
        with pytest.raises(m.PfError, match="not implemented"):
            p.field_refine([8], np.zeros(8))
        with pytest.raises(m.PfError, match="not implemented"):
            p.field_coarsen([8], np.zeros(8))
        with pytest.raises(m.PfError, match="cannot coarsen odd number of cells"):
            p.field_coarsen([6, 3], np.zeros((3, 6)))
        with pytest.raises(m.PfError, match="outside the domain"):
            p.field_evaluate([4, 4], [1.0, 1.0], np.zeros((4, 4)), np.array([[0.5, 1.5]]))
        with pytest.raises(m.PfError, match="outside the domain"):
            p.field_evaluate([4, 4], [1.0, 1.0], np.zeros((4, 4)), np.array([[-0.1, 0.5]]))
        # vector space operations on host arrays run on the device too
        x = np.arange(10.0)
        y = np.ones(10)
        p.axpy(2.0, x, y)
        assert np.array_equal(y, 1.0 + 2.0 * x)
        p.scale(0.5, y)
        assert np.array_equal(y, 0.5 * (1.0 + 2.0 * x))


@pytest.mark.gpu
def test_cuda_full_size_properties(m):
    """256^3 -> 512^3 -> 256^3 (K5 -> K3 cell counts): every child equals its
    parent, the mean of 8 equal children returns the parent within 2 ulp, and
    evaluating at every fine cell centre reproduces the refined field."""
    import torch

    n = 256
    dev = torch.device("cuda")
    g = torch.Generator(device=dev).manual_seed(5)
    f = torch.randn((n, n, n), dtype=torch.float64, device=dev, generator=g)
    fine = torch.empty((2 * n, 2 * n, 2 * n), dtype=torch.float64, device=dev)
    with plan_for(m) as p:
        p.field_refine([n, n, n], f, fine)
        p.synchronize()
        for a in (0, 1):
            for b in (0, 1):
                for c in (0, 1):
                    assert torch.equal(fine[a::2, b::2, c::2], f)
        back = torch.empty_like(f)
        p.field_coarsen([2 * n] * 3, fine, back)
        p.synchronize()
        assert float(((back - f).abs() / f.abs().clamp_min(1e-300)).max()) < 5e-16
        # bulkEvaluate's locations (randomfield.hh:604-631) for one fine plane
        k = 2 * n
        c = (torch.arange(k, dtype=torch.float64, device=dev) + 0.5) * (1.0 / k)
        loc = torch.stack(torch.meshgrid(c[37:38], c, c, indexing="ij"), dim=-1).reshape(-1, 3).flip(-1).contiguous()
        out = torch.empty(loc.shape[0], dtype=torch.float64, device=dev)
        p.field_evaluate([k, k, k], [1.0, 1.0, 1.0], fine, loc, apply_transform=False, out=out)
        p.synchronize()
        assert torch.equal(out.reshape(k, k), fine[37])


# ---------------------------------------------------------------------------
# trend part + bulkEvaluate (trend.hh:601-634,1491-1505; randomfield.hh:585-631)
# ---------------------------------------------------------------------------
TREND_FIXTURES = sorted(glob.glob(os.path.join(HERE, "golden", "trend_*.npz")))


def load_trend(f):
    import json

    g, cells, ext, transform = load(f)
    trend = [(k, c) for k, c in json.loads(str(g["trend"]))]
    return g, cells, ext, transform, trend


def close(a, b):
    """exp() of the disk component / logNormal: libm against the device."""
    return rel(a, b) < EXP_TOL


def test_trend_fixtures_present():
    assert len(TREND_FIXTURES) == 3


@pytest.mark.parametrize("fixture", TREND_FIXTURES, ids=lambda f: os.path.basename(f)[:-4])
def test_oracle_trend_matches_reference_fixtures(fixture):
    g, cells, ext, transform, trend = load_trend(fixture)
    bulk = ol.field_bulk_evaluate(cells, ext, g["field"], trend, transform)
    assert np.array_equal(bulk, g["bulk"])
    v = ol.field_evaluate_trend(cells, ext, g["field"], g["locations"], trend, transform)
    assert np.array_equal(v, g["values"])
    # the trend really contributes
    assert rel(ol.field_bulk_evaluate(cells, ext, g["field"], [], transform), g["bulk"]) > 1e-3


@pytest.mark.gpu
@pytest.mark.parametrize("fixture", TREND_FIXTURES, ids=lambda f: os.path.basename(f)[:-4])
def test_cuda_trend_matches_reference_fixtures(m, fixture):
    g, cells, ext, transform, trend = load_trend(fixture)
    with plan_for(m, transform) as p:
        assert close(p.field_bulk_evaluate(cells, ext, g["field"], trend), g["bulk"])
        assert close(p.field_evaluate(cells, ext, g["field"], g["locations"], trend=trend), g["values"])
        # components without exp() are bit-exact
        plain = [t for t in trend if t[0] != "disk"]
        got = p.field_bulk_evaluate(cells, ext, g["field"], plain, apply_transform=False)
        assert np.array_equal(got, ol.field_bulk_evaluate(cells, ext, g["field"], plain))
        got = p.field_evaluate(cells, ext, g["field"], g["locations"], apply_transform=False, trend=plain)
        assert np.array_equal(got, ol.field_evaluate_trend(cells, ext, g["field"], g["locations"], plain))
        # in place
        buf = g["field"].copy()
        p.field_bulk_evaluate(cells, ext, buf, trend, out=buf)
        assert close(buf, g["bulk"])


@pytest.mark.gpu
def test_cuda_trend_errors(m):
    with plan_for(m) as p:
        f = np.zeros((4, 4))
        with pytest.raises(m.PfError, match="at most 16 trend components"):
            p.field_bulk_evaluate([4, 4], [1.0, 1.0], f, [("mean", [1.0])] * 17)
        d = m.PfTrendComponent()
        d.kind = 9
        import ctypes as C

        rc = m.lib().pf_field_bulk_evaluate(p._h, 2, p._cells([4, 4]), (C.c_double * 3)(1, 1, 1), f.ctypes.data, 1, C.byref(d), 0, f.ctypes.data)
        assert rc == m.PF_ERR_INVALID and m.lib().pf_last_error() == b"Trend component type not found!"


@pytest.mark.gpu
def test_cuda_bulk_evaluate_full_size(m):
    """512^3 (K3's cells): the fused sweep against torch on the same buffer."""
    import torch

    n = 512
    dev = torch.device("cuda")
    gen = torch.Generator(device=dev).manual_seed(11)
    f = torch.randn((n, n, n), dtype=torch.float64, device=dev, generator=gen)
    out = torch.empty_like(f)
    with plan_for(m, "logNormal") as p:
        p.field_bulk_evaluate([n] * 3, [1.0] * 3, f, out=out)
        p.synchronize()
        assert float(((out - torch.exp(f)).abs() / torch.exp(f)).max()) < 1e-15
        trend = [("mean", [0.25]), ("slope", [1.0, 2.0, -3.0])]
        p.field_bulk_evaluate([n] * 3, [1.0] * 3, f, trend, apply_transform=False, out=out)
        p.synchronize()
        c = (torch.arange(n, dtype=torch.float64, device=dev) + 0.5) * (1.0 / n) - 0.5
        want = f + (0.25 + ((1.0 * c)[None, None, :] + (2.0 * c)[None, :, None] + (-3.0 * c)[:, None, None]))
        assert float((out - want).abs().max()) < 1e-14
