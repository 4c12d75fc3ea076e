#!/usr/bin/env python
"""CPU restatement of the slab-mode data flow on 2 gloo ranks (no GPU):
rank r builds sigma*noise for its pass-A lines, inverse-transforms them along
the last axis (numpy), keeps the physical layers and sends each layer block to
the rank owning it; every rank then finishes passes B and C locally.  The
assembled slabs must equal the oracle's single-rank field.  This checks the
ownership arithmetic of `slab_layout` / pf_api.cu plan_resize and that the
post-exchange layout is the reference's field distribution."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run_case(dist, torch, pfb, ol, cells, **kw):
    rank, world = dist.get_rank(), dist.get_world_size()
    o = ol.Oracle(cells, **kw)
    dim = o.dim
    n = o.n
    m0 = n[0] // 2 + 1
    lay = pfb.slab_layout(cells, 2, world, rank)
    lam = o.lambda_natural()  # [k2][k1][k0] (2D: [1][k1][k0])
    noise = o.reference_noise(4242).reshape(-1, 2)
    if dim == 3:
        # stream order s = k0 + m0*(k2 + n2*k1)  
        xi = (noise[:, 0] + 1j * noise[:, 1]).reshape(n[1], n[2], m0)  # [k1][k2][k0]
        sig = np.sqrt(lam).transpose(1, 0, 2)  # [k1][k2][k0]
    else:
        xi = (noise[:, 0] + 1j * noise[:, 1]).reshape(n[1], m0)  # [k1][k0]
        sig = np.sqrt(lam[0])
    X = sig * xi / np.sqrt(2.0)
    # self-conjugate entries: real, full weight (r2cfieldbackend.hh:305-325)
    idx = np.indices(X.shape)
    if dim == 3:
        selfc = ((2 * idx[0]) % n[1] == 0) & ((2 * idx[1]) % n[2] == 0) & ((2 * idx[2]) % n[0] == 0)
    else:
        selfc = ((2 * idx[0]) % n[1] == 0) & ((2 * idx[1]) % n[0] == 0)
    X = np.where(selfc, (sig * xi.real).astype(complex), X)
    a0, a1 = lay["a_begin"], lay["a_begin"] + lay["a_count"]
    L = lay["layers"]
    if dim == 3:
        # pass A on my k1 lines: along k2, keep x2 < c2 -> [k1 mine][x2][k0]
        mine = np.fft.ifft(X[a0:a1], axis=1)[:, : cells[2], :] * n[2]
        shape_full = (n[1], L, m0)
    else:
        # pass A on my k0 lines: along k1, keep x1 < c1 -> [x1][k0 mine]
        mine = np.fft.ifft(X[:, a0:a1], axis=0)[: cells[1], :] * n[1]
        shape_full = (L, m0)
    # exchange: block (my lines, layers of rank q) goes to rank q
    recv = np.zeros(shape_full, dtype=complex)
    for src in range(world):
        lay_s = pfb.slab_layout(cells, 2, world, src)
        s0, s1 = lay_s["a_begin"], lay_s["a_begin"] + lay_s["a_count"]
        for dst in range(world):
            if dim == 3:
                shp = (s1 - s0, L, m0)
            else:
                shp = (L, s1 - s0)
            buf = torch.zeros(shp, dtype=torch.complex128)
            if src == rank:
                blk = mine[:, dst * L : (dst + 1) * L, :] if dim == 3 else mine[dst * L : (dst + 1) * L, :]
                buf = torch.from_numpy(np.ascontiguousarray(blk))
            if src == dst:
                got = buf if rank == src else None
            else:
                if rank == src:
                    dist.send(torch.view_as_real(buf).contiguous(), dst)
                    got = None
                elif rank == dst:
                    t = torch.zeros(shp + (2,), dtype=torch.float64)
                    dist.recv(t, src)
                    got = torch.view_as_complex(t)
                else:
                    got = None
            if got is not None and rank == dst and s1 > s0:
                if dim == 3:
                    recv[s0:s1] = got.numpy()
                else:
                    recv[:, s0:s1] = got.numpy()
    # local passes B and C
    if dim == 3:
        yb = np.fft.ifft(recv, axis=0)[: cells[1]] * n[1]  # [x1][x2 local][k0]
        field = np.fft.irfft(yb, n=n[0], axis=2)[:, :, : cells[0]] * n[0]
        field = field.transpose(1, 0, 2)  # [x2 local][x1][x0]
    else:
        field = np.fft.irfft(recv, n=n[0], axis=1)[:, : cells[0]] * n[0]
    full = o.generate_from_noise(o.reference_noise(4242))
    ref = full[rank * L : (rank + 1) * L]
    err = float(np.max(np.abs(field - ref)) / np.max(np.abs(full)))
    return err, lay


def plane_split_arithmetic(pfb):
    """ownership arithmetic of the streaming route: the planes of all ranks
    tile [0, m0) exactly once, whatever the remainder"""
    for cells in ([256, 256, 256], [512, 512, 512]):
        for world in (2, 4, 8):
            m0 = cells[0] + 1
            nxt = 0
            sent = 0
            for r in range(world):
                lay = pfb.stream_slab_layout(cells, 2, world, r)
                assert lay["plane_begin"] == nxt and lay["planes"] in (m0 // world, m0 // world + 1)
                nxt += lay["planes"]
                assert lay["layers"] * world == cells[2] and lay["layer_begin"] == r * lay["layers"]
                sent += lay["sent_bytes"]
            assert nxt == m0
            # all ranks together move (P-1)/P of pass B's output: HALF of what the
            # transposed transform moves after pass A
            assert sent == 16 * m0 * cells[1] * cells[2] * (world - 1) // world
    assert pfb.stream_slab_layout([128, 128, 128], 2, 2, 0) is None
    assert pfb.stream_slab_layout([256, 256, 256], 2, 64, 0) is None  # 4 layers < one B-item


def run_plane_split(dist, torch, ol, cells, **kw):
    """the streaming multi-GPU route on gloo ranks (numpy FFTs): rank r takes
    its planes of k0 through the first TWO inverse axes, sends every line
    (k0, x2, x1 < c1) to the owner of layer x2, and all ranks finish with the
    last axis -- the result must be the oracle's field on the rank's layers"""
    rank, world = dist.get_rank(), dist.get_world_size()
    o = ol.Oracle(cells, **kw)
    n = o.n
    m0 = n[0] // 2 + 1
    base, extra = divmod(m0, world)
    planes = [base + (1 if r < extra else 0) for r in range(world)]
    begin = [base * r + min(r, extra) for r in range(world)]
    L = cells[2] // world
    lam = o.lambda_natural()
    noise = o.reference_noise(77).reshape(-1, 2)
    xi = (noise[:, 0] + 1j * noise[:, 1]).reshape(n[1], n[2], m0)  # [k1][k2][k0]
    sig = np.sqrt(lam).transpose(1, 0, 2)
    X = sig * xi / np.sqrt(2.0)
    idx = np.indices(X.shape)
    selfc = ((2 * idx[0]) % n[1] == 0) & ((2 * idx[1]) % n[2] == 0) & ((2 * idx[2]) % n[0] == 0)
    X = np.where(selfc, (sig * xi.real).astype(complex), X)
    p0, p1 = begin[rank], begin[rank] + planes[rank]
    mine = X[:, :, p0:p1]  # my planes
    ya = np.fft.ifft(mine, axis=1)[:, : cells[2], :] * n[2]  # [k1][x2][k0 mine]
    yb = np.fft.ifft(ya, axis=0)[: cells[1]] * n[1]  # [x1][x2][k0 mine]
    zt = np.zeros((cells[1], L, m0), dtype=complex)  # my layers, all planes
    for src in range(world):
        for dst in range(world):
            shp = (cells[1], L, planes[src])
            if src == rank:
                blk = np.ascontiguousarray(yb[:, dst * L : (dst + 1) * L, :])
                if dst == rank:
                    zt[:, :, begin[src] : begin[src] + planes[src]] = blk
                else:
                    dist.send(torch.view_as_real(torch.from_numpy(blk)).contiguous(), dst)
            elif dst == rank:
                t = torch.zeros(shp + (2,), dtype=torch.float64)
                dist.recv(t, src)
                zt[:, :, begin[src] : begin[src] + planes[src]] = torch.view_as_complex(t).numpy()
    field = np.fft.irfft(zt, n=n[0], axis=2)[:, :, : cells[0]] * n[0]
    field = field.transpose(1, 0, 2)  # [x2 local][x1][x0]
    full = o.generate_from_noise(o.reference_noise(77))
    ref = full[rank * L : (rank + 1) * L]
    return float(np.max(np.abs(field - ref)) / np.max(np.abs(full)))


def main():
    import torch
    import torch.distributed as dist

    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("gloo")
    pfb = importlib.import_module("dune-parafields_b200")
    from oracle import oracle_lib as ol

    worst = 0.0
    for cells, kw in [
        ([4, 128, 4], dict(covariance="exponential", corr_length=0.1)),
        ([512, 4], dict(covariance="exponential", corr_length=0.05)),
        ([8, 4, 8], dict(covariance="spherical", corr_length=0.2)),
        ([16, 8], dict(covariance="gaussian", corr_length=0.05, approximate=True)),
    ]:
        err, lay = run_case(dist, torch, pfb, ol, cells, **kw)
        print("rank %d cells %s lines [%d,%d) layers %d err %.2e" % (dist.get_rank(), cells, lay["a_begin"], lay["a_begin"] + lay["a_count"], lay["layers"], err), flush=True)
        worst = max(worst, err)
    plane_split_arithmetic(pfb)
    for cells, kw in [([8, 4, 8], dict(covariance="spherical", corr_length=0.2)), ([4, 16, 6], dict(covariance="exponential", corr_length=0.1))]:
        err = run_plane_split(dist, torch, ol, cells, **kw)
        print("rank %d plane split cells %s err %.2e" % (dist.get_rank(), cells, err), flush=True)
        worst = max(worst, err)
    t = torch.tensor([worst], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ok = t.item() < 1e-10
    if dist.get_rank() == 0:
        print("GLOO SLAB %s worst %.2e" % ("OK" if ok else "FAIL", t.item()), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
