# K4-size covariance operators (2D 8192^2 cells, 16384^2 torus): time of one matvec
import importlib, json, sys, time
sys.path.insert(0, ".")
import torch
m = importlib.import_module("dune-parafields_b200")
case = dict(cells=[8192, 8192], covariance="cubic", corr_length=[0.08, 0.03, 0.03, 0.05], anisotropy="geometric", approximate=True)
with m.Plan(**case) as p:
    st = p.compute_eigenvalues()
    x = torch.empty(p.domain_size, dtype=torch.float64, device="cuda")
    p.generate(1, out=x)
    y = torch.empty_like(x)
    for op in ("matrix", "root"):
        p.apply(op, x, out=y); p.synchronize()
        t0 = time.perf_counter()
        for _ in range(3):
            p.apply(op, x, out=y)
        p.synchronize()
        ms = (time.perf_counter() - t0) / 3 * 1e3
        print(json.dumps({"workload": "K4", "what": "timesMatrix" if op == "matrix" else "timesMatrixRoot", "ms": ms,
                          "finite": bool(torch.isfinite(y).all()), "norm": float(y.norm())}))
