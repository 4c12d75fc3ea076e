#!/usr/bin/env python
"""bench.py -- realizations/s of the circulant-embedding sampler on B200.

Contract (see the task statement): `python bench.py --gpus N --steps K
--warmup W` prints ONE JSON line from rank 0.  A "step" is one realization of
the workload (K3 of SURVEY.md section 8: 3D 512^3 cells, spherical covariance,
corrLength 0.1, embedding factor 2 -> 1024^3 fp64 torus).

  value        whole-job realizations/s, fields stay in HBM, timed with CUDA
               events on the launching stream, max over ranks
  e2e          same metric through the public API with a HOST (pinned) output
               buffer: the device->host copy of every field is inside the
               timed region (the only per-step input is the (seed,
               realization) pair passed by value)
  roofline     N = 1: dominant kernel, algorithmic bytes / measured launch
               duration against MEASURED_PEAKS.json (HBM);  N > 1 (slab mode):
               bytes every GPU must send over NVLink per realization / step
               time against the measured 770 GB/s peer-copy figure, the HBM
               view nested below it
  rng_fp64     the same K3 step with rng_mode fp64 (10 Philox rounds + fp64
               Box-Muller, the reference's precision of the normals); the
               headline uses rng_mode fast (7 rounds + fp32 Box-Muller)
  other_workloads  K1, K2, K4 (one by one and as a batch of 64), K5 matvec and
               one PCG iteration: a few steps each with their roofline fraction
  cpu_baseline the CPU oracle (single thread = one MPI rank of the reference)
               on a bounded sample of the same workload
  parity       N > 1: one 64^3 slab-vs-single comparison per rank before the
               timed region; the run fails (exit 1) unless it is exactly 0.0

Multi-GPU (N > 1, launched by torch.distributed.run): `--mode slab` (default)
shards ONE realization over the GPUs (BASELINE.json's configuration, strong
scaling); `--mode replicas` gives every GPU its own realizations (weak).

`--impl reference` times the CPU restatement of the reference path (oracle/,
every host thread, the FULL workload: same config as the GPU arm) and, next to
it, a best-effort all-core variant on a real FFT library (scipy / pocketfft).
The reference itself cannot be built with its own FFT (needs FFTW3-MPI, MPI,
dune-common).
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # SURVEY.md 8(d) synthetic inputs
    "K3": dict(cells=[512, 512, 512], covariance="spherical", corr_length=0.1, label="3D 512^3 spherical l=0.1, factor 2 (1024^3 fp64 torus)"),
    "K2": dict(cells=[256, 256, 256], covariance="gaussian", corr_length=[0.05, 0.03, 0.02], anisotropy="axiparallel", transform="logNormal", approximate=True, label="3D 256^3 gaussian axiparallel + logNormal (512^3 torus)"),
    "K4": dict(cells=[8192, 8192], covariance="cubic", corr_length=[0.08, 0.03, 0.03, 0.05], anisotropy="geometric", approximate=True, label="2D 8192^2 cubic, geometric anisotropy (16384^2 torus, four-step lines)"),
    "K1": dict(cells=[512, 512], covariance="exponential", corr_length=0.05, label="2D 512^2 exponential l=0.05 (1024^2 torus)"),
    "K5": dict(cells=[256, 256, 256], covariance="exponential", corr_length=0.05, label="3D 256^3 exponential (512^3 torus)"),
    "S3": dict(cells=[128, 128, 128], covariance="spherical", corr_length=0.1, label="3D 128^3 spherical (256^3 torus)"),
}
SEED = 20240611


def plan_kwargs(name):
    return {k: v for k, v in WORKLOADS[name].items() if k != "label"}


def algorithmic_bytes(cells, factor=2, compact=True, batch=1):
    """bytes per realization of the pruned pass pipeline and
    of its passes (pass A = noise + first inverse axis, the dominant kernel).
    batch > 1: the sigma table is read once per batch."""
    c = list(cells)
    n = [factor * x for x in c]
    m0 = n[0] // 2 + 1
    if len(c) == 3:
        lam = 8 * ((n[0] // 2 + 1) * (n[1] // 2 + 1) * (n[2] // 2 + 1) if compact else m0 * n[1] * n[2])
        a_w = 16 * m0 * n[1] * c[2]
        b_w = 16 * m0 * c[1] * c[2]
        c_w = 8 * c[0] * c[1] * c[2]
        total = lam // batch + 2 * a_w + 2 * b_w + c_w
        return dict(total=total, pass_a=lam // batch + a_w, pass_b=a_w + b_w, pass_c=b_w + c_w)
    if len(c) == 2:
        lam = 8 * ((n[0] // 2 + 1) * (n[1] // 2 + 1) if compact else m0 * n[1])
        a_w = 16 * m0 * c[1]
        c_w = 8 * c[0] * c[1]
        return dict(total=lam // batch + 2 * a_w + c_w, pass_a=lam // batch + a_w, pass_b=0, pass_c=a_w + c_w)
    return dict(total=8 * n[0] + 16 * c[0], pass_a=8 * n[0] + 16 * c[0], pass_b=0, pass_c=0)


def matvec_bytes(cells, factor=2):
    """algorithmic bytes of one covariance matvec"""
    c = list(cells)
    n = [factor * x for x in c]
    m0 = n[0] // 2 + 1
    lam = 8 * (n[0] // 2 + 1) * (n[1] // 2 + 1) * (n[2] // 2 + 1)
    f = 8 * c[0] * c[1] * c[2]
    u = 16 * m0 * c[1] * c[2]
    v = 16 * m0 * n[1] * c[2]
    return 2 * f + 4 * u + 4 * v + lam


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""

    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=subprocess.PIPE,
                stderr=subprocess.DEVNULL,
                text=True,
            )
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                smax.append(float(f[1]))
                power.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return dict(
            sm_mhz=float(np.median(sm)) if sm else None,
            sm_max_mhz=max(smax) if smax else None,
            power_w_max=max(power) if power else None,
            samples=len(sm),
            reasons=sorted(reasons),
        )


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)), "measured"
        except Exception:
            pass
    return dict(hbm_gbs=6650.0), "fallback"


# ---------------------------------------------------------------------------
# CPU legs (the only places that may execute oracle/)
# ---------------------------------------------------------------------------
def host_threads():
    """threads this process may use: its CPU affinity, NOT OMP_NUM_THREADS
    (torch.distributed.run exports OMP_NUM_THREADS=1 to every rank)"""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


def host_mem_gb():
    try:
        for ln in open("/proc/meminfo"):
            if ln.startswith("MemAvailable"):
                return float(ln.split()[1]) / 1e6
    except Exception:
        pass
    return 0.0


def cpu_config(wl, div):
    """the workload with every dimension divided by `div` (1 = full size)"""
    cfg = {k: v for k, v in WORKLOADS[wl].items() if k not in ("label", "transform")}
    cells = cfg["cells"]
    cfg["cells"] = [c // div for c in cells]
    scale = float(np.prod(cells)) / float(np.prod(cfg["cells"]))
    return cfg, scale


def time_oracle(wl, threads, reps, div, warm=0):
    """seconds per realization of the CPU port (oracle/pf_oracle.cpp) on the
    workload with cells / div per dimension: RNG stream(s) + sqrt(lambda) loop
    + c2r + restriction, eigenvalue setup timed separately"""
    from oracle import oracle_lib as ol

    cfg, scale = cpu_config(wl, div)
    o = ol.Oracle(threads=threads, **cfg)
    t0 = time.perf_counter()
    o.eigenvalues()
    t_eig = time.perf_counter() - t0
    # `threads` emulates an MPI run of the reference with that many ranks:
    # rank r draws its slab of the noise stream with seed + r
    ranks = threads
    while ranks > 1 and o.noise_entries % ranks:
        ranks -= 1
    times = []
    for r in range(warm + reps):
        t0 = time.perf_counter()
        noise = o.reference_noise_ranks(SEED + r, ranks)  # RNG stream(s)
        o.generate_from_noise(noise)  # sqrt(lambda) loop + c2r + restrict
        del noise
        times.append(time.perf_counter() - t0)
    times = times[warm:]
    return dict(mean=float(np.mean(times)), best=float(np.min(times)), steps=len(times), scale=scale, cells=cfg["cells"], t_eig=t_eig, total=float(np.sum(times)))


def time_fftlib(wl, threads, reps, div, lam_stream=None):
    """Best-effort all-core CPU variant on a real FFT library:
    numpy Generator normals drawn by `threads` workers, sigma scaling with the
    Hermitian rule of r2cfieldbackend.hh:305-325, scipy.fft.irfftn (pocketfft,
    workers = threads), restriction.  Same arithmetic as the reference's
    generateField, not its RNG stream.  Returns seconds per realization."""
    import concurrent.futures as cf

    import scipy.fft

    cfg, scale = cpu_config(wl, div)
    c = cfg["cells"]
    n = [2 * x for x in c]
    dim = len(c)
    m0 = n[0] // 2 + 1
    # sigma on the half spectrum (k_last, ..., k0): the spectrum's values do not
    # matter for the timing, its shape and the arithmetic do.  Use the analytic
    # torus covariance -> rfftn once (setup, untimed).
    t0 = time.perf_counter()
    ax = [np.minimum(np.arange(k), k - np.arange(k)) / (k / 2.0) for k in n]
    if dim == 3:
        h = np.sqrt(ax[2][:, None, None] ** 2 + ax[1][None, :, None] ** 2 + ax[0][None, None, :] ** 2)
    else:
        h = np.sqrt(ax[1][:, None] ** 2 + ax[0][None, :] ** 2)
    ell = np.ravel(cfg["corr_length"])[0]
    h /= ell
    cov = np.where(h < 1.0, 1.0 - 1.5 * h + 0.5 * h**3, 0.0) if cfg["covariance"] == "spherical" else np.exp(-h)
    del h
    lam = scipy.fft.rfftn(cov, workers=threads).real / cov.size
    del cov
    np.maximum(lam, 0.0, out=lam)
    sig = np.sqrt(lam) * np.sqrt(0.5)
    del lam
    t_eig = time.perf_counter() - t0
    shape = sig.shape
    X = np.empty(shape, dtype=np.complex128)
    Xr = X.view(np.float64).reshape(-1)
    sigr = sig.reshape(-1)
    bounds = sorted(set(int(b) for b in np.linspace(0, sigr.size, 8 * threads + 1)))
    pool = cf.ThreadPoolExecutor(threads)

    def fill(args):
        i, lo, hi, seed = args
        g = np.random.Generator(np.random.SFC64([seed, i]))
        v = Xr[2 * lo : 2 * hi]
        g.standard_normal(out=v)
        v.reshape(-1, 2)[:] *= sigr[lo:hi, None]

    times = []
    for r in range(reps):
        t0 = time.perf_counter()
        list(pool.map(fill, [(i, bounds[i], bounds[i + 1], SEED + r) for i in range(len(bounds) - 1)]))
        # self-conjugate entries: real, full weight (r2cfieldbackend.hh:318-321)
        idx = [(0, n[d] // 2) for d in range(dim)]
        for sel in np.array(np.meshgrid(*idx[::-1], indexing="ij")).reshape(dim, -1).T:
            X[tuple(sel)] = X[tuple(sel)].real * np.sqrt(2.0)
        y = scipy.fft.irfftn(X, s=tuple(n[::-1]), workers=threads, overwrite_x=True)
        y *= float(y.size)  # FFTW's backward transform is unnormalised
        field = np.ascontiguousarray(y[tuple(slice(0, k) for k in c[::-1])])
        times.append(time.perf_counter() - t0)
        del y, field
    pool.shutdown()
    return dict(mean=float(np.mean(times)), best=float(np.min(times)), steps=len(times), scale=scale, cells=c, t_eig=t_eig)


def run_reference(args, rank, world):
    """--impl reference: the CPU restatement of the reference path, every host
    thread, the FULL workload (same config as the GPU arm) when the host has
    the memory for it, rank 0 only."""
    if rank != 0:
        return
    threads = host_threads()
    cells = WORKLOADS[args.workload]["cells"]
    torus = float(np.prod([2 * c for c in cells]))
    need_gb = torus * 8 * (4.6 if len(cells) == 3 else 4.0) / 1e9  # noise + X + real + eigen scratch
    div = 1
    if host_mem_gb() < need_gb:
        div = 2
    steps = max(1, min(args.steps, 3))
    warm = 1 if args.warmup > 0 else 0
    t0 = time.perf_counter()
    res = time_oracle(args.workload, threads, steps, div, warm=warm)
    wall = time.perf_counter() - t0
    same = div == 1
    t_step = res["mean"] * res["scale"]
    value = 1.0 / t_step
    sample = "%d timed realization(s) (+%d warm-up) of %s cells%s; eigenvalue setup %.1f s excluded (one-time)" % (
        res["steps"],
        warm,
        "x".join(map(str, res["cells"])),
        " = the full workload" if same else " (1/%d of the workload's cells, host memory %.0f GB < %.0f GB needed; linear scaling)" % (int(res["scale"]), host_mem_gb(), need_gb),
        res["t_eig"],
    )
    fftlib = None
    if not args.no_fftlib:
        try:
            fr = time_fftlib(args.workload, threads, max(1, min(steps, 2)), div)
            fftlib = dict(
                value=1.0 / (fr["mean"] * fr["scale"]),
                unit="realizations/s",
                cores=threads,
                kind="port",
                library="scipy.fft (pocketfft), numpy Generator normals",
                ms_per_step=1e3 * fr["mean"] * fr["scale"],
                sample="%d realization(s) of %s cells%s; spectrum setup %.1f s excluded" % (fr["steps"], "x".join(map(str, fr["cells"])), " = the full workload" if same else " (scaled)", fr["t_eig"]),
            )
        except Exception as e:  # never lose the main figure to the extra one
            fftlib = dict(unavailable="%s: %s" % (type(e).__name__, e))
    line = dict(
        impl="reference",
        metric="realizations/sec",
        value=value,
        unit="realizations/s",
        n_gpus=args.gpus,
        steps=res["steps"],
        steps_requested=args.steps,
        warmup=warm,
        ms_per_step=1e3 * t_step,
        higher_is_better=True,
        scaling="strong" if args.gpus > 1 else "weak",
        vs_baseline=None,
        dtype="f64",
        data="synthetic",
        config=dict(workload=WORKLOADS[args.workload]["label"], name=args.workload, same_config=same, timed_wall_s=res["total"], run_wall_s=wall),
        cpu_baseline=dict(value=value, unit="realizations/s", cores=threads, kind="port", sample=sample),
        cpu_baseline_fftlib=fftlib,
        e2e=dict(value=value, unit="realizations/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0),
        note="reference arm = CPU restatement of parafields-core's generateField (oracle/pf_oracle.cpp: reference RNG streams of `cores` MPI ranks, sqrt(lambda) loop, c2r, restriction; own plain FFT, OpenMP over FFT lines); cpu_baseline_fftlib = the same arithmetic on scipy/pocketfft with all cores.  The reference's own FFT (FFTW3-MPI) is absent from this image.",
    )
    print(json.dumps(line))


# ---------------------------------------------------------------------------
# GPU legs
# ---------------------------------------------------------------------------
def timed_generate(torch, plan, stream, K, W, apply_tr, first=0, batch=None):
    """K realizations after W warm-ups, CUDA events on the plan's stream"""
    D = plan.local_domain_size
    field = torch.empty(D if batch is None else (batch, D), dtype=torch.float64, device="cuda")
    for s in range(W):
        if batch is None:
            plan.generate(SEED, first + s, out=field, apply_transform=apply_tr)
        else:
            plan.generate_batch(SEED, first + s * batch, batch, out=field, apply_transform=apply_tr)
    plan.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for s in range(K):
        if batch is None:
            plan.generate(SEED, first + W + s, out=field, apply_transform=apply_tr)
        else:
            plan.generate_batch(SEED, first + (W + s) * batch, batch, out=field, apply_transform=apply_tr)
    e1.record(stream)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (K * (batch or 1))


def other_workloads(torch, pfb, stream, peak):
    """K1, K2, K4 (single and batch 64), K5 matvec / PCG iteration: a few steps
    each, fields resident in HBM, roofline fraction against the same peak"""
    out = []

    def entry(name, what, ms, nbytes, **kw):
        d = dict(workload=name, what=what, ms=ms, per_s=1e3 / ms, algorithmic_bytes=nbytes, achieved_gbs=nbytes / ms / 1e6, frac=nbytes / ms / 1e6 / peak)
        d.update(kw)
        out.append(d)

    for name in ("K1", "K2", "K4"):
        try:
            wl = plan_kwargs(name)
            compact = wl.get("anisotropy", "none") != "geometric"
            tr = wl.get("transform", "none") != "none"
            with pfb.Plan(device=torch.cuda.current_device(), **wl) as p:
                p.set_stream(stream)
                p.compute_eigenvalues()
                ms = timed_generate(torch, p, stream, 20 if name != "K4" else 10, 3, tr)
                entry(name, "generate, one call per realization", ms, algorithmic_bytes(wl["cells"], compact=compact)["total"], label=WORKLOADS[name]["label"], note="K1 is L2-resident (12.6 MB): informational" if name == "K1" else None)
                nb = 64 if name != "K2" else 16
                msb = timed_generate(torch, p, stream, 3, 1, tr, first=1000, batch=nb if name != "K4" else 16)
                entry(name, "generate_batch (%d realizations per call; bytes of a batch of 64: sigma read once per batch)" % (nb if name != "K4" else 16), msb, algorithmic_bytes(wl["cells"], compact=compact, batch=64)["total"])
                if name == "K4":
                    # timesMatrix at K4's size (two-step complex transforms of the padded
                    # field, apply_long2d).  Algorithmic bytes, 2D analogue of matvec_bytes:
                    # field in + out, U[c1][m0] and V[n1][m0] written and read twice, lambda
                    c0, c1 = wl["cells"]
                    m0 = c0 + 1
                    nb2 = 2 * 8 * c0 * c1 + 4 * 16 * m0 * c1 + 4 * 16 * m0 * 2 * c1 + 8 * m0 * 2 * c1
                    x = torch.empty(p.domain_size, dtype=torch.float64, device="cuda")
                    y = torch.empty_like(x)
                    p.generate(1, out=x)
                    p.apply("matrix", x, out=y)
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    stream.synchronize()
                    e0.record(stream)
                    for _ in range(3):
                        p.apply("matrix", x, out=y)
                    e1.record(stream)
                    e1.synchronize()
                    entry(name, "timesMatrix (one matvec, long lines)", e0.elapsed_time(e1) / 3, nb2)
        except Exception as e:
            out.append(dict(workload=name, error="%s: %s" % (type(e).__name__, e)))
    try:
        wl = plan_kwargs("K5")
        nbytes = matvec_bytes(wl["cells"])
        with pfb.Plan(device=torch.cuda.current_device(), **wl) as p:
            p.set_stream(stream)
            p.compute_eigenvalues()
            x = torch.empty(p.domain_size, dtype=torch.float64, device="cuda")
            y = torch.empty_like(x)
            p.generate(1, 0, out=x)
            for op in ("matrix", "root", "inverse_ext"):
                for _ in range(3):
                    p.apply(op, x, out=y)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for _ in range(20):
                    p.apply(op, x, out=y)
                e1.record(stream)
                torch.cuda.synchronize()
                entry("K5", "times%s (one matvec)" % {"matrix": "Matrix", "root": "MatrixRoot", "inverse_ext": "InverseExtended"}[op], e0.elapsed_time(e1) / 20, nbytes)
        # one PCG iteration of timesInverseMatrix: difference of two pinned counts
        t = {}
        for its in (2, 12):
            with pfb.Plan(device=torch.cuda.current_device(), cg_iterations=its, **wl) as p:
                p.set_stream(stream)
                p.compute_eigenvalues()
                x = torch.empty(p.domain_size, dtype=torch.float64, device="cuda")
                y = torch.empty_like(x)
                p.generate(1, 0, out=x)
                p.solve(x, out=y)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                p.solve(x, out=y)
                torch.cuda.synchronize()
                t[its] = (time.perf_counter() - t0, p.cg_iterations_used)
        if t[12][1] > t[2][1]:
            ms_it = 1e3 * (t[12][0] - t[2][0]) / (t[12][1] - t[2][1])
            D8 = 8 * int(np.prod(wl["cells"]))
            entry("K5", "one PCG iteration of timesInverseMatrix (2 matvecs + 15 vector passes)", ms_it, 2 * nbytes + 15 * D8, label=WORKLOADS["K5"]["label"])
    except Exception as e:
        out.append(dict(workload="K5", error="%s: %s" % (type(e).__name__, e)))
    return out


def slab_parity(torch, dist, pfb, local_rank, rank, world):
    """slab-vs-single comparisons on this rank before the timed region (device
    noise, single calls and a pipelined batch): a 64^3 field (exchange route)
    and a 256^3 field (512^3 torus: the streaming route the K3 run takes);
    max over ranks of max|diff| / max|field|"""
    worst = 0.0
    for n in (64, 256):
        case = dict(cells=[n, n, n], covariance="exponential", corr_length=0.05)
        with pfb.Plan(device=local_rank, **case) as single, pfb.Plan(device=local_rank, n_ranks=world, rank=rank, **case) as slab:
            slab.connect()
            layers = n // world
            for r in range(3):
                full = single.generate(SEED, r)
                mine = slab.generate(SEED, r)
                ref = full[rank * layers : (rank + 1) * layers]
                worst = max(worst, float(np.max(np.abs(mine - ref)) / np.max(np.abs(full))))
            dev = torch.empty((3, slab.local_domain_size), dtype=torch.float64, device="cuda")
            slab.generate_batch(SEED, 10, 3, out=dev)
            slab.synchronize()
            for r in range(3):
                full = single.generate(SEED, 10 + r)
                ref = full[rank * layers : (rank + 1) * layers].ravel()
                worst = max(worst, float(np.max(np.abs(dev[r].cpu().numpy() - ref)) / np.max(np.abs(full))))
            del dev
            dist.barrier()
    t = torch.tensor([worst], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()), "3D 64^3 (exchange route) and 256^3 (streaming route) exponential: 3 single calls + a pipelined batch of 3 per rank each, device noise"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="K3", choices=sorted(WORKLOADS))
    ap.add_argument("--rng", default="fast", choices=["fast", "fp64"])
    ap.add_argument("--mode", default="auto", choices=["auto", "slab", "replicas"], help="N>1: slab = one realization sharded over the GPUs (BASELINE config, default), replicas = independent realizations per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-others", action="store_true", help="skip the other_workloads array and the rng_fp64 figure")
    ap.add_argument("--no-fftlib", action="store_true", help="reference arm: skip the scipy/pocketfft variant")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    distributed = world > 1
    if distributed:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    pfb = importlib.import_module("dune-parafields_b200")
    wl = plan_kwargs(args.workload)
    slab = distributed and args.mode in ("auto", "slab")

    parity = None
    if slab:
        err, what = slab_parity(torch, dist, pfb, local_rank, rank, world)
        parity = dict(slab_vs_single_max_rel=err, case=what, ranks=world)
        if err != 0.0:
            if rank == 0:
                print(json.dumps(dict(metric="realizations/sec", value=None, n_gpus=world, parity=parity, error="slab mode does not reproduce the single-GPU field bit for bit")))
            dist.destroy_process_group()
            sys.exit(1)

    if slab:
        plan = pfb.Plan(device=local_rank, rng_mode=args.rng, n_ranks=world, rank=rank, **wl)
    else:
        plan = pfb.Plan(device=local_rank, rng_mode=args.rng, **wl)
    stream = torch.cuda.Stream()
    plan.set_stream(stream)
    if slab:
        plan.connect()

    t0 = time.perf_counter()
    stats = plan.compute_eigenvalues()
    plan.synchronize()
    t_eig = time.perf_counter() - t0

    D = plan.local_domain_size
    # realization index of step s on this rank: slab mode = every rank works
    # on the same realization; replicas = disjoint indices
    ridx = (lambda s: s) if slab else (lambda s: rank + world * s)
    per_step = 1 if slab else world  # realizations finished per step, whole job
    apply_tr = wl.get("transform", "none") != "none"
    field = torch.empty(D, dtype=torch.float64, device="cuda")
    K, W = args.steps, args.warmup

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up -----------------------------------------------------------
    for s in range(W):
        plan.generate(SEED, ridx(s), out=field, apply_transform=apply_tr)
    plan.synchronize()

    # ---- timed region: K realizations per rank, fields resident in HBM -----
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    plan.set_profiling(True)
    l0 = plan.launch_count
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if slab:
        # one batch call: the plan pipelines pass A of realization r+1 with
        # the exchange and tail of realization r
        fields = torch.empty((K, D), dtype=torch.float64, device="cuda")
        barrier()
    ev0.record(stream)
    if slab:
        plan.generate_batch(SEED, W, K, out=fields, apply_transform=apply_tr)
    else:
        for s in range(K):
            plan.generate(SEED, ridx(W + s), out=field, apply_transform=apply_tr)
    ev1.record(stream)
    barrier()
    ms = ev0.elapsed_time(ev1)
    launches = plan.launch_count - l0
    times = plan.pass_times()
    plan.set_profiling(False)
    if distributed:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = per_step * K / (ms * 1e-3)
    # nvidia-smi samples every 20 ms: a timed region shorter than that (8 GPUs:
    # 20 steps take 13 ms) would leave no clock sample, so the same work is
    # repeated UNTIMED with the sampler still running (every rank takes the same
    # decision from the all-reduced time: the slab calls are collective)
    extra = 0
    if ms < 250.0:
        extra = int(250.0 / max(ms, 1e-3)) + 1
        for _ in range(extra):
            if slab:
                plan.generate_batch(SEED, W, K, out=fields, apply_transform=apply_tr)
            else:
                for s_ in range(K):
                    plan.generate(SEED, ridx(W + s_), out=field, apply_transform=apply_tr)
        plan.synchronize()
        barrier()
    clocks = sampler.stop() if rank == 0 else None
    if clocks is not None and extra:
        clocks["note"] = "timed region %.1f ms: sampled over it and %d untimed repeats of it" % (ms, extra)
    if slab:
        del fields

    # ---- e2e: host (pinned) output, D2H copy inside the timed region -------
    e2e = None
    if not args.no_e2e:
        Ke = max(2, min(K, 8))
        host = torch.empty((Ke, D), dtype=torch.float64).pin_memory()
        plan.generate_batch(SEED, 0, 2, out=host[:2], apply_transform=apply_tr)  # warm-up
        barrier()
        t0 = time.perf_counter()
        plan.generate_batch(SEED, 0 if slab else rank * Ke, Ke, out=host, apply_transform=apply_tr)
        torch.cuda.synchronize()
        te = time.perf_counter() - t0
        if distributed:
            t = torch.tensor([te], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            te = float(t.item())
        e2e = dict(
            value=per_step * Ke / te,
            unit="realizations/s",
            h2d_bytes_per_step=16,
            d2h_bytes_per_step=8 * D,
            d2h_gbs_per_gpu=8 * D * Ke / te / 1e9,
            steps=Ke,
            api="pf_generate_batch with a pinned host output buffer",
        )
        del host

    # ---- the same step with fp64 normals (N = 1) ----------------------------
    rng_other = None
    if world == 1 and not args.no_others:
        other = "fp64" if args.rng == "fast" else "fast"
        plan.close()
        plan = pfb.Plan(device=local_rank, rng_mode=other, **wl)
        plan.set_stream(stream)
        plan.compute_eigenvalues()
        ms_o = timed_generate(torch, plan, stream, max(3, min(K, 10)), 3, apply_tr)
        rng_other = dict(rng=other, value=1e3 / ms_o, unit="realizations/s", ms_per_step=ms_o, note="10 Philox rounds + fp64 Box-Muller (log, sqrt, sincospi)" if other == "fp64" else "7 Philox rounds + fp32 MUFU Box-Muller")

    if rank == 0:
        peaks, peak_kind = measured_peaks()
        peak = float(peaks["hbm_gbs"])
        ab = algorithmic_bytes(wl["cells"], compact=wl.get("anisotropy", "none") != "geometric")
        # dominant kernel = the pass with the largest share of the step
        # 12: the fused kernel of the streaming route (first two axes in one
        # persistent kernel, pf_stream.cuh): its algorithmic bytes are those of
        # pass A + pass B
        names = {0: "pass_a", 2: "pass_b", 4: "pass_c", 1: "pass_a", 12: "pass_ab"}
        ab["pass_ab"] = ab["pass_a"] + ab["pass_b"]
        shares = {names[k]: v for k, v in times.items() if k in names}
        dom = max(shares, key=lambda k: shares[k][0]) if shares else "pass_a"
        dom_ms = shares[dom][0] / max(shares[dom][1], 1) if shares else float("nan")
        share = world if slab else 1  # slab mode: every rank moves 1/P of the bytes
        if ab.get(dom, 0) == 0:
            # long-line route (K4): the passes are split differently; judge the
            # whole step against the algorithmic bytes of one realization
            dom, dom_ms = "step", ms / K
            ab = dict(ab, step=ab["total"])
        achieved = ab[dom] / share / (dom_ms * 1e-3) / 1e9
        traffic = None
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            if world == 1:
                traffic = tr.get(args.workload, {}).get(dom)
        except Exception:
            pass
        step_s = ms / K * 1e-3
        hbm = dict(
            bound="hbm",
            kernel=dom,
            achieved=achieved,
            peak=peak,
            unit="GB/s",
            frac=achieved / peak,
            traffic=traffic,
            traffic_source="profiles/traffic.json (ncu --set full capture of the same workload)" if traffic else None,
            peak_source=peak_kind + " copy bandwidth (MEASURED_PEAKS.json hbm_gbs)",
            algorithmic_bytes_per_launch=ab[dom] // share,
            ms_per_launch=dom_ms,
            step=dict(
                algorithmic_bytes=ab["total"],
                achieved=ab["total"] / step_s / 1e9,
                frac=ab["total"] / step_s / 1e9 / (peak * world),
                note="whole job: bytes of one realization / step time, against N x the measured HBM peak",
            ),
            passes_ms={k: v[0] / max(v[1], 1) for k, v in shares.items()},
        )
        roofline = hbm
        if slab:
            lay = pfb.slab_layout(wl["cells"], 2, world, 0)
            c = wl["cells"]
            m0 = c[0] + 1
            row = 16 * lay["a_count"] * (m0 if len(c) == 3 else 1)  # bytes per layer written by rank 0
            sent = row * c[-1] * (world - 1) // world  # layers owned by other ranks
            roofline = dict(
                bound="nvlink",
                kernel="pass_a (tile rows stored to the layer owners) / exchange",
                achieved=sent / step_s / 1e9,
                peak=770.0,
                unit="GB/s",
                frac=sent / step_s / 1e9 / 770.0,
                traffic=None,
                sent_bytes_per_gpu_per_realization=sent,
                peak_source="measured peer copy per direction per GPU (B200_PROFILING.md)",
                note="slab mode is NVLink-bound at every N (SURVEY 8(d)): roofline time = bytes each GPU must send / 770 GB/s, overlapped with the local passes",
                hbm=hbm,
            )
            st = pfb.stream_slab_layout(wl["cells"], 2, world, 0)
            if st and os.environ.get("PF_STREAM", "1") != "0":
                roofline["route"] = "streaming route: every rank takes its planes of k0 through both first axes (kp_ab) and stores the twice-pruned lines to the layer owners; no exchange pass"
                roofline["actual_sent_bytes_per_gpu_per_realization"] = st["sent_bytes"]
                roofline["actual_nvlink_gbs"] = st["sent_bytes"] / step_s / 1e9
                roofline["note"] += "; the streaming route moves HALF of SURVEY's bytes (the exchange happens after both prunings), so `frac` -- kept on SURVEY's byte count -- can exceed what the links would allow the transposed transform"
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            r = time_oracle(args.workload, 1, 1, 2 if len(wl["cells"]) == 3 else 1)
            cpu = dict(
                value=1.0 / (r["mean"] * r["scale"]),
                unit="realizations/s",
                cores=1,
                kind="port",
                sample="one realization of %s cells (1/%d of the workload's cells, %.1f s), linear scaling; single thread = one MPI rank of the reference; the all-core figures at full size are in the --impl reference line"
                % ("x".join(map(str, r["cells"])), int(r["scale"]), r["mean"]),
            )
        others = None
        if world == 1 and not args.no_others:
            plan.close()
            others = other_workloads(torch, pfb, stream, peak)
        line = dict(
            metric="realizations/sec",
            value=value,
            unit="realizations/s",
            n_gpus=world,
            steps=K,
            warmup=W,
            ms_per_step=ms / K,
            higher_is_better=True,
            scaling="strong" if slab else "weak",
            vs_baseline=None,
            dtype="f64",
            data="synthetic",
            config=dict(
                workload=WORKLOADS[args.workload]["label"],
                name=args.workload,
                parallelism=(
                    "slab x%d: one realization sharded over the GPUs, pass A scatters tile rows to the layer owners by peer stores over NVLink (CUDA IPC; no NCCL on the data path)" % world
                    if slab
                    else "replicas x%d (independent realizations, no collective)" % world
                ),
                l2="per-step working set (>9 GB) exceeds the 126 MB L2",
                rng=args.rng,
                eigenvalue_setup_s=t_eig,
                eigenvalue_census=stats,
            ),
            clocks=clocks,
            e2e=e2e,
            gpu_launches=launches,
            roofline=roofline,
            cpu_baseline=cpu,
        )
        if rng_other:
            line["rng_" + rng_other["rng"]] = rng_other
        if parity:
            line["parity"] = parity
        if others is not None:
            line["other_workloads"] = others
        print(json.dumps(line))
    plan.close()
    if distributed:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
