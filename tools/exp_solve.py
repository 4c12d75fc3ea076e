# experiment: where does the time of pf_solve go at K5 size with device operands?
import ctypes as C, importlib, sys, time
sys.path.insert(0, ".")
import torch
m = importlib.import_module("dune-parafields_b200")
L = m.lib()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
p = m.Plan([n, n, n], covariance="exponential", corr_length=0.05, cg_iterations=10, rng_mode="fast")
D = p.domain_size
x = torch.empty(D, dtype=torch.float64, device="cuda")
p.generate(42, out=x)
y = torch.empty_like(x); z = torch.empty_like(x)
def t(f, reps=3):
    f(); p.synchronize(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): f()
    p.synchronize(); torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3
print("apply matrix ms", t(lambda: p.apply("matrix", x, out=y)))
print("solve(M x) ms", t(lambda: p.solve(y, out=z)), "iterations", p.cg_iterations_used)
print("solve(x) ms", t(lambda: p.solve(x, out=z)), "iterations", p.cg_iterations_used)
def alloc_cycle():
    ptr = C.c_void_p()
    L.pf_device_alloc(p._h, D * 8, C.byref(ptr))
    L.pf_solve(p._h, y.data_ptr(), ptr, None)
    L.pf_memcpy(p._h, z.data_ptr(), ptr, D * 8)
    L.pf_device_free(p._h, ptr)
print("alloc + solve(M x) + copy + free ms", t(alloc_cycle))
def alloc_only():
    ptr = C.c_void_p()
    L.pf_device_alloc(p._h, D * 8, C.byref(ptr))
    L.pf_device_free(p._h, ptr)
print("alloc + free ms", t(alloc_only))
