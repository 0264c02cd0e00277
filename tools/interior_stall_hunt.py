"""Which call inside optimize() stalls when an iteration of the interior loop takes > 0.2 s?  (sporadic)"""
import sys, time
sys.path.insert(0, ".")
import torch
from sdmm_mitsuba_b200 import scenes, mixture
from sdmm_mitsuba_b200.integrator import SDMMIntegrator
log = []
def wrap(obj, name):
    f = getattr(obj, name)
    def g(*a, **k):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        r = f(*a, **k)
        torch.cuda.synchronize(); log.append((name, time.perf_counter() - t0))
        return r
    setattr(obj, name, g)
for n in ("set_samples", "em_step", "initialize", "copy_cells", "set_tree", "find_cells"):
    wrap(mixture.SDMMCells, n)
for n in ("sort", "empty", "as_tensor", "nonzero", "bincount"):
    wrap(torch, n)
quads, cam = scenes.interior(640, 360)
for rep in range(8):
    g = SDMMIntegrator(quads, cam, K=16, max_depth=6, samples_per_iteration=8, seed=5, pool_capacity=96_000_000)
    spp, rendered = 8, 0
    while rendered < 504:
        st, new = g.render_pass(spp, collect=True)
        torch.cuda.synchronize(); del log[:]; t1 = time.perf_counter()
        info = g.optimize(new)
        torch.cuda.synchronize(); dt = time.perf_counter() - t1
        if dt > 0.2:
            print("rep %d spp %d optimize %.3f s:" % (rep, spp, dt), [(n, round(t, 3)) for n, t in log if t > 0.01])
        g.iteration += 1; g.total_spp += spp; rendered += spp; spp *= 2
    g.close()
print("done")
