"""Device progressive loop vs the CPU restatement built on the reference's own headers (oracle/ref_render.inc),
same scene, same per-path random streams.  Prints the comparisons tests/test_gpu_render.py asserts."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import refbind
from sdmm_mitsuba_b200 import scenes
from sdmm_mitsuba_b200.integrator import SDMMIntegrator
from parity import joint_param_errs

W, H = int(os.environ.get("W", 160)), int(os.environ.get("H", 90))
K, MD, SPI, NIT, NTRAIN = 16, 5, int(os.environ.get("SPI", 32)), int(os.environ.get("NIT", 6)), int(os.environ.get("NTRAIN", 3))
quads, cam = scenes.cornell_box(W, H)
lo, hi, norm = scenes.scene_bounds(quads)
rmin, rmax = scenes.root_aabb(lo, hi, norm)
g = SDMMIntegrator(quads, cam, K=K, max_depth=MD, samples_per_iteration=SPI, seed=7, pool_capacity=4_000_000)
r = refbind.RefRenderer(quads, cam, lo, norm, rmin, rmax, K=K, max_depth=MD, capacity=150000)
def compare_state(tag):
    nr = r.nodes(); t = g.tree
    same = t.size() == nr.shape[0] and all(bool(nr[i, 0] > 0) == t.is_leaf[i] for i in range(t.size()))
    print("   [%s] nodes gpu/ref %d %d  same topology: %s" % (tag, t.size(), nr.shape[0], same))
    if not same:
        return
    j_gpu, _, s_gpu = g.cells.download(want_em=False)
    init_g = g.initialized.cpu().numpy()
    errs = []
    for node in t.leaves():
        jr, sr = r.leaf_joint(node)
        if jr is None: continue
        e = joint_param_errs(j_gpu[t.slot[node]][None], jr[None])
        errs.append((node, e, int(s_gpu[t.slot[node], 1]), int(sr[1]), int(s_gpu[t.slot[node], 2]), int(sr[2]), int(s_gpu[t.slot[node], 7])))
    init_r = nr[:, 12]
    print("   leaves with mixtures: %d; initialized flags equal: %s" % (len(errs), all(bool(init_r[n_]) == bool(init_g[t.slot[n_]]) for n_ in t.leaves())))
    if errs:
        for k in ("pmf", "pos", "ang", "cov"):
            v = np.array([e[1][k] for e in errs]); print("     %s: median %.2e  90%% %.2e  max %.2e (node %d)" % (k, np.median(v), np.quantile(v, 0.9), v.max(), errs[int(v.argmax())][0]))
        print("     samples_seen gpu-ref where different:", [(e[0], e[4] - e[5], e[4]) for e in errs if e[4] != e[5]][:12])
        print("     iterations_run equal:", all(e[2] == e[3] for e in errs), " samples_seen equal:", sum(e[4] == e[5] for e in errs), "/", len(errs), " status bits:", sorted(set(e[6] for e in errs)))

tg = tr = 0.0
for it in range(NIT):
    train = it < NTRAIN
    seed = (g.seed * 0x9E3779B97F4A7C15 + g.iteration) & 0xFFFFFFFFFFFFFFFF
    t0 = time.time(); st, new = g.render_pass(SPI, collect=train); info = g.optimize(new) if train else {}
    g.iteration += 1; g.total_spp += SPI; tg += time.time() - t0
    t0 = time.time(); rs = r.iteration(SPI, seed, train); tr += time.time() - t0
    print("it", it, "gpu", dict(guided=int(st.guided_queries), mis=int(st.mis_queries), **info))
    print("      ref", rs)
    ig, ir = g.image(), r.image()
    compare_state("it %d" % it)
    print("      mean radiance gpu %.5f ref %.5f   mean|diff| %.5f" % (ig.mean(), ir.mean(), np.abs(ig - ir).mean()))
print("time gpu %.2f s   cpu restatement %.2f s" % (tg, tr))
