"""Diagnostic: per-iteration log-likelihood error of the GPU path vs the oracles, and at iteration
0 (identical model on all sides) vs a float64 evaluation of the same formulas."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import portbind, refbind
from sdmm_mitsuba_b200 import synth
from sdmm_mitsuba_b200.mixture import SDMMCells


def exact_ll(joint, point6):
    """float64 log sum_k pi_k p_k(x) for one cell (joint [32,K]) with the reference's formulas."""
    K = joint.shape[1]
    j = joint.astype(np.float64)
    x = point6.astype(np.float64)
    N = x.shape[1]
    out = np.zeros((K, N))
    pmf = j[0] / j[0].sum()
    for k in range(K):
        mu = j[1:7, k]
        n = mu[3:6] / np.linalg.norm(mu[3:6])
        sign = np.copysign(1.0, n[2]); a = -1.0 / (sign + n[2]); b = n[0] * n[1] * a
        s = np.array([1 + sign * n[0] * n[0] * a, sign * b, -sign * n[0]])
        t = np.array([b, sign + n[1] * n[1] * a, -n[1]])
        to = np.stack([s, t, n])
        cov = j[7:32, k].reshape(5, 5)
        L = np.linalg.cholesky(cov + 0 * np.eye(5))
        # reference adds 1e-5 to each pivot: emulate in f64
        L = np.zeros((5, 5))
        for r in range(5):
            for c in range(r):
                L[r, c] = (cov[r, c] - L[r, :c] @ L[c, :c]) / L[c, c]
            L[r, r] = np.sqrt(cov[r, r] - L[r, :r] @ L[r, :r] + 1e-5)
        inv_det = 1.0 / np.prod(np.diag(L))
        loc = to @ x[3:6]
        c = np.clip(loc[2], -1, 1)
        ang = np.arccos(c); sn = np.sqrt(np.maximum(1 - c * c, 0))
        r = np.where(sn < 1e-4, 1.0, ang / np.maximum(sn, 1e-300))
        tang = np.concatenate([x[0:3] - mu[0:3, None], loc[0:2] * r])
        z = np.linalg.solve(L, tang)
        out[k] = pmf[k] * inv_det * r * (2 * np.pi) ** -2.5 * np.exp(-0.5 * (z * z).sum(0))
    return np.log(out.sum(0))


C, K, N, iters = 4, int(sys.argv[1]) if len(sys.argv) > 1 else 16, 777, 4
point6, weight, off = synth.make_samples(C, N, seed=K)
joint, em, scal = synth.make_initial_model(C, K, seed=K + 1)
cells = SDMMCells(C, K); cells.upload(joint, em, scal); cells.set_samples(point6, weight, off)
ll_gpu = cells.em_step(iters, want_loglik=True).cpu().numpy()
res = {}
for name, cls in [("port", portbind.PortCells)] + ([("ref", refbind.RefCells)] if refbind.available() else []):
    o = cls(C, K); o.set(joint, em, scal)
    res[name] = np.concatenate(o.em_step(off, point6, weight, iters, want_loglik=True), axis=1)
ex = np.concatenate([exact_ll(joint[c], point6[:, off[c]:off[c + 1]]) for c in range(C)])
rel = lambda a, b: np.abs(a - b) / np.maximum(1, np.abs(b))
print("iteration 0 vs float64: gpu %.3g" % rel(ll_gpu[0], ex).max(), {n: "%.3g" % rel(v[0], ex).max() for n, v in res.items()})
for it in range(iters):
    for n, v in res.items():
        e = rel(ll_gpu[it], v[it]); i = int(e.argmax())
        print("it %d vs %s: max %.3g (ll=%.4f gpu=%.4f) median %.3g  p99 %.3g" % (it, n, e.max(), v[it][i], ll_gpu[it][i], np.median(e), np.quantile(e, 0.99)))
if "ref" in res:
    for it in range(iters):
        e = rel(res["port"][it], res["ref"][it])
        print("it %d port vs ref: max %.3g median %.3g" % (it, e.max(), np.median(e)))
