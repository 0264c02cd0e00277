"""Shared parity metrics (tolerances from BASELINE.json north_star: relative 1e-4 on per-sample
log-likelihood, 1e-3 on mixture parameters)."""
import numpy as np

LL_RTOL = 1e-4
PARAM_RTOL = 1e-3


def loglik_err(a, b):
    """max |a-b| / max(1, |b|) over finite entries; also checks the non-finite patterns agree."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    fin = np.isfinite(a) & np.isfinite(b)
    mismatch = np.mean(np.isfinite(a) != np.isfinite(b))
    err = np.abs(a[fin] - b[fin]) / np.maximum(1.0, np.abs(b[fin]))
    return float(err.max()) if err.size else 0.0, float(mismatch)


def joint_param_errs(j_a, j_b):
    """j_*: [C,32,K] joint planes.  Returns dict of max errors:
    pmf  : |d pmf| / max_k pmf           (per cell)
    pos  : |d mean_xyz|                  (cells are unit cubes)
    ang  : angle between direction means (radians)
    cov  : max_ij |d cov_ij| / max_ij |cov_ij|  per component
    Components whose reference weight is 0 are ignored (their parameters are frozen)."""
    a = np.asarray(j_a, np.float64)
    b = np.asarray(j_b, np.float64)
    live = b[:, 0, :] > 0
    pmf = np.abs(a[:, 0] - b[:, 0]) / b[:, 0].max(axis=1, keepdims=True)
    pos = np.abs(a[:, 1:4] - b[:, 1:4]).max(axis=1)
    da = a[:, 4:7] / np.linalg.norm(a[:, 4:7], axis=1, keepdims=True)
    db = b[:, 4:7] / np.linalg.norm(b[:, 4:7], axis=1, keepdims=True)
    cr = np.linalg.norm(np.cross(da, db, axis=1), axis=1)
    ang = np.arctan2(cr, (da * db).sum(axis=1))
    cov = np.abs(a[:, 7:32] - b[:, 7:32]).max(axis=1) / np.abs(b[:, 7:32]).max(axis=1)
    out = {}
    for k, v in dict(pmf=pmf, pos=pos, ang=ang, cov=cov).items():
        out[k] = float(np.where(live, v, 0.0).max())
    return out


def assert_joint_close(j_a, j_b, tol=PARAM_RTOL, what=""):
    e = joint_param_errs(j_a, j_b)
    bad = {k: v for k, v in e.items() if not v <= tol}
    assert not bad, "%s parameter mismatch beyond %g: %s (all: %s)" % (what, tol, bad, e)
    return e
