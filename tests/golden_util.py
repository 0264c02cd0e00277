"""Loader + comparison helpers for tests/golden/*.npz (made by tools/make_golden.py from the compiled
reference).  Used by the CPU test that pins the plain-C oracle and the GPU test that pins the CUDA path."""
import os

import numpy as np

from parity import LL_RTOL, PARAM_RTOL, assert_joint_close, loglik_err

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["em_k16", "em_k64", "em_k128", "em_k64_n4096", "em_k128_n8192"]
TIE_DELTA = 2e-5


def load(name):
    g = dict(np.load(os.path.join(GOLDEN_DIR, name + ".npz")))
    C, K, N, iters, nq = [int(v) for v in g["meta"]]
    g.update(C=C, K=K, N=N, iters=iters, nq=nq)
    return g


def check_em_iteration(g, it, ll, joint, em, scal, what):
    """State after iteration `it` (0-based) started from the golden state entering it."""
    err, mism = loglik_err(ll, g["loglik%d" % it])
    assert mism == 0.0 and err <= LL_RTOL, "%s it %d: log-likelihood error %.3g" % (what, it, err)
    j1, e1, s1 = g["joint%d" % (it + 1)], g["em%d" % (it + 1)], g["scal%d" % (it + 1)]
    assert_joint_close(joint, j1, PARAM_RTOL, "%s it %d" % (what, it))
    w_err = np.abs(em[:, 0] - e1[:, 0]).max() / np.abs(e1[:, 0]).max()
    c_err = np.abs(em[:, 6:31] - e1[:, 6:31]).max() / np.abs(e1[:, 6:31]).max()
    assert w_err <= PARAM_RTOL and c_err <= PARAM_RTOL, (what, it, w_err, c_err)
    assert (scal[:, 1] == s1[:, 1]).all() and (scal[:, 2] == s1[:, 2]).all(), (what, it, scal[:, 1:3], s1[:, 1:3])
    np.testing.assert_allclose(scal[:, 0].view(np.float32), s1[:, 0].view(np.float32), rtol=1e-5)


def check_queries(g, r, what):
    """r: dict(dir, pdf, inv_jac, comp, valid) for the golden queries."""
    np.testing.assert_array_equal(np.asarray(r["valid"]), g["q_valid"], err_msg=what)
    v = g["q_valid"].astype(bool)
    diff = v & (np.asarray(r["comp"]) != g["q_comp"])
    # component selection is bit-exact away from CDF ties
    assert (g["q_tie_margin"][diff] < TIE_DELTA).all(), (what, np.nonzero(diff)[0][:10], g["q_tie_margin"][diff][:10])
    same = v & ~diff
    assert same.sum() >= 0.999 * v.sum()
    ang = np.arccos(np.clip((r["dir"][:, same] * g["q_dir"][:, same]).sum(axis=0), -1, 1))
    assert ang.max() < 2e-3, (what, ang.max())
    np.testing.assert_allclose(r["inv_jac"][same], g["q_inv_jac"][same], rtol=1e-4, atol=1e-6, err_msg=what)
    big = same & (g["q_pdf"] > 1e-6)
    rel = np.abs(r["pdf"][big] - g["q_pdf"][big]) / g["q_pdf"][big]
    assert np.median(rel) < 1e-4 and np.quantile(rel, 0.999) < 5e-3, (what, np.quantile(rel, [0.5, 0.999, 1]))


def check_pdf(g, p, v, what, rtol=1e-3):
    np.testing.assert_array_equal(np.asarray(v), g["p_valid"], err_msg=what)
    big = g["p_pdf"] > 1e-6
    rel = np.abs(p[big] - g["p_pdf"][big]) / g["p_pdf"][big]
    assert rel.max() < rtol, (what, rel.max())
