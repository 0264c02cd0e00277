"""Pins the plain-C oracle restatement (oracle/sdmm_oracle.c) against golden vectors produced by the
reference's own code (tests/golden/*.npz, tools/make_golden.py).  CPU only; this is the check that
still holds on the GPU box where /root/reference does not exist."""
import numpy as np
import pytest

import golden_util as G


@pytest.mark.parametrize("name", G.CASES)
def test_port_em_matches_golden(oracle_port, name):
    g = G.load(name)
    o = oracle_port.PortCells(g["C"], g["K"])
    d0 = None
    for it in range(g["iters"]):
        tag = "%d" % it
        o.set(g["joint" + tag], g["em" + tag], g["scal" + tag])
        if it == 0:
            d0 = o.derived()
        ll = np.concatenate(o.em_step(g["offsets"], g["point6"], g["weight"], 1, want_loglik=True), axis=1)[0]
        j, e, s = o.get()
        G.check_em_iteration(g, it, ll, j, e, s, "port " + name)
    # prepare_vectorized of the initial model: frames, Cholesky, inv_det, cdf, pmf, psd
    r = g["derived0"]
    np.testing.assert_allclose(d0[:, 0:9], r[:, 0:9], atol=2e-6)
    np.testing.assert_allclose(d0[:, 9:34], r[:, 9:34], rtol=2e-5, atol=1e-6)
    np.testing.assert_allclose(d0[:, 34], r[:, 34], rtol=5e-5)
    np.testing.assert_allclose(d0[:, 35:37], r[:, 35:37], rtol=1e-6, atol=1e-7)
    np.testing.assert_array_equal(d0[:, 37], r[:, 37])


@pytest.mark.parametrize("name", G.CASES)
def test_port_free_running_matches_golden(oracle_port, name):
    """All iterations without re-synchronising: final parameters within 1e-3."""
    g = G.load(name)
    o = oracle_port.PortCells(g["C"], g["K"])
    o.set(g["joint0"], g["em0"], g["scal0"])
    o.em_step(g["offsets"], g["point6"], g["weight"], g["iters"])
    j, e, s = o.get()
    G.assert_joint_close(j, g["joint%d" % g["iters"]], G.PARAM_RTOL, "port free-running " + name)
    sf = g["scal%d" % g["iters"]]
    assert (s[:, 1] == sf[:, 1]).all() and (s[:, 2] == sf[:, 2]).all()


@pytest.mark.parametrize("name", G.CASES)
def test_port_queries_match_golden(oracle_port, name):
    g = G.load(name)
    it = "%d" % g["iters"]
    o = oracle_port.PortCells(g["C"], g["K"])
    o.set(g["joint" + it], g["em" + it], g["scal" + it])
    o.prepare_conditioner()
    c, r = o.conditioner(), g["conditioner"]
    np.testing.assert_allclose(c[:, 0:3], r[:, 0:3], atol=1e-6)
    np.testing.assert_allclose(c[:, 3:14], r[:, 3:14], rtol=1e-4, atol=1e-6)
    for sl in (slice(14, 20), slice(29, 33), slice(34, 38)):
        scale = np.abs(r[:, sl]).max(axis=1, keepdims=True)
        assert (np.abs(c[:, sl] - r[:, sl]) / scale).max() < 1e-3
    np.testing.assert_allclose(c[:, 20:29], r[:, 20:29], atol=2e-6)
    q, _ = o.sample_pdf(g["q_cell"], g["q_pos3"], g["q_u3"])
    G.check_queries(g, q, "port " + name)
    p, v, _ = o.pdf(g["q_cell"], g["q_pos3"], g["p_dir"])
    G.check_pdf(g, p, v, "port " + name)


def test_golden_fixtures_are_sane():
    """The fixtures themselves: finite, skipped-sample bookkeeping visible, likelihood improves."""
    for name in G.CASES:
        g = G.load(name)
        assert np.isfinite(g["loglik0"]).all()
        # two invalid-weight samples were planted in cell 0: the loop skips them but samples_seen
        # still counts them (sdmm/opt/em.h:264 adds data.size; only a failed normaliser decrements)
        assert int(g["scal1"][0, 2]) == g["N"]
        assert int(g["scal%d" % g["iters"]][0, 1]) == g["iters"]
        w = np.where(g["weight"] >= 1e-8, g["weight"], 0).astype(np.float64)
        first = (g["loglik0"] * w).sum() / w.sum()
        last = (g["loglik%d" % (g["iters"] - 1)] * w).sum() / w.sum()
        assert last > first
