"""GPU parity against golden vectors produced by the reference's own code (tests/golden/*.npz,
tools/make_golden.py): the check that does not depend on any oracle library being present."""
import numpy as np
import pytest

import golden_util as G

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", G.CASES)
def test_gpu_em_matches_golden(name):
    from sdmm_mitsuba_b200.mixture import SDMMCells
    g = G.load(name)
    cells = SDMMCells(g["C"], g["K"])
    cells.upload(g["joint0"], g["em0"], g["scal0"])
    cells.set_samples(g["point6"], g["weight"], g["offsets"])
    for it in range(g["iters"]):
        tag = "%d" % it
        cells.upload(g["joint" + tag], g["em" + tag], g["scal" + tag])
        ll = cells.em_step(1, with_conditioner=False, want_loglik=True).cpu().numpy()[0]
        j, e, s = cells.download()
        G.check_em_iteration(g, it, ll, j, e, s, "gpu " + name)
        assert (s[:, 7] == 0).all()


@pytest.mark.parametrize("name", G.CASES)
def test_gpu_free_running_matches_golden(name):
    from sdmm_mitsuba_b200.mixture import SDMMCells
    g = G.load(name)
    cells = SDMMCells(g["C"], g["K"])
    cells.upload(g["joint0"], g["em0"], g["scal0"])
    cells.set_samples(g["point6"], g["weight"], g["offsets"])
    cells.em_step(g["iters"], with_conditioner=True)
    j, e, s = cells.download()
    G.assert_joint_close(j, g["joint%d" % g["iters"]], G.PARAM_RTOL, "gpu free-running " + name)
    sf = g["scal%d" % g["iters"]]
    assert (s[:, 1] == sf[:, 1]).all() and (s[:, 2] == sf[:, 2]).all()


@pytest.mark.parametrize("name", G.CASES)
@pytest.mark.parametrize("binned", [True, False])
def test_gpu_queries_match_golden(name, binned):
    from sdmm_mitsuba_b200.mixture import SDMMCells
    g = G.load(name)
    it = "%d" % g["iters"]
    cells = SDMMCells(g["C"], g["K"])
    cells.upload(g["joint" + it], g["em" + it], g["scal" + it])
    cells.prepare()
    q = cells.sample_pdf(g["q_cell"], g["q_pos3"], g["q_u3"], binned=binned)
    G.check_queries(g, q, "gpu " + name)
    G.check_query_pdf_cause(g, q, "gpu " + name)
    p, v = cells.pdf(g["q_cell"], g["q_pos3"], g["p_dir"], binned=binned)
    G.check_pdf(g, p, v, "gpu " + name)
