"""CPU checks of the product-guiding checker itself (oracle/_ref/libsdmm_ref_product.so, the reference's own
sdmm::product driven per vertex): zero-weight BSDF components are skipped, the product is normalised, and its
samples follow its pdf.  The GPU parity tests proper are in test_gpu_product.py."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def fitted_ref():
    from oracle import portbind, refbind
    from sdmm_mitsuba_b200 import synth
    if not (refbind.available() and refbind.product_available()):
        pytest.skip("oracle/_ref not built")
    C, K = 2, 16
    point6, weight, off = synth.make_samples(C, 1024, seed=0)
    joint, em, scal = synth.make_initial_model(C, K, seed=1)
    o = portbind.PortCells(C, K)
    o.set(joint, em, scal)
    o.em_step(off, point6, weight, 3)
    r = refbind.RefCells(C, K)
    r.set(*o.get())
    r.prepare_conditioner()
    return r, C, K


BSDF = np.asarray([[0.6, 0, 0, 1, 0.45, 0.02, 0.5], [0.25, 0.3, -0.2, 0.9, 0.08, -0.01, 0.05],
                   [0.0, 1, 0, 0, 0.1, 0, 0.1], [0.15, -0.5, 0.4, 0.7, 0.03, 0.005, 0.06]], np.float32)


def test_product_layout_and_normalisation(fitted_ref):
    r, C, K = fitted_ref
    x = np.array([0.4, 0.5, 0.6], np.float32)
    comps = r.product_components(0, x, BSDF)
    assert comps.shape == (K * 3, 7)               # the pmf-0 BSDF component is skipped (sdmm_product.h:83-88)
    assert abs(comps[:, 0].sum() - 1) < 1e-5 and abs(comps[-1, 6] - 1) < 1e-5
    np.testing.assert_allclose(np.cumsum(comps[:, 0].astype(np.float64)), comps[:, 6], atol=2e-6)
    assert (comps[:, 3] > 0).all() and (comps[:, 3] * comps[:, 5] - comps[:, 4] ** 2 > 0).all()
    # dropping the zero-weight row changes nothing
    comps2 = r.product_components(0, x, BSDF[[0, 1, 3]])
    np.testing.assert_array_equal(comps, comps2)


def test_product_samples_follow_product_pdf(fitted_ref):
    r, C, K = fitted_ref
    n = 60000
    rng = np.random.default_rng(2)
    cell = np.zeros(n, np.uint32)
    pos3 = np.ascontiguousarray(np.repeat(np.array([[0.45], [0.5], [0.55]], np.float32), n, axis=1))
    u3 = rng.random(size=(3, n), dtype=np.float32)
    g, _ = r.product_sample_pdf(cell, pos3, BSDF, u3=u3)
    assert g["valid"].all()
    ok = g["inv_jac"] > 0
    d = rng.normal(size=(3, n)).astype(np.float32)
    d /= np.linalg.norm(d, axis=0, keepdims=True)
    d = np.ascontiguousarray(d)
    q, _ = r.product_sample_pdf(cell, pos3, BSDF, dir3=d)
    edges = np.linspace(-1, 1, 7)
    freq = np.histogram(g["dir"][2, ok], bins=edges)[0] / n
    mass = np.asarray([4 * np.pi * np.mean(q["pdf"].astype(np.float64) * ((d[2] >= lo) & (d[2] < hi)))
                       for lo, hi in zip(edges[:-1], edges[1:])])
    np.testing.assert_allclose(freq, mass, atol=0.015, rtol=0.1)
