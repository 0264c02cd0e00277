"""GPU parity of product guiding: conditional x learned-BSDF mixture, sample + pdf per
vertex, against the reference's own sdmm::product / SDMM::sample / posterior driven through oracle/_ref
(sdmm_product.h:62-203; sdmm_proc.cpp:327-339, 375-381).  Component selection must be identical for identical
uniforms away from CDF ties; pdf within PDF_RTOL."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

# Tolerances.  A product weight carries exp(-sq/2) with sq = |L^-1 (BSDF mean - component mean)|^2 up to ~30 for
# components that still matter, so float32 rounding of the two Cholesky factors is amplified ~30x: evaluating the
# same formulas in float64 from either side's conditioner planes (tools/product_diag2.py) agrees to 1e-6 at typical
# vertices, while at the worst vertices the REFERENCE's own float32 result is 5e-3 off that float64 value and the
# GPU's 6e-3 off, in opposite directions.  Hence: 99% of the vertices within PDF_RTOL, all within PDF_RTOL_MAX,
# and a component choice may differ only where u0 is within TIE_DELTA of a reference CDF step.
PDF_RTOL = 2e-3
PDF_RTOL_MAX = 5e-2
TIE_DELTA = 1e-3


def _bsdf(M_zero=True):
    """A learned-BSDF style directional mixture in world space: a wide lobe (re-centred on the normal in the
    diffuse case), two narrower ones and, optionally, a zero-weight component the product must skip."""
    rows = [[0.6, 0.0, 0.0, 1.0, 0.45, 0.02, 0.5],
            [0.25, 0.3, -0.2, 0.9, 0.08, -0.01, 0.05]]
    if M_zero:
        rows.append([0.0, 1.0, 0.0, 0.0, 0.1, 0.0, 0.1])
    rows.append([0.15, -0.5, 0.4, 0.7, 0.03, 0.005, 0.06])
    return np.asarray(rows, np.float32)


def _setup(C, K, n, seed):
    from oracle import refbind
    from sdmm_mitsuba_b200 import synth
    from sdmm_mitsuba_b200.mixture import SDMMCells
    from test_gpu_query import _fitted
    if not (refbind.available() and refbind.product_available()):
        pytest.skip("oracle/_ref not built")
    joint, em, scal = _fitted(C, K, seed=seed)
    cells = SDMMCells(C, K)
    cells.upload(joint, em, scal)
    cells.prepare()
    ref = refbind.RefCells(C, K)
    ref.set(joint, em, scal)
    ref.prepare_conditioner()
    cell, pos3, u3 = synth.make_queries(n, C, seed=seed + 1, binned=False)
    rng = np.random.default_rng(seed + 2)
    nrm = rng.normal(size=(3, n)).astype(np.float32)
    nrm /= np.linalg.norm(nrm, axis=0, keepdims=True)
    return cells, ref, cell, pos3, u3, np.ascontiguousarray(nrm)


def _compare(g, r, ref, cell, pos3, u3, bsdf, normal3):
    np.testing.assert_array_equal(g["valid"], r["valid"])
    v = r["valid"].astype(bool)
    assert v.sum() > 0.5 * v.size
    diff = np.nonzero((g["comp"] != r["comp"]) & v)[0]
    for i in diff[:100]:
        comps = ref.product_components(int(cell[i]), pos3[:, i], bsdf, None if normal3 is None else normal3[:, i])
        assert np.abs(comps[:, 6] - u3[0, i]).min() < TIE_DELTA, (i, u3[0, i], g["comp"][i], r["comp"][i])
    same = v & (g["comp"] == r["comp"])
    assert same.sum() >= 0.998 * v.sum(), (same.sum(), v.sum())
    ang = np.arccos(np.clip((g["dir"][:, same] * r["dir"][:, same]).sum(axis=0), -1, 1))
    assert np.quantile(ang, 0.999) < 2e-3 and ang.max() < 2e-2, ("direction", np.quantile(ang, [0.5, 0.999, 1.0]))
    np.testing.assert_allclose(g["inv_jac"][same], r["inv_jac"][same], rtol=1e-3, atol=1e-5)
    big = same & (r["pdf"] > 1e-6)
    rel = np.abs(g["pdf"][big] - r["pdf"][big]) / r["pdf"][big]
    assert rel.max() < PDF_RTOL_MAX and np.quantile(rel, 0.99) < PDF_RTOL and np.median(rel) < 1e-4, np.quantile(rel, [0.5, 0.99, 0.999, 1.0])


@pytest.mark.parametrize("K,diffuse", [(16, True), (64, True), (64, False), (128, True)])
def test_product_sample_pdf_parity(K, diffuse):
    C, n = 4, 6000
    cells, ref, cell, pos3, u3, nrm = _setup(C, K, n, seed=50 + K)
    bsdf = _bsdf()
    normal3 = nrm if diffuse else None
    g = cells.sample_pdf_product(cell, pos3, u3, bsdf, normal3=normal3)
    r, _ = ref.product_sample_pdf(cell, pos3, bsdf, u3=u3, normal3=normal3)
    _compare(g, r, ref, cell, pos3, u3, bsdf, normal3)


def test_product_components_match_reference():
    """The K*M product components themselves (weights, tangent means, covariances) at single vertices, via the
    pdf: the product pdf at a fixed direction is a weighted sum over all of them."""
    C, K, n = 3, 32, 4000
    cells, ref, cell, pos3, u3, nrm = _setup(C, K, n, seed=7)
    bsdf = _bsdf(M_zero=False)
    rng = np.random.default_rng(3)
    d = rng.normal(size=(3, n)).astype(np.float32)
    d /= np.linalg.norm(d, axis=0, keepdims=True)
    d = np.ascontiguousarray(d)
    g = cells.sample_pdf_product(cell, pos3, None, bsdf, normal3=nrm, dir3=d)
    r, _ = ref.product_sample_pdf(cell, pos3, bsdf, dir3=d, normal3=nrm)
    np.testing.assert_array_equal(g["valid"], r["valid"])
    big = r["valid"].astype(bool) & (r["pdf"] > 1e-6)
    assert big.sum() > 100
    rel = np.abs(g["pdf"][big] - r["pdf"][big]) / r["pdf"][big]
    assert np.quantile(rel, 0.99) < PDF_RTOL and rel.max() < PDF_RTOL_MAX and np.median(rel) < 1e-4, np.quantile(rel, [0.5, 0.99, 1.0])


def _shading_frames(n, seed):
    """Random orthonormal shading frames (to-world matrix [s t n], row-major planes) and local incident directions."""
    rng = np.random.default_rng(seed)
    nn = rng.normal(size=(n, 3))
    nn /= np.linalg.norm(nn, axis=1, keepdims=True)
    a = rng.normal(size=(n, 3))
    ss = a - (a * nn).sum(axis=1, keepdims=True) * nn
    ss /= np.linalg.norm(ss, axis=1, keepdims=True)
    tt = np.cross(nn, ss)
    W = np.stack([ss, tt, nn], axis=2)                      # [n, row, col]: columns s, t, n
    frame9 = np.ascontiguousarray(W.reshape(n, 9).T.astype(np.float32))
    wi = rng.normal(size=(n, 3))
    wi[:, 2] = np.abs(wi[:, 2]) + 0.1
    wi /= np.linalg.norm(wi, axis=1, keepdims=True)
    return frame9, np.ascontiguousarray(wi.T.astype(np.float32))


def test_product_pdf_outliers_are_float32_conditioning():
    """Why the product pdf is held to 2e-3 at the 99th percentile and 5e-2 at the maximum instead of 1e-4: at the
    vertices where GPU and reference differ most the product pdf amplifies float32 rounding several-hundred-fold.
    Three differences are measured against the FLOAT64 evaluation of the reference's formulas (product_f64.py):
    the two sides' conditioner planes agree to float32 rounding, yet that alone moves the float64 pdf by ~6e-3
    there; the REFERENCE's own float32 result is as far from float64 as the GPU's; and the GPU is never further
    from float64 than 3 x the reference's worst.  At typical vertices all three are below 1e-4."""
    from product_f64 import product_pdf64
    K, C, n = 128, 4, 6000
    cells, ref, cell, pos3, u3, nrm = _setup(C, K, n, seed=50 + K)
    bsdf = _bsdf()
    rng = np.random.default_rng(3)
    d = rng.normal(size=(3, n)).astype(np.float32)
    d /= np.linalg.norm(d, axis=0, keepdims=True)
    d = np.ascontiguousarray(d)                       # pdf at GIVEN directions: no sampled-direction difference
    gp = cells.sample_pdf_product(cell, pos3, None, bsdf, normal3=nrm, dir3=d)
    rp, _ = ref.product_sample_pdf(cell, pos3, bsdf, dir3=d, normal3=nrm)
    gpdf = gp["pdf"].cpu().numpy() if hasattr(gp["pdf"], "cpu") else np.asarray(gp["pdf"])
    Pg, Pr = cells.conditioner(), ref.conditioner()
    big = rp["valid"].astype(bool) & (rp["pdf"] > 1e-6)
    rel = np.where(big, np.abs(gpdf - rp["pdf"]) / np.maximum(rp["pdf"], 1e-30), 0)
    worst = list(np.argsort(-rel)[:16])
    typical = list(np.nonzero(big)[0][:16])
    e_gpu, e_ref, e_in = [], [], []
    for i in worst + typical:
        c = int(cell[i])
        fg = product_pdf64(Pg[c], pos3[:, i], bsdf, nrm[:, i], d[:, i])
        fr = product_pdf64(Pr[c], pos3[:, i], bsdf, nrm[:, i], d[:, i])
        e_in.append(abs(fg - fr) / fr)
        e_gpu.append(abs(gpdf[i] - fg) / fg)
        e_ref.append(abs(rp["pdf"][i] - fr) / fr)
    e_gpu, e_ref, e_in = np.array(e_gpu), np.array(e_ref), np.array(e_in)
    print("product pdf vs float64: inputs %.2e | worst 16: gpu %.2e ref %.2e | typical: gpu %.2e ref %.2e | gpu-ref worst %.2e"
          % (e_in.max(), e_gpu[:16].max(), e_ref[:16].max(), np.median(e_gpu[16:]), np.median(e_ref[16:]), rel[worst[0]]))
    # the two sides' conditioner planes (float32 prepare(): Cholesky of Sigma_aa, Schur complement) agree to float32
    # rounding; at the worst vertices that alone moves the FLOAT64 pdf by up to ~6e-3 -- the amplification that also
    # separates each side's float32 evaluation from float64
    cs = sorted(set(int(cell[i]) for i in worst))
    plane = max(float((np.abs(Pg[c] - Pr[c]).max(axis=1) / np.maximum(np.abs(Pr[c]).max(axis=1), 1e-30)).max())
                for c in cs)
    print("conditioner planes: worst relative difference %.2e -> float64 pdf moves %.2e (x%.0f)" % (plane, e_in.max(), e_in.max() / max(plane, 1e-30)))
    assert plane < 5e-6, plane                        # measured 4.4e-7 (a few float32 ulp) -> pdf moves 6e-3: x14000
    assert np.median(e_in[16:]) < 1e-4 and np.median(e_gpu[16:]) < 1e-4 and np.median(e_ref[16:]) < 1e-4   # typical vertices: exact
    assert e_ref[:16].max() > 0.3 * rel[worst[0]], (e_ref[:16].max(), rel[worst[0]])   # the reference's own float32 is off as far
    assert e_gpu.max() <= 3 * e_ref.max() + 1e-4, (e_gpu.max(), e_ref.max())


@pytest.mark.parametrize("K", [16, 64])
def test_product_glossy_branch_parity(K):
    """sdmm_proc.cpp:340-355: the BSDF mixture is given in the local shading frame; per vertex it is rotated to the
    incident direction (rotate_to_wo) and taken to world space before the product."""
    C, n = 4, 6000
    cells, ref, cell, pos3, u3, _ = _setup(C, K, n, seed=80 + K)
    frame9, wi3 = _shading_frames(n, seed=K)
    bsdf = _bsdf()
    g = cells.sample_pdf_product(cell, pos3, u3, bsdf, frame9=frame9, wi3=wi3)
    r, _ = ref.product_sample_pdf(cell, pos3, bsdf, u3=u3, frame9=frame9, wi3=wi3)
    np.testing.assert_array_equal(g["valid"], r["valid"])
    v = r["valid"].astype(bool)
    same = v & (g["comp"] == r["comp"])
    assert v.sum() > 0.5 * n and same.sum() >= 0.998 * v.sum(), (v.sum(), same.sum())
    ang = np.arccos(np.clip((g["dir"][:, same] * r["dir"][:, same]).sum(axis=0), -1, 1))
    assert np.quantile(ang, 0.999) < 2e-3 and ang.max() < 2e-2, np.quantile(ang, [0.5, 0.999, 1.0])
    big = same & (r["pdf"] > 1e-6)
    rel = np.abs(g["pdf"][big] - r["pdf"][big]) / r["pdf"][big]
    assert rel.max() < PDF_RTOL_MAX and np.quantile(rel, 0.99) < PDF_RTOL and np.median(rel) < 1e-4, np.quantile(rel, [0.5, 0.99, 1.0])
    # the transform matters: without it the result is a different distribution
    plain = cells.sample_pdf_product(cell, pos3, u3, bsdf)
    assert np.abs(plain["pdf"][big] - g["pdf"][big]).max() > 0.1


def test_product_samples_follow_the_product_pdf():
    """Size-independent self-consistency: the fraction of the product's own samples falling in a region equals
    the integral of the product pdf over that region (uniform quadrature through the pdf-only entry point), and
    the product pulls the samples towards the BSDF lobe."""
    C, K, n = 1, 16, 200000
    cells, ref, _, _, _, _ = _setup(C, K, 16, seed=11)
    rng = np.random.default_rng(5)
    cell = np.zeros(n, np.uint32)
    x = np.array([0.45, 0.5, 0.55], np.float32)
    pos3 = np.ascontiguousarray(np.repeat(x[:, None], n, axis=1))
    u3 = rng.random(size=(3, n), dtype=np.float32)
    bsdf = np.asarray([[1.0, 0.0, 0.0, 1.0, 0.3, 0.0, 0.3]], np.float32)
    g = cells.sample_pdf_product(cell, pos3, u3, bsdf)
    assert g["valid"].all()
    ok = g["inv_jac"] > 0
    assert ok.mean() > 0.95
    d = rng.normal(size=(3, n)).astype(np.float32)
    d /= np.linalg.norm(d, axis=0, keepdims=True)
    d = np.ascontiguousarray(d)
    q = cells.sample_pdf_product(cell, pos3, None, bsdf, dir3=d)
    edges = np.linspace(-1, 1, 9)
    freq = np.histogram(g["dir"][2, ok], bins=edges)[0] / n
    mass = np.asarray([4 * np.pi * np.mean(q["pdf"].astype(np.float64) * ((d[2] >= lo) & (d[2] < hi)))
                       for lo, hi in zip(edges[:-1], edges[1:])])
    assert abs(mass.sum() - ok.mean()) < 0.03, (mass.sum(), ok.mean())
    np.testing.assert_allclose(freq, mass, atol=0.01, rtol=0.08)
    c = cells.sample_pdf(cell, pos3, u3)
    okc = c["valid"].astype(bool) & (c["inv_jac"] > 0)
    assert g["dir"][2, ok].mean() > c["dir"][2, okc].mean()


def test_product_argument_errors():
    from sdmm_mitsuba_b200 import capi
    cells, ref, cell, pos3, u3, nrm = _setup(2, 16, 64, seed=3)
    bad = _bsdf().copy()
    bad[1, 1:4] = 0    # zero mean direction
    with pytest.raises(capi.SdmmError):
        cells.sample_pdf_product(cell, pos3, u3, bad)
    with pytest.raises(capi.SdmmError):
        cells.sample_pdf_product(cell, pos3, u3, np.zeros((17, 7), np.float32) + 1)
    with pytest.raises(capi.SdmmError):   # frame without wi
        cells.sample_pdf_product(cell, pos3, u3, _bsdf(), frame9=np.zeros((9, cell.size), np.float32))
    # cells outside the context come back invalid, not out-of-bounds reads
    far = cell.copy()
    far[::2] = 1000
    g = cells.sample_pdf_product(far, pos3, u3, _bsdf())
    assert not g["valid"][::2].any() and (g["pdf"][::2] == 0).all()
