"""Pins the oracle: known-answer literals held by the reference's OWN doctest files
(sdmm/tests/distributions.cpp, linalg.cpp -- SURVEY.md 8c) checked against (a) the plain-C
restatement oracle/sdmm_oracle.c and (b) the reference headers compiled as oracle/_ref (when
/root/reference was present at build time)."""
import numpy as np
import pytest

# the +1e-5 Cholesky pivot jitter (sdmm/linalg/cholesky.h:32) moves results by ~1e-5 relative; the
# reference's own tests compare with sdmm::epsilon<float> = 1e-5 (sdmm/tests/utils.h:7-14)
KAT_RTOL = 2e-5


def test_pdf_scalar_kat_port(oracle_port):
    # sdmm/tests/distributions.cpp:16-36: 2-D, cov diag(1,2), x=(1,2) -> 0.025110965476047437
    v = oracle_port.gauss_pdf_euclid([0, 0], [[1, 0], [0, 2]], [1, 2])
    assert abs(v - 0.025110965476047437) / 0.025110965476047437 < KAT_RTOL


def test_pdf_scalar_kat_ref(oracle_ref):
    v = oracle_ref.kat_pdf_scalar()
    assert abs(v - 0.025110965476047437) / 0.025110965476047437 < KAT_RTOL


def test_pdf_dynamic_kat(oracle_port, oracle_ref):
    # sdmm/tests/distributions.cpp:148-248: two 2-D components, NumPy literals at :59-62, :69-72
    expected_12 = [0.05207269256276517, 0.0018674935212148857]
    expected_00 = [0.054777174730721315, 0.002397909146739601]
    means = [[0, 0], [1, 1]]  # set_mean(Value(0, 0), Value(1, 1)): both components sit at (0, 1)
    covs = [np.array([[3, 0.5], [0.5, 1.4]]), np.array([[2, 0], [0, 0.1]])]
    ref = oracle_ref.kat_pdf_dynamic()
    for k in range(2):
        m = [means[0][k], means[1][k]]
        p12 = oracle_port.gauss_pdf_euclid(m, covs[k], [1, 2])
        p00 = oracle_port.gauss_pdf_euclid(m, covs[k], [0, 0])
        # the reference's approx_equals is ABSOLUTE sdmm::epsilon<float> = 1e-5 (sdmm/tests/utils.h:7-14);
        # component 1 (variance 0.1 + the 1e-5 pivot jitter) is off by 4.5e-4 relative, 8e-7 absolute
        assert abs(p12 - expected_12[k]) <= 1e-5
        assert abs(p00 - expected_00[k]) <= 1e-5
        assert abs(ref[k] - expected_12[k]) <= 1e-5
        assert abs(ref[2 + k] - expected_00[k]) <= 1e-5
        assert abs(ref[k] - p12) <= 1e-6 * p12 + 1e-9 and abs(ref[2 + k] - p00) <= 1e-6 * p00 + 1e-9
    # posterior = weight x pdf (:243-246), weights 0.2 / 0.8
    np.testing.assert_allclose(ref[4:6], np.array([0.2, 0.8]) * ref[0:2], rtol=1e-6)


def test_conditioner_kat(oracle_port, oracle_ref):
    # sdmm/tests/distributions.cpp:251-332; expected literals at :302-307 and :326-329
    exp_cov0 = np.array([[0.99645569, 0.18835443], [0.18835443, 2.93316455]])
    exp_cov1 = np.array([[0.995, 0.18], [0.18, 2.84]])
    exp_mean = np.array([[2.02278481, -0.78227848], [2.05, 1.6]])
    cov0 = np.array([[3, 0.5, 0.1, 0.4], [0.5, 1.4, 0, 0.2], [0.1, 0, 1, 0.2], [0.4, 0.2, 0.2, 3]], np.float32)
    cov1 = np.array([[2, 0, 0.1, 0.4], [0, 0.5, 0, 0.2], [0.1, 0, 1, 0.2], [0.4, 0.2, 0.2, 3]], np.float32)
    mean = np.array([[0, 1, 2, -1], [0, 1, 2, 1]], np.float32)  # per component
    x = np.array([1, 2], np.float32)
    for k, (cov, ec) in enumerate([(cov0, exp_cov0), (cov1, exp_cov1)]):
        La, M, cc, Lc = oracle_port.schur(cov, 2)
        np.testing.assert_allclose(cc, ec, rtol=1e-4, atol=2e-5)
        cm = mean[k, 2:] + M @ (x - mean[k, :2])
        np.testing.assert_allclose(cm, exp_mean[k], rtol=1e-4, atol=2e-5)
    ref = oracle_ref.kat_conditioner()
    np.testing.assert_allclose(ref[0:4].reshape(2, 2), exp_cov0, rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(ref[4:8].reshape(2, 2), exp_cov1, rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(ref[8:12].reshape(2, 2), exp_mean, rtol=1e-4, atol=2e-5)


REF_CHOLESKY_MATS = [
    # sdmm/tests/linalg.cpp:24-32: the two (nearly rank-1) SPD matrices the reference factors
    [[0.0352478, 0.294885, 0.13223], [0.294885, 2.46774, 1.10656], [0.13223, 1.10656, 0.496206]],
    [[0.0752945, 0.203741, 0.0132939], [0.203741, 0.551391, 0.0359771], [0.0132939, 0.0359771, 0.00235748]],
]
EXTRA_CHOLESKY_MATS = [
    # well-conditioned extras (not from the reference): let the factor itself be compared with LAPACK
    [[4, 12, -16], [12, 37, -43], [-16, -43, 98]],
    [[2, -1, 0], [-1, 2, -1], [0, -1, 2]],
]


@pytest.mark.parametrize("m", REF_CHOLESKY_MATS + EXTRA_CHOLESKY_MATS)
def test_cholesky_kat(oracle_port, oracle_ref, m):
    well_conditioned = m in EXTRA_CHOLESKY_MATS
    m = np.array(m, np.float32)
    L_port, psd = oracle_port.cholesky(m)
    L_ref, psd_r = oracle_ref.kat_cholesky3(m)
    assert psd and psd_r
    # same algorithm, same +1e-5 pivot jitter (sdmm/linalg/cholesky.h:32) as the compiled reference
    np.testing.assert_allclose(L_port, L_ref, rtol=1e-6, atol=1e-6)
    # "Checking A = L * L^T for consistency" (linalg.cpp:55-59): the reference's bound is absolute
    # sdmm::epsilon = 1e-5 -- which the jitter alone uses up on the diagonal, so 2e-5 here, scaled
    # by the largest entry for the extras
    scale = max(1.0, float(np.abs(m).max()))
    np.testing.assert_allclose(L_port @ L_port.T, m + 1e-5 * np.eye(3), rtol=0, atol=1e-5 * scale)
    np.testing.assert_allclose(L_port @ L_port.T, m, rtol=0, atol=2e-5 * scale)
    if well_conditioned:
        L_np = np.linalg.cholesky(m.astype(np.float64) + 1e-5 * np.eye(3))
        np.testing.assert_allclose(L_port, L_np, rtol=1e-5, atol=1e-5)
    # solve / inverse consistency (linalg.cpp:74-125): b = (1, 2, 0.5)
    b = np.array([1, 2, 0.5], np.float32)
    x = oracle_port.solve_lower(L_port, b)
    np.testing.assert_allclose(L_port @ x, b, rtol=1e-5, atol=1e-5)
    inv = oracle_port.inverse_lower_tri(L_port)
    np.testing.assert_allclose(L_port @ inv, np.eye(3), atol=1e-5)


def test_categorical_kat(oracle_port, oracle_ref):
    # sdmm/tests/distributions.cpp:335-364: prepare normalises pmf, cdf ends at 1; all-zero invalid
    pmf = np.array([1, 2, 3, 4], np.float32)
    p, c, ok = oracle_port.categorical_prepare(pmf)
    pr, cr, okr = oracle_ref.kat_categorical(pmf)
    assert ok and okr
    np.testing.assert_allclose(p, [0.1, 0.2, 0.3, 0.4], rtol=1e-6)
    np.testing.assert_allclose(c, [0.1, 0.3, 0.6, 1.0], rtol=1e-6)
    np.testing.assert_array_equal(p, pr)
    np.testing.assert_array_equal(c, cr)
    _, _, ok0 = oracle_port.categorical_prepare(np.zeros(4, np.float32))
    _, _, ok0r = oracle_ref.kat_categorical(np.zeros(4, np.float32))
    assert not ok0 and not ok0r


@pytest.mark.parametrize("n", [[0, 0, 1], [0, 0, -1], [1, 0, 0], [0.6, 0.0, 0.8], [-0.36, 0.48, -0.8]])
def test_frame_kat(oracle_port, oracle_ref, n):
    # sdmm/tests/linalg.cpp:142-159: to * n = +z, rows orthonormal
    n = np.array(n, np.float32)
    to, fr = oracle_port.frame(n)
    np.testing.assert_allclose(to @ n, [0, 0, 1], atol=1e-6)
    np.testing.assert_allclose(to @ to.T, np.eye(3), atol=1e-6)
    np.testing.assert_allclose(fr, to.T, atol=0)
    np.testing.assert_allclose(to, oracle_ref.kat_frame(n), atol=1e-7)
