"""sdmm::initialize (sdmm/include/sdmm/opt/init.h:17-193).

CPU: the plain-C restatement against golden vectors made by the compiled reference
(tests/golden/init_*.npz, tools/make_golden.py) and, where /root/reference is mounted, against the
reference on fresh inputs.  GPU: sdmmb_initialize against the golden vectors and the restatement.
Seeds (which samples become component centres) must match EXACTLY: they depend on a sequential float32
CDF and a replayed PCG32 stream; everything else within the parameter tolerance."""
import os

import numpy as np
import pytest

import golden_util as G
from parity import PARAM_RTOL, assert_joint_close

INIT_CASES = ["init_k64", "init_k16"]


def _load(name):
    g = dict(np.load(os.path.join(G.GOLDEN_DIR, name + ".npz")))
    g["C"], g["K"], g["N"] = [int(v) for v in g["meta"]]
    return g


def _check_against(j, e, s, g, what):
    # the seeds: position means are copies of sample positions -> bit-exact
    np.testing.assert_array_equal(j[:, 1:4], g["joint"][:, 1:4], err_msg=what + ": seed positions")
    assert_joint_close(j, g["joint"], PARAM_RTOL, what)
    np.testing.assert_allclose(j[:, 4:7], g["joint"][:, 4:7], atol=2e-6, err_msg=what + ": lobe directions")
    np.testing.assert_allclose(j[:, 7:], g["joint"][:, 7:], rtol=1e-5, atol=1e-7, err_msg=what + ": covariances")
    np.testing.assert_allclose(e, g["em"], rtol=1e-5, atol=1e-12, err_msg=what + ": EM state / priors")
    np.testing.assert_array_equal(s[:, :7], g["scal"][:, :7], err_msg=what + ": scalars")


def test_pcg32_matches_reference_generator(oracle_port):
    g = _load("init_k64")
    np.testing.assert_array_equal(oracle_port.pcg32_floats(64), g["pcg32"])


@pytest.mark.parametrize("name", INIT_CASES)
def test_port_initialize_matches_golden(oracle_port, name):
    g = _load(name)
    o = oracle_port.PortCells(g["C"], g["K"])
    seeds = o.initialize(g["offsets"], g["point6"], g["normal3"], g["weight"], g["spatial_distance"])
    j, e, s = o.get()
    _check_against(j, e, s, g, "port " + name)
    # every seed index points at the sample whose position became the component mean
    for c in range(g["C"]):
        b = int(g["offsets"][c])
        for sc in range(g["K"] // 8):
            np.testing.assert_array_equal(g["point6"][0:3, b + seeds[c, sc]], g["joint"][c, 1:4, 8 * sc])
    d = o.derived()
    np.testing.assert_allclose(d[:, 9:34], g["derived"][:, 9:34], rtol=2e-5, atol=1e-6)
    np.testing.assert_allclose(d[:, 34], g["derived"][:, 34], rtol=5e-5)


def test_port_initialize_matches_reference_on_fresh_inputs(oracle_port, oracle_ref):
    from sdmm_mitsuba_b200 import synth
    for (C, K, N, seed) in [(6, 64, 2500, 11), (6, 32, 900, 12), (3, 128, 4000, 13)]:
        p6, w, off = synth.make_samples(C, N, seed=seed)
        n3 = synth.make_normals(C, N, seed=seed)
        sd = np.full(C, 3 * 0.6 / (K // 8), np.float32)
        r = oracle_ref.RefCells(C, K)
        r.initialize(off, p6, n3, w, sd)
        o = oracle_port.PortCells(C, K)
        o.initialize(off, p6, n3, w, sd)
        jr, er, sr = r.get()
        jo, eo, so = o.get()
        _check_against(jo, eo, so, dict(joint=jr, em=er, scal=sr), "port vs ref K=%d" % K)


def test_degenerate_cell_falls_back_to_uniform_cdf(oracle_port):
    """All samples identical: after the first seed every sampling weight is 0, Categorical::prepare fails
    and the uniform fallback (init.h:127-139) is taken; the result must stay finite."""
    K, N = 16, 64
    p6 = np.zeros((6, N), np.float32)
    p6[0:3] = 0.5
    p6[5] = 1.0
    n3 = np.zeros((3, N), np.float32)
    n3[2] = 1.0
    w = np.ones(N, np.float32)
    o = oracle_port.PortCells(1, K)
    o.initialize(np.array([0, N], np.uint32), p6, n3, w, np.array([0.3], np.float32))
    j, _, _ = o.get()
    assert np.isfinite(j).all() and (j[0, 1:4] == 0.5).all()


@pytest.mark.gpu
@pytest.mark.parametrize("name", INIT_CASES)
def test_gpu_initialize_matches_golden(name):
    from sdmm_mitsuba_b200.mixture import SDMMCells
    g = _load(name)
    cells = SDMMCells(g["C"], g["K"])
    cells.initialize(g["point6"], g["normal3"], g["weight"], g["offsets"], g["spatial_distance"])
    j, e, s = cells.download()
    _check_against(j, e, s, g, "gpu " + name)
    d = cells.derived()
    np.testing.assert_allclose(d[:, 9:34], g["derived"][:, 9:34], rtol=2e-5, atol=1e-6)
    np.testing.assert_allclose(d[:, 34], g["derived"][:, 34], rtol=5e-5)
    np.testing.assert_allclose(d[:, 35:37], g["derived"][:, 35:37], rtol=1e-6, atol=1e-7)


@pytest.mark.gpu
@pytest.mark.parametrize("C,K,N,ragged", [(64, 64, 4096, False), (97, 16, 700, True), (8, 128, 9000, True)])
def test_gpu_initialize_matches_port(oracle_port, C, K, N, ragged):
    """Device tensors, many cells, ragged cell sizes, multi-tile CDF (N > 2048); seeds must agree exactly."""
    import torch
    from sdmm_mitsuba_b200 import synth
    from sdmm_mitsuba_b200.mixture import SDMMCells
    p6, w, off = synth.make_samples(C, N, seed=21 + K)
    n3 = synth.make_normals(C, N, seed=21 + K)
    if ragged:  # carve ragged cells out of the same sample pool
        rng = np.random.default_rng(5)
        sizes = rng.integers(N // 3, N, C)
        off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint32)
    sd = (3 * np.random.default_rng(6).uniform(0.2, 1.0, C) / (K // 8)).astype(np.float32)
    o = oracle_port.PortCells(C, K)
    seeds_o = o.initialize(off, p6, n3, w, sd)
    jo, eo, so = o.get()
    cells = SDMMCells(C, K)
    dev = torch.device("cuda", 0)
    seeds_g = cells.initialize(torch.from_numpy(p6).to(dev), torch.from_numpy(n3).to(dev),
                               torch.from_numpy(w).to(dev), off, sd)
    np.testing.assert_array_equal(seeds_g, seeds_o)
    j, e, s = cells.download()
    _check_against(j, e, s, dict(joint=jo, em=eo, scal=so), "gpu vs port K=%d" % K)


@pytest.mark.gpu
def test_gpu_initialize_then_fit_matches_port(oracle_port):
    """The integrator's first optimize() of a new cell: initialize, then 2 x em_step + prepare
    (volpath_sdmm.cpp:292-307)."""
    from sdmm_mitsuba_b200 import synth
    from sdmm_mitsuba_b200.mixture import SDMMCells
    C, K, N = 8, 64, 3000
    p6, n3, w, off = synth.make_surface_samples(C, N, seed=31)
    sd = np.full(C, 3 * 1.0 / (K // 8), np.float32)
    o = oracle_port.PortCells(C, K)
    o.initialize(off, p6, n3, w, sd)
    o.em_step(off, p6, w, 2)
    jo, _, so = o.get()
    cells = SDMMCells(C, K)
    cells.initialize(p6, n3, w, off, sd)
    cells.set_samples(p6, w, off)
    cells.em_step(2, with_conditioner=True)
    j, _, s = cells.download()
    assert_joint_close(j, jo, PARAM_RTOL, "initialize + 2 em_step")
    assert (s[:, 1:3] == so[:, 1:3]).all()


@pytest.mark.gpu
def test_gpu_outlier_samples_do_not_poison_the_fit(oracle_port):
    """Samples many sigma away from every component (mixture density near the float32 underflow
    threshold) make w_n / sum_k overflow unless gamma = p / sum is formed first (em.h:295-309): the fit must
    stay finite and skip exactly the samples whose density underflows to 0 like the reference's exp does."""
    from sdmm_mitsuba_b200 import synth
    from sdmm_mitsuba_b200.mixture import SDMMCells
    C, K, N = 8, 64, 3000
    p6, w, off = synth.make_samples(C, N, seed=31)       # volumetric lobes, NOT on the seeds' thin manifolds
    n3 = synth.make_normals(C, N, seed=31)
    sd = np.full(C, 3 * 1.0 / (K // 8), np.float32)
    o = oracle_port.PortCells(C, K)
    o.initialize(off, p6, n3, w, sd)
    o.em_step(off, p6, w, 1)
    _, _, so = o.get()
    cells = SDMMCells(C, K)
    cells.initialize(p6, n3, w, off, sd)
    cells.set_samples(p6, w, off)
    ll = cells.em_step(3, with_conditioner=True, want_loglik=True).cpu().numpy()
    j, e, s = cells.download()
    assert np.isfinite(j).all() and np.isfinite(e).all()
    assert np.isinf(ll[0]).sum() > 100                    # the regime is actually exercised
    # skipped-sample bookkeeping within 1% of the restatement (the underflow cut is a band)
    seen1 = SDMMCells(C, K)
    seen1.initialize(p6, n3, w, off, sd)
    seen1.set_samples(p6, w, off)
    seen1.em_step(1)
    s1 = seen1.download()[2]
    assert np.abs(s1[:, 2].astype(np.int64) - so[:, 2].astype(np.int64)).max() <= 0.01 * N


def test_port_initialize_then_fit_matches_reference(oracle_port, oracle_ref):
    from sdmm_mitsuba_b200 import synth
    C, K, N = 4, 64, 2000
    p6, n3, w, off = synth.make_surface_samples(C, N, seed=41)
    sd = np.full(C, 3 * 1.0 / (K // 8), np.float32)
    o = oracle_port.PortCells(C, K)
    o.initialize(off, p6, n3, w, sd)
    o.em_step(off, p6, w, 2)
    r = oracle_ref.RefCells(C, K)
    r.initialize(off, p6, n3, w, sd)
    r.em_step(off, p6, w, 2)
    assert_joint_close(o.get()[0], r.get()[0], PARAM_RTOL, "port vs ref: initialize + 2 em_step")
    assert (o.get()[2][:, 1:3] == r.get()[2][:, 1:3]).all()
