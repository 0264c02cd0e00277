"""Progressive training loop: the device path tracer + host loop against the CPU restatement
of the reference integrator built on the reference's own sdmm headers (oracle/ref_render.inc), on the Cornell
box of test-suite/scenes/cornell-box (BASELINE.json configs[0]).  Both sides draw from identical per-path PCG32
streams, so paths, collected samples, tree splits and fitted mixtures are comparable one to one."""
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

from parity import PARAM_RTOL, joint_param_errs


def test_host_tree_matches_reference_stree(oracle_ref):
    """HostTree (sdmm_mitsuba_b200/tree.py) vs the reference's STree: bootstrap split_to_depth(2) and the
    data-driven split_leaf_recurse on clustered points (spatial_tree.h:133-292)."""
    import torch  # noqa: F401  (only to keep import order stable on the GPU box)
    from sdmm_mitsuba_b200.tree import HostTree
    rng = np.random.default_rng(3)
    n = 20000
    centres = rng.uniform(0.15, 0.85, (4, 3))
    p = np.clip(centres[rng.integers(0, 4, n)] + rng.normal(0, 0.06, (n, 3)), 1e-4, 1 - 1e-4).astype(np.float32)
    r = oracle_ref.RefTree(capacity=n + 16)
    r.split_to_depth(2)
    assert r.push(np.ascontiguousarray(p.T)) == 0
    r.split_leaves(300)
    nodes = r.nodes()
    t = HostTree([0, 0, 0], [1, 1, 1], 8192)
    t.split_to_depth(2)
    assert t.size() == 127 and t.n_leaves == 64
    # SpatialStats of the bootstrap leaves from the same points, then the same splits
    leaf_of = np.array([t.find(x) for x in p])
    cell_of = np.array([t.slot[i] for i in leaf_of])
    for node in t.leaves():
        m = leaf_of == node
        t.st_size[node] = int(m.sum())
        t.st_sum[node] = p[m].astype(np.float64).sum(0)
        t.st_sqr[node] = (p[m].astype(np.float64) ** 2).sum(0)

    def on_split(parent, s0, s1, axis, c0_min, c1_max):
        m = cell_of == parent
        upper = p[:, axis] > np.float32(c0_min)
        cell_of[m & upper] = s0
        cell_of[m & ~upper] = s1

    def child_stats(slot, amin, amax):
        q = p[cell_of == slot]
        inside = ((q >= amin) & (q <= amax)).all(1)
        q = q[inside].astype(np.float64)
        return int(inside.sum()), q.sum(0), (q * q).sum(0)

    t.split_leaves(300, child_stats, on_split)
    assert t.size() == nodes.shape[0]
    for i in range(t.size()):
        assert bool(nodes[i, 0] > 0) == t.is_leaf[i], i
        if not t.is_leaf[i]:
            assert int(nodes[i, 1]) == t.axis[i] and (int(nodes[i, 2]), int(nodes[i, 3])) == t.children[i], i
        else:
            assert int(nodes[i, 4]) == t.st_size[i], (i, nodes[i, 4], t.st_size[i])
    np.testing.assert_allclose(np.array(t.amin), nodes[:, 5:8], rtol=0, atol=1e-6)
    np.testing.assert_allclose(np.array(t.amax), nodes[:, 8:11], rtol=0, atol=1e-6)


def test_cornell_box_scene_is_closed_and_oriented():
    """All wall / box normals face the interior, the light faces down, bounds as the XML's geometry."""
    from sdmm_mitsuba_b200 import scenes
    quads, cam = scenes.cornell_box()
    assert len(quads) == 18
    lo, hi, norm = scenes.scene_bounds(quads)
    np.testing.assert_allclose(lo, [-1, 0, -1], atol=1e-5)
    np.testing.assert_allclose(hi, [1, 2, 1], atol=1e-5)
    assert abs(norm - 2.0) < 1e-5
    centre = np.array([0, 1, 0.0])
    for q in quads[:5]:
        n = np.cross(q["e1"], q["e2"])
        mid = q["p0"] + 0.5 * q["e1"] + 0.5 * q["e2"]
        assert np.dot(n, centre - mid) > 0
    light = quads[-1]
    assert np.cross(light["e1"], light["e2"])[1] < 0 and light["emission"][0] > 0


@pytest.mark.gpu
def test_progressive_loop_matches_cpu_restatement(oracle_ref):
    from sdmm_mitsuba_b200 import scenes
    from sdmm_mitsuba_b200.integrator import SDMMIntegrator
    W, H, K, MD, SPI = 320, 180, 16, 5, 64
    quads, cam = scenes.cornell_box(W, H)
    lo, hi, norm = scenes.scene_bounds(quads)
    rmin, rmax = scenes.root_aabb(lo, hi, norm)
    g = SDMMIntegrator(quads, cam, K=K, max_depth=MD, samples_per_iteration=SPI, seed=7, pool_capacity=4_000_000)
    r = oracle_ref.RefRenderer(quads, cam, lo, norm, rmin, rmax, K=K, max_depth=MD, capacity=150000)
    infos = []
    for it in range(3):
        seed = (g.seed * 0x9E3779B97F4A7C15 + g.iteration) & 0xFFFFFFFFFFFFFFFF
        st, new = g.render_pass(SPI, collect=True)
        info = g.optimize(new)
        g.iteration += 1
        g.total_spp += SPI
        rs = r.iteration(SPI, seed, True)
        infos.append((info, rs, int(st.guided_queries), int(st.mis_queries)))
        ig, ir = g.image(), r.image()
        # identical random streams -> identical paths: the images agree to float rounding; once guided sampling is
        # on, the rare mixtures that differ (below) perturb a few paths
        assert abs(ig.mean() - ir.mean()) <= 2e-3 * ir.mean(), (it, ig.mean(), ir.mean())
        if it < 2:
            assert np.abs(ig - ir).mean() <= 1e-5 * ir.mean(), (it, np.abs(ig - ir).mean())
    # iteration 1 trained the same cells on the same samples; iteration 2 was rendered with guiding
    assert infos[1][0]["trained"] == infos[1][1]["trained"] > 10
    assert infos[1][0]["train_samples"] == infos[1][1]["train_samples"]
    assert infos[2][2] > 100000 and infos[2][3] > 100000
    assert abs(infos[2][2] - infos[2][1]["guided_queries"]) <= 1e-3 * infos[2][2]
    # same tree
    nr, t = r.nodes(), g.tree
    assert t.size() == nr.shape[0] and t.size() > 127
    assert all(bool(nr[i, 0] > 0) == t.is_leaf[i] for i in range(t.size()))
    np.testing.assert_allclose(np.array(t.amin), nr[:, 5:8], atol=2e-5)   # split = mean of the leaf's samples
    # fitted mixtures leaf by leaf within the parameter tolerance (a seed drawn on a CDF tie or a sample on the
    # underflow edge can change a leaf completely: allow a few)
    j_gpu, _, s_gpu = g.cells.download(want_em=False)
    init_g = g.initialized.cpu().numpy()
    n_mix = n_ok = 0
    for node in t.leaves():
        jr, sr = r.leaf_joint(node)
        assert bool(nr[node, 12]) == bool(init_g[t.slot[node]])
        if jr is None:
            continue
        n_mix += 1
        e = joint_param_errs(j_gpu[t.slot[node]][None], jr[None])
        assert int(s_gpu[t.slot[node], 1]) == int(sr[1])
        n_ok += all(v <= PARAM_RTOL for v in e.values())
    assert n_mix > 50 and n_ok >= 0.9 * n_mix, (n_mix, n_ok)


@pytest.mark.gpu
def test_product_guiding_in_the_tracer_matches_cpu_restatement(oracle_ref):
    """sampleProduct inside the device path tracer (sdmm_proc.cpp:326-392, 409-417, 476-487): guided vertices sample
    conditional x learned-BSDF mixture (component 0 re-centred on the shading normal: all surfaces are diffuse), the
    BSDF fraction of the MIS is 0.3.  Oracle: the same loop on the CPU calling the reference's own sdmm::product
    (oracle/ref_render.inc compiled into libsdmm_ref_product.so); learned-BSDF files are LFS-only, so the mixture is
    synthetic (a cosine-like lobe about the normal + two broad side lobes)."""
    if not oracle_ref.product_available():
        pytest.skip("oracle/_ref/libsdmm_ref_product.so not built")
    from sdmm_mitsuba_b200 import scenes
    from sdmm_mitsuba_b200.integrator import SDMMIntegrator
    W, H, K, MD, SPI = 320, 180, 16, 5, 64
    bsdf = np.array([[0.6, 0, 0, 1, 0.45, 0.0, 0.45],
                     [0.2, 0.6, 0.0, 0.8, 0.30, 0.05, 0.25],
                     [0.2, -0.6, 0.0, 0.8, 0.30, -0.05, 0.25]], np.float32)
    quads, cam = scenes.cornell_box(W, H)
    lo, hi, norm = scenes.scene_bounds(quads)
    rmin, rmax = scenes.root_aabb(lo, hi, norm)
    g = SDMMIntegrator(quads, cam, K=K, max_depth=MD, samples_per_iteration=SPI, seed=7, pool_capacity=4_000_000)
    g.set_product(bsdf)
    r = oracle_ref.RefRenderer(quads, cam, lo, norm, rmin, rmax, K=K, max_depth=MD, capacity=150000, product_bsdf=bsdf)
    plain = SDMMIntegrator(quads, cam, K=K, max_depth=MD, samples_per_iteration=SPI, seed=7, pool_capacity=4_000_000)
    stats = []
    for it in range(3):
        seed = (g.seed * 0x9E3779B97F4A7C15 + g.iteration) & 0xFFFFFFFFFFFFFFFF
        st, new = g.render_pass(SPI, collect=True)
        g.optimize(new)
        g.iteration += 1
        g.total_spp += SPI
        rs = r.iteration(SPI, seed, True)
        stp, newp = plain.render_pass(SPI, collect=True)
        plain.optimize(newp)
        plain.iteration += 1
        plain.total_spp += SPI
        stats.append((int(st.guided_queries), int(st.mis_queries), rs, int(stp.guided_queries), int(stp.mis_queries)))
        ig, ir = g.image(), r.image()
        assert abs(ig.mean() - ir.mean()) <= 2e-3 * ir.mean(), (it, ig.mean(), ir.mean())
        if it < 2:
            assert np.abs(ig - ir).mean() <= 1e-5 * ir.mean(), (it, np.abs(ig - ir).mean())
    ga, gb, rs, pa, pb = stats[2]
    assert ga > 100000 and gb > 50000
    # the same vertices were guided as in the oracle, and with the 0.3 / 0.7 split instead of 0.5 / 0.5
    assert abs(ga - rs["guided_queries"]) <= 2e-3 * ga and abs(gb - rs["mis_queries"]) <= 2e-3 * gb
    frac = gb / float(ga + gb)
    assert 0.28 < frac < 0.32, frac
    assert 0.48 < pb / float(pa + pb) < 0.52
    # guided iteration: images agree with the CPU restatement far below the noise of the estimate
    ig, ir = g.image(), r.image()
    assert np.abs(ig - ir).mean() <= 2e-2 * ir.mean(), np.abs(ig - ir).mean() / ir.mean()


@pytest.mark.gpu
def test_async_schedule_matches_cpu_restatement(oracle_ref):
    """optimizeAsync, the reference's default schedule (volpath_sdmm.cpp:65, 180-242): fits become visible one pass
    later (double-buffered records + deferred `initialized` flags), one em_step per fit, seeds spaced from 0.1 x the
    leaf diagonal."""
    from sdmm_mitsuba_b200 import scenes
    from sdmm_mitsuba_b200.integrator import SDMMIntegrator
    W, H, K, MD, SPI = 320, 180, 16, 5, 64
    quads, cam = scenes.cornell_box(W, H)
    lo, hi, norm = scenes.scene_bounds(quads)
    rmin, rmax = scenes.root_aabb(lo, hi, norm)
    g = SDMMIntegrator(quads, cam, K=K, max_depth=MD, samples_per_iteration=SPI, seed=11, pool_capacity=4_000_000,
                       optimize_async=True)
    r = oracle_ref.RefRenderer(quads, cam, lo, norm, rmin, rmax, K=K, max_depth=MD, capacity=150000)
    r.set_async(True)
    infos = []
    for it in range(4):
        seed = (g.seed * 0x9E3779B97F4A7C15 + g.iteration) & 0xFFFFFFFFFFFFFFFF
        st, new = g.render_pass(SPI, collect=True)
        g.publish_pending()
        n_init = int(g.initialized.sum())
        info = g.optimize(new)
        g.iteration += 1
        g.total_spp += SPI
        rs = r.iteration(SPI, seed, True)
        infos.append((info, rs, int(st.guided_queries), n_init))
        ig, ir = g.image(), r.image()
        assert abs(ig.mean() - ir.mean()) <= 2e-3 * ir.mean(), (it, ig.mean(), ir.mean())
        if it < 3:
            assert np.abs(ig - ir).mean() <= 1e-5 * ir.mean(), (it, np.abs(ig - ir).mean())
    # the lag: the fit started after pass 1 is first used in pass 3 (pass 2 still ran unguided)
    assert infos[1][0]["trained"] == infos[1][1]["trained"] > 10
    assert infos[1][3] == 0 and infos[2][2] == 0 and infos[2][3] > 10 and infos[3][2] > 100000
    assert abs(infos[3][2] - infos[3][1]["guided_queries"]) <= 1e-3 * infos[3][2]
    for it, (a, b, _, _) in enumerate(infos):
        assert a["trained"] == b["trained"], (it, a, b)
        # passes 0-2 are unguided: identical paths, identical samples; a guided pass may differ in a few paths
        tol = 0 if it < 3 else 2e-3 * b["train_samples"]
        assert abs(a.get("train_samples", 0) - b["train_samples"]) <= tol, (it, a, b)
    nr, t = r.nodes(), g.tree
    assert t.size() == nr.shape[0]
    assert all(bool(nr[i, 0] > 0) == t.is_leaf[i] for i in range(t.size()))
    j_gpu, _, s_gpu = g.cells.download(want_em=False)
    n_mix = n_ok = 0
    for node in t.leaves():
        jr, sr = r.leaf_joint(node)
        if jr is None:
            continue
        n_mix += 1
        e = joint_param_errs(j_gpu[t.slot[node]][None], jr[None])
        assert int(s_gpu[t.slot[node], 1]) == int(sr[1])
        n_ok += all(v <= PARAM_RTOL for v in e.values())
    assert n_mix > 10 and n_ok >= 0.9 * n_mix, (n_mix, n_ok)


@pytest.mark.gpu
def test_interior_doubling_schedule_matches_cpu_restatement(oracle_ref):
    """BASELINE.json configs[3]: the progressive loop on the three-room interior with the doubling schedule
    (8, 16, 32, 64 spp) and a refit after every iteration, against the CPU restatement on the same schedule."""
    from sdmm_mitsuba_b200 import scenes
    from sdmm_mitsuba_b200.integrator import SDMMIntegrator
    W, H, K, MD = 160, 90, 16, 6
    quads, cam = scenes.interior(W, H)
    assert len(quads) <= 64
    lo, hi, norm = scenes.scene_bounds(quads)
    rmin, rmax = scenes.root_aabb(lo, hi, norm)
    g = SDMMIntegrator(quads, cam, K=K, max_depth=MD, samples_per_iteration=8, seed=3, pool_capacity=4_000_000)
    r = oracle_ref.RefRenderer(quads, cam, lo, norm, rmin, rmax, K=K, max_depth=MD, capacity=200000)
    seeds = []
    g.render(120, train_fraction=1.0, schedule="doubling",
             log=lambda info: seeds.append(((g.seed * 0x9E3779B97F4A7C15 + info["iteration"]) & 0xFFFFFFFFFFFFFFFF, info["spp"])))
    assert [s[1] for s in seeds] == [8, 16, 32, 64]
    rstats = [r.iteration(spp, seed, True) for seed, spp in seeds]
    ig, ir = g.image(), r.image()
    assert ir.mean() > 0 and abs(ig.mean() - ir.mean()) <= 5e-3 * ir.mean(), (ig.mean(), ir.mean())
    for it, (a, b) in enumerate(zip(g.stats, rstats)):
        assert a["trained"] == b["trained"], (it, a, b)
        tol = 0 if a["guided_queries"] == 0 else 5e-3 * b["train_samples"]
        assert abs(a.get("train_samples", 0) - b["train_samples"]) <= tol, (it, a, b)
    assert g.stats[-1]["guided_queries"] > 10000 and g.stats[2]["trained"] > 4
    assert abs(g.stats[-1]["guided_queries"] - rstats[-1]["guided_queries"]) <= 5e-3 * rstats[-1]["guided_queries"]
    nr, t = r.nodes(), g.tree
    assert t.size() == nr.shape[0]
    assert all(bool(nr[i, 0] > 0) == t.is_leaf[i] for i in range(t.size()))


@pytest.mark.gpu
def test_guiding_is_unbiased_and_reduces_variance():
    """MIS-combined guided sampling must converge to the same image as BSDF sampling (sdmm_proc.cpp:383-393,
    587-589) and, on this scene (tiny light, no next-event estimation), with a much smaller error per sample."""
    from sdmm_mitsuba_b200 import scenes
    from sdmm_mitsuba_b200.integrator import SDMMIntegrator
    W, H, K, MD = 320, 180, 16, 5
    quads, cam = scenes.cornell_box(W, H)
    ref = SDMMIntegrator(quads, cam, K=K, max_depth=MD, seed=99)
    for i in range(64):
        ref.iteration = 1000 + i
        ref.render_pass(64, collect=False, guided=False)
        ref.total_spp += 64
    img_ref = ref.image()
    b = SDMMIntegrator(quads, cam, K=K, max_depth=MD, samples_per_iteration=32, seed=5)
    for i in range(8):
        b.iteration = i
        b.render_pass(32, collect=False, guided=False)
        b.total_spp += 32
    gi = SDMMIntegrator(quads, cam, K=K, max_depth=MD, samples_per_iteration=32, seed=5)
    img_g = gi.render(256)
    img_b = b.image()
    assert abs(img_g.mean() - img_ref.mean()) < 0.02 * img_ref.mean()
    assert abs(img_b.mean() - img_ref.mean()) < 0.05 * img_ref.mean()
    mse_b, mse_g = ((img_b - img_ref) ** 2).mean(), ((img_g - img_ref) ** 2).mean()
    assert mse_g < 0.8 * mse_b, (mse_g, mse_b)
    assert int(gi.initialized.sum()) > 5


@pytest.mark.gpu
def test_render_writes_a_reference_layout_checkpoint(tmp_path):
    """saveCheckpoint (volpath_sdmm.cpp:121-130, 513): after render() the accelerator is written as
    checkpoints/model_%05i.asdmm; reading it back gives the same tree and, uploaded into a fresh context, the same
    query records for every trained leaf."""
    import torch
    from sdmm_mitsuba_b200 import checkpoint, scenes
    from sdmm_mitsuba_b200.integrator import SDMMIntegrator
    from sdmm_mitsuba_b200.mixture import SDMMCells
    quads, cam = scenes.cornell_box(320, 180)
    gi = SDMMIntegrator(quads, cam, K=16, max_depth=5, samples_per_iteration=32, seed=5, max_cells=256)
    gi.render(256)
    path = gi.save_checkpoint(str(tmp_path))
    assert path.endswith("checkpoints/model_%05i.asdmm" % gi.iteration)
    tree, joint, ex = checkpoint.load_asdmm(path, max_cells=256)
    assert tree.size() == gi.tree.size() and tree.n_leaves == gi.tree.n_leaves
    assert tree.is_leaf == gi.tree.is_leaf and tree.children == gi.tree.children
    trained = [i for i in range(tree.size()) if tree.is_leaf[i] and gi.has_mixture[gi.tree.slot[i]]]
    assert len(trained) > 5 and ex["K"] == 16
    assert sorted(ex["empty_slots"]) == sorted(tree.slot[i] for i in range(tree.size())
                                               if tree.is_leaf[i] and i not in trained)
    fresh = SDMMCells(256, 16)
    n = joint.shape[0]
    em = np.zeros((n, 81, 16), np.float32)
    sc = np.zeros((n, 8), np.uint32)
    for s in ex["em"]:
        em[s], sc[s] = ex["em"][s], ex["scalars"][s]
    # untrained leaves: any valid mixture (not compared)
    for s in ex["empty_slots"]:
        joint[s, 0] = 1.0 / 16
        joint[s, 6] = 1.0
        for d in range(5):
            joint[s, 7 + 6 * d] = 1.0
    fresh.upload(joint, em, sc)
    fresh.prepare(0, n)
    torch.cuda.synchronize()
    r_old, r_new = gi.cells.query_records(), fresh.query_records()
    for i in trained:
        # upload normalises the (already normalised) weights once more, like the reference's from_json -> prepare
        assert torch.allclose(r_old[gi.tree.slot[i]], r_new[tree.slot[i]], rtol=2e-5, atol=2e-6), i


def _pool_sorted_by_seq(scene, env, spp=64, batch=None):
    """One unguided collecting pass in a fresh process (the tracer reads its switches once); returns the pooled
    samples sorted by their push sequence number (the pool itself is filled in scheduling order)."""
    import os
    import subprocess
    import sys
    import tempfile
    out = tempfile.mktemp(suffix=".npy")
    code = (
        "import sys, numpy as np, torch\n"
        "sys.path.insert(0, %r)\n"
        "from sdmm_mitsuba_b200 import scenes\n"
        "from sdmm_mitsuba_b200.integrator import SDMMIntegrator\n"
        "quads, cam = getattr(scenes, %r)(160, 90)\n"
        "g = SDMMIntegrator(quads, cam, K=16, max_depth=6, samples_per_iteration=%d, seed=7, pool_capacity=4_000_000)\n"
        "st, new = g.render_pass(%d, collect=True)\n"
        "p = g.pool[:, :g.pool_count].cpu().numpy()\n"
        "order = np.argsort(p[14].view(np.uint32), kind='stable')\n"
        "np.save(%r, np.concatenate([p[:, order], g.image_sum.cpu().numpy().sum(0, keepdims=True)[:, :1].repeat(15, 0)], axis=1))\n"
        % (ROOT, scene, spp, spp, out))
    e = dict(os.environ, **env)
    if batch:
        e["SDMMB_PT_BATCH"] = str(batch)
    subprocess.check_call([sys.executable, "-c", code], env=e)
    a = np.load(out)
    os.remove(out)
    return a[:, :-1], float(a[0, -1])


@pytest.mark.gpu
@pytest.mark.parametrize("scene", ["interior", "cornell_box"])
def test_cluster_culling_and_batching_do_not_change_a_single_vertex(scene):
    """The bounce kernel only looks at the quads of clusters whose (enlarged) box the ray passes, and a pass is traced
    in batches of paths; both are pure reorganisations: every pooled vertex (position, direction, normal, weight, key)
    is bit-identical to the plain loop over all quads in the caller's order, traced in one batch."""
    plain, img_plain = _pool_sorted_by_seq(scene, {"SDMMB_PT_NO_CULL": "1"})
    culled, img_culled = _pool_sorted_by_seq(scene, {"SDMMB_PT_NO_CULL": "0"})
    batched, img_batched = _pool_sorted_by_seq(scene, {"SDMMB_PT_NO_CULL": "0"}, batch=50000)
    assert plain.shape[1] > 1000
    assert plain.shape == culled.shape == batched.shape
    assert np.array_equal(plain.view(np.uint32), culled.view(np.uint32))
    assert np.array_equal(plain.view(np.uint32), batched.view(np.uint32))
    assert abs(img_culled - img_plain) <= 1e-5 * abs(img_plain) and abs(img_batched - img_plain) <= 1e-5 * abs(img_plain)


@pytest.mark.gpu
def test_every_fit_of_the_loop_is_the_reference_em_from_its_own_state(oracle_ref):
    """Why the loop-level comparisons above allow a few leaves to differ: a free-running trajectory amplifies float32
    differences (a seed on a CDF tie, summation order), it is not an error of the fit.  Proof: the inputs of the LAST
    fit of the loop are captured (every trained cell's samples in training order and its state entering the fit); the
    reference's own em_step (oracle/_ref), started from that state on those samples with the integrator's 2|1 iteration
    rule (volpath_sdmm.cpp:299-305), reproduces the device's fitted mixture of EVERY trained leaf within the parameter
    tolerance, and the iteration / sample counters exactly."""
    from sdmm_mitsuba_b200 import scenes
    from sdmm_mitsuba_b200.integrator import SDMMIntegrator
    W, H, K, MD, SPI = 320, 180, 16, 5, 64
    quads, cam = scenes.cornell_box(W, H)
    g = SDMMIntegrator(quads, cam, K=K, max_depth=MD, samples_per_iteration=SPI, seed=7, pool_capacity=4_000_000)
    cap = {}

    def hook(d):
        cap["pre"] = g.cells.download()
        cap["point6"] = d["point6"].cpu().numpy().copy()
        cap["weight"] = d["weight"].cpu().numpy().copy()
        cap["off"] = np.asarray(d["offsets"]).astype(np.int64)
        cap["can"] = np.asarray(d["can"]).copy()

    for it in range(4):
        st, new = g.render_pass(SPI, collect=True)
        g.capture_fit = hook if it == 3 else None
        g.optimize(new)
        g.iteration += 1
        g.total_spp += SPI
    assert cap, "the last iteration trained nothing"
    j0, e0, s0 = cap["pre"]
    j1, e1, s1 = g.cells.download()
    off = cap["off"]
    idx = np.nonzero(cap["can"])[0]
    runs = s0[idx, 1]
    assert len(idx) > 30 and (runs < 4).any()
    worst, checked = {}, 0
    for young in (True, False):
        sub = idx[(runs < 4) == young]
        if sub.size == 0:
            continue
        cols = np.concatenate([np.arange(off[c], off[c + 1]) for c in sub])
        sub_off = np.concatenate([[0], np.cumsum(off[sub + 1] - off[sub])]).astype(np.uint32)
        o = oracle_ref.RefCells(len(sub), K)
        o.set(j0[sub], e0[sub], s0[sub])
        o.em_step(sub_off, np.ascontiguousarray(cap["point6"][:, cols]), np.ascontiguousarray(cap["weight"][cols]),
                  2 if young else 1)
        jr, er, sr = o.get()
        np.testing.assert_array_equal(s1[sub, 1], sr[:, 1])      # iterations_run
        np.testing.assert_array_equal(s1[sub, 2], sr[:, 2])      # samples_seen
        for k, c in enumerate(sub):
            e = joint_param_errs(j1[c][None], jr[k][None])
            checked += 1
            for name, v in e.items():
                if v > worst.get(name, (0.0, -1))[0]:
                    worst[name] = (v, int(c), int(off[c + 1] - off[c]))
    assert checked == len(idx)
    assert all(v[0] <= PARAM_RTOL for v in worst.values()), worst
