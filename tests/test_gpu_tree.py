"""Spatial cell structure on the device: STree::find and the renderer's sample
collection, against the REFERENCE's own STree (oracle/_ref)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ref_tree(oracle_ref, n_points=30000, threshold=400, seed=0):
    rng = np.random.default_rng(seed)
    t = oracle_ref.RefTree(capacity=n_points + 16)
    t.split_to_depth(2)
    # clustered points so that the variance-driven splits are unbalanced
    centres = rng.uniform(0.15, 0.85, (5, 3))
    which = rng.integers(0, 5, n_points)
    p = np.clip(centres[which] + rng.normal(0, 0.07, (n_points, 3)), 1e-4, 1 - 1e-4).astype(np.float32).T.copy()
    assert t.push(p) == 0
    t.split_leaves(threshold)
    return t, p


def _flatten(nodes):
    from sdmm_mitsuba_b200 import tree
    return tree.flatten(nodes[:, 0] > 0, nodes[:, 1], nodes[:, 2], nodes[:, 3], nodes[:, 5:8])


def test_find_cells_matches_reference_tree(oracle_ref):
    from sdmm_mitsuba_b200.mixture import SDMMCells
    t, p = _ref_tree(oracle_ref)
    nodes = t.nodes()
    flat, leaf_slot = _flatten(nodes)
    n_leaves = int((nodes[:, 0] > 0).sum())
    assert n_leaves > 64                      # the data-driven splits happened
    cells = SDMMCells(n_leaves, 16)
    cells.set_tree(flat)
    rng = np.random.default_rng(1)
    q = rng.uniform(0, 1, (3, 20000)).astype(np.float32)
    # points exactly ON split planes take the reference's strict `<` branch
    inner = np.nonzero(nodes[:, 0] == 0)[0]
    for j, i in enumerate(inner[:200]):
        ax = int(nodes[i, 1])
        q[ax, j] = nodes[int(nodes[i, 2]), 5 + ax]
    ref_node = t.find(q)
    assert (ref_node >= 0).all()
    got = cells.find_cells(q)
    np.testing.assert_array_equal(got.astype(np.int64), leaf_slot[ref_node])
    import torch
    got_d = cells.find_cells(torch.from_numpy(q).cuda()).cpu().numpy().view(np.uint32)
    np.testing.assert_array_equal(got_d, got)


def test_collect_samples_bins_like_push_back(oracle_ref, oracle_port):
    """Device-side collection == the reference's find + push_back per sample: same per-cell counts,
    invalid weights dropped (Data::push_back), and the EM fit on the collected samples matches the fit
    on the same samples handed over pre-sorted (order inside a cell is arbitrary on both sides)."""
    import torch
    from parity import PARAM_RTOL, assert_joint_close
    from sdmm_mitsuba_b200 import synth
    from sdmm_mitsuba_b200.mixture import SDMMCells
    t, p = _ref_tree(oracle_ref, n_points=20000, threshold=600, seed=3)
    nodes = t.nodes()
    flat, leaf_slot = _flatten(nodes)
    C = int((nodes[:, 0] > 0).sum())
    K, n = 16, p.shape[1]
    rng = np.random.default_rng(4)
    d = rng.normal(size=(3, n)).astype(np.float32)
    d /= np.linalg.norm(d, axis=0, keepdims=True)
    point6 = np.ascontiguousarray(np.concatenate([p, d], axis=0))
    weight = np.exp(rng.normal(size=n)).astype(np.float32)
    weight[::97] = 0.0
    weight[5] = np.nan
    weight[11] = np.inf
    valid = np.isfinite(weight) & (weight >= 1e-8)
    slot = leaf_slot[t.find(p)]
    expect_counts = np.bincount(slot[valid], minlength=C)

    joint, em, scal = synth.make_initial_model(C, K, seed=5)
    a = SDMMCells(C, K)
    a.upload(joint, em, scal)
    a.set_tree(flat)
    counts, dropped = a.collect_samples(torch.from_numpy(point6).cuda(), torch.from_numpy(weight).cuda())
    np.testing.assert_array_equal(counts, expect_counts)
    assert dropped == int((~valid).sum())
    a.set_min_samples(64)
    a.em_step(2, with_conditioner=True)
    ja, _, sa = a.download()

    order = np.argsort(slot[valid], kind="stable")
    p6s = np.ascontiguousarray(point6[:, valid][:, order])
    ws = np.ascontiguousarray(weight[valid][order])
    off = np.concatenate([[0], np.cumsum(expect_counts)]).astype(np.uint32)
    b = SDMMCells(C, K)
    b.upload(joint, em, scal)
    b.set_samples(p6s, ws, off)
    b.set_min_samples(64)
    b.em_step(2, with_conditioner=True)
    jb, _, sb = b.download()
    np.testing.assert_array_equal(sa[:, 1:3], sb[:, 1:3])
    assert_joint_close(ja, jb, PARAM_RTOL, "collected vs pre-sorted")
    # cells below the canBeOptimized threshold were left untouched, the others trained twice
    small = expect_counts < 64
    assert (sa[small, 1] == 0).all() and (sa[~small, 1] == 2).all()
    np.testing.assert_array_equal(ja[small][:, 1:], joint[small][:, 1:])   # means and covariances as uploaded


def test_find_cells_in_a_65536_leaf_tree():
    """The reference caps the tree at 10000 nodes with uint16 child indices (spatial_tree.h:96, 311-312); BASELINE.json
    configs[4] needs 65536 cells.  131071 nodes, 65536 leaves: device find == host find, every slot reachable."""
    from sdmm_mitsuba_b200.mixture import SDMMCells
    from sdmm_mitsuba_b200.tree import HostTree
    eps = np.float32(1e-5)
    t = HostTree(np.zeros(3, np.float32) - eps, np.ones(3, np.float32) + eps, 65536)
    t.split_to_depth(5)                      # 8^5 = 32768 leaves
    for node in t.leaves():
        assert t._split(node, 0.5) is not None
    assert t.n_leaves == 65536 and t.size() == 131071
    flat = t.flatten()
    assert int(flat["child1"].max()) > 65535          # indices beyond uint16
    cells = SDMMCells(65536, 16)
    cells.set_tree(flat)
    rng = np.random.default_rng(4)
    q = rng.uniform(0, 1, (3, 40000)).astype(np.float32)
    got = cells.find_cells(q).astype(np.int64)
    want = np.array([t.slot[t.find(q[:, i])] for i in range(q.shape[1])], np.int64)
    np.testing.assert_array_equal(got, want)
    assert sorted(t.slot[n] for n in t.leaves()) == list(range(65536))
    # leaf centres hit every slot exactly once
    centres = np.array([0.5 * (t.amin[n] + t.amax[n]) for n in t.leaves()], np.float32).T.copy()
    hit = cells.find_cells(centres).astype(np.int64)
    np.testing.assert_array_equal(hit, np.array([t.slot[n] for n in t.leaves()]))


def test_copy_cells_is_sequential_whatever_path_it_takes():
    """sdmmb_copy_cells: children inherit the parent's complete state on a split (spatial_tree.h:187-196).  The
    copies are sequential by contract (a fresh child can be the source of a later copy); the library resolves chains
    and runs them as one kernel when no origin is overwritten, else pair by pair -- same result as a host loop."""
    from sdmm_mitsuba_b200 import synth
    from sdmm_mitsuba_b200.mixture import SDMMCells
    C, K = 12, 16
    joint, em, scal = synth.make_initial_model(C, K, seed=3)
    scal[:, 1] = np.arange(C)            # make the scalar words distinguishable too
    for src, dst in (([0, 8, 1, 9], [8, 9, 10, 11]),        # chains through fresh slots: one kernel
                     ([1, 0], [2, 1]),                      # an origin is overwritten later: ordered launches
                     ([3, 3, 5], [3, 4, 4])):               # self copy, a destination written twice
        cells = SDMMCells(C, K)
        cells.upload(joint, em, scal)
        cells.prepare()
        j0, e0, s0 = cells.download()
        q0 = cells.query_records().clone().cpu().numpy().reshape(C, -1)
        cells.copy_cells(src, dst)
        j1, e1, s1 = cells.download()
        q1 = cells.query_records().clone().cpu().numpy().reshape(C, -1)
        jw, ew, sw, qw = j0.copy(), e0.copy(), s0.copy(), q0.copy()
        for a, b in zip(src, dst):
            jw[b], ew[b], sw[b], qw[b] = jw[a], ew[a], sw[a], qw[a]
        np.testing.assert_array_equal(j1, jw)
        np.testing.assert_array_equal(e1, ew)
        np.testing.assert_array_equal(s1, sw)
        np.testing.assert_array_equal(q1, qw)
