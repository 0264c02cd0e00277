"""`.asdmm` checkpoints (sdmm_mitsuba_b200/checkpoint.py) against the reference's own JSON serializers:
oracle/_ref/libsdmm_ref_ckpt.so (where it was built) and a golden file the reference wrote
(tests/golden/ref_tree.asdmm.gz, made by tools/make_golden.py --checkpoint)."""
import gzip
import json
import os

import numpy as np
import pytest

from sdmm_mitsuba_b200 import checkpoint, synth
from sdmm_mitsuba_b200.tree import HostTree

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_tree.asdmm.gz")
AABB = ((0.0, -1.0, 0.5), (1.0, 2.0, 3.0))
K = 8


def _models(n, K, seed=4):
    joint, _, _ = synth.make_initial_model(n, K, seed=seed)
    rng = np.random.default_rng(seed)
    # non-trivial SPD covariances and weights that are already normalised (the reference normalises on set)
    for c in range(n):
        A = rng.normal(0, 0.05, (5, 5, K)).astype(np.float32)
        cov = np.einsum("ijk,ljk->ilk", A, A) + 0.01 * np.eye(5, dtype=np.float32)[:, :, None]
        joint[c, 7:32] = cov.reshape(25, K)
        w = rng.uniform(0.2, 1.0, K).astype(np.float32)
        joint[c, 0] = w / w.sum(dtype=np.float32)
    return joint


def _host_tree(depth=1, max_cells=64):
    t = HostTree(AABB[0], AABB[1], max_cells)
    t.split_to_depth(depth)
    return t


def _ckpt():
    from oracle import ckptbind
    if not ckptbind.available():
        pytest.skip("oracle/_ref/libsdmm_ref_ckpt.so not built (needs /root/reference)")
    return ckptbind


def _same_structure(a, b, path="$"):
    """same keys / list lengths / scalar kinds all the way down"""
    if isinstance(a, dict):
        assert isinstance(b, dict) and sorted(a) == sorted(b), path
        for k in a:
            _same_structure(a[k], b[k], path + "." + k)
    elif isinstance(a, list):
        assert isinstance(b, list) and len(a) == len(b), path
        for i, (x, y) in enumerate(zip(a, b)):
            _same_structure(x, y, "%s[%d]" % (path, i))
    else:
        assert not isinstance(b, (dict, list)), path


def test_roundtrip_bit_exact(tmp_path):
    tree = _host_tree(1)
    joint = _models(tree.n_leaves, K)
    em = np.random.default_rng(1).normal(size=(tree.n_leaves, 81, K)).astype(np.float32)
    scal = np.arange(tree.n_leaves * 8, dtype=np.uint32).reshape(-1, 8)
    p = str(tmp_path / "model_00001.asdmm")
    checkpoint.save_asdmm(p, tree, joint, em_by_slot=em, scal_by_slot=scal, pad_nodes=0)
    t2, j2, ex = checkpoint.load_asdmm(p)
    assert t2.size() == tree.size() and t2.n_leaves == tree.n_leaves
    assert t2.is_leaf == tree.is_leaf and t2.axis == tree.axis and t2.children == tree.children
    assert t2.depth == tree.depth
    for i in range(tree.size()):
        np.testing.assert_array_equal(t2.amin[i], tree.amin[i])
        np.testing.assert_array_equal(t2.amax[i], tree.amax[i])
    # slots are renumbered depth-first on load: compare leaf by leaf through the node index
    for i in range(tree.size()):
        if tree.is_leaf[i]:
            np.testing.assert_array_equal(j2[t2.slot[i]], joint[tree.slot[i]])
            np.testing.assert_array_equal(ex["em"][t2.slot[i]], em[tree.slot[i]])
            np.testing.assert_array_equal(ex["scalars"][t2.slot[i]], scal[tree.slot[i]])
    assert ex["empty_slots"] == [] and ex["K"] == K


def test_frames_match_reference_kat():
    """frames_from_mean_dir vs the reference's CoordinateSystem::prepare through the plain-C restatement's frame
    (itself pinned to the reference in tests/test_oracle_kat.py)."""
    from oracle import portbind
    rng = np.random.default_rng(0)
    n = rng.normal(size=(3, 64)).astype(np.float32)
    n /= np.linalg.norm(n, axis=0, keepdims=True)
    n[:, 0] = (0, 0, 1)
    n[:, 1] = (0, 0, -1)
    to, frm = checkpoint.frames_from_mean_dir(n)
    for k in range(n.shape[1]):
        ref_to = np.asarray(portbind.frame(n[:, k])).reshape(-1)[:9]
        np.testing.assert_allclose(to[:, k], ref_to, atol=2e-7)
    np.testing.assert_array_equal(to.reshape(3, 3, -1), frm.reshape(3, 3, -1).transpose(1, 0, 2))


def test_reference_reads_our_file(tmp_path):
    ck = _ckpt()
    tree = _host_tree(1)
    joint = _models(tree.n_leaves, K)
    p = str(tmp_path / "ours.asdmm")
    checkpoint.save_asdmm(p, tree, joint)
    rt = ck.RefCheckpointTree(AABB[0], AABB[1], 16)
    rt.load(p)
    nodes = rt.stored_nodes(tree.size())
    for i in range(tree.size()):
        assert bool(nodes[i, 0]) == tree.is_leaf[i]
        assert int(nodes[i, 11]) == i and int(nodes[i, 12]) == tree.depth[i]
        if not tree.is_leaf[i]:
            assert int(nodes[i, 1]) == tree.axis[i] and (int(nodes[i, 2]), int(nodes[i, 3])) == tree.children[i]
        np.testing.assert_array_equal(nodes[i, 5:8].astype(np.float32), tree.amin[i])
        np.testing.assert_array_equal(nodes[i, 8:11].astype(np.float32), tree.amax[i])
        if tree.is_leaf[i]:
            got, frame = rt.leaf_model(i)
            want = joint[tree.slot[i]]
            np.testing.assert_allclose(got[0], want[0], rtol=2e-6)       # from_json re-normalises (sdmm.h:157-160)
            np.testing.assert_array_equal(got[1:], want[1:])
            to, frm = checkpoint.frames_from_mean_dir(want[4:7])
            np.testing.assert_allclose(frame[:9], to, atol=5e-7)          # set_mean recomputed them
            np.testing.assert_allclose(frame[9:], frm, atol=5e-7)
    # and the reference's find walks the loaded table like ours
    rng = np.random.default_rng(2)
    pos = (np.asarray(AABB[0], np.float32)[:, None] +
           rng.uniform(0.01, 0.99, (3, 500)).astype(np.float32) * (np.asarray(AABB[1]) - np.asarray(AABB[0]))[:, None]).astype(np.float32)
    found = rt.find(pos)
    assert (found >= 0).all() and all(tree.is_leaf[i] for i in found)


def test_non_finite_entries_do_not_break_the_file(tmp_path):
    """One NaN / inf in one mixture (a cell flagged ST_NON_PSD, an aborted fit) must not make the whole .asdmm
    unreadable: nlohmann::json would write null and then refuse to read it back as a float, Python's json would
    write a bare NaN token.  The writer switches the broken component off instead (weight 0, identity covariance);
    the reference's own from_json parses the file and every other component is untouched."""
    ck = _ckpt()
    tree = _host_tree(1)
    joint = _models(tree.n_leaves, K)
    good = joint.copy()
    joint[0, 7, 1] = np.nan
    joint[1, 2, 0] = np.inf
    p = str(tmp_path / "nonfinite.asdmm")
    checkpoint.save_asdmm(p, tree, joint)
    text = open(p).read()
    assert "NaN" not in text and "Infinity" not in text
    rt = ck.RefCheckpointTree(AABB[0], AABB[1], 16)
    rt.load(p)                                              # the reference's own from_json parses it
    nodes = rt.stored_nodes(tree.size())
    assert all(bool(nodes[i, 0]) == tree.is_leaf[i] for i in range(tree.size()))
    _, j, _ = checkpoint.load_asdmm(p)
    assert np.isfinite(j).all()
    assert (j[:, 0, :] == 0).sum() == 2                     # the two broken components are switched off
    assert (good[:, 0, :] == 0).sum() == 0


def test_we_read_the_references_file(tmp_path):
    ck = _ckpt()
    rt = ck.RefCheckpointTree(AABB[0], AABB[1], 16)
    rt.split_to_depth(1)
    nodes = rt.stored_nodes(rt.size())
    leaves = [i for i in range(rt.size()) if nodes[i, 4]]
    joint = _models(len(leaves), K)
    for c, i in enumerate(leaves):
        assert rt.set_leaf_model(i, joint[c]) == 1
    p = str(tmp_path / "ref.asdmm")
    rt.save(p)
    tree, j2, ex = checkpoint.load_asdmm(p)
    assert tree.size() == rt.size() and tree.n_leaves == len(leaves)
    for c, i in enumerate(leaves):
        got_ref, _ = rt.leaf_model(i)
        np.testing.assert_array_equal(j2[tree.slot[i]], got_ref)
        np.testing.assert_allclose(j2[tree.slot[i]][0], joint[c][0], rtol=2e-6)
        np.testing.assert_array_equal(j2[tree.slot[i]][1:], joint[c][1:])
    # our writer produces the same document shape as the reference's, value for value
    p2 = str(tmp_path / "ours.asdmm")
    checkpoint.save_asdmm(p2, tree, j2)
    a, b = json.load(open(p)), json.load(open(p2))
    assert len(a["m_nodes"]) == len(b["m_nodes"]) == checkpoint.REFERENCE_NODE_COUNT
    a["m_nodes"], b["m_nodes"] = a["m_nodes"][:rt.size()], b["m_nodes"][:rt.size()]
    _same_structure(a, b)
    for na, nb in zip(a["m_nodes"], b["m_nodes"]):
        for key in ("is_leaf", "idx", "depth", "aabb"):
            assert na[key] == nb[key]
        if na["is_leaf"]:
            sa, sb = na["value"]["sdmm"], nb["value"]["sdmm"]
            assert sa["weight.pmf"] == sb["weight.pmf"] and sa["cov"] == sb["cov"]
            assert sa["tangent_space"]["tangent_space.mean"] == sb["tangent_space"]["tangent_space.mean"]
            for which in ("coordinate_system.to", "coordinate_system.from"):
                fa = np.asarray([pl["data"] for pl in sa["tangent_space"]["tangent_space.coordinate_system"][which]["data"]])
                fb = np.asarray([pl["data"] for pl in sb["tangent_space"]["tangent_space.coordinate_system"][which]["data"]])
                np.testing.assert_allclose(fa, fb, atol=5e-7)
        else:
            assert na["axis"] == nb["axis"] and na["children"] == nb["children"]


def test_golden_reference_file():
    """A checkpoint the reference itself wrote (committed, gzip): travels to boxes without /root/reference."""
    tree, joint, ex = checkpoint.load_asdmm(GOLDEN)
    meta = json.load(open(GOLDEN.replace(".asdmm.gz", ".meta.json")))
    assert tree.size() == meta["nodes"] and tree.n_leaves == meta["leaves"] and ex["K"] == meta["K"]
    want = _models(meta["leaves"], meta["K"], seed=meta["seed"])
    leaves = [i for i in range(tree.size()) if tree.is_leaf[i]]
    for c, i in enumerate(leaves):
        np.testing.assert_allclose(joint[tree.slot[i]][0], want[c][0], rtol=2e-6)
        np.testing.assert_array_equal(joint[tree.slot[i]][1:], want[c][1:])
    with gzip.open(GOLDEN, "rt") as f:
        doc = json.load(f)
    assert len(doc["m_nodes"]) == checkpoint.REFERENCE_NODE_COUNT
    assert doc["m_aabb"] == {"max": {"data": list(AABB[1])}, "min": {"data": list(AABB[0])}}


@pytest.mark.gpu
def test_device_fit_survives_a_checkpoint(tmp_path):
    """fit on the device -> download -> .asdmm -> load -> upload into a fresh context: identical query records."""
    import torch
    from sdmm_mitsuba_b200.mixture import SDMMCells
    tree = _host_tree(1)
    C, Kd, N = tree.n_leaves, 16, 600
    point6, weight, off = synth.make_samples(C, N, seed=3)
    joint, em, scal = synth.make_initial_model(C, Kd, seed=4)
    cells = SDMMCells(C, Kd)
    cells.upload(joint, em, scal)
    cells.set_samples(point6, weight, off)
    cells.em_step(3, with_conditioner=True)
    fitted, fem, fscal = cells.download()
    p = str(tmp_path / "model_00003.asdmm")
    checkpoint.save_asdmm(p, tree, fitted, em_by_slot=fem, scal_by_slot=fscal)
    t2, j2, ex = checkpoint.load_asdmm(p)
    perm = [tree.slot[i] for i in range(tree.size()) if tree.is_leaf[i]]   # file order -> original slots
    new_slot = [t2.slot[i] for i in range(t2.size()) if t2.is_leaf[i]]
    cells2 = SDMMCells(C, Kd)
    em2 = np.stack([ex["em"][s] for s in range(C)])
    sc2 = np.stack([ex["scalars"][s] for s in range(C)])
    cells2.upload(j2, em2, sc2)
    cells2.prepare(0, C)
    torch.cuda.synchronize()
    r1, r2 = cells.query_records(), cells2.query_records()
    for old, new in zip(perm, new_slot):
        # upload normalises the (already normalised) weights once more, like the reference's from_json -> prepare
        assert torch.allclose(r1[old], r2[new], rtol=2e-5, atol=2e-6)
