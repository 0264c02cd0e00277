"""tests/golden/tree.npz (node table + STree::find results produced by the reference's own STree,
tools/make_golden.py): CPU check of the host-side flattening (a numpy walk over the flattened table
must reproduce the reference's find), GPU check of the device walk."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tree.npz")


def _load():
    g = dict(np.load(GOLD))
    from sdmm_mitsuba_b200 import tree
    nodes = g["nodes"]
    flat, leaf_slot = tree.flatten(nodes[:, 0] > 0, nodes[:, 1], nodes[:, 2], nodes[:, 3], nodes[:, 5:8])
    return g, flat, leaf_slot


def _walk(flat, q):
    """numpy restatement of STree::find (spatial_tree.h:19-51) over the flattened table."""
    n = q.shape[1]
    node = np.zeros(n, np.int64)
    for _ in range(96):
        ax = flat["axis"][node]
        leaf = ax < 0
        if leaf.all():
            break
        p = q[np.clip(ax, 0, 2), np.arange(n)]
        nxt = np.where(flat["split"][node] < p, flat["child0"][node], flat["child1"][node])
        node = np.where(leaf, node, nxt)
    return flat["child0"][node].astype(np.int64)


def test_flattened_tree_reproduces_reference_find():
    g, flat, leaf_slot = _load()
    for pts, ref in ((g["queries"], g["queries_node"]), (g["points"], g["points_node"])):
        assert (ref >= 0).all()
        np.testing.assert_array_equal(_walk(flat, pts), leaf_slot[ref])
    # every leaf got exactly one slot, slots are dense
    leaves = g["nodes"][:, 0] > 0
    assert sorted(leaf_slot[leaves].tolist()) == list(range(int(leaves.sum())))
    # the children of an inner node tile the parent's box along `axis` at the stored split
    nodes = g["nodes"]
    for i in np.nonzero(~leaves)[0]:
        ax, c0, c1 = int(nodes[i, 1]), int(nodes[i, 2]), int(nodes[i, 3])
        assert flat["split"][i] == nodes[c0, 5 + ax]
        assert abs(nodes[c1, 8 + ax] - nodes[c0, 5 + ax]) <= 1e-6 * max(1.0, abs(nodes[c0, 5 + ax]))


@pytest.mark.gpu
def test_gpu_find_matches_golden():
    from sdmm_mitsuba_b200.mixture import SDMMCells
    g, flat, leaf_slot = _load()
    cells = SDMMCells(int((g["nodes"][:, 0] > 0).sum()), 16)
    cells.set_tree(flat)
    for pts, ref in ((g["queries"], g["queries_node"]), (g["points"], g["points_node"])):
        np.testing.assert_array_equal(cells.find_cells(pts).astype(np.int64), leaf_slot[ref])
