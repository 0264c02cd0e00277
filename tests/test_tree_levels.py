"""HostTree.split_leaves_levels (data work batched per recursion level, then renumbered) must build exactly the node
table, slots and statistics of the depth-first split_leaves that restates the reference's split_leaf_recurse."""
import numpy as np

from sdmm_mitsuba_b200.tree import HostTree


def _run(levels, seed, n=60000, threshold=900):
    rng = np.random.default_rng(seed)
    pos = np.concatenate([rng.random((3, n // 2)) * 0.3 + 0.1, rng.random((3, n - n // 2))], axis=1).astype(np.float32)
    t = HostTree(np.zeros(3, np.float32) - 1e-5, np.ones(3, np.float32) + 1e-5, 4096)
    t.split_to_depth(1)
    cell = np.array([t.slot[t.find(pos[:, i])] for i in range(n)], np.int64)
    for node in t.leaves():
        m = cell == t.slot[node]
        q = pos[:, m].astype(np.float64)
        t.st_size[node], t.st_sum[node], t.st_sqr[node] = int(m.sum()), q.sum(axis=1), (q * q).sum(axis=1)

    def stats(slot, amin, amax):
        p = pos[:, cell == slot]
        inside = ((p >= amin[:, None]) & (p <= amax[:, None])).all(axis=0)
        q = p[:, inside].astype(np.float64)
        return int(inside.sum()), q.sum(axis=1), (q * q).sum(axis=1)

    def rebucket_one(parent_slot, slot0, slot1, axis, c0_min, c1_max):
        m = cell == parent_slot
        upper = pos[axis] > np.float32(c0_min)
        lower = pos[axis] <= np.float32(c1_max)
        cell[m & upper] = slot0
        cell[m & ~upper & lower] = slot1
        cell[m & ~upper & ~lower] = -1

    if not levels:
        splits = t.split_leaves(threshold, stats, rebucket_one)
        return t, cell, splits

    def rebucket(level):
        out = []
        for e in level:
            rebucket_one(e["parent_slot"], e["slot0"], e["slot1"], e["axis"], e["c0_min"], e["c1_max"])
        for e in level:
            out.append((stats(e["slot0"], e["amin0"], e["amax0"]), stats(e["slot1"], e["amin1"], e["amax1"])))
        return out

    splits, slot_map = t.split_leaves_levels(threshold, rebucket)
    lut = np.arange(4096)
    for a, b in slot_map.items():
        lut[a] = b
    cell = np.where(cell >= 0, lut[np.maximum(cell, 0)], -1)
    return t, cell, splits


def test_level_synchronous_split_equals_depth_first():
    for seed in (1, 2):
        a, cell_a, splits_a = _run(False, seed)
        b, cell_b, splits_b = _run(True, seed)
        assert a.size() == b.size() > 60
        assert a.is_leaf == b.is_leaf and a.axis == b.axis and a.children == b.children and a.slot == b.slot
        assert a.depth == b.depth and a.st_size == b.st_size and a.n_leaves == b.n_leaves
        np.testing.assert_array_equal(np.array(a.amin), np.array(b.amin))
        np.testing.assert_array_equal(np.array(a.amax), np.array(b.amax))
        np.testing.assert_allclose(np.array(a.st_sum), np.array(b.st_sum), rtol=1e-12)
        np.testing.assert_array_equal(cell_a, cell_b)
        assert splits_a == splits_b
        assert sorted(a.free_slots) == sorted(b.free_slots)
