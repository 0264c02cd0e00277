"""The data side of optimize() over the renderer's sample pool -- sdmmb_pt_pool_stats (cell of every pooled sample,
Data::size, SpatialStats of the main pushes: sdmm_proc.cpp:880-895, volpath_sdmm.cpp:140-149) and sdmmb_pt_rebucket (one
recursion level of split_leaf_recurse / create_child, spatial_tree.h:169-214) -- against plain numpy on random pools."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

PLANES = 15


def _setup(n, cap, seed, depth=3):
    import torch
    from sdmm_mitsuba_b200.mixture import SDMMCells
    from sdmm_mitsuba_b200.tree import HostTree
    rng = np.random.default_rng(seed)
    t = HostTree(np.zeros(3, np.float32), np.ones(3, np.float32), 4096)
    t.split_to_depth(depth)
    cells = SDMMCells(4096, 16)
    cells.set_tree(t.flatten())
    pool = np.zeros((PLANES, cap), np.float32)
    pool[0:3, :n] = rng.uniform(0.02, 0.98, (3, n))
    pool[10:13, :n] = np.clip(pool[0:3, :n] + rng.normal(0, 0.05, (3, n)), 0.001, 0.999)      # keys: jittered copies
    pool[13, :n] = (rng.uniform(size=n) < 0.6) + 2.0 * rng.integers(0, 5, n)                   # is_main + 2 * pass
    dev = torch.device("cuda", 0)
    return t, cells, pool, torch.from_numpy(pool).to(dev), rng


def _host_find(t, p):
    return np.array([t.slot[t.find(p[:, i])] for i in range(p.shape[1])], np.int64)


def test_pool_stats_matches_numpy():
    import torch
    n, cap, lo = 50000, 60000, 12345
    t, cells, pool, pool_d, _ = _setup(n, cap, seed=1)
    cell = torch.full((n,), -7, dtype=torch.int32, device=pool_d.device)
    stats = torch.full((8, 4096), 99.0, dtype=torch.float64, device=pool_d.device)      # the call zeroes it
    rc = cells._L.sdmmb_pt_pool_stats(cells._h, C.c_void_p(pool_d.data_ptr()), cap, n, lo, C.c_void_p(cell.data_ptr()),
                                      C.c_void_p(stats.data_ptr()), None)
    assert rc == 0
    torch.cuda.synchronize()
    want = _host_find(t, pool[10:13, :n])
    np.testing.assert_array_equal(cell.cpu().numpy(), want)
    s = stats.cpu().numpy()
    np.testing.assert_array_equal(s[7], np.bincount(want, minlength=4096))                      # Data::size
    main = (np.arange(n) >= lo) & ((pool[13, :n].astype(np.int64) & 1) == 1)
    np.testing.assert_array_equal(s[0], np.bincount(want[main], minlength=4096))                # SpatialStats size
    p = pool[0:3, :n].astype(np.float64)
    for k in range(3):
        np.testing.assert_allclose(s[1 + k], np.bincount(want[main], weights=p[k, main], minlength=4096), rtol=1e-12, atol=1e-12)
        np.testing.assert_allclose(s[4 + k], np.bincount(want[main], weights=p[k, main] ** 2, minlength=4096), rtol=1e-12, atol=1e-12)
    # n = 0: only the zeroing
    rc = cells._L.sdmmb_pt_pool_stats(cells._h, C.c_void_p(pool_d.data_ptr()), cap, 0, 0, None, C.c_void_p(stats.data_ptr()), None)
    torch.cuda.synchronize()
    assert rc == 0 and float(stats.abs().sum()) == 0.0
    assert cells._L.sdmmb_pt_pool_stats(cells._h, C.c_void_p(pool_d.data_ptr()), cap, cap + 1, 0, C.c_void_p(cell.data_ptr()),
                                        C.c_void_p(stats.data_ptr()), None) == -1          # n beyond the capacity


def test_rebucket_level_matches_numpy():
    import torch
    n, cap = 40000, 40000
    t, cells, pool, pool_d, rng = _setup(n, cap, seed=2, depth=2)
    NC = 4096
    cell0 = _host_find(t, pool[0:3, :n]).astype(np.int32)      # samples sit in the leaf of their position here
    cell0[rng.uniform(size=n) < 0.01] = -1                      # a few without a cell
    # split a third of the leaves at a random location along a random axis
    leaves = [i for i in t.leaves() if rng.uniform() < 0.33]
    tab_i = np.full((3, NC), -1, np.int32)
    tab_f = np.zeros((2, NC), np.float32)
    box = np.zeros((2, 3, NC), np.float32)
    free = 3000
    for node in leaves:
        ps, ax = t.slot[node], int(rng.integers(0, 3))
        lo, hi = t.amin[node].copy(), t.amax[node].copy()
        c0_min = np.float32(lo[ax] + np.float32(rng.uniform(0.3, 0.7)) * (hi[ax] - lo[ax]))
        c1_max = np.float32(c0_min - np.float32(0.01) if rng.uniform() < 0.5 else c0_min)   # a (wide) gap between the boxes for some
        s0, s1 = ps, free
        free += 1
        tab_i[:, ps] = (ax, s0, s1)
        tab_f[:, ps] = (c0_min, c1_max)
        a0, b1 = lo.copy(), hi.copy()
        a0[ax], b1[ax] = c0_min, c1_max
        box[0, :, s0], box[1, :, s0] = a0, hi
        box[0, :, s1], box[1, :, s1] = lo, b1
    dev = pool_d.device
    cell_d = torch.from_numpy(cell0.copy()).to(dev)
    stats = torch.empty((7, NC), dtype=torch.float64, device=dev)
    tab_i_d, tab_f_d, box_d = torch.from_numpy(tab_i).to(dev), torch.from_numpy(tab_f).to(dev), torch.from_numpy(box).to(dev)
    rc = cells._L.sdmmb_pt_rebucket(cells._h, C.c_void_p(pool_d.data_ptr()), cap, n, C.c_void_p(cell_d.data_ptr()),
                                    C.c_void_p(tab_i_d.data_ptr()), C.c_void_p(tab_f_d.data_ptr()), C.c_void_p(box_d.data_ptr()),
                                    C.c_void_p(stats.data_ptr()), None)
    assert rc == 0
    torch.cuda.synchronize()
    # numpy: create_child's rule (child 0 takes min < p, child 1 takes max >= p, else nobody) + AABB::contains
    p = pool[0:3, :n]
    want = cell0.copy()
    ok = (cell0 >= 0) & (tab_i[0][np.maximum(cell0, 0)] >= 0)
    idx = np.nonzero(ok)[0]
    c = cell0[idx]
    pa = p[tab_i[0][c], idx]
    want[idx] = np.where(pa > tab_f[0][c], tab_i[1][c], np.where(pa <= tab_f[1][c], tab_i[2][c], -1))
    np.testing.assert_array_equal(cell_d.cpu().numpy(), want)
    assert (want[idx] == -1).any(), "the rounding-gap case must occur"
    moved = idx[want[idx] >= 0]
    nc = want[moved]
    inside = ((p[:, moved] >= box[0][:, nc]) & (p[:, moved] <= box[1][:, nc])).all(axis=0)
    ci, q = nc[inside], p[:, moved][:, inside].astype(np.float64)
    s = stats.cpu().numpy()
    np.testing.assert_array_equal(s[0], np.bincount(ci, minlength=NC))
    for k in range(3):
        np.testing.assert_allclose(s[1 + k], np.bincount(ci, weights=q[k], minlength=NC), rtol=1e-12, atol=1e-12)
        np.testing.assert_allclose(s[4 + k], np.bincount(ci, weights=q[k] ** 2, minlength=NC), rtol=1e-12, atol=1e-12)
