"""Host side of the spatial cell structure (reference: sdmm/include/sdmm/accelerators/spatial_tree.h).
The tree is built and split by the host (the reference's own serial code in the plugin); this module
flattens its node table into the device records of include/sdmm_b200.h (sdmmb_tree_node) and assigns
device cell slots to the leaves."""
import numpy as np

TREE_NODE_DTYPE = np.dtype([("axis", "<i4"), ("split", "<f4"), ("child0", "<u4"), ("child1", "<u4")])


def flatten(is_leaf, axis, child0, child1, aabb_min, leaf_slot=None):
    """is_leaf/axis/child0/child1: [n_nodes]; aabb_min: [n_nodes, 3] (STreeNode::aabb.min).
    leaf_slot: optional [n_nodes] device slot per leaf node (default: leaves numbered in node order).
    Returns (nodes: structured array of sdmmb_tree_node, leaf_slot: [n_nodes] int64, -1 for inner nodes).

    find() at an inner node takes children[0] iff nodes[children[0]].aabb.min[axis] < p[axis]
    (spatial_tree.h:36-45), so the only number an inner node needs is that coordinate."""
    is_leaf = np.asarray(is_leaf).astype(bool)
    n = is_leaf.shape[0]
    axis = np.asarray(axis).astype(np.int64)
    child0 = np.asarray(child0).astype(np.int64)
    child1 = np.asarray(child1).astype(np.int64)
    aabb_min = np.asarray(aabb_min, np.float32).reshape(n, 3)
    if leaf_slot is None:
        leaf_slot = np.full(n, -1, np.int64)
        leaf_slot[is_leaf] = np.arange(int(is_leaf.sum()))
    leaf_slot = np.asarray(leaf_slot, np.int64)
    nodes = np.zeros(n, TREE_NODE_DTYPE)
    inner = ~is_leaf
    nodes["axis"][is_leaf] = -1
    nodes["child0"][is_leaf] = leaf_slot[is_leaf]
    nodes["axis"][inner] = axis[inner]
    nodes["child0"][inner] = child0[inner]
    nodes["child1"][inner] = child1[inner]
    nodes["split"][inner] = aabb_min[child0[inner], axis[inner]]
    return nodes, leaf_slot


_ZERO3 = np.zeros(3)
_ZERO3.flags.writeable = False


class HostTree:
    """Host restatement of the reference's STree bookkeeping (sdmm/include/sdmm/accelerators/spatial_tree.h):
    node table, split_to_depth (:216-252), data-driven split_leaf_recurse (:254-292) with the split position at
    the mean of the leaf's SpatialStats along the axis of largest variance (get_split_location, :133-167).
    Leaves own a device cell slot; the samples themselves stay on the device (they are re-bucketed there with
    find), so a split only needs the children's SpatialStats, which the caller supplies through `child_stats`.
    The reference's limits (10000 nodes, uint16 indices, :96, 311-312) are lifted."""

    def __init__(self, aabb_min, aabb_max, max_cells):
        self.is_leaf = [True]
        self.axis = [0]
        self.children = [(0, 0)]
        self.amin = [np.asarray(aabb_min, np.float32).copy()]
        self.amax = [np.asarray(aabb_max, np.float32).copy()]
        self.depth = [0]
        self.slot = [0]
        # SpatialStats (accelerators/spatial_stats.h): size, sum of points, sum of squared points (double)
        self.st_size = [0]
        self.st_sum = [np.zeros(3)]
        self.st_sqr = [np.zeros(3)]
        self.max_cells = int(max_cells)
        self.free_slots = list(range(self.max_cells - 1, 0, -1))
        self.n_leaves = 1

    def size(self):
        return len(self.is_leaf)

    def leaves(self):
        return [i for i in range(self.size()) if self.is_leaf[i]]

    def _create_child(self, node, child_i, loc):
        """create_child (:169-214): child 0 is the UPPER part (min raised), child 1 the lower part."""
        ax = self.axis[node]
        amin, amax = self.amin[node].copy(), self.amax[node].copy()
        diag = np.float32(amax[ax] - amin[ax])
        if child_i == 0:
            amin[ax] = np.float32(amin[ax] + np.float32(loc) * diag)
        else:
            amax[ax] = np.float32(amax[ax] - np.float32(np.float32(1.0) - np.float32(loc)) * diag)
        self.is_leaf.append(True)
        self.axis.append((ax + 1) % 3)
        self.children.append((0, 0))
        self.amin.append(amin)
        self.amax.append(amax)
        self.depth.append(self.depth[node] + 1)
        self.slot.append(-1)
        self.st_size.append(0)
        self.st_sum.append(np.zeros(3))
        self.st_sqr.append(np.zeros(3))
        return self.size() - 1

    def _split(self, node, loc):
        """Returns (child0, child1); child 0 keeps the parent's slot, child 1 takes a fresh one."""
        if not self.free_slots:
            return None
        self.is_leaf[node] = False
        c0 = self._create_child(node, 0, loc)
        c1 = self._create_child(node, 1, loc)
        self.children[node] = (c0, c1)
        self.slot[c0] = self.slot[node]
        self.slot[c1] = self.free_slots.pop()
        self.slot[node] = -1
        self.n_leaves += 1
        return c0, c1

    def split_to_depth(self, max_depth):
        def rec(node, depth):
            max_axis = 3
            next_depth = depth + 1 if self.axis[node] == max_axis - 1 else depth
            if not self.is_leaf[node]:
                for c in self.children[node]:
                    rec(c, next_depth)
                return
            if depth >= max_depth:
                return
            ch = self._split(node, 0.5)
            if ch is None:
                return
            for c in ch:
                rec(c, next_depth)
        rec(0, 0)

    def split_location(self, node):
        """get_split_location (:133-167), in double like SpatialStats."""
        n = float(self.st_size[node])
        mean, sqr = self.st_sum[node], self.st_sqr[node]
        var = (sqr - (mean / n) ** 2 * n) / (n - 1.0)
        ax = 0
        for i in range(3):
            if var[i] > var[ax]:
                ax = i
        loc = np.float32((mean[ax] / n - float(self.amin[node][ax])) / float(self.amax[node][ax] - self.amin[node][ax]))
        return ax, loc

    def split_leaves(self, threshold, child_stats, on_split=None):
        """split_leaf_recurse over every node (volpath_sdmm.cpp:253-259).  child_stats(parent_slot, amin, amax) ->
        (size, sum[3], sqr[3]) of the parent's CURRENT data samples inside the child's box (create_child :204-206).
        on_split(parent_slot, slot0, slot1, axis, child0_min, child1_max) is called right after a split so the caller
        can re-bucket the parent's data the way create_child does (:190-203: by the sample POSITION along the split axis).
        Returns the list of (parent_slot, child0_slot, child1_slot) splits performed, in order."""
        done = []

        def rec(node):
            if not self.is_leaf[node] or self.st_size[node] <= threshold:
                return
            ax, loc = self.split_location(node)
            if not (0.0 < loc < 1.0):
                return
            parent_slot = self.slot[node]
            self.axis[node] = ax
            ch = self._split(node, loc)
            if ch is None:
                return
            if on_split is not None:
                on_split(parent_slot, self.slot[ch[0]], self.slot[ch[1]], ax, float(self.amin[ch[0]][ax]),
                         float(self.amax[ch[1]][ax]))
            for c in ch:
                self.st_size[c], self.st_sum[c], self.st_sqr[c] = child_stats(self.slot[c], self.amin[c], self.amax[c])
            self.st_size[node] = 0
            done.append((parent_slot, self.slot[ch[0]], self.slot[ch[1]]))
            for c in ch:
                rec(c)

        for node in range(self.size()):
            rec(node)
        return done

    def split_leaves_levels(self, threshold, rebucket):
        """split_leaves with the data work batched per recursion level (the device does it for all leaves of a level
        at once), then the nodes and slots created in this pass are renumbered into the order the reference's
        depth-first split_leaf_recurse would have created them in, so the node table is identical.
        rebucket(level) with level = list of dicts (parent_slot, slot0, slot1, axis, c0_min, c1_max, amin0, amax0,
        amin1, amax1) re-buckets the parents' samples (create_child, :190-203) and returns, per entry, the two
        children's SpatialStats ((size, sum, sqr), (size, sum, sqr)) over the samples inside their boxes (:204-206).
        Returns (splits, slot_map): splits = [(parent_slot, child0_slot, child1_slot)] in depth-first order with
        the FINAL slots; slot_map = {temporary slot used during the pass: final slot}."""
        S0 = self.size()
        created, created_slots, popped = {}, {}, []
        frontier = [n for n in range(S0) if self.is_leaf[n] and self.st_size[n] > threshold]
        f32, f64 = np.float32, np.float64
        while frontier:
            # One recursion level.  The arithmetic of split_location() and _create_child() for all candidates at once
            # (same float32 / float64 operations, element by element); only the bookkeeping stays a Python loop.
            cand = [n for n in frontier if self.st_size[n] > threshold]
            level, nodes = [], []
            if cand:
                m = len(cand)
                rows = np.arange(m)
                cnt = np.array([self.st_size[n] for n in cand], f64)[:, None]
                mean = np.array([self.st_sum[n] for n in cand], f64).reshape(m, 3)
                sqr = np.array([self.st_sqr[n] for n in cand], f64).reshape(m, 3)
                lo = np.array([self.amin[n] for n in cand], f32).reshape(m, 3)
                hi = np.array([self.amax[n] for n in cand], f32).reshape(m, 3)
                with np.errstate(all="ignore"):
                    var = (sqr - (mean / cnt) ** 2 * cnt) / (cnt - 1.0)
                    ax = np.zeros(m, np.int64)
                    for i in (1, 2):
                        ax = np.where(var[:, i] > var[rows, ax], i, ax)
                    lo_ax, hi_ax = lo[rows, ax], hi[rows, ax]
                    loc = ((mean[rows, ax] / cnt[:, 0] - lo_ax.astype(f64)) / (hi_ax - lo_ax).astype(f64)).astype(f32)
                    ok = (loc > 0.0) & (loc < 1.0)
                    diag = hi_ax - lo_ax                                               # float32
                    c0_min = lo_ax + loc * diag                                        # child 0: min raised
                    c1_max = hi_ax - (f32(1.0) - loc) * diag                           # child 1: max lowered
                lo0, hi1 = lo.copy(), hi.copy()
                lo0[rows, ax] = c0_min
                hi1[rows, ax] = c1_max
                for r, node in enumerate(cand):
                    if not ok[r]:
                        continue
                    a_ = int(ax[r])
                    self.axis[node] = a_
                    if not self.free_slots:       # out of cells: the leaf stays (its axis is already the split's)
                        continue
                    parent_slot = self.slot[node]
                    c0 = len(self.is_leaf)
                    c1 = c0 + 1
                    s1 = self.free_slots.pop()
                    d1 = self.depth[node] + 1
                    self.is_leaf[node] = False
                    self.children[node] = (c0, c1)
                    self.slot[node] = -1
                    self.is_leaf += (True, True)
                    self.axis += ((a_ + 1) % 3, (a_ + 1) % 3)
                    self.children += ((0, 0), (0, 0))
                    self.amin += (lo0[r], lo[r])
                    self.amax += (hi[r], hi1[r])
                    self.depth += (d1, d1)
                    self.slot += (parent_slot, s1)
                    self.st_size += (0, 0)
                    self.st_sum += (_ZERO3, _ZERO3)
                    self.st_sqr += (_ZERO3, _ZERO3)
                    self.n_leaves += 1
                    ch = (c0, c1)
                    popped.append(s1)
                    created[node] = ch
                    created_slots[node] = (parent_slot, s1)   # a later split clears slot[child]
                    nodes.append((node, ch))
                    level.append(dict(parent_slot=parent_slot, slot0=parent_slot, slot1=s1, axis=a_,
                                      c0_min=float(c0_min[r]), c1_max=float(c1_max[r]),
                                      amin0=lo0[r], amax0=hi[r], amin1=lo[r], amax1=hi1[r]))
            if not level:
                break
            stats = rebucket(level)
            frontier = []
            for (node, ch), st in zip(nodes, stats):
                for c, (size, ssum, ssqr) in zip(ch, st):
                    self.st_size[c], self.st_sum[c], self.st_sqr[c] = int(size), ssum, ssqr
                    frontier.append(c)
                self.st_size[node] = 0
        if not created:
            return [], {}
        # depth-first creation order of the new nodes and of the new slots
        order, split_seq = [], []

        def visit(node):
            if node in created:
                c0, c1 = created[node]
                order.extend((c0, c1))
                split_seq.append(node)
                visit(c0)
                visit(c1)

        for node in range(S0):
            visit(node)
        new_id = {b: S0 + i for i, b in enumerate(order)}
        ident = lambda b: new_id.get(b, b)
        slot_map = {created_slots[node][1]: popped[j] for j, node in enumerate(split_seq)}
        smap = lambda sl: slot_map.get(sl, sl)
        splits = []
        for node in split_seq:     # child 0 carries the slot its parent had
            s0, s1 = created_slots[node]
            splits.append((smap(s0), smap(s0), smap(s1)))
        perm = list(range(S0)) + order
        for name in ("is_leaf", "axis", "children", "amin", "amax", "depth", "slot", "st_size", "st_sum", "st_sqr"):
            old = getattr(self, name)
            setattr(self, name, [old[b] for b in perm])
        self.children = [(ident(a), ident(b)) for a, b in self.children]
        self.slot = [smap(sl) if sl >= 0 else sl for sl in self.slot]
        return splits, slot_map

    def find(self, p):
        node = 0
        while not self.is_leaf[node]:
            c0, c1 = self.children[node]
            ax = self.axis[node]
            node = c0 if self.amin[c0][ax] < p[ax] else c1
        return node

    def flatten(self):
        n = self.size()
        ch = np.array(self.children, np.int64).reshape(n, 2)
        nodes, _ = flatten(np.array(self.is_leaf), np.array(self.axis), ch[:, 0], ch[:, 1], np.array(self.amin),
                           leaf_slot=np.array(self.slot, np.int64))
        return nodes
