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
