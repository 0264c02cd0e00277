"""`.asdmm` checkpoints in the reference's JSON layout, so that fitted trees interchange with the reference
integrator and the `pysdmm` tooling (SURVEY §8 f, F3).

The reference writes `checkpoints/model_%05i.asdmm` with `sdmm::save_json(m_accelerator, path)`
(mitsuba/src/integrators/sdmm/volpath_sdmm.cpp:120-130; sdmm/include/sdmm/core/utils.h:130-137:
nlohmann::json, `std::setw(4)`).  Layout, from the serializers:

  file   = {"m_nodes": [node, ...], "m_aabb": box}      NLOHMANN_DEFINE_TYPE_INTRUSIVE(STree, m_nodes, m_aabb)
                                                        (accelerators/spatial_tree.h:314); the reference
                                                        always stores 10000 nodes (:89), unused ones default
  node   = {"is_leaf", "axis", "idx", "children": [c0, c1], "value": null | {"sdmm": mix}, "aabb": box,
            "depth"}                                     (spatial_tree.h:322-349; SDMMContext keeps only
                                                        `sdmm`, distributions/sdmm_context.h:44)
  box    = {"min": {"data": [x, y, z]}, "max": {...}}    (linalg/aabb.h:32, core/utils.h:86-101)
  mix    = {"weight.pmf": dyn, "tangent_space": {"tangent_space.mean": {"data": [dyn x 6]},
            "tangent_space.coordinate_system": {"coordinate_system.to": {"data": [dyn x 9]},
            "coordinate_system.from": {"data": [dyn x 9]}}}, "cov": {"data": [dyn x 25]}}
                                                        (distributions/sdmm.h:138-144,
                                                        spaces/spatio_directional.h:264-272,
                                                        linalg/coordinate_system.h:46-52)
  dyn    = {"size": K, "data": [K floats]}               one enoki DynamicArray = one SoA plane
                                                        (core/utils.h:53-70); matrices list their planes
                                                        ROW-major (core/utils.h:103-124 calls array(c, r)
                                                        with c outer, and Matrix::operator()(i, j) is
                                                        (row, col), linalg/matrix.h:111-116)

which is exactly this package's joint plane order (pmf, mean[6], cov[25] row-major; K floats per plane).
On load the reference recomputes the frames and the Cholesky factors (`set_mean` + `prepare`,
sdmm.h:157-160); so does `SDMMCells.upload`.  The EM statistics are NOT part of the reference's format
(a resumed reference run restarts them); `save_asdmm(..., em=...)` can store them under an extra
"b200.em" key that the reference's loader ignores.

Checked against the reference's own serializers by tests/test_checkpoint.py (oracle/_ref/libsdmm_ref_ckpt.so)
and, where that checker cannot travel, against a golden file the reference wrote (tests/golden/ref_tree.asdmm)."""
import json
import math

import numpy as np

from .tree import HostTree

REFERENCE_NODE_COUNT = 10000   # spatial_tree.h:89


def _num(v):
    """nlohmann::json writes a non-finite float as null and rejects bare NaN / Infinity tokens when it parses
    (Python's json would emit those): one non-finite entry -- a cell flagged ST_NON_PSD, an aborted fit -- must
    not make the whole .asdmm unreadable by the reference.  The loaders map null back to NaN."""
    v = float(v)
    return v if math.isfinite(v) else None


def _dyn(plane):
    return {"data": [_num(v) for v in plane], "size": int(len(plane))}


def _planes(planes):
    return {"data": [_dyn(p) for p in planes]}


def _vec3(v):
    return {"data": [_num(v[0]), _num(v[1]), _num(v[2])]}


def _box(amin, amax):
    return {"max": _vec3(amax), "min": _vec3(amin)}


def frames_from_mean_dir(n):
    """CoordinateSystem::prepare (linalg/coordinate_system.h:21-38, Duff et al. 2017) in float32 for
    n[3][K]: returns (to[9][K], from[9][K]) row-major; `to` = transpose(`from`), columns of `from` = s, t, n."""
    n = np.asarray(n, np.float32)
    nx, ny, nz = n[0], n[1], n[2]
    sign = np.where(np.signbit(nz), np.float32(-1), np.float32(1)).astype(np.float32)
    a = (np.float32(-1) / (sign + nz)).astype(np.float32)
    b = (nx * ny * a).astype(np.float32)
    s = np.stack([sign * (nx * nx * a) + np.float32(1), sign * b, -sign * nx]).astype(np.float32)
    t = np.stack([b, sign + ny * ny * a, -ny]).astype(np.float32)
    frm = np.empty((9,) + nx.shape, np.float32)     # from(r, c): column 0 = s, 1 = t, 2 = n
    for r in range(3):
        frm[3 * r + 0], frm[3 * r + 1], frm[3 * r + 2] = s[r], t[r], n[r]
    to = np.empty_like(frm)
    for r in range(3):
        for c in range(3):
            to[3 * r + c] = frm[3 * c + r]
    return to, frm


def mixture_to_json(joint, K=None):
    """joint: [32, >=K] planes (pmf, mean[6], cov[25] row-major) -> the reference's SDMM object."""
    joint = np.asarray(joint, np.float32)
    K = joint.shape[1] if K is None else int(K)
    j = joint[:, :K]
    # A component with a non-finite entry (a cell flagged ST_NON_PSD, an aborted fit) has no usable parameters,
    # and ONE non-finite number makes the whole file unreadable by the reference: nlohmann::json writes it as
    # null and then refuses to read null as a float.  Such components are written switched off -- weight 0,
    # direction +z, identity covariance: the convention of this package's padding components.
    bad = ~np.isfinite(j).all(axis=0)
    if bad.any():
        j = j.copy()
        j[:, bad] = 0.0
        j[6, bad] = 1.0
        for i in range(5):
            j[7 + 6 * i, bad] = 1.0
    to, frm = frames_from_mean_dir(j[4:7])
    return {
        "cov": _planes(j[7:32]),
        "tangent_space": {
            "tangent_space.coordinate_system": {"coordinate_system.from": _planes(frm), "coordinate_system.to": _planes(to)},
            "tangent_space.mean": _planes(j[1:7]),
        },
        "weight.pmf": _dyn(j[0]),
    }


def _read_planes(obj, n):
    planes = obj["data"]
    if len(planes) != n:
        raise ValueError("expected %d planes, found %d" % (n, len(planes)))
    return np.asarray([[np.nan if v is None else v for v in p["data"]] for p in planes], np.float32)


def mixture_from_json(obj):
    """The reference's SDMM object -> joint [32, K] (None for an empty mixture).  Accepts the old format that
    stored only `tangent_space.mean` at the top level (sdmm.h:150-155)."""
    pmf = obj["weight.pmf"]
    K = int(pmf["size"])
    if K == 0:
        return None
    ts = obj["tangent_space"] if "tangent_space" in obj else obj
    out = np.zeros((32, K), np.float32)
    out[0] = np.asarray(pmf["data"], np.float32)
    out[1:7] = _read_planes(ts["tangent_space.mean"], 6)
    out[7:32] = _read_planes(obj["cov"], 25)
    return out


def save_asdmm(path, tree, joint_by_slot, K=None, em_by_slot=None, scal_by_slot=None, pad_nodes=REFERENCE_NODE_COUNT,
               indent=4):
    """Write `tree` (a HostTree) and its leaves' mixtures.  joint_by_slot: [slots, 32, K_padded] as
    SDMMCells.download() returns them, indexed by the leaf's cell slot; K = components to store (default all).
    pad_nodes: the reference stores a fixed table of 10000 nodes; shorter tables are padded with default nodes so
    that index-based tools see the same shape (larger trees are written in full; 0 = no padding)."""
    nodes = []
    for i in range(tree.size()):
        leaf = bool(tree.is_leaf[i])
        value = None
        if leaf and tree.slot[i] >= 0 and joint_by_slot is not None:
            value = {"sdmm": mixture_to_json(joint_by_slot[tree.slot[i]], K)}
            if em_by_slot is not None:
                value["b200.em"] = _planes(np.asarray(em_by_slot[tree.slot[i]], np.float32)[:, :K])
            if scal_by_slot is not None:
                value["b200.scalars"] = [int(v) for v in np.asarray(scal_by_slot[tree.slot[i]], np.uint32)]
        nodes.append({
            "aabb": _box(tree.amin[i], tree.amax[i]), "axis": int(tree.axis[i]),
            "children": [int(tree.children[i][0]), int(tree.children[i][1])], "depth": int(tree.depth[i]),
            "idx": i, "is_leaf": leaf, "value": value})
    while len(nodes) < pad_nodes:
        nodes.append({"aabb": _box((0, 0, 0), (0, 0, 0)), "axis": 0, "children": [0, 0], "depth": 0, "idx": 0,
                      "is_leaf": True, "value": None})
    doc = {"m_aabb": _box(tree.amin[0], tree.amax[0]), "m_nodes": nodes}
    with open(path, "w") as f:
        json.dump(doc, f, indent=indent, sort_keys=True, allow_nan=False)   # nlohmann's std::map orders keys the same way
        f.write("\n")


def load_asdmm(path, max_cells=None):
    """Read a checkpoint written by the reference or by save_asdmm.  Returns (tree, joint, extras):
    tree = HostTree with the reachable nodes in file order and one cell slot per leaf (depth-first from the
    root, the order STree::find visits them); joint = [n_leaves, 32, K] by slot (leaves without a mixture get
    zero planes and are listed in extras["empty_slots"]); extras["em"] / extras["scalars"] when present."""
    if str(path).endswith(".gz"):
        import gzip
        with gzip.open(path, "rt") as f:
            doc = json.load(f)
    else:
        with open(path) as f:
            doc = json.load(f)
    nodes = doc["m_nodes"]
    if not nodes:
        raise ValueError("checkpoint holds no nodes")

    def vec(o):
        return np.asarray([np.nan if v is None else v for v in o["data"]], np.float32)

    # reachable nodes, renumbered in file order
    reach, stack = set(), [0]
    while stack:
        i = stack.pop()
        if i in reach:
            continue
        reach.add(i)
        if not nodes[i]["is_leaf"]:
            stack.extend(int(c) for c in nodes[i]["children"])
    order = sorted(reach)
    renum = {old: new for new, old in enumerate(order)}
    n_leaves = sum(1 for i in order if nodes[i]["is_leaf"])
    root = nodes[0]
    tree = HostTree(vec(root["aabb"]["min"]), vec(root["aabb"]["max"]), max(n_leaves, max_cells or 0, 1))
    tree.is_leaf, tree.axis, tree.children, tree.amin, tree.amax, tree.depth, tree.slot = [], [], [], [], [], [], []
    tree.st_size, tree.st_sum, tree.st_sqr = [], [], []
    for i in order:
        nd = nodes[i]
        leaf = bool(nd["is_leaf"])
        tree.is_leaf.append(leaf)
        tree.axis.append(int(nd["axis"]))
        tree.children.append((0, 0) if leaf else (renum[int(nd["children"][0])], renum[int(nd["children"][1])]))
        tree.amin.append(vec(nd["aabb"]["min"]))
        tree.amax.append(vec(nd["aabb"]["max"]))
        tree.depth.append(int(nd["depth"]))
        tree.slot.append(-1)
        tree.st_size.append(0)
        tree.st_sum.append(np.zeros(3))
        tree.st_sqr.append(np.zeros(3))
    # slots: depth-first, child 0 before child 1
    leaf_nodes, stack = [], [0]
    while stack:
        i = stack.pop()
        if tree.is_leaf[i]:
            leaf_nodes.append(i)
        else:
            stack.extend(reversed(tree.children[i]))
    mixes, ems, scals, empty, K = [], {}, {}, [], 0
    for slot, i in enumerate(leaf_nodes):
        tree.slot[i] = slot
        value = nodes[order[i]]["value"]
        mix = mixture_from_json(value["sdmm"]) if value is not None else None
        mixes.append(mix)
        if mix is None:
            empty.append(slot)
        else:
            K = max(K, mix.shape[1])
            if "b200.em" in value:
                ems[slot] = np.asarray([p["data"] for p in value["b200.em"]["data"]], np.float32)
            if "b200.scalars" in value:
                scals[slot] = np.asarray(value["b200.scalars"], np.uint32)
    tree.n_leaves = n_leaves
    tree.free_slots = list(range(tree.max_cells - 1, n_leaves - 1, -1))
    joint = np.zeros((n_leaves, 32, max(K, 1)), np.float32)
    for slot, mix in enumerate(mixes):
        if mix is not None:
            joint[slot, :, :mix.shape[1]] = mix
    extras = {"empty_slots": empty, "K": K}
    if ems:
        extras["em"] = ems
    if scals:
        extras["scalars"] = scals
    return tree, joint, extras
