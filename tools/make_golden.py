#!/usr/bin/env python
"""Generates tests/golden/*.npz by running the REFERENCE ITSELF (its sdmm headers compiled in place
from /root/reference as oracle/_ref/libsdmm_ref.so) on small seeded inputs.  The fixtures travel to
the GPU box (where /root/reference does not exist) and pin both the plain-C oracle restatement and
the CUDA path.  Run in the build container:  python tools/make_golden.py

Each fixture holds: the inputs (samples, initial joint/em/scal), and for every one of `iters`
stepwise iterations the reference's per-sample log-likelihood and its state AFTER sdmm::em_step;
then the conditioner internals after sdmm::prepare and the results of create_conditional + sample +
posterior for seeded queries."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import refbind  # noqa: E402
from sdmm_mitsuba_b200 import synth  # noqa: E402

CASES = {
    # name: (cells, K, samples per cell, EM iterations, queries)
    "em_k16": (2, 16, 300, 4, 2000),
    "em_k64": (1, 64, 640, 4, 2000),
    "em_k128": (1, 128, 512, 2, 1000),
    # the per-cell shapes bench.py times (cfg2: K=64, N=4096, 4 iterations; cfg5: K=128, N=8192)
    "em_k64_n4096": (1, 64, 4096, 4, 500),
    "em_k128_n8192": (1, 128, 8192, 2, 500),
}


def make(name, C, K, N, iters, nq):
    point6, weight, off = synth.make_samples(C, N, seed=1000 + K)
    # a few samples the EM loop must skip (sdmm/opt/data.h:20-23, em.h:267-269)
    weight[3] = 0.0
    weight[N // 2] = 1e-9
    joint, em, scal = synth.make_initial_model(C, K, seed=2000 + K)
    o = refbind.RefCells(C, K)
    assert o.set(joint, em, scal).all()
    out = dict(point6=point6, weight=weight, offsets=off, joint0=joint, em0=em, scal0=scal)
    out["derived0"] = o.derived()
    for it in range(iters):
        ll = np.concatenate(o.em_step(off, point6, weight, 1, want_loglik=True), axis=1)[0]
        j, e, s = o.get()
        out["loglik%d" % it] = ll
        out["joint%d" % (it + 1)] = j
        out["em%d" % (it + 1)] = e
        out["scal%d" % (it + 1)] = s
    out["derived_final"] = o.derived()
    o.prepare_conditioner()
    out["conditioner"] = o.conditioner()
    cell, pos3, u3 = synth.make_queries(nq, C, seed=3000 + K, binned=True)
    r, _ = o.sample_pdf(cell, pos3, u3)
    out.update(q_cell=cell, q_pos3=pos3, q_u3=u3, q_dir=r["dir"], q_pdf=r["pdf"], q_inv_jac=r["inv_jac"],
               q_comp=r["comp"], q_valid=r["valid"])
    # distance of u0 to the nearest conditional-CDF entry: the tie margin of each query
    margin = np.zeros(nq, np.float32)
    for i in range(nq):
        pmf, cdf, ok = o.conditional_weights(int(cell[i]), pos3[:, i])
        margin[i] = np.abs(cdf - u3[0, i]).min() if ok else 0.0
    out["q_tie_margin"] = margin
    rng = np.random.default_rng(4000 + K)
    d = rng.normal(size=(3, nq)).astype(np.float32)
    d /= np.sqrt((d * d).sum(axis=0, keepdims=True, dtype=np.float32))
    p, v, _ = o.pdf(cell, pos3, d)
    out.update(p_dir=d, p_pdf=p, p_valid=v)
    out["meta"] = np.array([C, K, N, iters, nq], np.int64)
    path = os.path.join(ROOT, "tests", "golden", name + ".npz")
    np.savez_compressed(path, **out)
    print(path, "%.0f KB" % (os.path.getsize(path) / 1e3))


def make_tree():
    """The reference's STree: bootstrap split_to_depth(2) (volpath_sdmm.cpp:398), 30000 clustered points
    pushed through find + stats, data-driven splits at threshold 400, then the node table and find()
    results for fresh points (some exactly on split planes)."""
    rng = np.random.default_rng(77)
    n_points = 30000
    t = refbind.RefTree(capacity=n_points + 16)
    t.split_to_depth(2)
    centres = rng.uniform(0.15, 0.85, (5, 3))
    which = rng.integers(0, 5, n_points)
    p = np.clip(centres[which] + rng.normal(0, 0.07, (n_points, 3)), 1e-4, 1 - 1e-4).astype(np.float32).T.copy()
    assert t.push(p) == 0
    t.split_leaves(400)
    nodes = t.nodes()
    q = rng.uniform(0, 1, (3, 20000)).astype(np.float32)
    inner = np.nonzero(nodes[:, 0] == 0)[0]
    for j, i in enumerate(inner[:300]):
        ax = int(nodes[i, 1])
        q[ax, j] = nodes[int(nodes[i, 2]), 5 + ax]
    path = os.path.join(ROOT, "tests", "golden", "tree.npz")
    np.savez_compressed(path, nodes=nodes, points=p[:, :4000], points_node=t.find(p[:, :4000]), queries=q,
                        queries_node=t.find(q))
    print(path, "%.0f KB" % (os.path.getsize(path) / 1e3), "nodes", nodes.shape[0], "leaves", int(nodes[:, 0].sum()))


INIT_CASES = {
    # name: (cells, K, samples per cell)
    "init_k64": (3, 64, 1500),
    "init_k16": (4, 16, 600),
}


def make_init(name, C, K, N):
    """sdmm::initialize (opt/init.h:17-193) run by the reference on seeded surface-like samples, with the
    default-seeded PCG32 copy of initializeSDMMContext (volpath_sdmm.cpp:132-137); plus the first draws of
    the reference's generator."""
    point6, weight, off = synth.make_samples(C, N, seed=5000 + K)
    normal3 = synth.make_normals(C, N, seed=5000 + K)
    weight[5] = 0.0            # clamped up to 1e-3 by the sampling weight
    weight[7] = 50.0           # clamped down to 3
    rng = np.random.default_rng(6000 + K)
    sd = (3.0 * rng.uniform(0.3, 1.0, C) / (K // 8)).astype(np.float32)
    o = refbind.RefCells(C, K)
    o.initialize(off, point6, normal3, weight, sd)
    j, e, s = o.get()
    path = os.path.join(ROOT, "tests", "golden", name + ".npz")
    np.savez_compressed(path, point6=point6, normal3=normal3, weight=weight, offsets=off, spatial_distance=sd,
                        joint=j, em=e, scal=s, derived=o.derived(), pcg32=refbind.pcg32_floats(64),
                        meta=np.array([C, K, N], np.int64))
    print(path, "%.0f KB" % (os.path.getsize(path) / 1e3))


def make_checkpoint():
    """A `.asdmm` checkpoint written by the reference's own serializers (oracle/ref_checkpoint.cpp): an STree
    split to depth 1 (8 leaves) whose leaves hold the mixtures of tests/test_checkpoint.py::_models; gzip'd
    (the reference always stores 10000 nodes)."""
    import gzip
    import json
    import shutil
    import tempfile
    from oracle import ckptbind
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import test_checkpoint as tc
    rt = ckptbind.RefCheckpointTree(tc.AABB[0], tc.AABB[1], 16)
    rt.split_to_depth(1)
    nodes = rt.stored_nodes(rt.size())
    leaves = [i for i in range(rt.size()) if nodes[i, 4]]
    seed = 11
    joint = tc._models(len(leaves), tc.K, seed=seed)
    for c, i in enumerate(leaves):
        assert rt.set_leaf_model(i, joint[c]) == 1
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "model_00000.asdmm")
        rt.save(p)
        out = os.path.join(ROOT, "tests", "golden", "ref_tree.asdmm.gz")
        with open(p, "rb") as fi, gzip.GzipFile(out, "wb", mtime=0) as fo:
            shutil.copyfileobj(fi, fo)
    with open(out.replace(".asdmm.gz", ".meta.json"), "w") as f:
        json.dump(dict(nodes=rt.size(), leaves=len(leaves), K=tc.K, seed=seed,
                       written_by="sdmm::save_json layout via oracle/_ref/libsdmm_ref_ckpt.so (tools/make_golden.py --checkpoint)"), f)
    print(out, "%.0f KB" % (os.path.getsize(out) / 1e3))


if __name__ == "__main__":
    if "--checkpoint" in sys.argv:
        make_checkpoint()
        sys.exit(0)
    assert refbind.available(), "build oracle/_ref first (make -C oracle ref)"
    _only = [a for a in sys.argv[1:] if a in CASES]
    if _only:   # regenerate just the named EM fixtures (leaves every other pinned file untouched)
        for name in _only:
            make(name, *CASES[name])
        sys.exit(0)
    for name, cfg in INIT_CASES.items():
        make_init(name, *cfg)
    if "--init-only" in sys.argv:
        sys.exit(0)
    make_tree()
    for name, cfg in CASES.items():
        make(name, *cfg)
