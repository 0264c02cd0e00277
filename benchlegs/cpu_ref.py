"""The CPU legs of bench.py -- the ONLY bench module that touches oracle/ (as the thing timed beside the
product, or as its checker; never on the product's path):
  cpu_baseline / cpu_query_baseline   the reference's own EM / per-vertex query code on the host cores
  run_reference                       `bench.py --impl reference`
  parity_block                        after the timed region: 16 cells of the timed batch re-fitted from a
                                      fresh model on the GPU and by oracle/_ref on the same planes
The thread count is always passed explicitly: under torchrun OMP_NUM_THREADS is 1, and a reference arm
that silently runs on one core inflates every ratio by the host's core count."""
import json
import os

import numpy as np

from . import data


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def _ref():
    from oracle import portbind, refbind
    if refbind.available():
        return refbind.RefCells, "reference"
    return portbind.PortCells, "port"


def cpu_baseline(K, N, iters, seconds_target=12.0, n_threads=0, seed_rank=0):
    """oracle/_ref (the reference's headers compiled in place; else the plain-C port) with an OpenMP
    parallel-for over cells -- ThreadPool::parallelFor's decomposition (volpath_sdmm.cpp:287-311) -- on a
    bounded sample of the GPU arm's own batch: the first C cells of the same seeded stream and model."""
    cls, kind = _ref()
    cores = n_threads if n_threads else host_threads()

    def run(C):
        point6, weight = data.host_cells(0, C, N, data.batch_seed(seed_rank))
        joint, em, scal = data.model_cells(0, C, K, data.model_seed(seed_rank))
        off = (np.arange(C + 1, dtype=np.uint64) * N).astype(np.uint32)
        o = cls(C, K)
        o.set(joint, em, scal)
        return o.em_batch(off, point6, weight, iters, with_prepare=True, n_threads=cores)

    run(cores)                                              # warm-up (thread pool, page faults)
    per_round = max(run(cores), 1e-3)                       # calibration: one cell per thread
    rounds = int(max(1, min(256, seconds_target / per_round)))
    C = cores * rounds
    secs = run(C)
    updates = float(C) * K * N * iters
    return dict(value=updates / secs, unit="updates/s", cores=int(cores), kind=kind,
                sample="the first %d cells of the GPU arm's batch (same seeded samples and initial model; K=%d, "
                       "N=%d, %d EM iterations + conditioner prepare), %.1f s, OpenMP over cells, %d threads"
                       % (C, K, N, iters, secs, cores)), secs


def cpu_query_baseline(K, n_queries=1 << 20, n_cells=64, n_threads=0):
    """The reference's per-vertex create_conditional + SDMM::sample + posterior (sdmm_proc.cpp:361-372,
    419-421, 539-545) with OpenMP over vertices and per-thread scratch, on mixtures fitted by the same
    CPU code; a bounded sample (n_queries over n_cells cells) of cfg3."""
    from sdmm_mitsuba_b200 import synth
    cls, kind = _ref()
    cores = n_threads if n_threads else host_threads()
    point6, weight, off = synth.make_samples(n_cells, 1024, seed=321)
    joint, em, scal = synth.make_initial_model(n_cells, K, seed=322)
    o = cls(n_cells, K)
    o.set(joint, em, scal)
    o.em_batch(off, point6, weight, 2, with_prepare=True, n_threads=cores)
    o.prepare_conditioner()
    cell, pos3, u3 = synth.make_queries(n_queries, n_cells, seed=323, binned=False)
    o.sample_pdf(cell[:4096].copy(), np.ascontiguousarray(pos3[:, :4096]), np.ascontiguousarray(u3[:, :4096]),
                 n_threads=cores)
    _, secs = o.sample_pdf(cell, pos3, u3, n_threads=cores)
    return dict(value=n_queries / secs, unit="queries/s", cores=int(cores), kind=kind,
                sample="%d vertices over %d fitted cells (K=%d), %.1f s, OpenMP over vertices, %d threads"
                       % (n_queries, n_cells, K, secs, cores))


def run_reference(args, cfg, rank, world):
    """`bench.py --impl reference`: the reference's CPU EM on all host threads; rank 0 alone works."""
    if rank != 0:
        return
    K, N, iters = cfg["K"], cfg["N"], cfg["iters"]
    threads = host_threads()
    vals = []
    for i in range(args.warmup + args.steps):
        cb, secs = cpu_baseline(K, N, iters, seconds_target=3.0, n_threads=threads)
        if i >= args.warmup:
            vals.append((cb, secs))
    v = float(np.mean([c["value"] for c, _ in vals]))
    cb = dict(vals[-1][0])
    cb["value"] = v
    line = dict(metric="EM sample-component updates/s", value=v, unit="updates/s", n_gpus=args.gpus,
                steps=args.steps, warmup=args.warmup, ms_per_step=1e3 * float(np.mean([s for _, s in vals])),
                higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32", data="synthetic",
                impl="reference",
                config=dict(workload="%s: batched stepwise tangent EM, K=%d, N=%d/cell, %d iterations; each step a "
                                     "bounded sample of the GPU arm's cells on the host cores (%d threads, set "
                                     "explicitly: torchrun's OMP_NUM_THREADS=1 is not inherited)"
                                     % (args.config, K, N, iters, threads)),
                cpu_baseline=cb,
                e2e=dict(value=v, unit="updates/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line), flush=True)


def _param_errs(j_a, j_b):
    """Max parameter errors (the metrics of tests/parity.py: pmf relative to the cell's max pmf, position
    means absolute, direction means by angle, covariance entrywise relative to max|cov| per component)."""
    a = np.asarray(j_a, np.float64)
    b = np.asarray(j_b, np.float64)
    live = b[:, 0, :] > 0
    pmf = np.abs(a[:, 0] - b[:, 0]) / b[:, 0].max(axis=1, keepdims=True)
    pos = np.abs(a[:, 1:4] - b[:, 1:4]).max(axis=1)
    da = a[:, 4:7] / np.linalg.norm(a[:, 4:7], axis=1, keepdims=True)
    db = b[:, 4:7] / np.linalg.norm(b[:, 4:7], axis=1, keepdims=True)
    ang = np.arctan2(np.linalg.norm(np.cross(da, db, axis=1), axis=1), (da * db).sum(axis=1))
    cov = np.abs(a[:, 7:32] - b[:, 7:32]).max(axis=1) / np.abs(b[:, 7:32]).max(axis=1)
    return {k: float(np.where(live, v, 0.0).max()) for k, v in dict(pmf=pmf, pos=pos, ang=ang, cov=cov).items()}


def parity_block(K, N, iters, point6_np, weight_np, joint, em, scal, device, n_cells=16):
    """Re-fit n_cells of the timed batch from the fresh model on the GPU (its own small context) and through
    oracle/_ref on the same planes.  Two comparisons: `stepwise` = every iteration starts from the
    reference's state (isolates one em_step: the tolerances of BASELINE.json apply to it), `free_running` =
    all iterations without re-synchronising (float32 drift of two implementations compounds)."""
    from sdmm_mitsuba_b200.mixture import SDMMCells
    cls, kind = _ref()
    C = n_cells
    off = (np.arange(C + 1, dtype=np.uint64) * N).astype(np.uint32)

    def ll_err(a, b):
        a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
        fin = np.isfinite(a) & np.isfinite(b)
        rel = np.abs(a[fin] - b[fin]) / np.maximum(1.0, np.abs(b[fin]))
        return rel, float(np.mean(np.isfinite(a) != np.isfinite(b)))

    o = cls(C, K)
    o.set(joint, em, scal)
    cells = SDMMCells(C, K, device=device)
    cells.upload(joint, em, scal)
    cells.set_samples(point6_np, weight_np, off)
    ll_free = cells.em_step(iters, with_conditioner=True, want_loglik=True).cpu().numpy()
    j_free, _, s_free = cells.download()
    step_ll, step_par = 0.0, {}
    ll_ref_all = []
    for it in range(iters):
        j0, e0, s0 = o.get()
        ll_ref = np.concatenate(o.em_step(off, point6_np, weight_np, 1, want_loglik=True), axis=1)[0]
        ll_ref_all.append(ll_ref)
        j1, _, _ = o.get()
        cells.upload(j0, e0, s0)
        ll_g = cells.em_step(1, with_conditioner=False, want_loglik=True).cpu().numpy()[0]
        jg, _, _ = cells.download()
        rel, mism = ll_err(ll_g, ll_ref)
        step_ll = max(step_ll, float(rel.max()) if rel.size else 0.0, mism)
        for k, v in _param_errs(jg, j1).items():
            step_par[k] = max(step_par.get(k, 0.0), v)
    j_ref, _, s_ref = o.get()
    # the reference against ITSELF with every cell's samples in a permuted order: the float32 summation-order
    # noise of sdmm::em_step on this very data, the yardstick for the free-running differences below
    rng = np.random.default_rng(0)
    perm = np.concatenate([c * N + rng.permutation(N) for c in range(C)])
    inv = np.empty_like(perm)
    inv[perm] = np.arange(perm.size)
    o2 = cls(C, K)
    o2.set(joint, em, scal)
    ll_perm = np.concatenate(o2.em_step(off, np.ascontiguousarray(point6_np[:, perm]), np.ascontiguousarray(weight_np[perm]),
                                        iters, want_loglik=True), axis=1)[:, inv]
    rel_n, _ = ll_err(ll_perm, np.stack(ll_ref_all))
    noise_par = _param_errs(o2.get()[0], j_ref)
    rel, mism = ll_err(ll_free, np.stack(ll_ref_all))
    free_par = _param_errs(j_free, j_ref)
    cells.close()
    return dict(oracle=kind, cells=C, K=K, samples_per_cell=N, iterations=iters,
                stepwise=dict(ll_err=step_ll, param_err=max(step_par.values()), param_errs=step_par,
                              tolerance=dict(ll=1e-4, param=1e-3),
                              ok=bool(step_ll <= 1e-4 and max(step_par.values()) <= 1e-3)),
                free_running=dict(ll_err_p999=float(np.quantile(rel, 0.999)), ll_err_max=float(rel.max()),
                                  ll_nonfinite_mismatch=mism, param_err=max(free_par.values()), param_errs=free_par,
                                  skipped_samples_equal=bool((s_free[:, 2] == s_ref[:, 2]).all()),
                                  reference_order_noise=dict(
                                      ll_err_p999=float(np.quantile(rel_n, 0.999)), ll_err_max=float(rel_n.max()),
                                      param_err=max(noise_par.values()),
                                      what="the reference against itself with each cell's samples fed in a permuted order "
                                           "(float32 summation-order noise of sdmm::em_step on the same data)")),
                note="inputs are the timed batch's own first cells (bit-identical planes on both sides)")
