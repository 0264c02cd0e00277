"""BASELINE.json configs[4], timed in every `bench.py --gpus N` run next to the headline: 65536 cells x 128
components x 8192 samples/cell, cells sharded over the ranks (STRONG scaling: the total is fixed), one EM
iteration + conditioner per step, the refitted query records exchanged every step.  N = 1 fits one B200
(15 GB of sample planes).  Reference decomposition: the cell-parallel loop of volpath_sdmm.cpp:287-311."""
import time

import numpy as np

from . import data
from .em_leg import attach_exchange, estep_roofline

CFG5 = dict(cells=65536, K=128, N=8192, iters=1)


def run(S, args, steps=5, warmup=2, e2e_steps=2):
    torch = S.torch
    from sdmm_mitsuba_b200 import synth
    from sdmm_mitsuba_b200.mixture import SDMMCells
    K, N, iters = CFG5["K"], CFG5["N"], CFG5["iters"]
    C_total = CFG5["cells"]
    C_local = C_total // S.world
    first = S.rank * C_local
    cells = SDMMCells(C_total, K, device=S.local)
    for c0 in range(0, C_local, 4096):
        n = min(4096, C_local - c0)
        joint, em, scal = synth.make_initial_model(n, K, seed=5000 + first + c0)
        cells.upload(joint, em, scal, first_cell=first + c0)
    del joint, em, scal
    point6, weight = data.device_batch_generated(torch, C_local, N, 9000 + 100000 * S.rank, S.device)
    n_local = point6.shape[1]
    off = (np.arange(C_local + 1, dtype=np.uint64) * N).astype(np.uint32)
    torch.cuda.synchronize()
    overlap = args.overlap if args.overlap is not None else (S.world > 1)

    def timed(step, n_warm, n_steps):
        for _ in range(n_warm):
            step()
        S.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n_steps):
            step()
        e1.record()
        S.barrier()
        return S.maxreduce(e0.elapsed_time(e1)) / n_steps

    def fit():
        cells.set_samples(point6, weight, off, first_cell=first, borrowed=True)
        cells.em_step(iters, with_conditioner=True)

    # (1) the fit alone, in order: per-launch E-step durations for the roofline, and the exchange-free step time
    cells.set_em_overlap(False)
    ms_fit_inorder = timed(fit, warmup, steps)
    estep_ms, estep_n = cells.last_estep_ms()
    cells.set_em_overlap(overlap)
    ms_fit = timed(fit, 1, steps) if overlap else ms_fit_inorder
    # (2) with the record exchange inside the step
    exchange = attach_exchange(S, cells, args.exchange)
    wait_ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]

    def step():
        fit()
        if exchange == "peer":
            wait_ev[0].record()
            cells.wait_peers()
            wait_ev[1].record()
        elif exchange == "nccl":
            wait_ev[0].record()
            cells.allgather_records(C_local)
            wait_ev[1].record()

    ms_step = timed(step, 1, steps) if S.world > 1 else ms_fit
    wait_ms = S.maxreduce(wait_ev[0].elapsed_time(wait_ev[1])) if S.world > 1 else 0.0
    peer_timeouts = cells.peer_status()[1] if exchange == "peer" else None
    if peer_timeouts:
        raise SystemExit("bench cfg5: %d peer-exchange timeouts on rank %d" % (peer_timeouts, S.rank))
    status = cells.status(first_cell=first, n_cells=C_local)
    bad = int(S.maxreduce(float((status != 0).sum())))
    updates = float(C_total) * K * N * iters
    est_ms = estep_ms / max(1, estep_n)
    out = dict(
        workload="cfg5: 65536 cells x 128 components x 8192 samples/cell, %d cells per GPU (strong scaling), one EM "
                 "iteration + conditioner per step, query records exchanged every step (%s)" % (C_local, exchange),
        scaling="strong", n_gpus=S.world, value=updates / (ms_step * 1e-3), unit="updates/s", ms_per_iteration=ms_step,
        ms_fit_only=ms_fit, exchange_ms=ms_step - ms_fit, exchange_wait_ms=wait_ms, exchange=exchange,
        peer_timeouts=peer_timeouts, steps=steps, warmup=warmup, cells_with_nonzero_status=bad,
        schedule="overlapped (sdmmb_set_em_overlap)" if overlap else "in order",
        samples="generated on the device (torch generator of the synth.make_samples family), %.1f GB of planes per GPU"
                % (28.0 * n_local / 1e9),
        roofline=estep_roofline(S, K, est_ms, C_local * N, True, None, None))
    # (3) end to end through host buffers: H2D of the rank's planes + fit + D2H of its fitted mixtures
    try:
        h_point6 = torch.empty((6, n_local), dtype=torch.float32, pin_memory=True)
        h_weight = torch.empty(n_local, dtype=torch.float32, pin_memory=True)
        h_point6.copy_(point6)
        h_weight.copy_(weight)
        h_joint = torch.empty((C_local, 32, K), dtype=torch.float32, pin_memory=True)
        h_scal = torch.empty((C_local, 8), dtype=torch.int32, pin_memory=True)
        np_p6, np_w = h_point6.numpy(), h_weight.numpy()
        np_joint, np_scal = h_joint.numpy(), h_scal.numpy().view(np.uint32)
        del point6, weight
        torch.cuda.empty_cache()

        def e2e_step():
            cells.optimize_host(np_p6, np_w, off, iters, first_cell=first, with_conditioner=True,
                                out=(np_joint, None, np_scal))
            if S.world > 1:
                cells.allgather_records(C_local)

        e2e_step()
        S.barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        torch.cuda.synchronize()
        e_ms = S.maxreduce((time.perf_counter() - t0) * 1e3 / e2e_steps)
        out["e2e"] = dict(value=updates / (e_ms * 1e-3), unit="updates/s", ms_per_iteration=e_ms, steps=e2e_steps,
                          h2d_bytes_per_step=int(28 * n_local), d2h_bytes_per_step=int(np_joint.nbytes + np_scal.nbytes),
                          api="sdmmb_optimize_host per rank (+ NCCL all-gather of the records for N > 1)")
        del h_point6, h_weight, h_joint, h_scal
    except Exception as e:   # pinned host memory for 15 GB of planes is the box's to refuse
        out["e2e"] = dict(value=None, unavailable="%s: %s" % (type(e).__name__, str(e)[:200]))
    cells.close()
    torch.cuda.empty_cache()
    return out
