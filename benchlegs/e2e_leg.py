"""End to end through the host-buffer API: every step copies that step's samples from pinned host memory to
the device, fits, and reads the fitted mixtures back (sdmmb_optimize_host: chunk-pipelined inside the
library).  Also measures what the host link itself delivers (the same H2D bytes, copy only), so that a
multi-GPU e2e number can be read against the link instead of against the kernels."""
import time

import numpy as np


def h2d_copy_rate(S, h_point6, n_rep=3):
    """GB/s of this rank copying its step's sample planes host -> device, all ranks at the same time."""
    torch = S.torch
    d = torch.empty_like(h_point6, device=S.device)
    d.copy_(h_point6, non_blocking=True)
    S.barrier()
    t0 = time.perf_counter()
    for _ in range(n_rep):
        d.copy_(h_point6, non_blocking=True)
    torch.cuda.synchronize()
    secs = (time.perf_counter() - t0) / n_rep
    worst = S.maxreduce(secs)
    del d
    return h_point6.numel() * 4 / worst / 1e9


def run(S, args, cfg, live, updates_per_step):
    torch = S.torch
    K, N, iters = cfg["K"], cfg["N"], cfg["iters"]
    cells, off, first, C_local, n_local = live["cells"], live["off"], live["first"], live["C_local"], live["n_local"]
    h_point6 = torch.empty((6, n_local), dtype=torch.float32, pin_memory=True)
    h_weight = torch.empty(n_local, dtype=torch.float32, pin_memory=True)
    h_point6.copy_(live["point6"])
    h_weight.copy_(live["weight"])
    h_joint = torch.empty((C_local, 32, K), dtype=torch.float32, pin_memory=True)
    h_scal = torch.empty((C_local, 8), dtype=torch.int32, pin_memory=True)
    np_p6, np_w = h_point6.numpy(), h_weight.numpy()
    np_joint, np_scal = h_joint.numpy(), h_scal.numpy().view(np.uint32)
    torch.cuda.synchronize()
    exchange = live["exchange"]

    def e2e_step():
        # one call = H2D of the step's samples + fit + D2H of the fitted mixtures, pipelined over
        # chunks of cells inside the library (sdmmb_optimize_host)
        cells.optimize_host(np_p6, np_w, off, iters, first_cell=first, with_conditioner=True,
                            out=(np_joint, None, np_scal))
        if S.world > 1:
            cells.allgather_records(C_local)

    e2e_step()
    S.barrier()
    e2e_step()
    n_e2e = max(1, args.steps)
    S.barrier()
    t0 = time.perf_counter()
    for _ in range(n_e2e):
        e2e_step()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    e_ms = S.maxreduce((t1 - t0) * 1e3 / n_e2e)
    assert np.isfinite(np_joint).all()
    bad = int((np_scal[:, 7] != 0).sum())
    link = h2d_copy_rate(S, h_point6)
    h2d = int(28 * n_local)
    return dict(value=updates_per_step / (e_ms * 1e-3), unit="updates/s", ms_per_step=e_ms, steps=n_e2e,
                h2d_bytes_per_step=h2d, d2h_bytes_per_step=int(np_joint.nbytes + np_scal.nbytes),
                cells_with_nonzero_status=bad,
                host_link=dict(h2d_gbs_per_gpu_all_ranks_copying=link, copy_only_ms_per_step=h2d / link / 1e6,
                               note="the step's H2D bytes copied alone by all %d rank(s) at once, slowest rank: the floor "
                                    "the host link puts under ms_per_step" % S.world),
                api="SDMMCells.optimize_host = sdmmb_optimize_host (bind host samples -> %d x em_step -> conditioner "
                    "-> fitted mixtures back to host; pinned buffers, copies pipelined against the fit)%s"
                    % (iters, "; records all-gathered over NCCL" if S.world > 1 else ""))
