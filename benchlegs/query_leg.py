"""Guided sample+pdf queries (second half of BASELINE.json's metric; cfg3 = 16 M path vertices, K = 64) on
the mixtures the EM leg just fitted: device-resident (binned by cell, and in random order with the
device-side binning inside the call) and end to end through host buffers."""
import time

import numpy as np


def run(S, args, cfg, live):
    torch = S.torch
    from sdmm_mitsuba_b200 import synth
    K = cfg["K"]
    cells, C_total = live["cells"], live["C_total"]
    nq = args.queries if args.queries else (1 << 24)
    cell_q, pos_q, u_q = synth.make_queries(nq, C_total, seed=5 + S.rank, binned=True)
    d_cell = torch.from_numpy(cell_q.view(np.int32)).to(S.device)
    d_pos = torch.from_numpy(pos_q).to(S.device)
    d_u = torch.from_numpy(u_q).to(S.device)
    perm = torch.randperm(nq, device=S.device)
    r_cell, r_pos, r_u = d_cell[perm].contiguous(), d_pos[:, perm].contiguous(), d_u[:, perm].contiguous()
    res = {}
    for name, (cc, pp, uu, binned) in dict(binned=(d_cell, d_pos, d_u, True),
                                           random=(r_cell, r_pos, r_u, False)).items():
        for _ in range(2):
            out = cells.sample_pdf(cc, pp, uu, binned=binned)
        torch.cuda.synchronize()
        q0, q1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        q0.record()
        for _ in range(3):
            out = cells.sample_pdf(cc, pp, uu, binned=binned)
        q1.record()
        torch.cuda.synchronize()
        qms = S.maxreduce(q0.elapsed_time(q1) / 3)
        res[name] = dict(value=S.world * nq / (qms * 1e-3), unit="queries/s", ms=qms)
        valid_frac = float(out["valid"].float().mean().item())
    flop_q = 80.0 * K + 120.0
    peak = max(S.peaks["fp32_ffma2_tflops"], S.peaks["fp32_ffma_tflops"])
    ach = flop_q * res["binned"]["value"] / S.world / 1e12
    queries = dict(metric="guided sample+pdf queries/s", n=nq * S.world, K=K, binned=res["binned"],
                   random_order=res["random"], valid_fraction=valid_frac,
                   roofline=dict(bound="fp32", kernel="query_pair_kernel", achieved=ach, peak=peak, unit="TFLOP/s",
                                 frac=ach / peak, flop_per_query=flop_q,
                                 peak_source="packed FFMA2 micro-benchmark (the kernel is packed)"))
    # end to end through the host-buffer API (pinned inputs and outputs, copies inside the call)
    hq = dict(cell=torch.from_numpy(cell_q.view(np.int32)).pin_memory(), pos=torch.from_numpy(pos_q).pin_memory(),
              u=torch.from_numpy(u_q).pin_memory())
    perm_h = perm.cpu()
    hr = dict(cell=hq["cell"][perm_h].contiguous().pin_memory(), pos=hq["pos"][:, perm_h].contiguous().pin_memory(),
              u=hq["u"][:, perm_h].contiguous().pin_memory())
    ho = dict(dir=torch.empty((3, nq), dtype=torch.float32).pin_memory().numpy(),
              pdf=torch.empty(nq, dtype=torch.float32).pin_memory().numpy(),
              inv_jac=torch.empty(nq, dtype=torch.float32).pin_memory().numpy(),
              comp=torch.empty(nq, dtype=torch.int32).pin_memory().numpy().view(np.uint32),
              valid=torch.empty(nq, dtype=torch.uint8).pin_memory().numpy())
    e2e_q = {}
    for name, (hh, binned) in dict(binned=(hq, True), random=(hr, False)).items():
        a_cell, a_pos, a_u = hh["cell"].numpy().view(np.uint32), hh["pos"].numpy(), hh["u"].numpy()
        cells.sample_pdf(a_cell, a_pos, a_u, binned=binned, out=ho)
        t0 = time.perf_counter()
        for _ in range(3):
            cells.sample_pdf(a_cell, a_pos, a_u, binned=binned, out=ho)
        qms = S.maxreduce((time.perf_counter() - t0) * 1e3 / 3)
        e2e_q[name] = dict(value=S.world * nq / (qms * 1e-3), unit="queries/s", ms=qms)
    queries["e2e"] = dict(binned=e2e_q["binned"], random_order=e2e_q["random"], h2d_bytes_per_query=28,
                          d2h_bytes_per_query=25, api="SDMMCells.sample_pdf(host arrays) = sdmmb_sample_pdf_host")
    return queries
