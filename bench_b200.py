"""Second half of bench.py: the B200 arm (one process per GPU; torchrun provides RANK/WORLD_SIZE)."""
import json
import os
import sys
import time

import numpy as np


def _maxreduce(torch, dist, world, v, device):
    if world == 1:
        return v
    t = torch.tensor([v], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def run_b200(args, cfg, rank, world, ClockSampler, make_device_samples, cpu_baseline, FLOP_PER_UPDATE,
             SFU_PER_UPDATE):
    import torch
    import torch.distributed as dist
    from sdmm_mitsuba_b200 import synth
    from sdmm_mitsuba_b200.mixture import SDMMCells, measure_peaks

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path "
                         "(use --impl reference for the CPU reference arm)")
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"   # keep stdout to the one JSON line (no "NCCL version" banner)
        # NCCL prints its version banner on stdout when the communicator comes up (whatever the box's
        # NCCL_DEBUG says): point fd 1 at stderr meanwhile so that stdout stays the one JSON line
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=device)
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)

    K, N, iters = cfg["K"], cfg["N"], cfg["iters"]
    strong = args.config == "cfg5"
    C_local = cfg["cells"] // world if strong else cfg["cells"]
    C_total = C_local * world
    first = rank * C_local

    peaks = measure_peaks(local)

    cells = SDMMCells(C_total, K, device=local)
    joint, em, scal = synth.make_initial_model(C_local, K, seed=1000 + rank)
    cells.upload(joint, em, scal, first_cell=first)
    chunks = []
    step_c = 1024
    for c0 in range(0, C_local, step_c):
        chunks.append(make_device_samples(torch, min(step_c, C_local - c0), N, 77 + 1000 * rank + c0, device))
    point6 = torch.cat([c[0] for c in chunks], dim=1).contiguous()
    weight = torch.cat([c[1] for c in chunks]).contiguous()
    del chunks
    n_local = point6.shape[1]
    off = (np.arange(C_local + 1, dtype=np.uint64) * N).astype(np.uint32)
    torch.cuda.synchronize()

    # record exchange after the fit: "peer" = stores into every peer's record array fused into the
    # kernel that writes the records (NVLink peer memory, no collective call); "nccl" = all-gather
    exchange = "none"
    if world > 1:
        exchange = args.exchange
        ok = 1
        if exchange == "peer":
            try:
                cells.attach_peers()
            except Exception as e:  # no P2P / IPC between the ranks: say so, use the collective
                print("bench: peer attach failed on rank %d (%s); falling back to the NCCL all-gather" % (rank, e),
                      file=sys.stderr, flush=True)
                ok = 0
        flag = torch.tensor([ok], device=device, dtype=torch.int32)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if exchange == "peer" and int(flag.item()) == 0:
            if ok:
                cells.detach_peers()
            exchange = "nccl"

    borrowed = args.binding == "borrowed"
    cells.set_em_overlap(args.overlap)   # --overlap: M-steps hidden behind the other half's E-steps

    def step():
        # the step's batch of samples: bound zero-copy (the E-step reads the SoA planes in place, like the
        # reference trains on the swapped-in training_data) or packed into records first (--binding packed)
        cells.set_samples(point6, weight, off, first_cell=first, borrowed=borrowed)
        cells.em_step(iters, with_conditioner=True)
        if exchange == "peer":
            cells.wait_peers()
        elif exchange == "nccl":
            cells.allgather_records(C_local)

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = cells.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms_local = e0.elapsed_time(e1)
    peer_timeouts = cells.peer_status()[1] if exchange == "peer" else None
    launches_timed = cells.launch_count() - l0
    if args.overlap:
        # per-launch E-step durations need launches that do not queue behind each other: two more (untimed)
        # steps with the in-order schedule, same kernels on the same data, CUDA events around every E-step
        cells.set_em_overlap(False)
        for _ in range(2):
            step()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
    estep_ms, estep_n = cells.last_estep_ms()
    clocks = sampler.stop() if rank == 0 else None
    launches = launches_timed
    ms_total = _maxreduce(torch, dist, world, ms_local, device)
    ms_per_step = ms_total / args.steps
    updates_per_step = float(C_total) * K * N * iters
    value = updates_per_step / (ms_per_step * 1e-3)

    # ---- roofline of the dominant kernel (K1, E-step + statistics), rank 0's launches ----------
    est_ms = estep_ms / max(1, estep_n)
    upd_per_launch = float(C_local) * K * N
    achieved_tf = FLOP_PER_UPDATE * upd_per_launch / (est_ms * 1e-3) / 1e12
    peak_tf = max(peaks["fp32_ffma2_tflops"], peaks["fp32_ffma_tflops"])
    hbm_peak = None
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "MEASURED_PEAKS.json")) as f:
            hbm_peak = json.load(f).get("hbm_gbs")
    except Exception:
        pass
    hbm_src = "MEASURED_PEAKS.json" if hbm_peak else "fallback (B200_PROFILING.md)"
    hbm_peak = hbm_peak or 6650.0
    bytes_per_sample = 28.0 if borrowed else 32.0   # SoA planes vs packed records
    hbm_gbs = bytes_per_sample * C_local * N / (est_ms * 1e-3) / 1e9
    traffic, traffic_src = None, None
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r1_estep_traffic.json")) as f:
            tj = json.load(f)
        if not borrowed:
            tj = tj.get("packed", tj)
        traffic = tj["dram_bytes_per_sample"] * C_local * N
        traffic_src = tj["source"]
    except Exception:
        pass
    roofline = dict(
        bound="fp32", kernel="em_estep_kernel", achieved=achieved_tf, peak=peak_tf, unit="TFLOP/s",
        frac=achieved_tf / peak_tf, traffic=traffic, traffic_source=traffic_src,
        peak_source="FFMA2 micro-benchmark run by this process on this GPU (sdmmb_measure_peaks); "
                    "MEASURED_PEAKS.json carries no FP32 figure",
        flop_per_update=FLOP_PER_UPDATE, updates_per_launch=upd_per_launch, ms_per_launch=est_ms,
        share_of_step=estep_ms / ms_per_step if world == 1 else None,
        sfu=dict(achieved_gops=SFU_PER_UPDATE * upd_per_launch / (est_ms * 1e-3) / 1e9, peak_gops=peaks["mufu_gops"]),
        hbm=dict(achieved=hbm_gbs, peak=hbm_peak, unit="GB/s", frac=hbm_gbs / hbm_peak, peak_source=hbm_src,
                 algorithmic_bytes_per_sample=bytes_per_sample),
        measured_peaks=peaks,
        note="no tensor cores: the contraction is 5-D; flop-based fraction is capped at ~71% by the "
             "instruction mix (72 FP32-pipe lane-ops per 102 algorithmic flop)")

    # ---- guided queries (second half of the metric), all ranks as replicas ----------------------
    queries = None
    nq = args.queries if args.queries else (1 << 24)
    if nq > 0:
        cell_q, pos_q, u_q = synth.make_queries(nq, C_total, seed=5 + rank, binned=True)
        d_cell = torch.from_numpy(cell_q.view(np.int32)).to(device)
        d_pos = torch.from_numpy(pos_q).to(device)
        d_u = torch.from_numpy(u_q).to(device)
        perm = torch.randperm(nq, device=device)
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
            qms = _maxreduce(torch, dist, world, q0.elapsed_time(q1) / 3, device)
            res[name] = dict(value=world * nq / (qms * 1e-3), unit="queries/s", ms=qms)
            valid_frac = float(out["valid"].float().mean().item())
        flop_q = 80.0 * K + 120.0
        queries = dict(metric="guided sample+pdf queries/s", n=nq * world, K=K, binned=res["binned"],
                       random_order=res["random"], valid_fraction=valid_frac,
                       roofline=dict(bound="fp32", achieved=flop_q * res["binned"]["value"] / world / 1e12,
                                     peak=peaks["fp32_ffma_tflops"], unit="TFLOP/s",
                                     frac=flop_q * res["binned"]["value"] / world / 1e12 / peaks["fp32_ffma_tflops"],
                                     flop_per_query=flop_q))
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
            qms = _maxreduce(torch, dist, world, (time.perf_counter() - t0) * 1e3 / 3, device)
            e2e_q[name] = dict(value=world * nq / (qms * 1e-3), unit="queries/s", ms=qms)
        queries["e2e"] = dict(binned=e2e_q["binned"], random_order=e2e_q["random"], h2d_bytes_per_query=28,
                              d2h_bytes_per_query=25, api="SDMMCells.sample_pdf(host arrays) = sdmmb_sample_pdf_host")
        del d_cell, d_pos, d_u, r_cell, r_pos, r_u, out, hq, hr

    # ---- end to end through the host-buffer API --------------------------------------------------
    e2e = None
    if not args.no_e2e:
        h_point6 = torch.empty((6, n_local), dtype=torch.float32, pin_memory=True)
        h_weight = torch.empty(n_local, dtype=torch.float32, pin_memory=True)
        h_point6.copy_(point6)
        h_weight.copy_(weight)
        h_joint = torch.empty((C_local, 32, K), dtype=torch.float32, pin_memory=True)
        h_scal = torch.empty((C_local, 8), dtype=torch.int32, pin_memory=True)
        np_p6, np_w = h_point6.numpy(), h_weight.numpy()
        np_joint, np_scal = h_joint.numpy(), h_scal.numpy().view(np.uint32)
        torch.cuda.synchronize()

        def e2e_step():
            # one call = H2D of the step's samples + fit + D2H of the fitted mixtures, pipelined over
            # chunks of cells inside the library (sdmmb_optimize_host)
            cells.optimize_host(np_p6, np_w, off, iters, first_cell=first, with_conditioner=True,
                                out=(np_joint, None, np_scal))
            if world > 1:
                cells.allgather_records(C_local)

        e2e_step()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e2e_step()
        n_e2e = max(1, args.steps)
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            e2e_step()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        e_ms = _maxreduce(torch, dist, world, (t1 - t0) * 1e3 / n_e2e, device)
        assert np.isfinite(np_joint).all()
        e2e = dict(value=updates_per_step / (e_ms * 1e-3), unit="updates/s", ms_per_step=e_ms, steps=n_e2e,
                   h2d_bytes_per_step=int(28 * n_local), d2h_bytes_per_step=int(np_joint.nbytes + np_scal.nbytes),
                   api="SDMMCells.optimize_host = sdmmb_optimize_host (bind host samples -> 4 x em_step -> conditioner "
                       "-> fitted mixtures back to host; pinned buffers, copies pipelined against the fit)")

    cb = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cb, _ = cpu_baseline(K, N, iters)
        one, _ = cpu_baseline(K, N, iters, seconds_target=4.0, n_threads=1)
        cb["single_core"] = dict(value=one["value"], unit=one["unit"], sample=one["sample"])
        if queries is not None:
            from bench import cpu_query_baseline
            queries["cpu_baseline"] = cpu_query_baseline(K, n_threads=cb["cores"])   # all host threads again

    render = None
    if rank == 0 and world == 1 and not getattr(args, "no_render", False):
        from bench_render import run_render
        del point6, weight
        torch.cuda.empty_cache()
        render = run_render(torch, cpu=not args.no_cpu_baseline)

    if rank == 0:
        line = dict(
            metric="EM sample-component updates/s", value=value, unit="updates/s", n_gpus=world,
            steps=args.steps, warmup=args.warmup, ms_per_step=ms_per_step, higher_is_better=True,
            scaling="strong" if strong else "weak", vs_baseline=None, dtype="f32", data="synthetic",
            config=dict(workload="%s: batched StepwiseTangentEM (sdmm::em_step), %d cells x %d components x %d "
                                 "5-D samples/cell, %d EM iterations per step%s" %
                                 (args.config, C_total, K, N, iters,
                                  (" (%d cells per GPU, cell-sharded, query records exchanged every step: %s)"
                                   % (C_local, "peer-memory stores fused into the M-step kernel" if exchange == "peer"
                                      else "NCCL all-gather")) if world > 1 else ""),
                        exchange=exchange, peer_timeouts=peer_timeouts,
                        schedule="two halves, E-steps alternating on two streams, M-steps on a high-priority stream "
                                 "(sdmmb_set_em_overlap); roofline E-step durations from two extra in-order steps"
                        if args.overlap else "in order on one stream (E-step durations: CUDA events around every E-step "
                                             "launch of the timed region's last step)",
                        sample_binding="borrowed SoA planes read in place by the E-step (sdmmb_set_samples_borrowed)"
                        if borrowed else "packed 32-byte records (sdmmb_set_samples: one packing pass per step)",
                        cells=C_total, components=K, samples_per_cell=N, em_iterations=iters,
                        parallelism="cells sharded over %d GPU(s)" % world,
                        l2_policy="inputs larger than L2 (%.0f MB of samples per GPU per iteration)"
                                  % (bytes_per_sample * n_local / 1e6),
                        model_state="EM state carries over between steps (same work per step)"),
            clocks=clocks, e2e=e2e, gpu_launches=int(launches), roofline=roofline, cpu_baseline=cb,
            queries=queries, progressive_render=render)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
