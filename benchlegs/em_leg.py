"""The headline leg: device-resident training steps of the named configuration (cfg2: 4096 cells x 64
components x 4096 samples x 4 EM iterations), one process per GPU, and the roofline of its dominant kernel
K1 (em_estep_kernel).  Weak scaling for cfg2 (every rank fits its own cells); the refitted query records
reach the other ranks inside the timed region (peer-memory stores fused into the M-step kernel, or NCCL)."""
import json
import os
import sys

import numpy as np

from . import data

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FLOP_PER_UPDATE = 102.0   # algorithmic FP32 flop per sample-component update
SFU_PER_UPDATE = 4.0


class Session:
    """Process-wide state of one bench run: device, rank, communicator, measured peaks."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device: the product has no CPU path "
                             "(use --impl reference for the CPU reference arm)")
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.device = torch.device("cuda", self.local)
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
                os.environ["NCCL_DEBUG"] = "WARN"   # keep stdout to the one JSON line (no "NCCL version" banner)
            # NCCL prints its version banner on stdout when the communicator comes up (whatever the box's
            # NCCL_DEBUG says): point fd 1 at stderr meanwhile so that stdout stays the one JSON line
            sys.stdout.flush()
            saved = os.dup(1)
            os.dup2(2, 1)
            try:
                dist.init_process_group("nccl", device_id=self.device)
                dist.barrier()
                torch.cuda.synchronize()
            finally:
                sys.stdout.flush()
                os.dup2(saved, 1)
                os.close(saved)
        from sdmm_mitsuba_b200.mixture import measure_peaks
        self.peaks = measure_peaks(self.local)

    def maxreduce(self, v):
        if self.world == 1:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()

    def close(self):
        if self.world > 1:
            self.dist.barrier()
            self.dist.destroy_process_group()


def attach_exchange(S, cells, want):
    """Record exchange after the fit: "peer" = stores into every peer's record array fused into the kernel
    that writes the records (NVLink peer memory, no collective call); "nccl" = all-gather."""
    if S.world == 1:
        return "none"
    torch, dist = S.torch, S.dist
    exchange, ok = want, 1
    if exchange == "peer":
        try:
            cells.attach_peers()
        except Exception as e:  # no P2P / IPC between the ranks: say so, use the collective
            print("bench: peer attach failed on rank %d (%s); falling back to the NCCL all-gather" % (S.rank, e),
                  file=sys.stderr, flush=True)
            ok = 0
    flag = torch.tensor([ok], device=S.device, dtype=torch.int32)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if exchange == "peer" and int(flag.item()) == 0:
        if ok:
            cells.detach_peers()
        exchange = "nccl"
    return exchange


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            v = json.load(f).get("hbm_gbs")
        if v:
            return float(v), "MEASURED_PEAKS.json"
    except Exception:
        pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def _profile_json(name):
    try:
        with open(os.path.join(ROOT, "profiles", name)) as f:
            return json.load(f)
    except Exception:
        return None


def estep_roofline(S, K, est_ms, n_samples, borrowed, share_of_step, sm_mhz):
    """achieved = algorithmic flop of one launch / its CUDA-event duration; peak = the packed-FFMA2
    micro-benchmark run by this process.  Beside the flop fraction: the issue-slot / pipe-active fractions
    of the committed ncu capture and the register-file bound that actually limits the kernel."""
    peaks = S.peaks
    upd = float(n_samples) * K
    achieved = FLOP_PER_UPDATE * upd / (est_ms * 1e-3) / 1e12
    peak = max(peaks["fp32_ffma2_tflops"], peaks["fp32_ffma_tflops"])
    hp, hp_src = hbm_peak()
    bps = 28.0 if borrowed else 32.0   # SoA planes vs packed records
    hbm = bps * n_samples / (est_ms * 1e-3) / 1e9
    traffic, traffic_src = None, None
    tj = _profile_json("r2_estep_traffic.json") or _profile_json("r1_estep_traffic.json")
    if tj:
        if not borrowed:
            tj = tj.get("packed", tj)
        traffic = tj["dram_bytes_per_sample"] * n_samples
        traffic_src = tj["source"]
    out = dict(
        bound="fp32", kernel="em_estep_kernel", achieved=achieved, peak=peak, unit="TFLOP/s", frac=achieved / peak,
        traffic=traffic, traffic_source=traffic_src,
        peak_source="FFMA2 micro-benchmark run by this process on this GPU (sdmmb_measure_peaks); "
                    "MEASURED_PEAKS.json carries no FP32 figure",
        flop_per_update=FLOP_PER_UPDATE, updates_per_launch=upd, ms_per_launch=est_ms, share_of_step=share_of_step,
        sfu=dict(achieved_gops=SFU_PER_UPDATE * upd / (est_ms * 1e-3) / 1e9, peak_gops=peaks["mufu_gops"]),
        hbm=dict(achieved=hbm, peak=hp, unit="GB/s", frac=hbm / hp, peak_source=hp_src,
                 algorithmic_bytes_per_sample=bps),
        measured_peaks=peaks)
    # the instruction-mix caveat and the bound that explains the number
    rf = _profile_json("r2_k1/rf_model.json")
    if rf and str(K) in rf.get("per_K", {}):
        m = rf["per_K"][str(K)]
        sms = rf.get("sms", 148)
        clk = (sm_mhz or peaks.get("sm_mhz") or 1965.0) * 1e6
        pair_steps_per_smsp = upd / 64.0 / (sms * 4)   # a warp pair-step = 32 lanes x 2 components x 1 sample
        cyc = est_ms * 1e-3 * clk / pair_steps_per_smsp
        out["instruction_mix"] = dict(
            packed_fp32_per_pair_step=m["packed_per_pair_step"], scalar_fp32_per_pair_step=m["scalar_fp32_per_pair_step"],
            fp32_lane_ops_per_update=m["packed_per_pair_step"] + m["scalar_fp32_per_pair_step"] / 2.0,
            fma_pipe_cycles_per_pair_step=m["fma_pipe_cycles"], instructions_per_pair_step=m["instructions"],
            flop_fraction_cap=FLOP_PER_UPDATE / m["fma_pipe_cycles"],
            measured_cycles_per_pair_step=cyc, fma_pipe_active=m["fma_pipe_cycles"] / cyc,
            issue_slot_active=m["instructions"] / cyc, source=rf["source"])
        out["register_file_bound"] = dict(
            reads_per_pair_step_even_odd=m["rf_reads_even_odd"], floor_cycles_per_pair_step=max(m["rf_reads_even_odd"]),
            frac_of_rf_floor=max(m["rf_reads_even_odd"]) / cyc,
            rule="each SM sub-partition reads one 32-bit register per cycle from its even and one from its odd bank, for "
                 "all pipes together (profiles/ubench_r2.txt): a three-source FFMA2 issues every 3 cycles, not 2")
    nc = _profile_json("r2_estep_ncu.json")
    if nc:
        out["ncu"] = nc
    out["note"] = ("no tensor cores: the contraction is 5-D.  The kernel is bound by register-file read bandwidth "
                   "(register_file_bound), not by the FMA pipe: see DESIGN.md section 5")
    return out


def run(S, args, cfg):
    """Returns the fields of the headline line plus the live objects the later legs reuse."""
    torch = S.torch
    from sdmm_mitsuba_b200.mixture import SDMMCells
    K, N, iters = cfg["K"], cfg["N"], cfg["iters"]
    C_local = cfg["cells"]
    C_total = C_local * S.world
    first = S.rank * C_local
    cells = SDMMCells(C_total, K, device=S.local)
    joint, em, scal = data.model_cells(0, C_local, K, data.model_seed(S.rank))
    cells.upload(joint, em, scal, first_cell=first)
    point6, weight = data.device_batch_from_host(torch, C_local, N, data.batch_seed(S.rank), S.device)
    n_local = point6.shape[1]
    off = (np.arange(C_local + 1, dtype=np.uint64) * N).astype(np.uint32)
    torch.cuda.synchronize()
    exchange = attach_exchange(S, cells, args.exchange)
    borrowed = args.binding == "borrowed"
    overlap = args.overlap if args.overlap is not None else (S.world > 1)
    cells.set_em_overlap(overlap)

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
    S.barrier()
    from .clocks import ClockSampler
    sampler = ClockSampler(S.local)
    if S.rank == 0:
        sampler.start()
    l0 = cells.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    S.barrier()
    ms_local = e0.elapsed_time(e1)
    peer_timeouts = cells.peer_status()[1] if exchange == "peer" else None
    if peer_timeouts:
        raise SystemExit("bench: %d peer-exchange timeouts on rank %d: the records the other ranks hold are stale"
                         % (peer_timeouts, S.rank))
    launches = cells.launch_count() - l0
    clocks = sampler.stop() if S.rank == 0 else None
    if overlap:
        # per-launch E-step durations need launches that do not queue behind each other: two more (untimed)
        # steps with the in-order schedule, same kernels on the same data, CUDA events around every E-step
        cells.set_em_overlap(False)
        for _ in range(2):
            step()
        S.barrier()
        cells.set_em_overlap(True)
    estep_ms, estep_n = cells.last_estep_ms()
    ms_per_step = S.maxreduce(ms_local) / args.steps
    updates_per_step = float(C_total) * K * N * iters
    value = updates_per_step / (ms_per_step * 1e-3)
    est_ms = estep_ms / max(1, estep_n)
    roofline = estep_roofline(S, K, est_ms, C_local * N, borrowed,
                              estep_ms / ms_per_step if (S.world == 1 and not overlap) else None,
                              clocks["sm_mhz"] if clocks else None)
    # per-cell status words after all the stepwise iterations of the run (aborted / CDF invalid / non-PSD)
    status = cells.status(first_cell=first, n_cells=C_local)
    bad = int((status != 0).sum())
    bad_total = int(S.maxreduce(float(bad)))
    config = dict(
        workload="%s: batched StepwiseTangentEM (sdmm::em_step), %d cells x %d components x %d 5-D samples/cell, "
                 "%d EM iterations per step%s"
                 % (args.config, C_total, K, N, iters,
                    (" (%d cells per GPU, cell-sharded, query records exchanged every step: %s)"
                     % (C_local, "peer-memory stores fused into the M-step kernel" if exchange == "peer"
                        else "NCCL all-gather")) if S.world > 1 else ""),
        exchange=exchange, peer_timeouts=peer_timeouts,
        schedule="two halves, E-steps alternating on two streams, M-steps (and the peer-memory push) on a high-priority "
                 "stream (sdmmb_set_em_overlap); roofline E-step durations from two extra in-order steps"
        if overlap else "in order on one stream (E-step durations: CUDA events around every E-step launch of the timed "
                        "region's last step)",
        sample_binding="borrowed SoA planes read in place by the E-step (sdmmb_set_samples_borrowed)"
        if borrowed else "packed 32-byte records (sdmmb_set_samples: one packing pass per step)",
        cells=C_total, components=K, samples_per_cell=N, em_iterations=iters,
        parallelism="cells sharded over %d GPU(s)" % S.world,
        l2_policy="inputs larger than L2 (%.0f MB of samples per GPU per iteration)" % ((28.0 if borrowed else 32.0) * n_local / 1e6),
        model_state="EM state carries over between steps (same work per step)",
        samples="seeded numpy stream in blocks of %d cells (benchlegs/data.py); the cpu_baseline leg, the parity block and "
                "--impl reference fit cells of the same stream" % data.BLOCK_CELLS,
        cells_with_nonzero_status=bad_total)
    return dict(value=value, ms_per_step=ms_per_step, clocks=clocks, launches=int(launches), roofline=roofline,
                config=config, updates_per_step=updates_per_step,
                live=dict(cells=cells, point6=point6, weight=weight, off=off, first=first, C_local=C_local,
                          C_total=C_total, n_local=n_local, exchange=exchange, joint0=joint, em0=em, scal0=scal))
