#!/usr/bin/env python
"""bench.py -- headline benchmark of the SDMM hot path on B200.

  python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun)
  python bench.py --impl reference ...                     (the reference's CPU EM on the host cores)

Metric (BASELINE.json): EM sample-component updates/s on the synthetic batched stepwise tangent EM
  cfg2 = 4096 cells x 64 components x 4096 5-D samples/cell, 4 EM iterations.
A "step" is one pass of the training hot path over that batch: bind the cell's sample planes
(device-resident SoA -> packed records) and run 4 x em_step + the conditioner prepare.
  value : samples already in HBM when the timed region starts (CUDA events, max over ranks)
  e2e   : the same through the host-buffer API (pinned host samples -> H2D -> fit -> D2H of the
          fitted mixtures), wall clock around synchronised calls
With N > 1 every rank fits its own 4096 cells (weak scaling, cells are independent) and the refitted
query records are all-gathered over NCCL after every step, inside the timed region.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOP_PER_UPDATE = 102.0   # algorithmic FP32 flop per sample-component update
SFU_PER_UPDATE = 4.0
CONFIGS = {
    "cfg2": dict(cells=4096, K=64, N=4096, iters=4),
    "cfg5": dict(cells=65536, K=128, N=8192, iters=1),   # total; sharded by cell over the ranks
}


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["no samples"])
        hi = [s for s in sm if s >= 0.5 * max(sm)]
        return dict(sm_mhz=float(np.median(hi)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons),
                    samples=len(sm))


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="cfg2", choices=sorted(CONFIGS))
    ap.add_argument("--cells", type=int, default=0, help="override the cell count (debug)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-render", action="store_true", help="skip the progressive-render leg")
    ap.add_argument("--overlap", action="store_true",
                    help="hide the M-steps behind E-steps (sdmmb_set_em_overlap; +0.8 %% at cfg2).  Off by default so "
                         "that the roofline's per-launch E-step durations come from the timed region itself")
    ap.add_argument("--binding", default="borrowed", choices=["borrowed", "packed"],
                    help="how a step's device-resident samples are bound: zero-copy SoA planes or packed records")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="N>1: how the refitted query records reach the other GPUs")
    ap.add_argument("--queries", type=int, default=0, help="also time the guided query kernel on this many vertices")
    return ap.parse_args()


def make_device_samples(torch, C, N, seed, device):
    """Synthetic lobe-mixture samples generated ON the device (same family as synth.make_samples:
    8 lobes per cell in [0,1]^3 x S^2, log-normal weights), SoA planes."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    G = 8
    f32 = torch.float32
    centre = torch.rand((C, G, 3), generator=g, device=device, dtype=f32) * 0.8 + 0.1
    pos_sigma = torch.rand((C, G, 1), generator=g, device=device, dtype=f32) * 0.13 + 0.02
    dmean = torch.randn((C, G, 3), generator=g, device=device, dtype=f32)
    dmean = dmean / dmean.norm(dim=-1, keepdim=True)
    dsig = torch.rand((C, G, 1), generator=g, device=device, dtype=f32) * 0.45 + 0.05
    lobe = torch.randint(0, G, (C, N), generator=g, device=device)
    ci = torch.arange(C, device=device)[:, None]
    x = centre[ci, lobe] + pos_sigma[ci, lobe] * torch.randn((C, N, 3), generator=g, device=device, dtype=f32)
    x = x.clamp_(0.0, 1.0)
    n = dmean[ci, lobe]
    a = torch.where(n[..., :1].abs() < 0.9, torch.tensor([1.0, 0, 0], device=device), torch.tensor([0, 1.0, 0], device=device))
    s = torch.linalg.cross(n, a)
    s = s / s.norm(dim=-1, keepdim=True)
    t = torch.linalg.cross(n, s)
    tt = dsig[ci, lobe] * torch.randn((C, N, 2), generator=g, device=device, dtype=f32)
    ln = tt.norm(dim=-1, keepdim=True).clamp_(1e-6, 3.0)
    tt = tt / tt.norm(dim=-1, keepdim=True).clamp_min(1e-6) * ln
    d = s * (tt[..., :1] * torch.sin(ln) / ln) + t * (tt[..., 1:2] * torch.sin(ln) / ln) + n * torch.cos(ln)
    d = d / d.norm(dim=-1, keepdim=True)
    w = torch.exp(torch.randn((C, N), generator=g, device=device, dtype=f32)).clamp_(1e-3, 1e3)
    point6 = torch.empty((6, C * N), device=device, dtype=f32)
    point6[0:3] = x.reshape(C * N, 3).t()
    point6[3:6] = d.reshape(C * N, 3).t()
    return point6.contiguous(), w.reshape(C * N).contiguous()


def cpu_baseline(K, N, iters, seconds_target=12.0, n_threads=0):
    """The reference's CPU EM (oracle/_ref when it was built from /root/reference, else the plain-C
    port) with an OpenMP parallel-for over cells -- ThreadPool::parallelFor's decomposition
    (volpath_sdmm.cpp:287-311) -- on a bounded sample of the same workload."""
    from oracle import portbind, refbind
    from sdmm_mitsuba_b200 import synth
    if refbind.available():
        mod, cls, kind = refbind, refbind.RefCells, "reference"
        cores = mod.lib().sdmm_ref_max_threads()
    else:
        mod, cls, kind = portbind, portbind.PortCells, "port"
        cores = mod.lib().oracle_max_threads()
    if n_threads:
        cores = n_threads

    def run(C):
        point6, weight, off = synth.make_samples(C, N, seed=123)
        joint, em, scal = synth.make_initial_model(C, K, seed=124)
        o = cls(C, K)
        o.set(joint, em, scal)
        return o.em_batch(off, point6, weight, iters, with_prepare=True, n_threads=n_threads)

    c0 = max(cores, 1)
    run(c0)                                                 # warm-up (thread pool, page faults)
    per_round = max(run(c0), 1e-3)                          # calibration: one cell per thread
    rounds = int(max(1, min(256, seconds_target / per_round)))
    C = c0 * rounds
    secs = run(C)
    updates = float(C) * K * N * iters
    return dict(value=updates / secs, unit="updates/s", cores=int(cores), kind=kind,
                sample="%d of the cells (K=%d, N=%d, %d EM iterations + conditioner prepare), %.1f s, "
                       "OpenMP over cells" % (C, K, N, iters, secs)), secs


def cpu_query_baseline(K, n_queries=1 << 20, n_cells=64, n_threads=0):
    """The reference's per-vertex create_conditional + SDMM::sample + posterior (sdmm_proc.cpp:361-372,
    419-421, 539-545) with OpenMP over vertices and per-thread scratch, on mixtures fitted by the same
    CPU code; a bounded sample (n_queries over n_cells cells) of cfg3."""
    from oracle import portbind, refbind
    from sdmm_mitsuba_b200 import synth
    if refbind.available():
        mod, cls, kind = refbind, refbind.RefCells, "reference"
        cores = mod.lib().sdmm_ref_max_threads()
    else:
        mod, cls, kind = portbind, portbind.PortCells, "port"
        cores = mod.lib().oracle_max_threads()
    if n_threads:
        cores = n_threads
    point6, weight, off = synth.make_samples(n_cells, 1024, seed=321)
    joint, em, scal = synth.make_initial_model(n_cells, K, seed=322)
    o = cls(n_cells, K)
    o.set(joint, em, scal)
    o.em_batch(off, point6, weight, 2, with_prepare=True, n_threads=n_threads)
    o.prepare_conditioner()
    cell, pos3, u3 = synth.make_queries(n_queries, n_cells, seed=323, binned=False)
    o.sample_pdf(cell[:4096].copy(), np.ascontiguousarray(pos3[:, :4096]), np.ascontiguousarray(u3[:, :4096]), n_threads=n_threads)
    _, secs = o.sample_pdf(cell, pos3, u3, n_threads=n_threads)
    return dict(value=n_queries / secs, unit="queries/s", cores=int(cores), kind=kind,
                sample="%d vertices over %d fitted cells (K=%d), %.1f s, OpenMP over vertices" % (n_queries, n_cells, K, secs))


def run_reference(args, cfg, rank, world):
    if rank != 0:
        return
    K, N, iters = cfg["K"], cfg["N"], cfg["iters"]
    vals = []
    for i in range(args.warmup + args.steps):
        cb, secs = cpu_baseline(K, N, iters, seconds_target=3.0)
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
                                     "bounded sample of cells on the host cores" % (args.config, K, N, iters)),
                cpu_baseline=cb,
                e2e=dict(value=v, unit="updates/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line), flush=True)


from bench_b200 import run_b200  # noqa: E402  (second half of this file, split only for length)


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cfg = dict(CONFIGS[args.config])
    if args.cells:
        cfg["cells"] = args.cells
    if args.impl == "reference":
        run_reference(args, cfg, rank, world)
        return
    run_b200(args, cfg, rank, world, ClockSampler, make_device_samples, cpu_baseline,
             FLOP_PER_UPDATE, SFU_PER_UPDATE)


if __name__ == "__main__":
    main()
