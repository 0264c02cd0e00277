#!/usr/bin/env python
"""bench.py -- headline benchmark of the SDMM hot path on B200.

  python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun)
  python bench.py --impl reference ...                     (the reference's CPU EM on the host cores)

Metric (BASELINE.json): EM sample-component updates/s on the synthetic batched stepwise tangent EM
  cfg2 = 4096 cells x 64 components x 4096 5-D samples/cell, 4 EM iterations.
A "step" is one pass of the training hot path over that batch: bind the cell's sample planes and run
4 x em_step + the conditioner prepare.
  value : samples already in HBM when the timed region starts (CUDA events, max over ranks)
  e2e   : the same through the host-buffer API (pinned host samples -> H2D -> fit -> D2H of the
          fitted mixtures), wall clock around synchronised calls
With N > 1 every rank fits its own 4096 cells (weak scaling, cells are independent) and the refitted
query records reach every rank inside the timed region.  Every run also times BASELINE.json's scaling
configuration (cfg5: 65536 x 128 x 8192, cells sharded over the N ranks, strong) in the `cfg5` block, the
guided-query half of the metric in `queries`, and re-fits 16 cells of the timed batch against the
reference's own CPU EM in `parity`.  The legs live in benchlegs/ (one module per responsibility).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    "cfg2": dict(cells=4096, K=64, N=4096, iters=4),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="cfg2", choices=sorted(CONFIGS))
    ap.add_argument("--cells", type=int, default=0, help="override the headline's cell count (debug)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-render", action="store_true", help="skip the progressive-render leg")
    ap.add_argument("--no-cfg5", action="store_true", help="skip the cfg5 (65536 x 128 x 8192) block")
    ap.add_argument("--no-queries", action="store_true")
    ap.add_argument("--overlap", dest="overlap", action="store_true", default=None,
                    help="hide the M-steps (and the peer-memory push) behind E-steps (sdmmb_set_em_overlap).  Default: "
                         "on for N > 1, off for N = 1 so that the roofline's per-launch E-step durations come from "
                         "the timed region itself")
    ap.add_argument("--no-overlap", dest="overlap", action="store_false")
    ap.add_argument("--binding", default="borrowed", choices=["borrowed", "packed"],
                    help="how a step's device-resident samples are bound: zero-copy SoA planes or packed records")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="N>1: how the refitted query records reach the other GPUs")
    ap.add_argument("--queries", type=int, default=0, help="vertices of the guided-query leg (default 2^24 = cfg3)")
    return ap.parse_args()


def run_b200(args, cfg):
    from benchlegs import cfg5_leg, cpu_ref, e2e_leg, em_leg, query_leg
    S = em_leg.Session()
    head = em_leg.run(S, args, cfg)
    live = head.pop("live")
    K, N, iters = cfg["K"], cfg["N"], cfg["iters"]
    queries = None if args.no_queries else query_leg.run(S, args, cfg, live)
    e2e = None if args.no_e2e else e2e_leg.run(S, args, cfg, live, head["updates_per_step"])
    cb = parity = None
    if S.rank == 0 and not args.no_cpu_baseline:
        # the reference's CPU EM beside it (rank 0's host cores, thread count set explicitly), and the parity
        # block: 16 cells of the timed batch re-fitted from the fresh model, GPU against oracle/_ref
        import numpy as np
        n_par = min(16, live["C_local"])
        p6 = live["point6"][:, :n_par * N].cpu().numpy()
        w = live["weight"][:n_par * N].cpu().numpy()
        parity = cpu_ref.parity_block(K, N, iters, np.ascontiguousarray(p6), np.ascontiguousarray(w),
                                      live["joint0"][:n_par], live["em0"][:n_par], live["scal0"][:n_par], S.local, n_par)
        if S.world == 1:
            cb, _ = cpu_ref.cpu_baseline(K, N, iters)
            one, _ = cpu_ref.cpu_baseline(K, N, iters, seconds_target=4.0, n_threads=1)
            cb["single_core"] = dict(value=one["value"], unit=one["unit"], sample=one["sample"])
            if queries is not None:
                queries["cpu_baseline"] = cpu_ref.cpu_query_baseline(K, n_threads=cb["cores"])
    S.barrier()
    live["cells"].close()
    del live
    S.torch.cuda.empty_cache()
    cfg5 = None if args.no_cfg5 else cfg5_leg.run(S, args)
    render = refbench = None
    if S.rank == 0 and S.world == 1 and not args.no_render:
        from benchlegs.render_leg import run_render
        S.torch.cuda.empty_cache()       # the cfg5 leg's blocks go back to the driver before the tracer allocates its own
        render = run_render(S.torch, cpu=not args.no_cpu_baseline)
    if S.rank == 0 and S.world == 1 and not args.no_cpu_baseline:
        from benchlegs import refbench_leg
        refbench = refbench_leg.run(S.torch, S.local)
    if S.rank == 0:
        line = dict(
            metric="EM sample-component updates/s", value=head["value"], unit="updates/s", n_gpus=S.world,
            steps=args.steps, warmup=args.warmup, ms_per_step=head["ms_per_step"], higher_is_better=True,
            scaling="weak", vs_baseline=None, dtype="f32", data="synthetic", config=head["config"],
            clocks=head["clocks"], e2e=e2e, gpu_launches=head["launches"], roofline=head["roofline"],
            cpu_baseline=cb, parity=parity, cfg5=cfg5, queries=queries, progressive_render=render,
            reference_benchmark_shape=refbench)
        print(json.dumps(line), flush=True)
    S.close()


def main():
    args = parse_args()
    cfg = dict(CONFIGS[args.config])
    if args.cells:
        cfg["cells"] = args.cells
    if args.impl == "reference":
        from benchlegs import cpu_ref
        cpu_ref.run_reference(args, cfg, int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")))
        return
    run_b200(args, cfg)


if __name__ == "__main__":
    main()
