"""world_size-2 tests of the host-side multi-GPU logic on CPU (gloo): cell partitioning, the in-place
all-gather of query records, and bench.py's reference arm under a 2-rank launch."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, cells_per_rank, K, result_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, ROOT)
    from sdmm_mitsuba_b200 import sharding
    slots = world * cells_per_rank + 3      # a few unused trailing slots must stay untouched
    rec = torch.full((slots, K, 32), -1.0)
    first, n = sharding.cell_range(rank, world, world * cells_per_rank)
    assert (first, n) == (rank * cells_per_rank, cells_per_rank)
    rec[first:first + n] = torch.arange(n * K * 32, dtype=torch.float32).reshape(n, K, 32) + 1000.0 * (rank + 1)
    sharding.allgather_records(rec, cells_per_rank)
    np.save(os.path.join(result_dir, "rec%d.npy" % rank), rec.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_allgather_records_world2(tmp_path):
    world, cpr, K = 2, 5, 16
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(world, port, cpr, K, str(tmp_path)), nprocs=world, join=True)
    a = np.load(tmp_path / "rec0.npy")
    b = np.load(tmp_path / "rec1.npy")
    np.testing.assert_array_equal(a, b)
    for r in range(world):
        expect = np.arange(cpr * K * 32, dtype=np.float32).reshape(cpr, K, 32) + 1000.0 * (r + 1)
        np.testing.assert_array_equal(a[r * cpr:(r + 1) * cpr], expect)
    assert (a[world * cpr:] == -1).all()


def test_cell_range_covers_all_cells_once():
    from sdmm_mitsuba_b200 import sharding
    for n_cells in (1, 7, 8, 4096, 65536, 1001):
        for world in (1, 2, 4, 8):
            seen = np.zeros(n_cells, int)
            per = sharding.slots_per_rank(n_cells, world)
            for r in range(world):
                first, n = sharding.cell_range(r, world, n_cells)
                assert first == min(r * per, n_cells) and n <= per
                seen[first:first + n] += 1
            assert (seen == 1).all()


def test_balance_by_samples():
    from sdmm_mitsuba_b200 import sharding
    rng = np.random.default_rng(0)
    counts = np.minimum(2_000_000, (8 + rng.pareto(1.2, 2048) * 500)).astype(np.int64)   # heavy-tailed cells
    for world in (2, 4, 8):
        order, tot = sharding.balance_by_samples(counts, world)
        assert sorted(order.tolist()) == list(range(2048))
        per = sharding.slots_per_rank(2048, world)
        for r in range(world):
            assert counts[order[r * per:(r + 1) * per]].sum() == tot[r]
        assert tot.max() <= 1.05 * tot.mean() + counts.max()


def test_bench_reference_arm_two_ranks():
    """Under a 2-rank launch rank 0 alone runs the CPU reference arm and prints ONE JSON line."""
    # torchrun forces OMP_NUM_THREADS=1 on its workers: the arm must set its thread count itself
    env = dict(os.environ, OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(29600 + os.getpid() % 300),
           os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["value"] > 0
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["e2e"]["h2d_bytes_per_step"] == 0
    n_host = len(os.sched_getaffinity(0))
    assert d["cpu_baseline"]["cores"] == n_host, (d["cpu_baseline"]["cores"], n_host)
    if n_host > 1:
        assert d["cpu_baseline"]["cores"] > 1
