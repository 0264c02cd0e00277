"""Record exchange over peer memory fused into the M-step (sdmmb_peer_*): two contexts of ONE process on
one GPU stand in for two ranks (same-process handles carry plain device pointers, so the whole protocol
-- ready flags, CTA-level stores into the peer's array, last-CTA arrival flag, device-side wait -- runs
as it does across GPUs over CUDA IPC; tools/mgpu_check.py covers the cross-process NVLink case).
Bit-exact bar: every rank must end with the records a single context fitting all cells produces."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ranks(C_total, K, N, world, seed=3):
    import torch
    from sdmm_mitsuba_b200 import sharding, synth
    from sdmm_mitsuba_b200.mixture import SDMMCells
    point6, weight, off = synth.make_samples(C_total, N, seed=seed)
    joint, em, scal = synth.make_initial_model(C_total, K, seed=seed + 1)
    ranks = []
    for r in range(world):
        first, n = sharding.cell_range(r, world, C_total)
        cells = SDMMCells(C_total, K)
        cells.upload(joint[first:first + n], em[first:first + n], scal[first:first + n], first_cell=first)
        lo, hi = int(off[first]), int(off[first + n])
        cells.set_samples(point6[:, lo:hi].copy(), weight[lo:hi].copy(), off[first:first + n + 1] - off[first],
                          first_cell=first)
        ranks.append((cells, torch.cuda.Stream()))
    whole = SDMMCells(C_total, K)
    whole.upload(joint, em, scal)
    whole.set_samples(point6, weight, off)
    torch.cuda.synchronize()
    return ranks, whole


def _attach(ranks):
    handles = [c.peer_export() for c, _ in ranks]
    for r, (c, _) in enumerate(ranks):
        c.peer_attach(r, handles)


@pytest.mark.parametrize("K,world,C_total", [(64, 2, 16), (16, 4, 36), (128, 2, 6)])
def test_peer_push_matches_single_context(K, world, C_total):
    import torch
    ranks, whole = _ranks(C_total, K, 512, world)
    _attach(ranks)
    for step in range(3):  # three exchanges (two epochs each): the ready/arrived handshake is reused
        for cells, st in ranks:
            with torch.cuda.stream(st):
                cells.em_step(2, with_conditioner=True)
        for cells, st in ranks:
            with torch.cuda.stream(st):
                cells.wait_peers()
        whole.em_step(2, with_conditioner=True)
        torch.cuda.synchronize()
        ref = whole.query_records()[:C_total]
        assert bool((ref != 0).any())
        for r, (cells, _) in enumerate(ranks):
            ep, timeouts = cells.peer_status()
            assert (ep, timeouts) == (2 * (step + 1), 0)   # every record-writing call consumes two epochs
            assert torch.equal(cells.query_records()[:C_total], ref), "rank %d step %d" % (r, step)


def test_peer_push_from_prepare_conditioner():
    """sdmmb_prepare_conditioner (K3 alone) pushes too."""
    import torch
    ranks, whole = _ranks(8, 32, 256, 2)
    _attach(ranks)
    from sdmm_mitsuba_b200 import sharding
    for r, (cells, st) in enumerate(ranks):
        first, n = sharding.cell_range(r, 2, 8)
        with torch.cuda.stream(st):
            cells.em_step(1, with_conditioner=False)
            cells.prepare(first, n)
    for cells, st in ranks:
        with torch.cuda.stream(st):
            cells.wait_peers()
    whole.em_step(1, with_conditioner=False)
    whole.prepare(0, 8)
    torch.cuda.synchronize()
    for cells, _ in ranks:
        assert cells.peer_status() == (2, 0)
        assert torch.equal(cells.query_records()[:8], whole.query_records()[:8])


def test_peer_wait_is_bounded_when_a_rank_never_arrives():
    """A rank whose peer never enters the step must not hang: the bounded wait times out, the stores to
    that peer are skipped and the timeout is reported."""
    import torch
    old = os.environ.get("SDMMB_PEER_TIMEOUT_MS")
    os.environ["SDMMB_PEER_TIMEOUT_MS"] = "20"
    try:
        ranks, _ = _ranks(4, 16, 128, 2)
        _attach(ranks)
    finally:
        if old is None:
            os.environ.pop("SDMMB_PEER_TIMEOUT_MS")
        else:
            os.environ["SDMMB_PEER_TIMEOUT_MS"] = old
    (a, sa), (b, _) = ranks
    before = b.query_records().clone()
    with torch.cuda.stream(sa):
        a.em_step(1, with_conditioner=True)   # b never signals ready
        a.wait_peers()                        # ... and never pushes
    torch.cuda.synchronize()
    ep, timeouts = a.peer_status()
    assert ep == 2 and timeouts >= 2
    assert torch.equal(b.query_records(), before)
    # the receiving rank is told that the records it holds for a's cells are stale (one count per failed step)
    assert b.peer_status()[1] == 1
    a.detach_peers()
    assert a.peer_status() == (0, 0)


def test_peer_push_with_overlapped_schedule():
    """em_step's overlapped schedule pushes each half from the (single, in-order) M-step stream: two exchange
    epochs per call, same records."""
    import torch
    ranks, whole = _ranks(24, 64, 700, 2)
    _attach(ranks)
    for cells, _ in ranks:
        cells.set_em_overlap(True)
    for step in range(2):
        for cells, st in ranks:
            with torch.cuda.stream(st):
                cells.em_step(2, with_conditioner=True)
        for cells, st in ranks:
            with torch.cuda.stream(st):
                cells.wait_peers()
        whole.em_step(2, with_conditioner=True)
        torch.cuda.synchronize()
        for cells, _ in ranks:
            assert cells.peer_status() == (2 * (step + 1), 0)
            assert torch.equal(cells.query_records()[:24], whole.query_records()[:24])


def test_peer_epochs_stay_in_step_when_ranks_pick_different_schedules():
    """One rank fits with the overlapped schedule (two pushes per call), the other in order (one push: what the
    overlapped schedule also falls back to for a shard of one cell or one item).  Both consume two epochs per
    call, so ready/arrived flags keep matching over several steps and nobody waits for an epoch that never
    comes (the epoch counters used to drift apart)."""
    import torch
    ranks, whole = _ranks(24, 64, 700, 2)
    _attach(ranks)
    ranks[0][0].set_em_overlap(True)
    ranks[1][0].set_em_overlap(False)
    for step in range(3):
        for cells, st in ranks:
            with torch.cuda.stream(st):
                cells.em_step(2, with_conditioner=True)
        for cells, st in ranks:
            with torch.cuda.stream(st):
                cells.wait_peers()
        whole.em_step(2, with_conditioner=True)
        torch.cuda.synchronize()
        for cells, _ in ranks:
            assert cells.peer_status() == (2 * (step + 1), 0)
            assert torch.equal(cells.query_records()[:24], whole.query_records()[:24])
