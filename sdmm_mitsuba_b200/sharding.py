"""Host-side multi-GPU plumbing of the hot path: cells are independent units (the reference fits them
from a thread pool, mitsuba/src/integrators/sdmm/volpath_sdmm.cpp:287-311), so they are partitioned
into contiguous slot ranges, one per rank, with no collective inside the fit; after a training
iteration the refitted QUERY RECORDS are all-gathered so that every rank can answer queries for any
cell.  Works on CUDA tensors over NCCL (the product path) and on CPU tensors over gloo (tests)."""
import numpy as np


def slots_per_rank(n_cells, world):
    """Equal contiguous slot ranges; the last rank's range may be partly unused."""
    return (int(n_cells) + world - 1) // world


def cell_range(rank, world, n_cells):
    """(first_slot, n_cells_of_this_rank) of rank `rank`."""
    per = slots_per_rank(n_cells, world)
    first = min(rank * per, n_cells)
    return first, max(0, min(per, n_cells - first))


def balance_by_samples(sample_counts, world):
    """For the real renderer, where cells hold 8 ... 2 M samples (volpath_sdmm.cpp:143-148, 371):
    an ORDER of the cells (a permutation) such that cutting it into `world` equal-length contiguous
    slot ranges gives every rank about the same number of samples.  Greedy longest-processing-time
    into `world` bins of equal cardinality.  Returns (order, per_rank_sample_totals)."""
    counts = np.asarray(sample_counts, dtype=np.int64)
    n = counts.shape[0]
    per = slots_per_rank(n, world)
    bins = [[] for _ in range(world)]
    tot = np.zeros(world, dtype=np.int64)
    for c in np.argsort(-counts, kind="stable"):
        open_bins = [r for r in range(world) if len(bins[r]) < per]
        r = min(open_bins, key=lambda b: (tot[b], b))
        bins[r].append(int(c))
        tot[r] += counts[c]
    order = [c for b in bins for c in sorted(b)]
    return np.asarray(order, dtype=np.int64), tot


def allgather_records(records, cells_per_rank, group=None):
    """In-place all-gather of equal contiguous slices of `records` ([slots, K, 32], slots >=
    world * cells_per_rank): rank r contributes records[r*cells_per_rank : (r+1)*cells_per_rank]."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if records.shape[0] < world * cells_per_rank:
        raise ValueError("records hold %d slots, need %d" % (records.shape[0], world * cells_per_rank))
    flat = records[: world * cells_per_rank].reshape(-1)
    per = flat.numel() // world
    mine = flat[rank * per:(rank + 1) * per]
    try:
        dist.all_gather_into_tensor(flat, mine, group=group)
    except (RuntimeError, NotImplementedError):
        # backends without the flat form: gather into views (copies for non-aliased outputs)
        outs = [flat[r * per:(r + 1) * per] for r in range(world)]
        tmp = [o if r != rank else mine.clone() for r, o in enumerate(outs)]
        dist.all_gather(tmp, mine.clone(), group=group)
        for r, (o, t) in enumerate(zip(outs, tmp)):
            if r != rank and o.data_ptr() != t.data_ptr():
                o.copy_(t)
    return records
