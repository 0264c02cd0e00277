"""Synthetic sample batches (8 lobes per cell in [0,1]^3 x S^2, log-normal weights).

cfg2 (the headline) comes from ONE seeded numpy stream, cut into blocks of BLOCK_CELLS cells: block b of a
batch seeded `seed` is synth.make_samples(BLOCK_CELLS, N, seed * 100003 + b).  Every arm -- the GPU arm, its
cpu_baseline leg, the parity block and `--impl reference` in its own process -- regenerates exactly the
cells it needs from that rule, so all of them fit IDENTICAL samples.  cfg5 (537 M samples) is generated on
the device by the torch generator of the same family (numpy would take minutes); its cpu/parity legs read
the planes back from the device."""
import numpy as np

BLOCK_CELLS = 256


def batch_seed(rank):
    return 77 + 1000 * rank


def model_seed(rank):
    return 1000 + rank


def host_block(b, N, seed):
    from sdmm_mitsuba_b200 import synth
    p6, w, _ = synth.make_samples(BLOCK_CELLS, N, seed=seed * 100003 + b)
    return p6, w


def host_cells(first, count, N, seed):
    """Cells [first, first + count) of the batch seeded `seed`: (point6 [6, count * N], weight)."""
    b0, b1 = first // BLOCK_CELLS, (first + count - 1) // BLOCK_CELLS
    p, w = [], []
    for b in range(b0, b1 + 1):
        pb, wb = host_block(b, N, seed)
        lo = max(first, b * BLOCK_CELLS) - b * BLOCK_CELLS
        hi = min(first + count, (b + 1) * BLOCK_CELLS) - b * BLOCK_CELLS
        p.append(pb[:, lo * N:hi * N])
        w.append(wb[lo * N:hi * N])
    return np.ascontiguousarray(np.concatenate(p, axis=1)), np.ascontiguousarray(np.concatenate(w))


def model_cells(first, count, K, seed):
    """Initial mixtures + EM state of cells [first, first + count) of the model seeded `seed` (same block rule)."""
    from sdmm_mitsuba_b200 import synth
    b0, b1 = first // BLOCK_CELLS, (first + count - 1) // BLOCK_CELLS
    parts = []
    for b in range(b0, b1 + 1):
        j, e, s = synth.make_initial_model(BLOCK_CELLS, K, seed=seed * 100003 + b)
        lo = max(first, b * BLOCK_CELLS) - b * BLOCK_CELLS
        hi = min(first + count, (b + 1) * BLOCK_CELLS) - b * BLOCK_CELLS
        parts.append((j[lo:hi], e[lo:hi], s[lo:hi]))
    return tuple(np.ascontiguousarray(np.concatenate([q[i] for q in parts])) for i in range(3))


def device_batch_from_host(torch, C, N, seed, device):
    """The whole batch on the device, uploaded block by block from the seeded numpy stream."""
    point6 = torch.empty((6, C * N), device=device, dtype=torch.float32)
    weight = torch.empty(C * N, device=device, dtype=torch.float32)
    for c0 in range(0, C, BLOCK_CELLS):
        n = min(BLOCK_CELLS, C - c0)
        p, w = host_cells(c0, n, N, seed)
        point6[:, c0 * N:(c0 + n) * N] = torch.from_numpy(p).to(device)
        weight[c0 * N:(c0 + n) * N] = torch.from_numpy(w).to(device)
    return point6, weight


def make_device_samples(torch, C, N, seed, device):
    """The same family generated ON the device (torch generator), SoA planes."""
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


def device_batch_generated(torch, C, N, seed, device, out_point6=None, out_weight=None, step_c=1024):
    """C cells generated on the device in slabs of step_c cells (bounded transient memory)."""
    point6 = out_point6 if out_point6 is not None else torch.empty((6, C * N), device=device, dtype=torch.float32)
    weight = out_weight if out_weight is not None else torch.empty(C * N, device=device, dtype=torch.float32)
    for c0 in range(0, C, step_c):
        n = min(step_c, C - c0)
        p, w = make_device_samples(torch, n, N, seed + c0, device)
        point6[:, c0 * N:(c0 + n) * N] = p
        weight[c0 * N:(c0 + n) * N] = w
        del p, w
    return point6, weight
