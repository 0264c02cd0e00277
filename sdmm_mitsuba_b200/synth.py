"""Seeded synthetic inputs for the SDMM hot path (SURVEY.md §8(d) "Synthetic inputs").

Everything is float32 and laid out the way the C ABI (include/sdmm_b200.h) takes it:
  samples : point6 [6, Ntot] SoA planes (x,y,z,dx,dy,dz), weight [Ntot], cell_offsets [C+1] u32
  model   : joint [C,32,K], em [C,81,K], scal [C,8] u32 words
  queries : cell_idx [n] u32, pos3 [3,n], u3 [3,n] (u by role: select, Box-Muller radius, angle)
"""
import numpy as np

JOINT_PLANES = 32
EM_PLANES = 81
SCALAR_WORDS = 8


def _unit(v):
    return v / np.linalg.norm(v, axis=-1, keepdims=True)


def _frame(n):
    """Any orthonormal (s, t) per row of n; only used to place synthetic lobes."""
    a = np.where(np.abs(n[..., :1]) < 0.9, np.array([1.0, 0, 0]), np.array([0, 1.0, 0]))
    s = _unit(np.cross(n, a))
    t = np.cross(n, s)
    return s, t


def make_samples(n_cells, n_per_cell, seed=0, n_lobes=8, dtype=np.float32):
    """Ground-truth generator per cell: n_lobes random spatio-directional lobes in [0,1]^3 x S^2,
    log-normal sample weights clamped to [1e-3, 1e3]; directions renormalised in float32."""
    rng = np.random.default_rng(seed)
    C, N, G = n_cells, n_per_cell, n_lobes
    centre = rng.uniform(0.1, 0.9, (C, G, 3))
    pos_sigma = rng.uniform(0.02, 0.15, (C, G, 1))
    dir_mean = _unit(rng.normal(size=(C, G, 3)))
    dir_sigma = rng.uniform(0.05, 0.5, (C, G, 1))
    lobe_w = rng.uniform(0.2, 1.0, (C, G))
    lobe_w /= lobe_w.sum(1, keepdims=True)
    cum = np.cumsum(lobe_w, 1)
    u = rng.uniform(size=(C, N))
    lobe = np.minimum((u[..., None] >= cum[:, None, :]).sum(-1), G - 1)           # [C,N]
    ci = np.arange(C)[:, None]
    x = centre[ci, lobe] + pos_sigma[ci, lobe] * rng.normal(size=(C, N, 3))
    x = np.clip(x, 0.0, 1.0)
    n = dir_mean[ci, lobe]
    s, t = _frame(n)
    tt = dir_sigma[ci, lobe] * rng.normal(size=(C, N, 2))
    ln = np.linalg.norm(tt, axis=-1, keepdims=True)
    ln = np.minimum(ln, np.pi - 1e-3)
    sinc = np.where(ln < 1e-6, 1.0, np.sin(ln) / np.maximum(ln, 1e-30))
    tt_unit_scale = np.where(ln < 1e-6, 1.0, np.minimum(1.0, (np.pi - 1e-3) / np.maximum(np.linalg.norm(tt, axis=-1, keepdims=True), 1e-30)))
    tt = tt * tt_unit_scale
    d = s * (tt[..., :1] * sinc) + t * (tt[..., 1:2] * sinc) + n * np.cos(ln)
    d = d.astype(np.float32)
    d /= np.sqrt((d * d).sum(-1, keepdims=True, dtype=np.float32))                # unit in float32
    w = np.clip(np.exp(rng.normal(size=(C, N))), 1e-3, 1e3)
    point6 = np.empty((6, C * N), dtype)
    point6[0:3] = x.reshape(C * N, 3).T
    point6[3:6] = d.reshape(C * N, 3).T
    weight = w.reshape(C * N).astype(dtype)
    offsets = (np.arange(C + 1, dtype=np.uint64) * N).astype(np.uint32)
    return np.ascontiguousarray(point6), weight, offsets


def make_normals(n_cells, n_per_cell, seed=0, n_patches=6, jitter=0.05):
    """Surface normals for sdmm::initialize inputs: each sample belongs to one of n_patches flat-ish
    patches per cell (patch normal + small jitter), unit length in float32.  Returns [3, C*N]."""
    rng = np.random.default_rng(seed + 7919)
    C, N, G = n_cells, n_per_cell, n_patches
    pn = _unit(rng.normal(size=(C, G, 3)))
    which = rng.integers(0, G, (C, N))
    n = pn[np.arange(C)[:, None], which] + jitter * rng.normal(size=(C, N, 3))
    n = n.astype(np.float32)
    n /= np.sqrt((n * n).sum(-1, keepdims=True, dtype=np.float32))
    return np.ascontiguousarray(n.reshape(C * N, 3).T)


def make_surface_samples(n_cells, n_per_cell, seed=0, n_patches=4, dtype=np.float32):
    """Renderer-like training samples for sdmm::initialize + the first fits: positions ON a few planar
    patches per cell (tiny noise along the normal), the patch normal as the sample normal, directions from
    two lobes in the patch's upper hemisphere, log-normal weights.  Returns point6 [6,C*N], normal3 [3,C*N],
    weight [C*N], offsets [C+1]."""
    rng = np.random.default_rng(seed + 104729)
    C, N, G = n_cells, n_per_cell, n_patches
    centre = rng.uniform(0.25, 0.75, (C, G, 3))
    pn = _unit(rng.normal(size=(C, G, 3)))
    ps, pt = _frame(pn)
    extent = rng.uniform(0.08, 0.25, (C, G, 2))
    which = rng.integers(0, G, (C, N))
    ci = np.arange(C)[:, None]
    ab = rng.uniform(-1, 1, (C, N, 2)) * extent[ci, which]
    n = pn[ci, which]
    x = centre[ci, which] + ab[..., :1] * ps[ci, which] + ab[..., 1:] * pt[ci, which] + 0.003 * rng.normal(size=(C, N, 1)) * n
    x = np.clip(x, 0.0, 1.0)
    # two directional lobes per patch, tilted away from the normal by up to ~60 degrees
    lobe_dir = _unit(pn[:, :, None, :] + 0.9 * rng.normal(size=(C, G, 2, 3)) * 0.6)
    lobe = rng.integers(0, 2, (C, N))
    m = lobe_dir[ci, which, lobe]
    d = _unit(m + 0.25 * rng.normal(size=(C, N, 3)))
    flip = (d * n).sum(-1, keepdims=True) < 0.05          # keep directions in the upper hemisphere
    d = np.where(flip, _unit(d - 2 * (d * n).sum(-1, keepdims=True) * n + 0.1 * n), d)
    d = d.astype(np.float32)
    d /= np.sqrt((d * d).sum(-1, keepdims=True, dtype=np.float32))
    nn = n.astype(np.float32)
    nn /= np.sqrt((nn * nn).sum(-1, keepdims=True, dtype=np.float32))
    w = np.clip(np.exp(rng.normal(size=(C, N))), 1e-3, 1e3)
    point6 = np.empty((6, C * N), dtype)
    point6[0:3] = x.reshape(C * N, 3).T
    point6[3:6] = d.reshape(C * N, 3).T
    normal3 = np.ascontiguousarray(nn.reshape(C * N, 3).T)
    offsets = (np.arange(C + 1, dtype=np.uint64) * N).astype(np.uint32)
    return np.ascontiguousarray(point6), normal3, w.reshape(C * N).astype(dtype), offsets


def make_initial_model(n_cells, K, seed=1):
    """Initial mixtures + EM state as the integrator sets them (sdmm/opt/init.h:178-192 priors):
    position means ~U[0,1]^3, direction means uniform on S^2, cov = 0.05 I_5, pmf = 1/K,
    weight_prior = 1/K, cov_prior_strength = 5/K, cov_prior = diag(2e-3 x3, 2e-4 x2),
    depth_prior = 0, learning_rate 0.2, alpha 0.5, iterations_run = 0."""
    rng = np.random.default_rng(seed)
    C = n_cells
    joint = np.zeros((C, JOINT_PLANES, K), np.float32)
    joint[:, 0, :] = 1.0 / K
    joint[:, 1:4, :] = rng.uniform(0, 1, (C, 3, K))
    d = _unit(rng.normal(size=(C, K, 3))).astype(np.float32)
    d /= np.sqrt((d * d).sum(-1, keepdims=True, dtype=np.float32))
    joint[:, 4:7, :] = d.transpose(0, 2, 1)
    for i in range(5):
        joint[:, 7 + 6 * i, :] = 0.05
    em = np.zeros((C, EM_PLANES, K), np.float32)
    for i, v in enumerate([2e-3, 2e-3, 2e-3, 2e-4, 2e-4]):
        em[:, 31 + 6 * i, :] = v
    scal = np.zeros((C, SCALAR_WORDS), np.uint32)
    f = scal.view(np.float32)
    f[:, 3] = 1.0 / K
    f[:, 4] = 5.0 / K
    f[:, 5] = 0.2
    f[:, 6] = 0.5
    return joint, em, scal


def make_queries(n, n_cells, seed=2, binned=True):
    """cfg3 queries: cell ~ U{0..C-1}, x ~ U[0,1]^3, u ~ U[0,1)^3.  binned=True sorts by cell."""
    rng = np.random.default_rng(seed)
    cell = rng.integers(0, n_cells, n).astype(np.uint32)
    if binned:
        cell.sort()
    pos3 = rng.uniform(0, 1, (3, n)).astype(np.float32)
    u3 = rng.uniform(0, 1, (3, n)).astype(np.float32)
    u3 = np.minimum(u3, np.float32(1.0 - 2 ** -24))
    return cell, np.ascontiguousarray(pos3), np.ascontiguousarray(u3)


def make_reference_benchmark(n_points=1_000_000, K=64):
    """The input of the reference's own EM micro-benchmark (sdmm/benchmarks/em.cpp:99-157): ONE cell, K = 64
    components that all sit at mean (1, 1, 1 | 1, 0, 0) with cov = I and pmf = 1/K, and N = 10^6 identical samples
    (1, 1, 1 | 1, 0, 0) of weight 1 -- every sample lies exactly on every component's mean (theta = 0, the
    `sin < 1e-4 -> 1` guard of spaces/spatio_directional.h:105-108); priors as em.cpp:124-149.
    Returns point6 [6, N], weight [N], offsets [2], joint, em, scal."""
    point6 = np.zeros((6, n_points), np.float32)
    point6[0:4] = 1.0
    weight = np.ones(n_points, np.float32)
    off = np.array([0, n_points], np.uint32)
    joint = np.zeros((1, JOINT_PLANES, K), np.float32)
    joint[:, 0, :] = 1.0 / K
    joint[:, 1:5, :] = 1.0
    for i in range(5):
        joint[:, 7 + 6 * i, :] = 1.0
    em = np.zeros((1, EM_PLANES, K), np.float32)
    for i, v in enumerate([2e-3, 2e-3, 2e-3, 2e-4, 2e-4]):
        em[:, 31 + 6 * i, :] = v
    scal = np.zeros((1, SCALAR_WORDS), np.uint32)
    f = scal.view(np.float32)
    f[:, 3] = 1.0 / K
    f[:, 4] = 5.0 / K
    f[:, 5] = 0.2
    f[:, 6] = 0.5
    return point6, weight, off, joint, em, scal
