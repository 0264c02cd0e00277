"""ctypes binding of oracle/_ref/libsdmm_ref.so -- the REFERENCE's own sdmm headers compiled
behind oracle/ref_driver.cpp.  TEST INFRASTRUCTURE ONLY: importable from tests/,
__graft_entry__.smoke() and bench.py's CPU-baseline legs, never from the product package.

Flat layouts are documented at the top of oracle/ref_driver.cpp (and include/sdmm_b200.h).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libsdmm_ref.so")

JOINT_PLANES = 32
EM_PLANES = 81
SCALAR_WORDS = 8
# enoki::PCG32 defaults (ext/enoki/include/enoki/random.h:33-35)
PCG32_DEFAULT_STATE = 0x853c49e6748fea9b
PCG32_DEFAULT_STREAM = 0xda3e39cb94b95bdb

_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref/libsdmm_ref.so missing: run `make -C oracle ref` where "
                               "/root/reference is mounted")
        L = C.CDLL(LIB_PATH)
        fp = C.POINTER(C.c_float)
        up = C.POINTER(C.c_uint32)
        bp = C.POINTER(C.c_uint8)
        vpp = C.POINTER(C.c_void_p)
        L.sdmm_ref_cell_create.restype = C.c_void_p
        L.sdmm_ref_cell_create.argtypes = [C.c_int]
        L.sdmm_ref_cell_destroy.argtypes = [C.c_void_p]
        L.sdmm_ref_cell_set.argtypes = [C.c_void_p, fp, fp, up]
        L.sdmm_ref_cell_get.argtypes = [C.c_void_p, fp, fp, up]
        L.sdmm_ref_cell_get_derived.argtypes = [C.c_void_p, fp]
        L.sdmm_ref_em_step.argtypes = [C.c_void_p, C.c_int, fp, C.c_long, fp, C.c_int, fp]
        L.sdmm_ref_initialize.argtypes = [C.c_void_p, C.c_int, fp, fp, C.c_long, fp, C.c_int, C.c_float,
                                          C.c_uint64, C.c_uint64]
        L.sdmm_ref_pcg32_floats.argtypes = [C.c_uint64, C.c_uint64, C.c_int, fp]
        L.sdmm_ref_em_batch.restype = C.c_double
        L.sdmm_ref_em_batch.argtypes = [vpp, C.c_int, up, fp, C.c_long, fp, C.c_int, C.c_int, C.c_int]
        L.sdmm_ref_prepare_conditioner.argtypes = [C.c_void_p]
        L.sdmm_ref_get_conditioner.argtypes = [C.c_void_p, fp]
        L.sdmm_ref_sample_pdf.restype = C.c_double
        L.sdmm_ref_sample_pdf.argtypes = [vpp, C.c_long, up, fp, fp, fp, fp, fp, up, bp, C.c_int]
        L.sdmm_ref_pdf.restype = C.c_double
        L.sdmm_ref_pdf.argtypes = [vpp, C.c_long, up, fp, fp, fp, bp, fp, fp, C.c_int]
        L.sdmm_ref_kat_pdf_scalar.restype = C.c_float
        L.sdmm_ref_kat_pdf_dynamic.argtypes = [fp]
        L.sdmm_ref_kat_conditioner.argtypes = [fp]
        L.sdmm_ref_kat_cholesky3.argtypes = [fp, fp]
        L.sdmm_ref_kat_categorical.argtypes = [C.c_int, fp, fp, fp]
        L.sdmm_ref_kat_frame.argtypes = [fp, fp]
        L.sdmm_ref_joint_to.argtypes = [fp, fp, fp, fp]
        L.sdmm_ref_tree_create.restype = C.c_void_p
        L.sdmm_ref_tree_create.argtypes = [fp, fp, C.c_long]
        L.sdmm_ref_tree_destroy.argtypes = [C.c_void_p]
        L.sdmm_ref_tree_split_to_depth.argtypes = [C.c_void_p, C.c_int]
        L.sdmm_ref_tree_push.restype = C.c_long
        L.sdmm_ref_tree_push.argtypes = [C.c_void_p, C.c_long, fp]
        L.sdmm_ref_tree_split_leaves.argtypes = [C.c_void_p, C.c_long]
        L.sdmm_ref_tree_size.argtypes = [C.c_void_p]
        L.sdmm_ref_tree_export.argtypes = [C.c_void_p, fp]
        L.sdmm_ref_tree_find.argtypes = [C.c_void_p, C.c_long, fp, C.POINTER(C.c_int32)]
        _lib = L
    return _lib


PRODUCT_LIB_PATH = os.path.join(os.path.dirname(LIB_PATH), "libsdmm_ref_product.so")
_plib = None


def product_available():
    return os.path.exists(PRODUCT_LIB_PATH)


def product_lib():
    """sdmm::product entry points; a separate shared object (see oracle/ref_product.cpp) that takes the RefCell
    handles created by lib()."""
    global _plib
    if _plib is None:
        if not product_available():
            raise RuntimeError("oracle/_ref/libsdmm_ref_product.so missing: run `make -C oracle ref`")
        L = C.CDLL(PRODUCT_LIB_PATH)
        fp, up, bp = C.POINTER(C.c_float), C.POINTER(C.c_uint32), C.POINTER(C.c_uint8)
        L.sdmm_ref_product_sample_pdf.restype = C.c_double
        L.sdmm_ref_product_sample_pdf.argtypes = [C.POINTER(C.c_void_p), C.c_long, up, fp, fp, fp, fp, fp, fp, C.c_int, fp,
                                                  fp, fp, fp, up, bp, C.c_int]
        L.sdmm_ref_product_components.argtypes = [C.c_void_p, fp, fp, C.c_int, fp, fp, C.c_int]
        L.sdmm_ref_pruned_sample_pdf.restype = C.c_double
        L.sdmm_ref_pruned_sample_pdf.argtypes = [C.POINTER(C.c_void_p), C.c_long, up, fp, fp, C.c_int, C.c_int, fp, fp, fp,
                                                 up, bp, fp]
        _plib = L
    return _plib


def _f(a):
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _u(a):
    assert a.dtype == np.uint32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.POINTER(C.c_uint32))


def _b(a):
    assert a.dtype == np.uint8 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


def _off(a, n):
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return C.cast(a.ctypes.data + 4 * n, C.POINTER(C.c_float))


class RefCells:
    """A batch of reference cells (one JointSDMM + EM + Conditioner each)."""

    def __init__(self, n_cells, K):
        self.C, self.K = n_cells, K
        L = lib()
        self.handles = (C.c_void_p * n_cells)(*[L.sdmm_ref_cell_create(K) for _ in range(n_cells)])

    def __del__(self):
        try:
            L = lib()
            for h in self.handles:
                L.sdmm_ref_cell_destroy(h)
        except Exception:
            pass

    def set(self, joint, em=None, scal=None):
        """joint [C,32,K] f32; em [C,81,K] f32; scal [C,8] u32 words.  Returns per-cell bool."""
        L = lib()
        ok = np.zeros(self.C, dtype=bool)
        for c in range(self.C):
            ok[c] = bool(L.sdmm_ref_cell_set(
                self.handles[c], _f(joint[c]),
                _f(em[c]) if em is not None else None,
                _u(scal[c]) if scal is not None else None))
        return ok

    def get(self):
        L = lib()
        joint = np.zeros((self.C, JOINT_PLANES, self.K), np.float32)
        em = np.zeros((self.C, EM_PLANES, self.K), np.float32)
        scal = np.zeros((self.C, SCALAR_WORDS), np.uint32)
        for c in range(self.C):
            L.sdmm_ref_cell_get(self.handles[c], _f(joint[c]), _f(em[c]), _u(scal[c]))
        return joint, em, scal

    def initialize(self, offsets, point6, normal3, weight, spatial_distance,
                   rng_state=PCG32_DEFAULT_STATE, rng_stream=PCG32_DEFAULT_STREAM):
        """sdmm::initialize per cell (opt/init.h:17-193) with n_spatial = K / 8 seeds, as
        initializeSDMMContext does (volpath_sdmm.cpp:132-137).  spatial_distance: per-cell array."""
        L = lib()
        stride = point6.shape[1]
        for c in range(self.C):
            b, e = int(offsets[c]), int(offsets[c + 1])
            k = L.sdmm_ref_initialize(self.handles[c], e - b, _off(point6, b), _off(normal3, b), stride,
                                      _off(weight, b), self.K // 8, float(spatial_distance[c]),
                                      rng_state, rng_stream)
            assert k == self.K, (k, self.K)

    def derived(self):
        L = lib()
        out = np.zeros((self.C, 38, self.K), np.float32)
        for c in range(self.C):
            L.sdmm_ref_cell_get_derived(self.handles[c], _f(out[c]))
        return out

    def em_step(self, offsets, point6, weight, n_iters, want_loglik=False):
        """point6 [6,Ntot] f32, weight [Ntot], offsets [C+1] u32.  Serial, per cell."""
        L = lib()
        ntot = point6.shape[1]
        ll = None
        if want_loglik:
            ll = [None] * self.C
        for c in range(self.C):
            b, e = int(offsets[c]), int(offsets[c + 1])
            out = np.zeros((n_iters, e - b), np.float32) if want_loglik else None
            L.sdmm_ref_em_step(self.handles[c], e - b, _off(point6, b), ntot, _off(weight, b),
                               n_iters, _f(out) if want_loglik else None)
            if want_loglik:
                ll[c] = out
        return ll

    def em_batch(self, offsets, point6, weight, n_iters, with_prepare=True, n_threads=0):
        """OpenMP over cells; returns seconds inside the em_step(+prepare) loop."""
        L = lib()
        return L.sdmm_ref_em_batch(C.cast(self.handles, C.POINTER(C.c_void_p)), self.C, _u(offsets),
                                   _f(point6), point6.shape[1], _f(weight), n_iters,
                                   1 if with_prepare else 0, n_threads)

    def prepare_conditioner(self):
        L = lib()
        for h in self.handles:
            L.sdmm_ref_prepare_conditioner(h)

    def conditioner(self):
        L = lib()
        out = np.zeros((self.C, 38, self.K), np.float32)
        for c in range(self.C):
            L.sdmm_ref_get_conditioner(self.handles[c], _f(out[c]))
        return out

    def sample_pdf(self, cell_idx, pos3, u3, n_threads=0):
        """pos3, u3: [3,n] f32 (u by role: select, radius, angle).  Returns dict + seconds."""
        L = lib()
        n = pos3.shape[1]
        out_dir = np.zeros((3, n), np.float32)
        out_pdf = np.zeros(n, np.float32)
        out_ij = np.zeros(n, np.float32)
        out_comp = np.zeros(n, np.uint32)
        out_valid = np.zeros(n, np.uint8)
        secs = L.sdmm_ref_sample_pdf(C.cast(self.handles, C.POINTER(C.c_void_p)), n, _u(cell_idx),
                                     _f(pos3), _f(u3), _f(out_dir), _f(out_pdf), _f(out_ij),
                                     _u(out_comp), _b(out_valid), n_threads)
        return dict(dir=out_dir, pdf=out_pdf, inv_jac=out_ij, comp=out_comp, valid=out_valid), secs

    def pdf(self, cell_idx, pos3, dir3, n_threads=0):
        L = lib()
        n = pos3.shape[1]
        out_pdf = np.zeros(n, np.float32)
        out_valid = np.zeros(n, np.uint8)
        secs = L.sdmm_ref_pdf(C.cast(self.handles, C.POINTER(C.c_void_p)), n, _u(cell_idx), _f(pos3),
                              _f(dir3), _f(out_pdf), _b(out_valid), None, None, n_threads)
        return out_pdf, out_valid, secs

    def product_sample_pdf(self, cell_idx, pos3, bsdf, u3=None, dir3=None, normal3=None, frame9=None, wi3=None,
                           n_threads=0):
        """sampleSurface's product branch (sdmm_proc.cpp:327-339, 375-381): bsdf [M,7] rows (pmf, mean3, c00, c10,
        c11); u3 by role (sample + pdf) or dir3 (pdf only); normal3 re-centres BSDF component 0 (diffuse branch);
        frame9 + wi3: the glossy branch (sdmm_proc.cpp:340-355)."""
        L = product_lib()
        n = pos3.shape[1]
        bsdf = np.ascontiguousarray(bsdf, np.float32).reshape(-1, 7)
        out_dir = np.zeros((3, n), np.float32)
        out_pdf = np.zeros(n, np.float32)
        out_ij = np.zeros(n, np.float32)
        out_comp = np.zeros(n, np.uint32)
        out_valid = np.zeros(n, np.uint8)
        opt = lambda a: _f(a) if a is not None else None
        secs = L.sdmm_ref_product_sample_pdf(C.cast(self.handles, C.POINTER(C.c_void_p)), n, _u(cell_idx), _f(pos3),
                                             opt(u3), opt(dir3), opt(normal3), opt(frame9), opt(wi3), bsdf.shape[0],
                                             _f(bsdf), _f(out_dir),
                                             _f(out_pdf), _f(out_ij), _u(out_comp), _b(out_valid), n_threads)
        return dict(dir=out_dir, pdf=out_pdf, inv_jac=out_ij, comp=out_comp, valid=out_valid), secs

    def pruned_sample_pdf(self, cell_idx, pos3, u3, max_components, preserve_idx=-1):
        """create_conditional_pruned (sdmm_conditioner.h:128-168) -> sample -> posterior, per vertex."""
        L = product_lib()
        n = pos3.shape[1]
        out = dict(dir=np.zeros((3, n), np.float32), pdf=np.zeros(n, np.float32), inv_jac=np.zeros(n, np.float32),
                   comp=np.zeros(n, np.uint32), valid=np.zeros(n, np.uint8), pmf=np.zeros((n, self.K), np.float32))
        L.sdmm_ref_pruned_sample_pdf(C.cast(self.handles, C.POINTER(C.c_void_p)), n, _u(cell_idx), _f(pos3), _f(u3),
                                     int(max_components), int(preserve_idx), _f(out["dir"]), _f(out["pdf"]),
                                     _f(out["inv_jac"]), _u(out["comp"]), _b(out["valid"]), _f(out["pmf"]))
        return out

    def product_components(self, cell, x, bsdf, normal=None):
        """The product mixture at one vertex: [P,7] rows (pmf, tangent_mean2, cov00, cov10, cov11, cdf)."""
        L = product_lib()
        bsdf = np.ascontiguousarray(bsdf, np.float32).reshape(-1, 7)
        out = np.zeros((self.K * bsdf.shape[0], 7), np.float32)
        xx = np.ascontiguousarray(x, np.float32)
        nn = np.ascontiguousarray(normal, np.float32) if normal is not None else None
        P = L.sdmm_ref_product_components(self.handles[cell], _f(xx), _f(nn) if nn is not None else None,
                                          bsdf.shape[0], _f(bsdf), _f(out), out.shape[0])
        return out[:P]

    def conditional_weights(self, cell, x):
        """Normalised conditional pmf/cdf at one vertex (for tie analysis)."""
        L = lib()
        pos = np.asarray(x, np.float32).reshape(3, 1).copy()
        d = np.zeros((3, 1), np.float32)
        d[2] = 1
        pmf = np.zeros(self.K, np.float32)
        cdf = np.zeros(self.K, np.float32)
        pdf = np.zeros(1, np.float32)
        valid = np.zeros(1, np.uint8)
        idx = np.array([cell], np.uint32)
        L.sdmm_ref_pdf(C.cast(self.handles, C.POINTER(C.c_void_p)), 1, _u(idx), _f(pos), _f(d),
                       _f(pdf), _b(valid), _f(pmf), _f(cdf), 1)
        return pmf, cdf, bool(valid[0])


def pcg32_floats(n, state=PCG32_DEFAULT_STATE, stream=PCG32_DEFAULT_STREAM):
    out = np.zeros(n, np.float32)
    lib().sdmm_ref_pcg32_floats(state, stream, n, _f(out))
    return out


def kat_pdf_scalar():
    return float(lib().sdmm_ref_kat_pdf_scalar())


def kat_pdf_dynamic():
    out = np.zeros(6, np.float32)
    lib().sdmm_ref_kat_pdf_dynamic(_f(out))
    return out


def kat_conditioner():
    out = np.zeros(12, np.float32)
    lib().sdmm_ref_kat_conditioner(_f(out))
    return out


def kat_cholesky3(m):
    m = np.ascontiguousarray(m, np.float32)
    out = np.zeros((3, 3), np.float32)
    psd = lib().sdmm_ref_kat_cholesky3(_f(m), _f(out))
    return out, bool(psd)


def kat_categorical(pmf):
    pmf = np.ascontiguousarray(pmf, np.float32)
    p = np.zeros_like(pmf)
    c = np.zeros_like(pmf)
    ok = lib().sdmm_ref_kat_categorical(len(pmf), _f(pmf), _f(p), _f(c))
    return p, c, bool(ok)


def kat_frame(n):
    n = np.ascontiguousarray(n, np.float32)
    out = np.zeros((3, 3), np.float32)
    lib().sdmm_ref_kat_frame(_f(n), _f(out))
    return out


def joint_to(mean6, point6):
    mean6 = np.ascontiguousarray(mean6, np.float32)
    point6 = np.ascontiguousarray(point6, np.float32)
    t = np.zeros(5, np.float32)
    ij = np.zeros(1, np.float32)
    lib().sdmm_ref_joint_to(_f(mean6), _f(point6), _f(t), _f(ij))
    return t, float(ij[0])


class RefTree:
    """The reference's STree (sdmm/accelerators/spatial_tree.h) over SDMMContext cells."""

    def __init__(self, aabb_min=(0, 0, 0), aabb_max=(1, 1, 1), capacity=100000):
        lo = np.asarray(aabb_min, np.float32)
        hi = np.asarray(aabb_max, np.float32)
        self.h = lib().sdmm_ref_tree_create(_f(lo), _f(hi), capacity)

    def __del__(self):
        try:
            lib().sdmm_ref_tree_destroy(self.h)
        except Exception:
            pass

    def split_to_depth(self, depth):
        lib().sdmm_ref_tree_split_to_depth(self.h, depth)

    def push(self, pos3):
        """pos3 [3,n]: find the leaf of every point, push it into the leaf's data + stats."""
        pos3 = np.ascontiguousarray(pos3, np.float32)
        return int(lib().sdmm_ref_tree_push(self.h, pos3.shape[1], _f(pos3)))

    def split_leaves(self, threshold):
        lib().sdmm_ref_tree_split_leaves(self.h, threshold)

    def nodes(self):
        """[n_nodes, 11]: is_leaf, axis, child0, child1, stats.size, aabb.min[3], aabb.max[3]."""
        n = lib().sdmm_ref_tree_size(self.h)
        out = np.zeros((n, 11), np.float32)
        lib().sdmm_ref_tree_export(self.h, _f(out))
        return out

    def find(self, pos3):
        pos3 = np.ascontiguousarray(pos3, np.float32)
        out = np.zeros(pos3.shape[1], np.int32)
        lib().sdmm_ref_tree_find(self.h, pos3.shape[1], _f(pos3), out.ctypes.data_as(C.POINTER(C.c_int32)))
        return out


class RefRenderer:
    """The CPU restatement of the reference integrator's progressive loop (oracle/ref_render.inc) for a quad
    scene; mixture, tree and EM work is the reference's own code."""

    def __init__(self, quads, camera, lo, norm, root_min, root_max, K=16, max_depth=5, saved_per_path=8,
                 capacity=200000, split_threshold=4000, product_bsdf=None):
        # the render loop with sampleSurface's product branch only exists in libsdmm_ref_product.so
        L = self._L = product_lib() if product_bsdf is not None else lib()
        fp = C.POINTER(C.c_float)
        L.sdmm_ref_renderer_create.restype = C.c_void_p
        L.sdmm_ref_renderer_create.argtypes = [fp, C.c_int, fp, C.c_int, C.c_int, fp, C.c_float, fp, fp, C.c_int, C.c_int,
                                               C.c_int, C.c_long, C.c_uint64, C.c_long]
        L.sdmm_ref_renderer_destroy.argtypes = [C.c_void_p]
        L.sdmm_ref_renderer_iteration.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.c_int, C.c_int, C.POINTER(C.c_double)]
        L.sdmm_ref_renderer_image.argtypes = [C.c_void_p, fp]
        L.sdmm_ref_renderer_set_async.argtypes = [C.c_void_p, C.c_int]
        L.sdmm_ref_renderer_tree_size.argtypes = [C.c_void_p]
        L.sdmm_ref_renderer_tree_export.argtypes = [C.c_void_p, fp]
        L.sdmm_ref_renderer_leaf_joint.argtypes = [C.c_void_p, C.c_int, fp, C.POINTER(C.c_uint32)]
        q = np.zeros((len(quads), 15), np.float32)
        for i, x in enumerate(quads):
            q[i] = np.concatenate([x["p0"], x["e1"], x["e2"], x["reflectance"], x["emission"]])
        cam = np.concatenate([camera["origin"], camera["right"], camera["up"], camera["forward"],
                              [camera["fov_x_degrees"]]]).astype(np.float32)
        self.K, self.W, self.H = K, int(camera["width"]), int(camera["height"])
        self._h = C.c_void_p(L.sdmm_ref_renderer_create(
            _f(q), len(quads), _f(cam), self.W, self.H, _f(np.asarray(lo, np.float32)), float(norm),
            _f(np.asarray(root_min, np.float32)), _f(np.asarray(root_max, np.float32)), max_depth, saved_per_path, K,
            capacity, 0, split_threshold))
        self.total_spp = 0
        if product_bsdf is not None:
            b = np.ascontiguousarray(product_bsdf, np.float32).reshape(-1, 7)
            L.sdmm_ref_renderer_set_product.argtypes = [C.c_void_p, C.c_int, fp]
            assert L.sdmm_ref_renderer_set_product(self._h, b.shape[0], _f(b)) == 1

    def __del__(self):
        try:
            self._L.sdmm_ref_renderer_destroy(self._h)
        except Exception:
            pass

    def set_async(self, enabled=True):
        """optimizeAsync schedule (volpath_sdmm.cpp:180-242): fits become visible one pass later."""
        self._L.sdmm_ref_renderer_set_async(self._h, 1 if enabled else 0)

    def iteration(self, spp, pass_seed, still_training, force_unguided=False):
        st = (C.c_double * 7)()
        self._L.sdmm_ref_renderer_iteration(self._h, spp, pass_seed, 1 if still_training else 0,
                                          1 if force_unguided else 0, st)
        self.total_spp += spp
        keys = ["trained", "leaves", "fresh", "train_samples", "data_samples", "guided_queries", "mis_queries"]
        return {k: int(v) for k, v in zip(keys, st)}

    def image(self):
        out = np.zeros((3, self.H, self.W), np.float32)
        self._L.sdmm_ref_renderer_image(self._h, _f(out))
        return out / max(1, self.total_spp)

    def nodes(self):
        n = self._L.sdmm_ref_renderer_tree_size(self._h)
        out = np.zeros((n, 13), np.float32)
        self._L.sdmm_ref_renderer_tree_export(self._h, _f(out))
        return out

    def leaf_joint(self, node):
        joint = np.zeros((JOINT_PLANES, self.K), np.float32)
        scal = np.zeros(SCALAR_WORDS, np.uint32)
        k = self._L.sdmm_ref_renderer_leaf_joint(self._h, node, _f(joint), _u(scal))
        return (joint, scal) if k else (None, None)
