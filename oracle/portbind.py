"""ctypes binding of oracle/_build/liboracle.so -- the plain-C restatement (sdmm_oracle.c) of the
reference algorithm.  TEST INFRASTRUCTURE ONLY (tests/, __graft_entry__.smoke(), bench.py's
cpu_baseline leg); the product package never imports this.  Same interface as refbind.RefCells.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")

JOINT_PLANES = 32
EM_PLANES = 81
SCALAR_WORDS = 8

_lib = None


def build():
    subprocess.check_call(["make", "-C", _HERE, "_build/liboracle.so"], stdout=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        fp = C.POINTER(C.c_float)
        up = C.POINTER(C.c_uint32)
        bp = C.POINTER(C.c_uint8)
        vpp = C.POINTER(C.c_void_p)
        L.oracle_cell_create.restype = C.c_void_p
        L.oracle_cell_create.argtypes = [C.c_int]
        L.oracle_cell_destroy.argtypes = [C.c_void_p]
        L.oracle_cell_set.argtypes = [C.c_void_p, fp, fp, up]
        L.oracle_cell_get.argtypes = [C.c_void_p, fp, fp, up]
        L.oracle_cell_get_derived.argtypes = [C.c_void_p, fp]
        L.oracle_em_step.argtypes = [C.c_void_p, C.c_int, fp, C.c_long, fp, C.c_int, fp, C.c_int]
        L.oracle_em_batch.restype = C.c_double
        L.oracle_em_batch.argtypes = [vpp, C.c_int, up, fp, C.c_long, fp, C.c_int, C.c_int, C.c_int]
        L.oracle_prepare_conditioner.argtypes = [C.c_void_p]
        L.oracle_get_conditioner.argtypes = [C.c_void_p, fp]
        L.oracle_sample_pdf.restype = C.c_double
        L.oracle_sample_pdf.argtypes = [vpp, C.c_long, up, fp, fp, fp, fp, fp, up, bp, C.c_int]
        L.oracle_pdf.restype = C.c_double
        L.oracle_pdf.argtypes = [vpp, C.c_long, up, fp, fp, fp, bp, fp, fp, C.c_int]
        L.oracle_cholesky.argtypes = [C.c_int, fp, fp]
        L.oracle_solve_lower.argtypes = [C.c_int, fp, fp, fp]
        L.oracle_inverse_lower_tri.argtypes = [C.c_int, fp, fp]
        L.oracle_gauss_pdf_euclid.restype = C.c_float
        L.oracle_gauss_pdf_euclid.argtypes = [C.c_int, fp, fp, fp]
        L.oracle_schur.argtypes = [C.c_int, C.c_int, fp, fp, fp, fp, fp]
        L.oracle_categorical_prepare.argtypes = [C.c_int, fp, fp]
        L.oracle_frame.argtypes = [fp, fp, fp]
        L.oracle_joint_to.argtypes = [fp, fp, fp, fp]
        L.oracle_initialize.argtypes = [C.c_void_p, C.c_int, fp, fp, C.c_long, fp, C.c_int, C.c_float,
                                        C.c_uint64, C.c_uint64, C.POINTER(C.c_int32)]
        L.oracle_pcg32_floats.argtypes = [C.c_uint64, C.c_uint64, C.c_int, fp]
        _lib = L
    return _lib


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
    return C.cast(a.ctypes.data + 4 * n, C.POINTER(C.c_float))

PCG32_DEFAULT_STATE = 0x853c49e6748fea9b
PCG32_DEFAULT_STREAM = 0xda3e39cb94b95bdb


class PortCells:
    def __init__(self, n_cells, K):
        self.C, self.K = n_cells, K
        L = lib()
        self.handles = (C.c_void_p * n_cells)(*[L.oracle_cell_create(K) for _ in range(n_cells)])

    def __del__(self):
        try:
            L = lib()
            for h in self.handles:
                L.oracle_cell_destroy(h)
        except Exception:
            pass

    def _hp(self):
        return C.cast(self.handles, C.POINTER(C.c_void_p))

    def initialize(self, offsets, point6, normal3, weight, spatial_distance,
                   rng_state=PCG32_DEFAULT_STATE, rng_stream=PCG32_DEFAULT_STREAM):
        """sdmm::initialize per cell with K/8 seeds; returns the seed sample indices [C, K/8]."""
        L = lib()
        stride = point6.shape[1]
        seeds = np.zeros((self.C, self.K // 8), np.int32)
        for c in range(self.C):
            b, e = int(offsets[c]), int(offsets[c + 1])
            k = L.oracle_initialize(self.handles[c], e - b, _off(point6, b), _off(normal3, b), stride,
                                    _off(weight, b), self.K // 8, float(spatial_distance[c]), rng_state,
                                    rng_stream, seeds[c].ctypes.data_as(C.POINTER(C.c_int32)))
            assert k == self.K, (k, self.K)
        return seeds

    def set(self, joint, em=None, scal=None):
        L = lib()
        ok = np.zeros(self.C, dtype=bool)
        for c in range(self.C):
            ok[c] = bool(L.oracle_cell_set(self.handles[c], _f(joint[c]),
                                           _f(em[c]) if em is not None else None,
                                           _u(scal[c]) if scal is not None else None))
        return ok

    def get(self):
        L = lib()
        joint = np.zeros((self.C, JOINT_PLANES, self.K), np.float32)
        em = np.zeros((self.C, EM_PLANES, self.K), np.float32)
        scal = np.zeros((self.C, SCALAR_WORDS), np.uint32)
        for c in range(self.C):
            L.oracle_cell_get(self.handles[c], _f(joint[c]), _f(em[c]), _u(scal[c]))
        return joint, em, scal

    def derived(self):
        L = lib()
        out = np.zeros((self.C, 38, self.K), np.float32)
        for c in range(self.C):
            L.oracle_cell_get_derived(self.handles[c], _f(out[c]))
        return out

    def em_step(self, offsets, point6, weight, n_iters, want_loglik=False, sum_mode=0):
        L = lib()
        ntot = point6.shape[1]
        ll = [None] * self.C if want_loglik else None
        for c in range(self.C):
            b, e = int(offsets[c]), int(offsets[c + 1])
            out = np.zeros((n_iters, e - b), np.float32) if want_loglik else None
            L.oracle_em_step(self.handles[c], e - b, _off(point6, b), ntot, _off(weight, b), n_iters,
                             _f(out) if want_loglik else None, sum_mode)
            if want_loglik:
                ll[c] = out
        return ll

    def em_batch(self, offsets, point6, weight, n_iters, with_prepare=True, n_threads=0):
        L = lib()
        return L.oracle_em_batch(self._hp(), self.C, _u(offsets), _f(point6), point6.shape[1],
                                 _f(weight), n_iters, 1 if with_prepare else 0, n_threads)

    def prepare_conditioner(self):
        L = lib()
        for h in self.handles:
            L.oracle_prepare_conditioner(h)

    def conditioner(self):
        L = lib()
        out = np.zeros((self.C, 38, self.K), np.float32)
        for c in range(self.C):
            L.oracle_get_conditioner(self.handles[c], _f(out[c]))
        return out

    def sample_pdf(self, cell_idx, pos3, u3, n_threads=0):
        L = lib()
        n = pos3.shape[1]
        out_dir = np.zeros((3, n), np.float32)
        out_pdf = np.zeros(n, np.float32)
        out_ij = np.zeros(n, np.float32)
        out_comp = np.zeros(n, np.uint32)
        out_valid = np.zeros(n, np.uint8)
        secs = L.oracle_sample_pdf(self._hp(), n, _u(cell_idx), _f(pos3), _f(u3), _f(out_dir),
                                   _f(out_pdf), _f(out_ij), _u(out_comp), _b(out_valid), n_threads)
        return dict(dir=out_dir, pdf=out_pdf, inv_jac=out_ij, comp=out_comp, valid=out_valid), secs

    def pdf(self, cell_idx, pos3, dir3, n_threads=0):
        L = lib()
        n = pos3.shape[1]
        out_pdf = np.zeros(n, np.float32)
        out_valid = np.zeros(n, np.uint8)
        secs = L.oracle_pdf(self._hp(), n, _u(cell_idx), _f(pos3), _f(dir3), _f(out_pdf),
                            _b(out_valid), None, None, n_threads)
        return out_pdf, out_valid, secs

    def conditional_weights(self, cell, x):
        L = lib()
        pos = np.asarray(x, np.float32).reshape(3, 1).copy()
        d = np.zeros((3, 1), np.float32)
        d[2] = 1
        pmf = np.zeros(self.K, np.float32)
        cdf = np.zeros(self.K, np.float32)
        pdf = np.zeros(1, np.float32)
        valid = np.zeros(1, np.uint8)
        idx = np.array([cell], np.uint32)
        L.oracle_pdf(self._hp(), 1, _u(idx), _f(pos), _f(d), _f(pdf), _b(valid), _f(pmf), _f(cdf), 1)
        return pmf, cdf, bool(valid[0])


def pcg32_floats(n, state=PCG32_DEFAULT_STATE, stream=PCG32_DEFAULT_STREAM):
    out = np.zeros(n, np.float32)
    lib().oracle_pcg32_floats(state, stream, n, _f(out))
    return out


def cholesky(m):
    m = np.ascontiguousarray(m, np.float32)
    D = m.shape[0]
    out = np.zeros((D, D), np.float32)
    psd = lib().oracle_cholesky(D, _f(m), _f(out))
    return out, bool(psd)


def solve_lower(L_, b):
    L_ = np.ascontiguousarray(L_, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    x = np.zeros_like(b)
    lib().oracle_solve_lower(L_.shape[0], _f(L_), _f(b), _f(x))
    return x


def inverse_lower_tri(L_):
    L_ = np.ascontiguousarray(L_, np.float32)
    out = np.zeros_like(L_)
    lib().oracle_inverse_lower_tri(L_.shape[0], _f(L_), _f(out))
    return out


def gauss_pdf_euclid(mean, cov, x):
    mean = np.ascontiguousarray(mean, np.float32)
    cov = np.ascontiguousarray(cov, np.float32)
    x = np.ascontiguousarray(x, np.float32)
    return float(lib().oracle_gauss_pdf_euclid(len(mean), _f(mean), _f(cov), _f(x)))


def schur(cov, Da):
    cov = np.ascontiguousarray(cov, np.float32)
    D = cov.shape[0]
    Db = D - Da
    La = np.zeros((Da, Da), np.float32)
    M = np.zeros((Db, Da), np.float32)
    cc = np.zeros((Db, Db), np.float32)
    Lc = np.zeros((Db, Db), np.float32)
    lib().oracle_schur(D, Da, _f(cov), _f(La), _f(M), _f(cc), _f(Lc))
    return La, M, cc, Lc


def categorical_prepare(pmf):
    p = np.ascontiguousarray(pmf, np.float32).copy()
    c = np.zeros_like(p)
    ok = lib().oracle_categorical_prepare(len(p), _f(p), _f(c))
    return p, c, bool(ok)


def frame(n):
    n = np.ascontiguousarray(n, np.float32)
    to = np.zeros((3, 3), np.float32)
    fr = np.zeros((3, 3), np.float32)
    lib().oracle_frame(_f(n), _f(to), _f(fr))
    return to, fr


def joint_to(mean6, point6):
    mean6 = np.ascontiguousarray(mean6, np.float32)
    point6 = np.ascontiguousarray(point6, np.float32)
    t = np.zeros(5, np.float32)
    ij = np.zeros(1, np.float32)
    lib().oracle_joint_to(_f(mean6), _f(point6), _f(t), _f(ij))
    return t, float(ij[0])
