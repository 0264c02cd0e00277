"""Host-side mirror of the reference's sdmm EM / mixture interface for the hot path, batched over
spatial cells and backed ONLY by the CUDA library (capi.py -> libsdmm_b200.so).

Reference interface mirrored (paths relative to /root/reference):
  SDMMContext bundle per cell            sdmm/include/sdmm/distributions/sdmm_context.h:11-45
  Data::push_back / clear                sdmm/include/sdmm/opt/data.h:59-102           -> set_samples
  sdmm::em_step                          sdmm/include/sdmm/opt/em.h:218-253            -> em_step
  sdmm::prepare(conditioner, joint)      sdmm/include/sdmm/distributions/sdmm_conditioner.h:89-100 -> prepare
  sdmm::create_conditional + SDMM::sample + SDMM::posterior (per vertex, sdmm_proc.cpp:361-372,
  419-421, 539-545)                                                                    -> sample_pdf / pdf

NumPy arrays take the host-buffer entry points (copies inside the call); torch CUDA tensors take the
device-pointer entry points on torch's current stream.  PyTorch is only used for device memory,
streams and torch.distributed plumbing.
"""
import ctypes as C

import numpy as np

from . import capi
from .capi import EM_PLANES, JOINT_PLANES, QUERY_RECORD_FLOATS, SCALAR_WORDS, check


def _is_torch(x):
    return type(x).__module__.startswith("torch")


def _np(a, dtype):
    a = np.ascontiguousarray(a, dtype=dtype)
    return a


def _hp(a):
    return C.c_void_p(a.ctypes.data) if a is not None else None


def _stream(device):
    import torch
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class SDMMCells:
    """A batch of per-cell SDMM contexts (joint mixture + EM state + conditioner) resident on one GPU."""

    def __init__(self, max_cells, K, device=0):
        self._L = capi.lib()
        self.max_cells, self.K, self.device = int(max_cells), int(K), int(device)
        h = C.c_void_p()
        check(self._L.sdmmb_create(C.byref(h), self.device, self.max_cells, self.K), "sdmmb_create")
        self._h = h
        self.n_samples = 0
        self._keep = None

    def close(self):
        if getattr(self, "_h", None):
            self._L.sdmmb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- model state -----------------------------------------------------------------------
    def upload(self, joint, em=None, scal=None, first_cell=0):
        """joint [C,32,K] f32, em [C,81,K] f32 or None, scal [C,8] u32 words or None.
        Runs set_mean + prepare_vectorized (sdmm.h:157-173) on the device."""
        joint = _np(joint, np.float32)
        n = joint.shape[0]
        assert joint.shape == (n, JOINT_PLANES, self.K), joint.shape
        em = _np(em, np.float32) if em is not None else None
        scal = _np(scal, np.uint32) if scal is not None else None
        if em is not None:
            assert em.shape == (n, EM_PLANES, self.K), em.shape
        if scal is not None:
            assert scal.shape == (n, SCALAR_WORDS), scal.shape
        check(self._L.sdmmb_upload_cells(self._h, first_cell, n, _hp(joint), _hp(em), _hp(scal), None),
              "sdmmb_upload_cells", self._h)

    def download(self, first_cell=0, n_cells=None, out=None, want_em=True):
        """Returns (joint, em, scal).  `out` = optional (joint, em-or-None, scal) NumPy buffers to fill
        (e.g. views of pinned memory); want_em=False skips the EM statistics (em is None)."""
        n = self.max_cells - first_cell if n_cells is None else n_cells
        if out is not None:
            joint, em, scal = out
        else:
            joint = np.zeros((n, JOINT_PLANES, self.K), np.float32)
            em = np.zeros((n, EM_PLANES, self.K), np.float32) if want_em else None
            scal = np.zeros((n, SCALAR_WORDS), np.uint32)
        check(self._L.sdmmb_download_cells(self._h, first_cell, n, _hp(joint), _hp(em), _hp(scal), None),
              "sdmmb_download_cells", self._h)
        return joint, em, scal

    def status(self, first_cell=0, n_cells=None):
        """Per-cell status words of the last training call (ST_* bits; 0 = ok): only the scalars travel."""
        n = self.max_cells - first_cell if n_cells is None else n_cells
        scal = np.zeros((n, SCALAR_WORDS), np.uint32)
        check(self._L.sdmmb_download_cells(self._h, first_cell, n, None, None, _hp(scal), None),
              "sdmmb_download_cells", self._h)
        return scal[:, 7].copy()

    def derived(self, first_cell=0, n_cells=None):
        n = self.max_cells - first_cell if n_cells is None else n_cells
        out = np.zeros((n, 38, self.K), np.float32)
        check(self._L.sdmmb_download_derived(self._h, first_cell, n, _hp(out), None), "sdmmb_download_derived", self._h)
        return out

    def conditioner(self, first_cell=0, n_cells=None):
        n = self.max_cells - first_cell if n_cells is None else n_cells
        out = np.zeros((n, 38, self.K), np.float32)
        check(self._L.sdmmb_download_conditioner(self._h, first_cell, n, _hp(out), None),
              "sdmmb_download_conditioner", self._h)
        return out

    # ---- training ---------------------------------------------------------------------------
    def set_samples(self, point6, weight, cell_offsets, first_cell=0, borrowed=False):
        """point6 [6,N] (x,y,z,dx,dy,dz planes), weight [N], cell_offsets [C+1] (host, u32).
        borrowed=True (device tensors only): zero-copy, the E-step reads the planes in place; they must not
        change until the next binding (the object keeps them alive)."""
        off = _np(np.asarray(cell_offsets), np.uint32)
        n_cells = off.shape[0] - 1
        if _is_torch(point6):
            assert point6.is_cuda and weight.is_cuda and point6.dtype.is_floating_point
            assert point6.is_contiguous() and weight.is_contiguous() and point6.shape[0] == 6
            self._keep = (point6, weight)
            fn = self._L.sdmmb_set_samples_borrowed if borrowed else self._L.sdmmb_set_samples
            check(fn(self._h, first_cell, n_cells, C.c_void_p(point6.data_ptr()),
                     point6.shape[1], C.c_void_p(weight.data_ptr()), _hp(off),
                     _stream(self.device)), "sdmmb_set_samples", self._h)
        else:
            if borrowed:
                raise ValueError("borrowed=True needs device tensors")
            point6 = _np(point6, np.float32)
            weight = _np(weight, np.float32)
            assert point6.shape[0] == 6 and weight.shape[0] == point6.shape[1]
            check(self._L.sdmmb_set_samples_host(self._h, first_cell, n_cells, _hp(point6), point6.shape[1],
                                                 _hp(weight), _hp(off), None), "sdmmb_set_samples_host", self._h)
        self.n_samples = int(off[-1] - off[0])
        self._bound = (first_cell, n_cells)

    def initialize(self, point6, normal3, weight, cell_offsets, spatial_distance, first_cell=0,
                   rng_state=capi.PCG32_DEFAULT_STATE, rng_stream=capi.PCG32_DEFAULT_STREAM):
        """sdmm::initialize (sdmm/include/sdmm/opt/init.h:17-193) for every cell of the range from its own
        samples (point6 [6,N], normal3 [3,N], weight [N]); spatial_distance: per-cell array.  Returns the seed
        sample indices [C, K/8] (relative to each cell's first sample)."""
        off = _np(np.asarray(cell_offsets), np.uint32)
        n_cells = off.shape[0] - 1
        sd = _np(np.broadcast_to(np.asarray(spatial_distance, np.float32), (n_cells,)), np.float32)
        seeds = np.zeros((n_cells, self.K // 8), np.int32)
        if _is_torch(point6):
            assert point6.is_cuda and normal3.is_cuda and weight.is_cuda
            assert point6.is_contiguous() and normal3.is_contiguous() and weight.is_contiguous()
            assert point6.shape[0] == 6 and normal3.shape[0] == 3 and normal3.shape[1] == point6.shape[1]
            check(self._L.sdmmb_initialize(self._h, first_cell, n_cells, C.c_void_p(point6.data_ptr()),
                                           C.c_void_p(normal3.data_ptr()), point6.shape[1],
                                           C.c_void_p(weight.data_ptr()), _hp(off), _hp(sd), rng_state, rng_stream,
                                           _hp(seeds), _stream(self.device)), "sdmmb_initialize", self._h)
        else:
            point6, normal3, weight = _np(point6, np.float32), _np(normal3, np.float32), _np(weight, np.float32)
            assert point6.shape[0] == 6 and normal3.shape == (3, point6.shape[1])
            check(self._L.sdmmb_initialize_host(self._h, first_cell, n_cells, _hp(point6), _hp(normal3),
                                                point6.shape[1], _hp(weight), _hp(off), _hp(sd), rng_state,
                                                rng_stream, _hp(seeds), None), "sdmmb_initialize_host", self._h)
        return seeds

    def em_step(self, n_iters=1, first_cell=None, n_cells=None, with_conditioner=True, want_loglik=False):
        """n_iters x sdmm::em_step on every bound cell (+ sdmm::prepare of the conditioner after the
        last one).  want_loglik -> torch tensor [n_iters, N] of per-sample log-likelihoods.
        n_iters = capi.ITERS_INTEGRATOR (-1): the integrator's rule per cell on the device -- two iterations for
        cells whose em.iterations_run < 4, one for the others (volpath_sdmm.cpp:299-302)."""
        if first_cell is None:
            first_cell, n_cells = self._bound
        ll = None
        ptr = None
        st = None
        if want_loglik:
            import torch
            ll = torch.full((2 if n_iters < 0 else n_iters, self.n_samples), float("nan"), dtype=torch.float32,
                            device="cuda:%d" % self.device)
            ptr = C.c_void_p(ll.data_ptr())
            st = _stream(self.device)
        else:
            try:
                st = _stream(self.device)
            except Exception:
                st = None
        check(self._L.sdmmb_em_step(self._h, first_cell, n_cells, n_iters, 1 if with_conditioner else 0, ptr, st),
              "sdmmb_em_step", self._h)
        return ll

    def set_em_overlap(self, enabled=True):
        """Hide em_step's M-steps behind E-steps (two halves on two streams + a high-priority M-step stream)."""
        check(self._L.sdmmb_set_em_overlap(self._h, 1 if enabled else 0), "sdmmb_set_em_overlap", self._h)

    def optimize_host(self, point6, weight, cell_offsets, n_iters, first_cell=0, with_conditioner=True, out=None):
        """The integrator's per-iteration training call for a batch of cells, host buffers in and out
        (volpath_sdmm.cpp:287-311): bind samples -> n_iters x em_step -> conditioner -> fitted mixtures.
        `out` = (joint, em-or-None, scal) NumPy buffers (pinned memory overlaps the copies)."""
        off = _np(np.asarray(cell_offsets), np.uint32)
        n_cells = off.shape[0] - 1
        point6 = _np(point6, np.float32)
        weight = _np(weight, np.float32)
        assert point6.shape[0] == 6 and weight.shape[0] == point6.shape[1]
        if out is None:
            out = (np.zeros((n_cells, JOINT_PLANES, self.K), np.float32),
                   np.zeros((n_cells, EM_PLANES, self.K), np.float32),
                   np.zeros((n_cells, SCALAR_WORDS), np.uint32))
        joint, em, scal = out
        check(self._L.sdmmb_optimize_host(self._h, first_cell, n_cells, _hp(point6), point6.shape[1], _hp(weight),
                                          _hp(off), n_iters, 1 if with_conditioner else 0, _hp(joint), _hp(em),
                                          _hp(scal), None), "sdmmb_optimize_host", self._h)
        self.n_samples = int(off[-1] - off[0])
        self._bound = (first_cell, n_cells)
        return joint, em, scal

    def prepare(self, first_cell=0, n_cells=None):
        """sdmm::prepare(conditioner, sdmm) for a cell range."""
        n = self.max_cells - first_cell if n_cells is None else n_cells
        try:
            st = _stream(self.device)
        except Exception:
            st = None
        check(self._L.sdmmb_prepare_conditioner(self._h, first_cell, n, st), "sdmmb_prepare_conditioner", self._h)

    # ---- spatial cell structure ---------------------------------------------------------------
    def set_tree(self, nodes):
        """nodes: structured array of tree.TREE_NODE_DTYPE (see tree.flatten)."""
        from .tree import TREE_NODE_DTYPE
        nodes = np.ascontiguousarray(nodes, dtype=TREE_NODE_DTYPE)
        check(self._L.sdmmb_set_tree(self._h, nodes.shape[0], C.c_void_p(nodes.ctypes.data)), "sdmmb_set_tree", self._h)

    def copy_cells(self, src, dst):
        """Child cells inherit the parent's complete state on a split (spatial_tree.h:187-196)."""
        src = _np(np.asarray(src), np.int32)
        dst = _np(np.asarray(dst), np.int32)
        assert src.shape == dst.shape
        check(self._L.sdmmb_copy_cells(self._h, src.shape[0], _hp(src), _hp(dst), None), "sdmmb_copy_cells", self._h)

    def find_cells(self, pos3):
        """STree::find for a batch of points: pos3 [3,n] -> cell slot per point (0xffffffff = none)."""
        if _is_torch(pos3):
            import torch
            out = torch.empty(pos3.shape[1], dtype=torch.int32, device=pos3.device)
            check(self._L.sdmmb_find_cells(self._h, pos3.shape[1], C.c_void_p(pos3.data_ptr()),
                                           C.c_void_p(out.data_ptr()), _stream(self.device)), "sdmmb_find_cells", self._h)
            return out
        pos3 = _np(pos3, np.float32)
        out = np.zeros(pos3.shape[1], np.uint32)
        check(self._L.sdmmb_find_cells_host(self._h, pos3.shape[1], _hp(pos3), _hp(out), None),
              "sdmmb_find_cells_host", self._h)
        return out

    def collect_samples(self, point6, weight, key3=None):
        """The renderer's sample collection for a batch of unsorted samples (torch CUDA tensors):
        find each sample's cell and bin it; binds the result as the training samples of all cells.
        Returns (counts [max_cells] u32, dropped)."""
        assert _is_torch(point6) and point6.is_cuda and point6.is_contiguous() and point6.shape[0] == 6
        n = point6.shape[1]
        counts = np.zeros(self.max_cells, np.uint32)
        dropped = C.c_uint64()
        kp, ks = None, 0
        if key3 is not None:
            assert key3.is_cuda and key3.is_contiguous() and key3.shape == (3, n)
            kp, ks = C.c_void_p(key3.data_ptr()), n
        self._keep = (point6, weight, key3)
        check(self._L.sdmmb_collect_samples(self._h, n, C.c_void_p(point6.data_ptr()), n, C.c_void_p(weight.data_ptr()),
                                            kp, ks, _hp(counts), C.byref(dropped), _stream(self.device)),
              "sdmmb_collect_samples", self._h)
        self.n_samples = int(counts.sum())
        self._bound = (0, self.max_cells)
        return counts, int(dropped.value)

    def set_min_samples(self, n):
        check(self._L.sdmmb_set_min_samples(self._h, int(n)), "sdmmb_set_min_samples", self._h)

    def set_async_records(self, enabled=True):
        """optimizeAsync: fits write a back copy of the query records until publish_records()."""
        check(self._L.sdmmb_set_async_records(self._h, 1 if enabled else 0), "sdmmb_set_async_records", self._h)

    def publish_records(self, first_cell=0, n_cells=None):
        """optimize_async_wait_and_update (volpath_sdmm.cpp:227-242) for a cell range."""
        n = self.max_cells - first_cell if n_cells is None else n_cells
        check(self._L.sdmmb_publish_records(self._h, first_cell, n, None), "sdmmb_publish_records", self._h)

    # ---- queries ------------------------------------------------------------------------------
    def sample_pdf(self, cell_idx, pos3, u3, binned=False, out=None):
        """Per vertex: create_conditional(x) -> sample(u) -> posterior(omega) summed.
        u3 planes by role: select, Box-Muller radius, Box-Muller angle."""
        if _is_torch(pos3):
            import torch
            n = pos3.shape[1]
            dev = pos3.device
            out = dict(dir=torch.empty((3, n), dtype=torch.float32, device=dev),
                       pdf=torch.empty(n, dtype=torch.float32, device=dev),
                       inv_jac=torch.empty(n, dtype=torch.float32, device=dev),
                       comp=torch.empty(n, dtype=torch.int32, device=dev),
                       valid=torch.empty(n, dtype=torch.uint8, device=dev))
            check(self._L.sdmmb_sample_pdf(self._h, n, C.c_void_p(cell_idx.data_ptr()), C.c_void_p(pos3.data_ptr()),
                                           C.c_void_p(u3.data_ptr()), C.c_void_p(out["dir"].data_ptr()),
                                           C.c_void_p(out["pdf"].data_ptr()), C.c_void_p(out["inv_jac"].data_ptr()),
                                           C.c_void_p(out["comp"].data_ptr()), C.c_void_p(out["valid"].data_ptr()),
                                           1 if binned else 0, _stream(self.device)), "sdmmb_sample_pdf", self._h)
            return out
        cell_idx = _np(cell_idx, np.uint32)
        pos3 = _np(pos3, np.float32)
        u3 = _np(u3, np.float32)
        n = pos3.shape[1]
        if out is None:   # `out`: caller-provided (e.g. pinned) host arrays of the same shapes
            out = dict(dir=np.zeros((3, n), np.float32), pdf=np.zeros(n, np.float32), inv_jac=np.zeros(n, np.float32),
                       comp=np.zeros(n, np.uint32), valid=np.zeros(n, np.uint8))
        check(self._L.sdmmb_sample_pdf_host(self._h, n, _hp(cell_idx), _hp(pos3), _hp(u3), _hp(out["dir"]),
                                            _hp(out["pdf"]), _hp(out["inv_jac"]), _hp(out["comp"]),
                                            _hp(out["valid"]), 1 if binned else 0, None),
              "sdmmb_sample_pdf_host", self._h)
        return out

    def pdf(self, cell_idx, pos3, dir3, binned=False):
        """pdfSurface's mixture term for caller-supplied directions."""
        cell_idx = _np(cell_idx, np.uint32)
        pos3 = _np(pos3, np.float32)
        dir3 = _np(dir3, np.float32)
        n = pos3.shape[1]
        out_pdf = np.zeros(n, np.float32)
        out_valid = np.zeros(n, np.uint8)
        check(self._L.sdmmb_pdf_host(self._h, n, _hp(cell_idx), _hp(pos3), _hp(dir3), _hp(out_pdf), _hp(out_valid),
                                     1 if binned else 0, None), "sdmmb_pdf_host", self._h)
        return out_pdf, out_valid

    def sample_pdf_pruned(self, cell_idx, pos3, u3, max_components, preserve_idx=-1, dir3=None):
        """create_conditional_pruned (sdmm_conditioner.h:128-168) -> sample -> posterior per vertex: only the
        max_components strongest conditional weights survive.  u3=None + dir3: pdf only.  Host arrays or device tensors."""
        import torch
        dev = torch.device("cuda", self.device)

        def dv(a, dt):
            if a is None:
                return None
            return a.contiguous() if _is_torch(a) else torch.from_numpy(np.ascontiguousarray(a, dt)).to(dev)

        cell_d, pos_d = dv(cell_idx if _is_torch(cell_idx) else np.asarray(cell_idx).view(np.int32), np.int32), dv(pos3, np.float32)
        u_d, d_d = dv(u3, np.float32), dv(dir3, np.float32)
        n = pos_d.shape[1]
        out = dict(dir=torch.zeros((3, n), dtype=torch.float32, device=dev), pdf=torch.zeros(n, dtype=torch.float32, device=dev),
                   inv_jac=torch.zeros(n, dtype=torch.float32, device=dev), comp=torch.zeros(n, dtype=torch.int32, device=dev),
                   valid=torch.zeros(n, dtype=torch.uint8, device=dev))
        p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
        check(self._L.sdmmb_sample_pdf_pruned(self._h, n, p(cell_d), p(pos_d), p(u_d), p(d_d), int(max_components),
                                              int(preserve_idx), p(out["dir"]), p(out["pdf"]), p(out["inv_jac"]),
                                              p(out["comp"]), p(out["valid"]), _stream(self.device)),
              "sdmmb_sample_pdf_pruned", self._h)
        return out

    def sample_pdf_product(self, cell_idx, pos3, u3, bsdf, normal3=None, dir3=None, frame9=None, wi3=None):
        """Product guiding per vertex (sdmm_proc.cpp:327-339, 375-381; sdmm_product.h:62-203): conditional x
        learned-BSDF mixture.  `bsdf`: [M][7] rows (pmf, mean3, cov00, cov10, cov11) in world space; `normal3`
        re-centres component 0 on the shading normal (diffuse surfaces); `frame9` [9,n] (shading frame to-world
        matrix, row-major) + `wi3` [3,n] (local incident direction) select the glossy branch.  With `dir3` instead of `u3` only the
        pdf at the given directions is evaluated.  Inputs: host arrays or torch device tensors."""
        import torch
        dev = torch.device("cuda", self.device)

        def dv(a, dt):
            if a is None:
                return None
            return a.contiguous() if _is_torch(a) else torch.from_numpy(np.ascontiguousarray(a, dt)).to(dev)

        as_torch = _is_torch(pos3)
        if _is_torch(cell_idx):
            cell_t = cell_idx.contiguous()
        else:
            cell_t = torch.from_numpy(np.ascontiguousarray(cell_idx, np.uint32).view(np.int32)).to(dev)
        pos_t, u_t, n_t, d_t = dv(pos3, np.float32), dv(u3, np.float32), dv(normal3, np.float32), dv(dir3, np.float32)
        f_t, w_t = dv(frame9, np.float32), dv(wi3, np.float32)
        bsdf = np.ascontiguousarray(bsdf, np.float32).reshape(-1, 7)
        n = pos_t.shape[1]
        ptr = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
        out = dict(pdf=torch.zeros(n, dtype=torch.float32, device=dev),
                   valid=torch.zeros(n, dtype=torch.uint8, device=dev))
        if u_t is not None:
            out.update(dir=torch.zeros((3, n), dtype=torch.float32, device=dev),
                       inv_jac=torch.zeros(n, dtype=torch.float32, device=dev),
                       comp=torch.zeros(n, dtype=torch.int32, device=dev))
            check(self._L.sdmmb_sample_pdf_product(self._h, n, ptr(cell_t), ptr(pos_t), ptr(u_t), ptr(n_t),
                                                   ptr(f_t), ptr(w_t), bsdf.shape[0], _hp(bsdf), ptr(out["dir"]), ptr(out["pdf"]),
                                                   ptr(out["inv_jac"]), ptr(out["comp"]), ptr(out["valid"]),
                                                   _stream(self.device)), "sdmmb_sample_pdf_product", self._h)
        else:
            check(self._L.sdmmb_pdf_product(self._h, n, ptr(cell_t), ptr(pos_t), ptr(d_t), ptr(n_t), ptr(f_t), ptr(w_t),
                                            bsdf.shape[0],
                                            _hp(bsdf), ptr(out["pdf"]), ptr(out["valid"]), _stream(self.device)),
                  "sdmmb_pdf_product", self._h)
        if not as_torch:
            out = {k: v.cpu().numpy() for k, v in out.items()}
            if "comp" in out:
                out["comp"] = out["comp"].view(np.uint32)
        return out

    # ---- multi-GPU / introspection --------------------------------------------------------------
    def query_records(self):
        """Zero-copy torch view [max_cells, K_padded, 32] of the device-resident query records."""
        import torch
        p, n = C.c_void_p(), C.c_size_t()
        check(self._L.sdmmb_query_records(self._h, C.byref(p), C.byref(n)), "sdmmb_query_records")
        kp = n.value // QUERY_RECORD_FLOATS

        class _Ext:
            pass

        ext = _Ext()
        ext.__cuda_array_interface__ = dict(shape=(self.max_cells, kp, QUERY_RECORD_FLOATS), typestr="<f4",
                                            data=(p.value, False), version=2)
        self._ext = ext
        return torch.as_tensor(ext, device="cuda:%d" % self.device)

    def allgather_records(self, cells_per_rank, group=None):
        """After a training iteration: every rank contributes the records of its contiguous slice
        (rank r owns cells [r*cells_per_rank, (r+1)*cells_per_rank)).  NCCL all-gather, in place."""
        from . import sharding
        sharding.allgather_records(self.query_records(), cells_per_rank, group)

    # ---- record exchange over peer memory, fused into the kernel that writes the records ----------
    def peer_export(self):
        """This context's opaque peer handle (bytes) for sdmmb_peer_attach on the other ranks."""
        buf = (C.c_ubyte * capi.PEER_HANDLE_BYTES)()
        check(self._L.sdmmb_peer_export(self._h, C.cast(buf, C.c_void_p)), "sdmmb_peer_export", self._h)
        return bytes(buf)

    def peer_attach(self, rank, handles):
        """handles: list of every rank's peer_export() bytes, rank r's at index r."""
        blob = b"".join(handles)
        buf = (C.c_ubyte * len(blob)).from_buffer_copy(blob)
        check(self._L.sdmmb_peer_attach(self._h, int(rank), len(handles), C.cast(buf, C.c_void_p)),
              "sdmmb_peer_attach", self._h)

    def attach_peers(self, group=None):
        """One process per GPU: exchange the handles over torch.distributed (any backend) and attach.
        Afterwards em_step(with_conditioner=True) / prepare() store the records they write straight into
        every peer's record array; finish a step with wait_peers() instead of allgather_records()."""
        import torch.distributed as dist
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        mine = self.peer_export()
        handles = [None] * world
        dist.all_gather_object(handles, mine, group=group)
        self.peer_attach(rank, handles)

    def detach_peers(self):
        check(self._L.sdmmb_peer_detach(self._h), "sdmmb_peer_detach", self._h)

    def wait_peers(self):
        """Stream-ordered device-side wait until every peer's records of the current step arrived."""
        check(self._L.sdmmb_peer_wait(self._h, _stream(self.device)), "sdmmb_peer_wait", self._h)

    def peer_status(self):
        """(exchange epochs issued, bounded waits that timed out)."""
        ep, to = C.c_uint32(), C.c_uint32()
        check(self._L.sdmmb_peer_status(self._h, C.byref(ep), C.byref(to), _stream(self.device)),
              "sdmmb_peer_status", self._h)
        return int(ep.value), int(to.value)

    def launch_count(self):
        v = C.c_uint64()
        check(self._L.sdmmb_launch_count(self._h, C.byref(v)), "sdmmb_launch_count")
        return int(v.value)

    def last_estep_ms(self):
        ms, n = C.c_float(), C.c_int()
        check(self._L.sdmmb_last_estep_ms(self._h, C.byref(ms), C.byref(n)), "sdmmb_last_estep_ms", self._h)
        return float(ms.value), int(n.value)


def measure_peaks(device=0):
    """FP32 scalar / packed TFLOP/s, MUFU Gop/s and SM MHz measured by on-device micro-benchmarks."""
    out = (C.c_float * 4)()
    check(capi.lib().sdmmb_measure_peaks(device, out), "sdmmb_measure_peaks")
    return dict(fp32_ffma_tflops=out[0], fp32_ffma2_tflops=out[1], mufu_gops=out[2], sm_mhz=out[3])


# reference-named free functions (sdmm::em_step / sdmm::prepare)
def em_step(cells, n_iters=1, **kw):
    return cells.em_step(n_iters=n_iters, **kw)


def prepare(cells, first_cell=0, n_cells=None):
    return cells.prepare(first_cell, n_cells)
