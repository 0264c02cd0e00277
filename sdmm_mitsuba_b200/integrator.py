"""Progressive training loop with device-resident mixtures: the host side of
the reference integrator's render() / optimize() for analytic scenes.

Reference mirrored (paths relative to /root/reference/mitsuba/src/integrators/sdmm):
  SDMMVolumetricPathTracer::render    volpath_sdmm.cpp:334-516   iteration loop, m_still_training, m_totalSpp
  optimize()                          volpath_sdmm.cpp:244-312   split, canBeOptimized, initialize, em_step x 2|1,
                                                                 prepare, data.clear()
  canBeOptimized                      volpath_sdmm.cpp:140-149
  initializeSDMMContext               volpath_sdmm.cpp:132-137
The render pass itself (Li, sampleSurface, sample collection) runs on the device: sdmmb_pt_render
(include/sdmm_b200_render.h).  Everything numeric runs in the CUDA library; torch is used here for device
buffers and for index glue (sorting the sample pool by cell, per-cell counts and SpatialStats sums).
Synchronous schedule (optimizeAsync = false, as scripts/run_tests.sh:29 runs the reference)."""
import contextlib
import ctypes as C

import numpy as np

from . import capi, scenes
from .capi import check
from .mixture import SDMMCells
from .tree import HostTree

SPLIT_THRESHOLD = 4000      # volpath_sdmm.cpp:528
MAX_LEAF_NODES = 2048       # volpath_sdmm.cpp:529
POOL_PLANES = 15


class SDMMIntegrator:
    def __init__(self, quads, camera, K=16, max_depth=5, samples_per_iteration=4, saved_per_path=8, max_cells=4096,
                 pool_capacity=8_000_000, device=0, seed=1234, split_threshold=SPLIT_THRESHOLD, optimize_async=False):
        import torch
        self.torch = torch
        self.dev = torch.device("cuda", device)
        self.K, self.spi, self.max_depth = int(K), int(samples_per_iteration), int(max_depth)
        self.seed = int(seed)
        self.split_threshold = int(split_threshold)
        assert max_cells <= 65536, "the (cell, pass, sequence) sort key of optimize() holds 16 bits of cell index"
        self.cells = SDMMCells(max_cells, K, device=device)
        self.cells.set_min_samples(1)
        # optimizeAsync (the reference's default, volpath_sdmm.cpp:65, 180-242): a fit's records and `initialized`
        # flags become visible to the renderer one pass later; the records are double-buffered on the device
        self.optimize_async = bool(optimize_async)
        self._pending = None
        self._em_stream = None
        self._pending_keep = None
        if self.optimize_async:
            self.cells.set_async_records(True)
            self._em_stream = torch.cuda.Stream(device=self.dev)   # the fit overlaps the next render pass
        lo, hi, norm = scenes.scene_bounds(quads)
        rmin, rmax = scenes.root_aabb(lo, hi, norm)
        self.tree = HostTree(rmin, rmax, max_cells)
        self.tree.split_to_depth(2)                                   # volpath_sdmm.cpp:398
        self.cells.set_tree(self.tree.flatten())
        self.width, self.height = int(camera["width"]), int(camera["height"])
        qs = (capi.Quad * len(quads))()
        for i, q in enumerate(quads):
            for name in ("p0", "e1", "e2", "reflectance", "emission"):
                getattr(qs[i], name)[:] = [float(v) for v in q[name]]
        cam = capi.Camera()
        for name in ("origin", "right", "up", "forward"):
            getattr(cam, name)[:] = [float(v) for v in camera[name]]
        cam.fov_x_degrees, cam.width, cam.height = float(camera["fov_x_degrees"]), self.width, self.height
        cfg = capi.PtConfig()
        cfg.aabb_min[:] = [float(v) for v in lo]
        cfg.spatial_norm = float(norm)
        cfg.root_min[:] = [float(v) for v in rmin]
        cfg.root_max[:] = [float(v) for v in rmax]
        cfg.max_depth, cfg.saved_per_path = self.max_depth, int(saved_per_path)
        self._pt = C.c_void_p()
        check(self.cells._L.sdmmb_pt_create(self.cells._h, C.byref(self._pt), qs, len(quads), C.byref(cam), C.byref(cfg)),
              "sdmmb_pt_create", self.cells._h)
        self.pool = torch.zeros((POOL_PLANES, pool_capacity), dtype=torch.float32, device=self.dev)
        self.pool_count = 0
        self.initialized = torch.zeros(max_cells, dtype=torch.uint8, device=self.dev)   # SDMMContext::initialized
        self.has_mixture = np.zeros(max_cells, bool)                                   # slices(context->sdmm) != 0
        self.image_sum = torch.zeros((3, self.height * self.width), dtype=torch.float32, device=self.dev)
        self.iteration = 0
        self.total_spp = 0
        self.stats = []
        self._acc_size = np.zeros(max_cells, np.int64)          # SpatialStats not yet written into the host tree, per slot
        self._acc_sum = np.zeros((3, max_cells), np.float64)
        self._acc_sqr = np.zeros((3, max_cells), np.float64)
        self._leaf_version = -1
        self.capture_fit = None     # optional callable(dict) invoked right before a fit with its samples (tests)

    def set_product(self, bsdf):
        """sampleProduct (sdmm_proc.cpp:326-392): guided vertices sample conditional x learned-BSDF mixture; `bsdf` =
        [M, 7] rows (pmf, mean3, c00, c10, c11), one mixture for the whole (diffuse) scene, component 0 re-centred on
        the shading normal per vertex; the BSDF fraction of the MIS drops to 0.3.  None / empty switches it off."""
        b = np.zeros((0, 7), np.float32) if bsdf is None else np.ascontiguousarray(bsdf, np.float32).reshape(-1, 7)
        check(self.cells._L.sdmmb_pt_set_product(self._pt, int(b.shape[0]), b.ctypes.data_as(C.POINTER(C.c_float))),
              "sdmmb_pt_set_product", self.cells._h)

    def close(self):
        if getattr(self, "_pt", None) is not None and self._pt:
            self.cells._L.sdmmb_pt_destroy(self._pt)
            self._pt = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- one render pass (SDMMProcess::process) ---------------------------------------------------
    def render_pass(self, spp, collect, guided=None, image_sum=None):
        torch = self.torch
        guided = (self.iteration != 0) if guided is None else guided        # sdmm_proc.cpp:316-323
        img = self.image_sum if image_sum is None else image_sum
        st = capi.PtStats()
        seed = (self.seed * 0x9E3779B97F4A7C15 + self.iteration) & 0xFFFFFFFFFFFFFFFF
        check(self.cells._L.sdmmb_pt_render(self._pt, int(spp), seed, 1 if guided else 0, 1 if collect else 0,
                                            C.c_void_p(self.initialized.data_ptr()), C.c_void_p(img.data_ptr()),
                                            C.c_void_p(self.pool.data_ptr()), self.pool.shape[1], self.pool_count,
                                            C.byref(st), C.c_void_p(torch.cuda.current_stream(self.dev).cuda_stream)),
              "sdmmb_pt_render", self.cells._h)
        new = (self.pool_count, int(st.pool_count))
        self.pool_count = int(st.pool_count)
        if collect and new[1] > new[0]:
            self.pool[13, new[0]:new[1]] += 2.0 * (self.iteration & 0x7FFF)   # plane 13 = is_main + 2 * pass number (exact in float32)
        # the pool is filled in scheduling order; optimize() sorts the training samples of a cell by their push
        # sequence number (plane 14), the order a single-threaded reference would have pushed them in
        return st, new

    # ---- optimize() -------------------------------------------------------------------------------
    def _refresh_leaf_cache(self):
        """Per-slot views of the host tree's leaves (SpatialStats size, largest box extent); rebuilt when the tree grew."""
        t, C_ = self.tree, self.cells.max_cells
        self._st_base = np.zeros(C_, np.int64)
        self._diag = np.zeros(C_, np.float32)
        for node in t.leaves():
            self._st_base[t.slot[node]] = t.st_size[node]
            self._diag[t.slot[node]] = float((t.amax[node] - t.amin[node]).max())
        self._leaf_version = t.size()

    def _flush_leaf_stats(self):
        """Write the per-slot SpatialStats accumulated since the last flush into the tree's leaves."""
        t = self.tree
        if self._acc_size.any():
            for node in t.leaves():
                sl = t.slot[node]
                if self._acc_size[sl]:
                    t.st_size[node] += int(self._acc_size[sl])
                    t.st_sum[node] = t.st_sum[node] + self._acc_sum[:, sl]
                    t.st_sqr[node] = t.st_sqr[node] + self._acc_sqr[:, sl]
            self._acc_size[:] = 0
            self._acc_sum[:] = 0
            self._acc_sqr[:] = 0
        self._leaf_version = -1     # st_size changed: the per-slot base is rebuilt before its next use

    def optimize(self, new_range):
        torch = self.torch
        t = self.tree
        n = self.pool_count
        if n == 0:
            return dict(trained=0, leaves=t.n_leaves)
        # cell of every pooled sample = the Data it was pushed into (key = vertex or its jittered copy), the per-cell
        # sample counts and the SpatialStats of the leaves (only the main pushes of this pass feed them,
        # sdmm_proc.cpp:893-895): one kernel over the pool, one read-back
        lo, hi = new_range
        cell32 = torch.empty(n, dtype=torch.int32, device=self.dev)
        stats_d = torch.empty((8, self.cells.max_cells), dtype=torch.float64, device=self.dev)
        check(self.cells._L.sdmmb_pt_pool_stats(self.cells._h, C.c_void_p(self.pool.data_ptr()), self.pool.shape[1], n,
                                                min(lo, n), C.c_void_p(cell32.data_ptr()), C.c_void_p(stats_d.data_ptr()),
                                                C.c_void_p(torch.cuda.current_stream(self.dev).cuda_stream)),
              "sdmmb_pt_pool_stats", self.cells._h)
        stats_h = stats_d.cpu().numpy()
        data_size = np.rint(stats_h[7]).astype(np.int64)
        # The SpatialStats are accumulated per SLOT (vector adds) and only written into the host tree's per-node records
        # when a split is still possible: a loop over thousands of leaves per iteration is otherwise the largest host cost.
        if hi > lo:
            self._acc_size += np.rint(stats_h[0]).astype(np.int64)
            self._acc_sum += stats_h[1:4]
            self._acc_sqr += stats_h[4:7]
        if t.n_leaves <= MAX_LEAF_NODES:
            self._flush_leaf_stats()
        # data-driven splits (volpath_sdmm.cpp:253-259); children inherit the parent's state
        if t.n_leaves <= MAX_LEAF_NODES and any(t.st_size[node] > self.split_threshold for node in t.leaves()):
            # Level-synchronous: the host decides which leaves split (HostTree, the reference's rule) and ONE kernel per
            # recursion level re-buckets the samples of all splitting leaves and reduces their children's SpatialStats;
            # HostTree then renumbers the new nodes into the reference's depth-first creation order.
            NC = self.cells.max_cells
            L, h = self.cells._L, self.cells._h
            ptr = lambda x: C.c_void_p(x.data_ptr())

            def rebucket(level):
                tab_i = np.full((3, NC), -1, np.int32)            # per parent slot: axis, slot0, slot1
                tab_f = np.zeros((2, NC), np.float32)             # c0_min, c1_max
                box = np.zeros((2, 3, NC), np.float32)            # per CHILD slot: aabb min / max
                ps = np.fromiter((e["parent_slot"] for e in level), np.int64, len(level))
                sl0 = np.fromiter((e["slot0"] for e in level), np.int64, len(level))
                sl1 = np.fromiter((e["slot1"] for e in level), np.int64, len(level))
                tab_i[0, ps] = np.fromiter((e["axis"] for e in level), np.int32, len(level))
                tab_i[1, ps], tab_i[2, ps] = sl0, sl1
                tab_f[0, ps] = np.fromiter((e["c0_min"] for e in level), np.float32, len(level))
                tab_f[1, ps] = np.fromiter((e["c1_max"] for e in level), np.float32, len(level))
                box[0][:, sl0] = np.array([e["amin0"] for e in level], np.float32).T
                box[1][:, sl0] = np.array([e["amax0"] for e in level], np.float32).T
                box[0][:, sl1] = np.array([e["amin1"] for e in level], np.float32).T
                box[1][:, sl1] = np.array([e["amax1"] for e in level], np.float32).T
                tab_i_t = torch.as_tensor(tab_i, device=self.dev)
                tab_f_t = torch.as_tensor(tab_f, device=self.dev)
                box_t = torch.as_tensor(box, device=self.dev)
                st_d = torch.empty((7, NC), dtype=torch.float64, device=self.dev)
                check(L.sdmmb_pt_rebucket(h, ptr(self.pool), self.pool.shape[1], n, ptr(cell32), ptr(tab_i_t), ptr(tab_f_t),
                                          ptr(box_t), ptr(st_d), C.c_void_p(torch.cuda.current_stream(self.dev).cuda_stream)),
                      "sdmmb_pt_rebucket", h)
                sh = st_d.cpu().numpy()
                size = np.rint(sh[0]).astype(np.int64)
                ssum, ssqr = np.ascontiguousarray(sh[1:4].T), np.ascontiguousarray(sh[4:7].T)    # [slot][3]
                return [((int(size[a]), ssum[a], ssqr[a]), (int(size[b]), ssum[b], ssqr[b])) for a, b in zip(sl0, sl1)]

            splits, slot_map = t.split_leaves_levels(self.split_threshold, rebucket)
            if slot_map:
                lut = np.arange(NC, dtype=np.int32)
                for a_, b_ in slot_map.items():
                    lut[a_] = b_
                lut_t = torch.as_tensor(lut, device=self.dev)
                cell32 = torch.where(cell32 >= 0, lut_t[cell32.clamp(min=0).long()], cell32)
            if splits:
                src = [s[0] for s in splits]
                dst = [s[2] for s in splits]                              # child 0 keeps the parent's slot
                self.cells.copy_cells(src, dst)
                init_h = self.initialized.cpu().numpy()
                for a_, b_ in zip(src, dst):        # sequential: a fresh child can itself be split in the same pass
                    init_h[b_] = init_h[a_]
                    self.has_mixture[b_] = self.has_mixture[a_]
                self.initialized.copy_(torch.as_tensor(init_h, device=self.dev))
                self.cells.set_tree(t.flatten())
                # keep find(key) consistent with the re-bucketing: move the key of a sample whose key now resolves to
                # a different leaf to the centre of the leaf that owns it
                found = self.cells.find_cells(self.pool[10:13, :n].contiguous())
                centre = np.zeros((NC, 3), np.float32)
                for node in t.leaves():
                    centre[t.slot[node]] = 0.5 * (t.amin[node] + t.amax[node])
                bad = (found != cell32) & (cell32 >= 0)
                if bool(bad.any()):
                    self.pool[10:13, :n][:, bad] = torch.as_tensor(centre, device=self.dev)[cell32[bad].long()].t()
                dropped = cell32 < 0
                if bool(dropped.any()):
                    keepm = ~dropped
                    kept0 = self.pool[:, :n][:, keepm]
                    n = int(kept0.shape[1])
                    self.pool[:, :n] = kept0
                    self.pool_count = n
                    cell32 = cell32[keepm]
                data_size = torch.bincount(cell32.long(), minlength=NC).cpu().numpy()
        # canBeOptimized (volpath_sdmm.cpp:140-149)
        cells_all = cell32[:n].to(torch.int64)
        if self._leaf_version != t.size():
            self._refresh_leaf_cache()
        stats_size = self._st_base + self._acc_size
        diag = self._diag
        can = ((self.total_spp > 12) | (data_size > 1000)) & (stats_size >= 64) & (data_size >= 8)
        if not can.any():
            return dict(trained=0, leaves=t.n_leaves)
        can_t = torch.as_tensor(can, device=self.dev)
        counts = np.where(can, data_size, 0)
        all_train = int(counts.sum()) == n              # every pooled sample belongs to a cell that trains now
        # the swapped training_data of every cell: ONE sort of the pool indices by (cell, pass, push sequence) -- older
        # passes first, then the order a single-threaded reference would have pushed in -- and one gather of the ten
        # planes the fit reads
        seq = self.pool[14, :n].view(torch.int32).to(torch.int64) & 0xFFFFFFFF
        passno = (self.pool[13, :n] * 0.5).to(torch.int64)
        key = (cells_all << 47) | (passno << 32) | seq                   # cell < 2^16, pass < 2^15, seq < 2^32
        if all_train:
            sel = None
            idx = torch.sort(key)[1]
        else:
            sel = (cells_all >= 0) & can_t[cells_all.clamp(min=0)]
            base = torch.nonzero(sel).squeeze(1)
            idx = base[torch.sort(key[base])[1]]
        del key, seq, passno
        fresh = np.nonzero(can & ~self.has_mixture)[0]                   # cells sdmm::initialize will seed below
        off = np.concatenate([[0], np.cumsum(counts)]).astype(np.uint32)
        if fresh.size:          # the normals are only read by sdmm::initialize
            train = self.pool[0:10, :n].index_select(1, idx)             # sorted by cell
            point6, normal3, weight = train[0:6], train[6:9], train[9]
        else:
            point6 = self.pool[0:6, :n].index_select(1, idx)
            weight = self.pool[9, :n].index_select(0, idx)
            train, normal3 = point6, None
        # async schedule: the fit runs on its own stream, ordered after everything queued so far, and is joined in
        # publish_pending() after the next pass; the training copies must outlive it
        if self.optimize_async:
            self._em_stream.wait_stream(torch.cuda.current_stream(self.dev))
            self._pending_keep = (train, point6, normal3, weight)
        with (torch.cuda.stream(self._em_stream) if self.optimize_async else contextlib.nullcontext()):
            # sdmm::initialize for cells without a mixture (volpath_sdmm.cpp:292-294)
            if fresh.size:
                runs = np.split(fresh, np.nonzero(np.diff(fresh) != 1)[0] + 1)
                for run in runs:
                    a, b = int(run[0]), int(run[-1]) + 1
                    # optimize_async_run seeds from 0.1 x the leaf diagonal (volpath_sdmm.cpp:215), optimize from all of it
                    sd = (3.0 * (0.1 if self.optimize_async else 1.0) * diag[a:b] / (self.K // 8)).astype(np.float32)
                    base = int(off[a])
                    sub_off = (off[a:b + 1] - off[a]).astype(np.uint32)
                    self.cells.initialize(point6[:, base:base + int(sub_off[-1])].contiguous(),
                                          normal3[:, base:base + int(sub_off[-1])].contiguous(),
                                          weight[base:base + int(sub_off[-1])].contiguous(), sub_off, sd, first_cell=a)
                self.has_mixture[fresh] = True
            # 2 em_steps while iterations_run < 4, else 1 (volpath_sdmm.cpp:299-305): the rule is applied per cell ON
            # the device (SDMMB_ITERS_INTEGRATOR), no read-back of the iteration counters and no second gather of the
            # young cells' samples; optimize_async_run: always 1 (:217)
            if self.capture_fit is not None:   # test hook: the inputs of this fit (state is read through self.cells)
                self.capture_fit(dict(point6=point6, weight=weight, offsets=off, can=can))
            self.cells.set_samples(point6, weight, off)
            self.cells.em_step(1 if self.optimize_async else capi.ITERS_INTEGRATOR, with_conditioner=True)
        if self.optimize_async:
            self._pending = can_t          # published by publish_pending() after the next pass
        else:
            self.initialized[can_t] = 1
        # data.clear() for the optimised cells; everything else stays in the pool
        if all_train:
            self.pool_count = 0
        else:
            kept = self.pool[:, :n][:, ~sel]
            self.pool_count = int(kept.shape[1])
            self.pool[:, :self.pool_count] = kept
        return dict(trained=int(can.sum()), leaves=t.n_leaves, fresh=int(fresh.size), train_samples=int(off[-1]))

    def publish_pending(self):
        """optimize_async_wait_and_update (volpath_sdmm.cpp:227-242): make the last fit visible to the renderer."""
        if self._pending is not None:
            self.torch.cuda.current_stream(self.dev).wait_stream(self._em_stream)
            self._pending_keep = None
            self.cells.publish_records()
            self.initialized[self._pending] = 1
            self._pending = None

    # ---- render() ---------------------------------------------------------------------------------
    def render(self, sample_count, train_fraction=0.25, log=None, schedule="fixed"):
        """The reference's schedule ("fixed"): iterations of samples_per_iteration spp; sample collection + training
        while samplesRendered < sampleCount * train_fraction (1/4 in the reference, volpath_sdmm.cpp:411-416).
        "doubling" (BASELINE.json configs[3]; PPG's schedule, test-suite/scenes/_integrators/ppg.xml): iteration i
        renders samples_per_iteration * 2^i spp, with a refit after every iteration when train_fraction = 1."""
        rendered, spp = 0, self.spi
        while rendered < sample_count:
            spp = min(spp, sample_count - rendered)
            still_training = rendered < sample_count * train_fraction
            st, new = self.render_pass(spp, collect=still_training)
            self.publish_pending()                                    # volpath_sdmm.cpp:446-448
            info = dict(iteration=self.iteration, spp=spp, guided_queries=int(st.guided_queries),
                        mis_queries=int(st.mis_queries), pool=self.pool_count, overflow=int(st.pool_overflow))
            if still_training:
                info.update(self.optimize(new))        # before m_totalSpp advances (volpath_sdmm.cpp:494-503)
            self.stats.append(info)
            if log:
                log(info)
            self.iteration += 1
            self.total_spp += spp
            rendered += spp
            if schedule == "doubling":
                spp *= 2
        self._flush_leaf_stats()      # the host tree's SpatialStats are complete again when render() returns
        return self.image()

    def save_checkpoint(self, experiment_path):
        """saveCheckpoint (volpath_sdmm.cpp:121-130, called at the end of render(), :513): the accelerator -- node
        table + every leaf's joint mixture -- as `<experiment_path>/checkpoints/model_%05i.asdmm` in the reference's
        JSON layout (checkpoint.py); leaves that never received a mixture are written with value = null like the
        reference's empty contexts would be with size 0.  Also stores the EM statistics (extra keys)."""
        import os
        from . import checkpoint
        d = os.path.join(str(experiment_path), "checkpoints")
        os.makedirs(d, exist_ok=True)
        path = os.path.join(d, "model_%05i.asdmm" % self.iteration)
        self.publish_pending()
        joint, em, scal = self.cells.download()
        slots = [self.tree.slot[i] for i in range(self.tree.size()) if self.tree.is_leaf[i]]
        keep = np.zeros(joint.shape[0], bool)
        keep[[s for s in slots if self.has_mixture[s]]] = True
        tree = self.tree
        if not keep[slots].all():       # hide the slots of mixture-less leaves from the writer
            import copy
            tree = copy.copy(self.tree)
            tree.slot = [(-1 if (s >= 0 and not keep[s]) else s) for s in self.tree.slot]
        checkpoint.save_asdmm(path, tree, joint, K=self.K, em_by_slot=em, scal_by_slot=scal)
        return path

    def image(self):
        img = (self.image_sum / max(1, self.total_spp)).reshape(3, self.height, self.width)
        return img.cpu().numpy()
