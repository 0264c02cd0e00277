"""bench.py's third leg: the progressive training loop on the Cornell box (BASELINE.json configs[0] / [3])
through sdmm_mitsuba_b200.integrator, with the CPU restatement of the reference integrator (oracle/ref_render.inc,
the reference's own sdmm code on the host cores) timed beside it on a bounded sample.  Reported, not the headline."""
import time


def run_render(torch, width=640, height=360, K=16, max_depth=5, spi=64, sample_count=512, cpu=True):
    from sdmm_mitsuba_b200 import scenes
    from sdmm_mitsuba_b200.integrator import SDMMIntegrator
    quads, cam = scenes.cornell_box(width, height)
    out = dict(workload="Cornell box (test-suite/scenes/cornell-box), %dx%d, K=%d, maxDepth=%d, %d spp per iteration, "
                        "%d spp, training while spp < 1/4 (the reference's schedule)" % (width, height, K, max_depth, spi, sample_count))

    def loop():
        g = SDMMIntegrator(quads, cam, K=K, max_depth=max_depth, samples_per_iteration=spi, seed=11)
        t_r = t_o = 0.0
        rendered, q, dropped, trained_on = 0, 0, 0, 0
        while rendered < sample_count:
            train = rendered < sample_count / 4
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            st, new = g.render_pass(spi, collect=train)
            torch.cuda.synchronize()
            t1 = time.perf_counter()
            if train:
                trained_on += g.optimize(new).get("train_samples", 0)
            dropped += int(st.pool_overflow)
            torch.cuda.synchronize()
            t2 = time.perf_counter()
            t_r += t1 - t0
            t_o += t2 - t1
            q += int(st.guided_queries) + int(st.mis_queries)
            g.iteration += 1
            g.total_spp += spi
            rendered += spi
        launches = g.cells.launch_count()
        leaves, trained = g.tree.n_leaves, int(g.initialized.sum())
        g.close()
        return t_r, t_o, q, launches, leaves, trained, dropped, trained_on

    def whole(optimize_async):
        # the whole loop through SDMMIntegrator.render(), no synchronisation between the phases
        g = SDMMIntegrator(quads, cam, K=K, max_depth=max_depth, samples_per_iteration=spi, seed=11,
                           optimize_async=optimize_async)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        g.render(sample_count)
        torch.cuda.synchronize()
        secs = time.perf_counter() - t0
        g.close()
        return secs

    loop()                                   # warm-up (allocations, first-launch costs)
    # every repetition builds a fresh integrator, i.e. allocates its device state anew: the best of three keeps a
    # slow allocation out of the number (as `whole` and the interior leg do)
    t_r, t_o, q, launches, leaves, trained, dropped, trained_on = min((loop() for _ in range(3)), key=lambda r: r[0] + r[1])
    paths = width * height * sample_count
    out.update(paths=paths, seconds=t_r + t_o, render_seconds=t_r, optimize_seconds=t_o,
               paths_per_s=paths / (t_r + t_o), mixture_queries=q, gpu_launches=int(launches), leaves=leaves,
               guided_cells=trained, train_samples=int(trained_on), samples_dropped=int(dropped))
    whole(True)
    out["schedules"] = dict(
        note="whole render() call, wall clock: optimize() as volpath_sdmm.cpp:244-312 vs the optimizeAsync schedule "
             "(:180-242: fits on a second stream, visible one pass later, one em_step per fit)",
        synchronous_seconds=min(whole(False) for _ in range(2)), async_seconds=min(whole(True) for _ in range(2)))
    out["interior_doubling"] = run_interior(torch, cpu=cpu)
    if cpu:
        try:
            from oracle import refbind
            if refbind.available():
                w2, h2 = width // 2, height // 2
                quads2, cam2 = scenes.cornell_box(w2, h2)
                lo, hi, norm = scenes.scene_bounds(quads2)
                rmin, rmax = scenes.root_aabb(lo, hi, norm)
                r = refbind.RefRenderer(quads2, cam2, lo, norm, rmin, rmax, K=K, max_depth=max_depth, capacity=700000)
                t0 = time.perf_counter()
                rendered = 0
                it = 0
                while rendered < sample_count:
                    r.iteration(spi, 1000 + it, rendered < sample_count / 4)
                    rendered += spi
                    it += 1
                secs = time.perf_counter() - t0
                p2 = w2 * h2 * sample_count
                out["cpu_baseline"] = dict(value=p2 / secs, unit="paths/s", cores=int(refbind.lib().sdmm_ref_max_threads()),
                                           kind="reference",
                                           sample="the same loop at %dx%d (1/4 of the pixels), %.1f s; mixture, tree and EM code is the "
                                                  "reference's own, the ray caster a restatement" % (w2, h2, secs))
        except Exception as e:   # the CPU leg is informative only
            out["cpu_baseline"] = dict(error=str(e))
    return out


def run_interior(torch, width=640, height=360, K=16, max_depth=6, first_spp=8, sample_count=504, cpu=True):
    """BASELINE.json configs[3]: progressive loop with the doubling schedule and a refit after every iteration on the
    procedurally generated three-room interior (scenes.interior; every test-suite mesh is an LFS pointer)."""
    from sdmm_mitsuba_b200 import scenes
    from sdmm_mitsuba_b200.integrator import SDMMIntegrator
    quads, cam = scenes.interior(width, height)
    out = dict(workload="three-room interior of analytic boxes (scenes.interior, %d quads), %dx%d, K=%d, maxDepth=%d, "
                        "doubling schedule from %d spp to %d spp in total, stepwise EM refit after every iteration"
                        % (len(quads), width, height, K, max_depth, first_spp, sample_count))

    def whole():
        g = SDMMIntegrator(quads, cam, K=K, max_depth=max_depth, samples_per_iteration=first_spp, seed=5,
                           pool_capacity=96_000_000)   # 5.8 GB: every sample of the largest pass is kept
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        g.render(sample_count, train_fraction=1.0, schedule="doubling")
        torch.cuda.synchronize()
        secs = time.perf_counter() - t0
        info = dict(iterations=len(g.stats), leaves=g.tree.n_leaves, guided_cells=int(g.initialized.sum()),
                    samples_dropped=int(sum(s["overflow"] for s in g.stats)),
                    train_samples=int(sum(s.get("train_samples", 0) for s in g.stats)),
                    mixture_queries=int(sum(s["guided_queries"] + s["mis_queries"] for s in g.stats)))
        g.close()
        return secs, info

    whole()
    secs, info = min((whole() for _ in range(3)), key=lambda r: r[0])
    paths = width * height * sample_count
    out.update(paths=paths, seconds=secs, paths_per_s=paths / secs, **info)
    if cpu:
        try:
            from oracle import refbind
            if refbind.available():
                w2, h2 = width // 4, height // 4
                quads2, cam2 = scenes.interior(w2, h2)
                lo, hi, norm = scenes.scene_bounds(quads2)
                rmin, rmax = scenes.root_aabb(lo, hi, norm)
                r = refbind.RefRenderer(quads2, cam2, lo, norm, rmin, rmax, K=K, max_depth=max_depth, capacity=400000)
                t0 = time.perf_counter()
                rendered, spp, it = 0, first_spp, 0
                while rendered < sample_count:
                    spp = min(spp, sample_count - rendered)
                    r.iteration(spp, 2000 + it, True)
                    rendered += spp
                    spp *= 2
                    it += 1
                secs2 = time.perf_counter() - t0
                p2 = w2 * h2 * sample_count
                out["cpu_baseline"] = dict(value=p2 / secs2, unit="paths/s", cores=int(refbind.lib().sdmm_ref_max_threads()),
                                           kind="reference", sample="same loop at %dx%d (1/16 of the pixels), %.1f s" % (w2, h2, secs2))
        except Exception as e:
            out["cpu_baseline"] = dict(error=str(e))
    return out
