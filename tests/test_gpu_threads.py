"""The C ABI is callable from concurrent host threads (SURVEY 8b "Threading"): the reference queries from Mitsuba's
LocalWorker threads with thread_local scratch while optimize() fits (sdmm_proc.cpp:1086-1106,
volpath_sdmm.cpp:180-242).  Two threads issue guided queries on their own CUDA streams while a third runs training
steps on the same context; with the double-buffered records of the optimizeAsync schedule the queries read the
published records, so every result must be bit-identical to the serial answer."""
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_two_threads_query_while_a_third_fits():
    import torch
    from sdmm_mitsuba_b200 import synth
    from sdmm_mitsuba_b200.mixture import SDMMCells
    C, K, N, nq = 32, 32, 1024, 50000
    point6, weight, off = synth.make_samples(C, N, seed=21)
    joint, em, scal = synth.make_initial_model(C, K, seed=22)
    cells = SDMMCells(C, K)
    cells.upload(joint, em, scal)
    cells.set_samples(point6, weight, off)
    cells.em_step(2, with_conditioner=True)          # records to query
    cells.set_async_records(True)                    # from here on fits write the back copy (optimizeAsync)
    torch.cuda.synchronize()
    dev = torch.device("cuda", 0)
    queries, serial = [], []
    for t in range(2):
        cell, pos3, u3 = synth.make_queries(nq, C, seed=30 + t, binned=False)   # unsorted: exercises the binning scratch
        q = (torch.from_numpy(cell.view(np.int32)).to(dev), torch.from_numpy(pos3).to(dev), torch.from_numpy(u3).to(dev))
        queries.append(q)
        r = cells.sample_pdf(*q, binned=False)
        serial.append({k: v.clone() for k, v in r.items()})
    torch.cuda.synchronize()
    errors, stop = [], threading.Event()

    def query_worker(t):
        try:
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                for _ in range(20):
                    r = cells.sample_pdf(*queries[t], binned=False)
                    st.synchronize()
                    for k, v in serial[t].items():
                        if not torch.equal(r[k], v):
                            errors.append("thread %d: %s differs from the serial answer" % (t, k))
                            return
        except Exception as e:  # noqa: BLE001
            errors.append("thread %d: %r" % (t, e))
        finally:
            stop.set() if t == 0 else None

    def fit_worker():
        try:
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                n = 0
                while not stop.is_set() and n < 200:
                    cells.em_step(1, with_conditioner=True)
                    n += 1
                st.synchronize()
        except Exception as e:  # noqa: BLE001
            errors.append("fit thread: %r" % (e,))

    threads = [threading.Thread(target=query_worker, args=(t,)) for t in range(2)] + [threading.Thread(target=fit_worker)]
    for th in threads:
        th.start()
    for th in threads:
        th.join(timeout=300)
    assert not any(th.is_alive() for th in threads), "a worker hung"
    assert not errors, errors
    # the fits really ran and changed the (unpublished) model
    j, _, s = cells.download()
    assert int(s[:, 1].min()) > 2 and (s[:, 7] == 0).all()
    cells.publish_records()
    torch.cuda.synchronize()
    r = cells.sample_pdf(*queries[0], binned=False)
    assert not torch.equal(r["pdf"], serial[0]["pdf"])   # after publishing, queries see the new records
