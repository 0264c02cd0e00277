"""sdmmb_sample_pdf_noalloc (static device scratch, no allocation inside the call) against the existing
sdmmb_sample_pdf(binned=0) through the C ABI, on two calls in a row.  Collected last on purpose
(file name)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _cells(n_cells, K, seed):
    import torch  # noqa: F401
    from sdmm_mitsuba_b200 import synth
    from sdmm_mitsuba_b200.mixture import SDMMCells
    point6, weight, off = synth.make_samples(n_cells, 512, seed=seed)
    joint, em, scal = synth.make_initial_model(n_cells, K, seed=seed + 1)
    cells = SDMMCells(n_cells, K, device=0)
    cells.upload(joint, em, scal)
    cells.set_samples(point6, weight, off)
    cells.em_step(2, with_conditioner=True)
    return cells


def _queries(n, n_cells, seed):
    import torch
    from sdmm_mitsuba_b200 import synth
    cell, pos3, u3 = synth.make_queries(n, n_cells, seed=seed, binned=False)
    dev = torch.device("cuda", 0)
    return (torch.from_numpy(np.ascontiguousarray(cell).view(np.int32)).to(dev),
            torch.from_numpy(np.ascontiguousarray(pos3, np.float32)).to(dev),
            torch.from_numpy(np.ascontiguousarray(u3, np.float32)).to(dev))


def _outputs(n):
    import torch
    dev = torch.device("cuda", 0)
    return dict(dir=torch.full((3, n), -7.0, dtype=torch.float32, device=dev),
                pdf=torch.full((n,), -7.0, dtype=torch.float32, device=dev),
                inv_jac=torch.full((n,), -7.0, dtype=torch.float32, device=dev),
                comp=torch.full((n,), -7, dtype=torch.int32, device=dev),
                valid=torch.full((n,), 9, dtype=torch.uint8, device=dev))


def _call_noalloc(cells, cell, pos3, u3, out, n=None):
    from sdmm_mitsuba_b200.mixture import _stream
    p = lambda t: C.c_void_p(t.data_ptr())
    return cells._L.sdmmb_sample_pdf_noalloc(cells._h, pos3.shape[1] if n is None else n, p(cell), p(pos3), p(u3),
                                             p(out["dir"]), p(out["pdf"]), p(out["inv_jac"]), p(out["comp"]),
                                             p(out["valid"]), _stream(cells.device))


@pytest.mark.parametrize("K", [64, 16])
def test_noalloc_query_equals_allocating_query_on_two_calls_in_a_row(K):
    import torch
    n_cells = 37
    cells = _cells(n_cells, K, seed=11 + K)
    # two different batches: the second is SHORTER and differently distributed, so stale words of the first
    # call (cursor, invperm, sorted records) sit behind and inside the region the second call uses
    batches = [_queries(50001, n_cells, seed=3), _queries(12345, n_cells, seed=4)]
    refs = [cells.sample_pdf(c, p, u, binned=False) for (c, p, u) in batches]
    torch.cuda.synchronize()
    # warm-up call: symbol look-up (module load) happens here, as the header says
    outs0 = _outputs(batches[0][1].shape[1])
    assert _call_noalloc(cells, *batches[0], outs0) == 0
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(0)[0]
    got = []
    for (c, p, u) in batches:
        out = _outputs(p.shape[1])   # torch allocations are the TEST's, made outside the entry point
        free_mid = torch.cuda.mem_get_info(0)[0]
        rc = _call_noalloc(cells, c, p, u, out)
        assert rc == 0, cells._L.sdmmb_get_error_string(rc)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info(0)[0] == free_mid, "the call changed the device's free memory"
        got.append(out)
    assert torch.cuda.mem_get_info(0)[0] <= free_before
    for ref, out in zip(refs, got):
        v = ref["valid"].cpu().numpy().astype(bool)
        assert v.mean() > 0.5
        np.testing.assert_array_equal(out["valid"].cpu().numpy(), ref["valid"].cpu().numpy())
        np.testing.assert_array_equal(out["comp"].cpu().numpy(), ref["comp"].cpu().numpy())
        # same kernels over a different scratch address: bit-identical (tolerance 0), as the queries of
        # tests/test_gpu_query.py::test_pdf_only_parity_and_binned_equals_gather are between binned and gathered
        np.testing.assert_array_equal(out["pdf"].cpu().numpy(), ref["pdf"].cpu().numpy())
        np.testing.assert_array_equal(out["inv_jac"].cpu().numpy(), ref["inv_jac"].cpu().numpy())
        np.testing.assert_array_equal(out["dir"].cpu().numpy(), ref["dir"].cpu().numpy())
        assert np.isfinite(out["pdf"].cpu().numpy()[v]).all()


def test_noalloc_query_rejects_sizes_beyond_the_static_scratch():
    from sdmm_mitsuba_b200 import capi
    from sdmm_mitsuba_b200.mixture import SDMMCells
    n_cells = 4
    cells = _cells(n_cells, 16, seed=5)
    c, p, u = _queries(64, n_cells, seed=6)
    out = _outputs(64)
    before = {k: v.clone() for k, v in out.items()}
    rc = _call_noalloc(cells, c, p, u, out, n=capi.NOALLOC_MAX_VERTICES + 1)   # rejected before any pointer is read
    assert rc == capi.E_LIMIT
    assert b"SDMMB_NOALLOC_MAX_VERTICES" in cells._L.sdmmb_get_error_string(rc)
    import torch
    torch.cuda.synchronize()
    for k in out:
        assert torch.equal(out[k], before[k]), k   # nothing was launched
    big = SDMMCells(capi.NOALLOC_MAX_CELLS + 1, 16, device=0)   # a context above the cell limit
    assert _call_noalloc(big, c, p, u, out) == capi.E_LIMIT
    assert _call_noalloc(cells, c, p, u, out, n=0) == 0
    assert _call_noalloc(cells, c, p, u, out) == 0   # at the limits' inside it still works after the rejections
    torch.cuda.synchronize()
    assert (out["comp"].cpu().numpy() >= 0).all()
