"""Timing of product guiding (F4): GPU kernel (device-resident inputs, CUDA events) vs the reference's own
sdmm::product path on the host cores (oracle/_ref/libsdmm_ref_product.so) on a bounded sample."""
import json
import os
import sys
import numpy as np
import torch
sys.path.insert(0, "tests")
sys.path.insert(0, ".")
import test_gpu_product as T

K, C, n = int(os.environ.get("K", 64)), 256, 1 << 20
cells, ref, _, _, _, _ = T._setup(8, K, 16, seed=1)
# replicate 8 fitted cells over C slots
from sdmm_mitsuba_b200.mixture import SDMMCells
j, e, s = cells.download()
big = SDMMCells(C, K)
rep = C // 8
big.upload(np.tile(j, (rep, 1, 1)), np.tile(e, (rep, 1, 1)), np.tile(s, (rep, 1)))
big.prepare()
rng = np.random.default_rng(0)
cell = rng.integers(0, C, n).astype(np.uint32)
pos3 = rng.random((3, n), dtype=np.float32)
u3 = rng.random((3, n), dtype=np.float32)
nrm = rng.normal(size=(3, n)).astype(np.float32)
nrm /= np.linalg.norm(nrm, axis=0, keepdims=True)
bsdf = T._bsdf()
dev = torch.device("cuda", 0)
tc = torch.from_numpy(cell.view(np.int32)).to(dev)
tp, tu, tn = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (pos3, u3, nrm)]
for _ in range(3):
    big.sample_pdf_product(tc, tp, tu, bsdf, normal3=tn)
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 5
torch.cuda.synchronize()
ev0.record()
for _ in range(reps):
    big.sample_pdf_product(tc, tp, tu, bsdf, normal3=tn)
ev1.record()
torch.cuda.synchronize()
ms = ev0.elapsed_time(ev1) / reps
ev0.record()
for _ in range(reps):
    big.sample_pdf(tc, tp, tu)
ev1.record()
torch.cuda.synchronize()
ms_plain = ev0.elapsed_time(ev1) / reps
m = 1 << 16
secs = min(ref.product_sample_pdf(cell[:m] % 8, np.ascontiguousarray(pos3[:, :m]), bsdf,
                                  u3=np.ascontiguousarray(u3[:, :m]), normal3=np.ascontiguousarray(nrm[:, :m]))[1]
           for _ in range(3))
print(json.dumps({"K": K, "bsdf_components": 3, "vertices": n, "gpu_ms": ms, "gpu_product_queries_per_s": n / ms * 1e3,
                  "gpu_plain_query_ms": ms_plain, "cpu_ref_queries_per_s": m / secs, "cpu_threads": os.cpu_count(),
                  "cpu_sample": m}))
