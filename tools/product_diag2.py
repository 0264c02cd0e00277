"""Diagnostic: float64 evaluation of the product pdf from the GPU's and from the reference's conditioner planes,
for the vertices where GPU and reference differ most (is the gap input differences or float32 evaluation?)."""
import sys
import numpy as np
sys.path.insert(0, "tests")
sys.path.insert(0, ".")
import test_gpu_product as T


from product_f64 import product_pdf64  # noqa: E402


K, C, n = 128, 4, 6000
cells, ref, cell, pos3, u3, nrm = T._setup(C, K, n, seed=50 + K)
bsdf = T._bsdf()
rng = np.random.default_rng(3)
d = rng.normal(size=(3, n)).astype(np.float32)
d /= np.linalg.norm(d, axis=0, keepdims=True)
d = np.ascontiguousarray(d)
gp = cells.sample_pdf_product(cell, pos3, None, bsdf, normal3=nrm, dir3=d)
rp, _ = ref.product_sample_pdf(cell, pos3, bsdf, dir3=d, normal3=nrm)
Pg = cells.conditioner()
Pr = ref.conditioner()
big = rp["valid"].astype(bool) & (rp["pdf"] > 1e-6)
rel = np.where(big, np.abs(gp["pdf"] - rp["pdf"]) / np.maximum(rp["pdf"], 1e-30), 0)
order = np.argsort(-rel)[:8]
for i in list(order) + list(np.nonzero(big)[0][:4]):
    c = int(cell[i])
    eg = product_pdf64(Pg[c], pos3[:, i], bsdf, nrm[:, i], d[:, i])
    er = product_pdf64(Pr[c], pos3[:, i], bsdf, nrm[:, i], d[:, i])
    print(i, "gpu %.6e ref %.6e | f64(gpu planes) %.6e f64(ref planes) %.6e | rel gpu-ref %.2e, gpu-f64g %.2e, ref-f64r %.2e, f64g-f64r %.2e"
          % (gp["pdf"][i], rp["pdf"][i], eg, er, rel[i], abs(gp["pdf"][i] - eg) / eg, abs(rp["pdf"][i] - er) / er, abs(eg - er) / er))
