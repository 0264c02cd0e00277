"""Diagnostic: product guiding GPU vs reference -- mismatch statistics."""
import sys
import numpy as np
sys.path.insert(0, "tests")
sys.path.insert(0, ".")
import test_gpu_product as T

for K, diffuse in [(16, True), (64, True), (64, False), (128, True)]:
    C, n = 4, 6000
    cells, ref, cell, pos3, u3, nrm = T._setup(C, K, n, seed=50 + K)
    bsdf = T._bsdf()
    normal3 = nrm if diffuse else None
    g = cells.sample_pdf_product(cell, pos3, u3, bsdf, normal3=normal3)
    r, _ = ref.product_sample_pdf(cell, pos3, bsdf, u3=u3, normal3=normal3)
    v = r["valid"].astype(bool)
    print(K, diffuse, "valid mismatch", (g["valid"] != r["valid"]).sum(), "valid", v.sum())
    diff = np.nonzero((g["comp"] != r["comp"]) & v)[0]
    dist = []
    for i in diff:
        comps = ref.product_components(int(cell[i]), pos3[:, i], bsdf, None if normal3 is None else normal3[:, i])
        dist.append(np.abs(comps[:, 6] - u3[0, i]).min())
    print("  comp diffs", len(diff), "cdf distances", np.sort(np.asarray(dist))[-8:])
    same = v & (g["comp"] == r["comp"])
    ang = np.arccos(np.clip((g["dir"][:, same] * r["dir"][:, same]).sum(axis=0), -1, 1))
    print("  angle q", np.quantile(ang, [0.5, 0.99, 0.999, 1.0]))
    big = same & (r["pdf"] > 1e-6)
    rel = np.abs(g["pdf"][big] - r["pdf"][big]) / r["pdf"][big]
    print("  pdf rel q", np.quantile(rel, [0.5, 0.9, 0.99, 0.999, 1.0]))
    ij = np.abs(g["inv_jac"][same] - r["inv_jac"][same])
    print("  inv_jac abs max", ij.max())
    # pdf-only at fixed directions isolates the product parameters from the sampled direction
    rng = np.random.default_rng(3)
    d = rng.normal(size=(3, n)).astype(np.float32)
    d /= np.linalg.norm(d, axis=0, keepdims=True)
    d = np.ascontiguousarray(d)
    gp = cells.sample_pdf_product(cell, pos3, None, bsdf, normal3=normal3, dir3=d)
    rp, _ = ref.product_sample_pdf(cell, pos3, bsdf, dir3=d, normal3=normal3)
    big = rp["valid"].astype(bool) & (rp["pdf"] > 1e-6)
    rel = np.abs(gp["pdf"][big] - rp["pdf"][big]) / rp["pdf"][big]
    print("  pdf-only rel q", np.quantile(rel, [0.5, 0.9, 0.99, 0.999, 1.0]), big.sum())
    # the plain conditional for scale
    gc = cells.sample_pdf(cell, pos3, u3)
    rc, _ = ref.sample_pdf(cell, pos3, u3)
    print("  conditional comp diffs", ((gc["comp"] != rc["comp"]) & rc["valid"].astype(bool)).sum())
    # worst sample: component-level comparison needs the GPU's components -> compare total via pdf only
    if len(diff):
        i = diff[int(np.argmax(dist))]
        comps = ref.product_components(int(cell[i]), pos3[:, i], bsdf, None if normal3 is None else normal3[:, i])
        print("  worst", i, "u0", u3[0, i], "g", g["comp"][i], "r", r["comp"][i], "ref pmf around",
              comps[max(0, r["comp"][i] - 2):r["comp"][i] + 3, 0], "cdf", comps[max(0, r["comp"][i] - 2):r["comp"][i] + 3, 6])
