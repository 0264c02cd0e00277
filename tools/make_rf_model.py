#!/usr/bin/env python
"""Static instruction / register-read model of the shipped E-step kernels -> profiles/r2_k1/rf_model.json
(read by bench.py's roofline block).  Runs tools/sass_cost.py's model over the hot loop of the default
shape of every K (borrowed SoA planes, no log-likelihood output) and normalises to one warp pair-step
(= 64 sample-component updates: 32 lanes x one component pair x one sample)."""
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import sass_cost as sc  # noqa: E402

SHAPES = {16: (1, 8, 4, 3), 32: (1, 16, 4, 3), 64: (1, 32, 4, 3), 128: (2, 32, 4, 2)}


def main():
    obj = os.path.join(ROOT, "sdmm_mitsuba_b200", "lib", "estep_kernels.o")
    ks = sc.kernels(obj)
    per_K = {}
    for K, (NP, LPS, W, MINB) in SHAPES.items():
        pat = "estep_kernelILi%dELi%dELi%dELi%dELb0ELb1E" % (NP, LPS, W, MINB)
        (name, body), = [(n, b) for n, b in ks.items() if re.search(pat, n)]
        loop = sc.hot_loop(body, 30 * NP)
        c = sc.cost(loop)
        steps = 2 * NP   # the loop is unrolled twice; NP pairs per lane per sample
        ops = c["ops"]
        packed = sum(ops.get(o, 0) for o in sc.FMA2)
        scalar = sum(ops.get(o, 0) for o in sc.FMA1)
        per_K[str(K)] = dict(
            kernel="em_estep_kernel<%d,%d,%d,%d,LL=0,SOA=1>" % (NP, LPS, W, MINB),
            instructions=c["instructions"] / steps, packed_per_pair_step=packed / steps,
            scalar_fp32_per_pair_step=scalar / steps, fma_pipe_cycles=c["fma_pipe_cycles"] / steps,
            rf_reads_even_odd=[r / steps for r in c["rf_reads_even_odd"]], mufu_per_pair_step=ops.get("MUFU", 0) / steps,
            shfl_per_pair_step=ops.get("SHFL", 0) / steps)
    out = dict(source="tools/make_rf_model.py: tools/sass_cost.py over cuobjdump -sass of sdmm_mitsuba_b200/lib/estep_kernels.o "
                      "(hot loop of the default shape of each K, per warp pair-step = 64 updates)",
               sms=148, per_K=per_K)
    dst = os.path.join(ROOT, "profiles", "r2_k1", "rf_model.json")
    with open(dst, "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
