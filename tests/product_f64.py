"""Float64 restatement of the product-guiding pdf (sdmm_product.h:62-203 with offset_directional.h:126-170,
directional.h:89-95, cholesky.h +1e-5 pivots, sdmm.h:284-339) from float32 conditioner planes ([38][K] of one cell, as
sdmmb_download_conditioner / sdmm_ref_get_conditioner write them): the yardstick that tells float32 conditioning
from a formula difference.  Used by tests/test_gpu_product.py and tools/product_diag2.py."""
import numpy as np


def frame(n):
    n = n / np.linalg.norm(n)
    sign = np.copysign(1.0, n[2])
    a = -1.0 / (sign + n[2])
    b = n[0] * n[1] * a
    s = np.array([(n[0] * n[0] * a) * sign + 1.0, b * sign, n[0] * np.copysign(1.0, -n[2])])
    t = np.array([b, sign + n[1] * n[1] * a, -n[1]])
    return n, s, t


def product_pdf64(P, x, bsdf, normal, wo, dt=np.float64):
    P = P.astype(dt)
    K = P.shape[1]
    d = x.astype(dt)[:, None] - P[0:3]
    L = P[3:12].reshape(3, 3, K)
    z0 = d[0] / L[0, 0]
    z1 = (d[1] - L[1, 0] * z0) / L[1, 1]
    z2 = (d[2] - L[2, 0] * z0 - L[2, 1] * z1) / L[2, 2]
    w = P[13] * P[12] * (2 * np.pi) ** -1.5 * np.exp(-0.5 * (z0 * z0 + z1 * z1 + z2 * z2))
    pmf1 = w / w.sum()
    MT = P[14:20].reshape(2, 3, K)
    tm = np.einsum("rck,ck->rk", MT, d)
    R = P[20:29].reshape(3, 3, K)
    c00, c10, c11 = P[34], P[36], P[37]
    tot = 0.0
    acc = 0.0
    for m in range(bsdf.shape[0]):
        bw = float(bsdf[m, 0])
        if bw == 0:
            continue
        mean = normal if (m == 0 and normal is not None) else bsdf[m, 1:4]
        mean, s, t = frame(mean.astype(np.float64))
        b00, b10, b11 = [float(v) for v in bsdf[m, 4:7]]
        l = np.einsum("rck,c->rk", R, mean)
        lz = np.clip(l[2], -1, 1)
        ang = np.arccos(lz)
        sin2 = 1 - lz * lz
        sn = np.sqrt(np.maximum(sin2, 0))
        small = (sn < 1e-2) | (lz < -0.99)
        rcp = np.where(small, 1.0, ang / np.where(small, 1, sn))
        ij = np.where(lz <= -0.99, 0.0, rcp)
        temp = (lz * rcp - 1) * np.where(small, 0.0, 1 / np.where(small, 1, sin2))
        Jto = np.stack([rcp * R[0] + l[0] * temp * R[2], rcp * R[1] + l[1] * temp * R[2]])  # 2,3,K
        m0 = l[0] * rcp - tm[0]
        m1 = l[1] * rcp - tm[1]
        J = np.stack([np.einsum("rck,c->rk", Jto, s), np.einsum("rck,c->rk", Jto, t)], axis=1)  # 2(r),2(c),K
        B = np.array([[b00, b10], [b10, b11]])
        S = np.einsum("ack,cd,bdk->abk", J, B, J)
        q00, q10, q11 = c00 + S[0, 0], c10 + S[1, 0], c11 + S[1, 1]
        L00 = np.sqrt(q00 + 1e-5); L10 = q10 / L00; L11 = np.sqrt(q11 - L10 * L10 + 1e-5)
        i00, i11 = 1 / L00, 1 / L11
        i10 = -L10 * i00 * i11
        f00, f01 = i00 * c00, i00 * c10
        f10, f11 = i10 * c00 + i11 * c10, i10 * c10 + i11 * c11
        g00, g01 = i00 * S[0, 0], i00 * S[0, 1]
        g10, g11 = i10 * S[0, 0] + i11 * S[1, 0], i10 * S[0, 1] + i11 * S[1, 1]
        n00 = f00 * g00 + f10 * g10
        n10 = f01 * g00 + f11 * g10
        n11 = f01 * g01 + f11 * g11
        p0, p1 = i00 * m0, i10 * m0 + i11 * m1
        t0 = tm[0] + f00 * p0 + f10 * p1
        t1 = tm[1] + f01 * p0 + f11 * p1
        st0 = m0 / L00; st1 = (m1 - L10 * st0) / L11
        wk = pmf1 * bw / (L00 * L11) * ij / (2 * np.pi) * np.exp(-0.5 * (st0 * st0 + st1 * st1))
        N00 = np.sqrt(n00 + 1e-5); N10 = n10 / N00; N11 = np.sqrt(n11 - N10 * N10 + 1e-5)
        lw = np.einsum("rck,c->rk", R, wo.astype(np.float64))
        lzw = np.clip(lw[2], -1, 1)
        sa = np.sqrt(np.maximum(1 - lzw * lzw, 0))
        rs = np.where(sa < 1e-4, 1.0, np.arccos(lzw) / np.where(sa < 1e-4, 1, sa))
        ijw = np.where(lzw < -0.98, 0.0, rs)
        a0 = lw[0] * rs - t0; a1 = lw[1] * rs - t1
        y0 = a0 / N00; y1 = (a1 - N10 * y0) / N11
        gk = 1 / (N00 * N11) / (2 * np.pi) * np.exp(-0.5 * (y0 * y0 + y1 * y1))
        tot += wk.sum()
        acc += (wk * gk * ijw).sum()
    return acc / tot
