// query_kernels.cu -- K4: fused guided query (condition on position -> CDF select -> sample -> pdf)
// for one path vertex per thread.  Replaces, per vertex, the reference's
//   sdmm::create_conditional   R: distributions/sdmm_conditioner.h:102-115, 232-248
//   SDMM::sample (D = 2)       R: distributions/sdmm.h:209-273 (+ categorical.h, offset_directional.h:62-82)
//   SDMM::posterior + hsum     R: distributions/sdmm.h:331-339, offset_directional.h:39-60
// as called from SDMMRenderer::sampleSurface / pdfSurface (mitsuba/src/integrators/sdmm/sdmm_proc.cpp:
// 361-372, 419-421, 491-501, 539-545).
//
// Mapping: thread <-> vertex, sequential loop over the K components so that the conditional CDF is
// the same left-to-right float32 prefix sum the reference builds (categorical.h:92-111) -- the
// component chosen for a given uniform then only depends on the pmf bits.  The cell's query
// records (K x 128 B) are staged in shared memory when the vertices are pre-binned by cell
// (a CTA works on a run of vertices of ONE cell; all lanes read the same record -> broadcast LDS),
// or gathered through L2 otherwise.  The K unnormalised conditional weights of a vertex live in a
// per-thread shared-memory column (conflict-free) between the three passes.
#include "kernels.h"
#include "sdmm_math.cuh"

namespace sdmmb {

namespace {

constexpr int QTHREADS = 128;

__device__ __forceinline__ float q_rsqrt(float x) {
    float y;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float q_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// theta / sin(theta) from cos(theta); same construction as the E-step (em_kernels.cu).
__device__ __forceinline__ float rcp_sinc_from_cos(float c) {
    float a = fminf(fabsf(c), 0.99999994f);
    float P = 2.295893482e-03f;
    P = fmaf(P, a, -1.114577623e-02f);
    P = fmaf(P, a, 2.689914581e-02f);
    P = fmaf(P, a, -4.880191494e-02f);
    P = fmaf(P, a, 8.875543537e-02f);
    P = fmaf(P, a, -2.145852374e-01f);
    P = fmaf(P, a, 1.570796172e+00f);
    float u = q_rsqrt(1.f + a);
    float v = q_rsqrt(1.f - a);
    float num = (c >= 0.f) ? P : fmaf(kPi, v, -P);
    return num * u;
}

struct QRec {
    float4 v[8];
    __device__ __forceinline__ float f(int i) const { return reinterpret_cast<const float*>(v)[i]; }
};

template <bool SMEM>
__device__ __forceinline__ QRec load_rec(const float4* __restrict__ base, int k) {
    QRec r;
#pragma unroll
    for (int i = 0; i < 8; ++i) r.v[i] = SMEM ? base[k * 8 + i] : __ldg(base + k * 8 + i);
    return r;
}

// marginal posterior pi_k N3(x; mu_a, Sigma_aa) and tangent_mean = M (x - mu_a)
// R: sdmm_conditioner.h:242-244
__device__ __forceinline__ float marginal_weight(const float4 r0, const float4 r1, const float4 r2,
                                                 float x, float y, float z, float& d0, float& d1, float& d2) {
    // r0 = (mu0, mu1, mu2, wa00), r1 = (wa10, wa11, wa20, wa21), r2 = (wa22, nlca, m00, m01)
    d0 = x - r0.x; d1 = y - r0.y; d2 = z - r0.z;
    float z0 = r0.w * d0;
    float z1 = fmaf(r1.y, d1, r1.x * d0);
    float z2 = fmaf(r2.x, d2, fmaf(r1.w, d1, r1.z * d0));
    float e = fmaf(z0, z0, r2.y);
    e = fmaf(z1, z1, e);
    e = fmaf(z2, z2, e);
    return q_ex2(-e);
}

template <bool SMEM, bool SAMPLE>
__device__ __forceinline__ void query_one(const float4* __restrict__ rec, int K, float* __restrict__ wcol,
                                          int wstride, float x, float y, float z, float u0, float u1,
                                          float u2, float& wx, float& wy, float& wz, float& out_pdf,
                                          float& out_ij, uint32_t& out_comp, uint8_t& out_valid) {
    // ---- pass 1: unnormalised conditional weights, sequential sum -------------------------------
    float total = 0.f;
    for (int k = 0; k < K; ++k) {
        const float4 r0 = SMEM ? rec[k * 8 + 0] : __ldg(rec + k * 8 + 0);
        const float4 r1 = SMEM ? rec[k * 8 + 1] : __ldg(rec + k * 8 + 1);
        const float4 r2 = SMEM ? rec[k * 8 + 2] : __ldg(rec + k * 8 + 2);
        float d0, d1, d2;
        float w = marginal_weight(r0, r1, r2, x, y, z, d0, d1, d2);
        wcol[k * wstride] = w;
        total = (k == 0) ? w : total + w;
    }
    // Categorical::prepare: valid iff sum > 1e-20                    R: categorical.h:97-101
    if (!(total > 1e-20f)) {
        out_valid = 0; out_pdf = 0.f; out_ij = 0.f; out_comp = 0u;
        if (SAMPLE) { wx = wy = wz = 0.f; }
        return;
    }
    out_valid = 1;
    const float inv = 1.f / total;
    if (SAMPLE) {
        // ---- pass 2: first k with cdf[k] * inv >= u0 (== enoki::binary_search on the monotone
        // cdf, clamped to K-1), then back off over zero-pmf components   R: sdmm.h:220-233
        int sel = K - 1;
        float run = 0.f;
        bool found = false;
        for (int k = 0; k < K; ++k) {
            float w = wcol[k * wstride];
            run = (k == 0) ? w : run + w;
            if (!found && !(run * inv < u0)) { sel = k; found = true; }
        }
        while (sel > 0 && wcol[sel * wstride] * inv == 0.f) --sel;
        out_comp = (uint32_t)sel;
        // Box-Muller                                                 R: sdmm.h:199-207, 239-245
        float radius = sqrtf(-2.f * logf(1.f - u1));
        float theta = 2.f * kPi * u2;
        float sn, cs;
        sincosf(theta, &sn, &cs);
        float g0 = sn * radius, g1 = cs * radius;
        const QRec r = load_rec<SMEM>(rec, sel);
        float dx0 = x - r.f(QREC_MU + 0), dx1 = y - r.f(QREC_MU + 1), dx2 = z - r.f(QREC_MU + 2);
        float tm0 = r.f(QREC_M + 0) * dx0 + r.f(QREC_M + 1) * dx1 + r.f(QREC_M + 2) * dx2;
        float tm1 = r.f(QREC_M + 3) * dx0 + r.f(QREC_M + 4) * dx1 + r.f(QREC_M + 5) * dx2;
        // tangent = cov_sqrt * z (lower triangular)                  R: sdmm.h:247-262
        float t0 = r.f(QREC_LC + 0) * g0;
        float t1 = r.f(QREC_LC + 1) * g0 + r.f(QREC_LC + 2) * g1;
        // OffsetDirectionalTangentSpace::from                        R: offset_directional.h:62-82
        const float max_length = kPi - 1e-1f;
        float ox = t0 + tm0, oy = t1 + tm1;
        float length = sqrtf(ox * ox + oy * oy);
        float sl, cl;
        sincosf(length, &sl, &cl);
        float sinc = ((sl < 1e-2f) || length > max_length) ? 1.f : sl / length;
        out_ij = (length > max_length) ? 0.f : sinc;
        float l0 = ox * sinc, l1 = oy * sinc, l2 = cl;
        // from = transpose(to)
        wx = r.f(QREC_R + 0) * l0 + r.f(QREC_R + 3) * l1 + r.f(QREC_R + 6) * l2;
        wy = r.f(QREC_R + 1) * l0 + r.f(QREC_R + 4) * l1 + r.f(QREC_R + 7) * l2;
        wz = r.f(QREC_R + 2) * l0 + r.f(QREC_R + 5) * l1 + r.f(QREC_R + 8) * l2;
    } else {
        out_ij = 0.f;
        out_comp = 0u;
    }
    // ---- pass 3: conditional pdf at omega, summed over components     R: sdmm.h:331-339, 284-294
    float pdf = 0.f;
    for (int k = 0; k < K; ++k) {
        const QRec r = load_rec<SMEM>(rec, k);
        float dx0 = x - r.f(QREC_MU + 0), dx1 = y - r.f(QREC_MU + 1), dx2 = z - r.f(QREC_MU + 2);
        float tm0 = fmaf(r.f(QREC_M + 2), dx2, fmaf(r.f(QREC_M + 1), dx1, r.f(QREC_M + 0) * dx0));
        float tm1 = fmaf(r.f(QREC_M + 5), dx2, fmaf(r.f(QREC_M + 4), dx1, r.f(QREC_M + 3) * dx0));
        // OffsetDirectionalTangentSpace::to                          R: offset_directional.h:39-60
        float lx = fmaf(r.f(QREC_R + 2), wz, fmaf(r.f(QREC_R + 1), wy, r.f(QREC_R + 0) * wx));
        float ly = fmaf(r.f(QREC_R + 5), wz, fmaf(r.f(QREC_R + 4), wy, r.f(QREC_R + 3) * wx));
        float lz = fmaf(r.f(QREC_R + 8), wz, fmaf(r.f(QREC_R + 7), wy, r.f(QREC_R + 6) * wx));
        float rs = rcp_sinc_from_cos(lz);
        float ij = (lz < -0.98f) ? 0.f : rs;
        float t0 = fmaf(lx, rs, -tm0), t1 = fmaf(ly, rs, -tm1);
        float z0 = r.f(QREC_WC + 0) * t0;
        float z1 = fmaf(r.f(QREC_WC + 2), t1, r.f(QREC_WC + 1) * t0);
        float e = fmaf(z0, z0, r.f(QREC_NLCC));
        e = fmaf(z1, z1, e);
        float pk = ij * q_ex2(-e);
        pdf = fmaf(pk, wcol[k * wstride] * inv, pdf);
    }
    out_pdf = pdf;
}

// Binned path: vertices sorted by cell.  A CTA owns a contiguous tile of QTILE vertices; it walks
// the cell runs inside the tile, staging each cell's records in shared memory once.
template <bool SAMPLE>
__global__ void __launch_bounds__(QTHREADS)
query_binned_kernel(QueryLaunch q, int tile) {
    extern __shared__ __align__(16) unsigned char qsmem[];
    const int K = q.K;
    float4* srec = reinterpret_cast<float4*>(qsmem);                                   // [K][8]
    float* wcols = reinterpret_cast<float*>(qsmem + (size_t)K * QREC_FLOATS * 4);       // [K][QTHREADS]
    __shared__ uint32_t s_cell;
    const size_t tbegin = (size_t)blockIdx.x * tile;
    const size_t tend = (tbegin + tile < q.n) ? tbegin + tile : q.n;
    size_t pos = tbegin;
    while (pos < tend) {
        __syncthreads();
        if (threadIdx.x == 0) s_cell = q.cell_idx[pos];
        __syncthreads();
        const uint32_t cell = s_cell;
        const bool cell_ok = cell < (uint32_t)q.max_cells;
        if (cell_ok) {
            const float4* g = reinterpret_cast<const float4*>(q.qrec + (size_t)cell * K * QREC_FLOATS);
            for (int i = threadIdx.x; i < K * 8; i += QTHREADS) srec[i] = __ldg(g + i);
        }
        __syncthreads();
        // process QTHREADS vertices at a time while they belong to `cell`
        bool run_over = false;
        while (!run_over && pos < tend) {
            const size_t i = pos + threadIdx.x;
            const bool mine = (i < tend) && (q.cell_idx[i] == cell);
            if (mine) {
                float wx = 0.f, wy = 0.f, wz = 0.f, pdf = 0.f, ij = 0.f;
                uint32_t comp = 0;
                uint8_t valid = 0;
                if (cell_ok) {
                    const float x = q.pos3[i], y = q.pos3[q.n + i], z = q.pos3[2 * q.n + i];
                    if (SAMPLE) {
                        query_one<true, true>(srec, K, wcols + threadIdx.x, QTHREADS, x, y, z, q.u3[i],
                                              q.u3[q.n + i], q.u3[2 * q.n + i], wx, wy, wz, pdf, ij, comp, valid);
                    } else {
                        wx = q.dir3[i]; wy = q.dir3[q.n + i]; wz = q.dir3[2 * q.n + i];
                        query_one<true, false>(srec, K, wcols + threadIdx.x, QTHREADS, x, y, z, 0.f, 0.f, 0.f,
                                               wx, wy, wz, pdf, ij, comp, valid);
                    }
                }
                if (SAMPLE) {
                    q.out_dir3[i] = wx; q.out_dir3[q.n + i] = wy; q.out_dir3[2 * q.n + i] = wz;
                    q.out_inv_jac[i] = ij;
                    q.out_comp[i] = comp;
                }
                q.out_pdf[i] = pdf;
                if (q.out_valid) q.out_valid[i] = valid;
            }
            // the run ends inside this batch iff some thread saw another cell (or the tile ended)
            const int n_mine = __syncthreads_count(mine ? 1 : 0);
            pos += n_mine;
            run_over = (n_mine < QTHREADS);
        }
    }
}

// Gather path: arbitrary vertex order; records come through L2 (one 128-byte line per component).
template <bool SAMPLE>
__global__ void __launch_bounds__(QTHREADS)
query_gather_kernel(QueryLaunch q) {
    extern __shared__ __align__(16) unsigned char qsmem[];
    float* wcols = reinterpret_cast<float*>(qsmem);  // [K][QTHREADS]
    const int K = q.K;
    const size_t i = (size_t)blockIdx.x * QTHREADS + threadIdx.x;
    if (i >= q.n) return;
    const uint32_t cell = q.cell_idx[i];
    float wx = 0.f, wy = 0.f, wz = 0.f, pdf = 0.f, ij = 0.f;
    uint32_t comp = 0;
    uint8_t valid = 0;
    if (cell < (uint32_t)q.max_cells) {
        const float4* rec = reinterpret_cast<const float4*>(q.qrec + (size_t)cell * K * QREC_FLOATS);
        const float x = q.pos3[i], y = q.pos3[q.n + i], z = q.pos3[2 * q.n + i];
        if (SAMPLE) {
            query_one<false, true>(rec, K, wcols + threadIdx.x, QTHREADS, x, y, z, q.u3[i], q.u3[q.n + i],
                                   q.u3[2 * q.n + i], wx, wy, wz, pdf, ij, comp, valid);
        } else {
            wx = q.dir3[i]; wy = q.dir3[q.n + i]; wz = q.dir3[2 * q.n + i];
            query_one<false, false>(rec, K, wcols + threadIdx.x, QTHREADS, x, y, z, 0.f, 0.f, 0.f, wx, wy, wz,
                                    pdf, ij, comp, valid);
        }
    }
    if (SAMPLE) {
        q.out_dir3[i] = wx; q.out_dir3[q.n + i] = wy; q.out_dir3[2 * q.n + i] = wz;
        q.out_inv_jac[i] = ij;
        q.out_comp[i] = comp;
    }
    q.out_pdf[i] = pdf;
    if (q.out_valid) q.out_valid[i] = valid;
}

template <typename Kern>
cudaError_t set_smem(Kern kern, size_t smem) {
    return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
}

}  // namespace

cudaError_t launch_query(const QueryLaunch& q, cudaStream_t st) {
    if (q.n == 0) return cudaSuccess;
    const bool sample = q.u3 != nullptr;
    const size_t wcol_bytes = (size_t)q.K * QTHREADS * sizeof(float);
    cudaError_t e;
    if (q.binned) {
        const size_t smem = (size_t)q.K * QREC_FLOATS * 4 + wcol_bytes;
        // tile: multiple of QTHREADS; ~8 tiles per SM-slot keeps the tail short without making
        // the per-tile record staging (K*128 B) significant
        int tile = 1024;
        const size_t grid = (q.n + tile - 1) / tile;
        if (sample) {
            if ((e = set_smem(query_binned_kernel<true>, smem)) != cudaSuccess) return e;
            query_binned_kernel<true><<<(unsigned)grid, QTHREADS, smem, st>>>(q, tile);
        } else {
            if ((e = set_smem(query_binned_kernel<false>, smem)) != cudaSuccess) return e;
            query_binned_kernel<false><<<(unsigned)grid, QTHREADS, smem, st>>>(q, tile);
        }
    } else {
        const size_t grid = (q.n + QTHREADS - 1) / QTHREADS;
        if (sample) {
            if ((e = set_smem(query_gather_kernel<true>, wcol_bytes)) != cudaSuccess) return e;
            query_gather_kernel<true><<<(unsigned)grid, QTHREADS, wcol_bytes, st>>>(q);
        } else {
            if ((e = set_smem(query_gather_kernel<false>, wcol_bytes)) != cudaSuccess) return e;
            query_gather_kernel<false><<<(unsigned)grid, QTHREADS, wcol_bytes, st>>>(q);
        }
    }
    return cudaGetLastError();
}

}  // namespace sdmmb
