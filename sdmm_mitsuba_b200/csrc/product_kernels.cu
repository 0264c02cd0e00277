// product_kernels.cu -- product guiding on the device: the guided query with
// the conditional mixture p(omega | x) multiplied by a learned-BSDF directional mixture.
//
// Reference ("R:" = /root/reference/sdmm/include/sdmm):
//   sdmm::product(conditional, learned_bsdf, product)     R: distributions/sdmm_product.h:62-203
//   to_centered_tangent_space (offset variant)            R: distributions/sdmm_product.h:42-60
//   OffsetDirectionalTangentSpace::to_jacobian            R: spaces/offset_directional.h:126-170
//   DirectionalTangentSpace::from_center_jacobian         R: spaces/directional.h:89-95
//   prepare_vectorized(result) + SDMM::sample + posterior R: distributions/sdmm.h:163-173, 209-273, 331-339
// called per vertex from sampleSurface when sampleProduct is on (mitsuba/src/integrators/sdmm/sdmm_proc.cpp:
// 327-356, 375-381, 419-421, 476-487).  The product of K conditional and M BSDF components has K*M components
// whose parameters exist only for the duration of one vertex, so they are never stored: one thread owns one
// vertex and RECOMPUTES the K*M components in each of its passes (total weight, component select, pdf at the
// sampled direction) -- ~150 flop per component pair against 9 floats of state each.
// Component order = the reference's packet layout (product_i = first_packet * M_nz + m, 8 lanes per packet,
// sdmm_product.h:101-108): the sequential float32 CDF, hence the selected component, follows it.
// Vertices are first sorted by cell on the device (the counting sort of the plain query) so that a warp's record
// loads are uniform.  First version: scalar IEEE arithmetic, not tuned further.
#include "kernels.h"
#include "sdmm_math.cuh"

namespace sdmmb {

namespace {

constexpr float kInv2Pi = 0.15915494309189533577f;   // gaussian_normalization<2>   R: core/constants.h:45-47
constexpr float kPiF = 3.14159265358979323846f;

struct CondK {      // one component of the conditional mixture at x (create_conditional_vectorized)
    float w;        // unnormalised pmf
    float tm0, tm1; // tangent_mean
    float R[9];     // frame.to, row-major
    float c00, c10, c11;   // conditional covariance (lower triangle)
};

struct BsdfM {      // one component of the learned-BSDF mixture in world space
    float w;
    float mean[3];
    float s[3], t[3];      // first two columns of frame.from
    float c00, c10, c11;
};

struct ProdC {      // one prepared product component
    float w;        // unnormalised weight
    float tm0, tm1;
    float L00, L10, L11;
    float r0, r1;   // 1 / L00, 1 / L11
    float inv_det;
    bool psd;
};

__device__ __forceinline__ void load_cond(const float* __restrict__ rec, float x, float y, float z, CondK& c) {
    const float d0 = x - rec[QREC_MU + 0], d1 = y - rec[QREC_MU + 1], d2 = z - rec[QREC_MU + 2];
    const float z0 = rec[QREC_WA + 0] * d0;
    const float z1 = fmaf(rec[QREC_WA + 2], d1, rec[QREC_WA + 1] * d0);
    const float z2 = fmaf(rec[QREC_WA + 5], d2, fmaf(rec[QREC_WA + 4], d1, rec[QREC_WA + 3] * d0));
    const float e = fmaf(z2, z2, fmaf(z1, z1, fmaf(z0, z0, rec[QREC_NLCA])));
    c.w = exp2f(-e);
    c.tm0 = rec[QREC_M + 0] * d0 + rec[QREC_M + 1] * d1 + rec[QREC_M + 2] * d2;
    c.tm1 = rec[QREC_M + 3] * d0 + rec[QREC_M + 4] * d1 + rec[QREC_M + 5] * d2;
#pragma unroll
    for (int i = 0; i < 9; ++i) c.R[i] = rec[QREC_R + i];
    // the record keeps cov_sqrt; undo linalg::cholesky's +1e-5 pivots (R: linalg/cholesky.h:20-35)
    const float L00 = rec[QREC_LC + 0], L10 = rec[QREC_LC + 1], L11 = rec[QREC_LC + 2];
    c.c00 = L00 * L00 - 1e-5f;
    c.c10 = L10 * L00;
    c.c11 = L11 * L11 + L10 * L10 - 1e-5f;
}

// R: sdmm_product.h:110-199 for one (first, second) pair, then prepare_cov of the result.
// Every 1/sqrt(x) pair of the reference (Cholesky pivot, then division by it) is one rsqrt here and the divisions
// become multiplications by it: the kernel is issue-bound on exactly these chains (ncu: 34 % XU pipe, 69 % issue
// slots busy at 25 % occupancy), and 2 ulp on a pivot is far below the conditioning of the weights.
__device__ __forceinline__ ProdC product_component(const CondK& c, float pmf1, const BsdfM& b) {
    ProdC o;
    // first_ts.to_jacobian(second mean)                                  R: offset_directional.h:126-170
    const float lx = c.R[0] * b.mean[0] + c.R[1] * b.mean[1] + c.R[2] * b.mean[2];
    const float ly = c.R[3] * b.mean[0] + c.R[4] * b.mean[1] + c.R[5] * b.mean[2];
    const float lz = c.R[6] * b.mean[0] + c.R[7] * b.mean[1] + c.R[8] * b.mean[2];
    const float angle = acosf(fminf(1.f, fmaxf(-1.f, lz)));
    const float sin2 = 1.f - lz * lz;
    const bool small = (sin2 < 1e-4f) || (lz < -0.99f);     // sin < 1e-2
    const float rsin = rsqrtf(small ? 1.f : sin2);
    const float rcp = small ? 1.f : angle * rsin;
    const float inv_jac = (lz <= -0.99f) ? 0.f : rcp;
    const float temp = small ? 0.f : (lz * rcp - 1.f) * (rsin * rsin);
    const float jl02 = lx * temp, jl12 = ly * temp;
    float Jto[6];   // (Jl * to): 2x3
#pragma unroll
    for (int col = 0; col < 3; ++col) {
        Jto[col] = rcp * c.R[col] + jl02 * c.R[6 + col];
        Jto[3 + col] = rcp * c.R[3 + col] + jl12 * c.R[6 + col];
    }
    const float m0 = lx * rcp - c.tm0, m1 = ly * rcp - c.tm1;   // second mean in the first's tangent space
    // J = to_jacobian * from_center_jacobian (2x2)                        R: sdmm_product.h:121-122
    const float J00 = Jto[0] * b.s[0] + Jto[1] * b.s[1] + Jto[2] * b.s[2];
    const float J01 = Jto[0] * b.t[0] + Jto[1] * b.t[1] + Jto[2] * b.t[2];
    const float J10 = Jto[3] * b.s[0] + Jto[4] * b.s[1] + Jto[5] * b.s[2];
    const float J11 = Jto[3] * b.t[0] + Jto[4] * b.t[1] + Jto[5] * b.t[2];
    // second_cov = J cov2 J^T
    const float a00 = J00 * b.c00 + J01 * b.c10, a01 = J00 * b.c10 + J01 * b.c11;
    const float a10 = J10 * b.c00 + J11 * b.c10, a11 = J10 * b.c10 + J11 * b.c11;
    const float s00 = a00 * J00 + a01 * J01, s01 = a00 * J10 + a01 * J11;
    const float s10 = a10 * J00 + a11 * J01, s11 = a10 * J10 + a11 * J11;
    // cov_sum = first + second; Cholesky with the +1e-5 pivots; inverse
    const float q00 = c.c00 + s00, q10 = c.c10 + s10, q11 = c.c11 + s11;
    const float x0 = q00 + 1e-5f;
    const float i00 = rsqrtf(x0);                  // 1 / L00
    const float L10 = q10 * i00;
    const float x1 = q11 - L10 * L10 + 1e-5f;
    const float i11 = rsqrtf(x1);                  // 1 / L11
    const float i10 = -L10 * i00 * i11;
    // premultiplied covariances (inv_cov_sum_sqrt * cov), 2x2 each; first is symmetric
    const float f00 = i00 * c.c00, f01 = i00 * c.c10;
    const float f10 = i10 * c.c00 + i11 * c.c10, f11 = i10 * c.c10 + i11 * c.c11;
    const float g00 = i00 * s00, g01 = i00 * s01;
    const float g10 = i10 * s00 + i11 * s10, g11 = i10 * s01 + i11 * s11;
    // new_cov_tangent = first_premult^T * second_premult; only (0,0), (1,0), (1,1) reach the Cholesky
    const float n00 = f00 * g00 + f10 * g10;
    const float n10 = f01 * g00 + f11 * g10;
    const float n11 = f01 * g01 + f11 * g11;
    // new_mean_tangent = first_premult^T * (inv_cov_sum_sqrt * second_mean)   (first mean == 0)
    const float p0 = i00 * m0, p1 = i10 * m0 + i11 * m1;
    o.tm0 = c.tm0 + (f00 * p0 + f10 * p1);                                 // R: sdmm_product.h:57-58
    o.tm1 = c.tm1 + (f01 * p0 + f11 * p1);
    // weight                                                               R: sdmm_product.h:180-198
    const float st0 = p0, st1 = (m1 - L10 * st0) * i11;                    // solve(L, second_mean)
    const float sq = st0 * st0 + st1 * st1;
    const float wgt = pmf1 * b.w * (i00 * i11) * inv_jac * kInv2Pi * expf(-0.5f * sq);
    o.w = (x0 == 0.f || x1 == 0.f) ? 0.f : wgt;                            // cov_det == 0
    // prepare_cov of the product component                                 R: distributions/sdmm.h:186-197
    const float v0 = n00 + 1e-5f;
    o.r0 = rsqrtf(v0);
    o.L00 = v0 * o.r0;
    o.L10 = n10 * o.r0;
    const float v1 = n11 - o.L10 * o.L10 + 1e-5f;
    o.r1 = rsqrtf(v1);
    o.L11 = v1 * o.r1;
    o.psd = (v0 > 0.f) && (v1 > 0.f);
    o.inv_det = o.r0 * o.r1;
    return o;
}

// world-space BSDF component from (pmf, mean, cov): frame from the normalised mean (DirectionalTangentSpace::set_mean,
// R: spaces/directional.h:197-200)
__device__ __forceinline__ void make_bsdf(float w, const float mean_in[3], float c00, float c10, float c11, BsdfM& b) {
    const float inv = 1.f / sqrtf(mean_in[0] * mean_in[0] + mean_in[1] * mean_in[1] + mean_in[2] * mean_in[2]);
    const float n[3] = {mean_in[0] * inv, mean_in[1] * inv, mean_in[2] * inv};
    float to[9], from[9];
    frame_from_normal(n, to, from);
    b.w = w;
#pragma unroll
    for (int r = 0; r < 3; ++r) { b.mean[r] = n[r]; b.s[r] = from[r * 3 + 0]; b.t[r] = from[r * 3 + 1]; }
    b.c00 = c00; b.c10 = c10; b.c11 = c11;
}

// glossy surfaces: the BSDF mixture lives in the local shading frame relative to the incident direction; its first
// packet (components 0..7) is rotated about the local z axis by -phi(wi) and taken to world space through the
// shading frame: mean, s, t -> W R^T (mean, s, t)            (rotate_to_wo + to_world_space, sdmm_proc.cpp:340-355,
// R: spaces/directional.h:202-218)
__device__ __forceinline__ void xform_bsdf(const float T[9], BsdfM& b) {
    float m[3], s[3], t[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        m[r] = T[3 * r] * b.mean[0] + T[3 * r + 1] * b.mean[1] + T[3 * r + 2] * b.mean[2];
        s[r] = T[3 * r] * b.s[0] + T[3 * r + 1] * b.s[1] + T[3 * r + 2] * b.s[2];
        t[r] = T[3 * r] * b.t[0] + T[3 * r + 1] * b.t[1] + T[3 * r + 2] * b.t[2];
    }
#pragma unroll
    for (int r = 0; r < 3; ++r) { b.mean[r] = m[r]; b.s[r] = s[r]; b.t[r] = t[r]; }
}

constexpr int PROD_THREADS = 128;
constexpr int PROD_MAX_GROUPS = 48;   // (K / 8) * non-zero BSDF components whose running sums fit in shared memory

// One thread per vertex.  Passes: (1) conditional total, (2) product total + PSD mask, (3) component select and
// sample, (4) product pdf at omega.  Every pass recomputes its components from the conditioner records.
template <bool SAMPLE>
__global__ void __launch_bounds__(PROD_THREADS, 6) product_query_kernel(ProductLaunch q, const float4* __restrict__ qin) {
    __shared__ BsdfM s_b[SDMMB_PRODUCT_MAX_BSDF];
    __shared__ float s_pref[SAMPLE ? PROD_MAX_GROUPS * PROD_THREADS : 1];   // running sum at the end of each group
    if ((int)threadIdx.x < q.M) {
        const float* src = q.bsdf + threadIdx.x * 7;   // pmf, mean3, c00, c10, c11
        make_bsdf(src[0], src + 1, src[4], src[5], src[6], s_b[threadIdx.x]);
    }
    __syncthreads();
    // vertices arrive sorted by cell (launch_bin_vertices): a warp reads the same records, so the record loads
    // below are warp-uniform (one transaction each) instead of 32 scattered ones
    const size_t p = (size_t)blockIdx.x * PROD_THREADS + threadIdx.x;
    if (p >= q.n) return;
    const size_t n = q.n;
    const float4 ra = __ldg(qin + 2 * p), rb = __ldg(qin + 2 * p + 1);
    const size_t i = __float_as_uint(rb.z);
    const uint32_t cell = __float_as_uint(rb.w);
    const float x = ra.x, y = ra.y, z = ra.z;
    float wo[3] = {0.f, 0.f, 0.f};
    float out_pdf = 0.f, out_ij = 0.f;
    uint32_t out_comp = 0u;
    bool valid = cell < (uint32_t)q.max_cells;
    const float* rec = q.qrec + (size_t)(valid ? cell : 0u) * q.K * QREC_FLOATS;
    const int K = q.K_user, M = q.M;

    // diffuse surfaces: component 0 is re-centred on the shading normal       (sdmm_proc.cpp:335-339)
    BsdfM b0 = s_b[0];
    if (q.normal3 != nullptr && !(q.frame9 != nullptr && q.wi3 != nullptr)) {
        const float nn[3] = {q.normal3[i], q.normal3[n + i], q.normal3[2 * n + i]};
        make_bsdf(b0.w, nn, b0.c00, b0.c10, b0.c11, b0);
    }

    // glossy surfaces: T = W R^T from the shading frame (columns s, t, n) and the local incident direction
    const bool glossy = (q.frame9 != nullptr) && (q.wi3 != nullptr);
    float T[9];
    if (glossy) {
        const float neg_phi = -atan2f(q.wi3[n + i], q.wi3[i]);
        float sn, cs;
        sincosf(neg_phi, &sn, &cs);
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const float w0 = q.frame9[(3 * r) * n + i], w1 = q.frame9[(3 * r + 1) * n + i], w2 = q.frame9[(3 * r + 2) * n + i];
            // R^T = [[c, s, 0], [-s, c, 0], [0, 0, 1]]
            T[3 * r] = w0 * cs - w1 * sn;
            T[3 * r + 1] = w0 * sn + w1 * cs;
            T[3 * r + 2] = w2;
        }
    }
    auto fetch = [&](int m) -> BsdfM {
        BsdfM b = (m == 0) ? b0 : s_b[m];
        if (glossy && m < 8) xform_bsdf(T, b);
        return b;
    };

    // ---- pass 1: the conditional's normalisation                               R: categorical.h:97-101
    float total1 = 0.f;
    CondK c;
    for (int k = 0; k < K; ++k) {
        load_cond(rec + k * QREC_FLOATS, x, y, z, c);
        total1 = (k == 0) ? c.w : total1 + c.w;
    }
    valid = valid && (total1 > 1e-20f);
    const float inv1 = 1.f / total1;

    // ---- pass 2: product total (sequential, in the reference's packet order) and the PSD mask.  A "group" is one
    // (packet of 8 conditional components) x (one non-zero BSDF component); the running sum at the end of every
    // group is kept so that pass 3 only has to redo the group the CDF crosses u0 in.
    int M_nz = 0;
    for (int m = 0; m < M; ++m) M_nz += (s_b[m].w != 0.f) ? 1 : 0;
    const int G = (K / 8) * M_nz;
    const bool use_pref = SAMPLE && G <= PROD_MAX_GROUPS;
    float* pref = s_pref + threadIdx.x;
    float total = 0.f;
    bool all_psd = true, first = true;
    {
        int gi = 0;
        for (int kp = 0; kp < K; kp += 8)
            for (int m = 0; m < M; ++m) {
                if (s_b[m].w == 0.f) continue;
                const BsdfM b = fetch(m);
                for (int k = kp; k < kp + 8; ++k) {
                    load_cond(rec + k * QREC_FLOATS, x, y, z, c);
                    const ProdC pc = product_component(c, c.w * inv1, b);
                    total = first ? pc.w : total + pc.w;
                    first = false;
                    all_psd = all_psd && pc.psd;
                }
                if (use_pref) pref[gi * PROD_THREADS] = total;
                ++gi;
            }
    }
    valid = valid && !first && (total > 1e-20f) && all_psd;   // R: sdmm.h:163-173 prepare_vectorized
    const float inv = 1.f / total;

    if (valid) {
        if (SAMPLE) {
            // ---- pass 3: first component with cdf >= u0, backing off over zero-pmf ones  R: sdmm.h:220-233
            const float u0 = ra.w, u1 = rb.x, u2 = rb.y;
            ProdC sel, best;
            int sel_k = 0, best_k = 0;
            uint32_t sel_i = 0, best_i = 0;
            // the group the crossing lies in: the same comparisons on the same floats as the full scan
            int g_start = 0;
            if (use_pref)
                while (g_start < G - 1 && pref[g_start * PROD_THREADS] * inv < u0) ++g_start;
            for (int attempt = 0; attempt < 2; ++attempt) {
                float run = (g_start > 0) ? pref[(g_start - 1) * PROD_THREADS] : 0.f;
                bool found = false, have_best = false, nz_hit = false;
                uint32_t idx = (uint32_t)g_start * 8u;
                int gi = 0;
                for (int kp = 0; kp < K && !found; kp += 8)
                    for (int m = 0; m < M && !found; ++m) {
                        if (s_b[m].w == 0.f) continue;
                        if (gi++ < g_start) continue;
                        const BsdfM b = fetch(m);
                        for (int k = kp; k < kp + 8; ++k, ++idx) {
                            load_cond(rec + k * QREC_FLOATS, x, y, z, c);
                            const ProdC pc = product_component(c, c.w * inv1, b);
                            run = (idx == 0) ? pc.w : run + pc.w;
                            const bool nz = pc.w * inv != 0.f;
                            if (nz || idx == 0) {   // index 0 is where the back-off stops even at pmf 0
                                best = pc; best_k = k; best_i = idx; have_best = true;
                            }
                            if (!(run * inv < u0)) {
                                found = true;
                                nz_hit = nz;
                                if (nz) { sel = pc; sel_k = k; sel_i = idx; }
                                break;
                            }
                        }
                    }
                // crossing on a zero-pmf component, or no crossing at all (binary_search clamps to the last index):
                // back off to the last non-zero one
                if (!(found && nz_hit)) {
                    if (!have_best && g_start > 0) { g_start = 0; continue; }   // it lies before this group: full scan
                    sel = best; sel_k = best_k; sel_i = best_i;
                }
                break;
            }
            out_comp = sel_i;
            // Box-Muller + offset exp map about the conditional component's frame   R: sdmm.h:234-273
            const float radius = sqrtf(-2.f * logf(1.f - u1));
            float sn, cs;
            sincosf(2.f * kPiF * u2, &sn, &cs);
            const float g0 = sn * radius, g1 = cs * radius;
            const float ox = sel.L00 * g0 + sel.tm0, oy = sel.L10 * g0 + sel.L11 * g1 + sel.tm1;
            const float max_length = kPiF - 1e-1f;
            const float length = sqrtf(ox * ox + oy * oy);
            float sl, cl;
            sincosf(length, &sl, &cl);
            const float sinc = ((sl < 1e-2f) || length > max_length) ? 1.f : sl / length;
            out_ij = (length > max_length) ? 0.f : sinc;
            const float l0 = ox * sinc, l1 = oy * sinc, l2 = cl;
            const float* R = rec + sel_k * QREC_FLOATS + QREC_R;
            wo[0] = R[0] * l0 + R[3] * l1 + R[6] * l2;
            wo[1] = R[1] * l0 + R[4] * l1 + R[7] * l2;
            wo[2] = R[2] * l0 + R[5] * l1 + R[8] * l2;
        } else {
            wo[0] = ra.w; wo[1] = rb.x; wo[2] = rb.y;
        }
        // ---- pass 4: posterior-weighted pdf at omega                                R: sdmm.h:331-339
        float pdf = 0.f;
        for (int kp = 0; kp < K; kp += 8)
            for (int m = 0; m < M; ++m) {
                const BsdfM b = fetch(m);
                if (b.w == 0.f) continue;
                for (int k = kp; k < kp + 8; ++k) {
                    load_cond(rec + k * QREC_FLOATS, x, y, z, c);
                    const ProdC pc = product_component(c, c.w * inv1, b);
                    const float lx = c.R[0] * wo[0] + c.R[1] * wo[1] + c.R[2] * wo[2];
                    const float ly = c.R[3] * wo[0] + c.R[4] * wo[1] + c.R[5] * wo[2];
                    const float lz = c.R[6] * wo[0] + c.R[7] * wo[1] + c.R[8] * wo[2];
                    const float ang = acosf(fminf(1.f, fmaxf(-1.f, lz)));
                    const float s2 = 1.f - lz * lz;
                    const bool tiny = !(s2 >= 1e-8f);                        // sin < 1e-4
                    const float rs = tiny ? 1.f : ang * rsqrtf(s2);          // R: offset_directional.h:40-60
                    const float ij = (lz < -0.98f) ? 0.f : rs;
                    const float t0 = lx * rs - pc.tm0, t1 = ly * rs - pc.tm1;
                    const float z0 = t0 * pc.r0, z1 = (t1 - pc.L10 * z0) * pc.r1;
                    const float g = pc.inv_det * kInv2Pi * expf(-0.5f * (z0 * z0 + z1 * z1));
                    pdf = fmaf(ij * g, pc.w * inv, pdf);
                }
            }
        out_pdf = pdf;
    }
    if (SAMPLE) {
        q.out_dir3[i] = wo[0]; q.out_dir3[n + i] = wo[1]; q.out_dir3[2 * n + i] = wo[2];
        if (q.out_inv_jac) q.out_inv_jac[i] = out_ij;
        if (q.out_comp) q.out_comp[i] = out_comp;
    }
    q.out_pdf[i] = out_pdf;
    if (q.out_valid) q.out_valid[i] = valid ? 1 : 0;
}

}  // namespace

cudaError_t launch_product_query(const ProductLaunch& q, cudaStream_t st) {
    if (q.n == 0) return cudaSuccess;
    if (q.M < 1 || q.M > SDMMB_PRODUCT_MAX_BSDF || q.K_user < 8 || (q.K_user % 8) != 0) return cudaErrorInvalidValue;
    float4* qin = nullptr;
    uint32_t* invperm = nullptr;
    cudaError_t e = launch_bin_vertices(q.cell_idx, q.n, q.max_cells, q.pos3, q.u3 ? q.u3 : q.dir3, q.scratch, &qin,
                                        &invperm, st);
    if (e != cudaSuccess) return e;
    const unsigned blocks = (unsigned)((q.n + PROD_THREADS - 1) / PROD_THREADS);
    if (q.u3 != nullptr) product_query_kernel<true><<<blocks, PROD_THREADS, 0, st>>>(q, qin);
    else product_query_kernel<false><<<blocks, PROD_THREADS, 0, st>>>(q, qin);
    return cudaGetLastError();
}

}  // namespace sdmmb
