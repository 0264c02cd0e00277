// em_kernels.cu -- the training half of the SDMM hot path on sm_100a:
//   K0 pack_samples_kernel     SoA sample planes -> packed 32-byte records
//   K1 em_estep_kernel<NP,LPS> E-step + sufficient statistics (the N*K hot loop)
//   K2 em_mstep_kernel         stats reduction over items, stepwise blend, normalise, update_model,
//                              prepare (Cholesky, records, CDF) and -- fused in its tail --
//   K3                         the conditioner prepare (Schur complement -> query records)
//   prepare_cells_kernel       K2's tail alone, for freshly uploaded models
//
// Reference for every formula: /root/reference/sdmm/include/sdmm ("R:" below), see DESIGN.md.
//
// K1 design (B200): the reference runs SIMD over components and sequentially over samples
// (R: opt/em.h:265-312).  K1 keeps that orientation because it makes the 21 sufficient statistics
// per component a pure register accumulation: a lane owns NP *pairs* of components, the two
// components of a pair ride in the two halves of Blackwell's packed FP32 instructions
// (fma.rn.f32x2 / FFMA2), and a group of LPS lanes covers all K = 2*NP*LPS components of one
// sample.  The only cross-lane traffic per sample is the log2(LPS)-step shuffle butterfly for the
// mixture normaliser sum_k pi_k p_k(x_n).  Samples are staged per warp through shared memory in a
// "duplicated pair" layout so that every broadcast operand is directly a 64-bit register pair.
#include "kernels.h"
#include "sdmm_math.cuh"

namespace sdmmb {

// ------------------------------------------------------------------------------------------------
// packed FP32x2 helpers (sm_100a: FFMA2 / FMUL2 / FADD2)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float2 mul2(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ float2 add2(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 splat(float v) { return make_float2(v, v); }
// flush-to-zero product: the reference integrator runs with FTZ/DAZ (flushDenormals = true by default,
// mitsuba/src/integrators/sdmm/volpath_sdmm.cpp:62, 85-87), so a component density below 2^-126 is exactly 0
__device__ __forceinline__ float2 mul2_ftz(float2 a, float2 b) {
    unsigned long long ra, rb, rc;
    ra = *reinterpret_cast<unsigned long long*>(&a);
    rb = *reinterpret_cast<unsigned long long*>(&b);
    asm("mul.rn.ftz.f32x2 %0, %1, %2;" : "=l"(rc) : "l"(ra), "l"(rb));
    return *reinterpret_cast<float2*>(&rc);
}

__device__ __forceinline__ float rsqrt_approx(float x) {
    float y;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// ------------------------------------------------------------------------------------------------
// K0: SoA planes -> packed records.  rec[2n] = (x, y, z, w), rec[2n+1] = (dx, dy, dz, w_eff).
// HBM-bound: 28 B read + 32 B written per sample, fully coalesced.
// ------------------------------------------------------------------------------------------------
__global__ void pack_samples_kernel(const float* __restrict__ point6, size_t stride,
                                    const float* __restrict__ weight, float4* __restrict__ rec,
                                    size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t step = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += step) {
        float4 a = make_float4(point6[i], point6[stride + i], point6[2 * stride + i], weight[i]);
        // b.w = effective weight: 0 for samples the EM loop skips (R: opt/data.h:20-23, em.h:267-269)
        const float w = a.w;
        const float w_eff = (w >= 1e-8f && w <= 3.402823466e+38f) ? w : 0.f;
        float4 b = make_float4(point6[3 * stride + i], point6[4 * stride + i],
                               point6[5 * stride + i], w_eff);
        rec[2 * i] = a;
        rec[2 * i + 1] = b;
    }
}

// ------------------------------------------------------------------------------------------------
// K1: E-step + sufficient statistics
// ------------------------------------------------------------------------------------------------
// theta/sin(theta) as a function of c = cos(theta) without acos, sqrt or a division:
//   acos(a) = sqrt(1-a) * P(a) on [0,1]  (degree-6 minimax fit of acos(a)/sqrt(1-a), max rel. error
//   1.9e-7 evaluated in float32),  acos(-a) = pi - acos(a),  sin = sqrt(1-a) sqrt(1+a)
//   => c >= 0: r = P(a) * rsqrt(1+a);   c < 0: r = (pi * rsqrt(1-a) - P(a)) * rsqrt(1+a)
// replaces safe_acos / safe_sqrt / divide of R: spaces/spatio_directional.h:101-108.  The
// reference's `sin < 1e-4 -> 1` guard only triggers for |c| == 1 in float32; clamping a to the
// largest float below 1 keeps both rsqrt arguments positive and gives r -> 1 at c = 1.
#define SDMMB_ACOS_C0 1.570796172e+00f
#define SDMMB_ACOS_C1 -2.145852374e-01f
#define SDMMB_ACOS_C2 8.875543537e-02f
#define SDMMB_ACOS_C3 -4.880191494e-02f
#define SDMMB_ACOS_C4 2.689914581e-02f
#define SDMMB_ACOS_C5 -1.114577623e-02f
#define SDMMB_ACOS_C6 2.295893482e-03f
#define SDMMB_ONE_MINUS 0.99999994f
// Sample weights enter the accumulation scaled by 2^-64 and the partial sums leave scaled by 2^64
// (exact powers of two: bit-identical statistics).  w_n / sum_k pi_k p_k overflows float32 when the
// mixture density at an outlier is tiny (the reference forms gamma = p / sum <= 1 first, em.h:295-309).
#define SDMMB_W_SCALE 5.421010862427522e-20f
#define SDMMB_W_UNSCALE 18446744073709551616.f
// Broadcast constants as 64-bit pairs in constant memory so FFMA2 can take them as a c[][] operand
// instead of burning two registers (or two MOVs per use) each.
__constant__ float2 kC2[11] = {
    {SDMMB_ACOS_C0, SDMMB_ACOS_C0}, {SDMMB_ACOS_C1, SDMMB_ACOS_C1}, {SDMMB_ACOS_C2, SDMMB_ACOS_C2},
    {SDMMB_ACOS_C3, SDMMB_ACOS_C3}, {SDMMB_ACOS_C4, SDMMB_ACOS_C4}, {SDMMB_ACOS_C5, SDMMB_ACOS_C5},
    {SDMMB_ACOS_C6, SDMMB_ACOS_C6}, {1.f, 1.f}, {-1.f, -1.f}, {-3.14159265358979323846f, -3.14159265358979323846f},
    {-1.5f, -1.5f}};

struct PairOut {
    float2 p;     // pi_k p_k(x) for the two components of the pair
    float2 t[5];  // tangent coordinates
};

// prm: 25 packed parameter pairs (kernels.h: EREC_*).  `ndz` is MINUS the sample's direction z.
//
// d_loc = R_to d without the 3x3 matrix.  The frame is always the Duff et al. basis of the unit mean
// direction n (R: linalg/coordinate_system.h:21-38): rows of R_to are
//   s = (1 + sign a nx^2, sign b, -sign nx),  t = (b, sign + a ny^2, -ny),  n,   b = nx ny a,
// so with q = nx dx + ny dy and h = a q - dz:
//   s.d = dx + sign nx h,   t.d = sign (dy + sign ny h),   n.d = q + nz dz.
// 6 FMA-pipe ops instead of 9 and 6 parameter registers instead of 9.  The kernel carries
// ly' = sign * (t.d); the factor is folded into W(4,4) for the Mahalanobis form and undone on the
// five t_4-odd statistics when the M-step reduces the partials.
__device__ __forceinline__ PairOut estep_pair(const float2* __restrict__ prm, float2 xx, float2 yy,
                                              float2 zz, float2 dx, float2 dy, float2 ndz) {
    PairOut o;
    const float2* F = prm + EREC_ROT;  // nx, ny, -nz, a, sign nx, sign ny
    float2 q = fma2(F[1], dy, mul2(F[0], dx));
    float2 lz = fma2(F[2], ndz, q);
    float2 h = fma2(F[3], q, ndz);
    float2 lx = fma2(F[4], h, dx);
    float2 ly = fma2(F[5], h, dy);
    float2 a = make_float2(fminf(fabsf(lz.x), SDMMB_ONE_MINUS), fminf(fabsf(lz.y), SDMMB_ONE_MINUS));
    float2 P = fma2(a, kC2[6], kC2[5]);
    P = fma2(P, a, kC2[4]);
    P = fma2(P, a, kC2[3]);
    P = fma2(P, a, kC2[2]);
    P = fma2(P, a, kC2[1]);
    P = fma2(P, a, kC2[0]);
    float2 opa = add2(a, kC2[7]);
    float2 oma = fma2(a, kC2[8], kC2[7]);
    float2 u = make_float2(rsqrt_approx(opa.x), rsqrt_approx(opa.y));
    float2 v = make_float2(rsqrt_approx(oma.x), rsqrt_approx(oma.y));
    float2 alt = fma2(v, kC2[9], P);  // P - pi*rsqrt(1-a) = -(numerator for c < 0)
    float2 sel = make_float2(lz.x >= 0.f ? P.x : -alt.x, lz.y >= 0.f ? P.y : -alt.y);
    float2 r = mul2(sel, u);  // theta / sin(theta) = inv_jacobian    R: spatio_directional.h:105-110
    o.t[0] = add2(xx, prm[EREC_NMU + 0]);
    o.t[1] = add2(yy, prm[EREC_NMU + 1]);
    o.t[2] = add2(zz, prm[EREC_NMU + 2]);
    o.t[3] = mul2(lx, r);
    o.t[4] = mul2(ly, r);
    // z = W t with W = sqrt(log2(e)/2) * L^-1 (replaces linalg::solve, R: linalg/cholesky.h:38-58),
    // column by column: the five rows are independent chains and consecutive instructions share t_c
    // (an operand repeated in the same slot comes from the reuse cache instead of the register file)
    const float2* W = prm + EREC_W;
    float2 z0 = mul2(W[0], o.t[0]);
    float2 z1 = mul2(W[1], o.t[0]);
    float2 z2 = mul2(W[3], o.t[0]);
    float2 z3 = mul2(W[6], o.t[0]);
    float2 z4 = mul2(W[10], o.t[0]);
    z1 = fma2(W[2], o.t[1], z1);
    z2 = fma2(W[4], o.t[1], z2);
    z3 = fma2(W[7], o.t[1], z3);
    z4 = fma2(W[11], o.t[1], z4);
    z2 = fma2(W[5], o.t[2], z2);
    z3 = fma2(W[8], o.t[2], z3);
    z4 = fma2(W[12], o.t[2], z4);
    z3 = fma2(W[9], o.t[3], z3);
    z4 = fma2(W[13], o.t[3], z4);
    z4 = fma2(W[14], o.t[4], z4);
    // e = |z|^2 (log2 units);  p = (inv_jac * c_k) * 2^-e                   R: distributions/sdmm.h:292-294, 337
    // The reference integrator runs with FTZ/DAZ (flushDenormals = true, volpath_sdmm.cpp:62, 85-87): its exp
    // result is dead below 2^-126 and so is every product below 2^-126.  ex2.ftz gives 0 for e > 126 whatever
    // the component's coefficient c_k, the FTZ product does the rest -- this decides which outlier samples the
    // E-step skips (hsum == 0, opt/em.h:289-293).
    float2 e = mul2(z0, z0);
    e = fma2(z1, z1, e);
    e = fma2(z2, z2, e);
    e = fma2(z3, z3, e);
    e = fma2(z4, z4, e);
    float2 ex = make_float2(ex2_approx(-e.x), ex2_approx(-e.y));
    o.p = mul2_ftz(mul2(r, prm[EREC_NLC]), ex);
    return o;
}

// The same E-step for S samples in lockstep, written operation by operation across the samples so
// that consecutive instructions share their per-component parameter operand: the register file
// delivers one 64-bit operand per cycle (profiles/ubench_r1.txt), an operand repeated in the same
// slot of the next instruction comes from the reuse cache for free.
template <int S>
__device__ __forceinline__ void estep_pair_multi(const float2* __restrict__ prm, const float4 (&A)[S],
                                                 const float4 (&B)[S], PairOut (&o)[S]) {
    const float2* F = prm + EREC_ROT;
    const float2* W = prm + EREC_W;
    float2 q[S], lz[S], h[S], lx[S], ly[S], a[S], P[S], opa[S], oma[S], u[S], v[S], r[S];
#define SDMMB_EACH for (int s = 0; s < S; ++s)
#pragma unroll
    SDMMB_EACH q[s] = mul2(F[0], splat(A[s].w));
#pragma unroll
    SDMMB_EACH q[s] = fma2(F[1], splat(B[s].x), q[s]);
#pragma unroll
    SDMMB_EACH lz[s] = fma2(F[2], splat(B[s].y), q[s]);
#pragma unroll
    SDMMB_EACH h[s] = fma2(F[3], q[s], splat(B[s].y));
#pragma unroll
    SDMMB_EACH lx[s] = fma2(F[4], h[s], splat(A[s].w));
#pragma unroll
    SDMMB_EACH ly[s] = fma2(F[5], h[s], splat(B[s].x));
#pragma unroll
    SDMMB_EACH a[s] = make_float2(fminf(fabsf(lz[s].x), SDMMB_ONE_MINUS), fminf(fabsf(lz[s].y), SDMMB_ONE_MINUS));
#pragma unroll
    SDMMB_EACH P[s] = fma2(a[s], kC2[6], kC2[5]);
#pragma unroll
    for (int c = 4; c >= 0; --c)
#pragma unroll
        SDMMB_EACH P[s] = fma2(P[s], a[s], kC2[c]);
#pragma unroll
    SDMMB_EACH {
        opa[s] = add2(a[s], kC2[7]);
        oma[s] = fma2(a[s], kC2[8], kC2[7]);
        u[s] = make_float2(rsqrt_approx(opa[s].x), rsqrt_approx(opa[s].y));
        v[s] = make_float2(rsqrt_approx(oma[s].x), rsqrt_approx(oma[s].y));
    }
#pragma unroll
    SDMMB_EACH {
        float2 alt = fma2(v[s], kC2[9], P[s]);
        float2 sel = make_float2(lz[s].x >= 0.f ? P[s].x : -alt.x, lz[s].y >= 0.f ? P[s].y : -alt.y);
        r[s] = mul2(sel, u[s]);
    }
#pragma unroll
    SDMMB_EACH o[s].t[0] = add2(splat(A[s].x), prm[EREC_NMU + 0]);
#pragma unroll
    SDMMB_EACH o[s].t[1] = add2(splat(A[s].y), prm[EREC_NMU + 1]);
#pragma unroll
    SDMMB_EACH o[s].t[2] = add2(splat(A[s].z), prm[EREC_NMU + 2]);
#pragma unroll
    SDMMB_EACH o[s].t[3] = mul2(lx[s], r[s]);
#pragma unroll
    SDMMB_EACH o[s].t[4] = mul2(ly[s], r[s]);
    // z = W t, row by row; within a row the W entry is the shared operand across the samples
    float2 z[S][5];
    int e = 0;
#pragma unroll
    for (int row = 0; row < 5; ++row) {
#pragma unroll
        for (int c = 0; c <= row; ++c) {
#pragma unroll
            SDMMB_EACH z[s][row] = (c == 0) ? mul2(W[e], o[s].t[0]) : fma2(W[e], o[s].t[c], z[s][row]);
            ++e;
        }
    }
    float2 ee[S];
#pragma unroll
    SDMMB_EACH ee[s] = mul2(z[s][0], z[s][0]);
#pragma unroll
    for (int row = 1; row < 5; ++row)
#pragma unroll
        SDMMB_EACH ee[s] = fma2(z[s][row], z[s][row], ee[s]);
#pragma unroll
    SDMMB_EACH {
        float2 ex = make_float2(ex2_approx(-ee[s].x), ex2_approx(-ee[s].y));
        o[s].p = mul2_ftz(mul2(r[s], prm[EREC_NLC]), ex);
    }
#undef SDMMB_EACH
}

struct PairAcc {
    float2 s0;
    float2 s1[5];
    float2 s2[15];
};

// Stats::plus_eq (R: opt/em.h:57-62) on the 15 unique covariance entries.
__device__ __forceinline__ void accumulate_pair(PairAcc& acc, const PairOut& o, float2 ss) {
    float2 g = mul2(o.p, ss);
    acc.s0 = add2(acc.s0, g);
    float2 gt[5];
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        gt[i] = mul2(g, o.t[i]);
        acc.s1[i] = add2(acc.s1[i], gt[i]);
    }
    int e = 0;
#pragma unroll
    for (int i = 0; i < 5; ++i)
#pragma unroll
        for (int j = 0; j <= i; ++j) {
            acc.s2[e] = fma2(gt[i], o.t[j], acc.s2[e]);
            ++e;
        }
}

// One sample-step of a lane: E-step of all its NP component pairs for the sample at smem slot `si`,
// plus the lane-local part of the mixture normaliser.
template <int NP>
struct StepOut {
    PairOut o[NP];
    float sum;
};

template <int NP>
__device__ __forceinline__ StepOut<NP> estep_sample(const float2 (&prm)[NP][EREC_PLANES],
                                                    const float4* __restrict__ wbuf, int si) {
    StepOut<NP> r;
    // (x, y, z, dx) (dy, -dz, w_eff, -): scalars, splat into both halves of the packed operands so
    // that the compiler can use the 32-bit broadcast operand form (half the register-file traffic)
    const float4 A = wbuf[si * 2 + 0], B = wbuf[si * 2 + 1];
#pragma unroll
    for (int q = 0; q < NP; ++q) {
        r.o[q] = estep_pair(prm[q], splat(A.x), splat(A.y), splat(A.z), splat(A.w), splat(B.x), splat(B.y));
        const float pq = r.o[q].p.x + r.o[q].p.y;
        r.sum = (q == 0) ? pq : r.sum + pq;
    }
    return r;
}

// One sample as the kernel stages it: a = (x, y, z, w), b = (dx, dy, dz, w_eff).  SOA reads the caller's
// planes directly (7 coalesced 4-byte loads per lane per 32-sample chunk) and applies pack_samples_kernel's
// weight rule on the fly, so the borrowed binding needs no packing pass and no second copy of the samples.
template <bool SOA>
__device__ __forceinline__ void load_sample(const float4* __restrict__ samples, const float* __restrict__ p6,
                                            size_t stride, const float* __restrict__ wt, uint32_t idx,
                                            float4& a, float4& b) {
    if constexpr (SOA) {
        const float w = wt[idx];
        const float w_eff = (w >= 1e-8f && w <= 3.402823466e+38f) ? w : 0.f;
        a = make_float4(p6[idx], p6[stride + idx], p6[2 * stride + idx], w);
        b = make_float4(p6[3 * stride + idx], p6[4 * stride + idx], p6[5 * stride + idx], w_eff);
    } else {
        a = samples[2 * (size_t)idx];
        b = samples[2 * (size_t)idx + 1];
    }
}

template <int NP, int LPS, int WARPS, int MINB, bool LL, int MODE, bool SOA = false>
__global__ void __launch_bounds__(WARPS * 32, MINB)
em_estep_kernel(const float4* __restrict__ samples, const EItem* __restrict__ items,
                const float* __restrict__ erec, float* __restrict__ partials,
                float* __restrict__ item_scal, float* __restrict__ out_loglik,
                const float* __restrict__ soa_p6 = nullptr, size_t soa_stride = 0,
                const float* __restrict__ soa_w = nullptr) {
    constexpr int K = 2 * NP * LPS;
    constexpr int SPW = 32 / LPS;      // samples a warp processes per step
    constexpr int STEPS = 32 / SPW;    // steps per 32-sample chunk
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4* sbuf = reinterpret_cast<float4*>(smem_raw);                     // [WARPS][32][2] float4
    float* red = reinterpret_cast<float*>(smem_raw + WARPS * 32 * 32);      // [WARPS][21][K]
    __shared__ float s_wsum[WARPS];
    __shared__ int s_skip[WARPS];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int grp = lane / LPS, l = lane % LPS;
    const EItem item = items[blockIdx.x];

    // this lane's component parameters -> registers (coalesced over lanes within a plane)
    float2 prm[NP][EREC_PLANES];
    {
        const float* base = erec + (size_t)item.cell * EREC_PLANES * K;
#pragma unroll
        for (int q = 0; q < NP; ++q)
#pragma unroll
            for (int pl = 0; pl < EREC_PLANES; ++pl) {
                prm[q][pl].x = base[pl * K + q * 2 * LPS + l];
                prm[q][pl].y = base[pl * K + q * 2 * LPS + LPS + l];
            }
    }
    PairAcc acc[NP];
#pragma unroll
    for (int q = 0; q < NP; ++q) {
        acc[q].s0 = splat(0.f);
#pragma unroll
        for (int i = 0; i < 5; ++i) acc[q].s1[i] = splat(0.f);
#pragma unroll
        for (int i = 0; i < 15; ++i) acc[q].s2[i] = splat(0.f);
    }
    // per-lane bookkeeping of the lane's OWN sample of each chunk (reduced once at the end)
    float wsum = 0.f;   // sum of all stored weights                      R: opt/data.h:104-113
    int n_wvalid = 0;   // samples with a valid weight                    R: opt/data.h:20-23
    int n_ok = 0;       // samples that contributed (valid weight, finite non-zero normaliser)

    float4* wbuf = sbuf + warp * 32 * 2;
    const uint32_t n_chunks = (item.end - item.begin + 31u) / 32u;
    // CTA-uniform trip count (every warp runs the same number of rounds; a warp whose chunk lies
    // beyond the item processes 32 zero-weight padding samples).  Keeping the loop bound free of
    // the warp index lets the compiler prove the warp converged at the shuffles below, so the
    // whole round stays one basic block that the scheduler can interleave.
    const uint32_t n_rounds = (n_chunks + WARPS - 1) / WARPS;
    // register prefetch of this lane's sample of the warp's first chunk; rb.w = effective weight
    float4 ra = make_float4(0.f, 0.f, 0.f, 0.f), rb = make_float4(0.f, 0.f, 1.f, 0.f);
    {
        uint32_t idx = item.begin + warp * 32u + lane;
        if (idx < item.end) load_sample<SOA>(samples, soa_p6, soa_stride, soa_w, idx, ra, rb);
    }
    for (uint32_t round = 0; round < n_rounds; ++round) {
        const uint32_t cbase = item.begin + (round * WARPS + warp) * 32u;
        __syncwarp();
        // per-warp staging slot of a sample: (x, y, z, dx) (dy, -dz, w_eff, -)
        wbuf[lane * 2 + 0] = make_float4(ra.x, ra.y, ra.z, rb.x);
        wbuf[lane * 2 + 1] = make_float4(rb.y, -rb.z, rb.w * SDMMB_W_SCALE, 0.f);
        wsum += ra.w;
        n_wvalid += (rb.w > 0.f) ? 1 : 0;
        __syncwarp();
        {   // prefetch the next chunk while this one is processed
            uint32_t nidx = cbase + WARPS * 32u + lane;
            ra = make_float4(0.f, 0.f, 0.f, 0.f);
            rb = make_float4(0.f, 0.f, 1.f, 0.f);
            if (nidx < item.end) load_sample<SOA>(samples, soa_p6, soa_stride, soa_w, nidx, ra, rb);
        }
        if constexpr (MODE == 1) {
            // phased: E-step of two samples in lockstep (shared parameter operands), both normaliser
            // butterflies, then both accumulations; latency is covered by the other resident warps
#pragma unroll 1
            for (int s = 0; s < STEPS; s += 2) {
                const int si0 = s * SPW + grp;
                float4 A[2], B[2];
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    A[j] = wbuf[(si0 + j * SPW) * 2 + 0];
                    B[j] = wbuf[(si0 + j * SPW) * 2 + 1];
                }
                PairOut o[NP][2];
                float sum[2];
#pragma unroll
                for (int q = 0; q < NP; ++q) {
                    estep_pair_multi<2>(prm[q], A, B, o[q]);
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        const float pq = o[q][j].p.x + o[q][j].p.y;
                        sum[j] = (q == 0) ? pq : sum[j] + pq;
                    }
                }
#pragma unroll
                for (int off = LPS / 2; off > 0; off >>= 1)
#pragma unroll
                    for (int j = 0; j < 2; ++j) sum[j] += __shfl_xor_sync(0xffffffffu, sum[j], off);
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const int si = si0 + j * SPW;
                    const float rs = rcp_approx(sum[j]);
                    const bool ok = (rs > 0.f) && (rs <= 3.402823466e+38f);
                    const float sc = ok ? B[j].z * rs : 0.f;
                    n_ok += (ok && B[j].z > 0.f) ? 1 : 0;
                    if (LL) {
                        if (l == 0 && cbase + si < item.end) out_loglik[cbase + si] = logf(sum[j]);
                    }
                    const float2 ss = splat(sc);
#pragma unroll
                    for (int q = 0; q < NP; ++q) {
                        PairOut oo = o[q][j];
                        oo.p.x = ok ? oo.p.x : 0.f;
                        oo.p.y = ok ? oo.p.y : 0.f;
                        accumulate_pair(acc[q], oo, ss);
                    }
                }
            }
        } else {
        // software pipeline over the chunk's steps: the E-step of step s+1 is independent of the
        // shuffle butterfly + accumulation of step s, so both sit in one basic block and overlap.
        StepOut<NP> cur = estep_sample<NP>(prm, wbuf, grp);
#pragma unroll 2
        for (int s = 0; s < STEPS; ++s) {
            const int si = s * SPW + grp;
            // the pipeline tail recomputes the last sample (result unused) to stay branch-free
            const int sn = (s + 1 < STEPS) ? si + SPW : si;
            StepOut<NP> nxt = estep_sample<NP>(prm, wbuf, sn);
            float sum = cur.sum;
#pragma unroll
            for (int off = LPS / 2; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
            const float w = reinterpret_cast<const float*>(wbuf + si * 2 + 1)[2];
            // hsum; skip if 0 / non-finite                           R: opt/em.h:288-293
            const float rs = rcp_approx(sum);
            const bool ok = (rs > 0.f) && (rs <= 3.402823466e+38f);   // sum finite, > 0, 1/sum finite
            const float sc = ok ? w * rs : 0.f;                       // w_n / sum   R: em.h:295-298, 309
            n_ok += (ok && w > 0.f) ? 1 : 0;
            if (LL) {
                if (l == 0 && cbase + si < item.end) out_loglik[cbase + si] = logf(sum);
            }
            const float2 ss = splat(sc);
#pragma unroll
            for (int q = 0; q < NP; ++q) {
                // a NaN/inf normaliser must not leak 0 * inf into the statistics
                PairOut o = cur.o[q];
                o.p.x = ok ? o.p.x : 0.f;
                o.p.y = ok ? o.p.y : 0.f;
                accumulate_pair(acc[q], o, ss);
            }
            cur = nxt;
        }
            }
    }
    // n_ok was counted by every lane of a sample's group (identical values): keep lane l == 0's
    int nskip = 0;
    // ---- reduce: sample groups within the warp, then warps within the CTA (fixed order) --------
    if (SPW > 1) {
#pragma unroll
        for (int q = 0; q < NP; ++q) {
            float* f = reinterpret_cast<float*>(&acc[q]);
#pragma unroll
            for (int i = 0; i < 42; ++i)
#pragma unroll
                for (int off = LPS; off < 32; off <<= 1) f[i] += __shfl_xor_sync(0xffffffffu, f[i], off);
        }
    }
    // wsum / n_wvalid are per-lane partials over the lane's own samples; n_ok is replicated across the
    // lanes of a sample group, so only lane l == 0 of each group contributes it
    {
        float ws = wsum;
        int ns = n_wvalid - ((l == 0) ? n_ok : 0);
        (void)nskip;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            ws += __shfl_xor_sync(0xffffffffu, ws, off);
            ns += __shfl_xor_sync(0xffffffffu, ns, off);
        }
        if (lane == 0) { s_wsum[warp] = ws; s_skip[warp] = ns; }
    }
    if (grp == 0) {
        float* wr = red + (size_t)warp * ESTAT_PLANES * K;
#pragma unroll
        for (int q = 0; q < NP; ++q) {
            const int ka = q * 2 * LPS + l, kb = ka + LPS;
            wr[0 * K + ka] = acc[q].s0.x;
            wr[0 * K + kb] = acc[q].s0.y;
#pragma unroll
            for (int i = 0; i < 5; ++i) {
                wr[(1 + i) * K + ka] = acc[q].s1[i].x;
                wr[(1 + i) * K + kb] = acc[q].s1[i].y;
            }
#pragma unroll
            for (int i = 0; i < 15; ++i) {
                wr[(6 + i) * K + ka] = acc[q].s2[i].x;
                wr[(6 + i) * K + kb] = acc[q].s2[i].y;
            }
        }
    }
    __syncthreads();
    float* out = partials + (size_t)blockIdx.x * ESTAT_PLANES * K;
    for (int i = threadIdx.x; i < ESTAT_PLANES * K; i += WARPS * 32) {
        float s = red[i];
#pragma unroll
        for (int wv = 1; wv < WARPS; ++wv) s += red[(size_t)wv * ESTAT_PLANES * K + i];
        out[i] = s * SDMMB_W_UNSCALE;
    }
    if (threadIdx.x == 0) {
        float ws = s_wsum[0];
        int ns = s_skip[0];
#pragma unroll
        for (int wv = 1; wv < WARPS; ++wv) { ws += s_wsum[wv]; ns += s_skip[wv]; }
        item_scal[2 * (size_t)blockIdx.x] = ws;
        item_scal[2 * (size_t)blockIdx.x + 1] = __int_as_float(ns);
    }
}

// ------------------------------------------------------------------------------------------------
// per-component prepare: frame, Cholesky, E-step record                R: distributions/sdmm.h:186-197
// ------------------------------------------------------------------------------------------------
struct ComponentPrep {
    float to[9], from[9];
    float L[25];
    float inv_det;
    bool psd;
};

__device__ inline void prepare_component(const float mean[6], const float cov[25], ComponentPrep& cp) {
    frame_from_mean_dir(mean + 3, cp.to, cp.from);
    cp.psd = cholesky<5>(cov, cp.L);
    float prod = 1.f;
#pragma unroll
    for (int i = 0; i < 5; ++i) prod *= cp.L[i * 5 + i];
    cp.inv_det = 1.f / prod;
}

// pmf must be the NORMALISED weight (after Categorical::prepare).
__device__ inline void write_erec(float* __restrict__ erec_cell, int K, int k, float pmf,
                                  const float mean[6], const ComponentPrep& cp) {
    float W[25];
    inverse_lower_tri<5>(cp.L, W);
    const float scale = 0.84932180028801904272f;  // sqrt(log2(e) / 2)
#pragma unroll
    for (int d = 0; d < 3; ++d) erec_cell[(EREC_NMU + d) * K + k] = -mean[d];
    {   // factored Duff frame of the unit mean direction (rows of cp.to are s, t, n)
        const float nx = cp.to[6], ny = cp.to[7], nz = cp.to[8];
        const float sign = copysignf(1.f, nz);
        float* f = erec_cell + (size_t)EREC_ROT * K + k;
        f[0 * K] = nx;
        f[1 * K] = ny;
        f[2 * K] = -nz;
        f[3 * K] = -1.f / (sign + nz);
        f[4 * K] = sign * nx;
        f[5 * K] = sign * ny;
        int e = 0;
#pragma unroll
        for (int r = 0; r < 5; ++r)
#pragma unroll
            for (int c = 0; c <= r; ++c) {
                const float w = W[r * 5 + c] * scale;
                erec_cell[(EREC_W + e++) * K + k] = (c == 4) ? sign * w : w;
            }
    }
    // pi_k * inv_det * (2 pi)^-5/2; 0 switches a zero-weight component off
    erec_cell[EREC_NLC * K + k] = pmf * cp.inv_det * gaussian_normalization(5);
}

// ------------------------------------------------------------------------------------------------
// K3: conditioner prepare for one component          R: distributions/sdmm_conditioner.h:170-212, 7-26
// ------------------------------------------------------------------------------------------------
struct ConditionerOut {
    float La[9], inv_det_a, M[6], cc[4], Lc[4], inv_det_c;
};

__device__ inline void conditioner_component(const float cov[25], ConditionerOut& o) {
    float cov_aa[9], A[9];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) cov_aa[r * 3 + c] = cov[r * 5 + c];
    cholesky<3>(cov_aa, o.La);
    o.inv_det_a = 1.f / (o.La[0] * o.La[4] * o.La[8]);
    inverse_lower_tri<3>(o.La, A);
    float cov_ab[6], cov_ba[6], cov_bb[4];  // ab: 3x2, ba: 2x3; both read joint.cov(3+r, c) (:199-201)
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            cov_ba[r * 3 + c] = cov[(3 + r) * 5 + c];
            cov_ab[c * 2 + r] = cov[(3 + r) * 5 + c];
        }
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int c = 0; c < 2; ++c) cov_bb[r * 2 + c] = cov[(3 + r) * 5 + 3 + c];
    float Aab[6];  // 3x2
    matmul<3, 3, 2>(A, cov_ab, Aab);
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            float s = 0.f;
#pragma unroll
            for (int k = 0; k < 3; ++k) s += Aab[k * 2 + i] * Aab[k * 2 + j];
            o.cc[i * 2 + j] = cov_bb[i * 2 + j] - s;
        }
    float At[9], tmp[6];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) At[r * 3 + c] = A[c * 3 + r];
    matmul<2, 3, 3>(cov_ba, At, tmp);  // (cov_ba * A^T) * A, left to right (:206-207)
    matmul<2, 3, 3>(tmp, A, o.M);
    cholesky<2>(o.cc, o.Lc);
    o.inv_det_c = 1.f / (o.Lc[0] * o.Lc[3]);
}

// Query record (kernels.h QREC_*).  pmf_a = normalised marginal weight.
__device__ inline void write_qrec(float* __restrict__ q, float pmf_a, const float mean[6],
                                  const float to[9], const ConditionerOut& o) {
    const float scale = 0.84932180028801904272f;  // sqrt(log2(e) / 2)
    float Wa[9];
    inverse_lower_tri<3>(o.La, Wa);
    q[QREC_MU + 0] = mean[0]; q[QREC_MU + 1] = mean[1]; q[QREC_MU + 2] = mean[2];
    q[QREC_WA + 0] = Wa[0] * scale; q[QREC_WA + 1] = Wa[3] * scale; q[QREC_WA + 2] = Wa[4] * scale;
    q[QREC_WA + 3] = Wa[6] * scale; q[QREC_WA + 4] = Wa[7] * scale; q[QREC_WA + 5] = Wa[8] * scale;
    float ca = pmf_a * o.inv_det_a * gaussian_normalization(3);
    q[QREC_NLCA] = (ca > 0.f) ? -log2f(ca) : (ca == 0.f ? INFINITY : NAN);
#pragma unroll
    for (int i = 0; i < 6; ++i) q[QREC_M + i] = o.M[i];
#pragma unroll
    for (int i = 0; i < 9; ++i) q[QREC_R + i] = to[i];
    q[QREC_LC + 0] = o.Lc[0]; q[QREC_LC + 1] = o.Lc[2]; q[QREC_LC + 2] = o.Lc[3];
    // inverse of the 2x2 lower-triangular factor, scaled
    float w00 = 1.f / o.Lc[0], w11 = 1.f / o.Lc[3];
    float w10 = (0.f - o.Lc[2] * w00) / o.Lc[3];
    q[QREC_WC + 0] = w00 * scale; q[QREC_WC + 1] = w10 * scale; q[QREC_WC + 2] = w11 * scale;
    float cc = o.inv_det_c * gaussian_normalization(2);
    q[QREC_NLCC] = (cc > 0.f) ? -log2f(cc) : (cc == 0.f ? INFINITY : NAN);
}

// Sequential Categorical::prepare over the K weights of one cell held in shared memory
// (R: distributions/categorical.h:86-114).  Called by ONE thread per cell; keeps the reference's
// left-to-right float32 association order.  Returns validity; on success pmf and cdf are scaled.
// Split in two so that only the order-dependent part is serial: ONE thread per cell walks the prefix sum
// left to right (the float32 association order decides the cdf bits), then EVERY thread scales its own
// entry by 1/total -- element for element the same operations as the reference's loop.
__device__ inline float categorical_prefix_seq(const float* pmf, float* cdf, int K) {
    float run = pmf[0];
    cdf[0] = run;
#pragma unroll 8
    for (int i = 1; i < K; ++i) {
        run += pmf[i];
        cdf[i] = run;
    }
    return run;
}
__device__ __forceinline__ bool categorical_scale_own(float* pmf_k, float* cdf_k, float total) {
    if (!(total > 1e-20f)) return false;
    const float inv = 1.f / total;
    *cdf_k *= inv;
    *pmf_k *= inv;
    return true;
}

// ------------------------------------------------------------------------------------------------
// K3 epilogue: record exchange over peer memory (replaces the all-gather that would follow).
// Called by every thread of every CTA after the CTA's records are in local memory (and after a
// __syncthreads()).  The CTA's records are one contiguous run of MSTEP_THREADS * 128 bytes: read
// once (L2-hot), stored coalesced to the same offset of every peer's array, peers visited in
// rotated order so that the ranks do not all hit the same destination at once.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// spin until *flag >= epoch (epochs only grow); false on timeout
__device__ __forceinline__ bool wait_flag(const uint32_t* flag, uint32_t epoch, uint32_t timeout_ms) {
    if ((int32_t)(ld_acquire_sys(flag) - epoch) >= 0) return true;
    const unsigned long long t0 = globaltimer_ns(), limit = (unsigned long long)timeout_ms * 1000000ull;
    while ((int32_t)(ld_acquire_sys(flag) - epoch) < 0) {
        __nanosleep(200);
        if (globaltimer_ns() - t0 > limit) return false;
    }
    return true;
}

__device__ __forceinline__ void peer_push_records(const MStepArgs& a) {
    const PeerPush& P = a.peer;
    const int K = a.K;
    const int cells_per_cta = MSTEP_THREADS / K;
    const int c0 = blockIdx.x * cells_per_cta;
    const int nc = min(cells_per_cta, a.n_cells - c0);
    uint32_t* lf = P.flags[P.rank];
    __shared__ int s_go;
    if (threadIdx.x == 0) s_go = 1;
    __syncthreads();
    // a peer's records may only be overwritten once it has entered this training step
    if (threadIdx.x < P.world && (int)threadIdx.x != P.rank) {
        if (!wait_flag(lf + PEER_READY + threadIdx.x, P.epoch, P.timeout_ms)) {
            atomicAdd(lf + PEER_ERROR, 1u);
            s_go = 0;
        }
    }
    __syncthreads();
    if (s_go && nc > 0) {
        constexpr int V = QREC_FLOATS / 4;  // float4 per record = per thread
        const size_t off = ((size_t)(a.first_cell + c0) * K) * V;
        const int n4 = nc * K * V;
        // the local array the records were written to may be the async back copy; peers always
        // receive into their published array
        const float4* src = reinterpret_cast<const float4*>(a.qrec) + off;
        float4 v[V];
#pragma unroll
        for (int j = 0; j < V; ++j) {
            const int idx = j * MSTEP_THREADS + threadIdx.x;
            v[j] = (idx < n4) ? __ldcg(src + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        for (int pp = 1; pp < P.world; ++pp) {
            const int p = (P.rank + pp) % P.world;
            float4* dst = reinterpret_cast<float4*>(P.qrec[p]) + off;
#pragma unroll
            for (int j = 0; j < V; ++j) {
                const int idx = j * MSTEP_THREADS + threadIdx.x;
                if (idx < n4) dst[idx] = v[j];
            }
        }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t done = atomicAdd(lf + PEER_COUNTER, 1u);
        if (done == gridDim.x - 1) {   // every CTA's stores are fenced: publish this rank's arrival
            lf[PEER_COUNTER] = 0;
            __threadfence_system();
            for (int pp = 1; pp < P.world; ++pp) {
                const int p = (P.rank + pp) % P.world;
                st_release_sys(P.flags[p] + PEER_ARRIVED + P.rank, P.epoch);
            }
        }
    }
}

__global__ void peer_signal_kernel(PeerPush P) {
    const int p = threadIdx.x;
    if (p < P.world && p != P.rank) st_release_sys(P.flags[p] + PEER_READY + P.rank, P.epoch);
}

__global__ void peer_wait_kernel(PeerPush P) {
    const int p = threadIdx.x;
    uint32_t* lf = P.flags[P.rank];
    if (p < P.world && p != P.rank) {
        if (!wait_flag(lf + PEER_ARRIVED + p, P.epoch, P.timeout_ms)) atomicAdd(lf + PEER_ERROR, 1u);
    }
}

// ------------------------------------------------------------------------------------------------
// K2 (+K3): M-step.  One thread per (cell, component); MSTEP_THREADS / K cells per CTA.
// mode bit 0: run the EM update (else only prepare from the stored model)
// mode bit 1: also write the conditioner's query records
// mode bit 2: normalise the stored pmf (upload path: prepare_vectorized on a fresh model)
// ------------------------------------------------------------------------------------------------
template <int MINB>
__global__ void __launch_bounds__(MSTEP_THREADS, MINB)
em_mstep_kernel(MStepArgs a) {
    const int K = a.K;
    const int cells_per_cta = MSTEP_THREADS / K;
    const int local = threadIdx.x / K, k = threadIdx.x % K;
    const int ci = blockIdx.x * cells_per_cta + local;  // index into the launch's cell range
    __shared__ float s_pmf[MSTEP_THREADS];
    __shared__ float s_cdf[MSTEP_THREADS];
    __shared__ int s_flag[MSTEP_THREADS / 16 + 1];
    __shared__ int s_psd[MSTEP_THREADS / 16 + 1];
    __shared__ float s_tot[MSTEP_THREADS / 16 + 1];
    const bool active = ci < a.n_cells;
    const int cell = a.first_cell + (active ? ci : 0);

    float* joint = a.joint + (size_t)cell * JOINT_PLANES * K;
    float* em = a.em + (size_t)cell * EM_PLANES * K;
    uint32_t* scal = a.scal + (size_t)cell * SCALAR_WORDS;

    float pmf = 0.f, mean[6], cov[25];
    bool do_update = false;
    uint32_t status = 0;
    if (threadIdx.x < MSTEP_THREADS / 16 + 1) { s_flag[threadIdx.x] = 0; s_psd[threadIdx.x] = 1; }
    __syncthreads();
    if (active) {
        pmf = joint[k];
#pragma unroll
        for (int d = 0; d < 6; ++d) mean[d] = joint[(1 + d) * K + k];
#pragma unroll
        for (int i = 0; i < 25; ++i) cov[i] = joint[(7 + i) * K + k];
    }

    // canBeOptimized (volpath_sdmm.cpp:140-149): cells that hold too few samples are not trained
    const bool trainable = active && (a.mode & 1) && a.cell_sample_count[ci] >= a.min_samples;
    if (trainable) {
        // ---- reduce the E-step partials of this cell's items (fixed order) ---------------------
        const uint32_t ib = a.cell_item_begin[ci] - a.item_base, ie = a.cell_item_begin[ci + 1] - a.item_base;
        float bt[ESTAT_PLANES];
#pragma unroll
        for (int i = 0; i < ESTAT_PLANES; ++i) bt[i] = 0.f;
        float weight_sum = 0.f;
        int nskip = 0;
        for (uint32_t it = ib; it < ie; ++it) {
            const float* p = a.partials + (size_t)it * ESTAT_PLANES * K;
#pragma unroll
            for (int i = 0; i < ESTAT_PLANES; ++i) bt[i] += p[i * K + k];
            weight_sum += a.item_scal[2 * (size_t)it];
            nskip += __float_as_int(a.item_scal[2 * (size_t)it + 1]);
        }
        {   // the E-step accumulated with sign * t_4 (see estep_pair): undo on the t_4-odd statistics
            const float sgn = copysignf(1.f, mean[5]);
            bt[5] *= sgn;
#pragma unroll
            for (int j = 0; j < 4; ++j) bt[16 + j] *= sgn;
        }
        const int n_samples = (int)(a.cell_sample_count[ci]);
        float total_weight = __uint_as_float(scal[0]);
        int iterations_run = (int)scal[1];
        int samples_seen = (int)scal[2];
        const float weight_prior = __uint_as_float(scal[3]);
        const float cov_prior_strength = __uint_as_float(scal[4]);
        const float learning_rate = __uint_as_float(scal[5]);
        const float alpha = __uint_as_float(scal[6]);
        samples_seen += n_samples - nskip;                               // R: opt/em.h:264, 290
        // eta = (learning_rate * iterations_run + 1)^-alpha             R: opt/em.h:318-319
        const float eta = powf(learning_rate * (float)iterations_run + 1.f, -alpha);
        // interpolate_stats                                            R: opt/em.h:316-328
        float st_w = em[k];
        st_w *= 1.f - eta;
        st_w = fmaf(eta, bt[0], st_w);
        float st_mean[5], st_cov[25];
#pragma unroll
        for (int i = 0; i < 5; ++i) {
            float v = em[(1 + i) * K + k];
            v *= 1.f - eta;
            v += eta * bt[1 + i];
            st_mean[i] = v;
        }
#pragma unroll
        for (int i = 0; i < 5; ++i)
#pragma unroll
            for (int j = 0; j < 5; ++j) {
                const int hi = i > j ? i : j, lo = i > j ? j : i;
                float v = em[(6 + i * 5 + j) * K + k];
                v *= 1.f - eta;
                v += eta * bt[6 + hi * (hi + 1) / 2 + lo];
                st_cov[i * 5 + j] = v;
            }
        // normalize_stats                                              R: opt/em.h:330-351
        total_weight = (1.f - eta) * total_weight + eta * weight_sum;
        const float rcp_weight_sum = 1.f / total_weight;
        const bool norm_ok = isfinite(rcp_weight_sum) && total_weight != 0.f;
        em[k] = st_w;
        if (norm_ok) {
            do_update = true;
            float snw = rcp_weight_sum * st_w, snm[5], snc[25];
#pragma unroll
            for (int i = 0; i < 5; ++i) snm[i] = rcp_weight_sum * st_mean[i];
#pragma unroll
            for (int i = 0; i < 25; ++i) snc[i] = rcp_weight_sum * st_cov[i];
            // update_model                                             R: opt/em.h:353-488
            const int it = iterations_run < 30 ? iterations_run : 30;
            const float weight_prior_decay = (float)(1.0 / pow(3.0, (double)it));
            const float cov_prior_strength_decay = (float)(1.0 / pow(2.0, (double)it));
            const float weight_prior_decayed = weight_prior * weight_prior_decay;
            const float cps_decayed = cov_prior_strength * cov_prior_strength_decay;
            const float rcp_weight = 1.f / snw;
            const float rcp_cov_weight = 1.f / (cps_decayed + snw);
            const bool non_zero = pmf > 0.f;
            const bool finite_rcp = isfinite(rcp_weight);
            const bool good = non_zero && finite_rcp;
            const float upd_weight =
                non_zero ? (finite_rcp ? snw + weight_prior_decayed : weight_prior_decayed) : 0.f;
            float upd_mean[5], upd_cov[25];
#pragma unroll
            for (int i = 0; i < 5; ++i) upd_mean[i] = good ? snm[i] * rcp_weight : 0.f;
#pragma unroll
            for (int i = 0; i < 5; ++i)
#pragma unroll
                for (int j = 0; j < 5; ++j) {
                    float mean_sub = (snm[i] * snm[j]) * rcp_weight;
                    float cov_unnorm = snc[i * 5 + j] - mean_sub + em[(31 + i * 5 + j) * K + k] * cps_decayed;
                    float cov_norm = cov_unnorm * rcp_cov_weight;
                    upd_cov[i * 5 + j] = good ? cov_norm + em[(56 + i * 5 + j) * K + k] : cov[i * 5 + j];
                }
            pmf = upd_weight;
            // re-centre the tangent space and transport covariance + statistics   R: em.h:470-487
            float to_old[9], from_old[9], mean_new[6], to_new[9], from_new[9], J[25];
            frame_from_mean_dir(mean + 3, to_old, from_old);
            recentre_jacobian(mean, from_old, upd_mean, mean_new, to_new, from_new, J);
#pragma unroll
            for (int d = 0; d < 6; ++d) mean[d] = mean_new[d];
            congruence5(J, upd_cov, cov);
#pragma unroll
            for (int i = 0; i < 5; ++i)
#pragma unroll
                for (int j = 0; j < 5; ++j) snc[i * 5 + j] -= (snm[i] * snm[j]) * rcp_weight;
            float carried[25];
            congruence5(J, snc, carried);
            if (good) {
#pragma unroll
                for (int i = 0; i < 25; ++i) st_cov[i] = carried[i] * total_weight;
#pragma unroll
                for (int i = 0; i < 5; ++i) st_mean[i] = 0.f;
            }
            ++iterations_run;
        } else {
            status |= ST_STEP_ABORTED;
        }
#pragma unroll
        for (int i = 0; i < 5; ++i) em[(1 + i) * K + k] = st_mean[i];
#pragma unroll
        for (int i = 0; i < 25; ++i) em[(6 + i) * K + k] = st_cov[i];
        if (k == 0) {
            scal[0] = __float_as_uint(total_weight);
            scal[1] = (uint32_t)iterations_run;
            scal[2] = (uint32_t)samples_seen;
        }
    }

    // ---- prepare_vectorized: Categorical::prepare over the cell, then per-component Cholesky ---
    // (all __syncthreads() below are at the top level: every thread of the CTA reaches them)
    const bool renorm = active && (do_update || (a.mode & 4));
    s_pmf[threadIdx.x] = pmf;
    __syncthreads();
    if (renorm && k == 0) s_tot[local] = categorical_prefix_seq(s_pmf + local * K, s_cdf + local * K, K);
    __syncthreads();
    if (renorm) {
        const bool ok = categorical_scale_own(&s_pmf[threadIdx.x], &s_cdf[threadIdx.x], s_tot[local]);
        if (k == 0) s_flag[local] = ok ? 0 : 1;
    }
    __syncthreads();
    ComponentPrep cp;
    if (active) {
        if (renorm) {
            pmf = s_pmf[threadIdx.x];
            if (s_flag[local] & 1) status |= ST_CDF_INVALID;
            a.cdf[(size_t)cell * K + k] = s_cdf[threadIdx.x];
        }
        prepare_component(mean, cov, cp);
        if (!cp.psd) atomicAnd(&s_psd[local], 0);
        if (renorm) {
            joint[k] = pmf;
#pragma unroll
            for (int d = 0; d < 6; ++d) joint[(1 + d) * K + k] = mean[d];
#pragma unroll
            for (int i = 0; i < 25; ++i) joint[(7 + i) * K + k] = cov[i];
        }
        write_erec(a.erec + (size_t)cell * EREC_PLANES * K, K, k, pmf, mean, cp);
    }
    if (a.mode & 2) {  // uniform over the CTA
        // sdmm::prepare(conditioner, joint): the marginal copies the joint's (already normalised)
        // weights and runs Categorical::prepare on them again (sdmm_conditioner.h:98)
        __syncthreads();
        s_pmf[threadIdx.x] = pmf;
        __syncthreads();
        if (active && k == 0) s_tot[local] = categorical_prefix_seq(s_pmf + local * K, s_cdf + local * K, K);
        __syncthreads();
        if (active) {
            const bool ok = categorical_scale_own(&s_pmf[threadIdx.x], &s_cdf[threadIdx.x], s_tot[local]);
            if (!ok && k == 0) s_flag[local] |= 2;
        }
        __syncthreads();
        if (active) {
            ConditionerOut co;
            conditioner_component(cov, co);
            write_qrec(a.qrec + ((size_t)cell * K + k) * QREC_FLOATS, s_pmf[threadIdx.x], mean, cp.to, co);
            if (s_flag[local] & 2) status |= ST_COND_INVALID;
        }
    }
    __syncthreads();
    if (active && k == 0) {
        if (!s_psd[local]) status |= ST_NON_PSD;
        scal[7] = status;
    }
    if ((a.mode & 2) && a.peer.world > 1) peer_push_records(a);  // uniform over the grid
}

// ------------------------------------------------------------------------------------------------
// parity read-outs (tests only; recomputed from the stored model, same code as the product path)
// ------------------------------------------------------------------------------------------------
__global__ void derived_readout_kernel(const float* __restrict__ joint_all, const float* __restrict__ cdf_all,
                                       int first_cell, int n_cells, int K, float* __restrict__ out) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_cells * K) return;
    int ci = idx / K, k = idx % K, cell = first_cell + ci;
    const float* joint = joint_all + (size_t)cell * JOINT_PLANES * K;
    float mean[6], cov[25];
    for (int d = 0; d < 6; ++d) mean[d] = joint[(1 + d) * K + k];
    for (int i = 0; i < 25; ++i) cov[i] = joint[(7 + i) * K + k];
    ComponentPrep cp;
    prepare_component(mean, cov, cp);
    float* o = out + (size_t)ci * 38 * K;
    for (int i = 0; i < 9; ++i) o[i * K + k] = cp.to[i];
    for (int i = 0; i < 25; ++i) o[(9 + i) * K + k] = cp.L[i];
    o[34 * K + k] = cp.inv_det;
    o[35 * K + k] = cdf_all[(size_t)cell * K + k];
    o[36 * K + k] = joint[k];
    o[37 * K + k] = cp.psd ? 1.f : 0.f;
}

__global__ void conditioner_readout_kernel(const float* __restrict__ joint_all, const float* __restrict__ qrec_all,
                                           int first_cell, int n_cells, int K, float* __restrict__ out) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_cells * K) return;
    int ci = idx / K, k = idx % K, cell = first_cell + ci;
    const float* joint = joint_all + (size_t)cell * JOINT_PLANES * K;
    float mean[6], cov[25];
    for (int d = 0; d < 6; ++d) mean[d] = joint[(1 + d) * K + k];
    for (int i = 0; i < 25; ++i) cov[i] = joint[(7 + i) * K + k];
    ConditionerOut co;
    conditioner_component(cov, co);
    float to[9], from[9];
    frame_from_mean_dir(mean + 3, to, from);
    const float* q = qrec_all + ((size_t)cell * K + k) * QREC_FLOATS;
    float* o = out + (size_t)ci * 38 * K;
    for (int d = 0; d < 3; ++d) o[d * K + k] = mean[d];
    for (int i = 0; i < 9; ++i) o[(3 + i) * K + k] = co.La[i];
    o[12 * K + k] = co.inv_det_a;
    // normalised marginal weight recovered from the record: 2^-nlca / (inv_det_a (2 pi)^-3/2)
    o[13 * K + k] = exp2f(-q[QREC_NLCA]) / (co.inv_det_a * gaussian_normalization(3));
    for (int i = 0; i < 6; ++i) o[(14 + i) * K + k] = co.M[i];
    for (int i = 0; i < 9; ++i) o[(20 + i) * K + k] = to[i];
    for (int i = 0; i < 4; ++i) { o[(29 + i) * K + k] = co.Lc[i]; o[(34 + i) * K + k] = co.cc[i]; }
    o[33 * K + k] = co.inv_det_c;
}

// ------------------------------------------------------------------------------------------------
// host-side launchers
// ------------------------------------------------------------------------------------------------
template <int NP, int LPS, int WARPS, int MINB, bool LL, int MODE>
static cudaError_t launch_estep_tl(const EStepLaunch& L, cudaStream_t st) {
    constexpr int K = 2 * NP * LPS;
    const size_t smem = (size_t)WARPS * 32 * 32 + (size_t)WARPS * ESTAT_PLANES * K * sizeof(float);
    auto kern = em_estep_kernel<NP, LPS, WARPS, MINB, LL, MODE>;
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    kern<<<L.n_items, WARPS * 32, smem, st>>>(L.samples, L.items, L.erec, L.partials, L.item_scal, L.out_loglik,
                                             nullptr, (size_t)0, nullptr);
    return cudaGetLastError();
}

template <int NP, int LPS, int WARPS, int MINB, int MODE = 0>
static cudaError_t launch_estep_t(const EStepLaunch& L, cudaStream_t st) {
    return L.out_loglik ? launch_estep_tl<NP, LPS, WARPS, MINB, true, MODE>(L, st)
                        : launch_estep_tl<NP, LPS, WARPS, MINB, false, MODE>(L, st);
}

// borrowed SoA planes: the default shape of each K only
template <int NP, int LPS, int WARPS, int MINB, bool LL>
static cudaError_t launch_estep_soa_l(const EStepLaunch& L, cudaStream_t st) {
    constexpr int K = 2 * NP * LPS;
    const size_t smem = (size_t)WARPS * 32 * 32 + (size_t)WARPS * ESTAT_PLANES * K * sizeof(float);
    auto kern = em_estep_kernel<NP, LPS, WARPS, MINB, LL, 0, true>;
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    kern<<<L.n_items, WARPS * 32, smem, st>>>(nullptr, L.items, L.erec, L.partials, L.item_scal, L.out_loglik,
                                             L.soa_point6, L.soa_stride, L.soa_weight);
    return cudaGetLastError();
}
template <int NP, int LPS, int WARPS, int MINB>
static cudaError_t launch_estep_soa(const EStepLaunch& L, cudaStream_t st) {
    return L.out_loglik ? launch_estep_soa_l<NP, LPS, WARPS, MINB, true>(L, st)
                        : launch_estep_soa_l<NP, LPS, WARPS, MINB, false>(L, st);
}

cudaError_t launch_estep(const EStepLaunch& L, cudaStream_t st) {
    if (L.n_items == 0) return cudaSuccess;
    if (L.soa_point6) {
        switch (L.K) {
            case 16: return launch_estep_soa<1, 8, 4, 3>(L, st);
            case 32: return launch_estep_soa<1, 16, 4, 3>(L, st);
            case 64: return launch_estep_soa<1, 32, 4, 3>(L, st);
            case 128: return launch_estep_soa<2, 32, 4, 2>(L, st);
            default: return cudaErrorInvalidValue;
        }
    }
    switch (L.K) {
        case 16: return launch_estep_t<1, 8, 4, 3>(L, st);
        case 32: return launch_estep_t<1, 16, 4, 3>(L, st);
        case 64: {
            // tuning hook (tools/estep_sweep.py): alternative lane/occupancy shapes of the same kernel
            static const int variant = [] { const char* v = getenv("SDMMB_ESTEP_VARIANT"); return v ? atoi(v) : 0; }();
            switch (variant) {
                case 1: return launch_estep_t<2, 16, 4, 2>(L, st);
                case 2: return launch_estep_t<1, 32, 4, 4>(L, st);
                case 3: return launch_estep_t<1, 32, 8, 1>(L, st);
                case 4: return launch_estep_t<2, 16, 8, 1>(L, st);
                case 5: return launch_estep_t<1, 32, 2, 6>(L, st);
                case 6: return launch_estep_t<1, 32, 4, 3, 1>(L, st);
                case 7: return launch_estep_t<1, 32, 4, 4, 1>(L, st);
                case 8: return launch_estep_t<1, 32, 4, 2>(L, st);   // 2 warps per sub-partition, 255 registers
                case 9: return launch_estep_t<1, 32, 2, 4>(L, st);
                default: return launch_estep_t<1, 32, 4, 3>(L, st);
            }
        }
        case 128: return launch_estep_t<2, 32, 4, 2>(L, st);
        default: return cudaErrorInvalidValue;
    }
}

cudaError_t launch_mstep(const MStepArgs& a, cudaStream_t st) {
    if (a.n_cells == 0) return cudaSuccess;
    const int cells_per_cta = MSTEP_THREADS / a.K;
    const int grid = (a.n_cells + cells_per_cta - 1) / cells_per_cta;
    // occupancy of this latency-bound kernel: MINB resident CTAs per SM (register cap 255 / 168 / 128)
    static const int minb = [] { const char* v = getenv("SDMMB_MSTEP_MINB"); return v ? atoi(v) : 2; }();
    if (minb >= 4) em_mstep_kernel<4><<<grid, MSTEP_THREADS, 0, st>>>(a);
    else if (minb == 3) em_mstep_kernel<3><<<grid, MSTEP_THREADS, 0, st>>>(a);
    else em_mstep_kernel<2><<<grid, MSTEP_THREADS, 0, st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_peer_signal(const PeerPush& p, cudaStream_t st) {
    peer_signal_kernel<<<1, 32, 0, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_peer_wait(const PeerPush& p, cudaStream_t st) {
    peer_wait_kernel<<<1, 32, 0, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_pack(const float* point6, size_t stride, const float* weight, float4* rec, size_t n,
                        cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    size_t blocks = (n + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    pack_samples_kernel<<<(unsigned)blocks, 256, 0, st>>>(point6, stride, weight, rec, n);
    return cudaGetLastError();
}

cudaError_t launch_derived_readout(const float* joint, const float* cdf, int first_cell, int n_cells, int K,
                                   float* out, cudaStream_t st) {
    int n = n_cells * K;
    if (n == 0) return cudaSuccess;
    derived_readout_kernel<<<(n + 127) / 128, 128, 0, st>>>(joint, cdf, first_cell, n_cells, K, out);
    return cudaGetLastError();
}

cudaError_t launch_conditioner_readout(const float* joint, const float* qrec, int first_cell, int n_cells,
                                       int K, float* out, cudaStream_t st) {
    int n = n_cells * K;
    if (n == 0) return cudaSuccess;
    conditioner_readout_kernel<<<(n + 127) / 128, 128, 0, st>>>(joint, qrec, first_cell, n_cells, K, out);
    return cudaGetLastError();
}

}  // namespace sdmmb
