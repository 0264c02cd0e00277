// estep_kernels.cu -- the N*K hot loop of the SDMM training path on sm_100a:
//   K0 pack_samples_kernel     SoA sample planes -> packed 32-byte records
//   K1 em_estep_kernel<NP,LPS> E-step + sufficient statistics
// (K2/K3, the per-component M-step and the conditioner, live in mstep_kernels.cu.)
//
// Reference for every formula: /root/reference/sdmm/include/sdmm ("R:" below), see DESIGN.md.
//
// K1 design (B200).  The reference runs SIMD over components and sequentially over samples
// (R: opt/em.h:265-312).  K1 keeps that orientation because it makes the 21 sufficient statistics
// per component a pure register accumulation: a lane owns NP *pairs* of components, the two
// components of a pair ride in the two halves of Blackwell's packed FP32 instructions
// (fma.rn.f32x2 / FFMA2), and a group of LPS lanes covers all K = 2*NP*LPS components of one
// sample.  The only cross-lane traffic per sample is the log2(LPS)-step shuffle butterfly for the
// mixture normaliser sum_k pi_k p_k(x_n).
//
// The register file, not the FMA pipe, bounds this loop (profiles/ubench_r2.txt):
// each SM sub-partition reads ONE 32-bit register per cycle from its even bank and one from its odd
// bank, for all pipes together.  A packed FFMA2 with three register sources reads 3 + 3 (three pairs) or
// 3 + 2 (pair, 32-bit broadcast, pair) and issues every 3 cycles instead of every 2; only an operand
// served by the reuse cache (same slot as in the previous instruction), a uniform register or a constant
// is free.  ptxas decides where reuse happens (source order does not survive its scheduler, block
// boundaries do not either: profiles/r2_k1/), so the arithmetic is arranged to READ LESS instead:
//   * positions are NOT centred per component.  With x' = x - c (c = the cell's mixture centre, one
//     value per cell) and the bias b = -W_pos (mu - c) folded into the record,
//         z = W (t_pos, t_dir) = W_pos x' + b + W_dir t_dir,
//     so 12 of the 15 forward-substitution FMAs take x' as a 32-bit broadcast operand (5 register reads
//     instead of 6) and the three subtractions x - mu disappear;
//   * the position block of the statistics is accumulated un-centred against the sample's own
//     products: sum g, sum g x', sum g x'x'^T (products x'x'^T staged once per sample) and
//     sum (g t_dir) x'; only the three direction-direction entries multiply two per-lane pairs.
//     The M-step re-centres them about the component mean in double precision
//     (sum g (x'-m)(x'-m)^T = U2 - m U1^T - U1 m^T + m m^T U0: exact algebra of R: opt/em.h:57-62).
// Per warp pair-step (64 updates): 64 packed instructions (was 70), 329 register reads (was 358) = 165
// cycles of register-file time against 135 of FMA-pipe time; 204 cycles measured (211 before).
// Accuracy: measured identical to the centred form on the test and bench batches (profiles/r2_k1/step_noise.txt);
// sum g x'x'^T carries (|mu - c| / sigma)^2 times the float32 rounding of its running sum, which is why the centre
// is the mixture's own weighted mean position (the component means of a cell lie within the cell).
#include "kernels.h"

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

// One staged sample: four float4 per-warp shared-memory slots, every value a warp- (or group-)
// uniform scalar that enters the packed arithmetic as a 32-bit broadcast operand.
//   q0 = (x', y', z', dx)   q1 = (dy, -dz, w * 2^-64, x'x')   q2 = (x'y', x'z', y'y', y'z')   q3 = (z'z', -, -, -)
constexpr int SLOT_F4 = 4;

struct PairOut {
    float2 p;       // pi_k p_k(x) for the two components of the pair
    float2 t3, t4;  // directional tangent coordinates (t4 carries the frame's sign, see below)
};

// prm: EREC_REG_PLANES packed parameter pairs (kernels.h: EREC_*).  `ndz` is MINUS the sample's direction z.
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
    const float2* W = prm + EREC_W;
    const float2* Bz = prm + EREC_B;
    float2 q = fma2(F[1], dy, mul2(F[0], dx));
    float2 lz = fma2(F[2], ndz, q);
    float2 h = fma2(F[3], q, ndz);
    float2 lx = fma2(F[4], h, dx);
    float2 ly = fma2(F[5], h, dy);
    // position part of z = W_pos x' + b: independent of the acos chain below (the scheduler overlaps them)
    float2 z0 = fma2(W[0], xx, Bz[0]);
    float2 z1 = fma2(W[1], xx, Bz[1]);
    float2 z2 = fma2(W[3], xx, Bz[2]);
    float2 z3 = fma2(W[6], xx, Bz[3]);
    float2 z4 = fma2(W[10], xx, Bz[4]);
    z1 = fma2(W[2], yy, z1);
    z2 = fma2(W[4], yy, z2);
    z3 = fma2(W[7], yy, z3);
    z4 = fma2(W[11], yy, z4);
    z2 = fma2(W[5], zz, z2);
    z3 = fma2(W[8], zz, z3);
    z4 = fma2(W[12], zz, z4);
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
    o.t3 = mul2(lx, r);
    o.t4 = mul2(ly, r);
    // z = W t with W = sqrt(log2(e)/2) * L^-1 (replaces linalg::solve, R: linalg/cholesky.h:38-58)
    z3 = fma2(W[9], o.t3, z3);
    z4 = fma2(W[13], o.t3, z4);
    z4 = fma2(W[14], o.t4, z4);
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

// 21 statistics per component, in the plane order of the partials (kernels.h ESTAT_*):
//   s0 = sum g | u1[i] = sum g x'_i | sd[d] = sum g t_d | cov (i >= j, row-major lower triangle):
//   (i, j < 3) sum g x'_i x'_j;  (3 + d, j < 3) sum g t_d x'_j;  (3 + d, 3 + d') sum g t_d t_d'
struct PairAcc {
    float2 s0;
    float2 s1[5];
    float2 s2[15];
};

// Stats::plus_eq (R: opt/em.h:57-62) on the 15 unique covariance entries, position block un-centred.
// q0..q3: the sample's staged slot (broadcast scalars); sc = w_n / sum_k pi_k p_k.
__device__ __forceinline__ void accumulate_pair(PairAcc& acc, const PairOut& o, float sc, const float4& q0,
                                                float xx, const float4& q2, float zz) {
    const float2 g = mul2(o.p, splat(sc));
    acc.s0 = add2(acc.s0, g);
    acc.s1[0] = fma2(g, splat(q0.x), acc.s1[0]);
    acc.s1[1] = fma2(g, splat(q0.y), acc.s1[1]);
    acc.s1[2] = fma2(g, splat(q0.z), acc.s1[2]);
    acc.s2[0] = fma2(g, splat(xx), acc.s2[0]);     // (0,0)
    acc.s2[1] = fma2(g, splat(q2.x), acc.s2[1]);   // (1,0) x'y'
    acc.s2[2] = fma2(g, splat(q2.z), acc.s2[2]);   // (1,1) y'y'
    acc.s2[3] = fma2(g, splat(q2.y), acc.s2[3]);   // (2,0) x'z'
    acc.s2[4] = fma2(g, splat(q2.w), acc.s2[4]);   // (2,1) y'z'
    acc.s2[5] = fma2(g, splat(zz), acc.s2[5]);     // (2,2)
    const float2 g3 = mul2(g, o.t3), g4 = mul2(g, o.t4);
    acc.s1[3] = add2(acc.s1[3], g3);
    acc.s1[4] = add2(acc.s1[4], g4);
    acc.s2[6] = fma2(g3, splat(q0.x), acc.s2[6]);   // (3,0..2)
    acc.s2[7] = fma2(g3, splat(q0.y), acc.s2[7]);
    acc.s2[8] = fma2(g3, splat(q0.z), acc.s2[8]);
    acc.s2[9] = fma2(g3, o.t3, acc.s2[9]);          // (3,3)
    acc.s2[10] = fma2(g4, splat(q0.x), acc.s2[10]); // (4,0..2)
    acc.s2[11] = fma2(g4, splat(q0.y), acc.s2[11]);
    acc.s2[12] = fma2(g4, splat(q0.z), acc.s2[12]);
    acc.s2[13] = fma2(g4, o.t3, acc.s2[13]);        // (4,3)
    acc.s2[14] = fma2(g4, o.t4, acc.s2[14]);        // (4,4)
}

// One sample-step of a lane: E-step of all its NP component pairs for the sample at smem slot `si`,
// plus the lane-local part of the mixture normaliser.
template <int NP>
struct StepOut {
    PairOut o[NP];
    float sum;
};

template <int NP>
__device__ __forceinline__ StepOut<NP> estep_sample(const float2 (&prm)[NP][EREC_REG_PLANES],
                                                    const float4* __restrict__ wbuf, int si) {
    StepOut<NP> r;
    const float4 A = wbuf[si * SLOT_F4 + 0];
    const float2 B = *reinterpret_cast<const float2*>(wbuf + si * SLOT_F4 + 1);
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

template <int NP, int LPS, int WARPS, int MINB, bool LL, bool SOA>
__global__ void __launch_bounds__(WARPS * 32, MINB)
em_estep_kernel(const float4* __restrict__ samples, const EItem* __restrict__ items,
                const float* __restrict__ erec, float* __restrict__ partials,
                float* __restrict__ item_scal, float* __restrict__ out_loglik,
                const float* __restrict__ soa_p6, size_t soa_stride, const float* __restrict__ soa_w,
                const uint32_t* __restrict__ scal, uint32_t skip_iterations_ge) {
    constexpr int K = 2 * NP * LPS;
    constexpr int SPW = 32 / LPS;      // samples a warp processes per step
    constexpr int STEPS = 32 / SPW;    // steps per 32-sample chunk
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4* sbuf = reinterpret_cast<float4*>(smem_raw);                                  // [WARPS][32][SLOT_F4] float4
    float* red = reinterpret_cast<float*>(smem_raw + WARPS * 32 * SLOT_F4 * 16);         // [WARPS][21][K]
    __shared__ float s_wsum[WARPS];
    __shared__ int s_skip[WARPS];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int grp = lane / LPS, l = lane % LPS;
    const EItem item = items[blockIdx.x];
    // second pass of the integrator's 2|1 iteration rule: cells that are done sit this launch out (CTA-uniform)
    if (skip_iterations_ge != 0u && scal[(size_t)item.cell * SCALAR_WORDS + 1] >= skip_iterations_ge) return;

    // this lane's component parameters -> registers (coalesced over lanes within a plane)
    float2 prm[NP][EREC_REG_PLANES];
    const float* base = erec + (size_t)item.cell * EREC_PLANES * K;
#pragma unroll
    for (int q = 0; q < NP; ++q)
#pragma unroll
        for (int pl = 0; pl < EREC_REG_PLANES; ++pl) {
            prm[q][pl].x = base[pl * K + q * 2 * LPS + l];
            prm[q][pl].y = base[pl * K + q * 2 * LPS + LPS + l];
        }
    // the cell's centre: positions are staged relative to it
    const float cx = base[(EREC_CEN + 0) * K], cy = base[(EREC_CEN + 1) * K], cz = base[(EREC_CEN + 2) * K];
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

    float4* wbuf = sbuf + warp * 32 * SLOT_F4;
    const uint32_t n_chunks = (item.end - item.begin + 31u) / 32u;
    // CTA-uniform trip count (every warp runs the same number of rounds; a warp whose chunk lies
    // beyond the item processes 32 zero-weight padding samples).  Keeping the loop bound free of
    // the warp index lets the compiler prove the warp converged at the shuffles below, so the
    // whole round stays one basic block that the scheduler can interleave.
    const uint32_t n_rounds = (n_chunks + WARPS - 1) / WARPS;
    // register prefetch of this lane's sample of the warp's first chunk; rb.w = effective weight.
    // Padding samples sit at the cell centre with direction +z and weight 0.
    float4 ra = make_float4(cx, cy, cz, 0.f), rb = make_float4(0.f, 0.f, 1.f, 0.f);
    {
        uint32_t idx = item.begin + warp * 32u + lane;
        if (idx < item.end) load_sample<SOA>(samples, soa_p6, soa_stride, soa_w, idx, ra, rb);
    }
    for (uint32_t round = 0; round < n_rounds; ++round) {
        const uint32_t cbase = item.begin + (round * WARPS + warp) * 32u;
        __syncwarp();
        {   // stage this lane's sample: centred position, its six products, direction, scaled weight
            const float x = ra.x - cx, y = ra.y - cy, z = ra.z - cz;
            wbuf[lane * SLOT_F4 + 0] = make_float4(x, y, z, rb.x);
            wbuf[lane * SLOT_F4 + 1] = make_float4(rb.y, -rb.z, rb.w * SDMMB_W_SCALE, x * x);
            wbuf[lane * SLOT_F4 + 2] = make_float4(x * y, x * z, y * y, y * z);
            wbuf[lane * SLOT_F4 + 3] = make_float4(z * z, 0.f, 0.f, 0.f);
        }
        wsum += ra.w;
        n_wvalid += (rb.w > 0.f) ? 1 : 0;
        __syncwarp();
        {   // prefetch the next chunk while this one is processed
            uint32_t nidx = cbase + WARPS * 32u + lane;
            ra = make_float4(cx, cy, cz, 0.f);
            rb = make_float4(0.f, 0.f, 1.f, 0.f);
            if (nidx < item.end) load_sample<SOA>(samples, soa_p6, soa_stride, soa_w, nidx, ra, rb);
        }
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
            const float4 q0 = wbuf[si * SLOT_F4 + 0];
            const float2 q1 = *reinterpret_cast<const float2*>(reinterpret_cast<const float*>(wbuf + si * SLOT_F4 + 1) + 2);
            const float4 q2 = wbuf[si * SLOT_F4 + 2];
            const float zz = reinterpret_cast<const float*>(wbuf + si * SLOT_F4 + 3)[0];
            const float w = q1.x;
            // hsum; skip if 0 / non-finite                           R: opt/em.h:288-293
            const float rs = rcp_approx(sum);
            const bool ok = (rs > 0.f) && (rs <= 3.402823466e+38f);   // sum finite, > 0, 1/sum finite
            const float sc = ok ? w * rs : 0.f;                       // w_n / sum   R: em.h:295-298, 309
            n_ok += (ok && w > 0.f) ? 1 : 0;
            if (LL) {
                if (l == 0 && cbase + si < item.end) out_loglik[cbase + si] = logf(sum);
            }
#pragma unroll
            for (int q = 0; q < NP; ++q) {
                // a NaN/inf normaliser must not leak 0 * inf into the statistics
                PairOut o = cur.o[q];
                o.p.x = ok ? o.p.x : 0.f;
                o.p.y = ok ? o.p.y : 0.f;
                accumulate_pair(acc[q], o, sc, q0, q1.y, q2, zz);
            }
            cur = nxt;
        }
    }
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
// host-side launchers
// ------------------------------------------------------------------------------------------------
template <int NP, int LPS, int WARPS, int MINB, bool LL, bool SOA>
static cudaError_t launch_estep_inst(const EStepLaunch& L, cudaStream_t st) {
    constexpr int K = 2 * NP * LPS;
    const size_t smem = (size_t)WARPS * 32 * SLOT_F4 * 16 + (size_t)WARPS * ESTAT_PLANES * K * sizeof(float);
    auto kern = em_estep_kernel<NP, LPS, WARPS, MINB, LL, SOA>;
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    kern<<<L.n_items, WARPS * 32, smem, st>>>(L.samples, L.items, L.erec, L.partials, L.item_scal, L.out_loglik,
                                             L.soa_point6, L.soa_stride, L.soa_weight, L.scal, L.skip_iterations_ge);
    return cudaGetLastError();
}

template <int NP, int LPS, int WARPS, int MINB>
static cudaError_t launch_estep_shape(const EStepLaunch& L, cudaStream_t st) {
    if (L.soa_point6)
        return L.out_loglik ? launch_estep_inst<NP, LPS, WARPS, MINB, true, true>(L, st)
                            : launch_estep_inst<NP, LPS, WARPS, MINB, false, true>(L, st);
    return L.out_loglik ? launch_estep_inst<NP, LPS, WARPS, MINB, true, false>(L, st)
                        : launch_estep_inst<NP, LPS, WARPS, MINB, false, false>(L, st);
}

// resident E-step CTAs per SM of the shape launch_estep picks for K (the host pipeline cuts its
// chunks at whole waves of this many CTAs)
int estep_ctas_per_sm(int K) { return K == 128 ? 2 : 3; }

cudaError_t launch_estep(const EStepLaunch& L, cudaStream_t st) {
    if (L.n_items == 0) return cudaSuccess;
    // tuning hook (tools/estep_sweep.py): alternative warp/occupancy shapes of the same kernel
    static const int variant = [] { const char* v = getenv("SDMMB_ESTEP_VARIANT"); return v ? atoi(v) : 0; }();
    switch (L.K) {
        case 16: return launch_estep_shape<1, 8, 4, 3>(L, st);
        case 32: return launch_estep_shape<1, 16, 4, 3>(L, st);
        case 64:
            switch (variant) {
                case 1: return launch_estep_shape<1, 32, 4, 2>(L, st);   // 255 registers, 2 warps per sub-partition
                case 2: return launch_estep_shape<1, 32, 4, 4>(L, st);   // 128 registers
                case 3: return launch_estep_shape<1, 32, 8, 1>(L, st);
                case 4: return launch_estep_shape<1, 32, 2, 6>(L, st);
                case 5: return launch_estep_shape<2, 16, 4, 2>(L, st);   // two pairs per lane, two samples per step
                default: return launch_estep_shape<1, 32, 4, 3>(L, st);
            }
        case 128:
            switch (variant) {
                case 1: return launch_estep_shape<2, 32, 8, 1>(L, st);
                case 2: return launch_estep_shape<2, 32, 2, 4>(L, st);   // 8 warps per SM in 4 CTAs
                default: return launch_estep_shape<2, 32, 4, 2>(L, st);
            }
        default: return cudaErrorInvalidValue;
    }
}

cudaError_t launch_pack(const float* point6, size_t stride, const float* weight, float4* rec, size_t n,
                        cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    size_t blocks = (n + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    pack_samples_kernel<<<(unsigned)blocks, 256, 0, st>>>(point6, stride, weight, rec, n);
    return cudaGetLastError();
}

}  // namespace sdmmb
