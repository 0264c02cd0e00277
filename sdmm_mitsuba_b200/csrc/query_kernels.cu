// query_kernels.cu -- K4: fused guided query (condition on position -> CDF select -> sample -> pdf)
// for one path vertex per thread.  Replaces, per vertex, the reference's
//   sdmm::create_conditional   R: distributions/sdmm_conditioner.h:102-115, 232-248
//   SDMM::sample (D = 2)       R: distributions/sdmm.h:209-273 (+ categorical.h, offset_directional.h:62-82)
//   SDMM::posterior + hsum     R: distributions/sdmm.h:331-339, offset_directional.h:39-60
// as called from SDMMRenderer::sampleSurface / pdfSurface (mitsuba/src/integrators/sdmm/sdmm_proc.cpp:
// 361-372, 419-421, 491-501, 539-545).
//
// Mapping: thread <-> a PAIR of vertices of the same cell (the two halves of the packed FP32
// instructions), sequential loop over the K components so that the conditional CDF is the same
// left-to-right float32 prefix sum the reference builds (categorical.h:92-111) -- the component chosen
// for a given uniform then only depends on the pmf bits.  The cell's query records (K x 128 B) are
// staged in shared memory once per run of vertices of that cell (all lanes read the same record ->
// broadcast LDS); vertices that arrive unsorted are binned on the device first (counting sort).  Three
// passes per vertex pair (sum -> select -> pdf).  For K <= 64 the K weights of the pair are parked in shared
// memory between the passes (QSTORE: 12 warps per SM at K = 64, but pass 3 sheds a quarter of its
// arithmetic); for K = 128 they are recomputed in each pass (QRECOMP: no per-thread storage, 32 warps per SM).
// Either way every pass sees the same weight bits.
#include <algorithm>
#include <cstdlib>

#include "kernels.h"
#include "sdmm_math.cuh"

namespace sdmmb {

namespace {

constexpr int QTHREADS = 128;
constexpr int SDMMB_QUERY_MAX_K = 128;

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

// the same for the two vertices of a pair: the polynomial and the final product ride in packed instructions
// (each half is the IEEE fma sequence of the scalar form above: bit-identical results)
__device__ __forceinline__ float2 rcp_sinc_from_cos2(float2 c) {
    const float2 a = make_float2(fminf(fabsf(c.x), 0.99999994f), fminf(fabsf(c.y), 0.99999994f));
    float2 P = make_float2(2.295893482e-03f, 2.295893482e-03f);
    P = __ffma2_rn(P, a, make_float2(-1.114577623e-02f, -1.114577623e-02f));
    P = __ffma2_rn(P, a, make_float2(2.689914581e-02f, 2.689914581e-02f));
    P = __ffma2_rn(P, a, make_float2(-4.880191494e-02f, -4.880191494e-02f));
    P = __ffma2_rn(P, a, make_float2(8.875543537e-02f, 8.875543537e-02f));
    P = __ffma2_rn(P, a, make_float2(-2.145852374e-01f, -2.145852374e-01f));
    P = __ffma2_rn(P, a, make_float2(1.570796172e+00f, 1.570796172e+00f));
    const float2 opa = __fadd2_rn(a, make_float2(1.f, 1.f));
    const float2 oma = __fadd2_rn(make_float2(1.f, 1.f), make_float2(-a.x, -a.y));
    const float2 u = make_float2(q_rsqrt(opa.x), q_rsqrt(opa.y));
    const float2 v = make_float2(q_rsqrt(oma.x), q_rsqrt(oma.y));
    const float2 alt = __ffma2_rn(make_float2(kPi, kPi), v, make_float2(-P.x, -P.y));
    const float2 num = make_float2(c.x >= 0.f ? P.x : alt.x, c.y >= 0.f ? P.y : alt.y);
    return __fmul2_rn(num, u);
}

struct QRec {
    float4 v[8];
    __device__ __forceinline__ float f(int i) const { return reinterpret_cast<const float*>(v)[i]; }
};

// Where record k starts (in float4) in a record array.  SW = false: the conditioner's own layout in global memory,
// 8 float4 per record.  SW = true: the shared-memory copy, skewed by one float4 per record and one more per block
// of QBLOCK records (k * 9 + k / 8): in the uniform loops every lane reads the same address (broadcast), but in the
// per-vertex parts (block re-scan of pass 2: lanes at the same j of DIFFERENT blocks; Box-Muller of the selected
// component: arbitrary k) lanes read different records, and with a 128-byte stride those all sit in the same
// 16-byte bank group -- a 32-way conflict per LDS.128 that made shared-memory wavefronts the busiest unit of the
// kernel (l1tex 67 %).  With the skew, bank group = (block + j + i) mod 8.
template <bool SW>
__device__ __forceinline__ int qoff(int k) { return SW ? k * 9 + (k >> 3) : k * 8; }

template <bool SW>
__device__ __forceinline__ QRec load_rec(const float4* __restrict__ base, int k) {
    QRec r;
    const float4* p = base + qoff<SW>(k);
#pragma unroll
    for (int i = 0; i < 8; ++i) r.v[i] = p[i];
    return r;
}

// ------------------------------------------------------------------------------------------------
// Two vertices of the same cell per thread, riding in the two halves of the packed FP32
// instructions (FFMA2/FMUL2/FADD2); the cell's record values enter as 32-bit broadcast operands, so
// one shared-memory read of a record serves both vertices.  Every half follows exactly the
// reference's operation order (left-to-right CDF sums), so results do not depend on its partner.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float2 q_splat(float v) { return make_float2(v, v); }
__device__ __forceinline__ float2 q_fma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float2 q_mul2(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ float2 q_add2(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 q_sub2(float2 a, float2 b) { return __fadd2_rn(a, make_float2(-b.x, -b.y)); }

struct PairIn {
    float2 x, y, z;      // positions of the two vertices
    float2 u0, u1, u2;   // uniforms (sample mode)
    float2 wx, wy, wz;   // directions (pdf-only mode: input; sample mode: output)
};
struct PairRes {
    float2 pdf, ij;
    uint32_t comp[2];
    uint8_t valid[2];
};

// unnormalised conditional weight pi_k N3(x; mu_a, Sigma_aa) of component k for both vertices
// (R: sdmm_conditioner.h:242-244).  12 packed instructions + 2 MUFU per component.  Where a pass recomputes it
// (QRECOMP), the same instruction sequence gives bit-identical weights every time.
__device__ __forceinline__ float2 pair_weight(const float4* __restrict__ rec, int k, const PairIn& in) {
    const float4* rk = rec + qoff<true>(k);
    const float4 r0 = rk[0], r1 = rk[1], r2 = rk[2];
    // r0 = (mu0, mu1, mu2, wa00), r1 = (wa10, wa11, wa20, wa21), r2 = (wa22, nlca, m00, m01)
    const float2 d0 = q_sub2(in.x, q_splat(r0.x)), d1 = q_sub2(in.y, q_splat(r0.y)), d2 = q_sub2(in.z, q_splat(r0.z));
    const float2 z0 = q_mul2(q_splat(r0.w), d0);
    const float2 z1 = q_fma2(q_splat(r1.y), d1, q_mul2(q_splat(r1.x), d0));
    const float2 z2 = q_fma2(q_splat(r2.x), d2, q_fma2(q_splat(r1.w), d1, q_mul2(q_splat(r1.z), d0)));
    float2 e = q_fma2(z0, z0, q_splat(r2.y));
    e = q_fma2(z1, z1, e);
    e = q_fma2(z2, z2, e);
    return make_float2(q_ex2(-e.x), q_ex2(-e.y));
}

// The same weight for ONE vertex with scalar instructions (each half of a packed FFMA2 is an IEEE fma,
// so this is bit-identical to the corresponding half of pair_weight).
template <bool SW>
__device__ __forceinline__ float single_weight(const float4* __restrict__ rec, int k, float x, float y, float z) {
    const float4* rk = rec + qoff<SW>(k);
    const float4 r0 = rk[0], r1 = rk[1], r2 = rk[2];
    const float d0 = x - r0.x, d1 = y - r0.y, d2 = z - r0.z;
    const float z0 = r0.w * d0;
    const float z1 = fmaf(r1.y, d1, r1.x * d0);
    const float z2 = fmaf(r2.x, d2, fmaf(r1.w, d1, r1.z * d0));
    float e = fmaf(z0, z0, r2.y);
    e = fmaf(z1, z1, e);
    e = fmaf(z2, z2, e);
    return q_ex2(-e);
}

// First k with cdf[k] = run_k * inv >= u0 for one vertex, given the running sums at the end of every
// block of QBLOCK components (`pref`, stride `ps`): find the block, then redo the sequential sum inside
// it -- the association order, hence every cdf bit, is that of the full left-to-right scan
// (== enoki::binary_search on the monotone cdf clamped to K-1; R: sdmm.h:220-233).
constexpr int QBLOCK = 8;
template <int K, int H>
__device__ __forceinline__ int select_component(const float4* __restrict__ rec, const float2* __restrict__ pref,
                                                int ps, float x, float y, float z, float u0, float inv) {
    int b = K / QBLOCK;
#pragma unroll
    for (int j = K / QBLOCK - 1; j >= 0; --j) {
        const float2 p = pref[j * ps];
        if (!((H ? p.y : p.x) * inv < u0)) b = j;
    }
    int sel = K - 1;
    if (b < K / QBLOCK) {
        float run = 0.f;
        if (b > 0) { const float2 p = pref[(b - 1) * ps]; run = H ? p.y : p.x; }
        bool found = false;
#pragma unroll
        for (int j = 0; j < QBLOCK; ++j) {
            const int k = b * QBLOCK + j;
            const float w = single_weight<true>(rec, k, x, y, z);
            run = (k == 0) ? w : run + w;
            if (!found && !(run * inv < u0)) { sel = k; found = true; }
        }
    }
    // back off over zero-pmf components
    while (sel > 0 && single_weight<true>(rec, sel, x, y, z) * inv == 0.f) --sel;
    return sel;
}

// The same selection when pass 1 parked the K weights of the pair in shared memory (`wst`, stride `ps`) and kept the
// block-end running sums in registers: the block re-scan reads the stored bits instead of recomputing them.
template <int K, int H>
__device__ __forceinline__ int select_stored(const float2* __restrict__ wst, int ps,
                                             const float2 (&bsum)[K / QBLOCK], float u0, float inv) {
    const float* w1 = reinterpret_cast<const float*>(wst) + H;
    int b = K / QBLOCK;
    float run = 0.f;   // running sum at the end of block b - 1 (static indices only: bsum stays in registers)
#pragma unroll
    for (int j = K / QBLOCK - 1; j >= 0; --j)
        if (!((H ? bsum[j].y : bsum[j].x) * inv < u0)) {
            b = j;
            run = (j > 0) ? (H ? bsum[j > 0 ? j - 1 : 0].y : bsum[j > 0 ? j - 1 : 0].x) : 0.f;
        }
    int sel = K - 1;
    if (b < K / QBLOCK) {
        bool found = false;
#pragma unroll
        for (int j = 0; j < QBLOCK; ++j) {
            const int k = b * QBLOCK + j;
            const float w = w1[(size_t)k * ps * 2];
            run = (k == 0) ? w : run + w;
            if (!found && !(run * inv < u0)) { sel = k; found = true; }
        }
    }
    while (sel > 0 && w1[(size_t)sel * ps * 2] * inv == 0.f) --sel;
    return sel;
}

// Box-Muller + `L_c z` + offset exp map of component `sel` for one vertex        R: sdmm.h:238-265,
// spaces/offset_directional.h:62-82
template <bool SW>
__device__ __forceinline__ void sample_direction(const float4* __restrict__ rec, int sel, float x, float y, float z,
                                                 float u1, float u2, float& wx, float& wy, float& wz, float& ij) {
    float radius = sqrtf(-2.f * logf(1.f - u1));
    float theta = 2.f * kPi * u2;
    float sn, cs;
    sincosf(theta, &sn, &cs);
    float g0 = sn * radius, g1 = cs * radius;
    const QRec r = load_rec<SW>(rec, sel);
    float dx0 = x - r.f(QREC_MU + 0), dx1 = y - r.f(QREC_MU + 1), dx2 = z - r.f(QREC_MU + 2);
    float tm0 = r.f(QREC_M + 0) * dx0 + r.f(QREC_M + 1) * dx1 + r.f(QREC_M + 2) * dx2;
    float tm1 = r.f(QREC_M + 3) * dx0 + r.f(QREC_M + 4) * dx1 + r.f(QREC_M + 5) * dx2;
    float t0 = r.f(QREC_LC + 0) * g0;
    float t1 = r.f(QREC_LC + 1) * g0 + r.f(QREC_LC + 2) * g1;
    const float max_length = kPi - 1e-1f;
    float ox = t0 + tm0, oy = t1 + tm1;
    float length = sqrtf(ox * ox + oy * oy);
    float sl, cl;
    sincosf(length, &sl, &cl);
    float sinc = ((sl < 1e-2f) || length > max_length) ? 1.f : sl / length;
    ij = (length > max_length) ? 0.f : sinc;
    float l0 = ox * sinc, l1 = oy * sinc, l2 = cl;
    wx = r.f(QREC_R + 0) * l0 + r.f(QREC_R + 3) * l1 + r.f(QREC_R + 6) * l2;
    wy = r.f(QREC_R + 1) * l0 + r.f(QREC_R + 4) * l1 + r.f(QREC_R + 7) * l2;
    wz = r.f(QREC_R + 2) * l0 + r.f(QREC_R + 5) * l1 + r.f(QREC_R + 8) * l2;
}

// conditional pdf of component k at direction w for ONE vertex: the scalar twin of pass 3 of query_pair (each half
// of a packed FFMA2 is an IEEE fma, so this reproduces that half bit for bit)          R: sdmm.h:284-313, 331-339
template <bool SW>
__device__ __forceinline__ float cond_pdf_single(const float4* __restrict__ rec, int k, float x, float y, float z,
                                                 float wx, float wy, float wz) {
    const QRec r = load_rec<SW>(rec, k);
    const float dx0 = x - r.f(QREC_MU + 0), dx1 = y - r.f(QREC_MU + 1), dx2 = z - r.f(QREC_MU + 2);
    const float tm0 = fmaf(r.f(QREC_M + 2), dx2, fmaf(r.f(QREC_M + 1), dx1, r.f(QREC_M + 0) * dx0));
    const float tm1 = fmaf(r.f(QREC_M + 5), dx2, fmaf(r.f(QREC_M + 4), dx1, r.f(QREC_M + 3) * dx0));
    const float lx = fmaf(r.f(QREC_R + 2), wz, fmaf(r.f(QREC_R + 1), wy, r.f(QREC_R + 0) * wx));
    const float ly = fmaf(r.f(QREC_R + 5), wz, fmaf(r.f(QREC_R + 4), wy, r.f(QREC_R + 3) * wx));
    const float lz = fmaf(r.f(QREC_R + 8), wz, fmaf(r.f(QREC_R + 7), wy, r.f(QREC_R + 6) * wx));
    const float rs = rcp_sinc_from_cos(lz);
    const float ij = (lz < -0.98f) ? 0.f : rs;
    const float t0 = fmaf(lx, rs, -tm0), t1 = fmaf(ly, rs, -tm1);
    const float z0 = r.f(QREC_WC + 0) * t0;
    const float z1 = fmaf(r.f(QREC_WC + 2), t1, r.f(QREC_WC + 1) * t0);
    float e = fmaf(z0, z0, r.f(QREC_NLCC));
    e = fmaf(z1, z1, e);
    return ij * q_ex2(-e);
}

// conditional density of component k at the pair's directions, times the log-map Jacobian (R: sdmm.h:284-313,
// offset_directional.h:39-60)
__device__ __forceinline__ float2 pair_cond_pdf(const float4* __restrict__ rec, int k, const PairIn& in) {
    const QRec r = load_rec<true>(rec, k);
    const float2 dx0 = q_sub2(in.x, q_splat(r.f(QREC_MU + 0))), dx1 = q_sub2(in.y, q_splat(r.f(QREC_MU + 1))),
                 dx2 = q_sub2(in.z, q_splat(r.f(QREC_MU + 2)));
    const float2 tm0 = q_fma2(q_splat(r.f(QREC_M + 2)), dx2, q_fma2(q_splat(r.f(QREC_M + 1)), dx1, q_mul2(q_splat(r.f(QREC_M + 0)), dx0)));
    const float2 tm1 = q_fma2(q_splat(r.f(QREC_M + 5)), dx2, q_fma2(q_splat(r.f(QREC_M + 4)), dx1, q_mul2(q_splat(r.f(QREC_M + 3)), dx0)));
    const float2 lx = q_fma2(q_splat(r.f(QREC_R + 2)), in.wz, q_fma2(q_splat(r.f(QREC_R + 1)), in.wy, q_mul2(q_splat(r.f(QREC_R + 0)), in.wx)));
    const float2 ly = q_fma2(q_splat(r.f(QREC_R + 5)), in.wz, q_fma2(q_splat(r.f(QREC_R + 4)), in.wy, q_mul2(q_splat(r.f(QREC_R + 3)), in.wx)));
    const float2 lz = q_fma2(q_splat(r.f(QREC_R + 8)), in.wz, q_fma2(q_splat(r.f(QREC_R + 7)), in.wy, q_mul2(q_splat(r.f(QREC_R + 6)), in.wx)));
    const float2 rs = rcp_sinc_from_cos2(lz);
    const float2 ij = make_float2((lz.x < -0.98f) ? 0.f : rs.x, (lz.y < -0.98f) ? 0.f : rs.y);
    const float2 t0 = q_fma2(lx, rs, make_float2(-tm0.x, -tm0.y)), t1 = q_fma2(ly, rs, make_float2(-tm1.x, -tm1.y));
    const float2 z0 = q_mul2(q_splat(r.f(QREC_WC + 0)), t0);
    const float2 z1 = q_fma2(q_splat(r.f(QREC_WC + 2)), t1, q_mul2(q_splat(r.f(QREC_WC + 1)), t0));
    float2 e = q_fma2(z0, z0, q_splat(r.f(QREC_NLCC)));
    e = q_fma2(z1, z1, e);
    return q_mul2(ij, make_float2(q_ex2(-e.x), q_ex2(-e.y)));
}

// MODE of the sample query: where the K conditional weights of a vertex pair live between the passes.
//   QSTORE   (K <= 64): pass 1 parks them in shared memory (K x 8 B per thread; 3 CTAs = 12 warps per SM at K = 64),
//            block-end running sums stay in registers; pass 2 and pass 3 read the stored bits.  Saves the 12 packed
//            instructions per component that recomputing the weight costs in pass 3, and the block re-scan of pass 2.
//   QRECOMP  (K = 128): nothing is parked but the block-end sums (`wst` = [K/8][threads]); every pass recomputes the
//            weights with the same instruction sequence (bit-identical), full occupancy.
// Both give the same bits: the selection depends only on the weight bits and the sequential float32 sums.
enum { QRECOMP = 0, QSTORE = 1 };

template <bool SAMPLE, int K, int MODE>
__device__ __forceinline__ void query_pair(const float4* __restrict__ rec, float2* __restrict__ wst, int ps,
                                           PairIn& in, PairRes& out) {
    float2 total = q_splat(0.f);
    float2 pdf = q_splat(0.f);
    if (!SAMPLE) {
        // pdf-only: the direction is known up front, so one pass gives sum_k w_k and sum_k p_k w_k; the division by
        // the total moves behind the sum (the reference normalises each weight first: a float32 rounding apart)
#pragma unroll 2
        for (int k = 0; k < K; ++k) {
            const float2 wk = pair_weight(rec, k, in);
            total = (k == 0) ? wk : q_add2(total, wk);
            pdf = q_fma2(pair_cond_pdf(rec, k, in), wk, pdf);
        }
    } else if (MODE == QSTORE) {
        // a block of QBLOCK weights is computed before any of it is stored: a store to `wst` between the record
        // loads would pin their order (the compiler cannot tell the two shared-memory arrays apart), and every
        // component would pay the full LDS and MUFU latency on its own
        float2 bsum[K / QBLOCK];
#pragma unroll
        for (int b = 0; b < K / QBLOCK; ++b) {
            float2 w[QBLOCK];
#pragma unroll
            for (int j = 0; j < QBLOCK; ++j) w[j] = pair_weight(rec, b * QBLOCK + j, in);
#pragma unroll
            for (int j = 0; j < QBLOCK; ++j) {
                total = (b == 0 && j == 0) ? w[j] : q_add2(total, w[j]);
                wst[(b * QBLOCK + j) * ps] = w[j];
            }
            bsum[b] = total;
        }
        const float2 inv = make_float2(1.f / total.x, 1.f / total.y);
        const int sel0 = select_stored<K, 0>(wst, ps, bsum, in.u0.x, inv.x);
        const int sel1 = select_stored<K, 1>(wst, ps, bsum, in.u0.y, inv.y);
        out.comp[0] = (uint32_t)sel0;
        out.comp[1] = (uint32_t)sel1;
        sample_direction<true>(rec, sel0, in.x.x, in.y.x, in.z.x, in.u1.x, in.u2.x, in.wx.x, in.wy.x, in.wz.x, out.ij.x);
        sample_direction<true>(rec, sel1, in.x.y, in.y.y, in.z.y, in.u1.y, in.u2.y, in.wx.y, in.wy.y, in.wz.y, out.ij.y);
#pragma unroll 4
        for (int k = 0; k < K; ++k) pdf = q_fma2(pair_cond_pdf(rec, k, in), q_mul2(wst[k * ps], inv), pdf);
    } else {
        // ---- pass 1: sequential sum of the unnormalised conditional weights -----------------------
#pragma unroll 1
        for (int b = 0; b < K / QBLOCK; ++b) {
#pragma unroll
            for (int j = 0; j < QBLOCK; ++j) {
                const float2 w = pair_weight(rec, b * QBLOCK + j, in);
                total = (b == 0 && j == 0) ? w : q_add2(total, w);
            }
            wst[b * ps] = total;   // running sums at the block ends, for pass 2
        }
        const float2 inv = make_float2(1.f / total.x, 1.f / total.y);
        // ---- pass 2: component select from the block prefix sums, Box-Muller + exp map per vertex ----
        const int sel0 = select_component<K, 0>(rec, wst, ps, in.x.x, in.y.x, in.z.x, in.u0.x, inv.x);
        const int sel1 = select_component<K, 1>(rec, wst, ps, in.x.y, in.y.y, in.z.y, in.u0.y, inv.y);
        out.comp[0] = (uint32_t)sel0;
        out.comp[1] = (uint32_t)sel1;
        sample_direction<true>(rec, sel0, in.x.x, in.y.x, in.z.x, in.u1.x, in.u2.x, in.wx.x, in.wy.x, in.wz.x, out.ij.x);
        sample_direction<true>(rec, sel1, in.x.y, in.y.y, in.z.y, in.u1.y, in.u2.y, in.wx.y, in.wy.y, in.wz.y, out.ij.y);
        // ---- pass 3: conditional pdf at omega, summed over components -----------------------------
#pragma unroll 2
        for (int k = 0; k < K; ++k)
            pdf = q_fma2(pair_cond_pdf(rec, k, in), q_mul2(pair_weight(rec, k, in), inv), pdf);
    }
    const bool v0 = total.x > 1e-20f, v1 = total.y > 1e-20f;      // R: categorical.h:97-101
    out.valid[0] = v0 ? 1 : 0;
    out.valid[1] = v1 ? 1 : 0;
    if (!SAMPLE) {
        pdf = q_mul2(pdf, make_float2(1.f / total.x, 1.f / total.y));
        out.comp[0] = out.comp[1] = 0u;
        out.ij = q_splat(0.f);
    }
    out.pdf = pdf;
    if (!v0) { out.pdf.x = 0.f; out.ij.x = 0.f; out.comp[0] = 0u; if (SAMPLE) { in.wx.x = in.wy.x = in.wz.x = 0.f; } }
    if (!v1) { out.pdf.y = 0.f; out.ij.y = 0.f; out.comp[1] = 0u; if (SAMPLE) { in.wx.y = in.wy.y = in.wz.y = 0.f; } }
}

// ------------------------------------------------------------------------------------------------
// Binning on the device for vertices that arrive in arbitrary order (the renderer's case): counting
// sort by cell -> sorted 32-byte input records + the inverse permutation.  Cell indices
// outside the context go to one extra trailing bucket.  The order inside a bucket is arbitrary
// (atomic cursor); every vertex is independent, so results do not depend on it.
// ------------------------------------------------------------------------------------------------
__global__ void bin_count_kernel(const uint32_t* __restrict__ cell_idx, size_t n, uint32_t n_buckets,
                                 uint32_t* __restrict__ count) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t step = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += step) {
        uint32_t c = cell_idx[i];
        atomicAdd(&count[c < n_buckets - 1 ? c : n_buckets - 1], 1u);
    }
}

// exclusive scan of count[0..n_buckets) into cursor[] by one CTA of 1024 threads
__global__ void __launch_bounds__(1024) bin_scan_kernel(const uint32_t* __restrict__ count, uint32_t n_buckets,
                                                        uint32_t* __restrict__ cursor) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n_buckets; base += 1024) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < n_buckets ? count[i] : 0u;
        uint32_t x = v;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, x, off);
            if ((threadIdx.x & 31) >= off) x += y;
        }
        if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = x;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = s_warp[threadIdx.x];
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                uint32_t y = __shfl_up_sync(0xffffffffu, w, off);
                if (threadIdx.x >= off) w += y;
            }
            s_warp[threadIdx.x] = w;
        }
        __syncthreads();
        const uint32_t warp_excl = (threadIdx.x >> 5) ? s_warp[(threadIdx.x >> 5) - 1] : 0u;
        const uint32_t carry = s_carry;
        if (i < n_buckets) cursor[i] = carry + warp_excl + x - v;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = carry + warp_excl + x;
        __syncthreads();
    }
}

// What the scatter pass moves besides the index: the vertex's inputs as ONE 32-byte record
//   (x, y, z, a0) (a1, a2, original index, cell)      a = uniforms (sample mode) or direction (pdf mode)
// so that random DRAM access happens in whole 32-byte sectors (one scattered write here, one scattered
// read in the un-permute pass) instead of thirteen scattered 4-byte accesses per vertex.
struct BinPayload {
    const float* pos3;   // 3 planes of n
    const float* in3;    // 3 planes of n: u3 or dir3
    float4* qin;         // [n][2] sorted input records
    uint32_t* invperm;   // [n] original index -> sorted position
};

__device__ __forceinline__ void write_sorted_record(const BinPayload& pl, size_t n, size_t i, uint32_t pos, uint32_t c) {
    const float4 a = make_float4(pl.pos3[i], pl.pos3[n + i], pl.pos3[2 * n + i], pl.in3[i]);
    const float4 b = make_float4(pl.in3[n + i], pl.in3[2 * n + i], __uint_as_float((uint32_t)i), __uint_as_float(c));
    pl.qin[2 * (size_t)pos] = a;
    pl.qin[2 * (size_t)pos + 1] = b;
    pl.invperm[i] = pos;
}

__global__ void bin_scatter_kernel(const uint32_t* __restrict__ cell_idx, size_t n, uint32_t n_buckets,
                                   uint32_t* __restrict__ cursor, BinPayload pl) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t step = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += step) {
        const uint32_t c = cell_idx[i];
        const uint32_t pos = atomicAdd(&cursor[c < n_buckets - 1 ? c : n_buckets - 1], 1u);
        write_sorted_record(pl, n, i, pos, c);
    }
}

// Shared-memory variants for bucket counts that fit on chip (2 * n_buckets words): a CTA histograms
// its contiguous chunk of vertices with shared-memory atomics and touches global memory once per
// non-empty bucket, instead of one contended global atomic per vertex.
__global__ void __launch_bounds__(512) bin_count_smem_kernel(const uint32_t* __restrict__ cell_idx, size_t n,
                                                             uint32_t n_buckets, uint32_t* __restrict__ count) {
    extern __shared__ uint32_t sh_bins[];
    for (uint32_t b = threadIdx.x; b < n_buckets; b += blockDim.x) sh_bins[b] = 0;
    __syncthreads();
    const size_t chunk = (n + gridDim.x - 1) / gridDim.x;
    const size_t begin = (size_t)blockIdx.x * chunk, end = begin + chunk < n ? begin + chunk : n;
#pragma unroll 4
    for (size_t i = begin + threadIdx.x; i < end; i += blockDim.x) {
        const uint32_t c = cell_idx[i];
        atomicAdd(&sh_bins[c < n_buckets - 1 ? c : n_buckets - 1], 1u);
    }
    __syncthreads();
    for (uint32_t b = threadIdx.x; b < n_buckets; b += blockDim.x) {
        const uint32_t v = sh_bins[b];
        if (v) atomicAdd(&count[b], v);
    }
}

__global__ void __launch_bounds__(512) bin_scatter_smem_kernel(const uint32_t* __restrict__ cell_idx, size_t n,
                                                               uint32_t n_buckets, uint32_t* __restrict__ cursor,
                                                               BinPayload pl) {
    extern __shared__ uint32_t sh_bins[];
    uint32_t* sh_cnt = sh_bins;               // local count, then local cursor
    uint32_t* sh_base = sh_bins + n_buckets;  // this CTA's reserved range per bucket
    for (uint32_t b = threadIdx.x; b < n_buckets; b += blockDim.x) sh_cnt[b] = 0;
    __syncthreads();
    const size_t chunk = (n + gridDim.x - 1) / gridDim.x;
    const size_t begin = (size_t)blockIdx.x * chunk, end = begin + chunk < n ? begin + chunk : n;
#pragma unroll 4
    for (size_t i = begin + threadIdx.x; i < end; i += blockDim.x) {
        const uint32_t c = cell_idx[i];
        atomicAdd(&sh_cnt[c < n_buckets - 1 ? c : n_buckets - 1], 1u);
    }
    __syncthreads();
    for (uint32_t b = threadIdx.x; b < n_buckets; b += blockDim.x) {
        const uint32_t v = sh_cnt[b];
        sh_base[b] = v ? atomicAdd(&cursor[b], v) : 0u;
        sh_cnt[b] = 0;
    }
    __syncthreads();
    // four vertices per thread and trip: their 28 loads are independent and in flight together (one vertex per trip
    // left the pass latency-bound at a fifth of the HBM rate)
    constexpr int U = 4;
    for (size_t i0 = begin + threadIdx.x; i0 < end; i0 += (size_t)U * blockDim.x) {
        uint32_t c[U];
        float4 a[U];
        float2 b2[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t i = i0 + (size_t)u * blockDim.x;
            if (i < end) {
                c[u] = cell_idx[i];
                a[u] = make_float4(pl.pos3[i], pl.pos3[n + i], pl.pos3[2 * n + i], pl.in3[i]);
                b2[u] = make_float2(pl.in3[n + i], pl.in3[2 * n + i]);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t i = i0 + (size_t)u * blockDim.x;
            if (i < end) {
                const uint32_t b = c[u] < n_buckets - 1 ? c[u] : n_buckets - 1;
                const uint32_t pos = sh_base[b] + atomicAdd(&sh_cnt[b], 1u);
                pl.qin[2 * (size_t)pos] = a[u];
                pl.qin[2 * (size_t)pos + 1] = make_float4(b2[u].x, b2[u].y, __uint_as_float((uint32_t)i), __uint_as_float(c[u]));
                pl.invperm[i] = pos;
            }
        }
    }
}

// A CTA owns a contiguous tile of the cell-sorted vertex order; it walks the cell runs inside the
// tile, staging each cell's records in shared memory once, and handles 2 * QTHREADS vertices of the run
// per round (thread t: sorted positions pos + t and pos + QTHREADS + t).
//   REC = false: the caller's arrays are already sorted by cell: SoA planes in, SoA planes out.
//   REC = true : device-binned vertices: packed 32-byte records in (`qin`, written by the scatter pass)
//                and out (`qout`: (dx, dy, dz, pdf) (inv_jac, comp, valid, -)), both in sorted order,
//                i.e. coalesced; query_unpermute_kernel puts the results back in the caller's order.
// One round's inputs of a thread: (x, y, z, in0) (in1, in2, -, -) for its two vertices -- the device-binned records
// as they are (REC), or gathered from the caller's SoA planes.
struct RoundIn { float4 a0, b0, a1, b1; };
template <bool SAMPLE, bool REC>
__device__ __forceinline__ void load_round(const QueryLaunch& q, const float4* __restrict__ qin, size_t base, size_t rend,
                                           RoundIn& r) {
    const size_t p0 = base + threadIdx.x, p1 = p0 + QTHREADS;
    r.a0 = r.b0 = r.a1 = r.b1 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (REC) {
        if (p0 < rend) { r.a0 = qin[2 * p0]; r.b0 = qin[2 * p0 + 1]; }
        if (p1 < rend) { r.a1 = qin[2 * p1]; r.b1 = qin[2 * p1 + 1]; }
    } else {
        const float* in3 = SAMPLE ? q.u3 : q.dir3;
        if (p0 < rend) {
            r.a0 = make_float4(q.pos3[p0], q.pos3[q.n + p0], q.pos3[2 * q.n + p0], in3[p0]);
            r.b0.x = in3[q.n + p0]; r.b0.y = in3[2 * q.n + p0];
        }
        if (p1 < rend) {
            r.a1 = make_float4(q.pos3[p1], q.pos3[q.n + p1], q.pos3[2 * q.n + p1], in3[p1]);
            r.b1.x = in3[q.n + p1]; r.b1.y = in3[2 * q.n + p1];
        }
    }
}

template <bool SAMPLE, int K>
struct QShape {
    static constexpr int MODE = (SAMPLE && K <= 64) ? QSTORE : QRECOMP;
    static constexpr int REC_F4 = K * 9 + K / 8;                                              // skewed [K][8], see qoff
    static constexpr int W_F2 = !SAMPLE ? 0 : (MODE == QSTORE ? K : K / QBLOCK) * QTHREADS;   // [K or K/8][QTHREADS]
    static constexpr size_t SMEM = (size_t)REC_F4 * sizeof(float4) + (size_t)W_F2 * sizeof(float2);
    // registers: the stored mode holds K/8 block sums and unrolls pass 3 by 4; its CTAs per SM are bounded by shared
    // memory anyway (3 at K = 64, 6 at K = 32)
    static constexpr int MIN_CTAS = MODE == QSTORE ? (K == 64 ? 3 : (K == 32 ? 6 : 8)) : 8;
};

template <bool SAMPLE, int K, bool REC>
__global__ void __launch_bounds__(QTHREADS, (QShape<SAMPLE, K>::MIN_CTAS))
query_pair_kernel(QueryLaunch q, const float4* __restrict__ qin, float4* __restrict__ qout, int tile) {
    using S = QShape<SAMPLE, K>;
    extern __shared__ __align__(16) unsigned char q_smem[];
    float4* srec = reinterpret_cast<float4*>(q_smem);
    float2* swst = reinterpret_cast<float2*>(q_smem + (size_t)S::REC_F4 * sizeof(float4));
    __shared__ uint32_t s_part[QTHREADS / 32];
    const size_t tbegin = (size_t)blockIdx.x * tile;
    const size_t tend = (tbegin + tile < q.n) ? tbegin + tile : q.n;
    size_t pos = tbegin;
    while (pos < tend) {
        // Start of a run of vertices of one cell.  Two global-memory round trips instead of four: every thread reads
        // the run's cell id itself (one address: a broadcast), then the cell's records, the cell ids of the rest of
        // the tile (to find where the run ends) and the inputs of the first round are all requested before anything
        // is waited for.
        const uint32_t cell = REC ? __float_as_uint(__ldg(&qin[2 * pos + 1].w)) : __ldg(q.cell_idx + pos);
        const bool cell_ok = cell < (uint32_t)q.max_cells;
        constexpr int G4 = K * 8 / QTHREADS;
        float4 g4[G4];
        if (cell_ok) {
            const float4* g = reinterpret_cast<const float4*>(q.qrec + (size_t)cell * K * QREC_FLOATS);
#pragma unroll
            for (int i = 0; i < G4; ++i) g4[i] = __ldg(g + threadIdx.x + i * QTHREADS);
        }
        // first position whose cell differs (no early exit: the loads are independent and stay in flight together).
        // Any order of vertices gives correct results -- unsorted input just degenerates into short runs.
        uint32_t first = (uint32_t)(tend - pos);
#pragma unroll 4
        for (size_t p = pos + threadIdx.x; p < tend; p += QTHREADS) {
            const uint32_t c = REC ? __float_as_uint(__ldg(&qin[2 * p + 1].w)) : __ldg(q.cell_idx + p);
            if (c != cell) first = min(first, (uint32_t)(p - pos));
        }
        // the inputs of round r + 1 are requested before round r is computed (12 warps per SM in the stored mode do
        // not hide a global-memory round trip by themselves); round 0's here, masked by the tile end
        RoundIn cur, nxt;
        load_round<SAMPLE, REC>(q, qin, pos, tend, cur);
        first = __reduce_min_sync(0xffffffffu, first);
        __syncthreads();   // the previous run's readers of srec and s_part are done
        if (cell_ok) {
#pragma unroll
            for (int i = 0; i < G4; ++i) {
                const int e = threadIdx.x + i * QTHREADS;
                srec[qoff<true>(e >> 3) + (e & 7)] = g4[i];
            }
        }
        if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = first;
        __syncthreads();
        uint32_t run = s_part[0];
#pragma unroll
        for (int w = 1; w < QTHREADS / 32; ++w) run = min(run, s_part[w]);
        const size_t rend = pos + run;
        for (size_t base = pos; base < rend; base += 2 * QTHREADS) {
            const size_t p0 = base + threadIdx.x, p1 = p0 + QTHREADS;
            const bool m0 = p0 < rend, m1 = p1 < rend;
            if (base + 2 * QTHREADS < rend) load_round<SAMPLE, REC>(q, qin, base + 2 * QTHREADS, rend, nxt);
            const float4 a0 = cur.a0, b0 = cur.b0, a1 = m1 ? cur.a1 : cur.a0, b1 = m1 ? cur.b1 : cur.b0;
            cur = nxt;
            if (m0) {   // a sorted run is contiguous: m1 implies m0
                PairIn in;
                PairRes res;
                res.pdf = res.ij = make_float2(0.f, 0.f);
                res.comp[0] = res.comp[1] = 0u;
                res.valid[0] = res.valid[1] = 0;
                in.wx = in.wy = in.wz = make_float2(0.f, 0.f);
                if (cell_ok) {
                    in.x = make_float2(a0.x, a1.x);
                    in.y = make_float2(a0.y, a1.y);
                    in.z = make_float2(a0.z, a1.z);
                    const float2 c0 = make_float2(a0.w, a1.w), c1 = make_float2(b0.x, b1.x), c2 = make_float2(b0.y, b1.y);
                    if (SAMPLE) { in.u0 = c0; in.u1 = c1; in.u2 = c2; }
                    else { in.wx = c0; in.wy = c1; in.wz = c2; }
                    query_pair<SAMPLE, K, S::MODE>(srec, swst + (SAMPLE ? threadIdx.x : 0), QTHREADS, in, res);
                }
                if (REC) {
                    qout[2 * p0] = make_float4(in.wx.x, in.wy.x, in.wz.x, res.pdf.x);
                    qout[2 * p0 + 1] = make_float4(res.ij.x, __uint_as_float(res.comp[0]), __uint_as_float((uint32_t)res.valid[0]), 0.f);
                    if (m1) {
                        qout[2 * p1] = make_float4(in.wx.y, in.wy.y, in.wz.y, res.pdf.y);
                        qout[2 * p1 + 1] = make_float4(res.ij.y, __uint_as_float(res.comp[1]), __uint_as_float((uint32_t)res.valid[1]), 0.f);
                    }
                } else {
                    if (SAMPLE) {
                        q.out_dir3[p0] = in.wx.x; q.out_dir3[q.n + p0] = in.wy.x; q.out_dir3[2 * q.n + p0] = in.wz.x;
                        q.out_inv_jac[p0] = res.ij.x;
                        q.out_comp[p0] = res.comp[0];
                    }
                    q.out_pdf[p0] = res.pdf.x;
                    if (q.out_valid) q.out_valid[p0] = res.valid[0];
                    if (m1) {
                        if (SAMPLE) {
                            q.out_dir3[p1] = in.wx.y; q.out_dir3[q.n + p1] = in.wy.y; q.out_dir3[2 * q.n + p1] = in.wz.y;
                            q.out_inv_jac[p1] = res.ij.y;
                            q.out_comp[p1] = res.comp[1];
                        }
                        q.out_pdf[p1] = res.pdf.y;
                        if (q.out_valid) q.out_valid[p1] = res.valid[1];
                    }
                }
            }
        }
        pos = rend;
    }
}

// Results of device-binned vertices back into the caller's order: thread i reads its sorted position
// (coalesced), fetches the 32-byte result record there (one scattered sector) and writes the output
// planes coalesced.
template <bool SAMPLE>
__global__ void query_unpermute_kernel(QueryLaunch q, const uint32_t* __restrict__ invperm,
                                       const float4* __restrict__ qout) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t step = (size_t)gridDim.x * blockDim.x;
    for (; i < q.n; i += step) {
        const size_t p = invperm[i];
        const float4 a = __ldg(qout + 2 * p), b = __ldg(qout + 2 * p + 1);
        if (SAMPLE) {
            q.out_dir3[i] = a.x; q.out_dir3[q.n + i] = a.y; q.out_dir3[2 * q.n + i] = a.z;
            q.out_inv_jac[i] = b.x;
            q.out_comp[i] = __float_as_uint(b.y);
        }
        q.out_pdf[i] = a.w;
        if (q.out_valid) q.out_valid[i] = (uint8_t)__float_as_uint(b.z);
    }
}

// ------------------------------------------------------------------------------------------------
// create_conditional_pruned (R: distributions/sdmm_conditioner.h:128-168) + SDMM::sample + posterior: only the
// `max_components` strongest conditional weights of a vertex survive (the reference walks the components in
// ascending order of weight and zeroes the first K_user - max_components, skipping `preserve_idx`, which shrinks
// the survivor budget by one), then Categorical::prepare, select, sample, pdf as in the full query.  The reference
// defines the function but no call site of its integrator uses it, so this is a plain one-thread-per-vertex kernel
// (the K weights of a vertex live in local memory), not a tuned one; with max_components >= K_user it returns
// the bits of query_pair_kernel.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
query_pruned_kernel(QueryLaunch q, int K_user, int max_components, int preserve_idx) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= q.n) return;
    const int K = q.K;
    const uint32_t cell = q.cell_idx[i];
    float wx = 0.f, wy = 0.f, wz = 0.f, pdf = 0.f, ij = 0.f;
    uint32_t comp = 0u;
    uint8_t valid = 0;
    if (cell < (uint32_t)q.max_cells) {
        const float4* rec = reinterpret_cast<const float4*>(q.qrec + (size_t)cell * K * QREC_FLOATS);
        const float x = q.pos3[i], y = q.pos3[q.n + i], z = q.pos3[2 * q.n + i];
        float w[SDMMB_QUERY_MAX_K];
        for (int k = 0; k < K; ++k) w[k] = single_weight<false>(rec, k, x, y, z);
        // kill the weakest components: ascending order of (weight, index), visited once each
        uint32_t seen[SDMMB_QUERY_MAX_K / 32] = {0u, 0u, 0u, 0u};
        int budget = max_components;
        for (int slice = 0; slice < K_user - budget; ++slice) {
            int best = -1;
            float bw = 0.f;
            for (int k = 0; k < K_user; ++k) {
                if (seen[k >> 5] >> (k & 31) & 1u) continue;
                const float wk = w[k];
                if (best < 0 || wk < bw) { best = k; bw = wk; }
            }
            if (best < 0) break;
            seen[best >> 5] |= 1u << (best & 31);
            if (best == preserve_idx) { --budget; continue; }
            w[best] = 0.f;
        }
        // a killed weight entered `seen` with its ORIGINAL value, so zeroing never changes the visiting order
        float total = 0.f;
        for (int k = 0; k < K; ++k) total = (k == 0) ? w[0] : total + w[k];
        valid = total > 1e-20f ? 1 : 0;                                // R: categorical.h:97-101
        if (valid) {
            const float inv = 1.f / total;
            if (q.u3) {
                const float u0 = q.u3[i], u1 = q.u3[q.n + i], u2 = q.u3[2 * q.n + i];
                int sel = K - 1;
                float run = 0.f;
                bool found = false;
                for (int k = 0; k < K; ++k) {
                    run = (k == 0) ? w[0] : run + w[k];
                    if (!found && !(run * inv < u0)) { sel = k; found = true; }
                }
                while (sel > 0 && w[sel] * inv == 0.f) --sel;
                comp = (uint32_t)sel;
                sample_direction<false>(rec, sel, x, y, z, u1, u2, wx, wy, wz, ij);
            } else {
                wx = q.dir3[i]; wy = q.dir3[q.n + i]; wz = q.dir3[2 * q.n + i];
            }
            for (int k = 0; k < K; ++k) pdf = fmaf(cond_pdf_single<false>(rec, k, x, y, z, wx, wy, wz), w[k] * inv, pdf);
        }
    }
    if (q.u3) {
        q.out_dir3[i] = wx; q.out_dir3[q.n + i] = wy; q.out_dir3[2 * q.n + i] = wz;
        q.out_inv_jac[i] = ij;
        q.out_comp[i] = comp;
    }
    q.out_pdf[i] = valid ? pdf : 0.f;
    if (q.out_valid) q.out_valid[i] = valid;
}

}  // namespace

cudaError_t launch_query_pruned(const QueryLaunch& q, int K_user, int max_components, int preserve_idx, cudaStream_t st) {
    if (q.n == 0) return cudaSuccess;
    if (q.K > SDMMB_QUERY_MAX_K) return cudaErrorInvalidValue;
    query_pruned_kernel<<<(unsigned)((q.n + 127) / 128), 128, 0, st>>>(q, K_user, max_components, preserve_idx);
    return cudaGetLastError();
}

namespace {
}  // namespace

// Counting sort of n vertices by cell into the caller-provided scratch (words):
//   count[B] | cursor[B] | invperm[n] | pad to 16 B | qin[8 n]      (B = max_cells + 1; the caller may append more)
// qin holds, per sorted position, (x, y, z, in0) (in1, in2, original index, cell).
cudaError_t launch_bin_vertices(const uint32_t* cell_idx, size_t n, int max_cells, const float* pos3,
                                const float* in3, uint32_t* scratch, float4** out_qin, uint32_t** out_invperm,
                                cudaStream_t st) {
    cudaError_t e;
    const uint32_t B = (uint32_t)max_cells + 1;
    uint32_t* count = scratch;
    uint32_t* cursor = count + B;
    uint32_t* invperm = cursor + B;
    size_t off = (size_t)2 * B + n;
    off = (off + 3) & ~(size_t)3;
    float4* qin = reinterpret_cast<float4*>(scratch + off);
    BinPayload pl{pos3, in3, qin, invperm};
    if ((e = cudaMemsetAsync(count, 0, (size_t)B * 4, st)) != cudaSuccess) return e;
    const size_t hist_smem = (size_t)2 * B * sizeof(uint32_t);
    if (hist_smem <= 160 * 1024) {
        static bool configured = false;
        if (!configured) {
            if ((e = cudaFuncSetAttribute(bin_count_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024)) != cudaSuccess) return e;
            if ((e = cudaFuncSetAttribute(bin_scatter_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024)) != cudaSuccess) return e;
            configured = true;
        }
        // one chunk per CTA; few, fat CTAs keep the per-bucket global atomics rare
        const unsigned grid = (unsigned)std::max<size_t>(1, std::min<size_t>((n + 8191) / 8192, 148 * 4));
        bin_count_smem_kernel<<<grid, 512, hist_smem / 2, st>>>(cell_idx, n, B, count);
        bin_scan_kernel<<<1, 1024, 0, st>>>(count, B, cursor);
        bin_scatter_smem_kernel<<<grid, 512, hist_smem, st>>>(cell_idx, n, B, cursor, pl);
    } else {
        const unsigned grid = (unsigned)std::min<size_t>((n + 255) / 256, 148 * 16);
        bin_count_kernel<<<grid, 256, 0, st>>>(cell_idx, n, B, count);
        bin_scan_kernel<<<1, 1024, 0, st>>>(count, B, cursor);
        bin_scatter_kernel<<<grid, 256, 0, st>>>(cell_idx, n, B, cursor, pl);
    }
    *out_qin = qin;
    *out_invperm = invperm;
    return cudaGetLastError();
}

cudaError_t launch_query(const QueryLaunch& q, cudaStream_t st) {
    if (q.n == 0) return cudaSuccess;
    const bool sample = q.u3 != nullptr;
    cudaError_t e;
    float4* qin = nullptr;
    float4* qout = nullptr;
    uint32_t* invperm = nullptr;
    if (!q.binned) {
        if ((e = launch_bin_vertices(q.cell_idx, q.n, q.max_cells, q.pos3, sample ? q.u3 : q.dir3, q.scratch, &qin,
                                     &invperm, st)) != cudaSuccess) return e;
        qout = qin + 2 * q.n;
    }
    // tile: multiple of 2 * QTHREADS; short enough that the last wave's tail is small, long enough that
    // the per-run record staging (K * 128 B) stays negligible
    static const int tile_env = [] {
        const char* s = getenv("SDMMB_QUERY_TILE");   // tuning knob (multiple of 2 * QTHREADS)
        const int v = s ? atoi(s) : 0;
        return (v >= 2 * QTHREADS && v % (2 * QTHREADS) == 0) ? v : 0;
    }();
    const int tile = tile_env ? tile_env : 2048;
    const unsigned grid = (unsigned)((q.n + tile - 1) / tile);
#define SDMMB_QLAUNCH1(S, KK, R, A, B)                                                                  \
    do {                                                                                                \
        auto kern = query_pair_kernel<S, KK, R>;                                                        \
        const size_t smem = QShape<S, KK>::SMEM;                                                        \
        /* per device, so set on every launch (a process may hold contexts on several GPUs); it is a host-side table write */ \
        if (smem > 48 * 1024 &&                                                                         \
            (e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e; \
        kern<<<grid, QTHREADS, smem, st>>>(q, A, B, tile);                                              \
    } while (0)
#define SDMMB_QLAUNCH(KK)                                                                               \
    do {                                                                                                \
        if (sample && qin) SDMMB_QLAUNCH1(true, KK, true, qin, qout);                                   \
        else if (sample) SDMMB_QLAUNCH1(true, KK, false, nullptr, nullptr);                             \
        else if (qin) SDMMB_QLAUNCH1(false, KK, true, qin, qout);                                       \
        else SDMMB_QLAUNCH1(false, KK, false, nullptr, nullptr);                                        \
    } while (0)
    switch (q.K) {
        case 16: SDMMB_QLAUNCH(16); break;
        case 32: SDMMB_QLAUNCH(32); break;
        case 64: SDMMB_QLAUNCH(64); break;
        case 128: SDMMB_QLAUNCH(128); break;
        default: return cudaErrorInvalidValue;
    }
#undef SDMMB_QLAUNCH1
#undef SDMMB_QLAUNCH
    if (qin) {
        const unsigned ug = (unsigned)std::min<size_t>((q.n + 255) / 256, 148 * 16);
        if (sample) query_unpermute_kernel<true><<<ug, 256, 0, st>>>(q, invperm, qout);
        else query_unpermute_kernel<false><<<ug, 256, 0, st>>>(q, invperm, qout);
    }
    return cudaGetLastError();
}

}  // namespace sdmmb
