// init_kernels.cu -- sdmm::initialize on the device.
//
// Reference: /root/reference/sdmm/include/sdmm/opt/init.h:17-193 ("R:" below), called once per new
// cell by initializeSDMMContext (mitsuba/src/integrators/sdmm/volpath_sdmm.cpp:132-137) with
// n_spatial = K/8 seeds and a COPY of the context's default-seeded enoki::PCG32.
//
// The algorithm is a farthest-point-style seeding that is sequential over the K/8 seeds but parallel
// over the cell's samples, so one CTA owns one cell:
//   * per seed, every thread updates the running minimum distances and the unnormalised sampling pmf
//     of its samples (R: init.h:96-126);
//   * Categorical::prepare (R: distributions/categorical.h:86-114) is a LEFT-TO-RIGHT float32 prefix
//     sum; the seed that gets drawn depends on the exact cdf bits, so thread 0 walks the pmf
//     sequentially out of shared-memory tiles that the whole CTA stages and writes back (the only
//     serial part: ~6 cycles per sample per seed);
//   * the seed draw replays PCG32 bit-exactly (R: ext/enoki/include/enoki/random.h:33-101) and
//     enoki::binary_search (array_utils.h:150-170) + the zero-pmf back-off (categorical.h:27-40).
// The initial components, priors and zeroed EM state are written straight into the context's device
// arrays; the caller then runs the prepare kernel (mode 4 of em_mstep_kernel) as after an upload.
#include "kernels.h"
#include "sdmm_math.cuh"

namespace sdmmb {

namespace {

constexpr int INIT_THREADS = 256;
constexpr int INIT_TILE = 2048;
constexpr int INIT_MAX_SEEDS = 16;  // K = 128

struct Pcg32 {
    uint64_t state, inc;
    __device__ uint32_t next_uint32() {
        uint64_t oldstate = state;
        state = oldstate * 0x5851f42d4c957f2dULL + inc;
        uint32_t xorshifted = (uint32_t)(((oldstate >> 18) ^ oldstate) >> 27);
        uint32_t rot = (uint32_t)(oldstate >> 59);
        return (xorshifted >> rot) | (xorshifted << ((32u - rot) & 31u));
    }
    __device__ void seed(uint64_t initstate, uint64_t initseq) {
        state = 0;
        inc = (initseq << 1) | 1u;
        next_uint32();
        state += initstate;
        next_uint32();
    }
    __device__ float next_float32() { return __uint_as_float((next_uint32() >> 9) | 0x3f800000u) - 1.f; }
};

// enoki::min(enoki::max(w, 1e-3), 3)                                         R: init.h:64-66, 111-112
__device__ __forceinline__ float clamp_weight(float w) {
    double m = (double)w > 1e-3 ? (double)w : 1e-3;
    m = m < 3.0 ? m : 3.0;
    return (float)m;
}

// Categorical::prepare over pmf[0..N): sequential float32 prefix sum, then pmf, cdf *= 1/sum.
// Returns (to every thread) whether the sum is valid (> 1e-20).
__device__ bool categorical_prepare(float* __restrict__ pmf, float* __restrict__ cdf, int N, float* tile,
                                    float* s_carry) {
    if (threadIdx.x == 0) *s_carry = 0.f;
    for (int base = 0; base < N; base += INIT_TILE) {
        const int n = min(INIT_TILE, N - base);
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += INIT_THREADS) tile[i] = pmf[base + i];
        __syncthreads();
        if (threadIdx.x == 0) {
            float run = *s_carry;
            int i = 0;
            if (base == 0) {  // cdf[0] = pmf[0] (not 0 + pmf[0]: identical in float, kept for the letter)
                run = tile[0];
                i = 1;
            }
#pragma unroll 8
            for (; i < n; ++i) {
                run = __fadd_rn(run, tile[i]);
                tile[i] = run;
            }
            *s_carry = run;
        }
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += INIT_THREADS) cdf[base + i] = tile[i];
    }
    __syncthreads();
    const float sum = *s_carry;
    if (!(sum > 1e-20f)) return false;
    const float inv = 1.f / sum;
    for (int i = threadIdx.x; i < N; i += INIT_THREADS) {
        cdf[i] *= inv;
        pmf[i] *= inv;
    }
    __syncthreads();
    return true;
}

// sdmm::sample(Categorical, rng) for one thread                       R: categorical.h:27-40
__device__ uint32_t categorical_sample(const float* __restrict__ pmf, const float* __restrict__ cdf, int N,
                                       float u) {
    uint32_t start = 0, end = (uint32_t)N - 1;
    uint32_t iterations = (start < end) ? (32u - __clz(end - start)) : 0u;  // log2i(end - start) + 1
    for (uint32_t i = 0; i < iterations; ++i) {
        uint32_t middle = (start + end) >> 1;
        if (cdf[middle] < u) start = min(middle + 1, end);
        else end = middle;
    }
    uint32_t idx = start;
    while (idx > 0 && pmf[idx] == 0.f) --idx;
    return idx;
}

}  // namespace

__global__ void __launch_bounds__(INIT_THREADS) init_cells_kernel(InitArgs a) {
    __shared__ float tile[INIT_TILE];
    __shared__ float s_carry;
    __shared__ uint32_t s_pick;
    __shared__ float s_pos[INIT_MAX_SEEDS][3];
    __shared__ float s_from[INIT_MAX_SEEDS][9];
    __shared__ float s_cov[INIT_MAX_SEEDS][9];
    __shared__ float s_dprior[INIT_MAX_SEEDS][9];

    const int ci = blockIdx.x;
    const int cell = a.first_cell + ci;
    const int K = a.K, Ku = a.K_user;
    const int n_spatial = Ku / 8;
    const uint32_t begin = a.offsets[ci], end = a.offsets[ci + 1];
    const int N = (int)(end - begin);
    const float* px = a.point6 + begin;
    const float* nx = a.normal3 + begin;
    const float* wt = a.weight + begin;
    const size_t S = a.stride;
    float* min_d = a.scratch + (size_t)begin * 5;   // 5 arrays of N floats, contiguous per cell
    float* min_sp = min_d + N;
    float* min_nd = min_sp + N;
    float* pmf = min_nd + N;
    float* cdf = pmf + N;

    float* joint = a.joint + (size_t)cell * JOINT_PLANES * K;
    float* em = a.em + (size_t)cell * EM_PLANES * K;
    uint32_t* scal = a.scal + (size_t)cell * SCALAR_WORDS;

    // em = zero<EM>(K); sdmm.cov = 0                                             R: init.h:36-39
    for (int i = threadIdx.x; i < EM_PLANES * K; i += INIT_THREADS) em[i] = 0.f;
    for (int i = threadIdx.x; i < JOINT_PLANES * K; i += INIT_THREADS) joint[i] = 0.f;
    if (N <= 0 || n_spatial <= 0) {
        if (threadIdx.x == 0) scal[7] = ST_STEP_ABORTED;
        return;
    }

    const float max_rad_sqr = (float)6.251388631170325;  // quantile(chi2(3), 0.90)   R: init.h:40-44
    const float sdist = a.spatial_distance[ci];
    const float width_var = sdist * sdist / max_rad_sqr;
    const float depth_var = (float)(3e-2 * 3e-2 / (double)max_rad_sqr);
    const float directional_var = (float)(1.f * 3.14159265358979323846 / 8.f);
    const float SPATIAL_THRESHOLD = (float)(2e-2 * 2e-2), NORMAL_THRESHOLD = (float)(2e-1 * 2e-1);

    for (int j = threadIdx.x; j < N; j += INIT_THREADS) {
        min_d[j] = min_sp[j] = min_nd[j] = INFINITY;
        pmf[j] = clamp_weight(wt[j]);                                          // R: init.h:64-66
    }
    categorical_prepare(pmf, cdf, N, tile, &s_carry);                          // R: init.h:68

    Pcg32 rng;
    rng.seed(a.rng_state, a.rng_stream);  // every thread replays the same stream (a few integer ops)

    for (int sc = 0; sc < n_spatial; ++sc) {
        const float u = rng.next_float32();
        if (threadIdx.x == 0) {
            const uint32_t pick = categorical_sample(pmf, cdf, N, u);          // R: init.h:77
            s_pick = pick;
            if (a.out_seed_idx) a.out_seed_idx[(size_t)ci * n_spatial + sc] = (int32_t)pick;
            const float n[3] = {nx[pick], nx[S + pick], nx[2 * S + pick]};
            float to[9], from[9];
            frame_from_normal(n, to, from);                                    // R: init.h:84-87
            const float s[3] = {from[0], from[3], from[6]}, t[3] = {from[1], from[4], from[7]};
            for (int r = 0; r < 3; ++r) s_pos[sc][r] = px[r * S + pick];
            for (int i = 0; i < 9; ++i) s_from[sc][i] = from[i];
            for (int r = 0; r < 3; ++r)
                for (int q = 0; q < 3; ++q) {
                    s_cov[sc][3 * r + q] = __fadd_rn(__fadd_rn(__fmul_rn(width_var, __fmul_rn(s[r], s[q])),
                                                               __fmul_rn(width_var, __fmul_rn(t[r], t[q]))),
                                                     __fmul_rn(depth_var, __fmul_rn(n[r], n[q])));
                    s_dprior[sc][3 * r + q] = (float)((double)__fmul_rn(n[r], n[q]) * 1e-6);  // R: init.h:88
                }
        }
        __syncthreads();
        const uint32_t pick = s_pick;
        const float n0 = nx[pick], n1 = nx[S + pick], n2 = nx[2 * S + pick];
        const float m0 = s_pos[sc][0], m1 = s_pos[sc][1], m2 = s_pos[sc][2];
        for (int j = threadIdx.x; j < N; j += INIT_THREADS) {                  // R: init.h:96-126
            const float ndot = __fadd_rn(__fadd_rn(__fmul_rn(n0, nx[j]), __fmul_rn(n1, nx[S + j])),
                                         __fmul_rn(n2, nx[2 * S + j]));
            const double ang = (double)acosf(fminf(1.f, fmaxf(-1.f, ndot))) / 3.14159265358979323846;
            const float normal_d = (float)(ang * ang);
            const float dx = m0 - px[j], dy = m1 - px[S + j], dz = m2 - px[2 * S + j];
            const float position_d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
            const float d = __fadd_rn(normal_d, position_d);
            float md = min_d[j], msp = min_sp[j], mnd = min_nd[j], p = pmf[j];
            if (d < md) {
                md = d;
                const float d2 = __fmul_rn(d, d);                               // enoki::pow(d, 5)
                p = __fmul_rn(clamp_weight(wt[j]), __fmul_rn(d, __fmul_rn(d2, d2)));
            }
            if (normal_d < NORMAL_THRESHOLD && position_d < msp) {
                msp = position_d;
                mnd = normal_d;
            }
            if (mnd < NORMAL_THRESHOLD && msp < SPATIAL_THRESHOLD) {
                md = 0.f;
                p = 0.f;
            }
            min_d[j] = md; min_sp[j] = msp; min_nd[j] = mnd; pmf[j] = p;
        }
        __syncthreads();
        if (!categorical_prepare(pmf, cdf, N, tile, &s_carry)) {               // R: init.h:127-139
            const float uni = 1.f / (float)N;
            for (int j = threadIdx.x; j < N; j += INIT_THREADS) pmf[j] = uni;
            categorical_prepare(pmf, cdf, N, tile, &s_carry);
        }
    }
    __syncthreads();

    // the 2 x 4 directional lobes of every seed                                R: init.h:141-179
    for (int k = threadIdx.x; k < K; k += INIT_THREADS) {
        float* jk = joint + k;
        if (k >= Ku) {  // padding component (as sdmmb_upload_cells): zero weight, +z, identity covariance
            jk[6 * K] = 1.f;
            for (int i = 0; i < 5; ++i) jk[(7 + 6 * i) * K] = 1.f;
            continue;
        }
        const int sc = k >> 3, lobe = k & 7, ti = lobe >> 2, pi2 = lobe & 3;
        float theta = 0.f, phi = 0.f;
        for (int i = 0; i <= ti; ++i) theta = (float)((double)theta + 0.5 * 3.14159265358979323846 / (double)(2.f + 1.f));
        for (int i = 0; i <= pi2; ++i) phi = (float)((double)phi + 2.0 * 3.14159265358979323846 / (double)4.f);
        const float cos_theta = cosf(theta);
        const float sin_theta = sqrtf(fmaxf(1.f - cos_theta * cos_theta, 0.f));
        float sin_phi, cos_phi;
        sincosf(phi, &sin_phi, &cos_phi);
        const float dl[3] = {sin_theta * cos_phi, sin_theta * sin_phi, cos_theta};
        const float* from = s_from[sc];
        jk[0] = 1.f / (float)Ku;
        for (int r = 0; r < 3; ++r) jk[(1 + r) * K] = s_pos[sc][r];
        for (int r = 0; r < 3; ++r)
            jk[(4 + r) * K] = from[3 * r] * dl[0] + from[3 * r + 1] * dl[1] + from[3 * r + 2] * dl[2];
        for (int r = 0; r < 3; ++r)
            for (int q = 0; q < 3; ++q) {
                jk[(7 + 5 * r + q) * K] = s_cov[sc][3 * r + q];
                em[(56 + 5 * r + q) * K + k] = s_dprior[sc][3 * r + q];
            }
        jk[(7 + 18) * K] = directional_var;
        jk[(7 + 24) * K] = directional_var;
        em[(31 + 0) * K + k] = 2e-3f;                                          // R: init.h:181-192
        em[(31 + 6) * K + k] = 2e-3f;
        em[(31 + 12) * K + k] = 2e-3f;
        em[(31 + 18) * K + k] = 2e-4f;
        em[(31 + 24) * K + k] = 2e-4f;
    }
    if (threadIdx.x == 0) {
        scal[0] = __float_as_uint(0.f);
        scal[1] = 0u;
        scal[2] = 0u;
        scal[3] = __float_as_uint(1.f / (float)Ku);
        scal[4] = __float_as_uint(5.f / (float)Ku);
        scal[5] = __float_as_uint(0.2f);   // learning_rate, alpha keep their defaults  R: em.h:137-138
        scal[6] = __float_as_uint(0.5f);
        scal[7] = 0u;
    }
}

// Children inherit their parent's complete state on a split (R: accelerators/spatial_tree.h:187-196): one CTA per
// (source, destination) pair copies the cell's block of every per-cell array.  The sources must not be destinations
// of the same launch (the host resolves copy chains to their original cell first).
__global__ void copy_cells_kernel(CopyCellsArgs a) {
    const size_t s = (size_t)a.pairs[2 * blockIdx.x], d = (size_t)a.pairs[2 * blockIdx.x + 1];
    if (s == d) return;
    for (int arr = 0; arr < a.n_arrays; ++arr) {
        const size_t len = a.floats_per_cell[arr];
        uint32_t* base = a.base[arr];
        for (size_t i = threadIdx.x; i < len; i += blockDim.x) base[d * len + i] = base[s * len + i];
    }
}

cudaError_t launch_copy_cells(const CopyCellsArgs& a, int n_pairs, cudaStream_t st) {
    if (n_pairs <= 0) return cudaSuccess;
    copy_cells_kernel<<<n_pairs, 256, 0, st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_init_cells(const InitArgs& a, cudaStream_t st) {
    if (a.n_cells <= 0) return cudaSuccess;
    init_cells_kernel<<<a.n_cells, INIT_THREADS, 0, st>>>(a);
    return cudaGetLastError();
}

}  // namespace sdmmb
