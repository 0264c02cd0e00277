// mstep_kernels.cu -- the per-component half of the SDMM training path on sm_100a:
//   K2 em_mstep_kernel         reduction of the E-step partials over a cell's items (re-centred about the
//                              component means in double), stepwise blend, normalise, update_model,
//                              prepare (Cholesky, E-step records, CDF) and -- fused in its tail --
//   K3                         the conditioner prepare (Schur complement -> query records) and
//   K5a                        the record exchange over peer memory (NVLink stores from K3's epilogue)
//   mode without bit 0         K2's tail alone, for freshly uploaded models
// (K0/K1, the N*K hot loop, live in estep_kernels.cu.)
//
// Reference for every formula: /root/reference/sdmm/include/sdmm ("R:" below), see DESIGN.md.
#include "kernels.h"
#include "sdmm_math.cuh"

namespace sdmmb {

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

// pmf must be the NORMALISED weight (after Categorical::prepare); cen = the cell's centre (the E-step
// stages positions relative to it and the record carries the matching bias, see estep_kernels.cu).
__device__ inline void write_erec(float* __restrict__ erec_cell, int K, int k, float pmf,
                                  const float mean[6], const ComponentPrep& cp, const float cen[3]) {
    float W[25];
    inverse_lower_tri<5>(cp.L, W);
    const float scale = 0.84932180028801904272f;  // sqrt(log2(e) / 2)
    // factored Duff frame of the unit mean direction (rows of cp.to are s, t, n)
    const float nx = cp.to[6], ny = cp.to[7], nz = cp.to[8];
    const float sign = copysignf(1.f, nz);
    float* f = erec_cell + (size_t)EREC_ROT * K + k;
    f[0 * K] = nx;
    f[1 * K] = ny;
    f[2 * K] = -nz;
    f[3 * K] = -1.f / (sign + nz);
    f[4 * K] = sign * nx;
    f[5 * K] = sign * ny;
    // z = W (x - mu, t_dir) = W_pos (x - cen) + b + W_dir t_dir with b = -W_pos (mu - cen): formed in double from
    // the float32 entries the E-step multiplies with, rounded once
    const double m[3] = {(double)mean[0] - (double)cen[0], (double)mean[1] - (double)cen[1],
                         (double)mean[2] - (double)cen[2]};
    int e = 0;
#pragma unroll
    for (int r = 0; r < 5; ++r) {
        double b = 0.0;
#pragma unroll
        for (int c = 0; c <= r; ++c) {
            const float w = W[r * 5 + c] * scale;
            erec_cell[(EREC_W + e++) * K + k] = (c == 4) ? sign * w : w;
            if (c < 3) b -= (double)w * m[c];
        }
        erec_cell[(EREC_B + r) * K + k] = (float)b;
    }
    // pi_k * inv_det * (2 pi)^-5/2; 0 switches a zero-weight component off
    erec_cell[EREC_NLC * K + k] = pmf * cp.inv_det * gaussian_normalization(5);
#pragma unroll
    for (int d = 0; d < 3; ++d) erec_cell[(EREC_CEN + d) * K + k] = cen[d];
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
            atomicExch(lf + PEER_FAIL, 1u);
            s_go = 0;   // several threads may store the same 0: benign
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
            // some CTA skipped its stores (a peer never raised its ready flag): the peers hold STALE records of
            // this rank's cells.  They are told so -- one count in THEIR error word, read by sdmmb_peer_status
            // on the receiving rank -- before the arrival flag lets their waits through (no 10 s stall per step).
            const bool failed = atomicExch(lf + PEER_FAIL, 0u) != 0u;
            __threadfence_system();
            for (int pp = 1; pp < P.world; ++pp) {
                const int p = (P.rank + pp) % P.world;
                if (failed) atomicAdd_system(P.flags[p] + PEER_ERROR, 1u);
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
    __shared__ float s_mpos[3][MSTEP_THREADS];           // component mean positions (after the update)
    __shared__ float s_cen[MSTEP_THREADS / 16 + 1][4];   // per cell: centre of the E-step's position frame
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
    const bool trainable = active && (a.mode & 1) && a.cell_sample_count[ci] >= a.min_samples &&
                           !(a.skip_iterations_ge != 0u && scal[1] >= a.skip_iterations_ge);
    if (trainable) {
        // ---- reduce the E-step partials of this cell's items (fixed order, double accumulators) --
        const uint32_t ib = a.cell_item_begin[ci] - a.item_base, ie = a.cell_item_begin[ci + 1] - a.item_base;
        double ut[ESTAT_PLANES];
#pragma unroll
        for (int i = 0; i < ESTAT_PLANES; ++i) ut[i] = 0.0;
        float weight_sum = 0.f;
        int nskip = 0;
        for (uint32_t it = ib; it < ie; ++it) {
            const float* p = a.partials + (size_t)it * ESTAT_PLANES * K;
#pragma unroll
            for (int i = 0; i < ESTAT_PLANES; ++i) ut[i] += (double)p[i * K + k];
            weight_sum += a.item_scal[2 * (size_t)it];
            nskip += __float_as_int(a.item_scal[2 * (size_t)it + 1]);
        }
        // The E-step accumulated the position entries about the cell's centre (v = x - cen), not about the
        // component mean: with m = mu - cen and t_pos = v - m (exact algebra of Stats::plus_eq, R: opt/em.h:57-62)
        //   sum g t_i       = U1_i - m_i U0
        //   sum g t_i t_j   = U2_ij - m_i U1_j - m_j U1_i + m_i m_j U0      (i, j < 3)
        //   sum g t_d t_j   = C_dj - m_j S1_d                               (d directional, j < 3)
        // The differences cancel to (component extent / distance to the centre)^2 of their terms: double.
        float bt[ESTAT_PLANES];
        {
            const float* er = a.erec + (size_t)cell * EREC_PLANES * K;   // the record the E-step read
            double m[3];
#pragma unroll
            for (int d = 0; d < 3; ++d) m[d] = (double)mean[d] - (double)er[(EREC_CEN + d) * K + k];
            const double u0 = ut[0];
            bt[0] = (float)u0;
#pragma unroll
            for (int i = 0; i < 3; ++i) bt[1 + i] = (float)(ut[1 + i] - m[i] * u0);
            bt[4] = (float)ut[4];
            bt[5] = (float)ut[5];
#pragma unroll
            for (int i = 0; i < 3; ++i)
#pragma unroll
                for (int j = 0; j <= i; ++j) {
                    const int e = 6 + i * (i + 1) / 2 + j;
                    bt[e] = (float)(ut[e] - m[i] * ut[1 + j] - m[j] * ut[1 + i] + m[i] * m[j] * u0);
                }
#pragma unroll
            for (int d = 3; d < 5; ++d) {
#pragma unroll
                for (int j = 0; j < 3; ++j) {
                    const int e = 6 + d * (d + 1) / 2 + j;
                    bt[e] = (float)(ut[e] - m[j] * ut[1 + d]);
                }
#pragma unroll
                for (int j = 3; j <= d; ++j) {
                    const int e = 6 + d * (d + 1) / 2 + j;
                    bt[e] = (float)ut[e];
                }
            }
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
            // 3^it and 2^it, it <= 30, are exact in double (3^30 < 2^53): a product loop gives what std::pow gives the
            // reference, without the several hundred instructions of a general double pow per thread
            double p3 = 1.0;
            for (int e = 0; e < it; ++e) p3 *= 3.0;
            const float weight_prior_decay = (float)(1.0 / p3);
            const float cov_prior_strength_decay = (float)(1.0 / (double)(1u << it));
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
#pragma unroll
    for (int d = 0; d < 3; ++d) s_mpos[d][threadIdx.x] = mean[d];
    __syncthreads();
    if (renorm && k == 0) s_tot[local] = categorical_prefix_seq(s_pmf + local * K, s_cdf + local * K, K);
    if (active && k >= 1 && k <= 3) {
        // centre of the cell for the next E-step: the mixture's weighted mean position (any point near
        // the data does; this one minimises sum_k pi_k |mu_k - cen|^2, the cancellation the un-centred
        // statistics have to survive).  Serial like the prefix sum next to it, one thread per coordinate.
        const float* w = s_pmf + local * K;
        const float* mp = s_mpos[k - 1] + local * K;
        float num = 0.f, den = 0.f;
        for (int i = 0; i < K; ++i) {
            const float wi = (w[i] > 0.f && fabsf(mp[i]) <= 3.402823466e+38f) ? w[i] : 0.f;
            num = fmaf(wi, mp[i], num);
            den += wi;
        }
        const float c = num / den;
        s_cen[local][k - 1] = (fabsf(c) <= 3.402823466e+38f) ? c : 0.f;
    }
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
        write_erec(a.erec + (size_t)cell * EREC_PLANES * K, K, k, pmf, mean, cp, s_cen[local]);
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
