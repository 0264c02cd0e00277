/* oracle/sdmm_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.  See sdmm_oracle.h.
 *
 * Scalar plain-C restatement of the reference's enoki SIMD code on the SDMM hot path.  It follows
 * the reference operation by operation (same guards, thresholds, priors and ordering of the
 * floating-point formulas) but loops over components one at a time instead of in packets of 8 and
 * uses libm (expf/acosf/...) where the reference uses enoki's polynomial approximations.
 * All paths "R:" below are relative to /root/reference/sdmm/include/sdmm/.
 */
#include "sdmm_oracle.h"

#include <math.h>
#include <xmmintrin.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define MAXD 8
#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif
#define PI_F 3.14159265358979323846f

/* enoki's packet exp (ext/enoki/include/enoki/array_math.h:711-770) returns exactly 0 below
 * -88.3762588501 (no gradual underflow); libm's expf would keep denormals down to e^-103.  The cut
 * decides which outlier samples the E-step skips (hsum == 0, opt/em.h:289-293). */
static float enoki_expf(float x) { return (x < -88.3762588501f) ? 0.f : expf(x); }

/* The reference INTEGRATOR runs with flush-to-zero / denormals-are-zero (flushDenormals = true by default,
 * mitsuba/src/integrators/sdmm/volpath_sdmm.cpp:62, 85-87 -> enoki::set_flush_denormals), so densities below
 * 2^-126 are exactly 0 there.  The EM / query entry points of this restatement run under the same mode. */
static unsigned ftz_enter(void) { unsigned old = _mm_getcsr(); _mm_setcsr(old | 0x8040u); return old; }
static void ftz_leave(unsigned old) { _mm_setcsr(old); }

/* R: core/constants.h:31-47  gaussian_normalization = (1/sqrt(2 pi))^D */
static float gaussian_normalization(int D) {
    return powf(0.39894228040143267793994605993438186847585863116492f, (float)D);
}

/* ------------------------------------------------------------------------------------------ */
/* linalg                                                                                      */
/* ------------------------------------------------------------------------------------------ */

/* R: linalg/cholesky.h:12-36.  Row-major DxD.  NOTE the +1e-5 added to every pivot (:32) and
 * the PSD flag = all pivots (with jitter) > 0 (:33).  sqrt of a negative pivot yields NaN exactly
 * as enoki::sqrt does. */
int oracle_cholesky(int D, const float* in, float* out) {
    int psd = 1;
    memset(out, 0, sizeof(float) * D * D);
    for (int r = 0; r < D; ++r) {
        for (int c = 0; c < r; ++c) {
            float ksum = 0.f;
            for (int k = 0; k < c; ++k) ksum += out[r * D + k] * out[c * D + k];
            out[r * D + c] = (in[r * D + c] - ksum) / out[c * D + c];
        }
        float row_sum = 0.f;
        for (int k = 0; k < D; ++k) row_sum += out[r * D + k] * out[r * D + k];
        float value = in[r * D + r] - row_sum + 1e-5f;
        psd = psd && (value > 0);
        out[r * D + r] = sqrtf(value);
    }
    return psd;
}

/* R: linalg/cholesky.h:38-58  forward substitution with divisions */
void oracle_solve_lower(int D, const float* L, const float* b, float* x) {
    x[0] = b[0] / L[0];
    for (int r = 1; r < D; ++r) {
        float numerator = b[r];
        for (int c = 0; c < r; ++c) numerator -= L[r * D + c] * x[c];
        x[r] = numerator / L[r * D + r];
    }
}

/* R: linalg/cholesky.h:60-78  column r of the inverse = solve(L, e_r) */
void oracle_inverse_lower_tri(int D, const float* L, float* inv) {
    float e[MAXD], col[MAXD];
    for (int r = 0; r < D; ++r) {
        for (int i = 0; i < D; ++i) e[i] = 0.f;
        e[r] = 1.f;
        oracle_solve_lower(D, L, e, col);
        for (int i = 0; i < D; ++i) inv[i * D + r] = col[i];
    }
}

/* R: linalg/coordinate_system.h:21-38 (Duff et al. branch-free ONB).  to = transpose(from);
 * to rows are (s, t, n).  enoki::sign = copysign(1, x) (ext/enoki array_router.h:370),
 * mulsign(a,b) = a*copysign(1,b), mulsign_neg(a,b) = a*copysign(1,-b) (:446, :472). */
void oracle_frame(const float* n, float* to, float* from) {
    float sign = copysignf(1.f, n[2]);
    float a = -1.f / (sign + n[2]);
    float b = n[0] * n[1] * a;
    float c0[3] = {(n[0] * n[0] * a) * sign + 1.f, b * sign, n[0] * copysignf(1.f, -n[2])};
    float c1[3] = {b, sign + n[1] * n[1] * a, -n[1]};
    for (int r = 0; r < 3; ++r) {
        from[r * 3 + 0] = c0[r];
        from[r * 3 + 1] = c1[r];
        from[r * 3 + 2] = n[r];
    }
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) to[r * 3 + c] = from[c * 3 + r];
}

static void matmul(int n, int m, int p, const float* A, const float* B, float* C) {
    /* C[n x p] = A[n x m] B[m x p], row-major (R: linalg/matrix.h:139-161) */
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < p; ++j) {
            float s = 0.f;
            for (int k = 0; k < m; ++k) s += A[i * m + k] * B[k * p + j];
            C[i * p + j] = s;
        }
}

static void transpose(int n, int m, const float* A, float* T) {
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < m; ++j) T[j * n + i] = A[i * m + j];
}

/* R: distributions/categorical.h:86-114  sequential prefix sum; valid iff sum > 1e-20;
 * when valid both pmf and cdf are multiplied by 1/sum; when invalid they are left untouched. */
int oracle_categorical_prepare(int n, float* pmf, float* cdf) {
    cdf[0] = pmf[0];
    for (int i = 1; i < n; ++i) cdf[i] = cdf[i - 1] + pmf[i];
    float pmf_sum = cdf[n - 1];
    int is_valid = pmf_sum > 1e-20f;
    if (!is_valid) return 0;
    float inv_normalizer = 1.f / pmf_sum;
    for (int i = 0; i < n; ++i) {
        cdf[i] *= inv_normalizer;
        pmf[i] *= inv_normalizer;
    }
    return 1;
}

/* R: distributions/sdmm.h:284-294 with EuclidianTangentSpace (spaces/euclidian.h:36-41):
 * tangent = x - mean, inv_jacobian = 1; z = solve(L, tangent);
 * pdf = inv_cov_sqrt_det * inv_jacobian * gaussian_normalization * exp(-0.5 |z|^2) */
float oracle_gauss_pdf_euclid(int D, const float* mean, const float* cov, const float* x) {
    float L[MAXD * MAXD], t[MAXD], z[MAXD];
    oracle_cholesky(D, cov, L);
    float prod = 1.f;
    for (int i = 0; i < D; ++i) prod *= L[i * D + i];
    float inv_det = 1.f / prod; /* sdmm.h:188 */
    for (int i = 0; i < D; ++i) t[i] = x[i] - mean[i];
    oracle_solve_lower(D, L, t, z);
    float sq = 0.f;
    for (int i = 0; i < D; ++i) sq += z[i] * z[i];
    return inv_det * 1.f * gaussian_normalization(D) * enoki_expf(-0.5f * sq);
}

/* R: distributions/sdmm_conditioner.h:170-212 (prepare_vectorized) + :7-26 (create_marginal),
 * linear-algebra part only: marginal Cholesky, A = L_aa^-1, cond_cov = cov_bb - (A cov_ab)^T (A cov_ab),
 * M = cov_ba A^T A, Cholesky of cond_cov. */
void oracle_schur(int D, int Da, const float* cov, float* L_a, float* M, float* cond_cov, float* L_c) {
    int Db = D - Da;
    float cov_aa[MAXD * MAXD], A[MAXD * MAXD], At[MAXD * MAXD];
    float cov_ab[MAXD * MAXD], cov_ba[MAXD * MAXD], cov_bb[MAXD * MAXD];
    float Aab[MAXD * MAXD], AabT[MAXD * MAXD], prod[MAXD * MAXD], tmp[MAXD * MAXD];
    for (int r = 0; r < Da; ++r)
        for (int c = 0; c < Da; ++c) cov_aa[r * Da + c] = cov[r * D + c];
    oracle_cholesky(Da, cov_aa, L_a);
    oracle_inverse_lower_tri(Da, L_a, A);
    for (int r = 0; r < Db; ++r)
        for (int c = 0; c < Db; ++c) cov_bb[r * Db + c] = cov[(Da + r) * D + (Da + c)];
    for (int r = 0; r < Db; ++r)
        for (int c = 0; c < Da; ++c) {
            cov_ba[r * Da + c] = cov[(Da + r) * D + c];
            cov_ab[c * Db + r] = cov[(Da + r) * D + c]; /* :199-201: both read joint.cov(r, c), r>=Da */
        }
    matmul(Da, Da, Db, A, cov_ab, Aab);       /* aa_sqrt_inv_ab  (Da x Db) */
    transpose(Da, Db, Aab, AabT);             /* (Db x Da) */
    matmul(Db, Da, Db, AabT, Aab, prod);
    for (int i = 0; i < Db * Db; ++i) cond_cov[i] = cov_bb[i] - prod[i];
    transpose(Da, Da, A, At);
    matmul(Db, Da, Da, cov_ba, At, tmp);      /* (cov_ba * A^T) * A, left to right (:206-207) */
    matmul(Db, Da, Da, tmp, A, M);
    oracle_cholesky(Db, cond_cov, L_c);
}

/* ------------------------------------------------------------------------------------------ */
/* per-cell state                                                                              */
/* ------------------------------------------------------------------------------------------ */

struct oracle_cell {
    int K;
    /* joint mixture (R: distributions/sdmm.h:118-124) */
    float *pmf, *cdf;        /* [K] */
    float* mean;             /* [6][K] */
    float *to, *from;        /* [9][K] row-major frames */
    float *cov, *cov_sqrt;   /* [25][K] row-major */
    float* inv_det;          /* [K] */
    unsigned char* psd;      /* [K] */
    /* EM (R: opt/em.h:124-142) */
    float *st_w, *st_mean, *st_cov;    /* stats            [K],[5][K],[25][K] */
    float *bt_w, *bt_mean, *bt_cov;    /* batch_stats */
    float *sn_w, *sn_mean, *sn_cov;    /* stats_normalized */
    float *cov_prior, *depth_prior;    /* [25][K] */
    float total_weight;
    int iterations_run, samples_seen;
    float weight_prior, cov_prior_strength, learning_rate, alpha;
    /* conditioner (R: distributions/sdmm_conditioner.h:70-76) */
    float *m_mean;           /* [3][K] marginal mean */
    float *m_L;              /* [9][K] marginal cov_sqrt */
    float *m_inv_det, *m_pmf, *m_cdf;
    float *M;                /* [6][K] mean_transform 2x3 */
    float *c_to, *c_from;    /* [9][K] conditional frame */
    float *c_cov, *c_L;      /* [4][K] */
    float *c_inv_det;
    int conditioner_ready;
};

#define P(arr, plane, k) ((arr)[(size_t)(plane) * (size_t)K + (size_t)(k)])

static float* falloc(size_t n) { return (float*)calloc(n, sizeof(float)); }

oracle_cell* oracle_cell_create(int K) {
    oracle_cell* c = (oracle_cell*)calloc(1, sizeof(oracle_cell));
    c->K = K;
    c->pmf = falloc(K); c->cdf = falloc(K); c->mean = falloc(6 * K);
    c->to = falloc(9 * K); c->from = falloc(9 * K);
    c->cov = falloc(25 * K); c->cov_sqrt = falloc(25 * K);
    c->inv_det = falloc(K); c->psd = (unsigned char*)calloc(K, 1);
    c->st_w = falloc(K); c->st_mean = falloc(5 * K); c->st_cov = falloc(25 * K);
    c->bt_w = falloc(K); c->bt_mean = falloc(5 * K); c->bt_cov = falloc(25 * K);
    c->sn_w = falloc(K); c->sn_mean = falloc(5 * K); c->sn_cov = falloc(25 * K);
    c->cov_prior = falloc(25 * K); c->depth_prior = falloc(25 * K);
    c->learning_rate = 0.2f; /* R: opt/em.h:137 */
    c->alpha = 0.5f;         /* R: opt/em.h:138 */
    c->m_mean = falloc(3 * K); c->m_L = falloc(9 * K);
    c->m_inv_det = falloc(K); c->m_pmf = falloc(K); c->m_cdf = falloc(K);
    c->M = falloc(6 * K); c->c_to = falloc(9 * K); c->c_from = falloc(9 * K);
    c->c_cov = falloc(4 * K); c->c_L = falloc(4 * K); c->c_inv_det = falloc(K);
    return c;
}

void oracle_cell_destroy(oracle_cell* c) {
    if (!c) return;
    free(c->pmf); free(c->cdf); free(c->mean); free(c->to); free(c->from); free(c->cov);
    free(c->cov_sqrt); free(c->inv_det); free(c->psd);
    free(c->st_w); free(c->st_mean); free(c->st_cov); free(c->bt_w); free(c->bt_mean);
    free(c->bt_cov); free(c->sn_w); free(c->sn_mean); free(c->sn_cov); free(c->cov_prior);
    free(c->depth_prior); free(c->m_mean); free(c->m_L); free(c->m_inv_det); free(c->m_pmf);
    free(c->m_cdf); free(c->M); free(c->c_to); free(c->c_from); free(c->c_cov); free(c->c_L);
    free(c->c_inv_det);
    free(c);
}

/* R: spaces/spatio_directional.h:247-255  set_mean: store the mean as given, build the frame from
 * the NORMALISED direction part. */
static void joint_set_mean(oracle_cell* c, int k, const float m[6]) {
    const int K = c->K;
    for (int d = 0; d < 6; ++d) P(c->mean, d, k) = m[d];
    float len = sqrtf(m[3] * m[3] + m[4] * m[4] + m[5] * m[5]);
    float inv = 1.f / len; /* enoki::normalize = v * rsqrt(|v|^2), ext/enoki array_generic */
    float n[3] = {m[3] * inv, m[4] * inv, m[5] * inv};
    float to[9], from[9];
    oracle_frame(n, to, from);
    for (int i = 0; i < 9; ++i) { P(c->to, i, k) = to[i]; P(c->from, i, k) = from[i]; }
}

/* R: distributions/sdmm.h:186-197 prepare_cov: Cholesky (+jitter), inv_cov_sqrt_det = 1/prod(diag) */
static void joint_prepare_cov(oracle_cell* c, int k) {
    const int K = c->K;
    float cov[25], L[25];
    for (int i = 0; i < 25; ++i) cov[i] = P(c->cov, i, k);
    c->psd[k] = (unsigned char)oracle_cholesky(5, cov, L);
    float prod = 1.f;
    for (int i = 0; i < 5; ++i) prod *= L[i * 5 + i];
    for (int i = 0; i < 25; ++i) P(c->cov_sqrt, i, k) = L[i];
    c->inv_det[k] = 1.f / prod;
}

/* R: distributions/sdmm.h:163-173 prepare_vectorized: prepare_cov on every component (weights of
 * non-PSD components are NOT zeroed here, unlike SDMM::prepare :176-184), then weight.prepare(). */
static int joint_prepare_vectorized(oracle_cell* c) {
    int all_psd = 1;
    for (int k = 0; k < c->K; ++k) {
        joint_prepare_cov(c, k);
        all_psd = all_psd && c->psd[k];
    }
    int ok = oracle_categorical_prepare(c->K, c->pmf, c->cdf);
    return ok && all_psd;
}

static float w2f(uint32_t w) { float f; memcpy(&f, &w, 4); return f; }
static uint32_t f2w(float f) { uint32_t w; memcpy(&w, &f, 4); return w; }

int oracle_cell_set(oracle_cell* c, const float* joint, const float* em, const uint32_t* scal) {
    const int K = c->K;
    for (int k = 0; k < K; ++k) {
        c->pmf[k] = joint[k];
        float m[6];
        for (int d = 0; d < 6; ++d) m[d] = P(joint, 1 + d, k);
        joint_set_mean(c, k, m);
        for (int i = 0; i < 25; ++i) P(c->cov, i, k) = P(joint, 7 + i, k);
    }
    int ok = joint_prepare_vectorized(c);
    if (em) {
        for (int k = 0; k < K; ++k) {
            c->st_w[k] = em[k];
            for (int d = 0; d < 5; ++d) P(c->st_mean, d, k) = P(em, 1 + d, k);
            for (int i = 0; i < 25; ++i) {
                P(c->st_cov, i, k) = P(em, 6 + i, k);
                P(c->cov_prior, i, k) = P(em, 31 + i, k);
                P(c->depth_prior, i, k) = P(em, 56 + i, k);
            }
        }
    }
    if (scal) {
        c->total_weight = w2f(scal[0]);
        c->iterations_run = (int)scal[1];
        c->samples_seen = (int)scal[2];
        c->weight_prior = w2f(scal[3]);
        c->cov_prior_strength = w2f(scal[4]);
        c->learning_rate = w2f(scal[5]);
        c->alpha = w2f(scal[6]);
    }
    c->conditioner_ready = 0;
    return ok;
}

void oracle_cell_get(oracle_cell* c, float* joint, float* em, uint32_t* scal) {
    const int K = c->K;
    if (joint)
        for (int k = 0; k < K; ++k) {
            joint[k] = c->pmf[k];
            for (int d = 0; d < 6; ++d) P(joint, 1 + d, k) = P(c->mean, d, k);
            for (int i = 0; i < 25; ++i) P(joint, 7 + i, k) = P(c->cov, i, k);
        }
    if (em)
        for (int k = 0; k < K; ++k) {
            em[k] = c->st_w[k];
            for (int d = 0; d < 5; ++d) P(em, 1 + d, k) = P(c->st_mean, d, k);
            for (int i = 0; i < 25; ++i) {
                P(em, 6 + i, k) = P(c->st_cov, i, k);
                P(em, 31 + i, k) = P(c->cov_prior, i, k);
                P(em, 56 + i, k) = P(c->depth_prior, i, k);
            }
        }
    if (scal) {
        scal[0] = f2w(c->total_weight);
        scal[1] = (uint32_t)c->iterations_run;
        scal[2] = (uint32_t)c->samples_seen;
        scal[3] = f2w(c->weight_prior);
        scal[4] = f2w(c->cov_prior_strength);
        scal[5] = f2w(c->learning_rate);
        scal[6] = f2w(c->alpha);
        scal[7] = 0;
    }
}

void oracle_cell_get_derived(oracle_cell* c, float* out) {
    const int K = c->K;
    for (int k = 0; k < K; ++k) {
        for (int i = 0; i < 9; ++i) P(out, i, k) = P(c->to, i, k);
        for (int i = 0; i < 25; ++i) P(out, 9 + i, k) = P(c->cov_sqrt, i, k);
        P(out, 34, k) = c->inv_det[k];
        P(out, 35, k) = c->cdf[k];
        P(out, 36, k) = c->pmf[k];
        P(out, 37, k) = c->psd[k] ? 1.f : 0.f;
    }
}

/* ------------------------------------------------------------------------------------------ */
/* tangent space of the joint mixture                                                          */
/* ------------------------------------------------------------------------------------------ */

static float safe_acosf(float a) { return acosf(fminf(1.f, fmaxf(-1.f, a))); } /* enoki array_math.h:1401 */
static float safe_sqrtf(float a) { return sqrtf(fmaxf(a, 0.f)); }               /* enoki array_math.h:1389 */

/* R: spaces/spatio_directional.h:86-114  SpatioDirectionalTangentSpace::to
 * (the |d|-1 >= 1e-4 check at :92-100 only prints; not restated) */
static void joint_to(const float mean[6], const float to[9], const float x[6], float t[5], float* inv_jac) {
    float lx = to[0] * x[3] + to[1] * x[4] + to[2] * x[5];
    float ly = to[3] * x[3] + to[4] * x[4] + to[5] * x[5];
    float lz = to[6] * x[3] + to[7] * x[4] + to[8] * x[5];
    float cos_angle = lz;
    float angle = safe_acosf(cos_angle);
    float sin_angle = safe_sqrtf(1.f - cos_angle * cos_angle);
    float rcp_sinc = (sin_angle < 1e-4f) ? 1.f : angle / sin_angle;
    *inv_jac = (cos_angle <= -1.f) ? 0.f : rcp_sinc;
    t[0] = x[0] - mean[0];
    t[1] = x[1] - mean[1];
    t[2] = x[2] - mean[2];
    t[3] = lx * rcp_sinc;
    t[4] = ly * rcp_sinc;
}

void oracle_joint_to(const float* mean6, const float* point6, float* tangent5, float* inv_jac) {
    float len = sqrtf(mean6[3] * mean6[3] + mean6[4] * mean6[4] + mean6[5] * mean6[5]);
    float n[3] = {mean6[3] / len, mean6[4] / len, mean6[5] / len};
    float to[9], from[9];
    oracle_frame(n, to, from);
    joint_to(mean6, to, point6, tangent5, inv_jac);
}

/* R: distributions/sdmm.h:331-339 posterior = pmf * pdf_gaussian (:284-294) */
static float joint_posterior(const oracle_cell* c, int k, const float x[6], float t[5]) {
    const int K = c->K;
    float mean[6], to[9], L[25], z[5], inv_jac;
    for (int d = 0; d < 6; ++d) mean[d] = P(c->mean, d, k);
    for (int i = 0; i < 9; ++i) to[i] = P(c->to, i, k);
    for (int i = 0; i < 25; ++i) L[i] = P(c->cov_sqrt, i, k);
    joint_to(mean, to, x, t, &inv_jac);
    oracle_solve_lower(5, L, t, z);
    float sq = 0.f;
    for (int i = 0; i < 5; ++i) sq += z[i] * z[i];
    float pdf = c->inv_det[k] * inv_jac * gaussian_normalization(5) * enoki_expf(-0.5f * sq);
    return pdf * c->pmf[k];
}

/* ------------------------------------------------------------------------------------------ */
/* EM                                                                                          */
/* ------------------------------------------------------------------------------------------ */

/* R: opt/data.h:20-23 */
static int is_valid_sample(float w) { return isfinite(w) && w >= 1e-8f; }

/* R: opt/em.h:255-314  EM::compute_stats_model_parallel.  sum_mode 1 accumulates the batch
 * statistics in float64 (oracle-only diagnostic, SURVEY.md §7 "summation order"). */
static void compute_stats(oracle_cell* c, int N, const float* point6, long stride, const float* weight,
                          float* out_loglik, int sum_mode) {
    const int K = c->K;
    float* post = (float*)malloc(sizeof(float) * K);
    float* tang = (float*)malloc(sizeof(float) * 5 * K);
    double* acc = sum_mode ? (double*)calloc((size_t)31 * K, sizeof(double)) : NULL;
    memset(c->bt_w, 0, sizeof(float) * K);
    memset(c->bt_mean, 0, sizeof(float) * 5 * K);
    memset(c->bt_cov, 0, sizeof(float) * 25 * K);
    c->samples_seen += N; /* :264 */
    for (int n = 0; n < N; ++n) {
        float w = weight[n];
        float x[6];
        for (int d = 0; d < 6; ++d) x[d] = point6[(size_t)d * stride + n];
        int valid = is_valid_sample(w);
        if (!valid && !out_loglik) continue; /* :267-269 */
        float sum = 0.f;
        for (int k = 0; k < K; ++k) {
            float t[5];
            post[k] = joint_posterior(c, k, x, t);
            for (int d = 0; d < 5; ++d) tang[d * K + k] = t[d];
            sum += post[k]; /* hsum, :288 */
        }
        if (out_loglik) out_loglik[n] = logf(sum);
        if (!valid) continue;
        if (sum == 0.f || !isfinite(1.f / sum)) { /* :289-293 */
            --c->samples_seen;
            continue;
        }
        float rcp = 1.f / sum;
        for (int k = 0; k < K; ++k) {
            float g = w * (post[k] * rcp); /* :295-298, :309 (data_slice.weight * posteriors) */
            /* Stats::plus_eq, :57-62: weight += g; mean += g*t; cov += g*outer(t) */
            if (!sum_mode) {
                c->bt_w[k] += g;
                for (int i = 0; i < 5; ++i) P(c->bt_mean, i, k) += g * tang[i * K + k];
                for (int i = 0; i < 5; ++i)
                    for (int j = 0; j < 5; ++j)
                        P(c->bt_cov, i * 5 + j, k) += g * (tang[i * K + k] * tang[j * K + k]);
            } else {
                acc[0 * K + k] += g;
                for (int i = 0; i < 5; ++i) acc[(1 + i) * K + k] += (double)g * tang[i * K + k];
                for (int i = 0; i < 5; ++i)
                    for (int j = 0; j < 5; ++j)
                        acc[(6 + i * 5 + j) * K + k] += (double)g * ((double)tang[i * K + k] * tang[j * K + k]);
            }
        }
    }
    if (sum_mode) {
        for (int k = 0; k < K; ++k) {
            c->bt_w[k] = (float)acc[k];
            for (int i = 0; i < 5; ++i) P(c->bt_mean, i, k) = (float)acc[(1 + i) * K + k];
            for (int i = 0; i < 25; ++i) P(c->bt_cov, i, k) = (float)acc[(6 + i) * K + k];
        }
        free(acc);
    }
    free(post);
    free(tang);
}

static float em_eta(const oracle_cell* c) {
    /* R: opt/em.h:318-319  eta = (learning_rate * iterations_run + 1)^(-alpha) */
    return powf(c->learning_rate * (float)c->iterations_run + 1.f, -c->alpha);
}

/* R: opt/em.h:316-328  stats = (1-eta) stats + eta batch_stats */
static void interpolate_stats(oracle_cell* c) {
    const int K = c->K;
    float eta = em_eta(c);
    for (int k = 0; k < K; ++k) {
        c->st_w[k] *= 1 - eta;
        c->st_w[k] = fmaf(eta, c->bt_w[k], c->st_w[k]); /* Stats::fmadd_eq :83 uses enoki::fmadd */
        for (int i = 0; i < 5; ++i) {
            P(c->st_mean, i, k) *= 1 - eta;
            P(c->st_mean, i, k) += eta * P(c->bt_mean, i, k);
        }
        for (int i = 0; i < 25; ++i) {
            P(c->st_cov, i, k) *= 1 - eta;
            P(c->st_cov, i, k) += eta * P(c->bt_cov, i, k);
        }
    }
}

/* R: opt/data.h:104-113  sum of ALL stored weights (no validity filter) */
static float sum_weights(int N, const float* weight) {
    float lanes[8] = {0, 0, 0, 0, 0, 0, 0, 0}; /* packets of 8, then hsum */
    for (int n = 0; n < N; ++n) lanes[n & 7] += weight[n];
    return ((lanes[0] + lanes[1]) + (lanes[2] + lanes[3])) + ((lanes[4] + lanes[5]) + (lanes[6] + lanes[7]));
}

/* R: opt/em.h:330-351 */
static int normalize_stats(oracle_cell* c, int N, const float* weight) {
    const int K = c->K;
    float eta = em_eta(c);
    float weight_sum = sum_weights(N, weight);
    c->total_weight = (1 - eta) * c->total_weight + eta * weight_sum;
    float rcp_weight_sum = 1 / c->total_weight;
    if (!isfinite(rcp_weight_sum)) return 0;
    if (c->total_weight == 0) return 0;
    for (int k = 0; k < K; ++k) {
        c->sn_w[k] = rcp_weight_sum * c->st_w[k];
        for (int i = 0; i < 5; ++i) P(c->sn_mean, i, k) = rcp_weight_sum * P(c->st_mean, i, k);
        for (int i = 0; i < 25; ++i) P(c->sn_cov, i, k) = rcp_weight_sum * P(c->st_cov, i, k);
    }
    return 1;
}

/* R: spaces/spatio_directional.h:173-245  from_jacobian: exp map of a tangent vector in the
 * CURRENT frame + 6x5 Jacobian (row-major). */
static void joint_from_jacobian(const float mean[6], const float from[9], const float t[5],
                                float embedded[6], float jac[30]) {
    const float cos_angle_min = -0.98f;
    float dj[6] = {0, 0, 0, 0, 0, 0}; /* 3x2 */
    memset(jac, 0, sizeof(float) * 30);
    float tx = t[3], ty = t[4];
    float length_sqr = tx * tx + ty * ty;
    float length = sqrtf(length_sqr);
    float sin_angle = sinf(length), cos_angle = cosf(length);
    int degenerate = (sin_angle < 1e-2f) || (cos_angle < cos_angle_min);
    float sinc_angle = degenerate ? 1.f : sin_angle / length;
    float cmsol = degenerate ? 0.f : (cos_angle - sinc_angle) / length_sqr;
    dj[0 * 2 + 0] = sinc_angle + tx * tx * cmsol;
    dj[1 * 2 + 1] = sinc_angle + ty * ty * cmsol;
    float off = tx * ty * cmsol;
    dj[0 * 2 + 1] = off;
    dj[1 * 2 + 0] = off;
    dj[2 * 2 + 0] = degenerate ? 0.f : -tx * sinc_angle;
    dj[2 * 2 + 1] = degenerate ? 0.f : -ty * sinc_angle;
    float fdj[6];
    matmul(3, 3, 2, from, dj, fdj); /* :217 directional_jacobian = from * directional_jacobian */
    for (int d = 0; d < 3; ++d) jac[d * 5 + d] = 1.f;
    for (int r = 0; r < 3; ++r)
        for (int cc = 0; cc < 2; ++cc) jac[(r + 3) * 5 + (cc + 3)] = fdj[r * 2 + cc];
    float local[3] = {tx * sinc_angle, ty * sinc_angle, cos_angle};
    for (int d = 0; d < 3; ++d) embedded[d] = t[d] + mean[d];
    for (int r = 0; r < 3; ++r)
        embedded[3 + r] = from[r * 3 + 0] * local[0] + from[r * 3 + 1] * local[1] + from[r * 3 + 2] * local[2];
}

/* R: spaces/spatio_directional.h:145-169  to_center_jacobian: 5x6, identity on position,
 * first two rows of `to` on direction. */
static void joint_to_center_jacobian(const float to[9], float jac[30]) {
    memset(jac, 0, sizeof(float) * 30);
    for (int d = 0; d < 3; ++d) jac[d * 6 + d] = 1.f;
    for (int r = 0; r < 2; ++r)
        for (int cc = 0; cc < 3; ++cc) jac[(r + 3) * 6 + (cc + 3)] = to[r * 3 + cc];
}

/* R: opt/em.h:353-488  update_model, one component */
static void update_model(oracle_cell* c, int k) {
    const int K = c->K;
    int it = c->iterations_run < 30 ? c->iterations_run : 30;
    float weight_prior_decay = (float)(1.0 / pow(3.0, (double)it));       /* :361-363 */
    float cov_prior_strength_decay = (float)(1.0 / pow(2.0, (double)it)); /* :364-366 */
    float weight_prior_decayed = c->weight_prior * weight_prior_decay;
    float cps_decayed = c->cov_prior_strength * cov_prior_strength_decay;

    float snw = c->sn_w[k];
    float snm[5], snc[25], upd_mean[5], upd_cov[25];
    for (int i = 0; i < 5; ++i) snm[i] = P(c->sn_mean, i, k);
    for (int i = 0; i < 25; ++i) snc[i] = P(c->sn_cov, i, k);

    float rcp_weight = 1.f / snw;
    float rcp_cov_weight = 1.f / (cps_decayed + snw);
    int non_zero = c->pmf[k] > 0;
    int finite_rcp = isfinite(rcp_weight);
    int good = non_zero && finite_rcp;

    float upd_weight = non_zero ? (finite_rcp ? snw + weight_prior_decayed : weight_prior_decayed) : 0.f;
    for (int i = 0; i < 5; ++i) upd_mean[i] = good ? snm[i] * rcp_weight : 0.f;

    for (int i = 0; i < 5; ++i)
        for (int j = 0; j < 5; ++j) {
            float mean_sub = (snm[i] * snm[j]) * rcp_weight;
            float cov_unnorm = snc[i * 5 + j] - mean_sub + P(c->cov_prior, i * 5 + j, k) * cps_decayed;
            float cov_norm = cov_unnorm * rcp_cov_weight;
            upd_cov[i * 5 + j] = good ? cov_norm + P(c->depth_prior, i * 5 + j, k) : P(c->cov, i * 5 + j, k);
        }
    c->pmf[k] = upd_weight; /* :469 */

    /* :470-476 re-centre the tangent space on the new mean and transport the covariance */
    float mean[6], from[9], to_new[9], embedded[6], fj[30], tj[30], J[25], Jt[25], tmp[25], out[25];
    for (int d = 0; d < 6; ++d) mean[d] = P(c->mean, d, k);
    for (int i = 0; i < 9; ++i) from[i] = P(c->from, i, k);
    float t5[5];
    for (int i = 0; i < 5; ++i) t5[i] = upd_mean[i];
    joint_from_jacobian(mean, from, t5, embedded, fj);
    joint_set_mean(c, k, embedded);
    for (int i = 0; i < 9; ++i) to_new[i] = P(c->to, i, k);
    joint_to_center_jacobian(to_new, tj);
    matmul(5, 6, 5, tj, fj, J);
    transpose(5, 5, J, Jt);
    matmul(5, 5, 5, J, upd_cov, tmp);
    matmul(5, 5, 5, tmp, Jt, out);
    for (int i = 0; i < 25; ++i) P(c->cov, i, k) = out[i];

    /* :478-487 carry the running statistics into the new tangent space */
    for (int i = 0; i < 5; ++i)
        for (int j = 0; j < 5; ++j) snc[i * 5 + j] -= (snm[i] * snm[j]) * rcp_weight;
    matmul(5, 5, 5, J, snc, tmp);
    matmul(5, 5, 5, tmp, Jt, out);
    for (int i = 0; i < 25; ++i) {
        P(c->sn_cov, i, k) = out[i];
        if (good) P(c->st_cov, i, k) = out[i] * c->total_weight;
    }
    if (good)
        for (int i = 0; i < 5; ++i) P(c->st_mean, i, k) = 0.f;
}

/* R: opt/em.h:218-253  em_step */
static void em_step(oracle_cell* c, int N, const float* point6, long stride, const float* weight,
                    float* out_loglik, int sum_mode) {
    compute_stats(c, N, point6, stride, weight, out_loglik, sum_mode);
    interpolate_stats(c);
    if (!normalize_stats(c, N, weight)) return; /* :237-240 step aborted, model untouched */
    for (int k = 0; k < c->K; ++k) update_model(c, k);
    (void)joint_prepare_vectorized(c); /* :250 (result only asserted) */
    ++c->iterations_run;
}

int oracle_em_step(oracle_cell* c, int N, const float* point6, long stride, const float* weight,
                   int n_iters, float* out_loglik, int sum_mode) {
    const unsigned csr = ftz_enter();
    for (int it = 0; it < n_iters; ++it)
        em_step(c, N, point6, stride, weight, out_loglik ? out_loglik + (size_t)it * N : NULL, sum_mode);
    ftz_leave(csr);
    c->conditioner_ready = 0;
    return 0;
}

int oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* one cell per task, as tev::ThreadPool::parallelFor at mitsuba/src/integrators/sdmm/volpath_sdmm.cpp:287-311 */
double oracle_em_batch(oracle_cell** cells, int C, const uint32_t* offsets, const float* point6,
                       long stride, const float* weight, int n_iters, int with_prepare, int n_threads) {
    double t0 = 0, t1 = 0;
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
    t0 = omp_get_wtime();
#pragma omp parallel for schedule(dynamic, 1)
#endif
    for (int ci = 0; ci < C; ++ci) {
        uint32_t b = offsets[ci], e = offsets[ci + 1];
        oracle_em_step(cells[ci], (int)(e - b), point6 + b, stride, weight + b, n_iters, NULL, 0);
        if (with_prepare) oracle_prepare_conditioner(cells[ci]);
    }
#ifdef _OPENMP
    t1 = omp_get_wtime();
#endif
    return t1 - t0;
}

/* ------------------------------------------------------------------------------------------ */
/* conditioner                                                                                 */
/* ------------------------------------------------------------------------------------------ */

/* R: distributions/sdmm_conditioner.h:89-100 prepare -> :170-212 prepare_vectorized per component
 * (+ :7-26 create_marginal), then marginal.weight.prepare(). */
void oracle_prepare_conditioner(oracle_cell* c) {
    const int K = c->K;
    for (int k = 0; k < K; ++k) {
        float cov[25], La[9], M[6], cc[4], Lc[4];
        for (int i = 0; i < 25; ++i) cov[i] = P(c->cov, i, k);
        oracle_schur(5, 3, cov, La, M, cc, Lc);
        c->m_pmf[k] = c->pmf[k];
        c->m_cdf[k] = c->cdf[k];
        for (int d = 0; d < 3; ++d) P(c->m_mean, d, k) = P(c->mean, d, k);
        for (int i = 0; i < 9; ++i) P(c->m_L, i, k) = La[i];
        c->m_inv_det[k] = 1.f / (La[0] * La[4] * La[8]);
        for (int i = 0; i < 6; ++i) P(c->M, i, k) = M[i];
        for (int i = 0; i < 4; ++i) { P(c->c_cov, i, k) = cc[i]; P(c->c_L, i, k) = Lc[i]; }
        c->c_inv_det[k] = 1.f / (Lc[0] * Lc[3]);
        /* :184-190 conditional frame from the joint's direction mean;
         * spaces/offset_directional.h:84-87 set_mean normalises */
        float m[3] = {P(c->mean, 3, k), P(c->mean, 4, k), P(c->mean, 5, k)};
        float len = sqrtf(m[0] * m[0] + m[1] * m[1] + m[2] * m[2]);
        float n[3] = {m[0] / len, m[1] / len, m[2] / len};
        float to[9], from[9];
        oracle_frame(n, to, from);
        for (int i = 0; i < 9; ++i) { P(c->c_to, i, k) = to[i]; P(c->c_from, i, k) = from[i]; }
    }
    (void)oracle_categorical_prepare(K, c->m_pmf, c->m_cdf); /* :98 */
    c->conditioner_ready = 1;
}

void oracle_get_conditioner(oracle_cell* c, float* out) {
    const int K = c->K;
    for (int k = 0; k < K; ++k) {
        for (int d = 0; d < 3; ++d) P(out, d, k) = P(c->m_mean, d, k);
        for (int i = 0; i < 9; ++i) P(out, 3 + i, k) = P(c->m_L, i, k);
        P(out, 12, k) = c->m_inv_det[k];
        P(out, 13, k) = c->m_pmf[k];
        for (int i = 0; i < 6; ++i) P(out, 14 + i, k) = P(c->M, i, k);
        for (int i = 0; i < 9; ++i) P(out, 20 + i, k) = P(c->c_to, i, k);
        for (int i = 0; i < 4; ++i) { P(out, 29 + i, k) = P(c->c_L, i, k); P(out, 34 + i, k) = P(c->c_cov, i, k); }
        P(out, 33, k) = c->c_inv_det[k];
    }
}

/* R: distributions/sdmm_conditioner.h:102-115 create_conditional -> :232-248 (Offset variant):
 * tangent_mean = mean_transform * (x - mu_a); pmf = marginal.posterior(x); then weight.prepare().
 * pmf/cdf/tm are caller scratch of K / K / 2K floats. */
static int create_conditional(const oracle_cell* c, const float x[3], float* pmf, float* cdf, float* tm) {
    const int K = c->K;
    for (int k = 0; k < K; ++k) {
        float t[3], z[3], L[9];
        for (int d = 0; d < 3; ++d) t[d] = x[d] - P(c->m_mean, d, k); /* spaces/euclidian.h:36-41 */
        tm[0 * K + k] = P(c->M, 0, k) * t[0] + P(c->M, 1, k) * t[1] + P(c->M, 2, k) * t[2];
        tm[1 * K + k] = P(c->M, 3, k) * t[0] + P(c->M, 4, k) * t[1] + P(c->M, 5, k) * t[2];
        for (int i = 0; i < 9; ++i) L[i] = P(c->m_L, i, k);
        oracle_solve_lower(3, L, t, z);
        float sq = z[0] * z[0] + z[1] * z[1] + z[2] * z[2];
        float pdf = c->m_inv_det[k] * 1.f * gaussian_normalization(3) * enoki_expf(-0.5f * sq);
        pmf[k] = pdf * c->m_pmf[k];
    }
    return oracle_categorical_prepare(K, pmf, cdf);
}

/* R: spaces/offset_directional.h:39-60  OffsetDirectionalTangentSpace::to */
static void cond_to(const float to[9], const float tm[2], const float w[3], float t[2], float* inv_jac) {
    const float cos_angle_min = -0.98f;
    float lx = to[0] * w[0] + to[1] * w[1] + to[2] * w[2];
    float ly = to[3] * w[0] + to[4] * w[1] + to[5] * w[2];
    float lz = to[6] * w[0] + to[7] * w[1] + to[8] * w[2];
    float cos_angle = lz;
    float angle = safe_acosf(cos_angle);
    float sin_angle = safe_sqrtf(1.f - cos_angle * cos_angle);
    float rcp_sinc = (sin_angle < 1e-4f) ? 1.f : angle / sin_angle;
    *inv_jac = (cos_angle < cos_angle_min) ? 0.f : rcp_sinc;
    t[0] = lx * rcp_sinc - tm[0];
    t[1] = ly * rcp_sinc - tm[1];
}

/* R: spaces/offset_directional.h:62-82  OffsetDirectionalTangentSpace::from */
static void cond_from(const float from[9], const float tm[2], const float t[2], float w[3], float* inv_jac) {
    const float max_length = PI_F - 1e-1f;
    float ox = t[0] + tm[0], oy = t[1] + tm[1];
    float length = sqrtf(ox * ox + oy * oy);
    float sin_angle = sinf(length), cos_angle = cosf(length);
    float sinc = ((sin_angle < 1e-2f) || length > max_length) ? 1.f : sin_angle / length;
    *inv_jac = (length > max_length) ? 0.f : sinc;
    float local[3] = {ox * sinc, oy * sinc, cos_angle};
    for (int r = 0; r < 3; ++r)
        w[r] = from[r * 3 + 0] * local[0] + from[r * 3 + 1] * local[1] + from[r * 3 + 2] * local[2];
}

/* R: distributions/sdmm.h:331-339 + :284-294 on the conditional (D = 2), summed over components
 * (mitsuba/src/integrators/sdmm/sdmm_proc.cpp:539-545 hsum_nested(posterior)). */
static float conditional_pdf(const oracle_cell* c, const float* pmf, const float* tm, const float w[3]) {
    const int K = c->K;
    float sum = 0.f;
    for (int k = 0; k < K; ++k) {
        float to[9], t[2], inv_jac, tmk[2] = {tm[0 * K + k], tm[1 * K + k]};
        for (int i = 0; i < 9; ++i) to[i] = P(c->c_to, i, k);
        cond_to(to, tmk, w, t, &inv_jac);
        float L[4] = {P(c->c_L, 0, k), P(c->c_L, 1, k), P(c->c_L, 2, k), P(c->c_L, 3, k)};
        float z[2];
        oracle_solve_lower(2, L, t, z);
        float sq = z[0] * z[0] + z[1] * z[1];
        float pdf = c->c_inv_det[k] * inv_jac * gaussian_normalization(2) * enoki_expf(-0.5f * sq);
        sum += pdf * pmf[k];
    }
    return sum;
}

/* R: distributions/sdmm.h:209-273 SDMM::sample on the conditional; uniforms BY ROLE:
 * u[0] component select, u[1] Box-Muller radius (first parameter of box_mueller_transform,
 * :199-207), u[2] Box-Muller angle.  Index search = enoki::binary_search over [0, K-1]
 * (ext/enoki/include/enoki/array_utils.h:150-170): first index whose cdf is not < u0, clamped
 * to K-1; then back off over zero-pmf components (:231-233). */
static uint32_t select_component(int K, const float* pmf, const float* cdf, float u0) {
    uint32_t start = 0, end = (uint32_t)K - 1;
    uint32_t iterations = 0;
    if (start < end) { /* log2i(end - start) + 1 */
        uint32_t v = end - start;
        while (v) { ++iterations; v >>= 1; }
    }
    for (uint32_t i = 0; i < iterations; ++i) {
        uint32_t middle = (start + end) >> 1;
        int cond = cdf[middle] < u0;
        if (cond) start = (middle + 1 < end) ? middle + 1 : end;
        else end = middle;
    }
    uint32_t idx = start;
    while (idx > 0 && pmf[idx] == 0) --idx;
    return idx;
}

static void sample_conditional(const oracle_cell* c, const float* pmf, const float* cdf, const float* tm,
                               const float u[3], float w[3], float* inv_jac, uint32_t* comp) {
    const int K = c->K;
    uint32_t k = select_component(K, pmf, cdf, u[0]);
    *comp = k;
    float radius = sqrtf(-2.f * logf(1.f - u[1]));
    float theta = 2.f * PI_F * u[2];
    float z0 = sinf(theta) * radius, z1 = cosf(theta) * radius;
    /* tangent = cov_sqrt[k] * z (:262) */
    float t[2] = {P(c->c_L, 0, k) * z0 + P(c->c_L, 1, k) * z1, P(c->c_L, 2, k) * z0 + P(c->c_L, 3, k) * z1};
    float from[9], tmk[2] = {tm[0 * K + k], tm[1 * K + k]};
    for (int i = 0; i < 9; ++i) from[i] = P(c->c_from, i, k);
    cond_from(from, tmk, t, w, inv_jac);
}

double oracle_sample_pdf(oracle_cell** cells, long n, const uint32_t* cell_idx, const float* pos3,
                         const float* u3, float* out_dir3, float* out_pdf, float* out_inv_jac,
                         uint32_t* out_comp, uint8_t* out_valid, int n_threads) {
    double t0 = 0, t1 = 0;
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
    t0 = omp_get_wtime();
#pragma omp parallel
#endif
    {
        float *pmf = NULL, *cdf = NULL, *tm = NULL;
        int cap = 0;
#ifdef _OPENMP
#pragma omp for schedule(static)
#endif
        for (long i = 0; i < n; ++i) {
            const oracle_cell* c = cells[cell_idx[i]];
            if (c->K > cap) {
                cap = c->K;
                pmf = (float*)realloc(pmf, sizeof(float) * cap);
                cdf = (float*)realloc(cdf, sizeof(float) * cap);
                tm = (float*)realloc(tm, sizeof(float) * 2 * cap);
            }
            float x[3] = {pos3[i], pos3[n + i], pos3[2 * n + i]};
            float u[3] = {u3[i], u3[n + i], u3[2 * n + i]};
            float w[3] = {0, 0, 0}, ij = 0.f, pdf = 0.f;
            uint32_t comp = 0;
            int valid = create_conditional(c, x, pmf, cdf, tm);
            if (valid) {
                sample_conditional(c, pmf, cdf, tm, u, w, &ij, &comp);
                pdf = conditional_pdf(c, pmf, tm, w);
            }
            out_dir3[i] = w[0]; out_dir3[n + i] = w[1]; out_dir3[2 * n + i] = w[2];
            out_pdf[i] = pdf; out_inv_jac[i] = ij; out_comp[i] = comp;
            if (out_valid) out_valid[i] = (uint8_t)valid;
        }
        free(pmf); free(cdf); free(tm);
    }
#ifdef _OPENMP
    t1 = omp_get_wtime();
#endif
    return t1 - t0;
}

double oracle_pdf(oracle_cell** cells, long n, const uint32_t* cell_idx, const float* pos3,
                  const float* dir3, float* out_pdf, uint8_t* out_valid, float* out_pmf, float* out_cdf,
                  int n_threads) {
    double t0 = 0, t1 = 0;
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
    t0 = omp_get_wtime();
#pragma omp parallel
#endif
    {
        float *pmf = NULL, *cdf = NULL, *tm = NULL;
        int cap = 0;
#ifdef _OPENMP
#pragma omp for schedule(static)
#endif
        for (long i = 0; i < n; ++i) {
            const oracle_cell* c = cells[cell_idx[i]];
            if (c->K > cap) {
                cap = c->K;
                pmf = (float*)realloc(pmf, sizeof(float) * cap);
                cdf = (float*)realloc(cdf, sizeof(float) * cap);
                tm = (float*)realloc(tm, sizeof(float) * 2 * cap);
            }
            float x[3] = {pos3[i], pos3[n + i], pos3[2 * n + i]};
            float w[3] = {dir3[i], dir3[n + i], dir3[2 * n + i]};
            int valid = create_conditional(c, x, pmf, cdf, tm);
            out_pdf[i] = valid ? conditional_pdf(c, pmf, tm, w) : 0.f;
            if (out_valid) out_valid[i] = (uint8_t)valid;
            if (out_pmf && out_cdf && n == 1) {
                memcpy(out_pmf, pmf, sizeof(float) * c->K);
                memcpy(out_cdf, cdf, sizeof(float) * c->K);
            }
        }
        free(pmf); free(cdf); free(tm);
    }
#ifdef _OPENMP
    t1 = omp_get_wtime();
#endif
    return t1 - t0;
}

/* ================================================================================================
 * sdmm::initialize                                                       R: opt/init.h:17-193
 * ================================================================================================
 * Called once per cell by initializeSDMMContext (mitsuba/src/integrators/sdmm/volpath_sdmm.cpp:
 * 132-137) with n_spatial = K/8 seeds and a COPY of the context's default-seeded PCG32. */

/* enoki::PCG32<float,1>                              R: ext/enoki/include/enoki/random.h:33-101 */
typedef struct { uint64_t state, inc; } oracle_pcg32;
static uint32_t pcg32_next_uint32(oracle_pcg32* r) {
    uint64_t oldstate = r->state;
    r->state = oldstate * 0x5851f42d4c957f2dULL + r->inc;
    uint32_t xorshifted = (uint32_t)(((oldstate >> 18) ^ oldstate) >> 27);
    uint32_t rot = (uint32_t)(oldstate >> 59);
    return (xorshifted >> rot) | (xorshifted << ((32u - rot) & 31u));
}
static void pcg32_seed(oracle_pcg32* r, uint64_t initstate, uint64_t initseq) {
    r->state = 0;
    r->inc = (initseq << 1) | 1u;
    pcg32_next_uint32(r);
    r->state += initstate;
    pcg32_next_uint32(r);
}
static float pcg32_next_float32(oracle_pcg32* r) {
    uint32_t w = (pcg32_next_uint32(r) >> 9) | 0x3f800000u;
    return w2f(w) - 1.f;
}
/* exposed so tests can pin the generator against the reference's */
void oracle_pcg32_floats(uint64_t state, uint64_t stream, int n, float* out) {
    oracle_pcg32 r;
    pcg32_seed(&r, state, stream);
    for (int i = 0; i < n; ++i) out[i] = pcg32_next_float32(&r);
}

/* sdmm::sample(Categorical&, RNG&)                         R: distributions/categorical.h:27-40 */
static uint32_t categorical_sample(int n, const float* pmf, const float* cdf, float u) {
    return select_component(n, pmf, cdf, u);
}

static float clamp_weight(float w) { /* enoki::min(enoki::max(w, 1e-3), 3)   R: init.h:64-66, 111-112 */
    double m = (double)w > 1e-3 ? (double)w : 1e-3;
    m = m < 3.0 ? m : 3.0;
    return (float)m;
}

/* enoki::pow(float, int 5): square-and-multiply in float  R: ext/enoki array_math.h:960-973 */
static float pow5f(float d) {
    float x2 = d * d;
    float x4 = x2 * x2;
    return d * x4;
}

int oracle_initialize(oracle_cell* c, int N, const float* point6, const float* normal3, long stride,
                      const float* weight, int n_spatial, float spatial_distance,
                      uint64_t rng_state, uint64_t rng_stream, int32_t* out_seed_idx) {
    const int K = c->K;
    const int n_thetas = 2, n_phis = 4;
    if (n_spatial * n_thetas * n_phis != K || N <= 0) return -1;
    oracle_pcg32 rng;
    pcg32_seed(&rng, rng_state, rng_stream);

    /* em = zero<EM>(K); sdmm.cov = 0                                       R: init.h:36-39 */
    memset(c->st_w, 0, sizeof(float) * K); memset(c->st_mean, 0, sizeof(float) * 5 * K);
    memset(c->st_cov, 0, sizeof(float) * 25 * K);
    memset(c->bt_w, 0, sizeof(float) * K); memset(c->bt_mean, 0, sizeof(float) * 5 * K);
    memset(c->bt_cov, 0, sizeof(float) * 25 * K);
    memset(c->sn_w, 0, sizeof(float) * K); memset(c->sn_mean, 0, sizeof(float) * 5 * K);
    memset(c->sn_cov, 0, sizeof(float) * 25 * K);
    memset(c->cov_prior, 0, sizeof(float) * 25 * K); memset(c->depth_prior, 0, sizeof(float) * 25 * K);
    memset(c->cov, 0, sizeof(float) * 25 * K);
    c->total_weight = 0.f; c->iterations_run = 0; c->samples_seen = 0;

    const float max_rad_sqr = (float)6.251388631170325;  /* quantile(chi2(3), 0.90)  R: init.h:40-44 */
    const float width_var = spatial_distance * spatial_distance / max_rad_sqr;
    const float depth_var = (float)(3e-2 * 3e-2 / (double)max_rad_sqr);
    const float directional_var = (float)(1.f * M_PI / 8.f);
    for (int k = 0; k < K; ++k) { P(c->cov, 18, k) = directional_var; P(c->cov, 24, k) = directional_var; }

    float* min_d = (float*)malloc(sizeof(float) * N);
    float* min_sp = (float*)malloc(sizeof(float) * N);
    float* min_nd = (float*)malloc(sizeof(float) * N);
    float* pmf = (float*)malloc(sizeof(float) * N);
    float* cdf = (float*)malloc(sizeof(float) * N);
    for (int j = 0; j < N; ++j) {
        min_d[j] = min_sp[j] = min_nd[j] = INFINITY;
        pmf[j] = clamp_weight(weight[j]);                                   /* R: init.h:64-66 */
    }
    oracle_categorical_prepare(N, pmf, cdf);                                /* R: init.h:68 */
    const float SPATIAL_THRESHOLD = (float)(2e-2 * 2e-2), NORMAL_THRESHOLD = (float)(2e-1 * 2e-1);

    int comp = 0;
    for (int sc = 0; sc < n_spatial; ++sc) {
        const uint32_t pi_ = categorical_sample(N, pmf, cdf, pcg32_next_float32(&rng));  /* R: init.h:77 */
        if (out_seed_idx) out_seed_idx[sc] = (int32_t)pi_;
        float mean[6];
        for (int d = 0; d < 6; ++d) mean[d] = point6[d * stride + pi_];
        const float n[3] = {normal3[pi_], normal3[stride + pi_], normal3[2 * stride + pi_]};
        float to[9], from[9];
        oracle_frame(n, to, from);                                          /* R: init.h:84-87 */
        const float s[3] = {from[0], from[3], from[6]}, t[3] = {from[1], from[4], from[7]};
        float manifold[9], dprior[9];
        for (int r = 0; r < 3; ++r)
            for (int q = 0; q < 3; ++q) {
                manifold[3 * r + q] = width_var * (s[r] * s[q]) + width_var * (t[r] * t[q]) + depth_var * (n[r] * n[q]);
                dprior[3 * r + q] = (float)((double)(n[r] * n[q]) * 1e-6);  /* R: init.h:88-92 */
            }
        for (int j = 0; j < N; ++j) {                                       /* R: init.h:96-126 */
            const float nj[3] = {normal3[j], normal3[stride + j], normal3[2 * stride + j]};
            const float ndot = n[0] * nj[0] + n[1] * nj[1] + n[2] * nj[2];
            const double a = (double)safe_acosf(ndot) / M_PI;
            const float normal_d = (float)(a * a);
            const float dx = mean[0] - point6[j], dy = mean[1] - point6[stride + j], dz = mean[2] - point6[2 * stride + j];
            const float position_d = dx * dx + dy * dy + dz * dz;
            const float d = normal_d + position_d;
            if (d < min_d[j]) {
                min_d[j] = d;
                pmf[j] = clamp_weight(weight[j]) * pow5f(d);
            }
            if (normal_d < NORMAL_THRESHOLD && position_d < min_sp[j]) {
                min_sp[j] = position_d;
                min_nd[j] = normal_d;
            }
            if (min_nd[j] < NORMAL_THRESHOLD && min_sp[j] < SPATIAL_THRESHOLD) {
                min_d[j] = 0;
                pmf[j] = 0;
            }
        }
        if (!oracle_categorical_prepare(N, pmf, cdf)) {                     /* R: init.h:127-139 */
            for (int j = 0; j < N; ++j) pmf[j] = 1.f / (float)N;
            oracle_categorical_prepare(N, pmf, cdf);
        }
        float theta = 0;
        for (int ti = 0; ti < n_thetas; ++ti) {                             /* R: init.h:141-174 */
            theta = (float)(theta + 0.5 * M_PI / ((float)n_thetas + 1.f));
            const float cos_theta = cosf(theta);
            const float sin_theta = safe_sqrtf(1.f - cos_theta * cos_theta);
            float phi = 0;
            for (int pi2 = 0; pi2 < n_phis; ++pi2) {
                phi = (float)(phi + 2 * M_PI / (float)n_phis);
                const float sin_phi = sinf(phi), cos_phi = cosf(phi);
                const float dl[3] = {sin_theta * cos_phi, sin_theta * sin_phi, cos_theta};
                float m6[6] = {mean[0], mean[1], mean[2], 0, 0, 0};
                for (int r = 0; r < 3; ++r) m6[3 + r] = from[3 * r] * dl[0] + from[3 * r + 1] * dl[1] + from[3 * r + 2] * dl[2];
                joint_set_mean(c, comp, m6);
                /* mean.coeff(3..5) of the NEXT lobe reuses `mean`, whose direction part was just
                 * overwritten (init.h:152-158) -- irrelevant, only the position part is read */
                for (int r = 0; r < 3; ++r)
                    for (int q = 0; q < 3; ++q) {
                        P(c->depth_prior, 5 * r + q, comp) = dprior[3 * r + q];
                        P(c->cov, 5 * r + q, comp) = manifold[3 * r + q];
                    }
                ++comp;
            }
        }
    }
    for (int k = 0; k < K; ++k) c->pmf[k] = 1.f / (float)K;                 /* R: init.h:177-179 */
    joint_prepare_vectorized(c);
    c->weight_prior = 1.f / (float)K;                                       /* R: init.h:181-192 */
    c->cov_prior_strength = 5.f / (float)K;
    const float cp[5] = {2e-3f, 2e-3f, 2e-3f, 2e-4f, 2e-4f};
    for (int k = 0; k < K; ++k)
        for (int i = 0; i < 5; ++i) P(c->cov_prior, 6 * i, k) = cp[i];
    c->conditioner_ready = 0;
    free(min_d); free(min_sp); free(min_nd); free(pmf); free(cdf);
    return K;
}
