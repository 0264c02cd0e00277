/* oracle/sdmm_oracle.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C CPU restatement ("port") of the reference algorithm on the SDMM hot path
 * (anadodik/sdmm-mitsuba, sdmm submodule @ 7fd406d).  Every function in sdmm_oracle.c cites the
 * reference file:line it follows.  Parity status: PINNED -- checked (tests/test_oracle_*.py)
 * against (a) the known-answer literals of the reference's own doctest files and (b) the
 * reference's own headers compiled here as oracle/_ref/libsdmm_ref.so on identical inputs, plus
 * committed fixtures under tests/golden/ generated from (b).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this library.
 * The product (sdmm_mitsuba_b200/, include/sdmm_b200.h) never links or calls it.
 *
 * Flat layouts (identical to oracle/ref_driver.cpp and include/sdmm_b200.h):
 *   joint : 32 planes of K floats: [0]=pmf, [1..6]=mean (x,y,z,dx,dy,dz), [7+5r+c]=cov(r,c)
 *   em    : 81 planes of K floats: [0]=stats.weight, [1..5]=stats.mean, [6+5r+c]=stats.cov,
 *           [31+5r+c]=cov_prior, [56+5r+c]=depth_prior
 *   scal  : 8 x 32-bit words: f32 total_weight, i32 iterations_run, i32 samples_seen,
 *           f32 weight_prior, f32 cov_prior_strength, f32 learning_rate, f32 alpha, i32 status
 */
#ifndef SDMM_ORACLE_H
#define SDMM_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct oracle_cell oracle_cell;

oracle_cell* oracle_cell_create(int K);
void oracle_cell_destroy(oracle_cell*);
int oracle_cell_set(oracle_cell*, const float* joint, const float* em, const uint32_t* scal);
void oracle_cell_get(oracle_cell*, float* joint, float* em, uint32_t* scal);
/* planes as sdmm_ref_cell_get_derived: [0..8] to, [9..33] cov_sqrt, [34] inv_det, [35] cdf,
 * [36] pmf, [37] psd */
void oracle_cell_get_derived(oracle_cell*, float* out);
/* sum_mode: 0 = float32 sequential accumulation (as the reference), 1 = float64 accumulators
 * for the sufficient statistics (separates ordering error from algorithmic error). */
int oracle_em_step(oracle_cell*, int N, const float* point6, long stride, const float* weight,
                   int n_iters, float* out_loglik, int sum_mode);
double oracle_em_batch(oracle_cell** cells, int C, const uint32_t* offsets, const float* point6,
                       long stride, const float* weight, int n_iters, int with_prepare,
                       int n_threads);
void oracle_prepare_conditioner(oracle_cell*);
/* planes as sdmm_ref_get_conditioner */
void oracle_get_conditioner(oracle_cell*, float* out);
double oracle_sample_pdf(oracle_cell** cells, long n, const uint32_t* cell_idx, const float* pos3,
                         const float* u3, float* out_dir3, float* out_pdf, float* out_inv_jac,
                         uint32_t* out_comp, uint8_t* out_valid, int n_threads);
double oracle_pdf(oracle_cell** cells, long n, const uint32_t* cell_idx, const float* pos3,
                  const float* dir3, float* out_pdf, uint8_t* out_valid, float* out_pmf,
                  float* out_cdf, int n_threads);

/* generic small-matrix pieces (row-major, D <= 8), exposed for the known-answer tests */
int oracle_cholesky(int D, const float* in, float* out);
void oracle_solve_lower(int D, const float* L, const float* b, float* x);
void oracle_inverse_lower_tri(int D, const float* L, float* inv);
float oracle_gauss_pdf_euclid(int D, const float* mean, const float* cov, const float* x);
/* Schur-complement decomposition of a DxD covariance into marginal (first Da dims) and
 * conditional (remaining Db = D-Da): outputs L_a (DaxDa), M = cov_ba cov_aa^-1 (DbxDa),
 * cond_cov (DbxDb), L_c (DbxDb). */
void oracle_schur(int D, int Da, const float* cov, float* L_a, float* M, float* cond_cov, float* L_c);
int oracle_categorical_prepare(int n, float* pmf, float* cdf);
void oracle_frame(const float* n, float* to, float* from);
void oracle_joint_to(const float* mean6, const float* point6, float* tangent5, float* inv_jac);
int oracle_max_threads(void);
/* sdmm::initialize (opt/init.h:17-193) with the default-seeded PCG32 copy of initializeSDMMContext;
 * out_seed_idx (may be NULL): the n_spatial sample indices drawn as seeds.  Returns K or -1. */
int oracle_initialize(oracle_cell*, int N, const float* point6, const float* normal3, long stride,
                      const float* weight, int n_spatial, float spatial_distance,
                      uint64_t rng_state, uint64_t rng_stream, int32_t* out_seed_idx);
void oracle_pcg32_floats(uint64_t state, uint64_t stream, int n, float* out);

#ifdef __cplusplus
}
#endif
#endif
