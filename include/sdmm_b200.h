/* include/sdmm_b200.h -- the drop-in boundary of sdmm-b200: a thin C ABI over hand-written
 * sm_100a CUDA kernels for the data-parallel hot path of SDMM path guiding.
 *
 * The reference (anadodik/sdmm-mitsuba) has NO FFI for this path: it is a header-only C++17
 * template library (sdmm/include/sdmm) called directly from the Mitsuba `sdmm` integrator plugin.
 * Each entry point below therefore names the reference C++ interface it replaces; the C++ shim
 * with the reference's own signatures (sdmm_mitsuba_b200/csrc/host_shim.hpp) and the plugin-side
 * call sites are shown in INTEGRATION.md.  All paths are relative to /root/reference.
 *
 * Conventions
 *   - every function returns 0 on success, a negative SDMMB_E_* code on failure, never throws;
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream); calls are asynchronous
 *     on that stream unless they take/return HOST data, in which case they synchronise it;
 *   - pointers named h_* are host pointers, d_* device pointers on the context's device;
 *   - there is NO CPU fallback: every entry point needs a CUDA device.
 *
 * Flat layouts (float32, identical to oracle/ref_driver.cpp so parity tests move the same bytes)
 *   joint : [cell][32][K]  plane 0 = weight.pmf, 1..6 = tangent_space.mean (x,y,z,dx,dy,dz),
 *                          7+5r+c = cov(r,c)                       (sdmm/distributions/sdmm.h:118-124)
 *   em    : [cell][81][K]  plane 0 = stats.weight, 1..5 = stats.mean, 6+5r+c = stats.cov,
 *                          31+5r+c = cov_prior, 56+5r+c = depth_prior     (sdmm/opt/em.h:124-142)
 *   scal  : [cell][8] 32-bit words: f32 total_weight, i32 iterations_run, i32 samples_seen,
 *                          f32 weight_prior, f32 cov_prior_strength, f32 learning_rate, f32 alpha,
 *                          u32 status (SDMMB_ST_* bits, written by the library)
 *   samples: point6 = 6 planes (x,y,z,dx,dy,dz) of n_total floats with a common plane stride,
 *            weight = n_total floats, cell_offsets = n_cells+1 u32 (CSR over cells)
 *                                                                  (sdmm/opt/data.h:25-57)
 *   queries: cell_idx[n] u32, pos3 = 3 planes of n, u3 = 3 planes of n with u BY ROLE
 *            (plane 0: component select, plane 1: Box-Muller radius, plane 2: Box-Muller angle;
 *            sdmm/distributions/sdmm.h:220, 239-240 leave the order of the two Box-Muller draws
 *            to the compiler), dir3 = 3 planes of n.
 */
#ifndef SDMM_B200_H
#define SDMM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SDMMB_JOINT_PLANES 32
#define SDMMB_EM_PLANES 81
#define SDMMB_SCALAR_WORDS 8
#define SDMMB_QUERY_RECORD_FLOATS 32 /* per component, 30 used */
#define SDMMB_ESTEP_RECORD_PLANES 30 /* per component planes of the E-step record */
#define SDMMB_MAX_K 128

/* error codes */
#define SDMMB_OK 0
#define SDMMB_E_INVALID (-1)   /* bad argument */
#define SDMMB_E_CUDA (-2)      /* CUDA runtime error (see sdmmb_last_cuda_error) */
#define SDMMB_E_NOMEM (-3)     /* device allocation failed */
#define SDMMB_E_RANGE (-4)     /* cell range / sample count outside the context's capacity */
#define SDMMB_E_STATE (-5)     /* call order: e.g. em_step before set_samples */
#define SDMMB_E_NCCL (-6)      /* NCCL missing or returned an error */
#define SDMMB_E_DATA_FULL (-7) /* as Data::push_back's "Data full." (sdmm/opt/data.h:80) */

/* per-cell status bits (scal word 7); a failed step leaves the model untouched like
 * sdmm/opt/em.h:237-240 */
#define SDMMB_ST_STEP_ABORTED 1u  /* normalize_stats returned false (total_weight 0/non-finite) */
#define SDMMB_ST_CDF_INVALID 2u   /* Categorical::prepare: sum(pmf) <= 1e-20 (categorical.h:97-101) */
#define SDMMB_ST_NON_PSD 4u       /* some component failed the Cholesky PSD mask (cholesky.h:33) */
#define SDMMB_ST_COND_INVALID 8u  /* marginal weights invalid after conditioner prepare */

typedef struct sdmmb_ctx sdmmb_ctx;

/* Owns all device memory for `max_cells` cells of K components (K <= 128; internally padded to
 * 16/32/64/128 with zero-weight components).  Replaces the per-cell SDMMContext bundle
 * (sdmm/distributions/sdmm_context.h:11-45) owned by the STree. */
int sdmmb_create(sdmmb_ctx** out, int device, int max_cells, int K);
int sdmmb_destroy(sdmmb_ctx* ctx);
const char* sdmmb_get_error_string(int code);
const char* sdmmb_last_cuda_error(sdmmb_ctx* ctx);
int sdmmb_num_components(sdmmb_ctx* ctx);
int sdmmb_max_cells(sdmmb_ctx* ctx);

/* Install mixtures + EM state for cells [first_cell, first_cell+n_cells): what the integrator does
 * after sdmm::initialize / from_json (sdmm/distributions/sdmm.h:157-160): set_mean (frame) +
 * prepare_vectorized (Cholesky with +1e-5 pivots, CDF, pmf normalisation; sdmm.h:163-173).
 * h_em / h_scal may be NULL (zero statistics, learning_rate 0.2, alpha 0.5). */
int sdmmb_upload_cells(sdmmb_ctx* ctx, int first_cell, int n_cells, const float* h_joint,
                       const float* h_em, const uint32_t* h_scal, void* stream);
int sdmmb_download_cells(sdmmb_ctx* ctx, int first_cell, int n_cells, float* h_joint, float* h_em,
                         uint32_t* h_scal, void* stream);
/* Derived per-component state for parity checks, [cell][38][K] planes exactly as
 * sdmm_ref_cell_get_derived in oracle/ref_driver.cpp: 0..8 frame.to, 9..33 cov_sqrt, 34 inv_det,
 * 35 cdf, 36 pmf, 37 psd. */
int sdmmb_download_derived(sdmmb_ctx* ctx, int first_cell, int n_cells, float* h_out, void* stream);

/* Zero-copy form of sdmmb_set_samples: the E-step reads the caller's planes in place (7 coalesced loads
 * per sample per EM iteration instead of a packing pass + a second copy), as the reference trains on the
 * `data` it swapped into `training_data` without copying (volpath_sdmm.cpp:290).  The planes are BORROWED:
 * they must stay valid and unchanged until the next set_samples* / collect call on this context.  Same
 * arguments and weight rule (opt/data.h:20-23) as sdmmb_set_samples. */
int sdmmb_set_samples_borrowed(sdmmb_ctx* ctx, int first_cell, int n_cells, const float* d_point6,
                               size_t plane_stride, const float* d_weight,
                               const uint32_t* h_cell_offsets, void* stream);

/* Bind the training samples of cells [first_cell, first_cell+n_cells): replaces Data::push_back /
 * the data <-> training_data swap (sdmm/opt/data.h:59-102; volpath_sdmm.cpp:290).  The device
 * variant reads SoA planes already resident in HBM; the host variant copies them first (that copy
 * is what bench.py's e2e number includes).  Samples are re-laid out on the device into the
 * kernel's packed 32-byte records by a pack kernel.  h_cell_offsets is always a HOST array of
 * n_cells+1 entries, relative to the first sample passed. */
int sdmmb_set_samples(sdmmb_ctx* ctx, int first_cell, int n_cells, const float* d_point6,
                      size_t plane_stride, const float* d_weight, const uint32_t* h_cell_offsets,
                      void* stream);
int sdmmb_set_samples_host(sdmmb_ctx* ctx, int first_cell, int n_cells, const float* h_point6,
                           size_t plane_stride, const float* h_weight,
                           const uint32_t* h_cell_offsets, void* stream);

/* sdmm::initialize (sdmm/include/sdmm/opt/init.h:17-193) for cells [first_cell, first_cell+n_cells), as
 * initializeSDMMContext calls it for a cell that has no mixture yet (volpath_sdmm.cpp:132-137, 292-294):
 * K/8 spatial seeds drawn farthest-point style from the cell's samples (sampling weight clamp(w, 1e-3, 3)
 * x distance^5 in position + normal space), 2 x 4 directional lobes per seed, manifold covariance from the
 * seed's normal, depth prior 1e-6 n n^T, priors weight 1/K, strength 5/K, cov diag(2e-3 x3, 2e-4 x2);
 * EM state zeroed; then set_mean + prepare_vectorized.  K must be a multiple of 8.
 * d_point6 / d_normal3: SoA planes `plane_stride` floats apart; h_cell_offsets: HOST, n_cells+1, relative
 * to the first sample passed; h_spatial_distance: HOST, per cell (the integrator passes
 * 3 * max AABB diagonal / (K/8)).  rng_state / rng_stream: the PCG32 seed of the context's generator,
 * which the reference passes BY VALUE (every cell replays the same stream; the defaults are
 * SDMMB_PCG32_DEFAULT_STATE / _STREAM).  h_out_seed_idx (nullable): HOST, [n_cells][K/8] sample index of
 * every seed (relative to the cell's first sample).  Synchronous. */
#define SDMMB_PCG32_DEFAULT_STATE 0x853c49e6748fea9bULL
#define SDMMB_PCG32_DEFAULT_STREAM 0xda3e39cb94b95bdbULL
int sdmmb_initialize(sdmmb_ctx* ctx, int first_cell, int n_cells, const float* d_point6,
                     const float* d_normal3, size_t plane_stride, const float* d_weight,
                     const uint32_t* h_cell_offsets, const float* h_spatial_distance,
                     uint64_t rng_state, uint64_t rng_stream, int32_t* h_out_seed_idx, void* stream);
int sdmmb_initialize_host(sdmmb_ctx* ctx, int first_cell, int n_cells, const float* h_point6,
                          const float* h_normal3, size_t plane_stride, const float* h_weight,
                          const uint32_t* h_cell_offsets, const float* h_spatial_distance,
                          uint64_t rng_state, uint64_t rng_stream, int32_t* h_out_seed_idx, void* stream);

/* STree::create_child's inheritance (sdmm/accelerators/spatial_tree.h:187-196): when a leaf is split, both
 * children start from copies of the parent's mixture, conditioner and EM state.  Copies the complete
 * per-cell device state of cell h_src[i] into cell h_dst[i] (in order).  Synchronous. */
int sdmmb_copy_cells(sdmmb_ctx* ctx, int n, const int32_t* h_src, const int32_t* h_dst, void* stream);

/* n_iters x sdmm::em_step (sdmm/opt/em.h:218-253) on every bound cell of the range, i.e. the body
 * of the ThreadPool::parallelFor in SDMMVolumetricPathTracer::optimize (volpath_sdmm.cpp:287-311):
 * E-step + sufficient statistics (em.h:255-314), stepwise blend (:316-328), normalise (:330-351),
 * update_model (:353-488), prepare_vectorized (sdmm.h:163-173).  If with_conditioner != 0 the
 * last iteration also runs sdmm::prepare(conditioner, sdmm) (sdmm_conditioner.h:89-100).
 * d_out_loglik (nullable): [n_iters][n_total] log sum_k pi_k p_k(x_n) under the model ENTERING
 * each iteration, indexed like the bound samples.
 * n_iters = SDMMB_ITERS_INTEGRATOR applies the integrator's own rule PER CELL on the device
 * (volpath_sdmm.cpp:299-302: `n_iterations = (em.iterations_run < 4) ? 2 : 1`): two passes, the second of
 * which only touches the cells that entered the call with iterations_run < 4 (d_out_loglik then holds two
 * rows; the second row's entries of the other cells are not written). */
#define SDMMB_ITERS_INTEGRATOR (-1)
int sdmmb_em_step(sdmmb_ctx* ctx, int first_cell, int n_cells, int n_iters, int with_conditioner,
                  float* d_out_loglik, void* stream);

/* The per-iteration training call of the integrator for a whole batch of cells in ONE entry point
 * with HOST buffers: the body of optimize()'s parallelFor (volpath_sdmm.cpp:287-311) = bind the cells'
 * samples, n_iters x sdmm::em_step, sdmm::prepare(conditioner, sdmm), read the fitted mixtures back.
 * Internally pipelined over chunks of cells (H2D of chunk c+1 and D2H of chunk c-1 overlap the fit of
 * chunk c on separate streams), which needs PINNED host buffers to take effect; pageable buffers work
 * but serialise.  h_out_* are nullable, laid out like sdmmb_download_cells.  Synchronous. */
int sdmmb_optimize_host(sdmmb_ctx* ctx, int first_cell, int n_cells, const float* h_point6,
                        size_t plane_stride, const float* h_weight, const uint32_t* h_cell_offsets,
                        int n_iters, int with_conditioner, float* h_out_joint, float* h_out_em,
                        uint32_t* h_out_scal, void* stream);

/* sdmm::prepare(conditioner, joint) alone (sdmm_conditioner.h:89-100, 170-212). */
int sdmmb_prepare_conditioner(sdmmb_ctx* ctx, int first_cell, int n_cells, void* stream);
/* Conditioner internals for parity checks, [cell][38][K] planes exactly as
 * sdmm_ref_get_conditioner in oracle/ref_driver.cpp. */
int sdmmb_download_conditioner(sdmmb_ctx* ctx, int first_cell, int n_cells, float* h_out,
                               void* stream);

/* Fused guided query, one per path vertex, as SDMMRenderer::sampleSurface does it
 * (mitsuba/src/integrators/sdmm/sdmm_proc.cpp:361-372, 419-421, 491-501 -> 539-545):
 * create_conditional(x) (sdmm_conditioner.h:102-115, 232-248) -> SDMM::sample (sdmm.h:209-273)
 * -> posterior + hsum (sdmm.h:331-339).  out_valid[i] = 0 when create_conditional failed (all
 * outputs zero); out_inv_jac[i] == 0 marks a rejected sample (sdmm_proc.cpp:436-439).
 * `binned` != 0 promises cell_idx is sorted (vertices pre-binned by cell).  Otherwise the library bins
 * them on the device first (counting sort by cell, 3 small kernels) -- either way a CTA stages each
 * cell's records in shared memory once per run of vertices and handles two vertices per thread. */
int sdmmb_sample_pdf(sdmmb_ctx* ctx, size_t n, const uint32_t* d_cell_idx, const float* d_pos3,
                     const float* d_u3, float* d_out_dir3, float* d_out_pdf, float* d_out_inv_jac,
                     uint32_t* d_out_comp, uint8_t* d_out_valid, int binned, void* stream);
/* pdfSurface for a caller-supplied direction (sdmm_proc.cpp:510-590). */
int sdmmb_pdf(sdmmb_ctx* ctx, size_t n, const uint32_t* d_cell_idx, const float* d_pos3,
              const float* d_dir3, float* d_out_pdf, uint8_t* d_out_valid, int binned, void* stream);
/* Host-buffer variants (H2D of inputs, kernel, D2H of outputs inside the call). */
/* The same query on a PRUNED conditional: sdmm::create_conditional_pruned
 * (sdmm/include/sdmm/distributions/sdmm_conditioner.h:128-168) keeps the max_components strongest conditional
 * weights of the vertex (the weakest K - max_components are zeroed; preserve_idx >= 0 is never zeroed and takes
 * one place of the budget; -1 = none), re-normalises, then sample + pdf as above.  d_u3 != NULL: sample mode;
 * d_u3 == NULL: pdf of d_dir3 (then d_out_dir3 / d_out_inv_jac / d_out_comp may be NULL).  The reference defines
 * the function for product guiding with many components; no call site of its integrator uses it. */
int sdmmb_sample_pdf_pruned(sdmmb_ctx* ctx, size_t n, const uint32_t* d_cell_idx, const float* d_pos3,
                            const float* d_u3, const float* d_dir3, int max_components, int preserve_idx,
                            float* d_out_dir3, float* d_out_pdf, float* d_out_inv_jac, uint32_t* d_out_comp,
                            uint8_t* d_out_valid, void* stream);

/* Allocation-free variant of sdmmb_sample_pdf for vertices in ANY order (binned = 0).  It has no counterpart in
 * the reference and was added: callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it) get the same binning + query kernels
 * with the binning scratch in a file-scope static __device__ array instead of the per-stream grow-only buffer.
 * The call makes no cudaMalloc / cudaFree / allocating library call; inputs and outputs are the caller's.
 * Limits: n <= SDMMB_NOALLOC_MAX_VERTICES per call and a context of <= SDMMB_NOALLOC_MAX_CELLS cells, else
 * SDMMB_E_LIMIT (nothing is launched).  The array (4.8 MB) is held by every GPU that loads the library until the
 * process exits.  ALL calls, contexts and streams of a device share it: calls that overlap on the GPU race --
 * keep them on one stream or order them with events.  Nothing in it needs clearing by the caller (the histogram
 * words are zeroed on `stream` at the start of every call).  The first call per device looks the array's device
 * address up (cudaGetSymbolAddress, which may load the module): make one warm-up call before a graph capture.
 * Results are bit-identical to sdmmb_sample_pdf(..., binned = 0, ...). */
#define SDMMB_NOALLOC_MAX_VERTICES 65536
#define SDMMB_NOALLOC_MAX_CELLS 65536
#define SDMMB_E_LIMIT (-10)    /* beyond the fixed limits of sdmmb_sample_pdf_noalloc */
int sdmmb_sample_pdf_noalloc(sdmmb_ctx* ctx, size_t n, const uint32_t* d_cell_idx, const float* d_pos3,
                             const float* d_u3, float* d_out_dir3, float* d_out_pdf, float* d_out_inv_jac,
                             uint32_t* d_out_comp, uint8_t* d_out_valid, void* stream);

int sdmmb_sample_pdf_host(sdmmb_ctx* ctx, size_t n, const uint32_t* h_cell_idx, const float* h_pos3,
                          const float* h_u3, float* h_out_dir3, float* h_out_pdf,
                          float* h_out_inv_jac, uint32_t* h_out_comp, uint8_t* h_out_valid,
                          int binned, void* stream);
int sdmmb_pdf_host(sdmmb_ctx* ctx, size_t n, const uint32_t* h_cell_idx, const float* h_pos3,
                   const float* h_dir3, float* h_out_pdf, uint8_t* h_out_valid, int binned,
                   void* stream);

/* ---- product guiding -----------------------------------------------------------
 * The guided query with the conditional mixture multiplied by a learned-BSDF directional mixture, per
 * vertex, as SDMMRenderer::sampleSurface does it when sampleProduct is on
 * (mitsuba/src/integrators/sdmm/sdmm_proc.cpp:327-339, 375-381, 415-417, 476-487):
 * create_conditional(x) -> sdmm::product(conditional, learned_bsdf, product)
 * (sdmm/include/sdmm/distributions/sdmm_product.h:62-203, incl. prepare_vectorized of the result) ->
 * product.sample (sdmm.h:209-273) -> hsum(posterior(product, omega)) (sdmm.h:331-339).
 * h_bsdf: n_bsdf x 7 floats (pmf, mean x y z, cov00, cov10, cov11) -- the BSDF::DMM of
 * mitsuba/include/mitsuba/render/bsdf.h:301-304 in world space; components with pmf == 0 are skipped as
 * the reference does.  d_normal3 (nullable, 3 planes of n): component 0 is re-centred on the shading normal
 * (the diffuse branch, sdmm_proc.cpp:335-339).  d_frame9 + d_wi3 (both or neither; 9 planes: the shading
 * frame's to-world matrix [s t n] row-major; 3 planes: the incident direction in that frame): the glossy
 * branch (:340-355) -- the first packet of the BSDF mixture (components 0..7), given in the local frame, is
 * rotated about the local z axis by -atan2(wi.y, wi.x) (DirectionalTangentSpace::rotate_to_wo,
 * sdmm/spaces/directional.h:202-218) and taken to world space; it takes precedence over d_normal3.
 * out_comp is the index in the product mixture: (k / 8 * n_nonzero_bsdf + m) * 8 + k % 8 for conditional
 * component k and the m-th non-zero BSDF component.  out_valid = 0: invalid conditional or product
 * (prepare_vectorized failed); all outputs zero.  n_bsdf <= 16; the context's K must be a multiple of 8.
 * Synchronous on `stream`. */
int sdmmb_sample_pdf_product(sdmmb_ctx* ctx, size_t n, const uint32_t* d_cell_idx, const float* d_pos3,
                             const float* d_u3, const float* d_normal3, const float* d_frame9,
                             const float* d_wi3, int n_bsdf, const float* h_bsdf, float* d_out_dir3,
                             float* d_out_pdf, float* d_out_inv_jac, uint32_t* d_out_comp,
                             uint8_t* d_out_valid, void* stream);
/* pdfSurface with the product mixture for a caller-supplied direction (sdmm_proc.cpp:476-487). */
int sdmmb_pdf_product(sdmmb_ctx* ctx, size_t n, const uint32_t* d_cell_idx, const float* d_pos3,
                      const float* d_dir3, const float* d_normal3, const float* d_frame9,
                      const float* d_wi3, int n_bsdf, const float* h_bsdf, float* d_out_pdf,
                      uint8_t* d_out_valid, void* stream);

/* ---- spatial cell structure on the device ------------------------------------
 * The STree (sdmm/accelerators/spatial_tree.h:83-349) is built and split on the host by the reference's
 * own code; the host flattens it into this node table (host_shim.hpp: flatten_tree).  Inner node: axis in
 * {0,1,2}, split = aabb.min[axis] of children[0], child0/child1 = node indices.  Leaf: axis = -1,
 * child0 = the cell's device slot. */
typedef struct sdmmb_tree_node {
    int32_t axis;
    float split;
    uint32_t child0;
    uint32_t child1;
} sdmmb_tree_node;
int sdmmb_set_tree(sdmmb_ctx* ctx, int n_nodes, const sdmmb_tree_node* h_nodes);
/* STree::find (spatial_tree.h:19-51, called from sdmm_proc.cpp:311-315, 917-920) for a batch of points:
 * pos3 = 3 planes of n; out = the leaf's cell slot (0xffffffff if the walk fails). */
int sdmmb_find_cells(sdmmb_ctx* ctx, size_t n, const float* d_pos3, uint32_t* d_out_cell, void* stream);
int sdmmb_find_cells_host(sdmmb_ctx* ctx, size_t n, const float* h_pos3, uint32_t* h_out_cell, void* stream);
/* The renderer's sample collection (sdmm_proc.cpp:880-928: find the vertex's cell, Data::push_back under
 * the cell's mutex) for a whole batch of UNSORTED samples: find + counting sort by cell straight into the
 * E-step's packed records.  Binds the result as the training samples of ALL cells [0, max_cells), i.e. it
 * replaces sdmmb_set_samples.  Samples Data::push_back rejects (weight < 1e-8 or non-finite,
 * opt/data.h:20-23) and samples without a cell are dropped and counted.  d_key3 (nullable) = the
 * positions used for the cell lookup when they differ from the sample's own position planes.
 * h_out_counts (nullable): [max_cells] samples per cell.  Synchronises the stream once (counts). */
int sdmmb_collect_samples(sdmmb_ctx* ctx, size_t n, const float* d_point6, size_t plane_stride,
                          const float* d_weight, const float* d_key3, size_t key_stride,
                          uint32_t* h_out_counts, uint64_t* h_out_dropped, void* stream);
/* canBeOptimized (volpath_sdmm.cpp:140-149): cells with fewer bound samples than this are left untouched
 * by sdmmb_em_step (default 0 = train every cell of the range). */
int sdmmb_set_min_samples(sdmmb_ctx* ctx, uint32_t min_samples);

/* Hide the M-steps of sdmmb_em_step behind E-steps: with overlap enabled the bound range is fitted as two
 * halves whose E-steps alternate on two internal streams while every M-step (and the peer-memory push of
 * the last one) runs on a third, high-priority stream; the caller's stream forks into them and joins at the
 * end, so the call is still stream-ordered for the caller and the results are bit-identical.  The reference
 * gets the same effect from its thread pool: a cell's M-step runs while other cells are still in their
 * E-step (volpath_sdmm.cpp:287-311).  Off by default; sdmmb_last_estep_ms reports nothing while it is on. */
int sdmmb_set_em_overlap(sdmmb_ctx* ctx, int enabled);

/* optimizeAsync (the plugin's default, volpath_sdmm.cpp:180-242, 446-448): the fit of iteration i runs
 * while iteration i+1 is rendered with the PREVIOUS conditioners, which are swapped in by
 * optimize_async_wait_and_update (:227-242).  With async records enabled the fit (sdmmb_em_step /
 * sdmmb_optimize_host / sdmmb_prepare_conditioner) writes a back copy of the query records and queries
 * keep reading the published ones until sdmmb_publish_records copies a cell range over (stream-ordered).
 * Run the fit and the queries on different streams to overlap them. */
int sdmmb_set_async_records(sdmmb_ctx* ctx, int enabled);
int sdmmb_publish_records(sdmmb_ctx* ctx, int first_cell, int n_cells, void* stream);

/* Multi-GPU: cells are partitioned across ranks; after a training iteration every rank needs every
 * cell's query records.  The context can hold records for a GLOBAL cell range while fitting only
 * its own slice: records live in one device array of max_cells * K * 32 floats. */
int sdmmb_query_records(sdmmb_ctx* ctx, float** d_records, size_t* floats_per_cell);
/* ncclAllGather of the records of equal contiguous slices (cells_per_rank each; rank r owns
 * cells [r*cells_per_rank, (r+1)*cells_per_rank)).  nccl_comm is an ncclComm_t; libnccl is
 * resolved at run time (dlopen), so the library has no link-time NCCL dependency. */
int sdmmb_allgather_records(sdmmb_ctx* ctx, void* nccl_comm, int cells_per_rank, void* stream);

/* The same exchange WITHOUT a collective call: fused into the kernel that writes the records.
 * After sdmmb_peer_attach the last M-step of sdmmb_em_step(with_conditioner = 1) and
 * sdmmb_prepare_conditioner store every freshly written record straight into the record arrays of
 * all peers (NVLink / NVSwitch peer memory, CTA by CTA as the cells finish) and raise an arrival flag
 * there; sdmmb_peer_wait enqueues a device-side wait for the flags of all peers, after which queries
 * on that stream see every cell.  Protocol (SPMD): every rank issues the same sequence of pushing
 * calls; a rank's records are only overwritten once it has entered the same step (its ready flag is
 * raised stream-ordered at the start of its own pushing call), so queries of the previous step that
 * precede that call on the stream are safe.  Waits are bounded (SDMMB_PEER_TIMEOUT_MS, default
 * 10000): a timeout skips the stores, is counted, and shows in sdmmb_peer_status.
 * With async records enabled (sdmmb_set_async_records) nothing is pushed.
 *   h_handle : SDMMB_PEER_HANDLE_BYTES, opaque (CUDA IPC handles + same-process pointers);
 *   h_handles: world * SDMMB_PEER_HANDLE_BYTES, rank r's handle at r (exchange them with any
 *              host-side all-gather: torch.distributed, MPI, a pipe).
 * This replaces the reference's conditioner swap loop for the multi-GPU case
 * (mitsuba/src/integrators/sdmm/volpath_sdmm.cpp:227-242 publishes refitted conditioners to the
 * render threads; here the "render threads" are the other GPUs). */
#define SDMMB_PEER_HANDLE_BYTES 192
#define SDMMB_E_PEER (-9)      /* peer memory not reachable (no P2P / IPC open failed) */
int sdmmb_peer_export(sdmmb_ctx* ctx, unsigned char* h_handle);
int sdmmb_peer_attach(sdmmb_ctx* ctx, int rank, int world, const unsigned char* h_handles);
int sdmmb_peer_detach(sdmmb_ctx* ctx);
int sdmmb_peer_wait(sdmmb_ctx* ctx, void* stream);
int sdmmb_peer_status(sdmmb_ctx* ctx, uint32_t* out_epoch, uint32_t* out_timeouts, void* stream);

/* Pinned (page-locked) host memory for the host-buffer entry points (sdmmb_optimize_host, sdmmb_*_host): their
 * chunk pipelines only overlap copies with kernels when the buffers are pinned.  The reference keeps its
 * samples in enoki DynamicArrays (opt/data.h:25-57); the C++ shim gathers them into such a block. */
int sdmmb_host_alloc(void** out, size_t bytes);
int sdmmb_host_free(void* p);

/* Threading (the reference queries from concurrent LocalWorker threads with thread_local scratch,
 * sdmm_proc.cpp:1086-1106, while optimize() fits, volpath_sdmm.cpp:180-242): every entry point locks its
 * context, so calls from any number of host threads are safe.  Device-pointer queries enqueued by different
 * threads on different streams run concurrently on the device (binning scratch is per stream); host-buffer
 * calls hold the context for their duration.  Queries concurrent with a fit need the double-buffered records
 * (sdmmb_set_async_records), exactly like the reference's optimizeAsync. */

/* Introspection used by bench.py */
int sdmmb_launch_count(sdmmb_ctx* ctx, uint64_t* out);           /* kernels launched by this ctx */
int sdmmb_last_estep_ms(sdmmb_ctx* ctx, float* total_ms, int* n_launches); /* CUDA-event time of
                                                                    the E-step launches of the last
                                                                    sdmmb_em_step (syncs stream) */
/* On-device micro-benchmarks that define the FP32 / SFU roofline denominators on THIS GPU:
 * out[0] = scalar FFMA TFLOP/s, out[1] = packed FFMA2 (fma.rn.f32x2) TFLOP/s,
 * out[2] = MUFU.EX2 Gop/s, out[3] = SM clock estimate in MHz during the FFMA2 run. */
int sdmmb_measure_peaks(int device, float* out4);

#ifdef __cplusplus
}
#endif
#endif /* SDMM_B200_H */
