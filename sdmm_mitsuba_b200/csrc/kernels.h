// kernels.h -- device data layouts and launcher declarations shared by the kernel translation
// units and the C ABI (capi.cu).  Internal; the public boundary is include/sdmm_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace sdmmb {

constexpr int JOINT_PLANES = 32;
constexpr int EM_PLANES = 81;
constexpr int SCALAR_WORDS = 8;

// E-step record: 30 planes of K floats per cell ([cell][plane][K], lanes read consecutive k).  The E-step
// stages a sample's position relative to the cell's centre `cen` (the mixture's weighted mean position):
// z = W (x - mu, t_dir) = W_pos (x - cen) + b + W_dir t_dir.
constexpr int EREC_B = 0;     // 5: bias b_r = -sum_{c<3} W_rc (mu_c - cen_c)
constexpr int EREC_ROT = 5;   // 6: the Duff frame in factored form: n.x, n.y, -n.z, a = -1/(sign + n.z),
                              //    sign * n.x, sign * n.y  (sign = copysign(1, n.z)); see estep_pair()
constexpr int EREC_W = 11;    // 15: sqrt(log2(e)/2) * L^-1, lower triangle row-major; the (4,4) entry
                              //    carries the factor `sign` because the kernel works with sign * t_4
constexpr int EREC_NLC = 26;  // 1: pmf * inv_cov_sqrt_det * (2 pi)^-5/2
constexpr int EREC_REG_PLANES = 27;  // planes a lane keeps in registers
constexpr int EREC_CEN = 27;  // 3: the cell's centre (the same value in all K lanes)
constexpr int EREC_PLANES = 30;

// E-step partial statistics per item: 21 planes of K floats, position entries UN-centred:
//   with v = (x - cen, t_dir):  0 = sum g, 1..5 = sum g v_i, 6 + i(i+1)/2 + j = sum g v_i v_j (i >= j);
//   the M-step re-centres them about the component mean (t_pos = x - mu)
constexpr int ESTAT_PLANES = 21;

// Query record: 32 floats per component, [cell][k][32] (one 128-byte line per component)
constexpr int QREC_MU = 0;     // 3: marginal mean
constexpr int QREC_WA = 3;     // 6: sqrt(log2(e)/2) * L_aa^-1 lower triangle (00,10,11,20,21,22)
constexpr int QREC_NLCA = 9;   // 1: -log2(pmf_marginal * inv_det_aa * (2 pi)^-3/2)
constexpr int QREC_M = 10;     // 6: mean_transform 2x3 row-major
constexpr int QREC_R = 16;     // 9: conditional frame.to row-major (from = transpose)
constexpr int QREC_LC = 25;    // 3: conditional cov_sqrt (00,10,11)
constexpr int QREC_WC = 28;    // 3: sqrt(log2(e)/2) * L_c^-1 (00,10,11)
constexpr int QREC_NLCC = 31;  // 1: -log2(inv_det_c * (2 pi)^-1)
constexpr int QREC_FLOATS = 32;

constexpr uint32_t ST_STEP_ABORTED = 1u, ST_CDF_INVALID = 2u, ST_NON_PSD = 4u, ST_COND_INVALID = 8u;

constexpr int MSTEP_THREADS = 128;

// One E-step work item = a contiguous run of one cell's samples, processed by one CTA.
struct EItem {
    uint32_t cell;   // absolute cell index (into erec)
    uint32_t begin;  // absolute sample index range in the packed sample array
    uint32_t end;
    uint32_t pad;
};

struct EStepLaunch {
    int K;
    int n_items;
    const float4* samples;  // packed records, 2 float4 per sample
    // borrowed SoA planes instead of packed records (sdmmb_set_samples_borrowed): 6 planes of the
    // point with `soa_stride` floats between them and the weight plane; null = use `samples`
    const float* soa_point6;
    size_t soa_stride;
    const float* soa_weight;
    const EItem* items;
    const float* erec;
    float* partials;   // [n_items][21][K]
    float* item_scal;  // [n_items][2]: sum of weights, skipped-sample count (int bits)
    float* out_loglik; // nullable, indexed by absolute sample index
    // the integrator's iteration rule (volpath_sdmm.cpp:299-302: two em_steps while em.iterations_run < 4, else
    // one) evaluated per cell on the device: in the rule's second pass the items of cells whose iterations_run
    // has reached `skip_iterations_ge` do nothing.  0 = every item runs.
    const uint32_t* scal;
    uint32_t skip_iterations_ge;
};

// Multi-GPU record exchange fused into K3 (sdmmb_peer_attach): the CTA that wrote the query records
// of its cells stores them straight into every peer's record array over NVLink, then the launch's
// last CTA raises this rank's arrival flag on every peer.  Flag words of one context (uint32):
//   [0 .. 7]  ready[r]   : rank r has entered training step `epoch` (its queries of the previous
//                          step are done, so its records may be overwritten)        -- written by r
//   [8 .. 15] arrived[r] : rank r's records of step `epoch` are complete here       -- written by r
//   [16] CTA counter of the pushing launch
//   [17] error count: bounded waits of THIS rank that timed out, plus one per step in which a PEER gave up
//        pushing its records here (written by that peer: the records of its cells are stale on this rank)
//   [18] "some CTA of the pushing launch timed out waiting for a peer's ready flag" (reset by the last CTA)
constexpr int PEER_MAX = 8;
constexpr int PEER_FLAG_WORDS = 32;
constexpr int PEER_READY = 0, PEER_ARRIVED = 8, PEER_COUNTER = 16, PEER_ERROR = 17, PEER_FAIL = 18;
struct PeerPush {
    int world;                       // 0 = disabled
    int rank;
    uint32_t epoch;                  // this launch's training step (1, 2, ...)
    uint32_t timeout_ms;
    float* qrec[PEER_MAX];           // every rank's record array (entry `rank` = the local one)
    uint32_t* flags[PEER_MAX];       // every rank's flag words
};

struct MStepArgs {
    int K;
    int first_cell, n_cells;
    int mode;  // bit0 EM update, bit1 conditioner records, bit2 normalise pmf (upload)
    float* joint;
    float* em;
    uint32_t* scal;
    float* cdf;
    float* erec;
    float* qrec;
    const float* partials;
    const float* item_scal;
    const uint32_t* cell_item_begin;    // [n_cells+1] absolute item indices, entry 0 = this launch's first cell
    uint32_t item_base;                 // item index that `partials` / `item_scal` start at
    const uint32_t* cell_sample_count;  // [n_cells]
    uint32_t min_samples;               // cells with fewer bound samples are left untouched by the EM update
    uint32_t skip_iterations_ge;        // != 0: cells whose iterations_run >= this skip the EM update (see EStepLaunch)
    PeerPush peer;                      // world == 0: no peer stores
};

struct QueryLaunch {
    int K;
    size_t n;
    const float* qrec;
    const uint32_t* cell_idx;
    const float* pos3;   // 3 planes of n
    const float* u3;     // 3 planes of n (sample mode) or nullptr
    const float* dir3;   // 3 planes of n (pdf-only mode) or nullptr
    float* out_dir3;
    float* out_pdf;
    float* out_inv_jac;
    uint32_t* out_comp;
    uint8_t* out_valid;
    int binned;
    int max_cells;
    uint32_t* scratch;  // unbinned only: 2 * (max_cells + 1) + 17 * n + 4 words (see launch_query)
};

// flattened STree node (tree_kernels.cu): inner node: axis 0..2, split = aabb.min[axis] of child 0;
// leaf: axis = -1, child0 = cell slot
struct TreeNode {
    int32_t axis;
    float split;
    uint32_t child0;
    uint32_t child1;
};

// sdmm::initialize on the device (init_kernels.cu): one CTA per cell
struct InitArgs {
    int K, K_user;
    int first_cell, n_cells;
    const float* point6;    // 6 planes, `stride` floats apart
    const float* normal3;   // 3 planes, same stride
    const float* weight;
    size_t stride;
    const uint32_t* offsets;          // device, [n_cells+1] sample ranges (relative to the plane starts)
    const float* spatial_distance;    // device, [n_cells]
    uint64_t rng_state, rng_stream;   // PCG32 seed every cell's COPY of the generator starts from
    float* scratch;                   // 5 floats per sample
    float* joint;
    float* em;
    uint32_t* scal;
    int32_t* out_seed_idx;            // nullable, device, [n_cells][K_user/8]
};
cudaError_t launch_init_cells(const InitArgs& a, cudaStream_t st);
// batched cell-to-cell copy of all per-cell state (sdmmb_copy_cells): pairs = [n][2] device words (source, destination)
struct CopyCellsArgs {
    const int32_t* pairs;
    int n_arrays;
    uint32_t* base[8];
    size_t floats_per_cell[8];
};
cudaError_t launch_copy_cells(const CopyCellsArgs& a, int n_pairs, cudaStream_t st);

cudaError_t launch_find_cells(const TreeNode* nodes, int n_nodes, const float* pos3, size_t stride, size_t n,
                              uint32_t* out, uint32_t miss, cudaStream_t st);
cudaError_t launch_collect_count(const TreeNode* nodes, int n_nodes, const float* key3, size_t key_stride,
                                 const float* weight, size_t n, uint32_t n_buckets, uint32_t* cell, uint32_t* count,
                                 uint32_t* cursor, cudaStream_t st);
cudaError_t launch_collect_scatter(const uint32_t* cell, const float* point6, size_t stride, const float* weight,
                                   size_t n, uint32_t dropped, uint32_t* cursor, float4* rec, cudaStream_t st);
cudaError_t launch_pack(const float* point6, size_t stride, const float* weight, float4* rec, size_t n,
                        cudaStream_t st);
cudaError_t launch_estep(const EStepLaunch& L, cudaStream_t st);
int estep_ctas_per_sm(int K);  // resident E-step CTAs per SM of the shape launch_estep picks
cudaError_t launch_mstep(const MStepArgs& a, cudaStream_t st);
// raises ready[rank] = epoch on every peer (stream-ordered after this rank's earlier work)
cudaError_t launch_peer_signal(const PeerPush& p, cudaStream_t st);
// device-side wait until arrived[r] >= epoch for every peer r (bounded by timeout_ms)
cudaError_t launch_peer_wait(const PeerPush& p, cudaStream_t st);
cudaError_t launch_derived_readout(const float* joint, const float* cdf, int first_cell, int n_cells, int K,
                                   float* out, cudaStream_t st);
cudaError_t launch_conditioner_readout(const float* joint, const float* qrec, int first_cell, int n_cells,
                                       int K, float* out, cudaStream_t st);
cudaError_t launch_query(const QueryLaunch& q, cudaStream_t st);
// create_conditional_pruned + sample + pdf, one thread per vertex (u3 != null: sample mode, else dir3)
cudaError_t launch_query_pruned(const QueryLaunch& q, int K_user, int max_components, int preserve_idx, cudaStream_t st);
// counting sort of vertices by cell (query_kernels.cu); scratch: 2 * (max_cells + 1) + 9 * n + 4 words
cudaError_t launch_bin_vertices(const uint32_t* cell_idx, size_t n, int max_cells, const float* pos3,
                                const float* in3, uint32_t* scratch, float4** out_qin, uint32_t** out_invperm,
                                cudaStream_t st);

// product guiding (product_kernels.cu): conditional x learned-BSDF mixture, one thread per vertex
constexpr int SDMMB_PRODUCT_MAX_BSDF = 16;
struct ProductLaunch {
    int K, K_user;
    int max_cells;
    size_t n;
    const float* qrec;
    const uint32_t* cell_idx;
    const float* pos3;     // 3 planes of n
    const float* u3;       // sample mode, or nullptr
    const float* dir3;     // pdf-only mode, or nullptr
    const float* normal3;  // nullable: BSDF component 0 is re-centred on it (diffuse surfaces)
    const float* frame9;   // nullable, 9 planes of n: shading frame to-world matrix row-major (glossy surfaces)
    const float* wi3;      // nullable, 3 planes of n: incident direction in the local shading frame
    int M;
    const float* bsdf;     // device, [M][7]: pmf, mean3, c00, c10, c11
    uint32_t* scratch;     // binning scratch: 2 * (max_cells + 1) + 9 * n + 4 words
    float* out_dir3;
    float* out_pdf;
    float* out_inv_jac;
    uint32_t* out_comp;
    uint8_t* out_valid;
};
cudaError_t launch_product_query(const ProductLaunch& q, cudaStream_t st);
cudaError_t measure_peaks(float* out4);

}  // namespace sdmmb
