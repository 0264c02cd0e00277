// capi_init.cu -- sdmm::initialize on the device (row F3 of SURVEY.md section 8f; sdmm/include/sdmm/opt/init.h:17-193)
// and cell copies (a split's children inherit the parent's mixture and EM state, spatial_tree.h:187-196).
#include "capi_internal.h"

using namespace sdmmb;
using namespace sdmmb_capi;

extern "C" {

int sdmmb_initialize(sdmmb_ctx* ctx, int first_cell, int n_cells, const float* d_point6,
                     const float* d_normal3, size_t plane_stride, const float* d_weight,
                     const uint32_t* h_cell_offsets, const float* h_spatial_distance,
                     uint64_t rng_state, uint64_t rng_stream, int32_t* h_out_seed_idx, void* stream) {
    if (!ctx || !d_point6 || !d_normal3 || !d_weight || !h_cell_offsets || !h_spatial_distance)
        return SDMMB_E_INVALID;
    if (!range_ok(ctx, first_cell, n_cells)) return SDMMB_E_RANGE;
    if (ctx->K_user % 8 != 0 || ctx->K_user / 8 > 16) return SDMMB_E_INVALID;  // 2 x 4 lobes per seed
    if (n_cells == 0) return SDMMB_OK;
    for (int c = 0; c < n_cells; ++c)
        if (h_cell_offsets[c + 1] < h_cell_offsets[c]) return SDMMB_E_INVALID;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const size_t n_total = h_cell_offsets[n_cells];
    const int n_spatial = ctx->K_user / 8;
    // staging: offsets | spatial_distance | seed indices | 5 scratch floats per sample
    const size_t off_bytes = ((size_t)(n_cells + 1) * 4 + 255) / 256 * 256;
    const size_t sd_bytes = ((size_t)n_cells * 4 + 255) / 256 * 256;
    const size_t seed_bytes = ((size_t)n_cells * n_spatial * 4 + 255) / 256 * 256;
    const size_t total = off_bytes + sd_bytes + seed_bytes + n_total * 5 * 4;
    int rc = ensure(ctx, &ctx->d_stage2, &ctx->stage2_cap, total);
    if (rc != SDMMB_OK) return rc;
    char* base = (char*)ctx->d_stage2;
    uint32_t* d_off = (uint32_t*)base;
    float* d_sd = (float*)(base + off_bytes);
    int32_t* d_seed = (int32_t*)(base + off_bytes + sd_bytes);
    float* d_scratch = (float*)(base + off_bytes + sd_bytes + seed_bytes);
    CK(cudaMemcpyAsync(d_off, h_cell_offsets, (size_t)(n_cells + 1) * 4, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_sd, h_spatial_distance, (size_t)n_cells * 4, cudaMemcpyHostToDevice, st));
    InitArgs a{};
    a.K = ctx->K;
    a.K_user = ctx->K_user;
    a.first_cell = first_cell;
    a.n_cells = n_cells;
    a.point6 = d_point6;
    a.normal3 = d_normal3;
    a.weight = d_weight;
    a.stride = plane_stride;
    a.offsets = d_off;
    a.spatial_distance = d_sd;
    a.rng_state = rng_state;
    a.rng_stream = rng_stream;
    a.scratch = d_scratch;
    a.joint = ctx->d_joint;
    a.em = ctx->d_em;
    a.scal = ctx->d_scal;
    a.out_seed_idx = d_seed;
    CK(launch_init_cells(a, st));
    // set_mean + prepare_vectorized on the fresh mixtures (R: init.h:177-179), as after an upload
    MStepArgs m = base_mstep_args(ctx, first_cell, n_cells, 4);
    CK(launch_mstep(m, st));
    ctx->launches += 2;
    if (h_out_seed_idx)
        CK(cudaMemcpyAsync(h_out_seed_idx, d_seed, (size_t)n_cells * n_spatial * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return SDMMB_OK;
}

int sdmmb_initialize_host(sdmmb_ctx* ctx, int first_cell, int n_cells, const float* h_point6,
                          const float* h_normal3, size_t plane_stride, const float* h_weight,
                          const uint32_t* h_cell_offsets, const float* h_spatial_distance,
                          uint64_t rng_state, uint64_t rng_stream, int32_t* h_out_seed_idx, void* stream) {
    if (!ctx || !h_point6 || !h_normal3 || !h_weight || !h_cell_offsets) return SDMMB_E_INVALID;
    if (!range_ok(ctx, first_cell, n_cells)) return SDMMB_E_RANGE;
    if (n_cells == 0) return SDMMB_OK;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const size_t n = h_cell_offsets[n_cells];
    int rc = ensure(ctx, &ctx->d_stage, &ctx->stage_cap, n * 10 * 4);
    if (rc != SDMMB_OK) return rc;
    float* d = (float*)ctx->d_stage;
    for (int p = 0; p < 6; ++p)
        CK(cudaMemcpyAsync(d + (size_t)p * n, h_point6 + (size_t)p * plane_stride, n * 4, cudaMemcpyHostToDevice, st));
    for (int p = 0; p < 3; ++p)
        CK(cudaMemcpyAsync(d + (size_t)(6 + p) * n, h_normal3 + (size_t)p * plane_stride, n * 4,
                           cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d + 9 * n, h_weight, n * 4, cudaMemcpyHostToDevice, st));
    return sdmmb_initialize(ctx, first_cell, n_cells, d, d + 6 * n, n, d + 9 * n, h_cell_offsets,
                            h_spatial_distance, rng_state, rng_stream, h_out_seed_idx, stream);
}

int sdmmb_copy_cells(sdmmb_ctx* ctx, int n, const int32_t* h_src, const int32_t* h_dst, void* stream) {
    if (!ctx || n < 0 || (n && (!h_src || !h_dst))) return SDMMB_E_INVALID;
    for (int i = 0; i < n; ++i)
        if (!range_ok(ctx, h_src[i], 1) || !range_ok(ctx, h_dst[i], 1)) return SDMMB_E_RANGE;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const size_t K = (size_t)ctx->K;
    for (int i = 0; i < n; ++i) {
        const size_t s = (size_t)h_src[i], d = (size_t)h_dst[i];
        if (s == d) continue;
        CK(cudaMemcpyAsync(ctx->d_joint + d * JOINT_PLANES * K, ctx->d_joint + s * JOINT_PLANES * K, JOINT_PLANES * K * 4,
                           cudaMemcpyDeviceToDevice, st));
        CK(cudaMemcpyAsync(ctx->d_em + d * EM_PLANES * K, ctx->d_em + s * EM_PLANES * K, EM_PLANES * K * 4,
                           cudaMemcpyDeviceToDevice, st));
        CK(cudaMemcpyAsync(ctx->d_scal + d * SCALAR_WORDS, ctx->d_scal + s * SCALAR_WORDS, SCALAR_WORDS * 4,
                           cudaMemcpyDeviceToDevice, st));
        CK(cudaMemcpyAsync(ctx->d_cdf + d * K, ctx->d_cdf + s * K, K * 4, cudaMemcpyDeviceToDevice, st));
        CK(cudaMemcpyAsync(ctx->d_erec + d * EREC_PLANES * K, ctx->d_erec + s * EREC_PLANES * K, EREC_PLANES * K * 4,
                           cudaMemcpyDeviceToDevice, st));
        CK(cudaMemcpyAsync(ctx->d_qrec + d * K * QREC_FLOATS, ctx->d_qrec + s * K * QREC_FLOATS, K * QREC_FLOATS * 4,
                           cudaMemcpyDeviceToDevice, st));
        if (ctx->d_qrec_back)
            CK(cudaMemcpyAsync(ctx->d_qrec_back + d * K * QREC_FLOATS, ctx->d_qrec_back + s * K * QREC_FLOATS,
                               K * QREC_FLOATS * 4, cudaMemcpyDeviceToDevice, st));
    }
    CK(cudaStreamSynchronize(st));
    return SDMMB_OK;
}

}  // extern "C"
