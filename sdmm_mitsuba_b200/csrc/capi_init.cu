// capi_init.cu -- sdmm::initialize on the device (row F3 of SURVEY.md section 8f; sdmm/include/sdmm/opt/init.h:17-193)
// and cell copies (a split's children inherit the parent's mixture and EM state, spatial_tree.h:187-196).
#include "capi_internal.h"

#include <vector>

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
    if (n == 0) return SDMMB_OK;
    // The copies are sequential by contract (a fresh child can be split again in the same pass: its slot is then the
    // source of a later copy).  Resolve every destination to the cell whose ORIGINAL content it ends up with; when no
    // such origin is itself overwritten -- always the case for splits, which only write fresh slots -- all copies run
    // in one kernel.
    std::map<int32_t, int32_t> origin;   // slot -> slot whose original content it holds now
    auto holds = [&](int32_t slot) { auto it = origin.find(slot); return it == origin.end() ? slot : it->second; };
    for (int i = 0; i < n; ++i) origin[h_dst[i]] = holds(h_src[i]);
    bool parallel_ok = true;
    for (const auto& kv : origin)
        if (kv.second != kv.first && origin.count(kv.second) && origin[kv.second] != kv.second) { parallel_ok = false; break; }
    CopyCellsArgs a{};
    int na = 0;
    auto add = [&](void* base, size_t floats) { a.base[na] = (uint32_t*)base; a.floats_per_cell[na] = floats; ++na; };
    add(ctx->d_joint, JOINT_PLANES * K);
    add(ctx->d_em, EM_PLANES * K);
    add(ctx->d_scal, SCALAR_WORDS);
    add(ctx->d_cdf, K);
    add(ctx->d_erec, EREC_PLANES * K);
    add(ctx->d_qrec, K * QREC_FLOATS);
    if (ctx->d_qrec_back) add(ctx->d_qrec_back, K * QREC_FLOATS);
    a.n_arrays = na;
    std::vector<int32_t> pairs;
    if (parallel_ok) {
        for (const auto& kv : origin)
            if (kv.second != kv.first) { pairs.push_back(kv.second); pairs.push_back(kv.first); }
    }
    const int launches = parallel_ok ? 1 : n;
    const size_t words = parallel_ok ? pairs.size() : (size_t)2 * n;
    int rc = ensure(ctx, &ctx->d_stage2, &ctx->stage2_cap, words * 4 + 16);
    if (rc) return rc;
    if (!parallel_ok) {
        pairs.clear();
        for (int i = 0; i < n; ++i) { pairs.push_back(h_src[i]); pairs.push_back(h_dst[i]); }
    }
    if (pairs.empty()) return SDMMB_OK;
    CK(cudaMemcpyAsync(ctx->d_stage2, pairs.data(), pairs.size() * 4, cudaMemcpyHostToDevice, st));
    if (parallel_ok) {
        a.pairs = (const int32_t*)ctx->d_stage2;
        CK(launch_copy_cells(a, (int)(pairs.size() / 2), st));
    } else {
        for (int i = 0; i < n; ++i) {      // one pair per launch, in order
            a.pairs = (const int32_t*)ctx->d_stage2 + 2 * i;
            CK(launch_copy_cells(a, 1, st));
        }
    }
    ctx->launches += launches;
    CK(cudaStreamSynchronize(st));
    return SDMMB_OK;
}

}  // extern "C"
