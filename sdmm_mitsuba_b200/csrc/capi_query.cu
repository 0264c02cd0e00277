// capi_query.cu -- the query half of the C ABI: guided sample + pdf at path vertices (device pointers, or host
// buffers pipelined over chunks), product guiding, the double-buffered records of the optimizeAsync schedule.
// Reference: SDMMRenderer::sampleSurface / pdfSurface (mitsuba/src/integrators/sdmm/sdmm_proc.cpp:361-372,
// 419-421, 539-545).
#include "capi_internal.h"

using namespace sdmmb;
using namespace sdmmb_capi;

namespace sdmmb_capi {

int query_common(sdmmb_ctx* ctx, size_t n, const uint32_t* d_cell, const float* d_pos3, const float* d_u3,
                 const float* d_dir3, float* d_out_dir3, float* d_out_pdf, float* d_out_ij,
                 uint32_t* d_out_comp, uint8_t* d_out_valid, int binned, cudaStream_t st) {
    SDMMB_NVTX("sdmmb guided query (K4)");
    QueryLaunch q{};
    q.K = ctx->K;
    q.n = n;
    q.qrec = ctx->d_qrec;
    q.cell_idx = d_cell;
    q.pos3 = d_pos3;
    q.u3 = d_u3;
    q.dir3 = d_dir3;
    q.out_dir3 = d_out_dir3;
    q.out_pdf = d_out_pdf;
    q.out_inv_jac = d_out_ij;
    q.out_comp = d_out_comp;
    q.out_valid = d_out_valid;
    q.binned = binned;
    q.max_cells = ctx->max_cells;
    q.scratch = nullptr;
    if (!binned && n) {
        // device-side binning scratch: count[B] | cursor[B] | invperm[n] | qin[8n] | qout[8n] (words)
        void* scratch = nullptr;
        int rc = stream_scratch(ctx, st, (2 * ((size_t)ctx->max_cells + 1) + 17 * n + 4) * 4, &scratch);
        if (rc) return rc;
        q.scratch = (uint32_t*)scratch;
    }
    CK(launch_query(q, st));
    if (n) ctx->launches += binned ? 1 : 5;
    return SDMMB_OK;
}

}  // namespace sdmmb_capi

namespace {

// Host-buffer queries, pipelined over chunks of vertices: H2D of chunk c+1 (copy stream) and D2H of
// chunk c-1 (third stream) overlap the kernels of chunk c.  Host planes have stride n; each chunk is
// staged as its own dense block:  cell[m] pos[3m] in[3m] | dir[3m] pdf[m] ij[m] comp[m] valid[m bytes].
int query_host(sdmmb_ctx* ctx, size_t n, const uint32_t* h_cell, const float* h_pos3, const float* h_in3,
               bool sample, float* h_dir3, float* h_pdf, float* h_ij, uint32_t* h_comp, uint8_t* h_valid,
               int binned, cudaStream_t st) {
    int rc = ensure_pipe(ctx);
    if (rc) return rc;
    const size_t Q = std::min(n, ctx->query_chunk);
    const size_t block = (Q * (14 * 4) + Q + 255) & ~(size_t)255;
    rc = ensure(ctx, &ctx->d_stage2, &ctx->stage2_cap, 2 * block);
    if (rc) return rc;
    cudaEvent_t* copied = ctx->pipe_ev;        // [0,1]
    cudaEvent_t* computed = ctx->pipe_ev + 2;  // [2,3]
    cudaEvent_t* drained = ctx->pipe_ev + 6;   // [6,7]
    CK(cudaEventRecord(ctx->pipe_ev[4], st));
    CK(cudaStreamWaitEvent(ctx->copy_stream, ctx->pipe_ev[4], 0));
    size_t c = 0;
    for (size_t off = 0; off < n; off += Q, ++c) {
        const size_t m = std::min(Q, n - off);
        const int buf = (int)(c & 1);
        float* f = (float*)((unsigned char*)ctx->d_stage2 + (size_t)buf * block);
        uint32_t* d_cell = (uint32_t*)f;
        float *d_pos = f + m, *d_in = f + 4 * m, *d_dir = f + 7 * m, *d_pdf = f + 10 * m, *d_ij = f + 11 * m;
        uint32_t* d_comp = (uint32_t*)(f + 12 * m);
        uint8_t* d_valid = (uint8_t*)(f + 13 * m);
        if (c >= 2) CK(cudaStreamWaitEvent(ctx->copy_stream, drained[buf], 0));
        CK(cudaMemcpyAsync(d_cell, h_cell + off, m * 4, cudaMemcpyHostToDevice, ctx->copy_stream));
        for (int p = 0; p < 3; ++p) {
            CK(cudaMemcpyAsync(d_pos + p * m, h_pos3 + p * n + off, m * 4, cudaMemcpyHostToDevice, ctx->copy_stream));
            CK(cudaMemcpyAsync(d_in + p * m, h_in3 + p * n + off, m * 4, cudaMemcpyHostToDevice, ctx->copy_stream));
        }
        CK(cudaEventRecord(copied[buf], ctx->copy_stream));
        CK(cudaStreamWaitEvent(st, copied[buf], 0));
        if (sample) rc = query_common(ctx, m, d_cell, d_pos, d_in, nullptr, d_dir, d_pdf, d_ij, d_comp, d_valid, binned, st);
        else rc = query_common(ctx, m, d_cell, d_pos, nullptr, d_in, nullptr, d_pdf, nullptr, nullptr, d_valid, binned, st);
        if (rc) return rc;
        CK(cudaEventRecord(computed[buf], st));
        CK(cudaStreamWaitEvent(ctx->d2h_stream, computed[buf], 0));
        if (sample) {
            for (int p = 0; p < 3; ++p)
                CK(cudaMemcpyAsync(h_dir3 + p * n + off, d_dir + p * m, m * 4, cudaMemcpyDeviceToHost, ctx->d2h_stream));
            CK(cudaMemcpyAsync(h_ij + off, d_ij, m * 4, cudaMemcpyDeviceToHost, ctx->d2h_stream));
            CK(cudaMemcpyAsync(h_comp + off, d_comp, m * 4, cudaMemcpyDeviceToHost, ctx->d2h_stream));
        }
        CK(cudaMemcpyAsync(h_pdf + off, d_pdf, m * 4, cudaMemcpyDeviceToHost, ctx->d2h_stream));
        if (h_valid) CK(cudaMemcpyAsync(h_valid + off, d_valid, m, cudaMemcpyDeviceToHost, ctx->d2h_stream));
        CK(cudaEventRecord(drained[buf], ctx->d2h_stream));
    }
    CK(cudaStreamSynchronize(st));
    CK(cudaStreamSynchronize(ctx->copy_stream));
    CK(cudaStreamSynchronize(ctx->d2h_stream));
    return SDMMB_OK;
}

}  // namespace

// ---- allocation-free guided query (no counterpart in the reference) ----
// The random-order query bins its vertices by cell on the device and needs scratch for that; the entry points
// above take it from a grow-only cudaMalloc'ed buffer per stream.  Callers that must not allocate (CUDA graph
// capture, tight serving loops) get the same kernels over ONE file-scope device array of a fixed size instead.
// Layout as in query_common: count[B] | cursor[B] | invperm[n] | qin[8n] | qout[8n] words, B = max_cells + 1.
#define NOALLOC_SCRATCH_WORDS (2 * ((size_t)SDMMB_NOALLOC_MAX_CELLS + 1) + 17 * (size_t)SDMMB_NOALLOC_MAX_VERTICES + 4)
static __device__ __align__(256) uint32_t g_noalloc_scratch[NOALLOC_SCRATCH_WORDS];

namespace {

// In host code `g_noalloc_scratch` is only the host-side stand-in of the symbol; the device address is looked up
// once per device, with that device current, and cached.
constexpr int NOALLOC_MAX_DEVICES = 64;
uint32_t* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};
std::mutex g_noalloc_mu;

int noalloc_scratch_for(sdmmb_ctx* ctx, uint32_t** out) {
    const int dev = ctx->device;
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return SDMMB_E_INVALID;
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        int cur = -1;
        CK(cudaGetDevice(&cur));
        if (cur != dev) return SDMMB_E_STATE;   // the caller holds a DeviceGuard; never look the symbol up elsewhere
        void* p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) return SDMMB_E_CUDA;
        g_noalloc_ptr[dev] = (uint32_t*)p;
    }
    *out = g_noalloc_ptr[dev];
    return SDMMB_OK;
}

}  // namespace

extern "C" {

int sdmmb_sample_pdf(sdmmb_ctx* ctx, size_t n, const uint32_t* d_cell_idx, const float* d_pos3,
                     const float* d_u3, float* d_out_dir3, float* d_out_pdf, float* d_out_inv_jac,
                     uint32_t* d_out_comp, uint8_t* d_out_valid, int binned, void* stream) {
    if (!ctx || (n && (!d_cell_idx || !d_pos3 || !d_u3 || !d_out_dir3 || !d_out_pdf || !d_out_inv_jac || !d_out_comp)))
        return SDMMB_E_INVALID;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    return query_common(ctx, n, d_cell_idx, d_pos3, d_u3, nullptr, d_out_dir3, d_out_pdf, d_out_inv_jac,
                        d_out_comp, d_out_valid, binned, (cudaStream_t)stream);
}

int sdmmb_pdf(sdmmb_ctx* ctx, size_t n, const uint32_t* d_cell_idx, const float* d_pos3,
              const float* d_dir3, float* d_out_pdf, uint8_t* d_out_valid, int binned, void* stream) {
    if (!ctx || (n && (!d_cell_idx || !d_pos3 || !d_dir3 || !d_out_pdf))) return SDMMB_E_INVALID;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    return query_common(ctx, n, d_cell_idx, d_pos3, nullptr, d_dir3, nullptr, d_out_pdf, nullptr, nullptr,
                        d_out_valid, binned, (cudaStream_t)stream);
}

int sdmmb_sample_pdf_host(sdmmb_ctx* ctx, size_t n, const uint32_t* h_cell_idx, const float* h_pos3,
                          const float* h_u3, float* h_out_dir3, float* h_out_pdf, float* h_out_inv_jac,
                          uint32_t* h_out_comp, uint8_t* h_out_valid, int binned, void* stream) {
    if (!ctx || (n && (!h_cell_idx || !h_pos3 || !h_u3 || !h_out_dir3 || !h_out_pdf || !h_out_inv_jac || !h_out_comp)))
        return SDMMB_E_INVALID;
    if (n == 0) return SDMMB_OK;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    return query_host(ctx, n, h_cell_idx, h_pos3, h_u3, true, h_out_dir3, h_out_pdf, h_out_inv_jac, h_out_comp,
                      h_out_valid, binned, (cudaStream_t)stream);
}

int sdmmb_pdf_host(sdmmb_ctx* ctx, size_t n, const uint32_t* h_cell_idx, const float* h_pos3,
                   const float* h_dir3, float* h_out_pdf, uint8_t* h_out_valid, int binned, void* stream) {
    if (!ctx || (n && (!h_cell_idx || !h_pos3 || !h_dir3 || !h_out_pdf))) return SDMMB_E_INVALID;
    if (n == 0) return SDMMB_OK;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    return query_host(ctx, n, h_cell_idx, h_pos3, h_dir3, false, nullptr, h_out_pdf, nullptr, nullptr, h_out_valid,
                      binned, (cudaStream_t)stream);
}

int sdmmb_set_async_records(sdmmb_ctx* ctx, int enabled) {
    if (!ctx) return SDMMB_E_INVALID;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    const size_t bytes = (size_t)ctx->max_cells * ctx->K * QREC_FLOATS * 4;
    if (enabled && !ctx->d_qrec_back) {
        CK(cudaDeviceSynchronize());
        CK(cudaMalloc(&ctx->d_qrec_back, bytes));
        CK(cudaMemcpy(ctx->d_qrec_back, ctx->d_qrec, bytes, cudaMemcpyDeviceToDevice));
    } else if (!enabled && ctx->d_qrec_back) {
        // leaving async mode publishes whatever is pending
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(ctx->d_qrec, ctx->d_qrec_back, bytes, cudaMemcpyDeviceToDevice));
        cudaFree(ctx->d_qrec_back);
        ctx->d_qrec_back = nullptr;
    }
    return SDMMB_OK;
}

int sdmmb_publish_records(sdmmb_ctx* ctx, int first_cell, int n_cells, void* stream) {
    if (!ctx) return SDMMB_E_INVALID;
    if (!range_ok(ctx, first_cell, n_cells)) return SDMMB_E_RANGE;
    if (!ctx->d_qrec_back || n_cells == 0) return SDMMB_OK;   // synchronous mode: already visible
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    const size_t per = (size_t)ctx->K * QREC_FLOATS;
    CK(cudaMemcpyAsync(ctx->d_qrec + (size_t)first_cell * per, ctx->d_qrec_back + (size_t)first_cell * per,
                       (size_t)n_cells * per * 4, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return SDMMB_OK;
}

int sdmmb_query_records(sdmmb_ctx* ctx, float** d_records, size_t* floats_per_cell) {
    if (!ctx || !d_records || !floats_per_cell) return SDMMB_E_INVALID;
    *d_records = ctx->d_qrec;
    *floats_per_cell = (size_t)ctx->K * QREC_FLOATS;
    return SDMMB_OK;
}

int sdmmb_sample_pdf_pruned(sdmmb_ctx* ctx, size_t n, const uint32_t* d_cell_idx, const float* d_pos3,
                            const float* d_u3, const float* d_dir3, int max_components, int preserve_idx,
                            float* d_out_dir3, float* d_out_pdf, float* d_out_inv_jac, uint32_t* d_out_comp,
                            uint8_t* d_out_valid, void* stream) {
    if (!ctx || max_components < 1 || (n && (!d_cell_idx || !d_pos3 || !d_out_pdf || (!d_u3 && !d_dir3)))) return SDMMB_E_INVALID;
    if (n && d_u3 && (!d_out_dir3 || !d_out_inv_jac || !d_out_comp)) return SDMMB_E_INVALID;
    if (preserve_idx >= ctx->K_user) return SDMMB_E_INVALID;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    SDMMB_NVTX("sdmmb pruned query");
    QueryLaunch q{};
    q.K = ctx->K;
    q.n = n;
    q.qrec = ctx->d_qrec;
    q.cell_idx = d_cell_idx;
    q.pos3 = d_pos3;
    q.u3 = d_u3;
    q.dir3 = d_dir3;
    q.out_dir3 = d_out_dir3;
    q.out_pdf = d_out_pdf;
    q.out_inv_jac = d_out_inv_jac;
    q.out_comp = d_out_comp;
    q.out_valid = d_out_valid;
    q.max_cells = ctx->max_cells;
    CK(launch_query_pruned(q, ctx->K_user, max_components, preserve_idx, (cudaStream_t)stream));
    if (n) ctx->launches += 1;
    return SDMMB_OK;
}

// ---- allocation-free guided query: static scratch and its address look-up are above extern "C" ----
int sdmmb_sample_pdf_noalloc(sdmmb_ctx* ctx, size_t n, const uint32_t* d_cell_idx, const float* d_pos3,
                             const float* d_u3, float* d_out_dir3, float* d_out_pdf, float* d_out_inv_jac,
                             uint32_t* d_out_comp, uint8_t* d_out_valid, void* stream) {
    if (!ctx || (n && (!d_cell_idx || !d_pos3 || !d_u3 || !d_out_dir3 || !d_out_pdf || !d_out_inv_jac || !d_out_comp)))
        return SDMMB_E_INVALID;
    if (n > (size_t)SDMMB_NOALLOC_MAX_VERTICES || ctx->max_cells > SDMMB_NOALLOC_MAX_CELLS) return SDMMB_E_LIMIT;
    if (n == 0) return SDMMB_OK;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    SDMMB_NVTX("sdmmb guided query (K4, static scratch)");
    uint32_t* scratch = nullptr;
    int rc = noalloc_scratch_for(ctx, &scratch);
    if (rc) return rc;
    QueryLaunch q{};
    q.K = ctx->K;
    q.n = n;
    q.qrec = ctx->d_qrec;
    q.cell_idx = d_cell_idx;
    q.pos3 = d_pos3;
    q.u3 = d_u3;
    q.dir3 = nullptr;
    q.out_dir3 = d_out_dir3;
    q.out_pdf = d_out_pdf;
    q.out_inv_jac = d_out_inv_jac;
    q.out_comp = d_out_comp;
    q.out_valid = d_out_valid;
    q.binned = 0;
    q.max_cells = ctx->max_cells;
    q.scratch = scratch;
    // the array keeps the previous call's contents: launch_bin_vertices zeroes the histogram words (count[B]) with
    // a memset node on `stream`; cursor / invperm / qin / qout are fully written for [0, n) before they are read
    CK(launch_query(q, (cudaStream_t)stream));
    ctx->launches += 5;
    return SDMMB_OK;
}

// ---- product guiding (row F4 of SURVEY.md section 8f) ----
static int product_common(sdmmb_ctx* ctx, size_t n, const uint32_t* d_cell_idx, const float* d_pos3,
                          const float* d_u3, const float* d_dir3, const float* d_normal3, const float* d_frame9,
                          const float* d_wi3, int n_bsdf,
                          const float* h_bsdf, float* d_out_dir3, float* d_out_pdf, float* d_out_inv_jac,
                          uint32_t* d_out_comp, uint8_t* d_out_valid, cudaStream_t st) {
    if ((d_frame9 == nullptr) != (d_wi3 == nullptr)) return SDMMB_E_INVALID;
    if (n_bsdf < 1 || n_bsdf > SDMMB_PRODUCT_MAX_BSDF || !h_bsdf) return SDMMB_E_INVALID;
    if (ctx->K_user % 8 != 0) return SDMMB_E_INVALID;   // the product's component order is packet-based
    if (n == 0) return SDMMB_OK;
    for (int m = 0; m < n_bsdf; ++m) {
        const float* b = h_bsdf + 7 * m;
        if (!(b[0] >= 0.f) || !(b[1] * b[1] + b[2] * b[2] + b[3] * b[3] > 0.f)) return SDMMB_E_INVALID;
    }
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    int rc = ensure(ctx, &ctx->d_stage2, &ctx->stage2_cap, (size_t)SDMMB_PRODUCT_MAX_BSDF * 7 * 4);
    if (rc != SDMMB_OK) return rc;
    CK(cudaMemcpyAsync(ctx->d_stage2, h_bsdf, (size_t)n_bsdf * 7 * 4, cudaMemcpyHostToDevice, st));
    void* pscratch = nullptr;
    rc = stream_scratch(ctx, st, (2 * ((size_t)ctx->max_cells + 1) + 9 * n + 4) * 4, &pscratch);
    if (rc != SDMMB_OK) return rc;
    ProductLaunch q{};
    q.scratch = (uint32_t*)pscratch;
    q.K = ctx->K;
    q.K_user = ctx->K_user;
    q.max_cells = ctx->max_cells;
    q.n = n;
    q.qrec = ctx->d_qrec;
    q.cell_idx = d_cell_idx;
    q.pos3 = d_pos3;
    q.u3 = d_u3;
    q.dir3 = d_dir3;
    q.normal3 = d_normal3;
    q.frame9 = d_frame9;
    q.wi3 = d_wi3;
    q.M = n_bsdf;
    q.bsdf = (const float*)ctx->d_stage2;
    q.out_dir3 = d_out_dir3;
    q.out_pdf = d_out_pdf;
    q.out_inv_jac = d_out_inv_jac;
    q.out_comp = d_out_comp;
    q.out_valid = d_out_valid;
    CK(launch_product_query(q, st));
    ctx->launches += 4;
    // the BSDF table lives in the context's staging buffer: do not let a later call overwrite it early
    CK(cudaStreamSynchronize(st));
    return SDMMB_OK;
}

int sdmmb_sample_pdf_product(sdmmb_ctx* ctx, size_t n, const uint32_t* d_cell_idx, const float* d_pos3,
                             const float* d_u3, const float* d_normal3, const float* d_frame9, const float* d_wi3,
                             int n_bsdf, const float* h_bsdf, float* d_out_dir3, float* d_out_pdf, float* d_out_inv_jac,
                             uint32_t* d_out_comp, uint8_t* d_out_valid, void* stream) {
    if (!ctx || (n && (!d_cell_idx || !d_pos3 || !d_u3 || !d_out_dir3 || !d_out_pdf))) return SDMMB_E_INVALID;
    return product_common(ctx, n, d_cell_idx, d_pos3, d_u3, nullptr, d_normal3, d_frame9, d_wi3, n_bsdf, h_bsdf, d_out_dir3,
                          d_out_pdf, d_out_inv_jac, d_out_comp, d_out_valid, (cudaStream_t)stream);
}

int sdmmb_pdf_product(sdmmb_ctx* ctx, size_t n, const uint32_t* d_cell_idx, const float* d_pos3,
                      const float* d_dir3, const float* d_normal3, const float* d_frame9, const float* d_wi3,
                      int n_bsdf, const float* h_bsdf, float* d_out_pdf, uint8_t* d_out_valid, void* stream) {
    if (!ctx || (n && (!d_cell_idx || !d_pos3 || !d_dir3 || !d_out_pdf))) return SDMMB_E_INVALID;
    return product_common(ctx, n, d_cell_idx, d_pos3, nullptr, d_dir3, d_normal3, d_frame9, d_wi3, n_bsdf, h_bsdf, nullptr,
                          d_out_pdf, nullptr, nullptr, d_out_valid, (cudaStream_t)stream);
}

}  // extern "C"
