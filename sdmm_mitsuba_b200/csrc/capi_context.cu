// capi_context.cu -- the context (device memory owner) of the C ABI of include/sdmm_b200.h: create / destroy,
// marshalling of the model state between the flat host layouts and the device layouts, parity read-outs,
// introspection.  No CPU fallback anywhere: every entry point drives the CUDA kernels of this library.
#include "capi_internal.h"

using namespace sdmmb;
using namespace sdmmb_capi;

namespace {

// host [cell][planes][K_user] -> padded [cell][planes][K]; `fill` per plane for the padding lanes
void pad_planes(const float* src, float* dst, int n_cells, int planes, int Ku, int K,
                const float* plane_fill) {
    for (int c = 0; c < n_cells; ++c)
        for (int p = 0; p < planes; ++p) {
            float* d = dst + ((size_t)c * planes + p) * K;
            if (src) std::memcpy(d, src + ((size_t)c * planes + p) * Ku, sizeof(float) * Ku);
            else for (int k = 0; k < Ku; ++k) d[k] = 0.f;
            for (int k = Ku; k < K; ++k) d[k] = plane_fill ? plane_fill[p] : 0.f;
        }
}

void strip_planes(const float* src, float* dst, int n_cells, int planes, int Ku, int K) {
    for (int c = 0; c < n_cells; ++c)
        for (int p = 0; p < planes; ++p)
            std::memcpy(dst + ((size_t)c * planes + p) * Ku, src + ((size_t)c * planes + p) * K,
                        sizeof(float) * Ku);
}

}  // namespace

extern "C" {

const char* sdmmb_get_error_string(int code) {
    switch (code) {
        case SDMMB_OK: return "ok";
        case SDMMB_E_INVALID: return "invalid argument";
        case SDMMB_E_CUDA: return "CUDA runtime error";
        case SDMMB_E_NOMEM: return "device allocation failed";
        case SDMMB_E_RANGE: return "cell range or sample count outside the context's capacity";
        case SDMMB_E_STATE: return "call order: samples not bound for this cell range";
        case SDMMB_E_NCCL: return "NCCL missing or returned an error";
        case SDMMB_E_DATA_FULL: return "Data full.";
        case SDMMB_E_PEER: return "peer memory not reachable (no P2P between the devices or CUDA IPC open failed)";
        case SDMMB_E_LIMIT: return "sdmmb_sample_pdf_noalloc: more than SDMMB_NOALLOC_MAX_VERTICES vertices per call or a context with more than SDMMB_NOALLOC_MAX_CELLS cells (fixed static scratch)";
        default: return "unknown error";
    }
}

const char* sdmmb_last_cuda_error(sdmmb_ctx* ctx) { return ctx ? ctx->last_cuda_error.c_str() : ""; }
int sdmmb_num_components(sdmmb_ctx* ctx) { return ctx ? ctx->K_user : SDMMB_E_INVALID; }
int sdmmb_max_cells(sdmmb_ctx* ctx) { return ctx ? ctx->max_cells : SDMMB_E_INVALID; }

int sdmmb_create(sdmmb_ctx** out, int device, int max_cells, int K) {
    if (!out || max_cells <= 0 || K <= 0 || K > SDMMB_MAX_K) return SDMMB_E_INVALID;
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || device < 0 || device >= n_dev) return SDMMB_E_CUDA;
    sdmmb_ctx* ctx = new (std::nothrow) sdmmb_ctx();
    if (!ctx) return SDMMB_E_NOMEM;
    ctx->device = device;
    ctx->max_cells = max_cells;
    ctx->K_user = K;
    ctx->K = pad_k(K);
    if (const char* s = std::getenv("SDMMB_ITEM_SAMPLES")) {
        long v = std::atol(s);
        if (v >= 32) ctx->item_samples = (uint32_t)v;
    }
    if (const char* s = std::getenv("SDMMB_QUERY_CHUNK")) {
        long v = std::atol(s);
        if (v >= 256) ctx->query_chunk = (size_t)v;
    }
    if (const char* s = std::getenv("SDMMB_PIPE_CHUNK_SAMPLES")) {
        long v = std::atol(s);
        if (v >= 1024) ctx->pipe_chunk_samples = (size_t)v;
    }
    DeviceGuard g(device);
    const size_t ck = (size_t)max_cells * ctx->K;
    cudaError_t e;
    if ((e = cudaMalloc(&ctx->d_joint, ck * JOINT_PLANES * 4)) != cudaSuccess ||
        (e = cudaMalloc(&ctx->d_em, ck * EM_PLANES * 4)) != cudaSuccess ||
        (e = cudaMalloc(&ctx->d_scal, (size_t)max_cells * SCALAR_WORDS * 4)) != cudaSuccess ||
        (e = cudaMalloc(&ctx->d_cdf, ck * 4)) != cudaSuccess ||
        (e = cudaMalloc(&ctx->d_erec, ck * EREC_PLANES * 4)) != cudaSuccess ||
        (e = cudaMalloc(&ctx->d_qrec, ck * QREC_FLOATS * 4)) != cudaSuccess) {
        int rc = cuda_fail(nullptr, e);
        sdmmb_destroy(ctx);
        return rc;
    }
    cudaMemset(ctx->d_joint, 0, ck * JOINT_PLANES * 4);
    cudaMemset(ctx->d_em, 0, ck * EM_PLANES * 4);
    cudaMemset(ctx->d_scal, 0, (size_t)max_cells * SCALAR_WORDS * 4);
    cudaMemset(ctx->d_cdf, 0, ck * 4);
    cudaMemset(ctx->d_erec, 0, ck * EREC_PLANES * 4);
    cudaMemset(ctx->d_qrec, 0, ck * QREC_FLOATS * 4);
    *out = ctx;
    return SDMMB_OK;
}

int sdmmb_destroy(sdmmb_ctx* ctx) {
    if (!ctx) return SDMMB_OK;
    // no lock: the caller guarantees that nothing else uses a context it destroys
    DeviceGuard g(ctx->device);
    cudaFree(ctx->d_joint); cudaFree(ctx->d_em); cudaFree(ctx->d_scal); cudaFree(ctx->d_cdf);
    cudaFree(ctx->d_erec); cudaFree(ctx->d_qrec); cudaFree(ctx->d_qrec_back); cudaFree(ctx->d_samples); cudaFree(ctx->d_items);
    cudaFree(ctx->d_cell_item_begin); cudaFree(ctx->d_cell_sample_count); cudaFree(ctx->d_partials);
    cudaFree(ctx->d_item_scal); cudaFree(ctx->d_stage); cudaFree(ctx->d_stage2); cudaFree(ctx->d_qscratch); cudaFree(ctx->d_tree);
    for (auto& kv : ctx->qscratch_by_stream) cudaFree(kv.second.p);
    for (void* m : ctx->peer_ipc_opened) cudaIpcCloseMemHandle(m);
    cudaFree(ctx->d_peer_flags);
    for (cudaStream_t ps : ctx->pipe_fit_stream) if (ps) cudaStreamDestroy(ps);
    if (ctx->ov_mstream) {
        cudaStreamDestroy(ctx->ov_stream[0]); cudaStreamDestroy(ctx->ov_stream[1]); cudaStreamDestroy(ctx->ov_mstream);
        for (cudaEvent_t e : ctx->ov_ev) cudaEventDestroy(e);
    }
    for (cudaEvent_t e : ctx->ev) cudaEventDestroy(e);
    if (ctx->copy_stream) {
        cudaStreamDestroy(ctx->copy_stream);
        cudaStreamDestroy(ctx->d2h_stream);
        for (cudaEvent_t e : ctx->pipe_ev) cudaEventDestroy(e);
    }
    delete ctx;
    return SDMMB_OK;
}

int sdmmb_upload_cells(sdmmb_ctx* ctx, int first_cell, int n_cells, const float* h_joint,
                       const float* h_em, const uint32_t* h_scal, void* stream) {
    if (!ctx || !h_joint) return SDMMB_E_INVALID;
    if (!range_ok(ctx, first_cell, n_cells)) return SDMMB_E_RANGE;
    if (n_cells == 0) return SDMMB_OK;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int K = ctx->K, Ku = ctx->K_user;
    // padding components: zero weight, direction +z, identity covariance (Cholesky stays finite)
    float jfill[JOINT_PLANES] = {0};
    jfill[6] = 1.f;
    for (int i = 0; i < 5; ++i) jfill[7 + 6 * i] = 1.f;
    // no padding needed (K already 16/32/64/128): copy straight from the caller's buffers
    const bool direct = (Ku == K);
    std::vector<float> hj, he;
    if (!direct) {
        hj.resize((size_t)n_cells * JOINT_PLANES * K);
        pad_planes(h_joint, hj.data(), n_cells, JOINT_PLANES, Ku, K, jfill);
    }
    if (!direct || !h_em) {
        he.resize((size_t)n_cells * EM_PLANES * K);
        pad_planes(h_em, he.data(), n_cells, EM_PLANES, Ku, K, nullptr);
    }
    const float* src_j = direct ? h_joint : hj.data();
    const float* src_e = (direct && h_em) ? h_em : he.data();
    std::vector<uint32_t> hs((size_t)n_cells * SCALAR_WORDS, 0u);
    for (int c = 0; c < n_cells; ++c) {
        uint32_t* s = hs.data() + (size_t)c * SCALAR_WORDS;
        if (h_scal) std::memcpy(s, h_scal + (size_t)c * SCALAR_WORDS, SCALAR_WORDS * 4);
        else {
            float lr = 0.2f, al = 0.5f;  // R: sdmm/opt/em.h:137-138
            std::memcpy(&s[5], &lr, 4);
            std::memcpy(&s[6], &al, 4);
        }
        s[7] = 0;
    }
    CK(cudaMemcpyAsync(ctx->d_joint + (size_t)first_cell * JOINT_PLANES * K, src_j,
                       (size_t)n_cells * JOINT_PLANES * K * 4, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(ctx->d_em + (size_t)first_cell * EM_PLANES * K, src_e,
                       (size_t)n_cells * EM_PLANES * K * 4, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(ctx->d_scal + (size_t)first_cell * SCALAR_WORDS, hs.data(), hs.size() * 4,
                       cudaMemcpyHostToDevice, st));
    MStepArgs a = base_mstep_args(ctx, first_cell, n_cells, 4);
    CK(launch_mstep(a, st));
    ctx->launches += 1;
    CK(cudaStreamSynchronize(st));  // the pageable staging vectors die at return
    return SDMMB_OK;
}

int sdmmb_download_cells(sdmmb_ctx* ctx, int first_cell, int n_cells, float* h_joint, float* h_em,
                         uint32_t* h_scal, void* stream) {
    if (!ctx) return SDMMB_E_INVALID;
    if (!range_ok(ctx, first_cell, n_cells)) return SDMMB_E_RANGE;
    if (n_cells == 0) return SDMMB_OK;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int K = ctx->K, Ku = ctx->K_user;
    std::vector<float> tmp;
    if (Ku == K) {
        if (h_joint)
            CK(cudaMemcpyAsync(h_joint, ctx->d_joint + (size_t)first_cell * JOINT_PLANES * K,
                               (size_t)n_cells * JOINT_PLANES * K * 4, cudaMemcpyDeviceToHost, st));
        if (h_em)
            CK(cudaMemcpyAsync(h_em, ctx->d_em + (size_t)first_cell * EM_PLANES * K,
                               (size_t)n_cells * EM_PLANES * K * 4, cudaMemcpyDeviceToHost, st));
        if (h_scal)
            CK(cudaMemcpyAsync(h_scal, ctx->d_scal + (size_t)first_cell * SCALAR_WORDS,
                               (size_t)n_cells * SCALAR_WORDS * 4, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        return SDMMB_OK;
    }
    if (h_joint) {
        tmp.resize((size_t)n_cells * JOINT_PLANES * K);
        CK(cudaMemcpyAsync(tmp.data(), ctx->d_joint + (size_t)first_cell * JOINT_PLANES * K, tmp.size() * 4,
                           cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        strip_planes(tmp.data(), h_joint, n_cells, JOINT_PLANES, Ku, K);
    }
    if (h_em) {
        tmp.resize((size_t)n_cells * EM_PLANES * K);
        CK(cudaMemcpyAsync(tmp.data(), ctx->d_em + (size_t)first_cell * EM_PLANES * K, tmp.size() * 4,
                           cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        strip_planes(tmp.data(), h_em, n_cells, EM_PLANES, Ku, K);
    }
    if (h_scal) {
        CK(cudaMemcpyAsync(h_scal, ctx->d_scal + (size_t)first_cell * SCALAR_WORDS,
                           (size_t)n_cells * SCALAR_WORDS * 4, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
    }
    return SDMMB_OK;
}

static int readout38(sdmmb_ctx* ctx, int first_cell, int n_cells, float* h_out, void* stream, bool cond) {
    if (!ctx || !h_out) return SDMMB_E_INVALID;
    if (!range_ok(ctx, first_cell, n_cells)) return SDMMB_E_RANGE;
    if (n_cells == 0) return SDMMB_OK;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int K = ctx->K, Ku = ctx->K_user;
    const size_t bytes = (size_t)n_cells * 38 * K * 4;
    int rc = ensure(ctx, &ctx->d_stage, &ctx->stage_cap, bytes);
    if (rc) return rc;
    if (cond) CK(launch_conditioner_readout(ctx->d_joint, ctx->d_qrec, first_cell, n_cells, K, (float*)ctx->d_stage, st));
    else CK(launch_derived_readout(ctx->d_joint, ctx->d_cdf, first_cell, n_cells, K, (float*)ctx->d_stage, st));
    ctx->launches += 1;
    std::vector<float> tmp((size_t)n_cells * 38 * K);
    CK(cudaMemcpyAsync(tmp.data(), ctx->d_stage, bytes, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    strip_planes(tmp.data(), h_out, n_cells, 38, Ku, K);
    return SDMMB_OK;
}

int sdmmb_download_derived(sdmmb_ctx* ctx, int first_cell, int n_cells, float* h_out, void* stream) {
    return readout38(ctx, first_cell, n_cells, h_out, stream, false);
}
int sdmmb_download_conditioner(sdmmb_ctx* ctx, int first_cell, int n_cells, float* h_out, void* stream) {
    return readout38(ctx, first_cell, n_cells, h_out, stream, true);
}


int sdmmb_host_alloc(void** out, size_t bytes) {
    if (!out) return SDMMB_E_INVALID;
    *out = nullptr;
    if (bytes == 0) return SDMMB_OK;
    cudaError_t e = cudaHostAlloc(out, bytes, cudaHostAllocPortable);
    if (e != cudaSuccess) { cudaGetLastError(); return e == cudaErrorMemoryAllocation ? SDMMB_E_NOMEM : SDMMB_E_CUDA; }
    return SDMMB_OK;
}

int sdmmb_host_free(void* p) {
    if (p && cudaFreeHost(p) != cudaSuccess) { cudaGetLastError(); return SDMMB_E_CUDA; }
    return SDMMB_OK;
}

int sdmmb_launch_count(sdmmb_ctx* ctx, uint64_t* out) {
    if (!ctx || !out) return SDMMB_E_INVALID;
    *out = ctx->launches;
    return SDMMB_OK;
}
int sdmmb_last_estep_ms(sdmmb_ctx* ctx, float* total_ms, int* n_launches) {
    if (!ctx || !total_ms) return SDMMB_E_INVALID;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    float tot = 0.f;
    for (int i = 0; i < ctx->ev_used; ++i) {
        CK(cudaEventSynchronize(ctx->ev[2 * i + 1]));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, ctx->ev[2 * i], ctx->ev[2 * i + 1]));
        tot += ms;
    }
    *total_ms = tot;
    if (n_launches) *n_launches = ctx->ev_used;
    return SDMMB_OK;
}
int sdmmb_measure_peaks(int device, float* out4) {
    if (!out4) return SDMMB_E_INVALID;
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || device < 0 || device >= n_dev) return SDMMB_E_CUDA;
    DeviceGuard g(device);
    return measure_peaks(out4) == cudaSuccess ? SDMMB_OK : SDMMB_E_CUDA;
}

}  // extern "C"
