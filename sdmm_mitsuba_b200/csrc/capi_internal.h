// capi_internal.h -- shared by the translation units of the C ABI (capi_*.cu): the context object (owner of all
// device memory), error plumbing, and the helpers one unit defines and another uses.  Internal; the public
// boundary is include/sdmm_b200.h.
//   capi_context.cu  create / destroy, model upload / download, read-outs, introspection
//   capi_fit.cu      sample binding, work items, em_step (in order / overlapped), optimize_host, conditioner
//   capi_query.cu    guided sample + pdf queries (device and host buffers), product guiding, async records
//   capi_tree.cu     spatial cell structure: node table, find, sample collection
//   capi_init.cu     sdmm::initialize on the device, cell copies (split inheritance)
//   capi_peer.cu     record exchange: peer memory (fused into K3) and the NCCL all-gather
//   capi_render.cu   the device path tracer's host side
#pragma once
#include <dlfcn.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/sdmm_b200.h"
#include "../../include/sdmm_b200_render.h"
#include "kernels.h"
#include "nvtx_range.h"
#include "render.h"

struct sdmmb_ctx {
    int device = 0;
    int max_cells = 0;
    int K_user = 0;  // components as the caller sees them
    int K = 0;       // padded to 16 / 32 / 64 / 128 with zero-weight components
    // model state
    float* d_joint = nullptr;    // [max_cells][32][K]
    float* d_em = nullptr;       // [max_cells][81][K]
    uint32_t* d_scal = nullptr;  // [max_cells][8]
    float* d_cdf = nullptr;      // [max_cells][K]
    float* d_erec = nullptr;     // [max_cells][EREC_PLANES][K]
    float* d_qrec = nullptr;     // [max_cells][K][32]  records the queries (and the all-gather) read
    float* d_qrec_back = nullptr;  // async mode: records the fit writes, published by sdmmb_publish_records
    // bound samples
    float4* d_samples = nullptr;
    size_t samples_cap = 0, n_samples = 0;
    // sdmmb_set_samples_borrowed: the caller's planes, read in place by the E-step (null = packed records)
    const float* soa_point6 = nullptr;
    size_t soa_stride = 0;
    const float* soa_weight = nullptr;
    int bound_first = 0, bound_cells = 0;
    std::vector<uint32_t> h_item_begin;  // [bound_cells+1] item index per bound cell
    sdmmb::EItem* d_items = nullptr;
    uint32_t* d_cell_item_begin = nullptr;
    uint32_t* d_cell_sample_count = nullptr;
    float* d_partials = nullptr;
    float* d_item_scal = nullptr;
    size_t items_cap = 0, cells_cap = 0, n_items = 0;
    // layout of the item table currently on the device (build_items skips an identical rebuild)
    std::vector<uint32_t> items_off;
    int items_first = -1;
    uint32_t items_target = 0;
    bool items_valid = false;
    // staging (grow-only)
    void* d_stage = nullptr;
    size_t stage_cap = 0;
    void* d_stage2 = nullptr;
    size_t stage2_cap = 0;
    sdmmb::TreeNode* d_tree = nullptr;  // flattened STree (breadth-first order)
    int n_tree_nodes = 0;
    std::vector<sdmmb::TreeNode> h_tree;   // host copy of d_tree (the tracer derives the leaves' boxes from it)
    uint64_t tree_version = 0;             // bumped by every sdmmb_set_tree
    uint32_t min_samples = 0;    // canBeOptimized threshold applied by the M-step
    void* d_qscratch = nullptr;  // binning scratch of the tree / product paths
    size_t qscratch_cap = 0;
    // query binning scratch, one buffer per caller stream: device-pointer queries enqueued on different
    // streams (two render threads) overlap on the device instead of sharing one scratch buffer
    struct QScratch { void* p = nullptr; size_t cap = 0; };
    std::map<cudaStream_t, QScratch> qscratch_by_stream;
    // bookkeeping
    std::atomic<uint64_t> launches{0};
    std::string last_cuda_error;
    // every entry point that takes the context locks it: calls from concurrent host threads (Mitsuba's
    // LocalWorkers query while optimize() fits, sdmm_proc.cpp:1086-1106 / volpath_sdmm.cpp:180-242) are
    // serialised at the API boundary; recursive because entry points call each other
    std::recursive_mutex mu;
    std::vector<cudaEvent_t> ev;  // pairs around the E-step launches of the last em_step
    int ev_used = 0;
    uint32_t item_samples = 1024;
    int sm_count = 0;
    // sdmmb_optimize_host pipeline: H2D copy stream, D2H stream, events
    // [0,1] chunk copied, [2,3] chunk packed (staging buffer reusable), [4] entry fence, [5] chunk fitted
    cudaStream_t copy_stream = nullptr, d2h_stream = nullptr;
    cudaStream_t pipe_fit_stream[4] = {nullptr, nullptr, nullptr, nullptr};   // optimize_host: alternating chunks
    int pipe_fit_streams = 2;
    cudaEvent_t pipe_ev[8] = {};  // [6,7] query_host: chunk's results copied out (staging reusable)
    size_t pipe_chunk_samples = 2u << 20;
    size_t query_chunk = 2u << 20;  // vertices per chunk of the host-buffer query pipeline
    // nccl (resolved lazily)
    void* nccl_handle = nullptr;
    int (*nccl_allgather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
    int (*nccl_user_rank)(void*, int*) = nullptr;
    // sdmmb_set_em_overlap: two E-step streams + one high-priority M-step stream, fork/E/M events
    bool em_overlap = false;
    cudaStream_t ov_stream[2] = {nullptr, nullptr}, ov_mstream = nullptr;
    cudaEvent_t ov_ev[17] = {};   // fork, E-step done [8], M-step done [8]
    // peer-memory record exchange (capi_part6.inc)
    uint32_t* d_peer_flags = nullptr;   // PEER_FLAG_WORDS words, exported to the peers
    sdmmb::PeerPush peer{};                    // world == 0 until sdmmb_peer_attach
    uint32_t peer_epoch = 0;            // pushing launches issued so far
    std::vector<void*> peer_ipc_opened; // mappings to close on detach
};

namespace sdmmb_capi {
using namespace sdmmb;

inline int pad_k(int k) {
    if (k <= 16) return 16;
    if (k <= 32) return 32;
    if (k <= 64) return 64;
    return 128;
}

inline int cuda_fail(sdmmb_ctx* c, cudaError_t e) {
    if (c) c->last_cuda_error = cudaGetErrorString(e);
    return (e == cudaErrorMemoryAllocation) ? SDMMB_E_NOMEM : SDMMB_E_CUDA;
}

#define CK(call)                                          \
    do {                                                  \
        cudaError_t _e = (call);                          \
        if (_e != cudaSuccess) return cuda_fail(ctx, _e); \
    } while (0)

// serialises the entry points of one context (see sdmmb_ctx::mu)
#define CTX_LOCK(c) std::lock_guard<std::recursive_mutex> _ctx_lock((c)->mu)

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
        else prev = -1;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

inline int ensure(sdmmb_ctx* ctx, void** p, size_t* cap, size_t bytes) {
    if (bytes <= *cap) return SDMMB_OK;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *cap = 0;
    size_t want = bytes + bytes / 8;
    CK(cudaMalloc(p, want));
    *cap = want;
    return SDMMB_OK;
}

// Binning scratch of the query / product / collection paths, one grow-only buffer per caller stream: work that
// different host threads enqueue on different streams never shares a buffer (the context lock only covers the
// enqueue, not the execution).
inline int stream_scratch(sdmmb_ctx* ctx, cudaStream_t st, size_t bytes, void** out) {
    sdmmb_ctx::QScratch& q = ctx->qscratch_by_stream[st];
    int rc = ensure(ctx, &q.p, &q.cap, bytes);
    *out = q.p;
    return rc;
}

inline bool range_ok(sdmmb_ctx* ctx, int first, int n) {
    return first >= 0 && n >= 0 && (long long)first + n <= ctx->max_cells;
}

inline MStepArgs base_mstep_args(sdmmb_ctx* ctx, int first_cell, int n_cells, int mode) {
    MStepArgs a{};
    a.K = ctx->K;
    a.first_cell = first_cell;
    a.n_cells = n_cells;
    a.mode = mode;
    a.joint = ctx->d_joint;
    a.em = ctx->d_em;
    a.scal = ctx->d_scal;
    a.cdf = ctx->d_cdf;
    a.erec = ctx->d_erec;
    a.qrec = ctx->d_qrec_back ? ctx->d_qrec_back : ctx->d_qrec;
    a.min_samples = ctx->min_samples;
    a.peer.world = 0;
    return a;
}

// capi_fit.cu
int bind_layout(sdmmb_ctx* ctx, int first_cell, int n_cells, const uint32_t* h_off, cudaStream_t st, bool reserve = true);
int arm_peer_push(sdmmb_ctx* ctx, PeerPush out[2], cudaStream_t st);
int ensure_pipe(sdmmb_ctx* ctx);
// capi_query.cu
int query_common(sdmmb_ctx* ctx, size_t n, const uint32_t* d_cell, const float* d_pos3, const float* d_u3,
                 const float* d_dir3, float* d_out_dir3, float* d_out_pdf, float* d_out_ij, uint32_t* d_out_comp,
                 uint8_t* d_out_valid, int binned, cudaStream_t st);

}  // namespace sdmmb_capi
