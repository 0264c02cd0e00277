// capi_peer.cu -- how the refitted query records reach the other GPUs of the box: (a) record exchange over peer
// memory (NVLink / NVSwitch P2P) fused into the kernel that writes the records -- one process per GPU shares the
// record array and the flag words of every context through CUDA IPC handles; contexts of the SAME process (a C++
// host driving all GPUs, or the single-GPU tests) use each other's pointers directly -- and (b) the in-place
// ncclAllGather (libnccl resolved with dlopen) for hosts without P2P and as the comparison arm.
#include "capi_internal.h"

using namespace sdmmb;
using namespace sdmmb_capi;

namespace {

int load_nccl(sdmmb_ctx* ctx) {
    if (ctx->nccl_allgather) return SDMMB_OK;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);  // the copy torch already loaded
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW);
    if (!h) return SDMMB_E_NCCL;
    ctx->nccl_handle = h;
    ctx->nccl_allgather = (decltype(ctx->nccl_allgather))dlsym(h, "ncclAllGather");
    ctx->nccl_user_rank = (decltype(ctx->nccl_user_rank))dlsym(h, "ncclCommUserRank");
    return (ctx->nccl_allgather && ctx->nccl_user_rank) ? SDMMB_OK : SDMMB_E_NCCL;
}

struct PeerBlob {               // SDMMB_PEER_HANDLE_BYTES, opaque to the caller
    uint64_t pid;
    int32_t device;
    int32_t max_cells;
    int32_t K;
    int32_t reserved;
    uint64_t qrec_ptr;
    uint64_t flags_ptr;
    cudaIpcMemHandle_t qrec;
    cudaIpcMemHandle_t flags;
};
static_assert(sizeof(PeerBlob) <= SDMMB_PEER_HANDLE_BYTES, "peer handle blob too large");
}  // namespace

extern "C" {

int sdmmb_peer_export(sdmmb_ctx* ctx, unsigned char* h_handle) {
    if (!ctx || !h_handle) return SDMMB_E_INVALID;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    if (!ctx->d_peer_flags) {
        CK(cudaMalloc(&ctx->d_peer_flags, PEER_FLAG_WORDS * sizeof(uint32_t)));
        CK(cudaMemset(ctx->d_peer_flags, 0, PEER_FLAG_WORDS * sizeof(uint32_t)));
    }
    PeerBlob b{};
    b.pid = (uint64_t)getpid();
    b.device = ctx->device;
    b.max_cells = ctx->max_cells;
    b.K = ctx->K;
    b.qrec_ptr = (uint64_t)(uintptr_t)ctx->d_qrec;
    b.flags_ptr = (uint64_t)(uintptr_t)ctx->d_peer_flags;
    CK(cudaIpcGetMemHandle(&b.qrec, ctx->d_qrec));
    CK(cudaIpcGetMemHandle(&b.flags, ctx->d_peer_flags));
    std::memset(h_handle, 0, SDMMB_PEER_HANDLE_BYTES);
    std::memcpy(h_handle, &b, sizeof(b));
    return SDMMB_OK;
}

int sdmmb_peer_detach(sdmmb_ctx* ctx) {
    if (!ctx) return SDMMB_E_INVALID;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    if (ctx->peer.world > 0) CK(cudaDeviceSynchronize());
    for (void* m : ctx->peer_ipc_opened) cudaIpcCloseMemHandle(m);
    ctx->peer_ipc_opened.clear();
    ctx->peer = PeerPush{};
    ctx->peer_epoch = 0;
    if (ctx->d_peer_flags) CK(cudaMemset(ctx->d_peer_flags, 0, PEER_FLAG_WORDS * sizeof(uint32_t)));
    return SDMMB_OK;
}

int sdmmb_peer_attach(sdmmb_ctx* ctx, int rank, int world, const unsigned char* h_handles) {
    if (!ctx || !h_handles || world < 1 || world > PEER_MAX || rank < 0 || rank >= world) return SDMMB_E_INVALID;
    if (!ctx->d_peer_flags) return SDMMB_E_STATE;   // sdmmb_peer_export first
    int rc = sdmmb_peer_detach(ctx);
    if (rc) return rc;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    PeerPush P{};
    P.world = world;
    P.rank = rank;
    const char* tmo = getenv("SDMMB_PEER_TIMEOUT_MS");
    P.timeout_ms = tmo ? (uint32_t)atoi(tmo) : 10000u;
    for (int r = 0; r < world; ++r) {
        PeerBlob b;
        std::memcpy(&b, h_handles + (size_t)r * SDMMB_PEER_HANDLE_BYTES, sizeof(b));
        if (b.K != ctx->K || b.max_cells != ctx->max_cells) return SDMMB_E_INVALID;
        if (r == rank) {
            P.qrec[r] = ctx->d_qrec;
            P.flags[r] = ctx->d_peer_flags;
        } else if (b.pid == (uint64_t)getpid()) {
            if (b.device != ctx->device) {
                int can = 0;
                CK(cudaDeviceCanAccessPeer(&can, ctx->device, b.device));
                if (!can) return SDMMB_E_PEER;
                cudaError_t e = cudaDeviceEnablePeerAccess(b.device, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return cuda_fail(ctx, e);
                cudaGetLastError();
            }
            P.qrec[r] = (float*)(uintptr_t)b.qrec_ptr;
            P.flags[r] = (uint32_t*)(uintptr_t)b.flags_ptr;
        } else {
            void *q = nullptr, *f = nullptr;
            cudaError_t e = cudaIpcOpenMemHandle(&q, b.qrec, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) { cuda_fail(ctx, e); cudaGetLastError(); return SDMMB_E_PEER; }
            ctx->peer_ipc_opened.push_back(q);
            e = cudaIpcOpenMemHandle(&f, b.flags, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) { cuda_fail(ctx, e); cudaGetLastError(); return SDMMB_E_PEER; }
            ctx->peer_ipc_opened.push_back(f);
            P.qrec[r] = (float*)q;
            P.flags[r] = (uint32_t*)f;
        }
    }
    ctx->peer = P;
    ctx->peer_epoch = 0;
    return SDMMB_OK;
}

int sdmmb_peer_wait(sdmmb_ctx* ctx, void* stream) {
    if (!ctx) return SDMMB_E_INVALID;
    if (ctx->peer.world <= 1 || ctx->peer_epoch == 0) return SDMMB_OK;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    SDMMB_NVTX("sdmmb peer wait");
    PeerPush P = ctx->peer;
    P.epoch = ctx->peer_epoch;
    CK(launch_peer_wait(P, (cudaStream_t)stream));
    ctx->launches += 1;
    return SDMMB_OK;
}

int sdmmb_peer_status(sdmmb_ctx* ctx, uint32_t* out_epoch, uint32_t* out_timeouts, void* stream) {
    if (!ctx) return SDMMB_E_INVALID;
    if (out_epoch) *out_epoch = ctx->peer_epoch;
    if (out_timeouts) {
        *out_timeouts = 0;
        if (ctx->d_peer_flags) {
            CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
            CK(cudaMemcpyAsync(out_timeouts, ctx->d_peer_flags + PEER_ERROR, sizeof(uint32_t),
                               cudaMemcpyDeviceToHost, (cudaStream_t)stream));
            CK(cudaStreamSynchronize((cudaStream_t)stream));
        }
    }
    return SDMMB_OK;
}

int sdmmb_allgather_records(sdmmb_ctx* ctx, void* nccl_comm, int cells_per_rank, void* stream) {
    if (!ctx || !nccl_comm || cells_per_rank <= 0) return SDMMB_E_INVALID;
    int rc = load_nccl(ctx);
    if (rc) return rc;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    int rank = -1;
    if (ctx->nccl_user_rank(nccl_comm, &rank) != 0 || rank < 0) return SDMMB_E_NCCL;
    const size_t per_rank = (size_t)cells_per_rank * ctx->K * QREC_FLOATS;
    if (((size_t)rank + 1) * cells_per_rank > (size_t)ctx->max_cells) return SDMMB_E_RANGE;
    // in-place all-gather: this rank's slice already sits at its final offset
    const int nccl_float = 7;
    if (ctx->nccl_allgather(ctx->d_qrec + (size_t)rank * per_rank, ctx->d_qrec, per_rank, nccl_float,
                            nccl_comm, (cudaStream_t)stream) != 0)
        return SDMMB_E_NCCL;
    return SDMMB_OK;
}

}  // extern "C"
