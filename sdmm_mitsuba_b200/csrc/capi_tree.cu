// capi_tree.cu -- spatial cell structure on the device (row F2 of SURVEY.md section 8f): the flattened STree node
// table, STree::find for batches of points, and the renderer's per-sample find + push_back as one counting sort
// into the E-step's sample records.  Reference: sdmm/include/sdmm/accelerators/spatial_tree.h:19-51,
// mitsuba/src/integrators/sdmm/sdmm_proc.cpp:880-928.
#include "capi_internal.h"

#include <vector>

using namespace sdmmb;
using namespace sdmmb_capi;

extern "C" {

int sdmmb_set_tree(sdmmb_ctx* ctx, int n_nodes, const sdmmb_tree_node* h_nodes) {
    if (!ctx || n_nodes <= 0 || !h_nodes) return SDMMB_E_INVALID;
    static_assert(sizeof(sdmmb_tree_node) == sizeof(TreeNode), "tree node layout");
    for (int i = 0; i < n_nodes; ++i) {
        const sdmmb_tree_node& nd = h_nodes[i];
        if (nd.axis > 2) return SDMMB_E_INVALID;
        if (nd.axis >= 0 && (nd.child0 >= (uint32_t)n_nodes || nd.child1 >= (uint32_t)n_nodes)) return SDMMB_E_INVALID;
        if (nd.axis < 0 && nd.child0 >= (uint32_t)ctx->max_cells) return SDMMB_E_RANGE;
    }
    // The device copy is stored BREADTH-FIRST (node 0 = root, then level by level): a descent is a chain of dependent
    // node loads, and the kernels that descend per path vertex stage the first nodes -- the top levels -- in shared
    // memory.  Node indices never leave the library (a find returns the leaf's cell slot), so the order is free.
    // Nodes not reachable from the root are dropped.
    std::vector<sdmmb_tree_node> bfs;
    bfs.reserve((size_t)n_nodes);
    {
        std::vector<int> order, new_id((size_t)n_nodes, -1);
        order.reserve((size_t)n_nodes);
        order.push_back(0);
        new_id[0] = 0;
        for (size_t at = 0; at < order.size(); ++at) {
            const sdmmb_tree_node& nd = h_nodes[order[at]];
            if (nd.axis < 0) continue;
            for (uint32_t c : {nd.child0, nd.child1}) {
                if (new_id[c] >= 0) return SDMMB_E_INVALID;   // a node with two parents: not a tree
                new_id[c] = (int)order.size();
                order.push_back((int)c);
            }
        }
        for (int old : order) {
            sdmmb_tree_node nd = h_nodes[old];
            if (nd.axis >= 0) { nd.child0 = (uint32_t)new_id[nd.child0]; nd.child1 = (uint32_t)new_id[nd.child1]; }
            bfs.push_back(nd);
        }
    }
    const int n_bfs = (int)bfs.size();
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    CK(cudaDeviceSynchronize());
    cudaFree(ctx->d_tree);
    ctx->d_tree = nullptr;
    ctx->n_tree_nodes = 0;
    CK(cudaMalloc(&ctx->d_tree, (size_t)n_bfs * sizeof(TreeNode)));
    CK(cudaMemcpy(ctx->d_tree, bfs.data(), (size_t)n_bfs * sizeof(TreeNode), cudaMemcpyHostToDevice));
    ctx->n_tree_nodes = n_bfs;
    ctx->h_tree.resize((size_t)n_bfs);
    std::memcpy(ctx->h_tree.data(), bfs.data(), (size_t)n_bfs * sizeof(TreeNode));
    ++ctx->tree_version;
    return SDMMB_OK;
}

int sdmmb_find_cells(sdmmb_ctx* ctx, size_t n, const float* d_pos3, uint32_t* d_out_cell, void* stream) {
    if (!ctx || (n && (!d_pos3 || !d_out_cell))) return SDMMB_E_INVALID;
    if (!ctx->d_tree) return SDMMB_E_STATE;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    CK(launch_find_cells(ctx->d_tree, ctx->n_tree_nodes, d_pos3, n, n, d_out_cell, 0xffffffffu, (cudaStream_t)stream));
    if (n) ctx->launches += 1;
    return SDMMB_OK;
}

int sdmmb_find_cells_host(sdmmb_ctx* ctx, size_t n, const float* h_pos3, uint32_t* h_out_cell, void* stream) {
    if (!ctx || (n && (!h_pos3 || !h_out_cell))) return SDMMB_E_INVALID;
    if (!ctx->d_tree) return SDMMB_E_STATE;
    if (n == 0) return SDMMB_OK;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    int rc = ensure(ctx, &ctx->d_stage2, &ctx->stage2_cap, n * 16);
    if (rc) return rc;
    float* d_pos = (float*)ctx->d_stage2;
    uint32_t* d_out = (uint32_t*)(d_pos + 3 * n);
    CK(cudaMemcpyAsync(d_pos, h_pos3, n * 12, cudaMemcpyHostToDevice, st));
    CK(launch_find_cells(ctx->d_tree, ctx->n_tree_nodes, d_pos, n, n, d_out, 0xffffffffu, st));
    ctx->launches += 1;
    CK(cudaMemcpyAsync(h_out_cell, d_out, n * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return SDMMB_OK;
}

int sdmmb_collect_samples(sdmmb_ctx* ctx, size_t n, const float* d_point6, size_t plane_stride,
                          const float* d_weight, const float* d_key3, size_t key_stride,
                          uint32_t* h_out_counts, uint64_t* h_out_dropped, void* stream) {
    if (!ctx || (n && (!d_point6 || !d_weight))) return SDMMB_E_INVALID;
    if (!ctx->d_tree) return SDMMB_E_STATE;
    if (n > 0x7fffffffu) return SDMMB_E_DATA_FULL;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t B = (uint32_t)ctx->max_cells + 1;   // trailing bucket = dropped samples
    // scratch (words): count[B] | cursor[B] | cell[n]
    int rc = ensure(ctx, &ctx->d_qscratch, &ctx->qscratch_cap, (2 * (size_t)B + n + 4) * 4);
    if (rc) return rc;
    uint32_t* count = (uint32_t*)ctx->d_qscratch;
    uint32_t* cursor = count + B;
    uint32_t* cell = cursor + B;
    const float* key = d_key3 ? d_key3 : d_point6;
    const size_t kstride = d_key3 ? key_stride : plane_stride;
    CK(launch_collect_count(ctx->d_tree, ctx->n_tree_nodes, key, kstride, d_weight, n, B, cell, count, cursor, st));
    ctx->launches += 2;
    std::vector<uint32_t> h_count(B);
    CK(cudaMemcpyAsync(h_count.data(), count, (size_t)B * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    std::vector<uint32_t> off((size_t)ctx->max_cells + 1, 0u);
    for (int c = 0; c < ctx->max_cells; ++c) off[c + 1] = off[c] + h_count[c];
    rc = bind_layout(ctx, 0, ctx->max_cells, off.data(), st);
    if (rc) return rc;
    CK(launch_collect_scatter(cell, d_point6, plane_stride, d_weight, n, B - 1, cursor, ctx->d_samples, st));
    if (n) ctx->launches += 1;
    if (h_out_counts) std::memcpy(h_out_counts, h_count.data(), (size_t)ctx->max_cells * 4);
    if (h_out_dropped) *h_out_dropped = h_count[B - 1];
    return SDMMB_OK;
}

}  // extern "C"
