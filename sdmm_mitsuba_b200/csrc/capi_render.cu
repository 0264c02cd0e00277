// capi_render.cu -- host side of the device path tracer (include/sdmm_b200_render.h; kernels in render_kernels.cu):
// the progressive training loop's render pass with device-resident mixtures (rows F1 of SURVEY.md section 8f).
#include "capi_internal.h"

#include <vector>

using namespace sdmmb;
using namespace sdmmb_capi;

struct sdmmb_pt {
    sdmmb_ctx* ctx = nullptr;
    PtScene h_scene{};
    PtScene* d_scene = nullptr;
    void* d_paths = nullptr;
    size_t paths_cap = 0;      // in paths
    void* d_lists = nullptr;
    size_t lists_cap = 0;      // in paths
    uint32_t* d_counters = nullptr;
    // sampleProduct: scene-wide learned-BSDF mixture [n_bsdf][7] on the device (0 = plain conditional guiding)
    float* d_bsdf = nullptr;
    int n_bsdf = 0;
    // per leaf (cell slot): the box STree::find arrives at, [2][max_cells] float4, rebuilt when the context's tree changed
    float4* d_leaf_box = nullptr;
    uint64_t box_version = 0;
};

namespace {

size_t pt_path_bytes() {
    // thr li (6 f) + rng (2 u64) + depth vcount pword (3 w) + pend (5 f) + verts
    return (6 + 4 + 3 + 5 + (size_t)PT_MAX_VERTS * PT_VERT_FLOATS) * 4;
}

void pt_carve_paths(sdmmb_pt* pt, size_t n, PtPaths& P) {
    char* b = (char*)pt->d_paths;
    auto take = [&](size_t bytes) { char* r = b; b += (bytes + 255) / 256 * 256; return r; };
    P.n = n;
    P.rng = (uint64_t*)take(2 * n * 8);
    P.thr = (float*)take(3 * n * 4);
    P.li = (float*)take(3 * n * 4);
    P.depth = (int32_t*)take(n * 4);
    P.vcount = (int32_t*)take(n * 4);
    P.pword = (uint32_t*)take(n * 4);
    P.pend = (float*)take(5 * n * 4);
    P.verts = (float*)take((size_t)PT_MAX_VERTS * PT_VERT_FLOATS * n * 4);
}

void pt_carve_lists(sdmmb_pt* pt, size_t n, PtLists& L) {
    char* b = (char*)pt->d_lists;
    auto take = [&](size_t bytes) { char* r = b; b += (bytes + 255) / 256 * 256; return r; };
    L.recA = (float*)take(8 * n * 4);
    L.recB = (float*)take(8 * n * 4);
    L.cellA = (uint32_t*)take(n * 4); L.posA = (float*)take(3 * n * 4); L.uA = (float*)take(3 * n * 4);
    L.dirA = (float*)take(3 * n * 4); L.pdfA = (float*)take(n * 4); L.ijA = (float*)take(n * 4);
    L.compA = (uint32_t*)take(n * 4); L.validA = (uint8_t*)take(n);
    L.cellB = (uint32_t*)take(n * 4); L.posB = (float*)take(3 * n * 4); L.dirB = (float*)take(3 * n * 4);
    L.pdfB = (float*)take(n * 4); L.validB = (uint8_t*)take(n);
    L.nrmA = (float*)take(3 * n * 4); L.nrmB = (float*)take(3 * n * 4);
    L.counters = pt->d_counters;
}

// Bottom-up clustering of the scene's quads for the bounce kernel's culling: start from one cluster per quad and keep
// merging the two clusters whose common bounding box has the smallest surface area, clusters of at most 6 quads, until
// about n / 4 clusters are left (the faces of a small box end up together, a wall spanning the scene stays alone).
// The quads are then stored in cluster order; boxes are enlarged by 1e-4 of the scene's extent.
void cluster_quads(PtScene& s) {
    const int n = s.n_quads;
    struct Box { float lo[3], hi[3]; };
    auto area = [](const Box& b) {
        const float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
        return 2.f * (dx * dy + dy * dz + dz * dx);
    };
    auto join = [](const Box& a, const Box& b) {
        Box r;
        for (int c = 0; c < 3; ++c) { r.lo[c] = std::min(a.lo[c], b.lo[c]); r.hi[c] = std::max(a.hi[c], b.hi[c]); }
        return r;
    };
    std::vector<std::vector<int>> members(n);
    std::vector<Box> box(n);
    Box scene{{3e38f, 3e38f, 3e38f}, {-3e38f, -3e38f, -3e38f}};
    for (int q = 0; q < n; ++q) {
        members[q] = {q};
        const PtQuad& Q = s.quads[q];
        Box b{{3e38f, 3e38f, 3e38f}, {-3e38f, -3e38f, -3e38f}};
        for (int corner = 0; corner < 4; ++corner)
            for (int c = 0; c < 3; ++c) {
                const float v = Q.p0[c] + ((corner & 1) ? Q.e1[c] : 0.f) + ((corner & 2) ? Q.e2[c] : 0.f);
                b.lo[c] = std::min(b.lo[c], v);
                b.hi[c] = std::max(b.hi[c], v);
            }
        box[q] = b;
        scene = join(scene, b);
    }
    // SDMMB_PT_NO_CULL=1 (tests): one cluster holding every quad in the caller's order = the plain loop over all quads
    const char* no_cull = getenv("SDMMB_PT_NO_CULL");
    if (no_cull && atoi(no_cull) != 0) {
        members.assign(1, std::vector<int>());
        for (int q = 0; q < n; ++q) members[0].push_back(q);
        box.assign(1, scene);
    }
    const size_t target = (size_t)std::max(1, (n + 3) / 4);
    while (members.size() > target) {
        int bi = -1, bj = -1;
        float best = 3e38f;
        for (size_t i = 0; i < members.size(); ++i)
            for (size_t j = i + 1; j < members.size(); ++j) {
                if (members[i].size() + members[j].size() > 6) continue;
                const float a = area(join(box[i], box[j]));
                if (a < best) { best = a; bi = (int)i; bj = (int)j; }
            }
        if (bi < 0) break;
        members[bi].insert(members[bi].end(), members[bj].begin(), members[bj].end());
        box[bi] = join(box[bi], box[bj]);
        members.erase(members.begin() + bj);
        box.erase(box.begin() + bj);
    }
    float extent = 0.f;
    for (int c = 0; c < 3; ++c) extent = std::max(extent, scene.hi[c] - scene.lo[c]);
    const float margin = 1e-4f * extent + 1e-6f;
    PtQuad sorted[PT_MAX_QUADS];
    int at = 0;
    s.n_groups = (int)members.size();
    for (size_t g = 0; g < members.size(); ++g) {
        std::sort(members[g].begin(), members[g].end());
        PtGroup& G = s.groups[g];
        G.first = at;
        G.count = (int)members[g].size();
        for (int c = 0; c < 3; ++c) { G.bmin[c] = box[g].lo[c] - margin; G.bmax[c] = box[g].hi[c] + margin; }
        for (int q : members[g]) {
            s.orig[at] = q;
            sorted[at++] = s.quads[q];
        }
    }
    for (int q = 0; q < n; ++q) s.quads[q] = sorted[q];
}

// The box of every leaf as the descent of STree::find rebuilds it from the root box (spatial_tree.h:19-51: child 0 is
// the upper half, its min along the split axis = the split; child 1 the lower half, its max = the split) -- pure
// copies of split values, so the table holds exactly what a per-vertex rebuild during the descent would produce.
int update_leaf_boxes(sdmmb_ctx* ctx, sdmmb_pt* pt, cudaStream_t st) {
    if (pt->d_leaf_box && pt->box_version == ctx->tree_version) return SDMMB_OK;
    const size_t C = (size_t)ctx->max_cells;
    if (!pt->d_leaf_box) CK(cudaMalloc(&pt->d_leaf_box, 2 * C * sizeof(float4)));
    std::vector<float4> box(2 * C, make_float4(0.f, 0.f, 0.f, 0.f));
    struct Item { uint32_t node; float lo[3], hi[3]; };
    std::vector<Item> stack;
    Item root{};
    for (int c = 0; c < 3; ++c) { root.lo[c] = pt->h_scene.root_min[c]; root.hi[c] = pt->h_scene.root_max[c]; }
    stack.push_back(root);
    while (!stack.empty()) {
        Item it = stack.back();
        stack.pop_back();
        if (it.node >= ctx->h_tree.size()) continue;
        const TreeNode& nd = ctx->h_tree[it.node];
        if (nd.axis < 0) {
            if (nd.child0 < C) {
                box[nd.child0] = make_float4(it.lo[0], it.lo[1], it.lo[2], 0.f);
                box[C + nd.child0] = make_float4(it.hi[0], it.hi[1], it.hi[2], 0.f);
            }
            continue;
        }
        Item up = it, down = it;
        up.node = nd.child0;
        up.lo[nd.axis] = nd.split;
        down.node = nd.child1;
        down.hi[nd.axis] = nd.split;
        stack.push_back(up);
        stack.push_back(down);
    }
    CK(cudaStreamSynchronize(st));
    CK(cudaMemcpy(pt->d_leaf_box, box.data(), 2 * C * sizeof(float4), cudaMemcpyHostToDevice));
    pt->box_version = ctx->tree_version;
    return SDMMB_OK;
}

// one product-guiding query launch over a list of the tracer (the BSDF table already lives on the device)
int product_query(sdmmb_ctx* ctx, sdmmb_pt* pt, size_t n, const uint32_t* cell, const float* pos3, const float* u3,
                  const float* dir3, const float* normal3, float* out_dir3, float* out_pdf, float* out_ij,
                  uint32_t* out_comp, uint8_t* out_valid, cudaStream_t st) {
    void* scratch = nullptr;
    int rc = stream_scratch(ctx, st, (2 * ((size_t)ctx->max_cells + 1) + 9 * n + 4) * 4, &scratch);
    if (rc) return rc;
    ProductLaunch q{};
    q.scratch = (uint32_t*)scratch;
    q.K = ctx->K;
    q.K_user = ctx->K_user;
    q.max_cells = ctx->max_cells;
    q.n = n;
    q.qrec = ctx->d_qrec;
    q.cell_idx = cell;
    q.pos3 = pos3;
    q.u3 = u3;
    q.dir3 = dir3;
    q.normal3 = normal3;
    q.M = pt->n_bsdf;
    q.bsdf = pt->d_bsdf;
    q.out_dir3 = out_dir3;
    q.out_pdf = out_pdf;
    q.out_inv_jac = out_ij;
    q.out_comp = out_comp;
    q.out_valid = out_valid;
    CK(launch_product_query(q, st));
    ctx->launches += 4;
    return SDMMB_OK;
}

}  // namespace

extern "C" {

int sdmmb_pt_set_product(sdmmb_pt* pt, int n_bsdf, const float* h_bsdf) {
    if (!pt || n_bsdf < 0 || n_bsdf > SDMMB_PRODUCT_MAX_BSDF || (n_bsdf > 0 && !h_bsdf)) return SDMMB_E_INVALID;
    sdmmb_ctx* ctx = pt->ctx;
    for (int m = 0; m < n_bsdf; ++m) {
        const float* b = h_bsdf + 7 * m;
        if (!(b[0] >= 0.f) || !(b[1] * b[1] + b[2] * b[2] + b[3] * b[3] > 0.f)) return SDMMB_E_INVALID;
    }
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    CK(cudaDeviceSynchronize());
    if (n_bsdf > 0) {
        if (!pt->d_bsdf) CK(cudaMalloc(&pt->d_bsdf, (size_t)SDMMB_PRODUCT_MAX_BSDF * 7 * 4));
        CK(cudaMemcpy(pt->d_bsdf, h_bsdf, (size_t)n_bsdf * 7 * 4, cudaMemcpyHostToDevice));
    }
    pt->n_bsdf = n_bsdf;
    return SDMMB_OK;
}

int sdmmb_pt_create(sdmmb_ctx* ctx, sdmmb_pt** out, const sdmmb_quad* h_quads, int n_quads,
                    const sdmmb_camera* cam, const sdmmb_pt_config* cfg) {
    if (!ctx || !out || !h_quads || !cam || !cfg) return SDMMB_E_INVALID;
    if (n_quads <= 0 || n_quads > PT_MAX_QUADS) return SDMMB_E_RANGE;
    if (cfg->max_depth < 1 || cfg->max_depth > PT_MAX_VERTS + 1 || cfg->saved_per_path < 0) return SDMMB_E_RANGE;
    if (cam->width <= 0 || cam->height <= 0 || !(cfg->spatial_norm > 0.f)) return SDMMB_E_INVALID;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    sdmmb_pt* pt = new (std::nothrow) sdmmb_pt();
    if (!pt) return SDMMB_E_NOMEM;
    pt->ctx = ctx;
    PtScene& s = pt->h_scene;
    s.n_quads = n_quads;
    for (int q = 0; q < n_quads; ++q) {
        PtQuad& Q = s.quads[q];
        const sdmmb_quad& I = h_quads[q];
        for (int c = 0; c < 3; ++c) {
            Q.p0[c] = I.p0[c]; Q.e1[c] = I.e1[c]; Q.e2[c] = I.e2[c];
            Q.reflectance[c] = I.reflectance[c]; Q.emission[c] = I.emission[c];
        }
        const float nx = I.e1[1] * I.e2[2] - I.e1[2] * I.e2[1], ny = I.e1[2] * I.e2[0] - I.e1[0] * I.e2[2],
                    nz = I.e1[0] * I.e2[1] - I.e1[1] * I.e2[0];
        const float len = std::sqrt(nx * nx + ny * ny + nz * nz);
        const float l1 = I.e1[0] * I.e1[0] + I.e1[1] * I.e1[1] + I.e1[2] * I.e1[2];
        const float l2 = I.e2[0] * I.e2[0] + I.e2[1] * I.e2[1] + I.e2[2] * I.e2[2];
        if (!(len > 0.f) || !(l1 > 0.f) || !(l2 > 0.f)) { delete pt; return SDMMB_E_INVALID; }
        Q.n[0] = nx / len; Q.n[1] = ny / len; Q.n[2] = nz / len;
        Q.inv_len2[0] = 1.f / l1; Q.inv_len2[1] = 1.f / l2;
    }
    cluster_quads(s);
    for (int c = 0; c < 3; ++c) {
        s.cam_origin[c] = cam->origin[c]; s.cam_right[c] = cam->right[c]; s.cam_up[c] = cam->up[c];
        s.cam_forward[c] = cam->forward[c];
        s.aabb_min[c] = cfg->aabb_min[c]; s.root_min[c] = cfg->root_min[c]; s.root_max[c] = cfg->root_max[c];
    }
    s.width = cam->width; s.height = cam->height;
    s.tan_x = std::tan(0.5f * cam->fov_x_degrees * 3.14159265358979323846f / 180.f);
    s.tan_y = s.tan_x * (float)cam->height / (float)cam->width;
    s.inv_spatial_norm = 1.f / cfg->spatial_norm;
    s.max_depth = cfg->max_depth;
    s.saved_per_path = cfg->saved_per_path;
    cudaError_t e = cudaMalloc(&pt->d_scene, sizeof(PtScene));
    if (e == cudaSuccess) e = cudaMemcpy(pt->d_scene, &s, sizeof(PtScene), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMalloc(&pt->d_counters, 16 * 4);
    if (e != cudaSuccess) { sdmmb_pt_destroy(pt); return cuda_fail(ctx, e); }
    *out = pt;
    return SDMMB_OK;
}

int sdmmb_pt_destroy(sdmmb_pt* pt) {
    if (!pt) return SDMMB_OK;
    DeviceGuard g(pt->ctx->device);
    cudaFree(pt->d_scene); cudaFree(pt->d_paths); cudaFree(pt->d_lists); cudaFree(pt->d_counters); cudaFree(pt->d_bsdf); cudaFree(pt->d_leaf_box);
    delete pt;
    return SDMMB_OK;
}

int sdmmb_pt_pool_stats(sdmmb_ctx* ctx, const float* d_pool, size_t pool_capacity, size_t n, size_t new_begin,
                        int32_t* d_out_cell, double* d_out_stats, void* stream) {
    if (!ctx || !d_out_stats || n > pool_capacity || (n && (!d_pool || !d_out_cell))) return SDMMB_E_INVALID;
    if (!ctx->d_tree) return SDMMB_E_STATE;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    SDMMB_NVTX("sdmmb pool cells + SpatialStats");
    CK(pt_launch_pool_stats(ctx->d_tree, ctx->n_tree_nodes, d_pool, pool_capacity, n, new_begin, (uint32_t)ctx->max_cells,
                            d_out_cell, d_out_stats, (cudaStream_t)stream));
    if (n) ctx->launches += 1;
    return SDMMB_OK;
}

int sdmmb_pt_rebucket(sdmmb_ctx* ctx, const float* d_pool, size_t pool_capacity, size_t n, int32_t* d_cell,
                      const int32_t* d_tab_i, const float* d_tab_f, const float* d_box, double* d_out_stats, void* stream) {
    if (!ctx || !d_out_stats || !d_tab_i || !d_tab_f || !d_box || n > pool_capacity || (n && (!d_pool || !d_cell)))
        return SDMMB_E_INVALID;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    SDMMB_NVTX("sdmmb split level: re-bucket + child SpatialStats");
    CK(pt_launch_rebucket(d_pool, pool_capacity, n, d_cell, d_tab_i, d_tab_f, d_box, (uint32_t)ctx->max_cells, d_out_stats,
                          (cudaStream_t)stream));
    if (n) ctx->launches += 1;
    return SDMMB_OK;
}

int sdmmb_pt_render(sdmmb_pt* pt, int spp, uint64_t seed, int guided, int collect,
                    const uint8_t* d_cell_initialized, float* d_image_sum, float* d_pool,
                    size_t pool_capacity, size_t pool_count, sdmmb_pt_stats* out, void* stream) {
    if (!pt || spp <= 0 || !d_image_sum) return SDMMB_E_INVALID;
    if (collect && (!d_pool || pool_count > pool_capacity)) return SDMMB_E_INVALID;
    sdmmb_ctx* ctx = pt->ctx;
    if ((guided || collect) && !ctx->d_tree) return SDMMB_E_STATE;
    if (guided && !d_cell_initialized) return SDMMB_E_INVALID;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const size_t n_total = (size_t)pt->h_scene.width * pt->h_scene.height * (size_t)spp;
    if (n_total >= (1u << 27)) return SDMMB_E_DATA_FULL;   // push-sequence ids are path * 32 + ... in 32 bits
    // A pass is traced in batches of at most `batch` paths: the per-path state (604 B) and the query lists (184 B) are
    // allocated once for a batch instead of growing with spp; a path's pixel, random stream and push-sequence ids
    // follow its GLOBAL index, so the result does not depend on the batch size.
    static const size_t batch_env = [] { const char* v = getenv("SDMMB_PT_BATCH"); return v ? (size_t)atoll(v) : (size_t)0; }();
    const size_t batch = std::min(std::min(n_total, batch_env ? batch_env : ((size_t)16 << 20)), (size_t)1 << 24);   // list slots are 24 bits
    if (batch > pt->paths_cap) {
        cudaFree(pt->d_paths); pt->d_paths = nullptr; pt->paths_cap = 0;
        CK(cudaMalloc(&pt->d_paths, batch * pt_path_bytes() + 64 * 256));
        pt->paths_cap = batch;
    }
    if (guided && batch > pt->lists_cap) {
        cudaFree(pt->d_lists); pt->d_lists = nullptr; pt->lists_cap = 0;
        CK(cudaMalloc(&pt->d_lists, batch * 184 + 64 * 256));  // 174 B of records + query planes per path
        pt->lists_cap = batch;
    }
    if (collect) {
        int rc = update_leaf_boxes(ctx, pt, st);
        if (rc) return rc;
    }
    PtFrame F{};
    F.leaf_box = pt->d_leaf_box;
    F.tree = ctx->d_tree;
    F.n_tree_nodes = ctx->n_tree_nodes;
    F.cell_initialized = d_cell_initialized;
    F.max_cells = (uint32_t)ctx->max_cells;
    F.guided = guided ? 1 : 0;
    F.w_bsdf = pt->n_bsdf > 0 ? 0.3f : 0.5f;   // heuristicConditionalWeight with / without the product  R: 383-390
    F.collect = collect ? 1 : 0;
    F.seed = seed;
    F.spp = spp;
    F.image = d_image_sum;
    F.pool = d_pool;
    F.pool_cap = pool_capacity;
    uint32_t h_cnt[4] = {0u, 0u, (uint32_t)pool_count, 0u};
    CK(cudaMemcpyAsync(pt->d_counters, h_cnt, sizeof(h_cnt), cudaMemcpyHostToDevice, st));
    uint64_t launches = 0, qa = 0, qb = 0;   // `launches` = this tracer's own kernels (K4's are counted by query_common)
    const uint64_t ctx_before = ctx->launches;
    for (size_t base = 0; base < n_total; base += batch) {
        const size_t n = std::min(batch, n_total - base);
        PtPaths P{};
        pt_carve_paths(pt, n, P);
        PtLists L{};
        if (guided) pt_carve_lists(pt, n, L);
        L.counters = pt->d_counters;
        F.path_base = base;
        size_t nA = 0, nB = 0;
        for (int bounce = 0; bounce < pt->h_scene.max_depth; ++bounce) {
            CK(cudaMemsetAsync(pt->d_counters, 0, 8, st));
            CK(pt_launch_bounce(pt->d_scene, P, L, F, bounce == 0 ? 1 : 0, nA, nB, st));
            ++launches;
            nA = nB = 0;
            if (!guided) continue;
            uint32_t c2[2];
            CK(cudaMemcpyAsync(c2, pt->d_counters, 8, cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            nA = c2[0];
            nB = c2[1];
            if (nA + nB == 0) continue;
            CK(pt_launch_pack_lists(pt->d_scene, L, nA, nB, st));
            ++launches;
            if (pt->n_bsdf > 0) {
                // sampleProduct: conditional x learned-BSDF mixture, component 0 re-centred on the shading normal
                // (every surface here is diffuse: sdmm_proc.cpp:335-339), sampled / evaluated per vertex (:375-381, 409-417)
                if (nA) {
                    int rc = product_query(ctx, pt, nA, L.cellA, L.posA, L.uA, nullptr, L.nrmA, L.dirA, L.pdfA, L.ijA, L.compA, L.validA, st);
                    if (rc) return rc;
                }
                if (nB) {
                    int rc = product_query(ctx, pt, nB, L.cellB, L.posB, nullptr, L.dirB, L.nrmB, nullptr, L.pdfB, nullptr, nullptr, L.validB, st);
                    if (rc) return rc;
                }
            } else {
                if (nA) {
                    int rc = query_common(ctx, nA, L.cellA, L.posA, L.uA, nullptr, L.dirA, L.pdfA, L.ijA, L.compA, L.validA, 0, st);
                    if (rc) return rc;
                }
                if (nB) {
                    int rc = query_common(ctx, nB, L.cellB, L.posB, nullptr, L.dirB, nullptr, L.pdfB, nullptr, nullptr, L.validB, 0, st);
                    if (rc) return rc;
                }
            }
            qa += nA;
            qb += nB;
        }
        CK(pt_launch_finish(pt->d_scene, P, L, F, st));
        ++launches;
    }
    CK(cudaMemcpyAsync(h_cnt, pt->d_counters, sizeof(h_cnt), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    ctx->launches += launches;
    if (out) {
        out->paths = n_total;
        out->guided_queries = qa;
        out->mis_queries = qb;
        out->pool_count = h_cnt[2] < pool_capacity ? h_cnt[2] : pool_capacity;
        out->pool_overflow = h_cnt[3];
        out->launches = ctx->launches - ctx_before;
    }
    return SDMMB_OK;
}

}  // extern "C"
