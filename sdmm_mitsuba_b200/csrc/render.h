// render.h -- internal layouts of the device path tracer (render_kernels.cu) shared with the C ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "kernels.h"

namespace sdmmb {

constexpr int PT_MAX_QUADS = 64;
constexpr int PT_MAX_VERTS = 8;          // vertices kept per path (savedSamplesPerPath default, volpath_sdmm.cpp:60)
constexpr int PT_POOL_PLANES = 15;       // point6 | normal3 | weight | key3 | is_main | push sequence (u32 bits)
constexpr int PT_VERT_FLOATS = 16;       // npos3 dir3 normal3 weight3 thr3 pdf

struct PtQuad {
    float p0[3], e1[3], e2[3], n[3];
    float inv_len2[2];
    float reflectance[3];
    float emission[3];
};

// Quads clustered by proximity on the host (sdmmb_pt_create); a ray tests a cluster's quads only when it passes
// the cluster's (slightly enlarged) bounding box.  quads[] is stored in cluster order; orig[] is the caller's index
// of each quad: ties in the hit distance go to the lowest ORIGINAL index, as a loop over the caller's order gives.
struct PtGroup {
    float bmin[3], bmax[3];
    int first, count;
};

struct PtScene {
    PtQuad quads[PT_MAX_QUADS];
    int n_quads;
    PtGroup groups[PT_MAX_QUADS];
    int n_groups;
    int orig[PT_MAX_QUADS];
    // perspective camera: ray dir = forward + (1 - 2 sx / w) tan_x right + (1 - 2 sy / h) tan_y up
    float cam_origin[3], cam_right[3], cam_up[3], cam_forward[3];
    float tan_x, tan_y;
    int width, height;
    // createCondition: npos = (p - aabb_min) / spatial_norm                      (sdmm_proc.cpp:263-273)
    float aabb_min[3];
    float inv_spatial_norm;
    float root_min[3], root_max[3];   // STree root AABB in normalised coordinates (volpath_sdmm.cpp:314-330)
    int max_depth;
    int saved_per_path;
};

// per-path state, SoA over n_paths
struct PtPaths {
    size_t n;
    float* thr;      // [3][n]
    float* li;       // [3][n]
    uint64_t* rng;   // [2][n]
    int32_t* depth;  // rRec.depth, <= 0 : path finished
    int32_t* vcount; // stored vertices
    uint32_t* pword; // pending hit and sampling decision: quad (6 bits) | mode (2: 0 BSDF only, 1 BSDF + MIS pdf query,
                     // 2 guided) | slot in list A or B (24 bits: a batch holds at most 2^24 paths)
    float* pend;     // [5][n]: hit p(3), b1, b2
    float* verts;    // [PT_MAX_VERTS][PT_VERT_FLOATS][n]
};

struct PtLists {
    // AoS staging records written by the bounce kernel (slot allocated with atomicAdd)
    float* recA;        // [n][8]: cell(bits), npos3, g0, g1, g2, quad(bits)
    float* recB;        // [n][8]: cell(bits), npos3, dir3, quad(bits)
    uint32_t* counters; // [0] nA, [1] nB, [2] pool count, [3] pool overflow
    // dense SoA query inputs/outputs (stride = list length)
    uint32_t* cellA; float* posA; float* uA; float* dirA; float* pdfA; float* ijA; uint32_t* compA; uint8_t* validA;
    uint32_t* cellB; float* posB; float* dirB; float* pdfB; uint8_t* validB;
    float* nrmA; float* nrmB;   // [3][n] shading normals of the vertices (product guiding re-centres the BSDF lobe on them)
};

struct PtFrame {
    const TreeNode* tree;
    int n_tree_nodes;
    const float4* leaf_box;           // [2][max_cells]: min / max corner of every leaf's box as STree::find rebuilds it
    const uint8_t* cell_initialized;  // [max_cells]
    uint32_t max_cells;
    int guided;
    float w_bsdf;     // heuristicConditionalWeight: 0.5, or 0.3 with product guiding (sdmm_proc.cpp:383-390)
    int collect;
    uint64_t seed;
    int spp;
    float* image;     // [3][w*h] running sums
    float* pool;      // [PT_POOL_PLANES][pool_cap]
    size_t pool_cap;
    size_t path_base; // a pass is traced in batches of paths: global path id = path_base + index in the batch
};

cudaError_t pt_launch_bounce(const PtScene* d_scene, const PtPaths& P, const PtLists& L, const PtFrame& F, int first,
                             size_t nA, size_t nB, cudaStream_t st);
cudaError_t pt_launch_pack_lists(const PtScene* d_scene, const PtLists& L, size_t nA, size_t nB, cudaStream_t st);
cudaError_t pt_launch_pool_stats(const TreeNode* nodes, int n_nodes, const float* pool, size_t cap, size_t n, size_t new_begin,
                                 uint32_t max_cells, int32_t* out_cell, double* stats, cudaStream_t st);
cudaError_t pt_launch_rebucket(const float* pool, size_t cap, size_t n, int32_t* cell_of, const int32_t* tab_i, const float* tab_f,
                               const float* box, uint32_t max_cells, double* stats, cudaStream_t st);
cudaError_t pt_launch_finish(const PtScene* d_scene, const PtPaths& P, const PtLists& L, const PtFrame& F, cudaStream_t st);

}  // namespace sdmmb
