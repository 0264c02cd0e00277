// tree_kernels.cu -- the spatial cell structure on the device:
//   find_cells_kernel        STree::find per point            R: accelerators/spatial_tree.h:19-51
//   collect_* kernels        the renderer's sample collection: find the vertex's leaf cell and push the
//                            sample into that cell's Data (mitsuba/src/integrators/sdmm/sdmm_proc.cpp:
//                            880-928, Data::push_back R: opt/data.h:59-87) -- done for a whole batch of
//                            samples as a counting sort by cell straight into the E-step's packed records,
//                            instead of one mutex-protected push_back per sample (9.3 % of the reference's
//                            run time, optimization_avenues.md:8).
// The tree itself is built and split on the host by the reference's own code (serial, once per
// iteration); the host flattens it into TreeNode records (host_shim.hpp: flatten_tree).
#include <algorithm>

#include "kernels.h"

namespace sdmmb {
namespace {

// STree::find: at an inner node go to child 0 iff child0.aabb.min[axis] < p[axis]  (child 0 is the UPPER
// half: create_child raises its min, spatial_tree.h:178-181), else child 1; stop at a leaf.
__device__ __forceinline__ uint32_t find_leaf(const TreeNode* __restrict__ nodes, int n_nodes, float x, float y,
                                              float z, uint32_t miss) {
    uint32_t node = 0;
    for (int depth = 0; depth < 96; ++depth) {
        if (node >= (uint32_t)n_nodes) return miss;
        const int4 raw = __ldg(reinterpret_cast<const int4*>(nodes) + node);
        const int axis = raw.x;
        if (axis < 0) return (uint32_t)raw.z;  // leaf: child0 holds the cell slot
        const float split = __int_as_float(raw.y);
        const float p = axis == 0 ? x : (axis == 1 ? y : z);
        node = (split < p) ? (uint32_t)raw.z : (uint32_t)raw.w;
    }
    return miss;
}

__global__ void find_cells_kernel(const TreeNode* __restrict__ nodes, int n_nodes, const float* __restrict__ pos3,
                                  size_t stride, size_t n, uint32_t* __restrict__ out, uint32_t miss) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t step = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += step) out[i] = find_leaf(nodes, n_nodes, pos3[i], pos3[stride + i], pos3[2 * stride + i], miss);
}

// cell of every sample; samples Data::push_back would reject (weight < 1e-8 or non-finite,
// R: opt/data.h:20-23, 68-71) and samples without a cell go to the trailing "dropped" bucket
__global__ void collect_find_kernel(const TreeNode* __restrict__ nodes, int n_nodes, const float* __restrict__ key3,
                                    size_t key_stride, const float* __restrict__ weight, size_t n,
                                    uint32_t n_buckets, uint32_t* __restrict__ cell, uint32_t* __restrict__ count) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t step = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += step) {
        const float w = weight[i];
        uint32_t c = n_buckets - 1;
        if (w >= 1e-8f && w <= 3.402823466e+38f) {
            c = find_leaf(nodes, n_nodes, key3[i], key3[key_stride + i], key3[2 * key_stride + i], n_buckets - 1);
            if (c > n_buckets - 1) c = n_buckets - 1;
        }
        cell[i] = c;
        atomicAdd(&count[c], 1u);
    }
}

__global__ void collect_scan_kernel(const uint32_t* __restrict__ count, uint32_t n_buckets, uint32_t* __restrict__ cursor) {
    // single CTA exclusive scan (n_buckets <= a few 10^4); simple and off the critical path
    __shared__ uint32_t s_part[1024];
    const uint32_t per = (n_buckets + blockDim.x - 1) / blockDim.x;
    const uint32_t b = threadIdx.x * per, e = min(b + per, n_buckets);
    uint32_t sum = 0;
    for (uint32_t i = b; i < e; ++i) sum += count[i];
    s_part[threadIdx.x] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t run = 0;
        for (uint32_t t = 0; t < blockDim.x; ++t) { const uint32_t v = s_part[t]; s_part[t] = run; run += v; }
    }
    __syncthreads();
    uint32_t run = s_part[threadIdx.x];
    for (uint32_t i = b; i < e; ++i) { cursor[i] = run; run += count[i]; }
}

// scatter into the E-step's packed records: rec[2p] = (x, y, z, w), rec[2p+1] = (dx, dy, dz, w_eff)
__global__ void collect_scatter_kernel(const uint32_t* __restrict__ cell, const float* __restrict__ point6,
                                       size_t stride, const float* __restrict__ weight, size_t n, uint32_t dropped,
                                       uint32_t* __restrict__ cursor, float4* __restrict__ rec) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t step = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += step) {
        const uint32_t c = cell[i];
        if (c == dropped) continue;
        const uint32_t p = atomicAdd(&cursor[c], 1u);
        const float w = weight[i];
        rec[2 * (size_t)p] = make_float4(point6[i], point6[stride + i], point6[2 * stride + i], w);
        rec[2 * (size_t)p + 1] = make_float4(point6[3 * stride + i], point6[4 * stride + i], point6[5 * stride + i], w);
    }
}

unsigned grid_for(size_t n) { return (unsigned)std::max<size_t>(1, std::min<size_t>((n + 255) / 256, 148 * 16)); }

}  // namespace

cudaError_t launch_find_cells(const TreeNode* nodes, int n_nodes, const float* pos3, size_t stride, size_t n,
                              uint32_t* out, uint32_t miss, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    find_cells_kernel<<<grid_for(n), 256, 0, st>>>(nodes, n_nodes, pos3, stride, n, out, miss);
    return cudaGetLastError();
}

cudaError_t launch_collect_count(const TreeNode* nodes, int n_nodes, const float* key3, size_t key_stride,
                                 const float* weight, size_t n, uint32_t n_buckets, uint32_t* cell, uint32_t* count,
                                 uint32_t* cursor, cudaStream_t st) {
    cudaError_t e = cudaMemsetAsync(count, 0, (size_t)n_buckets * 4, st);
    if (e != cudaSuccess) return e;
    if (n) collect_find_kernel<<<grid_for(n), 256, 0, st>>>(nodes, n_nodes, key3, key_stride, weight, n, n_buckets, cell, count);
    collect_scan_kernel<<<1, 1024, 0, st>>>(count, n_buckets, cursor);
    return cudaGetLastError();
}

cudaError_t launch_collect_scatter(const uint32_t* cell, const float* point6, size_t stride, const float* weight,
                                   size_t n, uint32_t dropped, uint32_t* cursor, float4* rec, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    collect_scatter_kernel<<<grid_for(n), 256, 0, st>>>(cell, point6, stride, weight, n, dropped, cursor, rec);
    return cudaGetLastError();
}

}  // namespace sdmmb
