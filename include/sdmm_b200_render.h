/* sdmm_b200_render.h -- C ABI of the device-resident guided path tracer for analytic scenes, the front
 * end of the progressive training loop.  It replaces, for scenes made of
 * diffuse quads and quad area lights, the render pass of one iteration of the reference integrator:
 *   SDMMProcess::process / SDMMRenderer::Li   mitsuba/src/integrators/sdmm/sdmm_proc.cpp:106-195, 592-968
 *   sampleSurface / pdfSurface                mitsuba/src/integrators/sdmm/sdmm_proc.cpp:275-590
 * The guided queries inside it go through the same fused query kernel as sdmmb_sample_pdf / sdmmb_pdf
 * (include/sdmm_b200.h).  The training half of an iteration (optimize(), volpath_sdmm.cpp:244-312) is the
 * existing sdmmb_initialize / sdmmb_set_samples / sdmmb_em_step; the loop around both lives in
 * sdmm_mitsuba_b200/integrator.py.  Mitsuba's scene loading, film and samplers are out of scope: the caller
 * passes quads and a pinhole camera.  All `d_` pointers are device pointers; return codes as sdmm_b200.h. */
#ifndef SDMM_B200_RENDER_H
#define SDMM_B200_RENDER_H
#include "sdmm_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct sdmmb_pt sdmmb_pt;

/* parallelogram p0 + u e1 + v e2, (u, v) in [0,1]^2, e1 orthogonal to e2; front side = e1 x e2.
 * One-sided diffuse reflectance (Mitsuba `diffuse`), one-sided emitted radiance (Mitsuba `area`). */
typedef struct {
    float p0[3], e1[3], e2[3];
    float reflectance[3];
    float emission[3];
} sdmmb_quad;

/* pinhole camera: direction through film position (sx, sy) in pixels =
 * forward + (1 - 2 sx / width) tan(fov_x / 2) right + (1 - 2 sy / height) tan(fov_x / 2) (height / width) up */
typedef struct {
    float origin[3], right[3], up[3], forward[3];
    float fov_x_degrees;
    int width, height;
} sdmmb_camera;

typedef struct {
    float aabb_min[3];      /* scene AABB minimum and largest extent: createCondition, sdmm_proc.cpp:263-273 */
    float spatial_norm;
    float root_min[3];      /* STree root box in normalised coordinates (getAABB, volpath_sdmm.cpp:314-330) */
    float root_max[3];
    int max_depth;          /* integrator maxDepth (== rrDepth, volpath_sdmm.cpp:78-79); <= 9 */
    int saved_per_path;     /* savedSamplesPerPath, default 8 (volpath_sdmm.cpp:60) */
} sdmmb_pt_config;

typedef struct {
    uint64_t paths;
    uint64_t guided_queries;   /* sample + pdf queries (mixture-sampled vertices) */
    uint64_t mis_queries;      /* pdf-only queries (BSDF-sampled vertices that need the mixture pdf) */
    uint64_t pool_count;       /* samples in the pool after this pass */
    uint64_t pool_overflow;    /* samples dropped because the pool was full ("Data full.", opt/data.h:60-66) */
    uint64_t launches;
} sdmmb_pt_stats;

int sdmmb_pt_create(sdmmb_ctx* ctx, sdmmb_pt** out, const sdmmb_quad* h_quads, int n_quads,
                    const sdmmb_camera* h_camera, const sdmmb_pt_config* h_config);
int sdmmb_pt_destroy(sdmmb_pt* pt);

/* sampleProduct (mitsuba/src/integrators/sdmm/sdmm_proc.cpp:326-392, 409-417, 476-487): guided vertices sample the
 * PRODUCT of the cell's conditional mixture and a learned-BSDF directional mixture (the BSDF::DMM of
 * mitsuba/include/mitsuba/render/bsdf.h:301-304) instead of the conditional alone, and the BSDF sampling fraction
 * of the one-sample MIS drops from 0.5 to 0.3 (heuristicConditionalWeight, :383-390).  h_bsdf: n_bsdf rows of
 * (pmf, mean3, c00, c10, c11), one mixture for the whole scene; all surfaces of this tracer are diffuse, so
 * component 0 is re-centred on the shading normal at every vertex (:335-339).  n_bsdf = 0 switches back to the
 * plain conditional.  A vertex whose product has no valid component is treated like one without a valid
 * conditional (BSDF sampling, weight 1); the reference falls back to the plain conditional there (:375-381). */
int sdmmb_pt_set_product(sdmmb_pt* pt, int n_bsdf, const float* h_bsdf);

/* One render pass = `spp` paths per pixel (one iteration's samplesPerIteration).  Uses the tree bound with
 * sdmmb_set_tree and the query records of ctx.  guided = 0 reproduces iteration 0 (BSDF sampling only,
 * sdmm_proc.cpp:316-323); d_cell_initialized [max_cells] marks cells whose conditioner may be used
 * (SDMMContext::initialized).  d_image_sum [3][height*width] accumulates radiance sums (divide by the total
 * spp).  If collect != 0, training samples are appended to d_pool [15][pool_capacity] planes
 * (x y z dx dy dz | nx ny nz | weight | key x y z | is_main | seq) starting at index pool_count:
 * key = the position whose leaf cell receives the sample (the vertex itself, or its jittered copy,
 * sdmm_proc.cpp:932-964); is_main = 1 for the pushes that also feed the cell's SpatialStats; seq (uint32 bits) = path * 32 +
 * vertex * 4 + jitter, the order in which a single-threaded reference would have pushed (the pool itself is
 * filled in scheduling order; sdmm::initialize depends on the order of a cell's samples).  Needs
 * width * height * spp < 2^27.  Synchronous. */
int sdmmb_pt_render(sdmmb_pt* pt, int spp, uint64_t seed, int guided, int collect,
                    const uint8_t* d_cell_initialized, float* d_image_sum, float* d_pool,
                    size_t pool_capacity, size_t pool_count, sdmmb_pt_stats* out, void* stream);

/* First pass of optimize() over the pool that sdmmb_pt_render filled (volpath_sdmm.cpp:244-259, 140-149): for the
 * samples [0, n) of d_pool [15][pool_capacity]: d_out_cell[i] = device slot of STree::find(key_i) (-1: none), and
 * d_out_stats [8][max_cells] doubles (zeroed by the call) = per cell: number, sum p, sum p^2 of the MAIN pushes with
 * index >= new_begin (the SpatialStats the renderer's push feeds, sdmm_proc.cpp:893-895; planes 0, 1-3, 4-6) and
 * the number of all pooled samples (Data::size of canBeOptimized; plane 7).  One kernel, asynchronous on `stream`.
 * Plane 13 is read as an integer whose bit 0 is is_main: a caller may add 2 * (pass number) to the values
 * sdmmb_pt_render wrote (integrator.py does, to keep older passes in front when it orders a cell's samples). */
int sdmmb_pt_pool_stats(sdmmb_ctx* ctx, const float* d_pool, size_t pool_capacity, size_t n, size_t new_begin,
                        int32_t* d_out_cell, double* d_out_stats, void* stream);

/* One recursion level of optimize()'s data-driven splits for all leaves that split at that level
 * (split_leaf_recurse -> create_child, sdmm/include/sdmm/accelerators/spatial_tree.h:169-214, 253-292): d_cell [n]
 * (in/out, the slots sdmmb_pt_pool_stats wrote) moves every sample of a splitting leaf to the child that takes it
 * (-1: rounding left it between the two boxes) and d_out_stats [7][max_cells] doubles (zeroed by the call) receives
 * the children's SpatialStats (n, sum p, sum p^2 over the samples inside the child's box, :204-206).
 * d_tab_i [3][max_cells] int32 per PARENT slot: split axis (-1 = does not split), child 0's slot, child 1's slot;
 * d_tab_f [2][max_cells]: child 0's min and child 1's max along the axis; d_box [2][3][max_cells]: min / max corner
 * of the box per CHILD slot.  The host decides which leaves split and where (HostTree, the reference's rule). */
int sdmmb_pt_rebucket(sdmmb_ctx* ctx, const float* d_pool, size_t pool_capacity, size_t n, int32_t* d_cell,
                      const int32_t* d_tab_i, const float* d_tab_f, const float* d_box, double* d_out_stats, void* stream);

#ifdef __cplusplus
}
#endif
#endif
