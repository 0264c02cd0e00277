// render_kernels.cu -- device-resident guided path tracer for analytic (quad) scenes: the front end of
// the progressive training loop.
//
// Reference ("R:" = /root/reference/mitsuba/src/integrators/sdmm/sdmm_proc.cpp):
//   SDMMRenderer::Li                R: 592-968   path loop, vertex records, sample weights, collection
//   SDMMRenderer::sampleSurface     R: 275-508   BSDF / guided one-sample MIS (fraction 0.5)
//   SDMMRenderer::pdfSurface        R: 510-590   0.5 bsdfPdf + 0.5 gmmPdf
//   createCondition                 R: 263-273   position normalised by the scene's largest extent
// specialised to what the analytic scenes contain: one-sided diffuse BSDFs, one-sided area emitters, no
// media, no next-event estimation (NEE is compiled out in the reference: `#ifdef NEE`, R: 696), maxDepth ==
// rrDepth as the integrator enforces (volpath_sdmm.cpp:78-79), which makes Russian roulette unreachable
// (R: 838: depth++ >= rrDepth never holds inside the loop).
//
// B200 design: WAVEFRONT, not a megakernel.  The guided query (conditional mixture of K components per
// vertex) is the expensive part and is served by the fused query kernel K4, which wants vertices batched and
// binned by cell.  So one bounce = one `pt_bounce_kernel` launch over all paths (finalise the previous
// bounce's sampling decision, trace, emission, vertex bookkeeping, decide BSDF vs guided and append a query
// record to list A (guided: sample + pdf) or list B (BSDF-sampled: pdf only)), then two K4 launches over the
// compacted lists.  Paths keep a private PCG32 stream (6 draws per bounce in a fixed order), so results do
// not depend on scheduling.
#include "kernels.h"
#include "render.h"

namespace sdmmb {

namespace {

constexpr float PT_RAY_EPS = 1e-4f;       // Mitsuba's ray epsilon
constexpr float PT_SPEC_EPS = 1e-4f;      // Mitsuba's Epsilon in Vertex::record          R: 615-619
constexpr float PT_INV_PI = 0.31830988618379067154f;
constexpr float PT_PI = 3.14159265358979323846f;

struct Rng {
    uint64_t state, inc;
    __device__ uint32_t next_uint32() {
        uint64_t oldstate = state;
        state = oldstate * 0x5851f42d4c957f2dULL + inc;
        uint32_t xorshifted = (uint32_t)(((oldstate >> 18) ^ oldstate) >> 27);
        uint32_t rot = (uint32_t)(oldstate >> 59);
        return (xorshifted >> rot) | (xorshifted << ((32u - rot) & 31u));
    }
    __device__ void seed(uint64_t initstate, uint64_t initseq) {
        state = 0;
        inc = (initseq << 1) | 1u;
        next_uint32();
        state += initstate;
        next_uint32();
    }
    __device__ float next() { return __uint_as_float((next_uint32() >> 9) | 0x3f800000u) - 1.f; }
};

struct V3 {
    float x, y, z;
};
__device__ __forceinline__ V3 v3(float x, float y, float z) { return V3{x, y, z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 ld3(const float* p) { return v3(p[0], p[1], p[2]); }

// The scene's quads and clusters as the bounce kernel stages them in shared memory (one copy per CTA):
//   quad q    : 4 float4 = (n, p0.x) (p0.yz, e1.xy) (e1.z, e2) (1/|e1|^2, 1/|e2|^2, original index, -)
//   cluster g : 2 float4 = (bmin, first) (bmax, count)
constexpr int PT_SQ_F4 = 5;   // 4 used + 1 of padding: lanes read DIFFERENT quads, a 64-byte stride would put them all in two bank groups
struct SceneSmem {
    float4 quad[PT_MAX_QUADS * PT_SQ_F4];
    float4 group[PT_MAX_QUADS * 2];
};

__device__ __forceinline__ void stage_scene(const PtScene& sc, SceneSmem& S) {
    for (int q = threadIdx.x; q < sc.n_quads; q += blockDim.x) {
        const PtQuad& Q = sc.quads[q];
        S.quad[q * PT_SQ_F4 + 0] = make_float4(Q.n[0], Q.n[1], Q.n[2], Q.p0[0]);
        S.quad[q * PT_SQ_F4 + 1] = make_float4(Q.p0[1], Q.p0[2], Q.e1[0], Q.e1[1]);
        S.quad[q * PT_SQ_F4 + 2] = make_float4(Q.e1[2], Q.e2[0], Q.e2[1], Q.e2[2]);
        S.quad[q * PT_SQ_F4 + 3] = make_float4(Q.inv_len2[0], Q.inv_len2[1], __int_as_float(sc.orig[q]), 0.f);
    }
    for (int g = threadIdx.x; g < sc.n_groups; g += blockDim.x) {
        const PtGroup& G = sc.groups[g];
        S.group[2 * g + 0] = make_float4(G.bmin[0], G.bmin[1], G.bmin[2], __int_as_float(G.first));
        S.group[2 * g + 1] = make_float4(G.bmax[0], G.bmax[1], G.bmax[2], __int_as_float(G.count));
    }
}

// nearest quad along (o, d); returns the quad's index (cluster order) or -1.  Per quad the arithmetic is the plain
// ray / parallelogram test; the cluster boxes only decide which quads are looked at (they are enlarged on the host
// by 1e-4 of the scene's extent, far above the rounding of a hit point that passed the (u, v) test), and a tie in t
// goes to the lowest original index -- the result equals a loop over all quads in the caller's order.
// Two phases, because the rays of a warp are incoherent after the first bounce: (1) every lane tests ALL cluster
// boxes and collects its own candidate quads in a 64-bit mask; (2) every lane walks its OWN candidates.  A loop
// "for each cluster: if my ray passes: for each quad" serialises the union of all lanes' clusters with a few lanes
// active in each (ncu: 3.3 of 32 lanes on the quad-test lines); here a warp does max_lane(#candidates) quad tests.
__device__ int intersect(const SceneSmem& S, int n_groups, V3 o, V3 d, float& t_hit) {
    const V3 inv = v3(__fdividef(1.f, d.x), __fdividef(1.f, d.y), __fdividef(1.f, d.z));   // culling only: 2 ulp is far inside the margin
    unsigned long long cand = 0ull;
    for (int g = 0; g < n_groups; ++g) {
        const float4 g0 = S.group[2 * g], g1 = S.group[2 * g + 1];
        // slab test.  For an axis the ray is parallel to (inv = +-inf) the two planes give -inf / +inf when the origin
        // lies between them (never culls) and the same infinity twice when it does not (culls, correctly); 0 * inf =
        // NaN (origin exactly on a plane) is dropped by fminf / fmaxf.  (The one-FMA form g * inv - o * inv is NOT
        // used: with inv = inf it turns "origin inside" into inf - inf.)
        const float ax = (g0.x - o.x) * inv.x, bx = (g1.x - o.x) * inv.x;
        const float ay = (g0.y - o.y) * inv.y, by = (g1.y - o.y) * inv.y;
        const float az = (g0.z - o.z) * inv.z, bz = (g1.z - o.z) * inv.z;
        const float t_in = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
        const float t_out = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
        if (t_in > t_out || t_out < 0.f) continue;
        const int first = __float_as_int(g0.w), count = __float_as_int(g1.w);
        cand |= ((count >= 64 ? ~0ull : ((1ull << count) - 1ull)) << first);
    }
    int best = -1, best_orig = 0x7fffffff;
    float tb = 3.0e38f;
    while (cand) {
        const int q = __ffsll((long long)cand) - 1;
        cand &= cand - 1ull;
        const float4 a = S.quad[q * PT_SQ_F4 + 0], b = S.quad[q * PT_SQ_F4 + 1];
        const V3 n = v3(a.x, a.y, a.z);
        const float dn = dot(n, d);
        if (dn == 0.f) continue;
        const V3 p0 = v3(a.w, b.x, b.y);
        const float t = dot(n, p0 - o) / dn;
        if (!(t > PT_RAY_EPS && t <= tb)) continue;
        const float4 c = S.quad[q * PT_SQ_F4 + 2], e = S.quad[q * PT_SQ_F4 + 3];
        const V3 h = o + d * t - p0;
        const float u = dot(h, v3(b.z, b.w, c.x)) * e.x;
        const float v = dot(h, v3(c.y, c.z, c.w)) * e.y;
        if (u < 0.f || u > 1.f || v < 0.f || v > 1.f) continue;
        const int orig = __float_as_int(e.z);
        if (t == tb && orig > best_orig) continue;
        tb = t;
        best = q;
        best_orig = orig;
    }
    t_hit = tb;
    return best;
}

// Mitsuba warp::squareToCosineHemisphere (concentric disk + projection)
__device__ V3 cosine_hemisphere(float u1, float u2) {
    const float r1 = 2.f * u1 - 1.f, r2 = 2.f * u2 - 1.f;
    float r, phi;
    if (r1 == 0.f && r2 == 0.f) {
        r = phi = 0.f;
    } else if (r1 * r1 > r2 * r2) {
        r = r1;
        phi = (PT_PI / 4.f) * (r2 / r1);
    } else {
        r = r2;
        phi = (PT_PI / 2.f) - (r1 / r2) * (PT_PI / 4.f);
    }
    float s, c;
    sincosf(phi, &s, &c);
    const float x = r * c, y = r * s;
    return v3(x, y, sqrtf(fmaxf(0.f, 1.f - x * x - y * y)));
}

// The first PT_TREE_TOP nodes of the device tree -- it is stored breadth-first (sdmmb_set_tree), so these are its top
// eight levels -- are staged in shared memory: a descent is a chain of DEPENDENT node loads, and the path state
// streaming through L1 keeps evicting the table.  Measured over the interior loop: pt_bounce_kernel 123 -> 89 ms with
// 256 nodes staged; 1024 nodes (ten levels) bring nothing more, so the cheap staging stays.
constexpr int PT_TREE_TOP = 256;
__device__ __forceinline__ void stage_tree_top(const TreeNode* __restrict__ nodes, int n_nodes, int4* s_top) {
    const int m = n_nodes < PT_TREE_TOP ? n_nodes : PT_TREE_TOP;
    for (int k = threadIdx.x; k < m; k += blockDim.x) s_top[k] = __ldg(reinterpret_cast<const int4*>(nodes) + k);
}
__device__ __forceinline__ int4 tree_node(const TreeNode* __restrict__ nodes, const int4* s_top, uint32_t node) {
    return node < (uint32_t)PT_TREE_TOP ? s_top[node] : __ldg(reinterpret_cast<const int4*>(nodes) + node);
}

// STree::find (R: sdmm/include/sdmm/accelerators/spatial_tree.h:19-51): the leaf's cell slot.  The leaf's box, which the
// reference rebuilds during the descent, comes from the per-leaf table F.leaf_box where it is needed.
__device__ __forceinline__ uint32_t find_leaf(const TreeNode* __restrict__ nodes, const int4* s_top, int n_nodes, V3 p, uint32_t miss) {
    uint32_t node = 0;
    for (int depth = 0; depth < 96; ++depth) {
        if (node >= (uint32_t)n_nodes) return miss;
        const int4 raw = tree_node(nodes, s_top, node);
        if (raw.x < 0) return (uint32_t)raw.z;
        const float c = raw.x == 0 ? p.x : (raw.x == 1 ? p.y : p.z);
        node = (__int_as_float(raw.y) < c) ? (uint32_t)raw.z : (uint32_t)raw.w;
    }
    return miss;
}

// A slot in one of the two query lists: the lanes of a warp that take this branch together reserve consecutive
// slots with ONE atomic (the list counter is a single word; neighbours' records also end up next to each other).
__device__ __forceinline__ uint32_t list_slot(uint32_t* counter) {
    const unsigned m = __activemask();
    const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(counter, (uint32_t)__popc(m));
    base = __shfl_sync(m, base, leader);
    return base + (uint32_t)__popc(m & ((1u << lane) - 1u));
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// one bounce of every live path
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128, 8) pt_bounce_kernel(const PtScene* __restrict__ scp, PtPaths P, PtLists L, PtFrame F,
                                                        int first, size_t nA, size_t nB) {
    __shared__ SceneSmem S;
    __shared__ int4 s_top[PT_TREE_TOP];
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    // first bounce: the path starts here (rRec.newQuery: depth = 1 at the first intersection), nothing to read
    int depth = (i < P.n) ? (first ? 1 : P.depth[i]) : 0;
    if (!__syncthreads_or(depth > 0)) return;   // a CTA of finished paths leaves at once
    const PtScene& sc = *scp;
    stage_scene(sc, S);
    if (F.guided) stage_tree_top(F.tree, F.n_tree_nodes, s_top);
    __syncthreads();
    if (depth <= 0) return;
    const int n_groups = sc.n_groups;
    const size_t n = P.n;
    Rng rng;
    V3 thr = v3(1.f, 1.f, 1.f), li = v3(0.f, 0.f, 0.f);
    // vcount word: stored vertices in the low byte, above it one bit per vertex whose outgoing ray found an emitter
    int vc = 0;
    uint32_t emit_mask = 0u;
    if (first) {
        rng.seed(F.seed, (uint64_t)(F.path_base + i));
    } else {
        rng.state = P.rng[i];
        rng.inc = P.rng[n + i];
        thr = v3(P.thr[i], P.thr[n + i], P.thr[2 * n + i]);
        li = v3(P.li[i], P.li[n + i], P.li[2 * n + i]);
        const int vword = P.vcount[i];
        vc = vword & 0xff;
        emit_mask = (uint32_t)vword >> 8;
    }
    V3 o, d;
    int hq;
    float t_hit;
    bool alive = true;

    if (!first) {
        // ---- finalise the sampling decision taken at the pending hit (sampleSurface, R: 383-507) ----
        // pending decision of the previous bounce, one word: quad (6 bits) | mode (2 bits) | list slot (24 bits)
        const uint32_t pword = P.pword[i];
        const PtQuad& Q = sc.quads[pword & 63u];
        const V3 nrm = ld3(Q.n), rho = ld3(Q.reflectance);
        const V3 p = v3(P.pend[i], P.pend[n + i], P.pend[2 * n + i]);
        const float b1 = P.pend[3 * n + i], b2 = P.pend[4 * n + i];
        const V3 npos = (p - ld3(sc.aabb_min)) * sc.inv_spatial_norm;    // createCondition, recomputed (same bits)
        const int mode = (int)((pword >> 6) & 3u);
        const uint32_t slot = pword >> 8;
        V3 wo = v3(0.f, 0.f, 0.f), weight = v3(0.f, 0.f, 0.f);
        float misPdf = 0.f;
        bool use_bsdf = (mode != 2);
        if (mode == 2) {
            if (L.validA[slot]) {                        // validConditional
                if (L.ijA[slot] == 0.f) alive = false;   // inv_jacobian == 0 -> Spectrum(0)   R: 436-439
                wo = v3(L.dirA[slot], L.dirA[nA + slot], L.dirA[2 * nA + slot]);
                const float cos_o = dot(nrm, wo);
                const float bsdfPdf = cos_o * PT_INV_PI;
                if (!(bsdfPdf > 0.f)) alive = false;     // pdfSurface: bsdfPdf <= 0 -> 0       R: 527-530
                misPdf = F.w_bsdf * bsdfPdf + (1.f - F.w_bsdf) * L.pdfA[slot];      // R: 587-589
                const float f = PT_INV_PI * cos_o / misPdf;   // bsdf->eval / pdf               R: 464, 507
                weight = rho * f;
            } else {
                use_bsdf = true;                         // !validConditional -> BSDF, weight 1 R: 388-393
            }
        }
        if (use_bsdf) {
            const V3 loc = cosine_hemisphere(b1, b2);
            V3 s = ld3(Q.e1);
            s = s * rsqrtf(dot(s, s));
            const V3 tt = v3(nrm.y * s.z - nrm.z * s.y, nrm.z * s.x - nrm.x * s.z, nrm.x * s.y - nrm.y * s.x);
            wo = s * loc.x + tt * loc.y + nrm * loc.z;
            const float bsdfPdf = loc.z * PT_INV_PI;
            if (!(bsdfPdf > 0.f)) alive = false;
            if (mode == 1 && L.validB[slot]) {
                misPdf = F.w_bsdf * bsdfPdf + (1.f - F.w_bsdf) * L.pdfB[slot];   // bsdfWeight *= pdf; / pdfSurface  R: 404, 507
                weight = rho * (bsdfPdf / misPdf);
            } else {
                misPdf = bsdfPdf;                        // pdf *= 1; bsdfWeight / 1            R: 398-402
                weight = rho;
            }
        }
        if (!(weight.x > 0.f || weight.y > 0.f || weight.z > 0.f) || !(misPdf > 0.f)) alive = false;  // R: 760
        if (alive) {
            thr = thr * weight;
            o = p;
            d = wo;
            hq = intersect(S, n_groups, o, d, t_hit);   // rayIntersectAndLookForEmitter      R: 796-798
            V3 value = v3(0.f, 0.f, 0.f);
            if (hq >= 0) {
                const PtQuad& H = sc.quads[hq];
                if (dot(ld3(H.n), d) < 0.f) value = ld3(H.emission);
            }
            // recordRadiance (R: 628-637, 810-812) adds thr * value / (thr_v * pdf_v) to the weight of EVERY earlier vertex
            // v.  That read-modify-write walk over the path's records does not happen here: the record of this vertex
            // keeps `value` and its throughput, and pt_finish_kernel replays the same sums in the same order.
            if (value.x > 0.f || value.y > 0.f || value.z > 0.f) {
                li = li + thr * value;
                if (vc < PT_MAX_VERTS) emit_mask |= 1u << vc;
            }
            if (vc < PT_MAX_VERTS) {                     // the cacheable vertex               R: 814-840
                float* vp = P.verts + ((size_t)vc * PT_VERT_FLOATS) * n + i;
                const float clamped = fmaxf(misPdf, 0.1f);
                vp[0] = npos.x; vp[n] = npos.y; vp[2 * n] = npos.z;
                vp[3 * n] = wo.x; vp[4 * n] = wo.y; vp[5 * n] = wo.z;
                vp[6 * n] = nrm.x; vp[7 * n] = nrm.y; vp[8 * n] = nrm.z;
                vp[9 * n] = value.x; vp[10 * n] = value.y; vp[11 * n] = value.z;   // emitted radiance found from here
                vp[12 * n] = thr.x; vp[13 * n] = thr.y; vp[14 * n] = thr.z;
                vp[15 * n] = clamped;
            }
            ++vc;
            ++depth;                                     // rRec.depth++                       R: 852
            if (hq < 0) alive = false;
        }
    } else {
        // camera ray (process(): samplePos = pixel + next2D, sensor ray; R: 146-176)
        const size_t npix = (size_t)sc.width * sc.height;
        const size_t pix = (F.path_base + i) % npix;
        const int px = (int)(pix % sc.width), py = (int)(pix / sc.width);
        const float sx = (float)px + rng.next(), sy = (float)py + rng.next();
        const float cx = (1.f - 2.f * sx / (float)sc.width) * sc.tan_x;
        const float cy = (1.f - 2.f * sy / (float)sc.height) * sc.tan_y;
        o = ld3(sc.cam_origin);
        d = ld3(sc.cam_forward) + ld3(sc.cam_right) * cx + ld3(sc.cam_up) * cy;
        d = d * rsqrtf(dot(d, d));
        hq = intersect(S, n_groups, o, d, t_hit);
        if (hq < 0) alive = false;
        else {
            const PtQuad& H = sc.quads[hq];              // emitted radiance seen directly   R: 672-675
            if (dot(ld3(H.n), d) < 0.f) li = li + thr * ld3(H.emission);
        }
    }

    if (alive && depth >= sc.max_depth) alive = false;   // R: 682-683
    if (alive) {
        const PtQuad& H = sc.quads[hq];
        const V3 hn = ld3(H.n);
        // one-sided diffuse: sample()/eval() return 0 from behind or for a black surface
        if (!(dot(hn, d) < 0.f) || !(H.reflectance[0] > 0.f || H.reflectance[1] > 0.f || H.reflectance[2] > 0.f))
            alive = false;
    }
    if (alive) {
        const V3 p = o + d * t_hit;
        const float u_mis = rng.next(), b1 = rng.next(), b2 = rng.next();
        const float g0 = rng.next(), g1 = rng.next(), g2 = rng.next();
        const V3 npos = (p - ld3(sc.aabb_min)) * sc.inv_spatial_norm;    // createCondition  R: 263-273
        int mode = 0;
        uint32_t slot = 0;
        if (F.guided) {
            const uint32_t cell = find_leaf(F.tree, s_top, F.n_tree_nodes, npos, 0xffffffffu);
            if (cell < F.max_cells && F.cell_initialized[cell]) {
                if (u_mis <= F.w_bsdf) {                 // heuristicConditionalWeight       R: 383-392
                    mode = 1;
                    const PtQuad& H = sc.quads[hq];
                    const V3 hn = ld3(H.n);
                    const V3 loc = cosine_hemisphere(b1, b2);
                    V3 s = ld3(H.e1);
                    s = s * rsqrtf(dot(s, s));
                    const V3 tt = v3(hn.y * s.z - hn.z * s.y, hn.z * s.x - hn.x * s.z, hn.x * s.y - hn.y * s.x);
                    const V3 wo = s * loc.x + tt * loc.y + hn * loc.z;
                    slot = list_slot(&L.counters[1]);
                    float4* r = reinterpret_cast<float4*>(L.recB + (size_t)slot * 8);
                    r[0] = make_float4(__uint_as_float(cell), npos.x, npos.y, npos.z);
                    r[1] = make_float4(wo.x, wo.y, wo.z, __int_as_float(hq));
                } else {
                    mode = 2;
                    slot = list_slot(&L.counters[0]);
                    float4* r = reinterpret_cast<float4*>(L.recA + (size_t)slot * 8);
                    r[0] = make_float4(__uint_as_float(cell), npos.x, npos.y, npos.z);
                    r[1] = make_float4(g0, g1, g2, __int_as_float(hq));
                }
            }
        }
        P.pword[i] = (uint32_t)hq | ((uint32_t)mode << 6) | (slot << 8);
        P.pend[i] = p.x; P.pend[n + i] = p.y; P.pend[2 * n + i] = p.z;
        P.pend[3 * n + i] = b1; P.pend[4 * n + i] = b2;
    }
    P.rng[i] = rng.state;
    if (first) P.rng[n + i] = rng.inc;
    P.thr[i] = thr.x; P.thr[n + i] = thr.y; P.thr[2 * n + i] = thr.z;
    P.li[i] = li.x; P.li[n + i] = li.y; P.li[2 * n + i] = li.z;
    P.vcount[i] = vc | (int)(emit_mask << 8);
    P.depth[i] = alive ? depth : 0;
}

// AoS staging records -> the dense SoA planes (stride = list length) the query kernels take
__global__ void pt_pack_lists_kernel(const PtScene* __restrict__ scp, PtLists L, size_t nA, size_t nB) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nA) {
        const float4* r = reinterpret_cast<const float4*>(L.recA + i * 8);
        const float4 a = r[0], b = r[1];
        L.cellA[i] = __float_as_uint(a.x);
        L.posA[i] = a.y; L.posA[nA + i] = a.z; L.posA[2 * nA + i] = a.w;
        L.uA[i] = b.x; L.uA[nA + i] = b.y; L.uA[2 * nA + i] = b.z;
        const float* nq = scp->quads[__float_as_int(b.w)].n;      // shading normal (product guiding)
        L.nrmA[i] = nq[0]; L.nrmA[nA + i] = nq[1]; L.nrmA[2 * nA + i] = nq[2];
    }
    if (i < nB) {
        const float4* r = reinterpret_cast<const float4*>(L.recB + i * 8);
        const float4 a = r[0], b = r[1];
        L.cellB[i] = __float_as_uint(a.x);
        L.posB[i] = a.y; L.posB[nB + i] = a.z; L.posB[2 * nB + i] = a.w;
        L.dirB[i] = b.x; L.dirB[nB + i] = b.y; L.dirB[2 * nB + i] = b.z;
        const float* nq = scp->quads[__float_as_int(b.w)].n;
        L.nrmB[i] = nq[0]; L.nrmB[nB + i] = nq[1]; L.nrmB[2 * nB + i] = nq[2];
    }
}

// The pushes of one vertex step -- the main sample and up to two jittered copies per lane -- executed by the WHOLE
// warp: the lanes reserve consecutive pool slots with one atomic per warp (the pool counter is a single word: per-lane
// atomics on it serialise across the GPU), the vertex record is read once, and the writes of neighbouring lanes are
// adjacent.  "Data full." (R: sdmm/include/sdmm/opt/data.h:60-66) is counted per sample.
__device__ __forceinline__ void pool_push_warp(const bool f[3], const V3 key[3], const uint32_t seq[3], const PtLists& L,
                                               const PtFrame& F, const float* vp, size_t n, float w) {
    const int cnt = (f[0] ? 1 : 0) + (f[1] ? 1 : 0) + (f[2] ? 1 : 0);
    if (!__any_sync(0xffffffffu, cnt > 0)) return;
    const int lane = threadIdx.x & 31;
    int incl = cnt;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, incl, off);
        if (lane >= off) incl += up;
    }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    uint32_t base = 0;
    if (lane == 0) base = atomicAdd(&L.counters[2], (uint32_t)total);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (cnt == 0) return;
    uint32_t idx = base + (uint32_t)(incl - cnt);
    float v9[9];
#pragma unroll
    for (int c = 0; c < 9; ++c) v9[c] = vp[c * n];   // point6 | normal3
    const size_t S = F.pool_cap;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        if (!f[r]) continue;
        if (idx >= F.pool_cap) {
            atomicAdd(&L.counters[3], 1u);
        } else {
            float* o = F.pool + idx;
#pragma unroll
            for (int c = 0; c < 9; ++c) o[c * S] = v9[c];
            o[9 * S] = w;
            o[10 * S] = key[r].x; o[11 * S] = key[r].y; o[12 * S] = key[r].z;
            o[13 * S] = r == 0 ? 1.f : 0.f;   // only the main push feeds the cell's SpatialStats
            o[14 * S] = __uint_as_float(seq[r]);   // order a single-threaded reference would push in: path, vertex, jitter
        }
        ++idx;
    }
}

// end of every path: film accumulation + the sample collection of Li's tail (R: 862-965).  The vertex loop is
// warp-uniform (a lane without a vertex at this step idles) so that every push site is reached by all 32 lanes.
__global__ void __launch_bounds__(256, 4) pt_finish_kernel(const PtScene* __restrict__ scp, PtPaths P, PtLists L, PtFrame F) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool in = i < P.n;
    const PtScene& sc = *scp;
    const size_t n = P.n;
    const size_t npix = (size_t)sc.width * sc.height;
    int depth = 0;
    uint32_t emit_mask = 0u;   // vertices whose outgoing ray found an emitter: usually none, or only the path's last one
    Rng rng{0, 1};
    if (in) {
        const size_t pix = (F.path_base + i) % npix;
        for (int c = 0; c < 3; ++c) atomicAdd(&F.image[c * npix + pix], P.li[c * n + i]);
        if (F.collect) {
            const int vword = P.vcount[i];   // vertex count | emitter bits (pt_bounce_kernel)
            depth = min(vword & 0xff, PT_MAX_VERTS);
            emit_mask = (uint32_t)vword >> 8;
            // a path that never found an emitter carries only zero weights: nothing of it is ever pushed, and the jitter
            // draws it would make feed nothing else (the path's random stream ends here)
            if (emit_mask == 0u) depth = 0;
            rng.state = P.rng[i];
            rng.inc = P.rng[n + i];
        }
    }
    if (!F.collect) return;
    __shared__ int4 s_top[PT_TREE_TOP];
    stage_tree_top(F.tree, F.n_tree_nodes, s_top);
    __syncthreads();
    const int first_saved = max(depth - sc.saved_per_path, 0);
    for (int dv = depth - 1;; --dv) {
        const bool act = dv >= first_saved && dv >= 0;
        if (!__any_sync(0xffffffffu, act)) break;
        const float* vp = P.verts + ((size_t)(act ? dv : 0) * PT_VERT_FLOATS) * n + (in ? i : 0);
        float w = 0.f;
        bool valid = false, cell_ok = false;
        V3 pos = v3(0.f, 0.f, 0.f);
        float smin[3] = {0.f, 0.f, 0.f}, smax[3] = {0.f, 0.f, 0.f};
        int n_jitters = 0;
        if (act) {
            // the vertex's sample weight: its own emitted term value / pdf (R: 836-838) plus, for every LATER vertex k of
            // the path that found an emitter, thr_k value_k / (thr_v pdf_v) per channel with thr_v above Mitsuba's Epsilon
            // (recordRadiance, R: 615-637) -- summed in path order, as the per-bounce updates were
            // No emitter at or after this vertex: every term is exactly 0, the sample is invalid (is_valid_sample).
            if ((emit_mask >> dv) != 0u) {
                const float vpdf = vp[15 * n];
                const float inv = 1.f / vpdf;
                float w3[3] = {vp[9 * n] * inv, vp[10 * n] * inv, vp[11 * n] * inv};
                if ((emit_mask >> (dv + 1)) != 0u) {
                    const float vthr[3] = {vp[12 * n], vp[13 * n], vp[14 * n]};
                    for (int k = dv + 1; k < depth; ++k) {
                        if (!((emit_mask >> k) & 1u)) continue;
                        const float* kp = P.verts + ((size_t)k * PT_VERT_FLOATS) * n + i;
#pragma unroll
                        for (int c = 0; c < 3; ++c) {
                            const float rad = kp[(12 + c) * n] * kp[(9 + c) * n];
                            if (vthr[c] > PT_SPEC_EPS) w3[c] += rad / (vthr[c] * vpdf);
                        }
                    }
                }
                w = (w3[0] + w3[1] + w3[2]) * (1.f / 3.f);                  // weight.average()
                valid = (w >= 1e-8f) && (w <= 3.402823466e+38f);           // is_valid_sample
            }
            n_jitters = (dv >= depth - 1) ? 1 : 0;
            if (w > 1000.f) ++n_jitters;
            // the cell (and its box) is only needed by a push or by a jitter draw
            if (valid || n_jitters > 0) {
                pos = v3(vp[0], vp[n], vp[2 * n]);
                const uint32_t cell = find_leaf(F.tree, s_top, F.n_tree_nodes, pos, 0xffffffffu);
                cell_ok = cell < F.max_cells;   // the reference throws otherwise; positions are inside the root AABB
                if (cell_ok) {                  // the leaf's box from the per-leaf table (what the descent would rebuild)
                    const float4 lo = __ldg(F.leaf_box + cell), hi = __ldg(F.leaf_box + F.max_cells + cell);
                    smin[0] = lo.x; smin[1] = lo.y; smin[2] = lo.z;
                    smax[0] = hi.x; smax[1] = hi.y; smax[2] = hi.z;
                }
            }
        }
        const uint32_t seq0 = ((uint32_t)(F.path_base + i) << 5) | ((uint32_t)(depth - 1 - dv) << 2);
        bool pf[3] = {act && cell_ok && valid, false, false};
        V3 pkey[3] = {pos, pos, pos};
        uint32_t pseq[3] = {seq0, seq0, seq0};
        uint32_t pushed = 0;
        int attempts = 0;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            bool want = act && cell_ok && j < n_jitters, got = false;
            V3 key = pos;
            while (want && !got) {
                const float u0 = rng.next(), u1 = rng.next(), u2 = rng.next();
                key = v3(pos.x + (u0 - 0.5f) * (smax[0] - smin[0]), pos.y + (u1 - 0.5f) * (smax[1] - smin[1]),
                         pos.z + (u2 - 0.5f) * (smax[2] - smin[2]));
                const uint32_t jc = find_leaf(F.tree, s_top, F.n_tree_nodes, key, 0xffffffffu);
                float4 jlo = make_float4(smin[0], smin[1], smin[2], 0.f);
                if (jc < F.max_cells) jlo = __ldg(F.leaf_box + jc);
                // find() never fails for a point outside the root box (it lands in a boundary leaf), as in the reference
                if (jc >= F.max_cells || (jlo.x == smin[0] && jlo.y == smin[1] && jlo.z == smin[2])) {
                    if (++attempts >= 8) want = false;   // this jitter is given up; a later one fails at its first miss
                } else {
                    got = true;
                }
            }
            const bool push = got && valid;
            if (push) ++pushed;
            pf[1 + j] = push;
            pkey[1 + j] = key;
            pseq[1 + j] = seq0 + pushed;
        }
        pool_push_warp(pf, pkey, pseq, L, F, vp, n, w);
    }
}

// ------------------------------------------------------------------------------------------------
// optimize(), first pass over the pool: cell of every pooled sample (STree::find of its key), per-cell sample
// counts (Data::size of canBeOptimized, volpath_sdmm.cpp:140-149) and the SpatialStats of the leaves -- count,
// sum p, sum p^2 of the MAIN pushes of this pass only (R: 893-895) -- in double.  A warp whose lanes all fall
// into one cell (neighbouring paths: the common case) reduces first and issues one atomic per quantity.
// stats: [8][max_cells] doubles = n_main, sum xyz, sum sq xyz, n_all.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) pt_pool_stats_kernel(const TreeNode* __restrict__ nodes, int n_nodes,
                                                            const float* __restrict__ pool, size_t cap, size_t n,
                                                            size_t new_begin, uint32_t max_cells, int32_t* __restrict__ out_cell,
                                                            double* __restrict__ stats) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool in = i < n;
    uint32_t cell = 0xffffffffu;
    bool is_main = false;
    double px = 0.0, py = 0.0, pz = 0.0;
    if (in) {
        const V3 key = v3(pool[10 * cap + i], pool[11 * cap + i], pool[12 * cap + i]);
        uint32_t node = 0;
        for (int depth = 0; depth < 96 && node < (uint32_t)n_nodes; ++depth) {
            const int4 raw = __ldg(reinterpret_cast<const int4*>(nodes) + node);
            if (raw.x < 0) { cell = (uint32_t)raw.z; break; }
            const float c = raw.x == 0 ? key.x : (raw.x == 1 ? key.y : key.z);
            node = (__int_as_float(raw.y) < c) ? (uint32_t)raw.z : (uint32_t)raw.w;
        }
        if (cell >= max_cells) cell = 0xffffffffu;
        out_cell[i] = (int32_t)cell;
        is_main = cell != 0xffffffffu && i >= new_begin && (((int)pool[13 * cap + i]) & 1);   // plane 13 = is_main + 2 * pass (integrator.py)
        if (is_main) { px = pool[i]; py = pool[cap + i]; pz = pool[2 * cap + i]; }
    }
    const unsigned full = 0xffffffffu;
    const bool counted = in && cell != 0xffffffffu;
    const unsigned mc = __ballot_sync(full, counted);
    if (mc == 0u) return;
    const int lane = threadIdx.x & 31;
    const int leader = __ffs(mc) - 1;
    const uint32_t cell0 = __shfl_sync(full, cell, leader);
    const bool uniform = __all_sync(full, !counted || cell == cell0);
    const size_t C = max_cells;
    if (uniform) {
        const unsigned mm = __ballot_sync(full, is_main);
        double v[6] = {px, py, pz, px * px, py * py, pz * pz};
        if (mm) {
#pragma unroll
            for (int k = 0; k < 6; ++k)
#pragma unroll
                for (int off = 16; off > 0; off >>= 1) v[k] += __shfl_xor_sync(full, v[k], off);
        }
        if (lane == leader) {
            atomicAdd(&stats[7 * C + cell0], (double)__popc(mc));
            if (mm) {
                atomicAdd(&stats[0 * C + cell0], (double)__popc(mm));
#pragma unroll
                for (int k = 0; k < 6; ++k) atomicAdd(&stats[(1 + k) * C + cell0], v[k]);
            }
        }
    } else if (counted) {
        atomicAdd(&stats[7 * C + cell], 1.0);
        if (is_main) {
            atomicAdd(&stats[0 * C + cell], 1.0);
            atomicAdd(&stats[1 * C + cell], px);
            atomicAdd(&stats[2 * C + cell], py);
            atomicAdd(&stats[3 * C + cell], pz);
            atomicAdd(&stats[4 * C + cell], px * px);
            atomicAdd(&stats[5 * C + cell], py * py);
            atomicAdd(&stats[6 * C + cell], pz * pz);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// optimize(), one recursion level of the data-driven splits (split_leaf_recurse -> create_child,
// R: sdmm/include/sdmm/accelerators/spatial_tree.h:169-214) for ALL leaves that split at this level: a sample of a
// splitting leaf moves to child 0 (position above the child's min along the split axis) or child 1 (at or below
// the child's max), or to -1 when rounding left it between the two boxes; the children's SpatialStats count the
// samples inside their boxes (AABB::contains, :204-206).  tab_i [3][C]: axis (-1: the slot does not split), slot of
// child 0, slot of child 1, per PARENT slot; tab_f [2][C]: child 0's min, child 1's max along the axis;
// box [2][3][C]: the children's boxes per CHILD slot.  stats [7][C] doubles: n, sum p, sum p^2.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) pt_rebucket_kernel(const float* __restrict__ pool, size_t cap, size_t n,
                                                          int32_t* __restrict__ cell_of, const int32_t* __restrict__ tab_i,
                                                          const float* __restrict__ tab_f, const float* __restrict__ box,
                                                          uint32_t max_cells, double* __restrict__ stats) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t c = cell_of[i];
    const size_t C = max_cells;
    if (c < 0 || (uint32_t)c >= max_cells) return;
    const int axis = tab_i[c];
    if (axis < 0) return;
    const float p[3] = {pool[i], pool[cap + i], pool[2 * cap + i]};
    const float pa = axis == 0 ? p[0] : (axis == 1 ? p[1] : p[2]);
    int32_t nc = -1;
    if (pa > tab_f[c]) nc = tab_i[C + c];
    else if (pa <= tab_f[C + c]) nc = tab_i[2 * C + c];
    cell_of[i] = nc;
    if (nc < 0) return;
    bool inside = true;
#pragma unroll
    for (int k = 0; k < 3; ++k) inside = inside && p[k] >= box[k * C + nc] && p[k] <= box[(3 + k) * C + nc];
    if (!inside) return;
    atomicAdd(&stats[nc], 1.0);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double v = (double)p[k];
        atomicAdd(&stats[(1 + k) * C + nc], v);
        atomicAdd(&stats[(4 + k) * C + nc], v * v);
    }
}

namespace {
unsigned blocks_for(size_t n, int t) { return (unsigned)((n + t - 1) / t); }
}  // namespace

cudaError_t pt_launch_bounce(const PtScene* d_scene, const PtPaths& P, const PtLists& L, const PtFrame& F, int first,
                             size_t nA, size_t nB, cudaStream_t st) {
    pt_bounce_kernel<<<blocks_for(P.n, 128), 128, 0, st>>>(d_scene, P, L, F, first, nA, nB);
    return cudaGetLastError();
}

cudaError_t pt_launch_pack_lists(const PtScene* d_scene, const PtLists& L, size_t nA, size_t nB, cudaStream_t st) {
    const size_t m = nA > nB ? nA : nB;
    if (m == 0) return cudaSuccess;
    pt_pack_lists_kernel<<<blocks_for(m, 256), 256, 0, st>>>(d_scene, L, nA, nB);
    return cudaGetLastError();
}

cudaError_t pt_launch_finish(const PtScene* d_scene, const PtPaths& P, const PtLists& L, const PtFrame& F, cudaStream_t st) {
    pt_finish_kernel<<<blocks_for(P.n, 256), 256, 0, st>>>(d_scene, P, L, F);
    return cudaGetLastError();
}

cudaError_t pt_launch_pool_stats(const TreeNode* nodes, int n_nodes, const float* pool, size_t cap, size_t n, size_t new_begin,
                                 uint32_t max_cells, int32_t* out_cell, double* stats, cudaStream_t st) {
    cudaError_t e = cudaMemsetAsync(stats, 0, (size_t)8 * max_cells * sizeof(double), st);
    if (e != cudaSuccess || n == 0) return e;
    pt_pool_stats_kernel<<<blocks_for(n, 256), 256, 0, st>>>(nodes, n_nodes, pool, cap, n, new_begin, max_cells, out_cell, stats);
    return cudaGetLastError();
}

cudaError_t pt_launch_rebucket(const float* pool, size_t cap, size_t n, int32_t* cell_of, const int32_t* tab_i, const float* tab_f,
                               const float* box, uint32_t max_cells, double* stats, cudaStream_t st) {
    cudaError_t e = cudaMemsetAsync(stats, 0, (size_t)7 * max_cells * sizeof(double), st);
    if (e != cudaSuccess || n == 0) return e;
    pt_rebucket_kernel<<<blocks_for(n, 256), 256, 0, st>>>(pool, cap, n, cell_of, tab_i, tab_f, box, max_cells, stats);
    return cudaGetLastError();
}

}  // namespace sdmmb
