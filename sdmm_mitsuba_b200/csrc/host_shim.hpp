// host_shim.hpp -- the reference-facing side of the drop-in boundary: header-only C++17 templates with
// the REFERENCE's own signatures for the hot path, implemented on top of the C ABI
// (include/sdmm_b200.h -> libsdmm_b200.so).  The Mitsuba `sdmm` plugin includes this header next to
// the reference's sdmm headers and calls sdmm_b200::em_step / prepare / optimize where it calls
// sdmm::em_step / sdmm::prepare today (see INTEGRATION.md for the exact call-site diff).
//
// The templates only use the public members of the reference types (they are generic over
// JointSDMM / EM / Data / Conditioner as instantiated in
// mitsuba/src/integrators/sdmm/sdmm_proc.h:68-96), so this file contains no reference code and
// compiles against the unmodified headers under /root/reference/sdmm/include.
//
//   reference call                                              replacement
//   sdmm::em_step(sdmm, em, data)        opt/em.h:218-253       sdmm_b200::em_step(dev, slot, sdmm, em, data)
//   sdmm::prepare(conditioner, sdmm)     sdmm_conditioner.h:89  sdmm_b200::prepare(dev, slot, conditioner, sdmm)
//   the parallelFor body                 volpath_sdmm.cpp:287   sdmm_b200::optimize(dev, contexts, n_iterations)
//
// There is no CPU fallback: every function needs the CUDA library and a device; errors surface as
// std::runtime_error carrying the C ABI's error string (the reference throws std::runtime_error for
// "Data full." too, opt/data.h:80).
#pragma once
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "sdmm_b200.h"

namespace sdmm_b200 {

inline void check(int rc, const char* what, sdmmb_ctx* ctx = nullptr) {
    if (rc == SDMMB_OK) return;
    std::string msg = std::string(what) + ": " + sdmmb_get_error_string(rc);
    if (ctx && (rc == SDMMB_E_CUDA || rc == SDMMB_E_NOMEM)) msg += std::string(" (") + sdmmb_last_cuda_error(ctx) + ")";
    throw std::runtime_error(msg);
}

// Owns the device context for up to `max_cells` leaf cells of K components: the device-side
// counterpart of the STree's SDMMContext bundles (distributions/sdmm_context.h:11-45).
class Device {
public:
    Device(int device, int max_cells, int K) : K_(K), max_cells_(max_cells) {
        check(sdmmb_create(&ctx_, device, max_cells, K), "sdmmb_create");
    }
    ~Device() { sdmmb_destroy(ctx_); }
    Device(const Device&) = delete;
    Device& operator=(const Device&) = delete;
    sdmmb_ctx* ctx() const { return ctx_; }
    int K() const { return K_; }
    int max_cells() const { return max_cells_; }

private:
    sdmmb_ctx* ctx_ = nullptr;
    int K_, max_cells_;
};

inline uint32_t f2w(float f) { uint32_t w; std::memcpy(&w, &f, 4); return w; }
inline float w2f(uint32_t w) { float f; std::memcpy(&f, &w, 4); return f; }

// ---- marshalling between the reference's enoki SoA members and the flat layouts of sdmm_b200.h ----
template <typename SDMM>
void pack_joint(const SDMM& d, int K, float* joint /* [32][K] */) {
    for (int k = 0; k < K; ++k) joint[k] = d.weight.pmf.coeff(k);
    for (int i = 0; i < 6; ++i)
        for (int k = 0; k < K; ++k) joint[(1 + i) * K + k] = d.tangent_space.mean.coeff(i).coeff(k);
    for (int r = 0; r < 5; ++r)
        for (int c = 0; c < 5; ++c)
            for (int k = 0; k < K; ++k) joint[(7 + 5 * r + c) * K + k] = d.cov(r, c).coeff(k);
}

// Writes pmf / mean / cov and refreshes every derived member with the reference's OWN set_mean +
// prepare_vectorized (frame, cov_sqrt, inv_cov_sqrt_det, cdf; O(K) per cell) so host-side readers of
// the object (serialisation sdmm.h:138-161, STree split inheritance) stay consistent.
template <typename SDMM>
bool unpack_joint(SDMM& d, int K, const float* joint) {
    for (int k = 0; k < K; ++k) d.weight.pmf.coeff(k) = joint[k];
    auto mean = d.tangent_space.mean;
    for (int i = 0; i < 6; ++i)
        for (int k = 0; k < K; ++k) mean.coeff(i).coeff(k) = joint[(1 + i) * K + k];
    d.tangent_space.set_mean(mean);
    for (int r = 0; r < 5; ++r)
        for (int c = 0; c < 5; ++c)
            for (int k = 0; k < K; ++k) d.cov(r, c).coeff(k) = joint[(7 + 5 * r + c) * K + k];
    return sdmm::prepare_vectorized(d);
}

template <typename EM>
void pack_em(const EM& em, int K, float* e /* [81][K] */, uint32_t* scal /* [8] */) {
    for (int k = 0; k < K; ++k) e[k] = em.stats.weight.coeff(k);
    for (int i = 0; i < 5; ++i)
        for (int k = 0; k < K; ++k) e[(1 + i) * K + k] = em.stats.mean.coeff(i).coeff(k);
    for (int r = 0; r < 5; ++r)
        for (int c = 0; c < 5; ++c)
            for (int k = 0; k < K; ++k) {
                e[(6 + 5 * r + c) * K + k] = em.stats.cov(r, c).coeff(k);
                e[(31 + 5 * r + c) * K + k] = em.cov_prior(r, c).coeff(k);
                e[(56 + 5 * r + c) * K + k] = em.depth_prior(r, c).coeff(k);
            }
    scal[0] = f2w(em.total_weight);
    scal[1] = (uint32_t)em.iterations_run;
    scal[2] = (uint32_t)em.samples_seen;
    scal[3] = f2w(em.weight_prior);
    scal[4] = f2w(em.cov_prior_strength);
    scal[5] = f2w(em.learning_rate);
    scal[6] = f2w(em.alpha);
    scal[7] = 0;
}

template <typename EM>
void unpack_em(EM& em, int K, const float* e, const uint32_t* scal) {
    for (int k = 0; k < K; ++k) em.stats.weight.coeff(k) = e[k];
    for (int i = 0; i < 5; ++i)
        for (int k = 0; k < K; ++k) em.stats.mean.coeff(i).coeff(k) = e[(1 + i) * K + k];
    for (int r = 0; r < 5; ++r)
        for (int c = 0; c < 5; ++c)
            for (int k = 0; k < K; ++k) em.stats.cov(r, c).coeff(k) = e[(6 + 5 * r + c) * K + k];
    em.total_weight = w2f(scal[0]);
    em.iterations_run = (int)scal[1];
    em.samples_seen = (int)scal[2];
}

// ---- sample staging: the reference keeps one Data<> object per cell (opt/data.h:25-57); the device
// wants one SoA block per call.  Planes of all cells are gathered into one host block. --------------
struct SampleBlock {
    std::vector<float> point6;       // [6][n_total]
    std::vector<float> weight;       // [n_total]
    std::vector<uint32_t> offsets;   // [n_cells + 1]
    size_t n_total = 0;

    template <typename DataPtrIt>
    void gather(DataPtrIt first, DataPtrIt last) {
        offsets.assign(1, 0u);
        for (auto it = first; it != last; ++it) offsets.push_back(offsets.back() + (*it)->size);
        n_total = offsets.back();
        point6.resize(6 * n_total);
        weight.resize(n_total);
        size_t c = 0;
        for (auto it = first; it != last; ++it, ++c) {
            const auto& d = **it;
            const size_t b = offsets[c], n = d.size;
            for (int i = 0; i < 6; ++i) std::memcpy(point6.data() + i * n_total + b, d.point.coeff(i).data(), n * sizeof(float));
            std::memcpy(weight.data() + b, d.weight.data(), n * sizeof(float));
        }
    }
};

// ---- sdmm::em_step (opt/em.h:218-253) for ONE cell held in device slot `slot` ---------------------
// Uploads the cell's model + EM state, binds its samples, runs one stepwise EM iteration on the GPU
// and writes the result back into the reference objects.  Same observable behaviour as the
// reference: a step whose normalisation fails leaves the model untouched (em.h:237-240).
template <typename SDMM_, typename EM_, typename Data_>
void em_step(Device& dev, int slot, SDMM_& distribution, EM_& em, Data_& data, int n_iterations = 1) {
    const int K = dev.K();
    std::vector<float> joint(32 * (size_t)K), e(81 * (size_t)K);
    uint32_t scal[8];
    pack_joint(distribution, K, joint.data());
    pack_em(em, K, e.data(), scal);
    check(sdmmb_upload_cells(dev.ctx(), slot, 1, joint.data(), e.data(), scal, nullptr), "sdmmb_upload_cells", dev.ctx());
    SampleBlock blk;
    const Data_* one[1] = {&data};
    blk.gather(one, one + 1);
    check(sdmmb_set_samples_host(dev.ctx(), slot, 1, blk.point6.data(), blk.n_total, blk.weight.data(),
                                 blk.offsets.data(), nullptr), "sdmmb_set_samples_host", dev.ctx());
    check(sdmmb_em_step(dev.ctx(), slot, 1, n_iterations, 0, nullptr, nullptr), "sdmmb_em_step", dev.ctx());
    check(sdmmb_download_cells(dev.ctx(), slot, 1, joint.data(), e.data(), scal, nullptr), "sdmmb_download_cells", dev.ctx());
    unpack_joint(distribution, K, joint.data());
    unpack_em(em, K, e.data(), scal);
}

// ---- sdmm::prepare(conditioner, joint) (distributions/sdmm_conditioner.h:89-100) ------------------
// Builds the device query records of slot `slot` (what sdmmb_sample_pdf reads) and, so that host-side
// users of the conditioner object keep working, the reference's own host conditioner (O(K)).
template <typename Conditioner, typename SDMM_>
void prepare(Device& dev, int slot, Conditioner& conditioner, SDMM_& distribution, bool upload_model = true) {
    const int K = dev.K();
    if (upload_model) {
        std::vector<float> joint(32 * (size_t)K);
        pack_joint(distribution, K, joint.data());
        check(sdmmb_upload_cells(dev.ctx(), slot, 1, joint.data(), nullptr, nullptr, nullptr), "sdmmb_upload_cells", dev.ctx());
    }
    check(sdmmb_prepare_conditioner(dev.ctx(), slot, 1, nullptr), "sdmmb_prepare_conditioner", dev.ctx());
    enoki::set_slices(conditioner, enoki::slices(distribution));
    sdmm::prepare(conditioner, distribution);
}

// ---- the body of SDMMVolumetricPathTracer::optimize's parallelFor (volpath_sdmm.cpp:287-311) for ALL
// trainable cells in one go.  `contexts[i]` is the SDMMContext* of device slot first_slot + i; its
// `training_data` must already hold the samples (the swap at :290 stays at the call site).  Cells whose
// em.iterations_run < 4 get 2 iterations, the others 1 (:299-302); then the conditioner (:306-307).
template <typename ContextPtr>
void optimize(Device& dev, int first_slot, const std::vector<ContextPtr>& contexts) {
    const int K = dev.K();
    const int n = (int)contexts.size();
    if (n == 0) return;
    // device slots are ordered: cells that need two iterations first
    std::vector<int> order;
    int n_two = 0;
    for (int i = 0; i < n; ++i) if (contexts[i]->em.iterations_run < 4) { order.push_back(i); ++n_two; }
    for (int i = 0; i < n; ++i) if (!(contexts[i]->em.iterations_run < 4)) order.push_back(i);
    std::vector<float> joint((size_t)n * 32 * K), e((size_t)n * 81 * K);
    std::vector<uint32_t> scal((size_t)n * 8);
    std::vector<const decltype(contexts[0]->training_data)*> datas(n);
    for (int s = 0; s < n; ++s) {
        auto& c = *contexts[order[s]];
        pack_joint(c.sdmm, K, joint.data() + (size_t)s * 32 * K);
        pack_em(c.em, K, e.data() + (size_t)s * 81 * K, scal.data() + (size_t)s * 8);
        datas[s] = &c.training_data;
    }
    SampleBlock blk;
    blk.gather(datas.begin(), datas.end());
    sdmmb_ctx* ctx = dev.ctx();
    check(sdmmb_upload_cells(ctx, first_slot, n, joint.data(), e.data(), scal.data(), nullptr), "sdmmb_upload_cells", ctx);
    check(sdmmb_set_samples_host(ctx, first_slot, n, blk.point6.data(), blk.n_total, blk.weight.data(),
                                 blk.offsets.data(), nullptr), "sdmmb_set_samples_host", ctx);
    if (n_two > 0) check(sdmmb_em_step(ctx, first_slot, n_two, 1, 0, nullptr, nullptr), "sdmmb_em_step", ctx);
    check(sdmmb_em_step(ctx, first_slot, n, 1, 1, nullptr, nullptr), "sdmmb_em_step", ctx);
    check(sdmmb_download_cells(ctx, first_slot, n, joint.data(), e.data(), scal.data(), nullptr), "sdmmb_download_cells", ctx);
    for (int s = 0; s < n; ++s) {
        auto& c = *contexts[order[s]];
        unpack_joint(c.sdmm, K, joint.data() + (size_t)s * 32 * K);
        unpack_em(c.em, K, e.data() + (size_t)s * 81 * K, scal.data() + (size_t)s * 8);
        enoki::set_slices(c.conditioner, enoki::slices(c.sdmm));
        sdmm::prepare(c.conditioner, c.sdmm);
        c.training_data.clear();
        c.data.clear();
        c.initialized = true;
    }
}

}  // namespace sdmm_b200

// ---- spatial cell structure -------------------------------------------------------
// Flattens the reference's STree (accelerators/spatial_tree.h) into the device node table and assigns
// device slots to the leaves (in node order).  `slot_of_node[i]` = slot of leaf node i, -1 for inner
// nodes; `contexts[slot]` = the leaf's SDMMContext*, ready for sdmm_b200::optimize.
namespace sdmm_b200 {

template <typename Tree, typename ContextPtr>
void flatten_tree(Tree& tree, std::vector<sdmmb_tree_node>& nodes, std::vector<int>& slot_of_node,
                  std::vector<ContextPtr>& contexts) {
    const auto& src = tree.data();
    const int n = (int)tree.size();
    nodes.assign(n, sdmmb_tree_node{});
    slot_of_node.assign(n, -1);
    contexts.clear();
    for (int i = 0; i < n; ++i) {
        if (src[i].is_leaf) {
            slot_of_node[i] = (int)contexts.size();
            contexts.push_back(src[i].value.get());
            nodes[i] = sdmmb_tree_node{-1, 0.f, (uint32_t)slot_of_node[i], 0u};
        } else {
            const int axis = src[i].axis;
            const uint32_t c0 = src[i].children[0], c1 = src[i].children[1];
            // find() descends into children[0] iff nodes[children[0]].aabb.min[axis] < p[axis]  (:36-45)
            nodes[i] = sdmmb_tree_node{axis, (float)src[c0].aabb.min.coeff(axis), c0, c1};
        }
    }
}

template <typename Tree, typename ContextPtr>
void upload_tree(Device& dev, Tree& tree, std::vector<int>& slot_of_node, std::vector<ContextPtr>& contexts) {
    std::vector<sdmmb_tree_node> nodes;
    flatten_tree(tree, nodes, slot_of_node, contexts);
    check(sdmmb_set_tree(dev.ctx(), (int)nodes.size(), nodes.data()), "sdmmb_set_tree", dev.ctx());
}

}  // namespace sdmm_b200
