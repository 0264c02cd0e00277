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
//   sdmm::em_step(sdmm, em, data)        opt/em.h:218-253       sdmm_b200::em_step(sdmm, em, data)        [bound device]
//   sdmm::prepare(conditioner, sdmm)     sdmm_conditioner.h:89  sdmm_b200::prepare(conditioner, sdmm)     [bound device]
//   the parallelFor body                 volpath_sdmm.cpp:287   sdmm_b200::optimize(dev, first_slot, contexts)
//   create_conditional + sample + posterior per vertex          sdmm_b200::sample_pdf(dev, QueryBatch&) (batched),
//     sdmm_conditioner.h:102-107, sdmm.h:93-97, 112-113         sdmm_b200::sample(conditioner, x, u) (one vertex)
// The overloads without a Device argument have the reference's exact parameter lists: they use the Device
// bound with sdmm_b200::bind_default_device() and the slot the Device assigned to the object (its address) the
// first time it saw it, so call sites change only their namespace.
//
// There is no CPU fallback: every function needs the CUDA library and a device; errors surface as
// std::runtime_error carrying the C ABI's error string (the reference throws std::runtime_error for
// "Data full." too, opt/data.h:80).
#pragma once
#include <cstdint>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>
#include <unordered_map>
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
    Device(int device, int max_cells, int K) : K_(K), max_cells_(max_cells), resident_((size_t)max_cells) {
        check(sdmmb_create(&ctx_, device, max_cells, K), "sdmmb_create");
    }
    ~Device() { sdmmb_destroy(ctx_); }
    Device(const Device&) = delete;
    Device& operator=(const Device&) = delete;
    sdmmb_ctx* ctx() const { return ctx_; }
    int K() const { return K_; }
    int max_cells() const { return max_cells_; }

    // Device slot of a reference object (keyed by its address): assigned in order of first use.  What lets
    // em_step(sdmm, em, data) / prepare(conditioner, sdmm) keep the reference's parameter lists.
    int slot_of(const void* key) {
        std::lock_guard<std::mutex> lk(mu_);
        auto it = slots_.find(key);
        if (it != slots_.end()) return it->second;
        if (next_slot_ >= max_cells_) throw std::runtime_error("sdmm_b200::Device: more cells than slots");
        const int s = next_slot_++;
        slots_.emplace(key, s);
        return s;
    }
    void alias(const void* key, const void* existing) {   // a second object of the same cell (its conditioner)
        const int s = slot_of(existing);
        std::lock_guard<std::mutex> lk(mu_);
        slots_[key] = s;
    }

    // Which host object the device copy of a slot mirrors, and the EM counters it had when the shim last wrote
    // it back: optimize() uploads only the cells whose host state changed since (a freshly initialised cell, a
    // child a split created from its parent, a model somebody edited and invalidate()d).
    struct Resident { const void* owner = nullptr; int iterations_run = -1; int samples_seen = -1; uint32_t total_weight = 0; };
    bool is_resident(int slot, const void* owner, int iterations_run, int samples_seen, float total_weight) const {
        const Resident& r = resident_[(size_t)slot];
        uint32_t tw;
        std::memcpy(&tw, &total_weight, 4);
        return r.owner == owner && r.iterations_run == iterations_run && r.samples_seen == samples_seen && r.total_weight == tw;
    }
    void set_resident(int slot, const void* owner, int iterations_run, int samples_seen, float total_weight) {
        Resident& r = resident_[(size_t)slot];
        r.owner = owner; r.iterations_run = iterations_run; r.samples_seen = samples_seen;
        std::memcpy(&r.total_weight, &total_weight, 4);
    }
    void invalidate(int slot) { resident_[(size_t)slot] = Resident{}; }

private:
    sdmmb_ctx* ctx_ = nullptr;
    int K_, max_cells_;
    std::mutex mu_;
    int next_slot_ = 0;
    std::unordered_map<const void*, int> slots_;
    std::vector<Resident> resident_;
};

// The Device the reference-signature overloads below use (one per process is the integrator's case).
inline Device*& default_device_ref() { static Device* d = nullptr; return d; }
inline void bind_default_device(Device* d) { default_device_ref() = d; }
inline Device& default_device() {
    if (!default_device_ref()) throw std::runtime_error("sdmm_b200: no default device bound (bind_default_device)");
    return *default_device_ref();
}

// page-locked host block (sdmmb_host_alloc): the host-buffer pipelines only overlap copies and kernels from it
template <typename T>
class PinnedBuffer {
public:
    PinnedBuffer() = default;
    ~PinnedBuffer() { sdmmb_host_free(p_); }
    PinnedBuffer(const PinnedBuffer&) = delete;
    PinnedBuffer& operator=(const PinnedBuffer&) = delete;
    void resize(size_t n) {
        if (n <= cap_) { n_ = n; return; }
        sdmmb_host_free(p_);
        p_ = nullptr; cap_ = n_ = 0;
        void* q = nullptr;
        check(sdmmb_host_alloc(&q, (n + n / 4) * sizeof(T)), "sdmmb_host_alloc");
        p_ = (T*)q; cap_ = n + n / 4; n_ = n;
    }
    T* data() { return p_; }
    const T* data() const { return p_; }
    size_t size() const { return n_; }
private:
    T* p_ = nullptr;
    size_t n_ = 0, cap_ = 0;
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

// what sdmm::initialize sets besides the statistics: the priors (opt/init.h:88, 178-192)
template <typename EM>
void unpack_em_priors(EM& em, int K, const float* e, const uint32_t* scal) {
    for (int r = 0; r < 5; ++r)
        for (int c = 0; c < 5; ++c)
            for (int k = 0; k < K; ++k) {
                em.cov_prior(r, c).coeff(k) = e[(31 + 5 * r + c) * K + k];
                em.depth_prior(r, c).coeff(k) = e[(56 + 5 * r + c) * K + k];
            }
    em.weight_prior = w2f(scal[3]);
    em.cov_prior_strength = w2f(scal[4]);
    em.learning_rate = w2f(scal[5]);
    em.alpha = w2f(scal[6]);
}

// ---- sample staging: the reference keeps one Data<> object per cell (opt/data.h:25-57); the device
// wants one SoA block per call.  Planes of all cells are gathered into one host block. --------------
struct SampleBlock {
    PinnedBuffer<float> point6;      // [6][n_total]
    PinnedBuffer<float> weight;      // [n_total]
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

// ---- sdmm::initialize (opt/init.h:17-193) for ONE cell held in device slot `slot` -----------------
// The reference takes the generator BY VALUE (initializeSDMMContext passes the context's rng, volpath_sdmm.cpp:132-137):
// the copy's (state, inc) is turned back into the (initstate, initseq) pair of PCG32::seed that the C ABI takes,
//   state = (inc + initstate) * M + inc   =>   initstate = (state - inc) * M^-1 - inc,  initseq = inc >> 1,
// so the device replays exactly the draws the reference's copy would make.
inline uint64_t pcg32_mult_inverse() {
    const uint64_t m = 0x5851f42d4c957f2dULL;
    uint64_t x = m;                                   // Newton iteration for the inverse modulo 2^64
    for (int i = 0; i < 6; ++i) x *= 2 - m * x;
    return x;
}
template <typename SDMM_, typename EM_, typename Data_, typename RNG_>
void initialize(Device& dev, int slot, SDMM_& distribution, EM_& em, Data_& data, RNG_ rng, size_t n_spatial_components,
                float spatial_distance) {
    const int K = dev.K();
    if ((size_t)K != 8 * n_spatial_components)
        throw std::runtime_error("sdmm_b200::initialize: the device context holds K = 8 x n_spatial_components components");
    const size_t n = data.size;
    PinnedBuffer<float> point6, normal3, weight;
    point6.resize(6 * n);
    normal3.resize(3 * n);
    weight.resize(n);
    for (int i = 0; i < 6; ++i) std::memcpy(point6.data() + i * n, data.point.coeff(i).data(), n * sizeof(float));
    for (int i = 0; i < 3; ++i) std::memcpy(normal3.data() + i * n, data.normal.coeff(i).data(), n * sizeof(float));
    std::memcpy(weight.data(), data.weight.data(), n * sizeof(float));
    const uint32_t offsets[2] = {0u, (uint32_t)n};
    const uint64_t state = (uint64_t)rng.state, inc = (uint64_t)rng.inc;
    const uint64_t initstate = (state - inc) * pcg32_mult_inverse() - inc, initseq = inc >> 1;
    check(sdmmb_initialize_host(dev.ctx(), slot, 1, point6.data(), normal3.data(), n, weight.data(), offsets,
                                &spatial_distance, initstate, initseq, nullptr, nullptr), "sdmmb_initialize_host", dev.ctx());
    std::vector<float> joint(32 * (size_t)K), e(81 * (size_t)K);
    uint32_t scal[8];
    check(sdmmb_download_cells(dev.ctx(), slot, 1, joint.data(), e.data(), scal, nullptr), "sdmmb_download_cells", dev.ctx());
    enoki::set_slices(distribution, K);
    enoki::set_slices(em, K);
    unpack_joint(distribution, K, joint.data());
    unpack_em(em, K, e.data(), scal);
    unpack_em_priors(em, K, e.data(), scal);
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
// trainable cells in one go.  `contexts[i]` is the SDMMContext* of device slot first_slot + i (slots are stable:
// flatten_tree's leaf order); its `training_data` must already hold the samples (the swap at :290 stays at the
// call site).  One sdmmb_optimize_host call: the samples cross PCIe chunk by chunk while earlier chunks are being
// fitted and fitted mixtures travel back; cells whose em.iterations_run < 4 get 2 iterations, the others 1
// (:299-302, applied per cell on the device: SDMMB_ITERS_INTEGRATOR); then the conditioner (:306-307).  Only the
// cells whose host objects changed since the shim last wrote them are uploaded first.
struct OptimizeScratch {               // reused across calls (pinned)
    SampleBlock samples;
    PinnedBuffer<float> joint, em;
    PinnedBuffer<uint32_t> scal;
};

template <typename ContextPtr>
void optimize(Device& dev, int first_slot, const std::vector<ContextPtr>& contexts, OptimizeScratch* scratch = nullptr) {
    const int K = dev.K();
    const int n = (int)contexts.size();
    if (n == 0) return;
    OptimizeScratch local;
    OptimizeScratch& S = scratch ? *scratch : local;
    S.joint.resize((size_t)n * 32 * K);
    S.em.resize((size_t)n * 81 * K);
    S.scal.resize((size_t)n * 8);
    sdmmb_ctx* ctx = dev.ctx();
    // upload runs of cells the device does not mirror yet
    int run_begin = -1;
    auto flush = [&](int end) {
        if (run_begin < 0) return;
        check(sdmmb_upload_cells(ctx, first_slot + run_begin, end - run_begin, S.joint.data() + (size_t)run_begin * 32 * K,
                                 S.em.data() + (size_t)run_begin * 81 * K, S.scal.data() + (size_t)run_begin * 8, nullptr),
              "sdmmb_upload_cells", ctx);
        run_begin = -1;
    };
    std::vector<const decltype(contexts[0]->training_data)*> datas(n);
    for (int s = 0; s < n; ++s) {
        auto& c = *contexts[s];
        datas[s] = &c.training_data;
        if (dev.is_resident(first_slot + s, &c, c.em.iterations_run, c.em.samples_seen, c.em.total_weight)) {
            flush(s);
            continue;
        }
        pack_joint(c.sdmm, K, S.joint.data() + (size_t)s * 32 * K);
        pack_em(c.em, K, S.em.data() + (size_t)s * 81 * K, S.scal.data() + (size_t)s * 8);
        if (run_begin < 0) run_begin = s;
    }
    flush(n);
    S.samples.gather(datas.begin(), datas.end());
    check(sdmmb_optimize_host(ctx, first_slot, n, S.samples.point6.data(), S.samples.n_total, S.samples.weight.data(),
                              S.samples.offsets.data(), SDMMB_ITERS_INTEGRATOR, 1, S.joint.data(), S.em.data(),
                              S.scal.data(), nullptr), "sdmmb_optimize_host", ctx);
    for (int s = 0; s < n; ++s) {
        auto& c = *contexts[s];
        unpack_joint(c.sdmm, K, S.joint.data() + (size_t)s * 32 * K);
        unpack_em(c.em, K, S.em.data() + (size_t)s * 81 * K, S.scal.data() + (size_t)s * 8);
        enoki::set_slices(c.conditioner, enoki::slices(c.sdmm));
        sdmm::prepare(c.conditioner, c.sdmm);
        c.training_data.clear();
        c.data.clear();
        c.initialized = true;
        dev.set_resident(first_slot + s, &c, c.em.iterations_run, c.em.samples_seen, c.em.total_weight);
    }
}

// ---- the reference's exact parameter lists over the bound default device ---------------------------------
// void sdmm::em_step(SDMM&, EM<SDMM>&, Data<SDMM>&)            opt/em.h:218-219
template <typename SDMM_, typename EM_, typename Data_>
void em_step(SDMM_& distribution, EM_& em, Data_& data) {
    Device& dev = default_device();
    em_step(dev, dev.slot_of(&distribution), distribution, em, data, 1);
}
// void sdmm::initialize(SDMM&, EM&, Data&, RNG, size_t n_spatial_components, float spatial_distance)   opt/init.h:17-24
template <typename SDMM_, typename EM_, typename Data_, typename RNG_>
void initialize(SDMM_& distribution, EM_& em, Data_& data, RNG_ rng, size_t n_spatial_components, float spatial_distance) {
    Device& dev = default_device();
    initialize(dev, dev.slot_of(&distribution), distribution, em, data, rng, n_spatial_components, spatial_distance);
}
// void sdmm::prepare(Conditioner&, JointSDMM&)                 distributions/sdmm_conditioner.h:89-93
template <typename Conditioner, typename SDMM_>
void prepare(Conditioner& conditioner, SDMM_& distribution) {
    Device& dev = default_device();
    const int slot = dev.slot_of(&distribution);
    dev.alias(&conditioner, &distribution);
    prepare(dev, slot, conditioner, distribution);
}

// ---- guided queries ----------------------------------------------------------------------------------------
// The reference answers a path vertex with three calls on thread_local scratch objects (sdmm_proc.cpp:361-372,
// 419-421, 539-545): create_conditional(conditioner, x, conditional) (sdmm_conditioner.h:102-107),
// conditional.sample(rng, dir, inv_jacobian, tangent) (sdmm.h:93-97), conditional.posterior(dir, posterior) + hsum
// (sdmm.h:112-113).  The device answers a whole batch of vertices in one fused kernel; the conditional mixture
// never leaves the GPU.  `slot[i]` = device slot of the vertex's cell (Device::slot_of(&context->sdmm), or
// sdmmb_find_cells on the uploaded tree); u = the three uniforms the reference's RNG would have drawn.
struct QueryBatch {
    std::vector<uint32_t> slot;                    // [n]
    std::vector<float> pos3, u3, dir_in3;          // [3][n] planes (u3 for sample_pdf, dir_in3 for pdf)
    std::vector<float> dir3, pdf, inv_jacobian;    // results
    std::vector<uint32_t> component;
    std::vector<uint8_t> valid;                    // 0 = create_conditional returned false
    size_t size() const { return slot.size(); }
    void resize(size_t n) {
        slot.resize(n); pos3.resize(3 * n); u3.resize(3 * n); dir_in3.resize(3 * n);
        dir3.resize(3 * n); pdf.resize(n); inv_jacobian.resize(n); component.resize(n); valid.resize(n);
    }
};

// create_conditional + SDMM::sample + posterior/hsum for every vertex of the batch
inline void sample_pdf(Device& dev, QueryBatch& q, bool binned = false) {
    check(sdmmb_sample_pdf_host(dev.ctx(), q.size(), q.slot.data(), q.pos3.data(), q.u3.data(), q.dir3.data(), q.pdf.data(),
                                q.inv_jacobian.data(), q.component.data(), q.valid.data(), binned ? 1 : 0, nullptr),
          "sdmmb_sample_pdf_host", dev.ctx());
}
// create_conditional + posterior/hsum at given directions (pdfSurface, sdmm_proc.cpp:510-590)
inline void pdf(Device& dev, QueryBatch& q, bool binned = false) {
    check(sdmmb_pdf_host(dev.ctx(), q.size(), q.slot.data(), q.pos3.data(), q.dir_in3.data(), q.pdf.data(), q.valid.data(),
                         binned ? 1 : 0, nullptr), "sdmmb_pdf_host", dev.ctx());
}

// One vertex with the reference's own types: fills the caller's scratch exactly where sampleSurface reads it.
// `conditioner` identifies the cell (prepare() registered it); point = the normalised position; u0,u1,u2 = the
// uniforms.  Returns create_conditional's bool; sample / inv_jacobian / pdf as SDMM::sample and hsum(posterior).
template <typename Conditioner, typename MarginalEmbedded, typename EmbeddedOut, typename Scalar>
bool sample(Conditioner& conditioner, const MarginalEmbedded& point, float u0, float u1, float u2, EmbeddedOut& sample_out,
            Scalar& inv_jacobian, Scalar& pdf_out, uint32_t* component = nullptr) {
    Device& dev = default_device();
    const uint32_t slot = (uint32_t)dev.slot_of(&conditioner);
    const float pos[3] = {(float)point.coeff(0), (float)point.coeff(1), (float)point.coeff(2)};
    const float u[3] = {u0, u1, u2};
    float dir[3], p = 0.f, ij = 0.f;
    uint32_t comp = 0;
    uint8_t ok = 0;
    check(sdmmb_sample_pdf_host(dev.ctx(), 1, &slot, pos, u, dir, &p, &ij, &comp, &ok, 1, nullptr), "sdmmb_sample_pdf_host",
          dev.ctx());
    for (int i = 0; i < 3; ++i) sample_out.coeff(i) = dir[i];
    inv_jacobian = ij;
    pdf_out = p;
    if (component) *component = comp;
    return ok != 0;
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
