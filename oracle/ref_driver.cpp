// oracle/ref_driver.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Thin extern "C" driver around the UNMODIFIED reference headers under
// /root/reference/sdmm/include (+ vendored enoki), compiled by oracle/Makefile into
// oracle/_ref/libsdmm_ref.so.  It instantiates exactly the types the shipping integrator
// uses (mitsuba/src/integrators/sdmm/sdmm_proc.h:55-96: Packet<float,8>, DynamicArray,
// JointSDMM / MarginalSDMM / ConditionalSDMM(OffsetDirectional) / Conditioner / EM / Data)
// and calls the reference's own functions:
//   sdmm::em_step                 sdmm/include/sdmm/opt/em.h:218-253
//   sdmm::prepare_vectorized      sdmm/include/sdmm/distributions/sdmm.h:163-173
//   sdmm::prepare(conditioner,..) sdmm/include/sdmm/distributions/sdmm_conditioner.h:89-100
//   sdmm::create_conditional      sdmm/include/sdmm/distributions/sdmm_conditioner.h:102-115
//   SDMM::sample / posterior      sdmm/include/sdmm/distributions/sdmm.h:209-273, 331-339
// Nothing of the reference is copied; this file only marshals flat float arrays in and out.
// Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load the result.
//
// Flat layouts (shared with oracle/sdmm_oracle.c and include/sdmm_b200.h):
//   joint : 32 planes of K floats: [0]=pmf, [1..6]=mean (x,y,z,dx,dy,dz), [7+5r+c]=cov(r,c)
//   em    : 81 planes of K floats: [0]=stats.weight, [1..5]=stats.mean, [6+5r+c]=stats.cov,
//           [31+5r+c]=cov_prior, [56+5r+c]=depth_prior
//   scal  : 8 x 32-bit words: f32 total_weight, i32 iterations_run, i32 samples_seen,
//           f32 weight_prior, f32 cov_prior_strength, f32 learning_rate, f32 alpha, i32 status
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include <enoki/array.h>
#include <enoki/dynamic.h>
#include <enoki/random.h>

#include "sdmm/core/utils.h"
#include "sdmm/distributions/sdmm.h"
#include "sdmm/distributions/sdmm_conditioner.h"
// the json.hpp vendored with the reference (ext/jmm) predates this macro; serialisation is not on the path
#ifndef NLOHMANN_DEFINE_TYPE_INTRUSIVE
#define NLOHMANN_DEFINE_TYPE_INTRUSIVE(...)
#endif
#include "sdmm/accelerators/spatial_tree.h"
#include "sdmm/distributions/sdmm_context.h"
#include "sdmm/linalg/cholesky.h"
#include "sdmm/linalg/coordinate_system.h"
#include "sdmm/opt/data.h"
#include "sdmm/opt/em.h"
#include "sdmm/opt/init.h"  // boost chi_squared comes from oracle/stubs (constant quantile)
#include "sdmm/spaces/directional.h"
#include "sdmm/spaces/euclidian.h"
#include "sdmm/spaces/offset_directional.h"
#include "sdmm/spaces/spatio_directional.h"

#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

constexpr size_t PacketSize = 8;
constexpr size_t JointSize = 5;
constexpr size_t MarginalSize = 3;
constexpr size_t ConditionalSize = 2;

using Packet = enoki::Packet<float, PacketSize>;
using Value = enoki::DynamicArray<Packet>;
using JointTangentSpace = sdmm::SpatioDirectionalTangentSpace<
    sdmm::Vector<Value, JointSize + 1>, sdmm::Vector<Value, JointSize>>;
using JointSDMM = sdmm::SDMM<sdmm::Matrix<Value, JointSize>, JointTangentSpace>;
using MarginalTangentSpace = sdmm::EuclidianTangentSpace<
    sdmm::Vector<Value, MarginalSize>, sdmm::Vector<Value, MarginalSize>>;
using MarginalSDMM = sdmm::SDMM<sdmm::Matrix<Value, MarginalSize>, MarginalTangentSpace>;
using ConditionalTangentSpace = sdmm::OffsetDirectionalTangentSpace<
    sdmm::Vector<Value, ConditionalSize + 1>, sdmm::Vector<Value, ConditionalSize>>;
using ConditionalSDMM = sdmm::SDMM<sdmm::Matrix<Value, ConditionalSize>, ConditionalTangentSpace>;
using Conditioner = sdmm::SDMMConditioner<JointSDMM, MarginalSDMM, ConditionalSDMM>;
using Data = sdmm::Data<JointSDMM>;
using EM = sdmm::EM<JointSDMM>;

// RNG that replays caller-supplied uniforms in draw order (SDMM::sample only needs
// next_float32(), Float32 and UInt32; sdmm.h:220-222).
struct ReplayRNG {
    using Float32 = float;
    using UInt32 = uint32_t;
    float draws[3];
    int next = 0;
    float next_float32() { return draws[next++ % 3]; }
};

// box_mueller_transform(rng.next_float32(), rng.next_float32()) (sdmm.h:239-240) has
// unspecified argument evaluation order; probe which draw lands in which parameter with
// THIS compiler so the by-role (u_radius, u_angle) interface is unambiguous.
//   returns true when the FIRST draw becomes the first parameter (u1 = radius).
bool first_draw_is_radius() {
    ReplayRNG probe;
    probe.draws[0] = 0.25f;
    probe.draws[1] = 0.75f;
    probe.next = 0;
    auto [a, b] = sdmm::box_mueller_transform(probe.next_float32(), probe.next_float32());
    (void)b;
    // radius from u1: sqrt(-2 log(1-u1)); angle 2 pi u2.  If u1=0.25,u2=0.75: a = sin(1.5pi)*r = -r(0.25)
    float r025 = std::sqrt(-2.f * std::log(1.f - 0.25f));
    return std::abs(a + r025) < 1e-4f;
}

struct RefCell {
    int K = 0;
    JointSDMM sdmm;
    EM em;
    Conditioner conditioner;
    // per-thread style scratch for queries (mirrors sdmm_proc.cpp:1086-1106 thread_locals)
    ConditionalSDMM conditional;
    Value posterior;
    bool conditioner_ready = false;
};

inline float word_to_f(uint32_t w) { float f; std::memcpy(&f, &w, 4); return f; }
inline uint32_t f_to_word(float f) { uint32_t w; std::memcpy(&w, &f, 4); return w; }

void load_data(Data& data, int N, const float* point6, size_t stride, const float* weight) {
    data.reserve((uint32_t)N);
    // Bulk fill (equivalent to N push_backs of valid samples; invalid ones are kept so the
    // EM loop's own is_valid_sample check (em.h:267-269) is exercised).
    for (int n = 0; n < N; ++n) {
        for (int d = 0; d < 6; ++d) data.point.coeff(d).coeff(n) = point6[d * stride + n];
        for (int d = 0; d < 3; ++d) data.normal.coeff(d).coeff(n) = 0.f;
        data.weight.coeff(n) = weight[n];
    }
    data.size = (uint32_t)N;
}

// log sum_k pi_k p_k(x_n) under the CURRENT model, using the same posterior call as the
// E-step (em.h:279-289).
void loglik(RefCell& c, Data& data, float* out) {
    Value post;
    enoki::set_slices(post, c.K);
    sdmm::tangent_t<JointSDMM> tangent;
    enoki::set_slices(tangent, c.K);
    for (uint32_t n = 0; n < data.size; ++n) {
        auto slice = enoki::slice(data, n);
        enoki::vectorize_safe(VECTORIZE_WRAP_MEMBER(posterior), c.sdmm, slice.point, post, tangent);
        out[n] = std::log(enoki::hsum(post));
    }
}

}  // namespace

extern "C" {

int sdmm_ref_packet_size() { return (int)PacketSize; }
int sdmm_ref_first_draw_is_radius() { return first_draw_is_radius() ? 1 : 0; }
int sdmm_ref_max_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void* sdmm_ref_cell_create(int K) {
    auto* c = new RefCell();
    c->K = K;
    c->sdmm = enoki::zero<JointSDMM>(K);
    enoki::set_slices(c->sdmm, K);
    c->em = enoki::zero<EM>(K);
    enoki::set_slices(c->em, K);
    enoki::set_slices(c->conditioner, K);
    enoki::set_slices(c->conditional, K);
    enoki::set_slices(c->posterior, K);
    return c;
}

void sdmm_ref_cell_destroy(void* h) { delete static_cast<RefCell*>(h); }

// Install a model: pmf, mean, cov -> set_mean (frame) + prepare_vectorized (Cholesky, CDF),
// i.e. what from_json does (sdmm.h:157-160) except that weights of non-PSD components are
// NOT zeroed (prepare_vectorized, the function em_step itself uses, sdmm.h:163-173).
// Returns the bool of prepare_vectorized.
int sdmm_ref_cell_set(void* h, const float* joint, const float* em, const uint32_t* scal) {
    auto& c = *static_cast<RefCell*>(h);
    const int K = c.K;
    for (int k = 0; k < K; ++k) c.sdmm.weight.pmf.coeff(k) = joint[0 * K + k];
    sdmm::embedded_t<JointSDMM> mean;
    enoki::set_slices(mean, K);
    for (int d = 0; d < 6; ++d)
        for (int k = 0; k < K; ++k) mean.coeff(d).coeff(k) = joint[(1 + d) * K + k];
    c.sdmm.tangent_space.set_mean(mean);
    for (int r = 0; r < 5; ++r)
        for (int col = 0; col < 5; ++col)
            for (int k = 0; k < K; ++k) c.sdmm.cov(r, col).coeff(k) = joint[(7 + 5 * r + col) * K + k];
    bool ok = sdmm::prepare_vectorized(c.sdmm);
    if (em) {
        for (int k = 0; k < K; ++k) c.em.stats.weight.coeff(k) = em[0 * K + k];
        for (int d = 0; d < 5; ++d)
            for (int k = 0; k < K; ++k) c.em.stats.mean.coeff(d).coeff(k) = em[(1 + d) * K + k];
        for (int r = 0; r < 5; ++r)
            for (int col = 0; col < 5; ++col)
                for (int k = 0; k < K; ++k) {
                    c.em.stats.cov(r, col).coeff(k) = em[(6 + 5 * r + col) * K + k];
                    c.em.cov_prior(r, col).coeff(k) = em[(31 + 5 * r + col) * K + k];
                    c.em.depth_prior(r, col).coeff(k) = em[(56 + 5 * r + col) * K + k];
                }
    }
    if (scal) {
        c.em.total_weight = word_to_f(scal[0]);
        c.em.iterations_run = (int)scal[1];
        c.em.samples_seen = (int)scal[2];
        c.em.weight_prior = word_to_f(scal[3]);
        c.em.cov_prior_strength = word_to_f(scal[4]);
        c.em.learning_rate = word_to_f(scal[5]);
        c.em.alpha = word_to_f(scal[6]);
    }
    c.conditioner_ready = false;
    return ok ? 1 : 0;
}

void sdmm_ref_cell_get(void* h, float* joint, float* em, uint32_t* scal) {
    auto& c = *static_cast<RefCell*>(h);
    const int K = c.K;
    if (joint) {
        for (int k = 0; k < K; ++k) joint[k] = c.sdmm.weight.pmf.coeff(k);
        for (int d = 0; d < 6; ++d)
            for (int k = 0; k < K; ++k) joint[(1 + d) * K + k] = c.sdmm.tangent_space.mean.coeff(d).coeff(k);
        for (int r = 0; r < 5; ++r)
            for (int col = 0; col < 5; ++col)
                for (int k = 0; k < K; ++k) joint[(7 + 5 * r + col) * K + k] = c.sdmm.cov(r, col).coeff(k);
    }
    if (em) {
        for (int k = 0; k < K; ++k) em[k] = c.em.stats.weight.coeff(k);
        for (int d = 0; d < 5; ++d)
            for (int k = 0; k < K; ++k) em[(1 + d) * K + k] = c.em.stats.mean.coeff(d).coeff(k);
        for (int r = 0; r < 5; ++r)
            for (int col = 0; col < 5; ++col)
                for (int k = 0; k < K; ++k) {
                    em[(6 + 5 * r + col) * K + k] = c.em.stats.cov(r, col).coeff(k);
                    em[(31 + 5 * r + col) * K + k] = c.em.cov_prior(r, col).coeff(k);
                    em[(56 + 5 * r + col) * K + k] = c.em.depth_prior(r, col).coeff(k);
                }
    }
    if (scal) {
        scal[0] = f_to_word(c.em.total_weight);
        scal[1] = (uint32_t)c.em.iterations_run;
        scal[2] = (uint32_t)c.em.samples_seen;
        scal[3] = f_to_word(c.em.weight_prior);
        scal[4] = f_to_word(c.em.cov_prior_strength);
        scal[5] = f_to_word(c.em.learning_rate);
        scal[6] = f_to_word(c.em.alpha);
        scal[7] = 0;
    }
}

// Derived per-component state for parity checks of the device records:
//   out planes (K floats each): [0..8] frame.to row-major, [9..33] cov_sqrt(r,c) row-major,
//   [34] inv_cov_sqrt_det, [35] cdf, [36] pmf (normalised), [37] cov_is_psd (0/1)
void sdmm_ref_cell_get_derived(void* h, float* out) {
    auto& c = *static_cast<RefCell*>(h);
    const int K = c.K;
    for (int r = 0; r < 3; ++r)
        for (int col = 0; col < 3; ++col)
            for (int k = 0; k < K; ++k)
                out[(3 * r + col) * K + k] = c.sdmm.tangent_space.coordinate_system.to(r, col).coeff(k);
    for (int r = 0; r < 5; ++r)
        for (int col = 0; col < 5; ++col)
            for (int k = 0; k < K; ++k) out[(9 + 5 * r + col) * K + k] = c.sdmm.cov_sqrt(r, col).coeff(k);
    for (int k = 0; k < K; ++k) {
        out[34 * K + k] = c.sdmm.inv_cov_sqrt_det.coeff(k);
        out[35 * K + k] = c.sdmm.weight.cdf.coeff(k);
        out[36 * K + k] = c.sdmm.weight.pmf.coeff(k);
        out[37 * K + k] = c.sdmm.cov_is_psd.coeff(k) ? 1.f : 0.f;
    }
}

// n_iters x sdmm::em_step on one cell.  point6: 6 planes with the given stride; loglik (may
// be null): n_iters x N, row i = log-likelihood of every sample under the model that
// ENTERS iteration i.
int sdmm_ref_em_step(void* h, int N, const float* point6, long stride, const float* weight,
                     int n_iters, float* out_loglik) {
    auto& c = *static_cast<RefCell*>(h);
    Data data;
    load_data(data, N, point6, (size_t)stride, weight);
    // the integrator's own floating-point mode: flushDenormals = true by default
    // (mitsuba/src/integrators/sdmm/volpath_sdmm.cpp:62, 85-87 -> enoki::set_flush_denormals(true))
    const bool was = enoki::flush_denormals();
    enoki::set_flush_denormals(true);
    for (int it = 0; it < n_iters; ++it) {
        if (out_loglik) loglik(c, data, out_loglik + (size_t)it * N);
        sdmm::em_step(c.sdmm, c.em, data);
    }
    enoki::set_flush_denormals(was);
    c.conditioner_ready = false;
    return 0;
}

// sdmm::initialize (opt/init.h:17-193) on one cell, exactly as initializeSDMMContext calls it
// (volpath_sdmm.cpp:132-137): the context's RNG is a default-seeded enoki::PCG32<float,1>
// passed BY VALUE.  The handle must have been created with K = 8 * n_spatial.  normal3: 3 planes
// with the same stride as point6.  Afterwards the cell holds the initial model + EM priors.
int sdmm_ref_initialize(void* h, int N, const float* point6, const float* normal3, long stride,
                        const float* weight, int n_spatial, float spatial_distance,
                        uint64_t rng_state, uint64_t rng_stream) {
    auto& c = *static_cast<RefCell*>(h);
    Data data;
    load_data(data, N, point6, (size_t)stride, weight);
    for (int n = 0; n < N; ++n)
        for (int d = 0; d < 3; ++d) data.normal.coeff(d).coeff(n) = normal3[d * stride + n];
    enoki::PCG32<float, 1> rng(rng_state, rng_stream);
    sdmm::initialize(c.sdmm, c.em, data, rng, (size_t)n_spatial, spatial_distance);
    c.K = (int)enoki::slices(c.sdmm);
    enoki::set_slices(c.conditioner, c.K);
    enoki::set_slices(c.conditional, c.K);
    enoki::set_slices(c.posterior, c.K);
    c.conditioner_ready = false;
    return c.K;
}

// enoki::PCG32<float,1>::next_float32 draws (ext/enoki/include/enoki/random.h:33-101), to pin the
// port's and the device's restatement of the generator.
void sdmm_ref_pcg32_floats(uint64_t state, uint64_t stream, int n, float* out) {
    enoki::PCG32<float, 1> rng(state, stream);
    for (int i = 0; i < n; ++i) out[i] = rng.next_float32();
}

// The timed CPU baseline: OpenMP parallel-for over cells, one cell per task -- the same
// decomposition as tev::ThreadPool::parallelFor at volpath_sdmm.cpp:287-311.  Sample
// marshalling into sdmm::Data happens before the timed region (returned seconds cover only
// the em_step calls + conditioner prepare, like optimize()).
double sdmm_ref_em_batch(void** cells, int C, const uint32_t* offsets, const float* point6,
                         long stride, const float* weight, int n_iters, int with_prepare,
                         int n_threads) {
    std::vector<Data> datas((size_t)C);
    for (int ci = 0; ci < C; ++ci) {
        uint32_t b = offsets[ci], e = offsets[ci + 1];
        load_data(datas[ci], (int)(e - b), point6 + b, (size_t)stride, weight + b);
    }
    double t0 = 0, t1 = 0;
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
    t0 = omp_get_wtime();
#pragma omp parallel for schedule(dynamic, 1)
#endif
    for (int ci = 0; ci < C; ++ci) {
        auto& c = *static_cast<RefCell*>(cells[ci]);
        const bool was = enoki::flush_denormals();
        enoki::set_flush_denormals(true);   // the integrator's mode (volpath_sdmm.cpp:62, 85-87)
        for (int it = 0; it < n_iters; ++it) sdmm::em_step(c.sdmm, c.em, datas[ci]);
        enoki::set_flush_denormals(was);
        if (with_prepare) {
            enoki::set_slices(c.conditioner, enoki::slices(c.sdmm));
            sdmm::prepare(c.conditioner, c.sdmm);
            c.conditioner_ready = true;
        }
    }
#ifdef _OPENMP
    t1 = omp_get_wtime();
#endif
    return t1 - t0;
}

void sdmm_ref_prepare_conditioner(void* h) {
    auto& c = *static_cast<RefCell*>(h);
    enoki::set_slices(c.conditioner, enoki::slices(c.sdmm));
    sdmm::prepare(c.conditioner, c.sdmm);
    c.conditioner_ready = true;
}

// Conditioner internals, planes of K floats:
//  [0..2] marginal mean, [3..11] marginal cov_sqrt(r,c) row-major, [12] marginal inv_det,
//  [13] marginal pmf (normalised by prepare), [14..19] mean_transform(r,c) (2x3 row-major),
//  [20..28] conditional frame.to (row-major), [29..32] conditional cov_sqrt (2x2 row-major),
//  [33] conditional inv_det, [34..37] conditional cov (2x2 row-major)
void sdmm_ref_get_conditioner(void* h, float* out) {
    auto& c = *static_cast<RefCell*>(h);
    const int K = c.K;
    auto& q = c.conditioner;
    for (int k = 0; k < K; ++k) {
        for (int d = 0; d < 3; ++d) out[d * K + k] = q.marginal.tangent_space.mean.coeff(d).coeff(k);
        for (int r = 0; r < 3; ++r)
            for (int col = 0; col < 3; ++col) out[(3 + 3 * r + col) * K + k] = q.marginal.cov_sqrt(r, col).coeff(k);
        out[12 * K + k] = q.marginal.inv_cov_sqrt_det.coeff(k);
        out[13 * K + k] = q.marginal.weight.pmf.coeff(k);
        for (int r = 0; r < 2; ++r)
            for (int col = 0; col < 3; ++col) out[(14 + 3 * r + col) * K + k] = q.mean_transform(r, col).coeff(k);
        for (int r = 0; r < 3; ++r)
            for (int col = 0; col < 3; ++col)
                out[(20 + 3 * r + col) * K + k] = q.tangent_space.coordinate_system.to(r, col).coeff(k);
        for (int r = 0; r < 2; ++r)
            for (int col = 0; col < 2; ++col) {
                out[(29 + 2 * r + col) * K + k] = q.conditional.cov_sqrt(r, col).coeff(k);
                out[(34 + 2 * r + col) * K + k] = q.conditional.cov(r, col).coeff(k);
            }
        out[33 * K + k] = q.conditional.inv_cov_sqrt_det.coeff(k);
    }
}

// One guided query on one cell, as sampleSurface does it (sdmm_proc.cpp:361-372, 419-421,
// 491-501 -> 539-545): create_conditional(x) -> conditional.sample(u) -> posterior(omega)
// -> hsum.  u is given BY ROLE: u[0] = component-select uniform, u[1] = Box-Muller radius
// uniform, u[2] = Box-Muller angle uniform.
// Returns 1 when create_conditional succeeded (valid CDF), else 0 (outputs zeroed).
static int query_one(RefCell& c, ConditionalSDMM& conditional, Value& posterior, bool radius_first,
                     const float x[3], const float u[3], float dir[3], float* pdf, float* inv_jac,
                     uint32_t* comp, float* pmf_sel) {
    sdmm::embedded_s_t<MarginalSDMM> condition({x[0], x[1], x[2]});
    if (enoki::slices(conditional) != enoki::slices(c.conditioner))
        enoki::set_slices(conditional, enoki::slices(c.conditioner));
    if (enoki::slices(posterior) != enoki::slices(c.conditioner))
        enoki::set_slices(posterior, enoki::slices(c.conditioner));
    bool valid = sdmm::create_conditional(c.conditioner, condition, conditional);
    if (!valid) {
        dir[0] = dir[1] = dir[2] = 0.f;
        *pdf = 0.f; *inv_jac = 0.f; *comp = 0u;
        if (pmf_sel) *pmf_sel = 0.f;
        return 0;
    }
    ReplayRNG rng;
    rng.draws[0] = u[0];
    if (radius_first) { rng.draws[1] = u[1]; rng.draws[2] = u[2]; }
    else              { rng.draws[1] = u[2]; rng.draws[2] = u[1]; }
    rng.next = 0;
    sdmm::embedded_s_t<ConditionalSDMM> embedded;
    sdmm::tangent_s_t<ConditionalSDMM> tangent;
    float ij = 0.f;
    conditional.sample(rng, embedded, ij, tangent);
    // recover the selected index the same way SDMM::sample does (sdmm.h:222-233)
    uint32_t idx = enoki::binary_search(0, (uint32_t)enoki::slices(conditional.weight.cdf) - 1,
                                        [&](uint32_t i) { return conditional.weight.cdf[i] < u[0]; });
    while (idx > 0 && conditional.weight.pmf[idx] == 0) --idx;
    *comp = idx;
    if (pmf_sel) *pmf_sel = conditional.weight.pmf[idx];
    dir[0] = embedded.coeff(0); dir[1] = embedded.coeff(1); dir[2] = embedded.coeff(2);
    *inv_jac = ij;
    enoki::vectorize_safe(VECTORIZE_WRAP_MEMBER(posterior), conditional, embedded, posterior);
    *pdf = enoki::hsum_nested(posterior);
    return 1;
}

// Batched guided sample+pdf.  pos3 / u3 / out_dir3 are SoA (3 planes of n).  Cells must have
// had sdmm_ref_prepare_conditioner called.  OpenMP over vertices with per-thread scratch
// (as sdmm_proc.cpp:1086-1106).  Returns seconds spent in the loop.
double sdmm_ref_sample_pdf(void** cells, long n, const uint32_t* cell_idx, const float* pos3,
                           const float* u3, float* out_dir3, float* out_pdf, float* out_inv_jac,
                           uint32_t* out_comp, uint8_t* out_valid, int n_threads) {
    const bool radius_first = first_draw_is_radius();
    double t0 = 0, t1 = 0;
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
    t0 = omp_get_wtime();
#pragma omp parallel
#endif
    {
        ConditionalSDMM conditional;
        Value posterior;
#ifdef _OPENMP
#pragma omp for schedule(static)
#endif
        for (long i = 0; i < n; ++i) {
            auto& c = *static_cast<RefCell*>(cells[cell_idx[i]]);
            float x[3] = {pos3[i], pos3[n + i], pos3[2 * n + i]};
            float u[3] = {u3[i], u3[n + i], u3[2 * n + i]};
            float dir[3], pdf, ij;
            uint32_t comp;
            int ok = query_one(c, conditional, posterior, radius_first, x, u, dir, &pdf, &ij, &comp, nullptr);
            out_dir3[i] = dir[0]; out_dir3[n + i] = dir[1]; out_dir3[2 * n + i] = dir[2];
            out_pdf[i] = pdf; out_inv_jac[i] = ij; out_comp[i] = comp;
            if (out_valid) out_valid[i] = (uint8_t)ok;
        }
    }
#ifdef _OPENMP
    t1 = omp_get_wtime();
#endif
    return t1 - t0;
}

// pdfSurface for a caller-supplied direction (BSDF-sampled branch, sdmm_proc.cpp:491-501):
// create_conditional(x) -> posterior(omega) -> hsum.  Also returns the normalised conditional
// pmf/cdf of one vertex when out_pmf/out_cdf are non-null (n must then be 1).
double sdmm_ref_pdf(void** cells, long n, const uint32_t* cell_idx, const float* pos3,
                    const float* dir3, float* out_pdf, uint8_t* out_valid, float* out_pmf,
                    float* out_cdf, int n_threads) {
    double t0 = 0, t1 = 0;
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
    t0 = omp_get_wtime();
#pragma omp parallel
#endif
    {
        ConditionalSDMM conditional;
        Value posterior;
#ifdef _OPENMP
#pragma omp for schedule(static)
#endif
        for (long i = 0; i < n; ++i) {
            auto& c = *static_cast<RefCell*>(cells[cell_idx[i]]);
            sdmm::embedded_s_t<MarginalSDMM> condition({pos3[i], pos3[n + i], pos3[2 * n + i]});
            if (enoki::slices(conditional) != enoki::slices(c.conditioner))
                enoki::set_slices(conditional, enoki::slices(c.conditioner));
            if (enoki::slices(posterior) != enoki::slices(c.conditioner))
                enoki::set_slices(posterior, enoki::slices(c.conditioner));
            bool valid = sdmm::create_conditional(c.conditioner, condition, conditional);
            float pdf = 0.f;
            if (valid) {
                sdmm::embedded_s_t<ConditionalSDMM> omega({dir3[i], dir3[n + i], dir3[2 * n + i]});
                enoki::vectorize_safe(VECTORIZE_WRAP_MEMBER(posterior), conditional, omega, posterior);
                pdf = enoki::hsum_nested(posterior);
            }
            out_pdf[i] = pdf;
            if (out_valid) out_valid[i] = valid ? 1 : 0;
            if (out_pmf && out_cdf && n == 1) {
                for (int k = 0; k < c.K; ++k) {
                    out_pmf[k] = conditional.weight.pmf.coeff(k);
                    out_cdf[k] = conditional.weight.cdf.coeff(k);
                }
            }
        }
    }
#ifdef _OPENMP
    t1 = omp_get_wtime();
#endif
    return t1 - t0;
}

// ---------------------------------------------------------------------------------------
// Known-answer entry points: each rebuilds the set-up of one of the reference's own doctest
// cases (which do not build at this commit) and returns the computed values so
// tests/ can compare them with the literals the reference test holds.
// ---------------------------------------------------------------------------------------

// sdmm/tests/distributions.cpp:16-36  "SDMM::pdf<float>": 2-D Euclidean, cov diag(1,2), x=(1,2)
float sdmm_ref_kat_pdf_scalar() {
    using V = float;
    using TS = sdmm::EuclidianTangentSpace<sdmm::Vector<V, 2>, sdmm::Vector<V, 2>>;
    using S = sdmm::SDMM<sdmm::Matrix<V, 2>, TS>;
    S d;
    d.weight.pmf = 1;
    d.tangent_space.set_mean(sdmm::embedded_t<S>(0));
    d.cov = enoki::diag<sdmm::matrix_t<S>>({1, 2});
    d.prepare_cov();
    sdmm::embedded_s_t<S> point({1, 2});
    V pdf(0);
    d.pdf_gaussian(point, pdf);
    return pdf;
}

// sdmm/tests/distributions.cpp:148-248 "SDMM::pdf<DynamicArray>": two 2-D components.
// out[0..1] = pdf at (1,2); out[2..3] = pdf at (0,0); out[4..5] = posterior at (1,2)
void sdmm_ref_kat_pdf_dynamic(float* out) {
    using P = enoki::Packet<float, 2>;
    using V = enoki::DynamicArray<P>;
    using TS = sdmm::EuclidianTangentSpace<sdmm::Vector<V, 2>, sdmm::Vector<V, 2>>;
    using S = sdmm::SDMM<sdmm::Matrix<V, 2>, TS>;
    S d;
    enoki::set_slices(d, 2);
    d.weight.pmf = V(0.2, 0.8);
    d.tangent_space.set_mean(sdmm::embedded_t<S>(V(0, 0), V(1, 1)));
    d.cov = sdmm::matrix_t<S>(V(3, 2), V(0.5, 0), V(0.5, 0), V(1.4, 0.1));
    // keep the test's un-normalised 0.2/0.8 weights for the posterior check: prepare only cov
    enoki::vectorize(VECTORIZE_WRAP_MEMBER(prepare_cov), d);
    V pdf(0);
    enoki::set_slices(pdf, 2);
    {
        sdmm::embedded_s_t<S> point({1, 2});
        enoki::vectorize(VECTORIZE_WRAP_MEMBER(pdf_gaussian), d, point, pdf);
        out[0] = pdf.coeff(0); out[1] = pdf.coeff(1);
    }
    {
        sdmm::embedded_s_t<S> point({0, 0});
        enoki::vectorize(VECTORIZE_WRAP_MEMBER(pdf_gaussian), d, point, pdf);
        out[2] = pdf.coeff(0); out[3] = pdf.coeff(1);
    }
    {
        V post(0);
        enoki::set_slices(post, 2);
        sdmm::embedded_t<S> tangents;
        enoki::set_slices(tangents, 2);
        sdmm::embedded_s_t<S> point({1, 2});
        enoki::vectorize_safe(VECTORIZE_WRAP_MEMBER(posterior), d, point, post, tangents);
        out[4] = post.coeff(0); out[5] = post.coeff(1);
    }
}

// sdmm/tests/distributions.cpp:251-332 "Conditioner": 4-D Euclidean joint, 2|2 split, x=(1,2).
// out[0..7]  = conditional cov (r,c) row-major for component 0 then component 1
// out[8..11] = conditional mean (component 0: x,y; component 1: x,y)
void sdmm_ref_kat_conditioner(float* out) {
    using P = enoki::Packet<float, 2>;
    using V = enoki::DynamicArray<P>;
    using JTS = sdmm::EuclidianTangentSpace<sdmm::Vector<V, 4>, sdmm::Vector<V, 4>>;
    using J = sdmm::SDMM<sdmm::Matrix<V, 4>, JTS>;
    using MTS = sdmm::EuclidianTangentSpace<sdmm::Vector<V, 2>, sdmm::Vector<V, 2>>;
    using M = sdmm::SDMM<sdmm::Matrix<V, 2>, MTS>;
    using C = sdmm::SDMM<sdmm::Matrix<V, 2>, MTS>;
    using Cond = sdmm::SDMMConditioner<J, M, C>;
    J d;
    enoki::set_slices(d, 2);
    d.weight.pmf = V(0.2, 0.8);
    d.tangent_space.set_mean(sdmm::embedded_t<J>(V(0, 0), V(1, 1), V(2, 2), V(-1, 1)));
    d.cov = sdmm::matrix_t<J>(
        V(3, 2), V(0.5, 0), V(0.1, 0.1), V(0.4, 0.4),
        V(0.5, 0), V(1.4, 0.5), V(0, 0), V(0.2, 0.2),
        V(0.1, 0.1), V(0, 0), V(1, 1), V(0.2, 0.2),
        V(0.4, 0.4), V(0.2, 0.2), V(0.2, 0.2), V(3, 3));
    bool ok = sdmm::prepare_vectorized(d);
    (void)ok;
    Cond conditioner;
    enoki::set_slices(conditioner, enoki::slices(d));
    sdmm::prepare(conditioner, d);
    C result;
    enoki::set_slices(result, 2);
    sdmm::embedded_s_t<M> point({1, 2});
    sdmm::create_conditional(conditioner, point, result);
    for (int k = 0; k < 2; ++k)
        for (int r = 0; r < 2; ++r)
            for (int c = 0; c < 2; ++c) out[4 * k + 2 * r + c] = conditioner.conditional.cov(r, c).coeff(k);
    for (int k = 0; k < 2; ++k)
        for (int r = 0; r < 2; ++r) out[8 + 2 * k + r] = result.tangent_space.mean.coeff(r).coeff(k);
}

// sdmm/tests/linalg.cpp Cholesky cases: factor a DxD SPD matrix (row-major in) with
// sdmm::linalg::cholesky; out = L row-major; returns is_psd.
int sdmm_ref_kat_cholesky3(const float* in, float* out) {
    using V = float;
    sdmm::Matrix<V, 3> m, l;
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) m(r, c) = in[3 * r + c];
    bool psd = true;
    sdmm::linalg::cholesky(m, l, psd);
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) out[3 * r + c] = l(r, c);
    return psd ? 1 : 0;
}

// sdmm/tests/distributions.cpp:354-363 "Categorical::prepare()": pmf -> normalised pmf/cdf.
int sdmm_ref_kat_categorical(int n, const float* pmf_in, float* pmf_out, float* cdf_out) {
    using P = enoki::Packet<float, 2>;
    using V = enoki::DynamicArray<P>;
    sdmm::Categorical<V> cat;
    enoki::set_slices(cat, n);
    for (int i = 0; i < n; ++i) cat.pmf.coeff(i) = pmf_in[i];
    bool ok = cat.prepare();
    for (int i = 0; i < n; ++i) { pmf_out[i] = cat.pmf.coeff(i); cdf_out[i] = cat.cdf.coeff(i); }
    return ok ? 1 : 0;
}

// Orthonormal frame (sdmm/tests/linalg.cpp:142-159): out = to (row-major 3x3) for normal n.
void sdmm_ref_kat_frame(const float* n, float* out) {
    sdmm::linalg::CoordinateSystem<sdmm::Vector<float, 3>> cs;
    cs.prepare(sdmm::Vector<float, 3>(n[0], n[1], n[2]));
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) out[3 * r + c] = cs.to(r, c);
}

// Tangent-space maps on single points (no stored vector in the reference; used to pin the C
// restatement against the reference itself).  mean6/point6: (x,y,z,dx,dy,dz).
void sdmm_ref_joint_to(const float* mean6, const float* point6, float* tangent5, float* inv_jac) {
    using TS = sdmm::SpatioDirectionalTangentSpace<sdmm::Vector<float, 6>, sdmm::Vector<float, 5>>;
    TS ts;
    ts.set_mean(sdmm::Vector<float, 6>(mean6[0], mean6[1], mean6[2], mean6[3], mean6[4], mean6[5]));
    float ij;
    auto t = ts.to(sdmm::Vector<float, 6>(point6[0], point6[1], point6[2], point6[3], point6[4], point6[5]), ij);
    for (int i = 0; i < 5; ++i) tangent5[i] = t.coeff(i);
    *inv_jac = ij;
}

}  // extern "C"

// ---- spatial cell structure: the reference's own STree (accelerators/spatial_tree.h) ---------------
// create -> push points (find leaf; stats + data like push_back_data, sdmm_proc.cpp:880-910) ->
// split_to_depth / split_leaf_recurse -> export the node table -> find.
namespace {
using TreeRNG = enoki::PCG32<float, 1>;
using TreeContext = sdmm::SDMMContext<JointSDMM, Conditioner, TreeRNG>;
using Tree = sdmm::accelerators::STree<float, 3, TreeContext>;
struct RefTree {
    std::unique_ptr<Tree> tree;
    size_t capacity;
};
}  // namespace

extern "C" {

void* sdmm_ref_tree_create(const float* aabb_min, const float* aabb_max, long capacity) {
    auto* t = new RefTree();
    t->capacity = (size_t)capacity;
    Tree::AABB aabb(Tree::Point(aabb_min[0], aabb_min[1], aabb_min[2]), Tree::Point(aabb_max[0], aabb_max[1], aabb_max[2]));
    t->tree = std::make_unique<Tree>(aabb, std::make_unique<TreeContext>((size_t)capacity));
    return t;
}
void sdmm_ref_tree_destroy(void* h) { delete static_cast<RefTree*>(h); }
void sdmm_ref_tree_split_to_depth(void* h, int depth) { static_cast<RefTree*>(h)->tree->split_to_depth(depth); }

// returns the number of points for which no leaf was found (the renderer throws there)
long sdmm_ref_tree_push(void* h, long n, const float* pos3) {
    auto& tree = *static_cast<RefTree*>(h)->tree;
    long missing = 0;
    for (long i = 0; i < n; ++i) {
        Tree::Point key(pos3[i], pos3[n + i], pos3[2 * n + i]);
        auto* ctx = tree.find(key);
        if (!ctx) { ++missing; continue; }
        sdmm::embedded_s_t<JointSDMM> point(pos3[i], pos3[n + i], pos3[2 * n + i], 0.f, 0.f, 1.f);
        ctx->data.push_back(point, sdmm::Vector<float, 3>(0.f, 0.f, 1.f), 1.f);
        ctx->stats.push_back(sdmm::SpatialStats::Point((double)pos3[i], (double)pos3[n + i], (double)pos3[2 * n + i]));
    }
    return missing;
}

void sdmm_ref_tree_split_leaves(void* h, long threshold) {
    auto& tree = *static_cast<RefTree*>(h)->tree;
    const int size = (int)tree.size();
    for (int node_i = 0; node_i < size; ++node_i) tree.split_leaf_recurse((uint32_t)node_i, (size_t)threshold);
}

int sdmm_ref_tree_size(void* h) { return (int)static_cast<RefTree*>(h)->tree->size(); }

// per node 11 values: is_leaf, axis, child0, child1, stats.size (leaf) , aabb.min[3], aabb.max[3]
void sdmm_ref_tree_export(void* h, float* out) {
    auto& tree = *static_cast<RefTree*>(h)->tree;
    const auto& nodes = tree.data();
    for (int i = 0; i < (int)tree.size(); ++i) {
        float* o = out + 11 * (size_t)i;
        o[0] = nodes[i].is_leaf ? 1.f : 0.f;
        o[1] = (float)nodes[i].axis;
        o[2] = (float)nodes[i].children[0];
        o[3] = (float)nodes[i].children[1];
        o[4] = (nodes[i].is_leaf && nodes[i].value) ? (float)nodes[i].value->stats.size : 0.f;
        for (int d = 0; d < 3; ++d) { o[5 + d] = nodes[i].aabb.min.coeff(d); o[8 + d] = nodes[i].aabb.max.coeff(d); }
    }
}

// node index of the leaf STree::find returns (-1 if none)
void sdmm_ref_tree_find(void* h, long n, const float* pos3, int32_t* out_node) {
    auto& tree = *static_cast<RefTree*>(h)->tree;
    const auto& nodes = tree.data();
    for (long i = 0; i < n; ++i) {
        Tree::Point key(pos3[i], pos3[n + i], pos3[2 * n + i]);
        auto* ctx = tree.find(key);
        int found = -1;
        if (ctx)
            for (int j = 0; j < (int)tree.size(); ++j)
                if (nodes[j].is_leaf && nodes[j].value.get() == ctx) { found = j; break; }
        out_node[i] = found;
    }
}

#ifdef SDMM_REF_WITH_JSON
// (compiled only by ref_checkpoint.cpp: needs nlohmann_json >= 3.9 for NLOHMANN_DEFINE_TYPE_INTRUSIVE; the json.hpp
// vendored with the reference's legacy jmm tree is 3.7.3 and the macro is stubbed out for the other drivers)
// ---- .asdmm checkpoints: the reference's own save_json of the accelerator (volpath_sdmm.cpp:120-130,
// core/utils.h:130-147) and the from_json NLOHMANN_DEFINE_TYPE_INTRUSIVE generates (spatial_tree.h:314).
// STree has no default constructor, so loading goes into an existing tree (m_nodes and m_aabb are
// replaced; the reference does not restore m_nodes_end, callers walk from the root).
int sdmm_ref_tree_save_json(void* h, const char* path) {
    // save_json's body (core/utils.h:130-137) with the two-member object NLOHMANN_DEFINE_TYPE_INTRUSIVE(STree,
    // m_nodes, m_aabb) (spatial_tree.h:314) written out by hand: json 3.11's conversion traits probe
    // STree::end(), whose body does not compile (spatial_tree.h:113-115; the reference's pinned 3.9.1 never
    // looks at it).  m_aabb is private; the constructor copies it to the root node (:91).  Nodes, mixtures,
    // tangent spaces, matrices and boxes go through the reference's own serializers.
    try {
        auto& tree = *static_cast<RefTree*>(h)->tree;
        nlohmann::json j;
        j["m_nodes"] = tree.data();
        j["m_aabb"] = tree.data()[0].aabb;
        std::ofstream file(path);
        file << std::setw(4) << j << std::endl;
    } catch (...) { return -1; }
    return 0;
}

int sdmm_ref_tree_load_json(void* h, const char* path) {
    try {
        auto& tree = *static_cast<RefTree*>(h)->tree;
        std::ifstream file(path);
        nlohmann::json j;
        file >> j;
        auto nodes = j.at("m_nodes").get<std::vector<Tree::Node>>();
        (void)j.at("m_aabb").get<Tree::AABB>();
        const_cast<std::vector<Tree::Node>&>(tree.data()) = std::move(nodes);
    } catch (...) { return -1; }
    return 0;
}

// Install / read back a leaf's joint mixture (planes: pmf, mean[6], cov[25] row-major, K floats each).
int sdmm_ref_tree_set_leaf_model(void* h, int node, int K, const float* joint) {
    auto& nodes = const_cast<std::vector<Tree::Node>&>(static_cast<RefTree*>(h)->tree->data());
    if (node < 0 || node >= (int)nodes.size() || !nodes[node].value) return -1;
    auto& m = nodes[node].value->sdmm;
    m = enoki::zero<JointSDMM>(K);
    enoki::set_slices(m, K);
    for (int k = 0; k < K; ++k) m.weight.pmf.coeff(k) = joint[k];
    sdmm::embedded_t<JointSDMM> mean;
    enoki::set_slices(mean, K);
    for (int d = 0; d < 6; ++d)
        for (int k = 0; k < K; ++k) mean.coeff(d).coeff(k) = joint[(1 + d) * K + k];
    m.tangent_space.set_mean(mean);
    for (int r = 0; r < 5; ++r)
        for (int col = 0; col < 5; ++col)
            for (int k = 0; k < K; ++k) m.cov(r, col).coeff(k) = joint[(7 + 5 * r + col) * K + k];
    return sdmm::prepare_vectorized(m) ? 1 : 0;
}

// returns K (0: leaf without a mixture, -1: not a node with a value); joint may be null to query K.
// frame (nullable): coordinate_system.to then .from, 9 planes each, row-major.
int sdmm_ref_tree_get_leaf_model(void* h, int node, int K_cap, float* joint, float* frame) {
    const auto& nodes = static_cast<RefTree*>(h)->tree->data();
    if (node < 0 || node >= (int)nodes.size() || !nodes[node].value) return -1;
    const auto& m = nodes[node].value->sdmm;
    const int K = (int)enoki::slices(m.weight.pmf);
    if (!joint || K > K_cap) return K;
    for (int k = 0; k < K; ++k) joint[k] = m.weight.pmf.coeff(k);
    for (int d = 0; d < 6; ++d)
        for (int k = 0; k < K; ++k) joint[(1 + d) * K + k] = m.tangent_space.mean.coeff(d).coeff(k);
    for (int r = 0; r < 5; ++r)
        for (int col = 0; col < 5; ++col)
            for (int k = 0; k < K; ++k) joint[(7 + 5 * r + col) * K + k] = m.cov(r, col).coeff(k);
    if (frame)
        for (int r = 0; r < 3; ++r)
            for (int col = 0; col < 3; ++col)
                for (int k = 0; k < K; ++k) {
                    frame[(3 * r + col) * K + k] = m.tangent_space.coordinate_system.to(r, col).coeff(k);
                    frame[(9 + 3 * r + col) * K + k] = m.tangent_space.coordinate_system.from(r, col).coeff(k);
                }
    return K;
}

// all stored nodes (the reference keeps 10000): per node 13 values: is_leaf, axis, child0, child1, has value,
// aabb.min[3], aabb.max[3], idx, depth
int sdmm_ref_tree_stored_nodes(void* h) { return (int)static_cast<RefTree*>(h)->tree->data().size(); }
void sdmm_ref_tree_export_stored(void* h, int first, int n, double* out) {
    const auto& nodes = static_cast<RefTree*>(h)->tree->data();
    for (int i = 0; i < n; ++i) {
        const auto& nd = nodes[first + i];
        double* o = out + 13 * (size_t)i;
        o[0] = nd.is_leaf ? 1 : 0;
        o[1] = nd.axis;
        o[2] = nd.children[0];
        o[3] = nd.children[1];
        o[4] = nd.value ? 1 : 0;
        for (int d = 0; d < 3; ++d) { o[5 + d] = nd.aabb.min.coeff(d); o[8 + d] = nd.aabb.max.coeff(d); }
        o[11] = nd.idx;
        o[12] = nd.depth;
    }
}
// STree::find on a LOADED tree (size() is not restored by from_json): leaf's node index over all stored nodes
void sdmm_ref_tree_find_stored(void* h, long n, const float* pos3, int32_t* out_node) {
    auto& tree = *static_cast<RefTree*>(h)->tree;
    const auto& nodes = tree.data();
    for (long i = 0; i < n; ++i) {
        Tree::Point key(pos3[i], pos3[n + i], pos3[2 * n + i]);
        auto* ctx = tree.find(key);
        int found = -1;
        if (ctx)
            for (int j = 0; j < (int)nodes.size(); ++j)
                if (nodes[j].is_leaf && nodes[j].value.get() == ctx) { found = j; break; }
        out_node[i] = found;
    }
}
#endif  // SDMM_REF_WITH_JSON

}  // extern "C"
// libsdmm_ref_product.so (ref_product.cpp) compiles this file with the product-guiding entry points in front of the
// render loop, so that the loop can take sampleSurface's product branch; libsdmm_ref.so itself stays as pinned.
#ifdef SDMM_REF_WITH_PRODUCT
#include "ref_product.inc"
#endif
#include "ref_render.inc"
