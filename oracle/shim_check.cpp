// oracle/shim_check.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Proves the drop-in claim at the C++ level: the reference's own types (exactly the aliases of
// mitsuba/src/integrators/sdmm/sdmm_proc.h:60-96) are fitted twice from identical state and samples,
//   (a) by the reference itself: the body of optimize()'s parallelFor, volpath_sdmm.cpp:287-311
//       (sdmm::em_step x n_iterations + sdmm::prepare), on the CPU, and
//   (b) by sdmm_b200::optimize (sdmm_mitsuba_b200/csrc/host_shim.hpp -> C ABI -> CUDA kernels),
// and the two sets of reference objects are compared member by member.
// Built by oracle/Makefile into oracle/_ref/shim_check (it includes the reference headers, so the
// binary lives with the other reference-derived artefacts); run by tests/test_gpu_shim.py on the GPU box.
#include <cmath>
#include <cstdio>
#include <memory>
#include <random>
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
#include "sdmm/distributions/sdmm_context.h"
#include "sdmm/opt/data.h"
#include "sdmm/opt/em.h"
#include "sdmm/spaces/euclidian.h"
#include "sdmm/spaces/offset_directional.h"
#include "sdmm/spaces/spatio_directional.h"

#include "../sdmm_mitsuba_b200/csrc/host_shim.hpp"

namespace {
constexpr size_t PacketSize = 8;
using Packet = enoki::Packet<float, PacketSize>;
using Value = enoki::DynamicArray<Packet>;
using JointTangentSpace = sdmm::SpatioDirectionalTangentSpace<sdmm::Vector<Value, 6>, sdmm::Vector<Value, 5>>;
using JointSDMM = sdmm::SDMM<sdmm::Matrix<Value, 5>, JointTangentSpace>;
using MarginalTangentSpace = sdmm::EuclidianTangentSpace<sdmm::Vector<Value, 3>, sdmm::Vector<Value, 3>>;
using MarginalSDMM = sdmm::SDMM<sdmm::Matrix<Value, 3>, MarginalTangentSpace>;
using ConditionalTangentSpace = sdmm::OffsetDirectionalTangentSpace<sdmm::Vector<Value, 3>, sdmm::Vector<Value, 2>>;
using ConditionalSDMM = sdmm::SDMM<sdmm::Matrix<Value, 2>, ConditionalTangentSpace>;
using Conditioner = sdmm::SDMMConditioner<JointSDMM, MarginalSDMM, ConditionalSDMM>;
using RNG = enoki::PCG32<float, 1>;
using Context = sdmm::SDMMContext<JointSDMM, Conditioner, RNG>;

// a fresh cell: K random components, priors as the integrator sets them (opt/init.h:178-192)
std::unique_ptr<Context> make_cell(int K, int N, int iterations_run, std::mt19937& gen) {
    auto c = std::make_unique<Context>((size_t)N);
    std::uniform_real_distribution<float> U(0.f, 1.f);
    std::normal_distribution<float> G(0.f, 1.f);
    c->sdmm = enoki::zero<JointSDMM>(K);
    enoki::set_slices(c->sdmm, K);
    c->em = enoki::zero<sdmm::EM<JointSDMM>>(K);
    enoki::set_slices(c->em, K);
    sdmm::embedded_t<JointSDMM> mean;
    enoki::set_slices(mean, K);
    for (int k = 0; k < K; ++k) {
        c->sdmm.weight.pmf.coeff(k) = 1.f / K;
        float d[3] = {G(gen), G(gen), G(gen)};
        float len = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        for (int i = 0; i < 3; ++i) mean.coeff(i).coeff(k) = U(gen);
        for (int i = 0; i < 3; ++i) mean.coeff(3 + i).coeff(k) = d[i] / len;
        for (int r = 0; r < 5; ++r)
            for (int col = 0; col < 5; ++col) {
                c->sdmm.cov(r, col).coeff(k) = (r == col) ? 0.05f : 0.f;
                c->em.cov_prior(r, col).coeff(k) = (r == col) ? (r < 3 ? 2e-3f : 2e-4f) : 0.f;
            }
    }
    c->sdmm.tangent_space.set_mean(mean);
    sdmm::prepare_vectorized(c->sdmm);
    c->em.weight_prior = 1.f / K;
    c->em.cov_prior_strength = 5.f / K;
    c->em.total_weight = 0.f;
    c->em.iterations_run = iterations_run;
    // samples: 4 lobes, directions on the sphere, log-normal weights
    float centre[4][3], dir[4][3];
    for (auto& l : centre) for (float& v : l) v = 0.1f + 0.8f * U(gen);
    for (auto& l : dir) {
        float len = 0.f;
        for (float& v : l) { v = G(gen); len += v * v; }
        for (float& v : l) v /= std::sqrt(len);
    }
    for (int n = 0; n < N; ++n) {
        int l = (int)(U(gen) * 4) & 3;
        float p[6], len = 0.f;
        for (int i = 0; i < 3; ++i) p[i] = std::min(1.f, std::max(0.f, centre[l][i] + 0.08f * G(gen)));
        for (int i = 0; i < 3; ++i) { p[3 + i] = dir[l][i] + 0.25f * G(gen); len += p[3 + i] * p[3 + i]; }
        for (int i = 0; i < 3; ++i) p[3 + i] /= std::sqrt(len);
        float w = std::exp(G(gen));
        sdmm::embedded_s_t<JointSDMM> point(p[0], p[1], p[2], p[3], p[4], p[5]);
        sdmm::Vector<float, 3> normal(0.f, 0.f, 1.f);
        c->training_data.push_back(point, normal, w);
    }
    return c;
}

double max_rel(const Value& a, const Value& b, int K, double floor_) {
    double scale = floor_, err = 0;
    for (int k = 0; k < K; ++k) scale = std::max(scale, (double)std::abs(b.coeff(k)));
    for (int k = 0; k < K; ++k) err = std::max(err, (double)std::abs(a.coeff(k) - b.coeff(k)));
    return err / scale;
}
}  // namespace

int main(int argc, char** argv) {
    const int K = argc > 1 ? atoi(argv[1]) : 16;
    const int C = argc > 2 ? atoi(argv[2]) : 6;
    const int N = argc > 3 ? atoi(argv[3]) : 700;
    std::vector<std::unique_ptr<Context>> ref, gpu;
    for (int c = 0; c < C; ++c) {
        // odd cells have already run 5 iterations -> 1 EM step, even cells 2 (volpath_sdmm.cpp:299-302)
        std::mt19937 g1(100 + c), g2(100 + c);
        ref.push_back(make_cell(K, N + 37 * c, (c & 1) ? 5 : 0, g1));
        gpu.push_back(make_cell(K, N + 37 * c, (c & 1) ? 5 : 0, g2));
    }
    // (a) the reference, verbatim loop body of volpath_sdmm.cpp:297-310
    for (auto& context : ref) {
        int n_iterations = 1;
        if (context->em.iterations_run < 4) n_iterations = 2;
        for (int it = 0; it < n_iterations; ++it) sdmm::em_step(context->sdmm, context->em, context->training_data);
        enoki::set_slices(context->conditioner, enoki::slices(context->sdmm));
        sdmm::prepare(context->conditioner, context->sdmm);
        context->training_data.clear();
        context->data.clear();
        context->initialized = true;
    }
    // (b) the drop-in
    try {
        sdmm_b200::Device dev(0, C, K);
        std::vector<Context*> ptrs;
        for (auto& c : gpu) ptrs.push_back(c.get());
        sdmm_b200::optimize(dev, 0, ptrs);
    } catch (const std::exception& e) {
        std::printf("FAIL exception: %s\n", e.what());
        return 2;
    }
    double worst = 0;
    bool ok = true;
    for (int c = 0; c < C; ++c) {
        auto &a = *gpu[c], &b = *ref[c];
        double e_pmf = max_rel(a.sdmm.weight.pmf, b.sdmm.weight.pmf, K, 1e-12);
        double e_cdf = max_rel(a.sdmm.weight.cdf, b.sdmm.weight.cdf, K, 1e-12);
        double e_mean = 0, e_cov = 0, e_stat = 0, e_cond = 0;
        for (int i = 0; i < 6; ++i) e_mean = std::max(e_mean, max_rel(a.sdmm.tangent_space.mean.coeff(i), b.sdmm.tangent_space.mean.coeff(i), K, 1.0));
        for (int r = 0; r < 5; ++r)
            for (int col = 0; col < 5; ++col) {
                e_cov = std::max(e_cov, max_rel(a.sdmm.cov(r, col), b.sdmm.cov(r, col), K, 1e-3));
                e_stat = std::max(e_stat, max_rel(a.em.stats.cov(r, col), b.em.stats.cov(r, col), K, 1e-3));
            }
        for (int r = 0; r < 2; ++r)
            for (int col = 0; col < 3; ++col)
                e_cond = std::max(e_cond, max_rel(a.conditioner.mean_transform(r, col), b.conditioner.mean_transform(r, col), K, 1.0));
        bool same_counts = a.em.iterations_run == b.em.iterations_run && a.em.samples_seen == b.em.samples_seen &&
                           a.initialized && a.training_data.size == 0;
        double e = std::max({e_pmf, e_cdf, e_mean, e_cov, e_stat, e_cond});
        worst = std::max(worst, e);
        std::printf("cell %d: iterations_run %d/%d samples_seen %d/%d  pmf %.2e cdf %.2e mean %.2e cov %.2e stats %.2e cond %.2e\n",
                    c, a.em.iterations_run, b.em.iterations_run, a.em.samples_seen, b.em.samples_seen, e_pmf, e_cdf,
                    e_mean, e_cov, e_stat, e_cond);
        ok = ok && same_counts && e <= 1e-3;
    }
    std::printf("%s worst %.3e (tolerance 1e-3)\n", ok ? "PASS" : "FAIL", worst);
    return ok ? 0 : 1;
}
