// capi_fit.cu -- the training half of the C ABI: sample binding and E-step work items, sdmmb_em_step (in order or
// with the M-steps hidden behind the other half's E-steps), the host-buffer pipeline sdmmb_optimize_host, the
// conditioner prepare.  Reference: the parallelFor body of SDMMVolumetricPathTracer::optimize
// (mitsuba/src/integrators/sdmm/volpath_sdmm.cpp:287-311) and sdmm::em_step (sdmm/opt/em.h:218-253).
#include "capi_internal.h"

using namespace sdmmb;
using namespace sdmmb_capi;

namespace sdmmb_capi {

// Build the E-step work items for the bound cells: each cell's samples are cut into runs of at
// most ctx->item_samples (a multiple of 32) so that (a) huge cells (the renderer caps a cell at
// 2 000 000 samples, volpath_sdmm.cpp:371) spread over many CTAs and (b) the grid is fine-grained
// enough for the hardware CTA scheduler to balance 148 SMs.
int build_items(sdmmb_ctx* ctx, int first_cell, int n_cells, const uint32_t* h_off, cudaStream_t st) {
    const uint32_t target = ctx->item_samples;
    // same layout as the previous binding (a progressive loop with a fixed budget per cell, the benchmark's
    // steps): the item table on the device is still valid -- skip the rebuild, the upload and its host sync
    if (ctx->items_valid && ctx->items_first == first_cell && ctx->items_target == target &&
        ctx->items_off.size() == (size_t)n_cells + 1) {
        bool same = true;
        const uint32_t base0 = h_off[0];
        for (int c = 0; c <= n_cells && same; ++c) same = (ctx->items_off[c] == h_off[c] - base0);
        if (same) return SDMMB_OK;
    }
    ctx->items_valid = false;
    std::vector<EItem> items;
    std::vector<uint32_t> item_begin((size_t)n_cells + 1), counts((size_t)n_cells);
    for (int c = 0; c < n_cells; ++c) {
        item_begin[c] = (uint32_t)items.size();
        const uint32_t b = h_off[c] - h_off[0], e = h_off[c + 1] - h_off[0];
        counts[c] = e - b;
        const uint32_t n = e - b;
        if (n == 0) continue;
        const uint32_t n_it = (n + target - 1) / target;
        // equal split rounded up to whole 32-sample chunks
        uint32_t per = ((n + n_it - 1) / n_it + 31u) & ~31u;
        for (uint32_t s = b; s < e; s += per)
            items.push_back(EItem{(uint32_t)(first_cell + c), s, std::min(e, s + per), 0u});
    }
    item_begin[n_cells] = (uint32_t)items.size();
    ctx->n_items = items.size();
    ctx->h_item_begin = item_begin;
    if (items.size() > ctx->items_cap) {
        cudaFree(ctx->d_items); cudaFree(ctx->d_partials); cudaFree(ctx->d_item_scal);
        ctx->d_items = nullptr; ctx->d_partials = nullptr; ctx->d_item_scal = nullptr;
        ctx->items_cap = 0;
        const size_t cap = items.size() + items.size() / 8 + 16;
        CK(cudaMalloc(&ctx->d_items, cap * sizeof(EItem)));
        CK(cudaMalloc(&ctx->d_partials, cap * ESTAT_PLANES * ctx->K * 4));
        CK(cudaMalloc(&ctx->d_item_scal, cap * 2 * 4));
        ctx->items_cap = cap;
    }
    if ((size_t)n_cells + 1 > ctx->cells_cap) {
        cudaFree(ctx->d_cell_item_begin); cudaFree(ctx->d_cell_sample_count);
        ctx->d_cell_item_begin = nullptr; ctx->d_cell_sample_count = nullptr;
        ctx->cells_cap = 0;
        CK(cudaMalloc(&ctx->d_cell_item_begin, ((size_t)n_cells + 1) * 4));
        CK(cudaMalloc(&ctx->d_cell_sample_count, ((size_t)n_cells + 1) * 4));
        ctx->cells_cap = (size_t)n_cells + 1;
    }
    if (!items.empty())
        CK(cudaMemcpyAsync(ctx->d_items, items.data(), items.size() * sizeof(EItem), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(ctx->d_cell_item_begin, item_begin.data(), item_begin.size() * 4, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(ctx->d_cell_sample_count, counts.data(), counts.size() * 4, cudaMemcpyHostToDevice, st));
    CK(cudaStreamSynchronize(st));  // host vectors die at return
    ctx->items_off.resize((size_t)n_cells + 1);
    for (int c = 0; c <= n_cells; ++c) ctx->items_off[c] = h_off[c] - h_off[0];
    ctx->items_first = first_cell;
    ctx->items_target = target;
    ctx->items_valid = true;
    return SDMMB_OK;
}

// Reserve the packed-sample array and build the work items for a cell range (no packing yet).
int bind_layout(sdmmb_ctx* ctx, int first_cell, int n_cells, const uint32_t* h_off, cudaStream_t st, bool reserve) {
    const size_t n = (size_t)h_off[n_cells] - h_off[0];
    if (n > 0x7fffffffu) return SDMMB_E_DATA_FULL;
    ctx->soa_point6 = nullptr;   // every binding that packs drops a borrowed one
    ctx->soa_weight = nullptr;
    if (reserve && n > ctx->samples_cap) {
        cudaFree(ctx->d_samples);
        ctx->d_samples = nullptr;
        ctx->samples_cap = 0;
        CK(cudaMalloc(&ctx->d_samples, n * 32));
        ctx->samples_cap = n;
    }
    ctx->n_samples = n;
    ctx->bound_first = first_cell;
    ctx->bound_cells = n_cells;
    return build_items(ctx, first_cell, n_cells, h_off, st);
}

int bind_common(sdmmb_ctx* ctx, int first_cell, int n_cells, const float* d_point6, size_t stride,
                const float* d_weight, const uint32_t* h_off, cudaStream_t st) {
    SDMMB_NVTX("sdmmb bind samples (K0)");
    int rc = bind_layout(ctx, first_cell, n_cells, h_off, st);
    if (rc) return rc;
    const size_t n = ctx->n_samples;
    CK(launch_pack(d_point6 + h_off[0], stride, d_weight + h_off[0], ctx->d_samples, n, st));
    if (n) ctx->launches += 1;
    return SDMMB_OK;
}

// Peers attached (sdmmb_peer_attach) and records written in place (not the async back copy): open the
// exchange of one record-writing call.  EVERY such call consumes exactly TWO epochs on every rank, whatever
// schedule it ends up using: the overlapped schedule pushes its first half with epoch base + 1 and its second
// with base + 2, the in-order schedule (and every fallback to it: one cell, one item, all items in one cell)
// pushes once with base + 2.  Flags only grow and are compared with >=, so one ready signal of base + 2 opens
// both pushes of a peer and one arrival of base + 2 satisfies its wait -- ranks whose shards make them choose
// different schedules stay in step.  out[0] = first half (overlapped only), out[1] = the last / only push.
int arm_peer_push(sdmmb_ctx* ctx, PeerPush out[2], cudaStream_t st) {
    out[0].world = out[1].world = 0;
    if (ctx->peer.world <= 1 || ctx->d_qrec_back) return SDMMB_OK;
    const uint32_t base = ctx->peer_epoch;
    ctx->peer_epoch = base + 2;
    out[0] = out[1] = ctx->peer;
    out[0].epoch = base + 1;
    out[1].epoch = base + 2;
    CK(launch_peer_signal(out[1], st));
    ctx->launches += 1;
    return SDMMB_OK;
}

// n_iters x (E-step + M-step) on cells [rel, rel + n_cells) of the bound range; the conditioner
// records are written by the last M-step when with_conditioner is set.
int run_em(sdmmb_ctx* ctx, int rel, int n_cells, int n_iters, int with_conditioner, float* d_out_loglik,
           bool timed, cudaStream_t st, bool peer_push = false) {
    const int first_cell = ctx->bound_first + rel;
    const uint32_t it0 = ctx->h_item_begin[rel], it1 = ctx->h_item_begin[rel + n_cells];
    // peers attached: tell them now (stream-ordered after this rank's queries of the previous step)
    // that this rank's records may be overwritten; the last M-step then stores into theirs
    PeerPush push2[2] = {};
    if (peer_push && with_conditioner && n_iters != 0) {
        int rc = arm_peer_push(ctx, push2, st);
        if (rc) return rc;
    }
    const PeerPush& push = push2[1];
    // n_iters == SDMMB_ITERS_INTEGRATOR: the integrator's rule, two passes of which the second skips (per cell, on
    // the device) the cells whose em.iterations_run was already >= 4 (volpath_sdmm.cpp:299-302)
    const bool rule = (n_iters == SDMMB_ITERS_INTEGRATOR);
    if (rule) n_iters = 2;
    for (int it = 0; it < n_iters; ++it) {
        EStepLaunch L{};
        L.K = ctx->K;
        L.scal = ctx->d_scal;
        L.skip_iterations_ge = (rule && it == 1) ? 5u : 0u;
        L.n_items = (int)(it1 - it0);
        L.samples = ctx->d_samples;
        L.soa_point6 = ctx->soa_point6;
        L.soa_stride = ctx->soa_stride;
        L.soa_weight = ctx->soa_weight;
        L.items = ctx->d_items + it0;
        L.erec = ctx->d_erec;
        L.partials = ctx->d_partials + (size_t)it0 * ESTAT_PLANES * ctx->K;
        L.item_scal = ctx->d_item_scal + (size_t)it0 * 2;
        L.out_loglik = d_out_loglik ? d_out_loglik + (size_t)it * ctx->n_samples : nullptr;
        if (timed) CK(cudaEventRecord(ctx->ev[2 * it], st));
        {
            SDMMB_NVTX("sdmmb E-step (K1)");
            CK(launch_estep(L, st));
        }
        if (timed) CK(cudaEventRecord(ctx->ev[2 * it + 1], st));
        if (L.n_items) ctx->launches += 1;
        const bool last = (it == n_iters - 1);
        MStepArgs a = base_mstep_args(ctx, first_cell, n_cells, 1 | ((last && with_conditioner) ? 2 : 0));
        // item indices inside the kernel are relative to L.partials
        a.partials = L.partials;
        a.item_scal = L.item_scal;
        a.cell_item_begin = ctx->d_cell_item_begin + rel;
        a.cell_sample_count = ctx->d_cell_sample_count + rel;
        a.item_base = it0;
        a.skip_iterations_ge = L.skip_iterations_ge;
        if (last && with_conditioner) a.peer = push;
        {
            SDMMB_NVTX((last && with_conditioner) ? (push.world > 1 ? "sdmmb M-step + conditioner + peer push (K2+K3+K5a)" : "sdmmb M-step + conditioner (K2+K3)") : "sdmmb M-step (K2)");
            CK(launch_mstep(a, st));
        }
        ctx->launches += 1;
        if (timed) ctx->ev_used = it + 1;
    }
    return SDMMB_OK;
}

// The same work with the M-steps hidden: the bound range is cut into two halves (by E-step items), the halves'
// E-steps alternate on two streams and every M-step runs on a third, high-priority stream, so M(half h,
// iteration i) executes while the E-step of the other half occupies the SMs (the M-step is latency bound: 8
// warps per SM waiting on DRAM, 4 % of a training step when serialised).  Dependencies per half:
// E(h, i) -> M(h, i) -> E(h, i + 1), carried by events; the caller's stream forks into the three and joins
// them at the end, so the call stays stream-ordered for the caller.  All M-steps (and with them every
// peer-memory push and its flags) stay in order on ONE stream.  Results are bit-identical to run_em.
int run_em_overlapped(sdmmb_ctx* ctx, int rel, int n_cells, int n_iters, int with_conditioner,
                      float* d_out_loglik, cudaStream_t st, bool* done) {
    *done = false;
    const uint32_t it0 = ctx->h_item_begin[rel], it1 = ctx->h_item_begin[rel + n_cells];
    if (n_cells < 2 || it1 - it0 < 2 || n_iters <= 0) return SDMMB_OK;
    // Parts of (nearly) equal item counts, cut at cell boundaries: 2 without peers (enough to hide the M-steps),
    // 4 with peers attached -- the record push of a part rides in its last M-step, so only the LAST part's push
    // (a quarter of the records) is exposed after the last E-step has drained.  SDMMB_OVERLAP_PARTS overrides.
    constexpr int MAXP = 8;
    static const int env_parts = [] { const char* v = getenv("SDMMB_OVERLAP_PARTS"); return v ? atoi(v) : 0; }();
    int want = env_parts > 0 ? env_parts : ((ctx->peer.world > 1 && !ctx->d_qrec_back) ? 4 : 2);
    want = std::max(2, std::min(MAXP, want));
    int cut[MAXP + 1];
    int P = 0;
    cut[0] = rel;
    for (int p = 1; p < want; ++p) {
        const uint32_t target = it0 + (uint32_t)(((uint64_t)(it1 - it0) * p) / want);
        int c = cut[P] + 1;
        while (c < rel + n_cells - 1 && ctx->h_item_begin[c] < target) ++c;
        if (c >= rel + n_cells) break;
        if (ctx->h_item_begin[c] == ctx->h_item_begin[cut[P]] || ctx->h_item_begin[c] == it1) continue;   // empty part
        cut[++P] = c;
    }
    cut[++P] = rel + n_cells;
    if (P < 2) return SDMMB_OK;
    if (!ctx->ov_mstream) {
        int lo = 0, hi = 0;
        CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));   // hi = numerically lowest = greatest priority
        for (int h = 0; h < 2; ++h) CK(cudaStreamCreateWithPriority(&ctx->ov_stream[h], cudaStreamNonBlocking, lo));
        CK(cudaStreamCreateWithPriority(&ctx->ov_mstream, cudaStreamNonBlocking, hi));
        for (cudaEvent_t& e : ctx->ov_ev) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    cudaEvent_t fork = ctx->ov_ev[0];
    cudaEvent_t* ev_e = ctx->ov_ev + 1;          // [MAXP] E-step of part p done
    cudaEvent_t* ev_m = ctx->ov_ev + 1 + MAXP;   // [MAXP] M-step of part p done
    CK(cudaEventRecord(fork, st));
    for (int h = 0; h < 2; ++h) CK(cudaStreamWaitEvent(ctx->ov_stream[h], fork, 0));
    CK(cudaStreamWaitEvent(ctx->ov_mstream, fork, 0));
    PeerPush push[2] = {};
    if (with_conditioner) {
        int rc = arm_peer_push(ctx, push, ctx->ov_mstream);
        if (rc) return rc;
    }
    for (int it = 0; it < n_iters; ++it) {
        const bool last = (it == n_iters - 1);
        for (int p = 0; p < P; ++p) {
            cudaStream_t es = ctx->ov_stream[p & 1];
            const int p_rel = cut[p], p_n = cut[p + 1] - cut[p];
            const uint32_t b = ctx->h_item_begin[p_rel], e = ctx->h_item_begin[p_rel + p_n];
            EStepLaunch L{};
            L.K = ctx->K;
            L.n_items = (int)(e - b);
            L.samples = ctx->d_samples;
            L.soa_point6 = ctx->soa_point6;
            L.soa_stride = ctx->soa_stride;
            L.soa_weight = ctx->soa_weight;
            L.items = ctx->d_items + b;
            L.erec = ctx->d_erec;
            L.partials = ctx->d_partials + (size_t)b * ESTAT_PLANES * ctx->K;
            L.item_scal = ctx->d_item_scal + (size_t)b * 2;
            L.out_loglik = d_out_loglik ? d_out_loglik + (size_t)it * ctx->n_samples : nullptr;
            if (it > 0) CK(cudaStreamWaitEvent(es, ev_m[p], 0));   // E(p, it) needs M(p, it - 1)
            CK(launch_estep(L, es));
            if (L.n_items) ctx->launches += 1;
            CK(cudaEventRecord(ev_e[p], es));
            CK(cudaStreamWaitEvent(ctx->ov_mstream, ev_e[p], 0));
            MStepArgs a = base_mstep_args(ctx, ctx->bound_first + p_rel, p_n, 1 | ((last && with_conditioner) ? 2 : 0));
            a.partials = L.partials;
            a.item_scal = L.item_scal;
            a.cell_item_begin = ctx->d_cell_item_begin + p_rel;
            a.cell_sample_count = ctx->d_cell_sample_count + p_rel;
            a.item_base = b;
            // every part but the last pushes with the call's first epoch, the last one with its second: the
            // peers' waits are satisfied by the second only (arm_peer_push)
            if (last && with_conditioner) a.peer = push[p == P - 1 ? 1 : 0];
            CK(launch_mstep(a, ctx->ov_mstream));
            ctx->launches += 1;
            if (!last) CK(cudaEventRecord(ev_m[p], ctx->ov_mstream));
        }
    }
    // join: the M stream has waited for every part's last E-step
    CK(cudaEventRecord(ev_m[0], ctx->ov_mstream));
    CK(cudaStreamWaitEvent(st, ev_m[0], 0));
    ctx->ev_used = 0;   // per-launch E-step timing is meaningless while launches queue behind each other
    *done = true;
    return SDMMB_OK;
}

int ensure_pipe(sdmmb_ctx* ctx) {
    if (ctx->copy_stream) return SDMMB_OK;
    CK(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&ctx->d2h_stream, cudaStreamNonBlocking));
    for (auto& e : ctx->pipe_ev) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    return SDMMB_OK;
}

}  // namespace sdmmb_capi

extern "C" {

int sdmmb_set_samples(sdmmb_ctx* ctx, int first_cell, int n_cells, const float* d_point6,
                      size_t plane_stride, const float* d_weight, const uint32_t* h_cell_offsets,
                      void* stream) {
    if (!ctx || !d_point6 || !d_weight || !h_cell_offsets) return SDMMB_E_INVALID;
    if (!range_ok(ctx, first_cell, n_cells)) return SDMMB_E_RANGE;
    for (int c = 0; c < n_cells; ++c)
        if (h_cell_offsets[c + 1] < h_cell_offsets[c]) return SDMMB_E_INVALID;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    return bind_common(ctx, first_cell, n_cells, d_point6, plane_stride, d_weight, h_cell_offsets,
                       (cudaStream_t)stream);
}

int sdmmb_set_samples_borrowed(sdmmb_ctx* ctx, int first_cell, int n_cells, const float* d_point6,
                               size_t plane_stride, const float* d_weight, const uint32_t* h_cell_offsets,
                               void* stream) {
    if (!ctx || !d_point6 || !d_weight || !h_cell_offsets) return SDMMB_E_INVALID;
    if (!range_ok(ctx, first_cell, n_cells)) return SDMMB_E_RANGE;
    for (int c = 0; c < n_cells; ++c)
        if (h_cell_offsets[c + 1] < h_cell_offsets[c]) return SDMMB_E_INVALID;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    int rc = bind_layout(ctx, first_cell, n_cells, h_cell_offsets, (cudaStream_t)stream, false);
    if (rc) return rc;
    ctx->soa_point6 = d_point6 + h_cell_offsets[0];
    ctx->soa_stride = plane_stride;
    ctx->soa_weight = d_weight + h_cell_offsets[0];
    return SDMMB_OK;
}

int sdmmb_set_samples_host(sdmmb_ctx* ctx, int first_cell, int n_cells, const float* h_point6,
                           size_t plane_stride, const float* h_weight, const uint32_t* h_cell_offsets,
                           void* stream) {
    if (!ctx || !h_point6 || !h_weight || !h_cell_offsets) return SDMMB_E_INVALID;
    if (!range_ok(ctx, first_cell, n_cells)) return SDMMB_E_RANGE;
    for (int c = 0; c < n_cells; ++c)
        if (h_cell_offsets[c + 1] < h_cell_offsets[c]) return SDMMB_E_INVALID;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t b = h_cell_offsets[0];
    const size_t n = (size_t)h_cell_offsets[n_cells] - b;
    int rc = ensure(ctx, &ctx->d_stage, &ctx->stage_cap, std::max<size_t>(n, 1) * 7 * 4);
    if (rc) return rc;
    float* d = (float*)ctx->d_stage;
    for (int p = 0; p < 6; ++p)
        CK(cudaMemcpyAsync(d + (size_t)p * n, h_point6 + (size_t)p * plane_stride + b, n * 4,
                           cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d + 6 * n, h_weight + b, n * 4, cudaMemcpyHostToDevice, st));
    std::vector<uint32_t> rel((size_t)n_cells + 1);
    for (int c = 0; c <= n_cells; ++c) rel[c] = h_cell_offsets[c] - b;
    return bind_common(ctx, first_cell, n_cells, d, n, d + 6 * n, rel.data(), st);
}

int sdmmb_em_step(sdmmb_ctx* ctx, int first_cell, int n_cells, int n_iters, int with_conditioner,
                  float* d_out_loglik, void* stream) {
    if (!ctx || (n_iters < 0 && n_iters != SDMMB_ITERS_INTEGRATOR)) return SDMMB_E_INVALID;
    if (!range_ok(ctx, first_cell, n_cells)) return SDMMB_E_RANGE;
    if (first_cell < ctx->bound_first || first_cell + n_cells > ctx->bound_first + ctx->bound_cells)
        return SDMMB_E_STATE;
    if (n_cells == 0 || n_iters == 0) return SDMMB_OK;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int rel = first_cell - ctx->bound_first;
    while ((int)ctx->ev.size() < 2 * std::max(n_iters, 2)) {
        cudaEvent_t e;
        CK(cudaEventCreate(&e));
        ctx->ev.push_back(e);
    }
    ctx->ev_used = 0;
    if (ctx->em_overlap) {
        bool done = false;
        int rc = run_em_overlapped(ctx, rel, n_cells, n_iters, with_conditioner, d_out_loglik, st, &done);
        if (rc || done) return rc;
    }
    return run_em(ctx, rel, n_cells, n_iters, with_conditioner, d_out_loglik, true, st, true);
}

// The whole per-iteration training call of the integrator in one entry point, pipelined over chunks
// of cells: while chunk c is being fitted, chunk c+1's sample planes cross PCIe on a copy stream and
// chunk c-1's fitted mixtures travel back on a third one.
int sdmmb_optimize_host(sdmmb_ctx* ctx, int first_cell, int n_cells, const float* h_point6,
                        size_t plane_stride, const float* h_weight, const uint32_t* h_cell_offsets,
                        int n_iters, int with_conditioner, float* h_out_joint, float* h_out_em,
                        uint32_t* h_out_scal, void* stream) {
    if (!ctx || !h_point6 || !h_weight || !h_cell_offsets || (n_iters < 0 && n_iters != SDMMB_ITERS_INTEGRATOR))
        return SDMMB_E_INVALID;
    if (!range_ok(ctx, first_cell, n_cells)) return SDMMB_E_RANGE;
    for (int c = 0; c < n_cells; ++c)
        if (h_cell_offsets[c + 1] < h_cell_offsets[c]) return SDMMB_E_INVALID;
    if (n_cells == 0) return SDMMB_OK;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    SDMMB_NVTX("sdmmb_optimize_host");
    const int K = ctx->K;
    const bool direct_out = (ctx->K_user == K);
    {
        int prc = ensure_pipe(ctx);
        if (prc) return prc;
    }
    const uint32_t base = h_cell_offsets[0];
    std::vector<uint32_t> rel((size_t)n_cells + 1);
    for (int c = 0; c <= n_cells; ++c) rel[c] = h_cell_offsets[c] - base;
    int rc = bind_layout(ctx, first_cell, n_cells, rel.data(), st);
    if (rc) return rc;
    // Chunks of whole cells.  The E-step of a chunk runs in waves of `resident` CTAs (one per item), so a chunk
    // is cut at a whole number of waves: the first one a single wave (the fit starts as soon as its 13 MB have
    // crossed PCIe), then growing -- chunk j may hold 1 + 3/4 of everything before it, which keeps its copy
    // hidden behind the fits already queued as long as PCIe moves a sample no slower than 4/3 of the time the
    // fit spends on it -- up to the number of waves closest to chunk_samples.
    const size_t n_total = rel[n_cells];
    if (ctx->sm_count == 0) {
        int v = 0;
        CK(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, ctx->device));
        ctx->sm_count = v > 0 ? v : 148;
    }
    const size_t resident = (size_t)ctx->sm_count * estep_ctas_per_sm(K);
    const size_t waves = std::max<size_t>(1, (ctx->pipe_chunk_samples / ctx->item_samples + resident / 2) / resident);
    std::vector<int> cuts(1, 0);
    {
        const std::vector<uint32_t>& ib = ctx->h_item_begin;
        size_t w = 1, w_sum = 0;
        for (int c = 1; c <= n_cells; ++c) {
            if (c == n_cells) { cuts.push_back(c); break; }
            // would the next cell overshoot the chunk's last wave?
            if ((size_t)(ib[c + 1] - ib[cuts.back()]) > w * resident && ib[c] > ib[cuts.back()]) {
                cuts.push_back(c);
                w_sum += w;
                w = std::min(waves, 1 + (3 * w_sum) / 4);
            }
        }
    }
    // The planes go straight to their final place: the context's sample array viewed as 7 planes of n_total
    // floats, read in place by the E-step's SoA instantiation -- no staging buffer, no packing pass.
    float* planes = reinterpret_cast<float*>(ctx->d_samples);
    ctx->soa_point6 = planes;
    ctx->soa_stride = n_total;
    ctx->soa_weight = planes + 6 * n_total;
    // Consecutive chunks are fitted on alternating streams (2; SDMMB_PIPE_FIT_STREAMS: 1 -> 14.02, 2 -> 13.67, 3 -> 13.64, 4 -> 13.74 ms per cfg2 step): chunks are independent, so the M-steps and the
    // ragged end of one chunk's E-step run under the other chunk's E-step instead of idling the SMs.
    if (!ctx->pipe_fit_stream[0]) {
        for (cudaStream_t& ps : ctx->pipe_fit_stream) CK(cudaStreamCreateWithFlags(&ps, cudaStreamNonBlocking));
        if (const char* v = std::getenv("SDMMB_PIPE_FIT_STREAMS")) ctx->pipe_fit_streams = std::min(4, std::max(1, atoi(v)));
    }
    // order the new work after whatever the caller queued on `st`
    CK(cudaEventRecord(ctx->pipe_ev[4], st));
    CK(cudaStreamWaitEvent(ctx->copy_stream, ctx->pipe_ev[4], 0));
    for (cudaStream_t ps : ctx->pipe_fit_stream) CK(cudaStreamWaitEvent(ps, ctx->pipe_ev[4], 0));
    for (size_t i = 1; i < cuts.size(); ++i) {
        const int c0 = cuts[i - 1], c1 = cuts[i], buf = (int)(i & 1);
        const size_t b = rel[c0], n = (size_t)rel[c1] - b;
        cudaEvent_t copied = ctx->pipe_ev[buf];
        cudaStream_t st = ctx->pipe_fit_stream[i % ctx->pipe_fit_streams];   // shadows the caller's stream in the loop
        if (n) {
            for (int p = 0; p < 6; ++p)
                CK(cudaMemcpyAsync(planes + (size_t)p * n_total + b, h_point6 + (size_t)p * plane_stride + base + b, n * 4,
                                   cudaMemcpyHostToDevice, ctx->copy_stream));
            CK(cudaMemcpyAsync(planes + 6 * n_total + b, h_weight + base + b, n * 4, cudaMemcpyHostToDevice,
                               ctx->copy_stream));
        }
        CK(cudaEventRecord(copied, ctx->copy_stream));
        CK(cudaStreamWaitEvent(st, copied, 0));
        if (n_iters != 0) {
            rc = run_em(ctx, c0, c1 - c0, n_iters, with_conditioner, nullptr, false, st);
            if (rc) return rc;
        }
        if (direct_out && (h_out_joint || h_out_em || h_out_scal)) {
            CK(cudaEventRecord(ctx->pipe_ev[5], st));
            CK(cudaStreamWaitEvent(ctx->d2h_stream, ctx->pipe_ev[5], 0));
            const size_t cell = (size_t)first_cell + c0, nc = (size_t)(c1 - c0);
            if (h_out_joint)
                CK(cudaMemcpyAsync(h_out_joint + (size_t)c0 * JOINT_PLANES * K, ctx->d_joint + cell * JOINT_PLANES * K,
                                   nc * JOINT_PLANES * K * 4, cudaMemcpyDeviceToHost, ctx->d2h_stream));
            if (h_out_em)
                CK(cudaMemcpyAsync(h_out_em + (size_t)c0 * EM_PLANES * K, ctx->d_em + cell * EM_PLANES * K,
                                   nc * EM_PLANES * K * 4, cudaMemcpyDeviceToHost, ctx->d2h_stream));
            if (h_out_scal)
                CK(cudaMemcpyAsync(h_out_scal + (size_t)c0 * SCALAR_WORDS, ctx->d_scal + cell * SCALAR_WORDS,
                                   nc * SCALAR_WORDS * 4, cudaMemcpyDeviceToHost, ctx->d2h_stream));
        }
    }
    for (cudaStream_t ps : ctx->pipe_fit_stream) CK(cudaStreamSynchronize(ps));
    CK(cudaStreamSynchronize(st));
    CK(cudaStreamSynchronize(ctx->copy_stream));
    CK(cudaStreamSynchronize(ctx->d2h_stream));
    if (!direct_out && (h_out_joint || h_out_em || h_out_scal))
        return sdmmb_download_cells(ctx, first_cell, n_cells, h_out_joint, h_out_em, h_out_scal, stream);
    return SDMMB_OK;
}

int sdmmb_set_em_overlap(sdmmb_ctx* ctx, int enabled) {
    if (!ctx) return SDMMB_E_INVALID;
    ctx->em_overlap = enabled != 0;
    return SDMMB_OK;
}

int sdmmb_prepare_conditioner(sdmmb_ctx* ctx, int first_cell, int n_cells, void* stream) {
    if (!ctx) return SDMMB_E_INVALID;
    if (!range_ok(ctx, first_cell, n_cells)) return SDMMB_E_RANGE;
    if (n_cells == 0) return SDMMB_OK;
    CTX_LOCK(ctx);
    DeviceGuard g(ctx->device);
    MStepArgs a = base_mstep_args(ctx, first_cell, n_cells, 2);
    PeerPush push2[2] = {};
    int rc = arm_peer_push(ctx, push2, (cudaStream_t)stream);
    if (rc) return rc;
    a.peer = push2[1];
    CK(launch_mstep(a, (cudaStream_t)stream));
    ctx->launches += 1;
    return SDMMB_OK;
}

int sdmmb_set_min_samples(sdmmb_ctx* ctx, uint32_t min_samples) {
    if (!ctx) return SDMMB_E_INVALID;
    ctx->min_samples = min_samples;
    return SDMMB_OK;
}


}  // extern "C"
