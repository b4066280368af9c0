// lr_wavefront_host.inl -- host driver of the wavefront `volpath` (kernels: lr_wavefront.cuh).
// Included by lr_backend_api.inl.  One render = GEN, then iterations of
//   TRACK -> SHADE_A (medium, surface buckets) -> NEE -> SHADE_B (medium, surface buckets) -> GEN
// until the active queue stays empty.  The host never waits for an iteration it has just enqueued: the live
// count of iteration i is read back asynchronously and checked three iterations later, so the stream stays full.

static const int WF_RING = 8, WF_LAG = 3;

static int wf_alloc(lr_ctx *c, WavefrontBuffers *&wp, uint64_t want) {
    if (!wp) wp = new WavefrontBuffers();
    WavefrontBuffers &w = *wp;
    if (w.cap < want) {
        w.cap = (uint32_t) want;
        LR_CUDA(c, w.states.alloc((size_t) w.cap * sizeof(WState)));
        LR_CUDA(c, w.queues.alloc((size_t) w.cap * 4 * (5 + WF_N_SURF)));
        LR_CUDA(c, w.counters.alloc(sizeof(WQueues)));
    }
    if (!w.live_host) {
        LR_CUDA(c, cudaHostAlloc((void **) &w.live_host, WF_RING * sizeof(uint32_t), cudaHostAllocDefault));
        w.ring_ev.resize(WF_RING);
        for (auto &e : w.ring_ev) LR_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    return LR_OK;
}

// `two_runs`: also set up the second pool + stream used when a lane range is rendered as overlapped batches
static int wf_prepare(lr_ctx *c, uint64_t n_lanes, bool two_runs = false) {
    uint64_t want = std::min<uint64_t>(c->wf_pool_slots, (n_lanes + 31) & ~31ull);
    want = std::max<uint64_t>(want, 1024);
    if (int r = wf_alloc(c, c->wf, want)) return r;
    if (two_runs && n_lanes > c->wf->cap) {
        uint64_t batches = (n_lanes + c->wf->cap - 1) / c->wf->cap;
        for (int k = 0; k + 1 < c->wf_runs && (uint64_t) (k + 1) < batches; ++k) {
            if (int r = wf_alloc(c, c->wf_more[k], c->wf->cap)) return r;
            if (!c->stream_more[k]) LR_CUDA(c, cudaStreamCreateWithFlags(&c->stream_more[k], cudaStreamNonBlocking));
        }
        if (!c->fork_ev) LR_CUDA(c, cudaEventCreateWithFlags(&c->fork_ev, cudaEventDisableTiming));
    }
    return LR_OK;
}

static void wf_bind(WFArgs &a, WavefrontBuffers &w) {
    a.st = w.states.as<WState>(); a.q = w.counters.as<WQueues>(); a.cap = w.cap;
    uint32_t *qb = w.queues.as<uint32_t>();
    a.q_active[0] = qb; a.q_active[1] = qb + w.cap; a.q_scatter = qb + 2 * (size_t) w.cap; a.q_regen = qb + 3 * (size_t) w.cap;
    for (int b = 0; b < WF_N_SURF; ++b) a.q_surf[b] = qb + (4 + b) * (size_t) w.cap;
    a.q_nee = qb + (4 + WF_N_SURF) * (size_t) w.cap;
}

// One batch of lanes in flight on one pool + stream.
struct WfRun {
    WavefrontBuffers *w = nullptr; cudaStream_t s = nullptr; WFArgs a{}; uint64_t iter = 0; bool busy = false;
    size_t kev_used = 0; std::vector<int> kev_kind;
};

// The lane range is cut into batches of one pool each.  A batch runs in lock step -- every lane of it starts in the same GEN
// launch, so iteration k holds the k-th vertex of paths whose queue order is still lane (= pixel) order: warps stay spatially
// coherent, which is worth 15-20 % over a steady-state pool that keeps refilling freed slots with unrelated lanes.  The price
// is the thin tail of a batch; two (LR_WF_RUNS, up to four) batches are kept in flight on separate streams so that the tail iterations (and the tail of
// every single launch) are filled by the other batch.  Lanes are independent, so results do not depend on the batching.
template <int C, bool TO_FILM>
static int wavefront_render_t(lr_ctx *c, const RenderArgs &ra) {
    const uint64_t cap = c->wf->cap, n_lanes = ra.lane_end - ra.lane_begin;
    WfRun runs[4]; runs[0].w = c->wf; runs[0].s = c->stream; int n_runs = 1;
    if (n_lanes > cap) {
        const uint64_t batches = (n_lanes + cap - 1) / cap;
        for (int k = 0; k < 3 && n_runs < c->wf_runs && (uint64_t) n_runs < batches && c->wf_more[k] && c->wf_more[k]->cap >= cap; ++k) {
            runs[n_runs].w = c->wf_more[k]; runs[n_runs].s = c->stream_more[k]; ++n_runs; }
    }
    if (n_runs > 1) {                                                  // the extra streams start after whatever is queued on the caller's stream
        LR_CUDA(c, cudaEventRecord(c->fork_ev, c->stream));
        for (int k = 1; k < n_runs; ++k) LR_CUDA(c, cudaStreamWaitEvent(runs[k].s, c->fork_ev, 0));
    }
    const uint32_t surf_mask = 1u | c->surf_bucket_mask;             // surface buckets present in this scene (bucket 0 also takes escaped rays)
    const int grid = c->n_sm * c->wf_grid_mult, grid128 = c->n_sm * c->wf_grid_mult128;   // blocks per SM handed to the grid-stride stage kernels (LR_WF_GRID / LR_WF_GRID128)
    uint64_t launches = 0, next = ra.lane_begin;
    auto timed = [&](WfRun &r, int kind) -> cudaEvent_t * {           // {start, stop} events for a TRACK (0) / NEE (1) launch
        auto &kev = r.w->kev;
        if (r.kev_used + 2 > kev.size()) { size_t n0 = kev.size(); kev.resize(n0 + 512); for (size_t i = n0; i < kev.size(); ++i) cudaEventCreate(&kev[i]); }
        r.kev_kind.push_back(kind); r.kev_used += 2; return &kev[r.kev_used - 2];
    };
    auto start = [&](WfRun &r) -> int {                                // 1: a batch was started, 0: no lanes left
        if (next >= ra.lane_end) return 0;
        WFArgs &a = r.a; a = WFArgs{}; wf_bind(a, *r.w);
        a.tris_gid = ra.tris_gid; a.film = ra.film; a.seed = ra.seed; a.spp_per_pass = ra.spp_per_pass; a.n_passes = ra.n_passes; a.stats = ra.stats;
        a.lane_begin = next; a.lane_end = std::min<uint64_t>(ra.lane_end, next + cap); next = a.lane_end;
        a.lanes_out = ra.lanes_out ? ra.lanes_out + (a.lane_begin - ra.lane_begin) * ra.n_passes * 4 : nullptr;
        k_wf_init<<<(r.w->cap + 255) / 256, 256, 0, r.s>>>(a.q, a.q_regen, r.w->cap, a.st);
        a.cur = 0; k_wf_gen<<<grid, 256, 0, r.s>>>(c->sc, a); a.cur = 1; launches += 2;
        r.iter = 0; r.busy = true;
        return 1;
    };
    auto step = [&](WfRun &r) -> int {                                 // enqueue one iteration; 1: the batch has drained
        WFArgs &a = r.a; cudaStream_t s = r.s; WavefrontBuffers &w = *r.w;
        if (r.iter >= (uint64_t) WF_LAG) {
            int slot = (int) ((r.iter - WF_LAG) % WF_RING);
            if (cudaEventSynchronize(w.ring_ev[slot]) != cudaSuccess) return -1;
            if (w.live_host[slot] == 0 || *c->cancel_host) { r.busy = false; return 1; }
        }
        k_wf_begin_iteration<<<1, 1, 0, s>>>(a.q, a.cur);
        { cudaEvent_t *e = timed(r, 0); cudaEventRecord(e[0], s); if (c->all_media_dda) k_wf_track_dda<C, TO_FILM><<<grid, 256, 0, s>>>(c->sc, a); else k_wf_track<C, TO_FILM><<<grid, 256, 0, s>>>(c->sc, a); cudaEventRecord(e[1], s); }
        k_wf_shade_a_medium<<<grid, 256, 0, s>>>(c->sc, a); launches += 3;
        for (int b = 0; b < WF_N_SURF; ++b) if ((surf_mask >> b) & 1u) { k_wf_shade_a_surface<C, TO_FILM><<<grid128, 128, 0, s>>>(c->sc, a, b); ++launches; }
        { cudaEvent_t *e = timed(r, 1); cudaEventRecord(e[0], s); k_wf_nee<<<grid, 256, 0, s>>>(c->sc, a, surf_mask); cudaEventRecord(e[1], s); }
        k_wf_shade_b_medium<<<grid, 256, 0, s>>>(c->sc, a); launches += 2;
        for (int b = 0; b < WF_N_SURF; ++b) if ((surf_mask >> b) & 1u) { k_wf_shade_b_surface<<<grid128, 128, 0, s>>>(c->sc, a, b); ++launches; }
        k_wf_gen<<<grid, 256, 0, s>>>(c->sc, a); ++launches;
        a.cur ^= 1;
        int slot = (int) (r.iter % WF_RING);
        if (cudaMemcpyAsync(&w.live_host[slot], &a.q->n_active[a.cur], 4, cudaMemcpyDeviceToHost, s) != cudaSuccess) return -1;
        if (cudaEventRecord(w.ring_ev[slot], s) != cudaSuccess) return -1;
        ++r.iter;
        return cudaGetLastError() == cudaSuccess ? 0 : -1;
    };
    for (bool any = true; any;) {
        any = false;
        for (int k = 0; k < n_runs; ++k) {
            WfRun &r = runs[k];
            if (!r.busy && (*c->cancel_host || !start(r))) continue;
            any = true;
            if (step(r) < 0) { LR_CUDA(c, cudaGetLastError()); LR_CUDA(c, cudaDeviceSynchronize()); c->err = "wavefront iteration failed"; return LR_ERR_CUDA; }
        }
    }
    LR_CUDA(c, cudaStreamSynchronize(c->stream));
    for (int k = 1; k < n_runs; ++k) LR_CUDA(c, cudaStreamSynchronize(runs[k].s));
    double ms_track = 0, ms_nee = 0; uint64_t n_track = 0, n_nee = 0;
    for (int k = 0; k < n_runs; ++k)
        for (size_t i = 0; i < runs[k].kev_kind.size(); ++i) {
            float ms = 0.f; if (cudaEventElapsedTime(&ms, runs[k].w->kev[2 * i], runs[k].w->kev[2 * i + 1]) != cudaSuccess) continue;
            if (runs[k].kev_kind[i] == 0) { ms_track += ms; ++n_track; } else { ms_nee += ms; ++n_nee; }
        }
    c->wf_launches = launches; c->wf_tracking_ms = ms_track + ms_nee; c->wf_tracking_launches = n_track + n_nee;
    c->wf_track_ms = ms_track; c->wf_nee_ms = ms_nee;
    return LR_OK;
}

// ---------------------------------------------------------------------------------------------
// wavefront prbvolpath (kernels: lr_wavefront_prb.cuh): the same iteration structure; ADJ selects the adjoint replay
// ---------------------------------------------------------------------------------------------
template <int C, bool TO_FILM, bool ADJ>
static int wavefront_prb_pass(lr_ctx *c, const RenderArgs &ra, const WFPrbArgs &pa) {
    WavefrontBuffers &w = *c->wf; cudaStream_t s = c->stream;
    WFArgs a{}; wf_bind(a, w);
    a.tris_gid = ra.tris_gid; a.film = ra.film; a.lanes_out = ra.lanes_out; a.lane_begin = ra.lane_begin; a.lane_end = ra.lane_end;
    a.seed = ra.seed; a.spp_per_pass = ra.spp_per_pass; a.n_passes = 1; a.stats = ra.stats;
    uint32_t surf_mask = 1u | c->surf_bucket_mask;
    const int grid = c->n_sm * 8;
    uint64_t launches = 0; size_t kev_used = 0;
    auto timed = [&]() -> cudaEvent_t * {
        if (kev_used + 2 > w.kev.size()) { size_t n0 = w.kev.size(); w.kev.resize(n0 + 512); for (size_t i = n0; i < w.kev.size(); ++i) cudaEventCreate(&w.kev[i]); }
        kev_used += 2; return &w.kev[kev_used - 2];
    };
    k_wf_init<<<(w.cap + 255) / 256, 256, 0, s>>>(a.q, a.q_regen, w.cap, a.st);
    a.cur = 0; k_wfp_gen<ADJ><<<grid, 256, 0, s>>>(c->sc, a, pa); a.cur = 1; launches += 2;
    LR_CUDA(c, cudaGetLastError());
    for (uint64_t iter = 0;; ++iter) {
        if (iter >= (uint64_t) WF_LAG) {
            int slot = (int) ((iter - WF_LAG) % WF_RING);
            LR_CUDA(c, cudaEventSynchronize(w.ring_ev[slot]));
            if (w.live_host[slot] == 0 || *c->cancel_host) break;
        }
        k_wf_begin_iteration<<<1, 1, 0, s>>>(a.q, a.cur);
        { cudaEvent_t *e = timed(); cudaEventRecord(e[0], s); k_wfp_track<C, TO_FILM, ADJ><<<grid, 256, 0, s>>>(c->sc, a); cudaEventRecord(e[1], s); }
        k_wfp_shade_a_medium<<<grid, 256, 0, s>>>(c->sc, a); launches += 3;
        for (int b = 0; b < WF_N_SURF; ++b) if ((surf_mask >> b) & 1u) { k_wfp_shade_a_surface<C, TO_FILM, ADJ><<<grid, 128, 0, s>>>(c->sc, a, b); ++launches; }
        { cudaEvent_t *e = timed(); cudaEventRecord(e[0], s); k_wfp_nee<ADJ><<<grid, 256, 0, s>>>(c->sc, a, surf_mask); cudaEventRecord(e[1], s); }
        k_wfp_shade_b_medium<<<grid, 256, 0, s>>>(c->sc, a); launches += 2;
        for (int b = 0; b < WF_N_SURF; ++b) if ((surf_mask >> b) & 1u) { k_wfp_shade_b_surface<C, TO_FILM, ADJ><<<grid, 128, 0, s>>>(c->sc, a, b); ++launches; }
        k_wfp_gen<ADJ><<<grid, 256, 0, s>>>(c->sc, a, pa); ++launches;
        a.cur ^= 1;
        int slot = (int) (iter % WF_RING);
        LR_CUDA(c, cudaMemcpyAsync(&w.live_host[slot], &a.q->n_active[a.cur], 4, cudaMemcpyDeviceToHost, s));
        LR_CUDA(c, cudaEventRecord(w.ring_ev[slot], s));
        LR_CUDA(c, cudaGetLastError());
    }
    LR_CUDA(c, cudaStreamSynchronize(s));
    double ms = 0; uint64_t n = 0;
    for (size_t i = 0; i + 1 < kev_used; i += 2) { float t = 0.f; if (cudaEventElapsedTime(&t, w.kev[i], w.kev[i + 1]) == cudaSuccess) { ms += t; ++n; } }
    c->wf_launches += launches; c->wf_tracking_ms += ms; c->wf_tracking_launches += n;
    return LR_OK;
}

// render_backward as two wavefront passes per lane chunk: primal radiance per lane -> buffer, then the adjoint replay
static int wavefront_prb_backward(lr_ctx *c, RenderArgs ra, const float *grad_dev, uint32_t grad_channels) {
    const uint64_t chunk = 1ull << 25;                                  // 32 Mi lanes: 512 MiB of per-lane radiance
    uint64_t n = ra.lane_end - ra.lane_begin;
    if (int r = wf_prepare(c, std::min(n, chunk))) return r;
    size_t need = (size_t) std::min(n, chunk) * 16;
    if (c->prb_L.bytes < need) LR_CUDA(c, c->prb_L.alloc(need));
    c->wf_launches = 0; c->wf_tracking_ms = 0; c->wf_tracking_launches = 0;
    WFPrbArgs pa{}; pa.grad_in = grad_dev; pa.weight_img = c->weight_img.as<float>(); pa.grad_channels = grad_channels; pa.primal_L = c->prb_L.as<float>();
    for (uint64_t b = ra.lane_begin; b < ra.lane_end; b += chunk) {
        RenderArgs r1 = ra; r1.lane_begin = b; r1.lane_end = std::min(ra.lane_end, b + chunk); r1.lanes_out = c->prb_L.as<float>(); r1.film = nullptr;
        if (int r = wavefront_prb_pass<4, false, false>(c, r1, pa)) return r;
        if (int r = wavefront_prb_pass<4, false, true>(c, r1, pa)) return r;
    }
    return LR_OK;
}
