// lr_wavefront_host.inl -- host driver of the wavefront `volpath` (kernels: lr_wavefront.cuh).
// Included by lr_backend_api.inl.  One render = GEN, then iterations of
//   TRACK -> SHADE_A (medium, surface buckets) -> NEE -> SHADE_B (medium, surface buckets) -> GEN
// until the active queue stays empty.  The host never waits for an iteration it has just enqueued: the live
// count of iteration i is read back asynchronously and checked three iterations later, so the stream stays full.

static const int WF_RING = 8, WF_LAG = 3;

static int wf_prepare(lr_ctx *c, uint64_t n_lanes) {
    if (!c->wf) c->wf = new WavefrontBuffers();
    WavefrontBuffers &w = *c->wf;
    uint64_t want = std::min<uint64_t>(c->wf_pool_slots, (n_lanes + 31) & ~31ull);
    want = std::max<uint64_t>(want, 1024);
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

template <int C, bool TO_FILM>
static int wavefront_render_t(lr_ctx *c, const RenderArgs &ra) {
    WavefrontBuffers &w = *c->wf; cudaStream_t s = c->stream;
    WFArgs a{}; a.st = w.states.as<WState>(); a.q = w.counters.as<WQueues>(); a.cap = w.cap;
    uint32_t *qb = w.queues.as<uint32_t>();
    a.q_active[0] = qb; a.q_active[1] = qb + w.cap; a.q_scatter = qb + 2 * (size_t) w.cap; a.q_regen = qb + 3 * (size_t) w.cap;
    for (int b = 0; b < WF_N_SURF; ++b) a.q_surf[b] = qb + (4 + b) * (size_t) w.cap;
    a.q_nee = qb + (4 + WF_N_SURF) * (size_t) w.cap;
    a.tris_gid = ra.tris_gid; a.film = ra.film; a.lanes_out = ra.lanes_out; a.lane_begin = ra.lane_begin; a.lane_end = ra.lane_end;
    a.seed = ra.seed; a.spp_per_pass = ra.spp_per_pass; a.n_passes = ra.n_passes; a.stats = ra.stats;
    // surface buckets present in this scene (bucket 0 also takes escaped rays)
    uint32_t surf_mask = 1u | c->surf_bucket_mask;
    const int grid = c->n_sm * 8, grid128 = c->n_sm * 8;
    uint64_t launches = 0; size_t kev_used = 0; std::vector<int> kev_kind;
    auto timed = [&](int kind) -> cudaEvent_t * {                                  // returns {start, stop} events for a TRACK (0) / NEE (1) launch
        if (kev_used + 2 > w.kev.size()) { size_t n0 = w.kev.size(); w.kev.resize(n0 + 512); for (size_t i = n0; i < w.kev.size(); ++i) cudaEventCreate(&w.kev[i]); }
        kev_kind.push_back(kind); kev_used += 2; return &w.kev[kev_used - 2];
    };
    k_wf_init<<<(w.cap + 255) / 256, 256, 0, s>>>(a.q, a.q_regen, w.cap, a.st);
    a.cur = 0; k_wf_gen<<<grid, 256, 0, s>>>(c->sc, a); a.cur = 1; launches += 2;
    LR_CUDA(c, cudaGetLastError());
    for (uint64_t iter = 0;; ++iter) {
        if (iter >= (uint64_t) WF_LAG) {
            int slot = (int) ((iter - WF_LAG) % WF_RING);
            LR_CUDA(c, cudaEventSynchronize(w.ring_ev[slot]));
            if (w.live_host[slot] == 0 || *c->cancel_host) break;
        }
        k_wf_begin_iteration<<<1, 1, 0, s>>>(a.q, a.cur);
        { cudaEvent_t *e = timed(0); cudaEventRecord(e[0], s); k_wf_track<C, TO_FILM><<<grid, 256, 0, s>>>(c->sc, a); cudaEventRecord(e[1], s); }
        k_wf_shade_a_medium<<<grid, 256, 0, s>>>(c->sc, a); launches += 3;
        for (int b = 0; b < WF_N_SURF; ++b) if ((surf_mask >> b) & 1u) { k_wf_shade_a_surface<C, TO_FILM><<<grid128, 128, 0, s>>>(c->sc, a, b); ++launches; }
        { cudaEvent_t *e = timed(1); cudaEventRecord(e[0], s); k_wf_nee<<<grid, 256, 0, s>>>(c->sc, a, surf_mask); cudaEventRecord(e[1], s); }
        k_wf_shade_b_medium<<<grid, 256, 0, s>>>(c->sc, a); launches += 2;
        for (int b = 0; b < WF_N_SURF; ++b) if ((surf_mask >> b) & 1u) { k_wf_shade_b_surface<<<grid128, 128, 0, s>>>(c->sc, a, b); ++launches; }
        k_wf_gen<<<grid, 256, 0, s>>>(c->sc, a); ++launches;
        a.cur ^= 1;
        int slot = (int) (iter % WF_RING);
        LR_CUDA(c, cudaMemcpyAsync(&w.live_host[slot], &a.q->n_active[a.cur], 4, cudaMemcpyDeviceToHost, s));
        LR_CUDA(c, cudaEventRecord(w.ring_ev[slot], s));
        LR_CUDA(c, cudaGetLastError());
    }
    LR_CUDA(c, cudaStreamSynchronize(s));
    double ms_track = 0, ms_nee = 0; uint64_t n_track = 0, n_nee = 0;
    for (size_t i = 0; i < kev_kind.size(); ++i) {
        float ms = 0.f; if (cudaEventElapsedTime(&ms, w.kev[2 * i], w.kev[2 * i + 1]) != cudaSuccess) continue;
        if (kev_kind[i] == 0) { ms_track += ms; ++n_track; } else { ms_nee += ms; ++n_nee; }
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
    WFArgs a{}; a.st = w.states.as<WState>(); a.q = w.counters.as<WQueues>(); a.cap = w.cap;
    uint32_t *qb = w.queues.as<uint32_t>();
    a.q_active[0] = qb; a.q_active[1] = qb + w.cap; a.q_scatter = qb + 2 * (size_t) w.cap; a.q_regen = qb + 3 * (size_t) w.cap;
    for (int b = 0; b < WF_N_SURF; ++b) a.q_surf[b] = qb + (4 + b) * (size_t) w.cap;
    a.q_nee = qb + (4 + WF_N_SURF) * (size_t) w.cap;
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
