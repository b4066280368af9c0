// lr_wavefront_host.inl -- host drivers of the wavefront `volpath` (kernels: lr_wavefront.cuh) and `prbvolpath`
// (lr_wavefront_prb.cuh).  Included by lr_backend_api.inl.  One forward batch = GEN, then iterations of
//   TRACK -> SHADE_A (medium, surface buckets) -> ESCAPE -> TRACE_S -> NEE -> SHADE_B (medium, surface buckets) -> TRACE_C [-> REGEN]
// until nothing is alive.  The host never waits for an iteration it has just enqueued.

static const int WF_RING = 8, WF_LAG = 3;

// ---------------------------------------------------------------------------------------------
// wavefront pools and host loop (kernels: lr_wavefront.cuh, lr_wavefront_prb.cuh)
// ---------------------------------------------------------------------------------------------
// stage kinds of the per-stage profile (lr_profile_get)
enum { ST_GEN = 0, ST_TRACE_C, ST_TRACK, ST_SHADE_A_MED, ST_SHADE_A_SURF, ST_ESCAPE, ST_TRACE_S, ST_NEE, ST_SHADE_B_MED, ST_SHADE_B_SURF, ST_REGEN, ST_OTHER, ST_COUNT };
static const char *const kStageNames[ST_COUNT] = { "gen", "trace_closest", "track", "shade_a_medium", "shade_a_surface", "escape", "trace_shadow", "nee",
                                                  "shade_b_medium", "shade_b_surface", "regen", "other" };

static int wv_alloc(lr_ctx *c, WVPool *&wp, uint64_t want) {
    if (!wp) wp = new WVPool();
    WVPool &w = *wp;
    if (w.cap < want) {
        w.cap = (uint32_t) want;
        LR_CUDA(c, w.planes.alloc((size_t) w.cap * 16 * WV_N_PLANES));
        LR_CUDA(c, w.queues.alloc((size_t) w.cap * 4 * WV_N_QUEUES));
        LR_CUDA(c, w.counters.alloc(sizeof(WVQueues)));
    }
    if (c->sc.integ.kind == LR_INTEGRATOR_VOLPATHMIS && w.mis.bytes < (size_t) w.cap * 128) LR_CUDA(c, w.mis.alloc((size_t) w.cap * 128));
    {   // stack entries beyond the shared-memory part of the trace kernels' stacks (one column per resident thread)
        const size_t need = (size_t) c->n_sm * LR_TRACE_THREADS * std::max<uint32_t>(c->sc.bvh4_stack > c->trace_sd ? c->sc.bvh4_stack - c->trace_sd : 0u, 1u) * sizeof(int);
        if (w.trace_ovf.bytes < need) LR_CUDA(c, w.trace_ovf.alloc(need));
    }
    if (!w.ring_host) {
        LR_CUDA(c, cudaHostAlloc((void **) &w.ring_host, WF_RING * sizeof(WVRing), cudaHostAllocMapped));
        LR_CUDA(c, cudaHostGetDevicePointer((void **) &w.ring_dev, w.ring_host, 0));
        w.ring_ev.resize(WF_RING);
        for (auto &e : w.ring_ev) LR_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    return LR_OK;
}

// pools for a render of `n_lanes` lanes: one pool of min(wf_pool_slots, n_lanes) slots, further ones when the lanes cannot all be
// resident at once (they may: primary rays that leave the scene never take a slot)
static int wv_prepare(lr_ctx *c, uint64_t n_lanes) {
    uint64_t want = std::min<uint64_t>(c->wf_pool_slots, (n_lanes + 31) & ~31ull);
    want = std::max<uint64_t>(want, 1024);
    if (int r = wv_alloc(c, c->wv[0], want)) return r;
    const int runs = c->profile_on ? 1 : c->wf_runs;
    if (n_lanes > c->wv[0]->cap) {
        uint64_t batches = (n_lanes + c->wv[0]->cap - 1) / c->wv[0]->cap;
        for (int k = 1; k < runs && (uint64_t) k < batches; ++k) {
            if (int r = wv_alloc(c, c->wv[k], c->wv[0]->cap)) return r;
            if (!c->stream_more[k - 1]) LR_CUDA(c, cudaStreamCreateWithFlags(&c->stream_more[k - 1], cudaStreamNonBlocking));
        }
        if (!c->fork_ev) LR_CUDA(c, cudaEventCreateWithFlags(&c->fork_ev, cudaEventDisableTiming));
    }
    if (!c->lane_cursor.p) LR_CUDA(c, c->lane_cursor.alloc(8 + 16 * 8));       // lane cursor + per-stage item totals
    return LR_OK;
}

static void wv_bind(WVArgs &a, WVPool &w) {
    char *pb = (char *) w.planes.p; const size_t u = (size_t) w.cap * 16;           // 16-byte units per slot: h0 2, h1 2, h2 2, hit 1, lane 1, n0 2, n1 2
    a.p.h0 = (float4 *) pb; a.p.h1 = (float4 *) (pb + 2 * u); a.p.h2 = (float4 *) (pb + 4 * u); a.p.hit_ = (float4 *) (pb + 6 * u); a.p.lane_ = (float4 *) (pb + 7 * u);
    a.p.n0 = (float4 *) (pb + 8 * u); a.p.n1 = (float4 *) (pb + 10 * u); a.p.adj_ = (float4 *) (pb + 12 * u);
    a.q = w.counters.as<WVQueues>(); a.cap = w.cap; a.mis = w.mis.as<float4>();
    uint32_t *qb = w.queues.as<uint32_t>(); const size_t qs = w.cap;
    a.q_trace = qb; a.q_track = qb + qs; a.q_scatter = qb + 2 * qs; a.q_escape = qb + 3 * qs; a.q_shadow = qb + 4 * qs; a.q_nee = qb + 5 * qs; a.q_regen = qb + 6 * qs;
    for (int b = 0; b < WF_N_SURF; ++b) a.q_surf[b] = qb + (7 + b) * qs;
}

// One batch of lanes in flight on one pool + stream.
struct WvRun {
    WVPool *w = nullptr; cudaStream_t s = nullptr; WVArgs a{}; TraceCfg tcfg{}; uint64_t iter = 0; bool busy = false; int gen_grid = 1;
};

// A render is a sequence of lock-step batches.  GEN fills one pool with the lanes whose primary ray hits something (or starts
// inside a medium) -- the others are finished inside GEN -- so iteration k of a batch holds the k-th vertex of paths whose queue
// order is still lane (= pixel) order: warps stay spatially coherent.  The price is the thin tail of a batch, so up to wf_runs
// batches are kept in flight on separate streams and pools: the tail iterations of one batch (and the tail of every single launch)
// are filled by the others.  All batches draw their lanes from one device-side cursor; lanes are independent, so results do not
// depend on the batching.  The host never waits for an iteration it has just enqueued: the live count of iteration i is written
// to mapped pinned memory by the iteration itself and checked WF_LAG iterations later.
// MODE 0: volpath (lr_wavefront.cuh); 1: prbvolpath primal, 2: prbvolpath adjoint replay (lr_wavefront_prb.cuh); 3: biovolpath
// (lr_wavefront_bio.cuh: no medium SHADE stages); 4 / 5: volpathmis with / without spectral MIS (lr_wavefront_mis.cuh) -- the same stage sequence, queues, pools and TRACE kernels; the stage kernels differ.
template <int C, bool TO_FILM, int MODE = 0>
static int wavefront_render_t(lr_ctx *c, const RenderArgs &ra, const WFPrbArgs pa = WFPrbArgs{}) {
    constexpr bool ADJ = MODE == 2, MIS = MODE == 4 || MODE == 5, SMIS = MODE == 4;
    const uint64_t cap = c->wv[0]->cap, n_lanes = ra.lane_end - ra.lane_begin;
    const bool prof = c->profile_on;
    WvRun runs[4]; int n_runs = 0;
    for (int k = 0; k < (prof ? 1 : c->wf_runs) && k < 4; ++k) {
        if (!c->wv[k] || c->wv[k]->cap < cap) break;
        if (k > 0 && n_lanes <= cap) break;
        runs[k].w = c->wv[k]; runs[k].s = k == 0 ? c->stream : c->stream_more[k - 1]; ++n_runs;
        runs[k].tcfg.n_smem = c->trace_n_smem; runs[k].tcfg.sd = c->trace_sd; runs[k].tcfg.ovf = c->wv[k]->trace_ovf.as<int>(); runs[k].tcfg.ovf_stride = (uint32_t) c->n_sm * LR_TRACE_THREADS;
    }
    if (n_runs > 1) {                                                  // the extra streams start after whatever is queued on the caller's stream
        LR_CUDA(c, cudaEventRecord(c->fork_ev, c->stream));
        for (int k = 1; k < n_runs; ++k) LR_CUDA(c, cudaStreamWaitEvent(runs[k].s, c->fork_ev, 0));
    }
    LR_CUDA(c, cudaMemsetAsync(c->lane_cursor.p, 0, 8 + 16 * 8, c->stream));
    unsigned long long *d_items = (unsigned long long *) c->lane_cursor.p + 1;
    const WVStageIds sids{ST_TRACK, ST_SHADE_A_MED, ST_SHADE_A_SURF, ST_ESCAPE, ST_TRACE_S, ST_NEE, ST_SHADE_B_MED, ST_SHADE_B_SURF, ST_TRACE_C, ST_REGEN};
    if (n_runs > 1) { LR_CUDA(c, cudaEventRecord(c->fork_ev, c->stream)); for (int k = 1; k < n_runs; ++k) LR_CUDA(c, cudaStreamWaitEvent(runs[k].s, c->fork_ev, 0)); }
    const uint32_t surf_mask = c->surf_bucket_mask;                    // surface buckets present in this scene
    const int grid = c->n_sm * c->wf_grid_mult, grid128 = c->n_sm * c->wf_grid_mult128, grid_tr = c->n_sm;      // trace: one block of LR_TRACE_THREADS per SM
    const int grid_mis = c->n_sm * LR_MIS_OCC, grid128_mis = c->n_sm * 2 * LR_MIS_OCC;       // the volpathmis stage kernels are compiled for 128 registers
    const bool stk16 = c->bvh4_small; const size_t smem_tr = c->trace_smem;
    const bool multipass = ra.n_passes > 1;
    const bool ref_persistent = c->sc.integ.tracking == LR_TRACKING_REFERENCE && !getenv("LR_REF_REFILL_OFF");   // reference-order TRACK / NEE as refill loops (A/B: the one-item-per-thread kernels)
    const int fuse = getenv("LR_TRACK_FUSE_OFF") ? 0 : 1;               // A/B: scatter vertices of sealed media finished inside TRACK
    bool all_sealed = !c->h_media.empty(); for (const auto &m : c->h_media) all_sealed = all_sealed && m.sealed != 0;
    const bool no_medium_shade = fuse && c->all_media_dda && all_sealed &&                 // nothing ever reaches the scatter queue
                                 (MODE == 0 || ((MODE == 1 || MODE == 2) && !getenv("LR_PRB_TRACK_SIMPLE")));
    const int shadow_mode = (MODE < 3 && c->sc.integ.tracking == LR_TRACKING_DDA && !c->sc.any_null_bsdf && !getenv("LR_SHADOW_CLOSEST")) ? 1 : 2;   // LR_SHADOW_CLOSEST: A/B of the any-hit query
    uint64_t launches = 0; bool lanes_done = false; const bool dbg_sync = getenv("LR_WF_SYNC") != nullptr;
    // per-stage CUDA events (profile mode: one batch in flight, so the intervals do not overlap)
    std::vector<cudaEvent_t> &kev = c->prof_ev; size_t kev_used = 0; std::vector<int> kev_kind;
    auto launch = [&](WvRun &r, int kind, auto &&fn) {
        cudaEvent_t *e = nullptr;
        if (prof) {
            if (kev_used + 2 > kev.size()) { size_t n0 = kev.size(); kev.resize(n0 + 1024); for (size_t i = n0; i < kev.size(); ++i) cudaEventCreate(&kev[i]); }
            e = &kev[kev_used]; kev_used += 2; kev_kind.push_back(kind); cudaEventRecord(e[0], r.s);
        }
        if (c->nvtx) nvtxRangePushA(kStageNames[kind]);                // NVTX range per stage launch (LR_NVTX=1), like the reference's ProfilerPhase scopes
        fn(); ++launches;
        if (c->nvtx) nvtxRangePop();
        if (prof) cudaEventRecord(e[1], r.s);
        if (dbg_sync) {                                                // LR_WF_SYNC=1: find the stage a device fault comes from
            cudaError_t err = cudaStreamSynchronize(r.s);
            if (err != cudaSuccess) { fprintf(stderr, "[lr] stage %s (iteration %llu): %s\n", kStageNames[kind], (unsigned long long) r.iter, cudaGetErrorString(err)); }
        }
    };
    auto readback = [&](WvRun &r) -> int {                             // live count + lane cursor of the work enqueued so far -> ring slot
        WVPool &w = *r.w; int slot = (int) (r.iter % WF_RING);
        k_wv_end_iteration<<<1, 1, 0, r.s>>>(r.a.q, &w.ring_dev[slot].live, (const unsigned long long *) c->lane_cursor.p, &w.ring_dev[slot].lane_next, d_items, sids);
        if (cudaEventRecord(w.ring_ev[slot], r.s) != cudaSuccess) return -1;
        ++r.iter; ++launches;
        return cudaGetLastError() == cudaSuccess ? 0 : -1;
    };
    auto start = [&](WvRun &r) -> int {                                // a new batch on this pool
        WVArgs &a = r.a; a = WVArgs{}; wv_bind(a, *r.w);
        a.film = ra.film; a.lanes_out = ra.lanes_out; a.seed = ra.seed; a.spp_per_pass = ra.spp_per_pass; a.n_passes = ra.n_passes; a.stats = ra.stats;
        a.lane_begin = ra.lane_begin; a.lane_end = ra.lane_end; a.map = ra.map; a.lane_next = (unsigned long long *) c->lane_cursor.p;
        // GEN hands out lanes while n_alloc + slack <= cap; slack = 32 slots per resident warp of the GEN grid bounds the overshoot
        r.gen_grid = (int) std::max<uint64_t>(1, std::min<uint64_t>((uint64_t) grid, a.cap / 512)); a.slack = (uint32_t) r.gen_grid * 256u;
        k_wv_init<<<1, 1, 0, r.s>>>(a.q); ++launches;
        launch(r, ST_GEN, [&] { if constexpr (MODE == 0 || MODE >= 3) k_wv_gen<C, TO_FILM><<<r.gen_grid, 256, 0, r.s>>>(c->sc, a);
                                else k_wvp_gen<C, TO_FILM, ADJ><<<r.gen_grid, 256, 0, r.s>>>(c->sc, a, pa); });
        r.iter = 0; r.busy = true;
        return readback(r);
    };
    auto step = [&](WvRun &r) -> int {                                 // enqueue one iteration; 1: the batch has drained
        WVArgs &a = r.a; cudaStream_t s = r.s; WVPool &w = *r.w;
        if (r.iter >= (uint64_t) WF_LAG) {
            int slot = (int) ((r.iter - WF_LAG) % WF_RING);
            if (cudaEventSynchronize(w.ring_ev[slot]) != cudaSuccess) return -1;
            if (w.ring_host[slot].lane_next >= n_lanes) lanes_done = true;
            if (w.ring_host[slot].live == 0 || *c->cancel_host) { r.busy = false; return 1; }
        }
        launch(r, ST_TRACK, [&] { if constexpr (MODE == 3) k_bio_track<C, TO_FILM><<<grid, 256, 0, s>>>(c->sc, a);
                                  else if constexpr (MIS) k_mis_track<C, TO_FILM, SMIS><<<grid_mis, 256, 0, s>>>(c->sc, a);
                                  else if constexpr (MODE != 0) { if (c->all_media_dda && !getenv("LR_PRB_TRACK_SIMPLE")) k_wvp_track_dda<C, TO_FILM, ADJ><<<grid, 256, 0, s>>>(c->sc, a, fuse); else k_wvp_track<C, TO_FILM, ADJ><<<grid, 256, 0, s>>>(c->sc, a); }
                                  else if (c->all_media_dda && c->smaj_floats) k_wv_track_dda<C, TO_FILM, true><<<c->n_sm, 1024, (size_t) c->smaj_floats * 4, s>>>(c->sc, a, TrackSmaj{c->smaj_grid, c->smaj_floats, fuse});
                                  else if (c->all_media_dda) k_wv_track_dda<C, TO_FILM, false><<<grid, 256, 0, s>>>(c->sc, a, TrackSmaj{-1, 0u, fuse});
                                  else if (ref_persistent) k_wv_track_ref<C, TO_FILM><<<grid, 256, 0, s>>>(c->sc, a); else k_wv_track<C, TO_FILM><<<grid, 256, 0, s>>>(c->sc, a); });
        if (MODE != 3 && !no_medium_shade) launch(r, ST_SHADE_A_MED, [&] { if constexpr (MODE == 0) k_wv_shade_a_medium<<<grid, 256, 0, s>>>(c->sc, a); else if constexpr (MIS) k_mis_shade_a_medium<SMIS><<<grid_mis, 256, 0, s>>>(c->sc, a);
                                                       else if constexpr (MODE != 3) k_wvp_shade_a_medium<ADJ><<<grid, 256, 0, s>>>(c->sc, a); });
        for (int b = 0; b < WF_N_SURF; ++b) if ((surf_mask >> b) & 1u) launch(r, ST_SHADE_A_SURF, [&] {
            if constexpr (MODE == 0) k_wv_shade_a_surface<C, TO_FILM><<<grid128, 128, 0, s>>>(c->sc, a, b);
            else if constexpr (MODE == 3) k_bio_shade_a<<<grid128, 128, 0, s>>>(c->sc, a, b);
            else if constexpr (MIS) k_mis_shade_a_surface<C, TO_FILM, SMIS><<<grid128_mis, 128, 0, s>>>(c->sc, a, b);
            else k_wvp_shade_a_surface<C, TO_FILM, ADJ><<<grid128, 128, 0, s>>>(c->sc, a, b); });
        if (MODE != 2) launch(r, ST_ESCAPE, [&] { if constexpr (MODE == 0) k_wv_escape<C, TO_FILM><<<grid, 256, 0, s>>>(c->sc, a);
                                                  else if constexpr (MODE == 3) k_bio_escape<C, TO_FILM><<<grid, 256, 0, s>>>(c->sc, a);
                                                  else if constexpr (MIS) k_mis_escape<C, TO_FILM, SMIS><<<grid_mis, 256, 0, s>>>(c->sc, a);
                                                  else k_wvp_escape<C, TO_FILM, ADJ><<<grid, 256, 0, s>>>(c->sc, a); });
        launch(r, ST_TRACE_S, [&] {
            if (shadow_mode == 1) { if (stk16) k_wv_trace<1, short><<<grid_tr, LR_TRACE_THREADS, smem_tr, s>>>(c->sc, a, r.tcfg); else k_wv_trace<1, int><<<grid_tr, LR_TRACE_THREADS, smem_tr, s>>>(c->sc, a, r.tcfg); }
            else { if (stk16) k_wv_trace<2, short><<<grid_tr, LR_TRACE_THREADS, smem_tr, s>>>(c->sc, a, r.tcfg); else k_wv_trace<2, int><<<grid_tr, LR_TRACE_THREADS, smem_tr, s>>>(c->sc, a, r.tcfg); } });
        launch(r, ST_NEE, [&] { if constexpr (MODE == 0) { if (ref_persistent && shadow_mode == 2) k_wv_nee_ref<<<grid, 256, 0, s>>>(c->sc, a); else k_wv_nee<<<grid, 256, 0, s>>>(c->sc, a, shadow_mode == 2 ? 1 : 0); }
                                else if constexpr (MODE == 3) k_bio_nee<<<grid, 256, 0, s>>>(c->sc, a);
                                else if constexpr (MIS) k_mis_nee<SMIS><<<grid_mis, 256, 0, s>>>(c->sc, a);
                                else k_wvp_nee<ADJ><<<grid, 256, 0, s>>>(c->sc, a, shadow_mode == 2 ? 1 : 0); });
        if (MODE != 3 && !no_medium_shade) launch(r, ST_SHADE_B_MED, [&] { if constexpr (MODE == 0) k_wv_shade_b_medium<C, TO_FILM><<<grid, 256, 0, s>>>(c->sc, a); else if constexpr (MIS) k_mis_shade_b_medium<C, TO_FILM, SMIS><<<grid_mis, 256, 0, s>>>(c->sc, a);
                                                       else if constexpr (MODE != 3) k_wvp_shade_b_medium<C, TO_FILM, ADJ><<<grid, 256, 0, s>>>(c->sc, a); });
        for (int b = 0; b < WF_N_SURF; ++b) if ((surf_mask >> b) & 1u) launch(r, ST_SHADE_B_SURF, [&] {
            if constexpr (MODE == 0) k_wv_shade_b_surface<C, TO_FILM><<<grid128, 128, 0, s>>>(c->sc, a, b);
            else if constexpr (MODE == 3) k_bio_shade_b<C, TO_FILM><<<grid128, 128, 0, s>>>(c->sc, a, b);
            else if constexpr (MIS) k_mis_shade_b_surface<C, TO_FILM, SMIS><<<grid128_mis, 128, 0, s>>>(c->sc, a, b);
            else k_wvp_shade_b_surface<C, TO_FILM, ADJ><<<grid128, 128, 0, s>>>(c->sc, a, b); });
        k_wv_mid_iteration<<<1, 1, 0, s>>>(a.q, d_items, sids); ++launches;
        launch(r, ST_TRACE_C, [&] { if (stk16) k_wv_trace<0, short><<<grid_tr, LR_TRACE_THREADS, smem_tr, s>>>(c->sc, a, r.tcfg); else k_wv_trace<0, int><<<grid_tr, LR_TRACE_THREADS, smem_tr, s>>>(c->sc, a, r.tcfg); });
        if constexpr (MODE == 0 || MODE >= 3) { if (multipass) launch(r, ST_REGEN, [&] { k_wv_regen<C, TO_FILM><<<grid, 256, 0, s>>>(c->sc, a); }); }
        return readback(r);
    };
    for (bool any = true; any;) {
        any = false;
        for (int k = 0; k < n_runs; ++k) {
            WvRun &r = runs[k];
            if (!r.busy) {
                if (*c->cancel_host || lanes_done) continue;
                if (start(r) < 0) { LR_CUDA(c, cudaGetLastError()); LR_CUDA(c, cudaDeviceSynchronize()); c->err = "wavefront batch start failed"; return LR_ERR_CUDA; }
            }
            any = true;
            if (step(r) < 0) { LR_CUDA(c, cudaGetLastError()); LR_CUDA(c, cudaDeviceSynchronize()); c->err = "wavefront iteration failed"; return LR_ERR_CUDA; }
        }
    }
    LR_CUDA(c, cudaStreamSynchronize(c->stream));
    for (int k = 1; k < n_runs; ++k) LR_CUDA(c, cudaStreamSynchronize(runs[k].s));
    LR_CUDA(c, cudaGetLastError());
    {   // queue entries each stage consumed in this call (GEN: the lanes of the call)
        unsigned long long h_items[16]; LR_CUDA(c, cudaMemcpy(h_items, d_items, sizeof h_items, cudaMemcpyDeviceToHost));
        for (int k = 0; k < ST_COUNT; ++k) c->prof_items[k] += h_items[k];
        c->prof_items[ST_GEN] += n_lanes;
    }
    if (prof) {
        for (size_t i = 0; i < kev_kind.size(); ++i) {
            float ms = 0.f; if (cudaEventElapsedTime(&ms, kev[2 * i], kev[2 * i + 1]) != cudaSuccess) continue;
            c->prof_ms[kev_kind[i]] += ms; c->prof_launches[kev_kind[i]] += 1;
        }
        c->wf_tracking_ms = c->prof_ms[ST_TRACK] + c->prof_ms[ST_NEE]; c->wf_tracking_launches = c->prof_launches[ST_TRACK] + c->prof_launches[ST_NEE];
    }
    c->wf_launches = launches;
    return LR_OK;
}

// ---------------------------------------------------------------------------------------------
// render_backward as two wavefront passes per lane chunk (kernels: lr_wavefront_prb.cuh): primal radiance per lane -> buffer,
// then the adjoint replay of the same streams
// ---------------------------------------------------------------------------------------------
static int wavefront_prb_backward(lr_ctx *c, RenderArgs ra, const float *grad_dev, uint32_t grad_channels) {
    const uint64_t chunk = 1ull << 27;                                  // 128 Mi lanes: 2 GiB of per-lane radiance
    uint64_t n = ra.lane_end - ra.lane_begin;
    if (int r = wv_prepare(c, std::min(n, chunk))) return r;
    size_t need = (size_t) std::min(n, chunk) * 16;
    if (c->prb_L.bytes < need) LR_CUDA(c, c->prb_L.alloc(need));
    uint64_t launches = 0; double t_ms = 0; uint64_t t_n = 0;
    WFPrbArgs pa{}; pa.grad_in = grad_dev; pa.weight_img = c->weight_img.as<float>(); pa.grad_channels = grad_channels; pa.primal_L = c->prb_L.as<float>();
    for (uint64_t b = ra.lane_begin; b < ra.lane_end; b += chunk) {
        RenderArgs r1 = ra; r1.lane_begin = b; r1.lane_end = std::min(ra.lane_end, b + chunk); r1.lanes_out = c->prb_L.as<float>(); r1.film = nullptr;
        if (int r = wavefront_render_t<4, false, 1>(c, r1, pa)) return r;
        launches += c->wf_launches; t_ms += c->wf_tracking_ms; t_n += c->wf_tracking_launches;
        r1.lanes_out = nullptr;
        if (int r = wavefront_render_t<4, false, 2>(c, r1, pa)) return r;
        launches += c->wf_launches; t_ms += c->wf_tracking_ms; t_n += c->wf_tracking_launches;
    }
    c->wf_launches = launches; c->wf_tracking_ms = t_ms; c->wf_tracking_launches = t_n;
    return LR_OK;
}
