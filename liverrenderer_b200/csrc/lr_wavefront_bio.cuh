// lr_wavefront_bio.cuh -- wavefront `biovolpath` (reference: src/integrators/biovolpath.cpp:95-551 with the tissue media of
// lr_biovolpath.cuh): the per-lane loop BioPath::step cut where lanes diverge, on the forward wavefront's machinery (lr_wavefront.cuh:
// SoA planes, ballot-compacted queues, GEN, the persistent TRACE stage with the BVH in shared memory, multi-batch host loop).
//
//   GEN        as volpath (first roulette draw, primary hit; a primary ray that leaves the scene outside any medium is finished there)
//   TRACE_C    closest hit of the path's ray over its WHOLE length; inside a tissue medium -> TRACK, else surface bucket / ESCAPE
//   TRACK      one tissue interaction (liver.cpp:483-547).  The reference shortens the ray to the sampled distance BEFORE it intersects
//              (:224-227), so "no surface found" is `hit.t > mei.t` here.  Scatter: phase sample, the roulette of the next iteration,
//              -> TRACE_C.  Boundary reached -> surface bucket / ESCAPE.  Tissue media have no emitter sampling at a scatter vertex
//   SHADE_A    surface: emitter hit + MIS, emitter sample + BSDF value -> shadow ray            ESCAPE  environment + MIS, film
//   TRACE_S    closest hit along the shadow ray (TRACE mode 2);  NEE: the walk of biovolpath.cpp:444-528 from that hit (bio_nee_walk)
//   SHADE_B    BSDF sample, medium transition, roulette of the next iteration (+ the dying-iteration rule, lr_biovolpath.cuh note (2))
//
// Extra state in the `adj` plane (unused by the forward integrators): { si.t of the previous intersection, tissue depth }.
// biovolpath.cpp:303-306 -- `result` cleared per channel where the iteration's MediumInteraction has transmittance 0, throughput
// multiplied by it -- runs in TRACK for lanes inside a medium and at the top of SHADE_A / ESCAPE for the others (their `mei` is zeros:
// note (1)).  Same PCG32 draws in the same order as BioPath::step: a wavefront render is lane-for-lane the per-lane kernel's.
#pragma once
#include "lr_wavefront.cuh"
#include "lr_biovolpath.cuh"

namespace lr {

// biovolpath.cpp:194-200 at the top of the next iteration, done by the stage that spawned the ray (no draw lies in between), plus the
// dying-iteration rule.  Returns true if the path goes on; otherwise `res` is final.
LR_DEV bool bio_next_iteration(const DScene &sc, Sampler &smp, V3 &thr, float eta, uint32_t depth, int medium, const Ray &ray, float si_t,
                               uint32_t channel, float tissue_depth, V3 &res, uint32_t &n_collisions) {
    bool active = any_nonzero(thr);
    const float q = fminf(hmax(thr) * sqr(eta), .95f);
    const bool perform_rr = depth > (uint32_t) sc.rr_depth;
    if (active) { const float u = smp.next_1d(); active &= (u < q) || !perform_rr; }
    if (perform_rr) thr *= 1.f / q;
    active &= depth < sc.max_depth;
    if (active) return true;
    BioInteraction mei;
    if (medium >= 0) { Sampler peek = smp; mei = bio_sample_interaction(sc.media[medium].m, ray, si_t, peek.next_1d(), channel, tissue_depth, false, n_collisions); }
    else mei = bio_zeros();
    bio_apply_transmittance(mei, medium, res, thr);
    return false;
}

template <int C, bool TO_FILM>
__global__ void __launch_bounds__(256, LR_WF_OCC) k_bio_track(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_track) {
        const bool item = qi < _n; const uint32_t slot = item ? a.q_track[qi] : 0;
        bool fin = false, to_trace = false, to_surface = false; uint32_t f = 0; Hit h; h.t = LR_INF; h.prim = 0xffffffffu; h.u = h.v = 0.f;
        V3 res = mk3(0.f);
        if (item) {
            Sampler smp = wv_rng(p, slot);
            const float4 o4 = p.o(slot), d4 = p.d(slot), t4 = p.thr(slot), r4 = p.res(slot), j4 = p.adj(slot);
            f = __float_as_uint(d4.w); uint32_t depth = __float_as_uint(t4.w); const int medium = wf_medium(f); const uint32_t channel = f & WF_CHANNEL_MASK;
            Ray ray; ray.o = mk3(o4.x, o4.y, o4.z); ray.d = mk3(d4.x, d4.y, d4.z); ray.maxt = o4.w;
            V3 thr = mk3(t4.x, t4.y, t4.z); res = mk3(r4.x, r4.y, r4.z); const float eta = r4.w;
            float tissue_depth = j4.y; h = wf_hit(p.hit(slot));
            const lr_medium &m = sc.media[medium].m;
            BioInteraction mei = bio_sample_interaction(m, ray, j4.x, smp.next_1d(), channel, tissue_depth, true, cnt.collisions);   // :222
            float si_t = h.t;
            if (bio_valid(mei) && !(h.t <= mei.t)) si_t = LR_INF;                     // ray.maxt = mei.t before the intersection (:224-227)
            if (si_t < mei.t) mei.t = LR_INF;
            if (m.has_spectral_extinction) {                                           // :231-235, medium.cpp:95-104
                const float t = fminf(mei.t, si_t);
                const V3 tr = exp3(mei.combined_extinction * (-t));
                const V3 pdf = si_t < mei.t ? tr : tr * mei.combined_extinction;
                const float tr_pdf = idx3(pdf, channel);
                thr *= tr_pdf > 0.f ? tr / tr_pdf : mk3(0.f);
            }
            if (bio_valid(mei)) {                                                      // scatter (the null-collision draw of :241 never succeeds)
                (void) smp.next_1d();
                depth += 1; cnt.real_scatters++;
                p.lsp(slot) = make_float4(mei.p.x, mei.p.y, mei.p.z, p.lsp(slot).w);
                if (depth < sc.max_depth) {
                    thr *= mei.transmittance;                                          // :266-275
                    tissue_depth += fabsf(__fmul_rn(-ray.d.z, mei.t));
                    Phase phase; phase.kind = m.phase; phase.g = m.phase_g;
                    (void) smp.next_1d(); const V2 s2 = smp.next_2d();
                    float phase_pdf; const V3 wo = phase.sample(mei.wi, s2, phase_pdf);
                    if (phase_pdf > 0.f) { ray.o = mei.p; ray.d = wo; ray.maxt = LR_LARGEST; reinterpret_cast<float *>(&p.lsp(slot))[3] = phase_pdf; }
                    bio_apply_transmittance(mei, medium, res, thr);                    // :303-306 (si_t: inf -- the shortened ray found nothing -- or, on a tie, mei.t)
                    to_trace = bio_next_iteration(sc, smp, thr, eta, depth, medium, ray, si_t, channel, tissue_depth, res, cnt.collisions);
                    fin = !to_trace;
                    if (to_trace) {
                        p.o(slot) = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.maxt);
                        p.d(slot) = make_float4(ray.d.x, ray.d.y, ray.d.z, d4.w);
                    }
                } else { bio_apply_transmittance(mei, medium, res, thr); fin = true; }  // depth limit (:256): the rest of the iteration is masked
            } else {                                                                   // the boundary (or nothing) is reached
                bio_apply_transmittance(mei, medium, res, thr);
                to_surface = true;
            }
            wv_rng_store(p, slot, smp);
            if (!fin) {
                p.thr(slot) = make_float4(thr.x, thr.y, thr.z, __uint_as_float(depth));
                p.res(slot) = make_float4(res.x, res.y, res.z, eta);
                p.adj(slot) = make_float4(si_t, tissue_depth, 0.f, 0.f);
            }
        }
        {
            float4 lp = make_float4(0.f, 0.f, 0.f, 0.f);
            if (fin) lp = p.lane(slot);
            wv_finish<C, TO_FILM>(sc, a, fin, lp, res, f);
        }
        const bool esc = to_surface && !hit_valid(h);
        const uint32_t b = (to_surface && !esc) ? wv_bucket(sc, h) : 0xffu;
        wf_push(a.q_regen, &a.q->n_regen, slot, fin && wv_more_passes(a, f));
        wf_push(a.q_trace, &a.q->n_trace, slot, to_trace);
        wf_push(a.q_escape, &a.q->n_escape, slot, esc);
#pragma unroll
        for (uint32_t k = 0; k < WF_N_SURF; ++k) wf_push(a.q_surf[k], &a.q->n_surf[k], slot, b == k);
    }
    flush_counters(cnt, 0, a.stats);
}

// surface (biovolpath.cpp:298-356): emitter hit + MIS, emitter sample + BSDF value; the shadow ray goes to TRACE_S / NEE
__global__ void __launch_bounds__(128, LR_WF_OCC_SURF) k_bio_shade_a(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a, int bucket) {
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_surf[bucket]) {
        const bool item = qi < _n; const uint32_t slot = item ? a.q_surf[bucket][qi] : 0; bool nee = false;
        if (item) {
            const float4 o4 = p.o(slot), d4 = p.d(slot), t4 = p.thr(slot);
            const uint32_t f = __float_as_uint(d4.w), depth = __float_as_uint(t4.w); const int medium = wf_medium(f);
            Ray ray; ray.o = mk3(o4.x, o4.y, o4.z); ray.d = mk3(d4.x, d4.y, d4.z); ray.maxt = o4.w;
            const V3 thr = mk3(t4.x, t4.y, t4.z);
            const SurfaceInteraction si = compute_si(sc, ray, wf_hit(p.hit(slot)), sc.tris_gid);
            float4 r4 = p.res(slot); bool res_dirty = false;
            if (medium < 0) { r4.x = r4.y = r4.z = 0.f; res_dirty = true; }            // :304 with mei = zeros (note (1)): what earlier iterations gathered is cleared
            const bool count_direct = depth == 0u || (f & WF_SPEC_CHAIN);
            const int emitter = emitter_of(sc, si);
            if (emitter >= 0 && !(depth == 0u && sc.hide_emitters)) {                  // :312-327
                const float4 l4 = p.lsp(slot);
                float emitter_pdf = 1.f;
                if (!count_direct) { DirectionSample ds = ds_from_si(sc, si, mk3(l4.x, l4.y, l4.z)); emitter_pdf = pdf_emitter_direction(sc, ds); }
                const V3 emitted = eval_emitter_at(sc, emitter, si);
                const V3 c = count_direct ? thr * emitted : thr * mis_weight(l4.w, emitter_pdf) * emitted;
                r4.x += c.x; r4.y += c.y; r4.z += c.z; res_dirty = true;
            }
            if (res_dirty) p.res(slot) = r4;
            BSDF bsdf(sc, si);
            if ((bsdf.flags() & F_SMOOTH) && (depth + 1 < sc.max_depth)) {             // :337-356
                Sampler smp = wv_rng(p, slot);
                V3 ev; const DirectionSample ds = sample_emitter_direction(sc, si.p, smp.next_2d(), ev);
                wv_rng_store(p, slot, smp);
                if (ds.pdf != 0.f) {
                    const Ray sr = spawn_ray_to(si.p, si.n, ds.p);
                    int walk_medium = medium;
                    if (mesh_is_transition(sc, si.mesh)) walk_medium = target_medium(sc, si, sr.d);
                    const V3 wo = si.sh.to_local(ds.d); float bsdf_pdf; const V3 bsdf_val = bsdf.eval_pdf(si.wi, wo, bsdf_pdf);
                    const V3 w = thr * bsdf_val * mis_weight(ds.pdf, ds.delta ? 0.f : bsdf_pdf) * ev;
                    p.no(slot) = make_float4(sr.o.x, sr.o.y, sr.o.z, sr.maxt);
                    p.nd(slot) = make_float4(sr.d.x, sr.d.y, sr.d.z, ds.dist);
                    p.nw(slot) = make_float4(w.x, w.y, w.z, __int_as_float(walk_medium));
                    nee = true;
                }
            }
        }
        wf_push(a.q_shadow, &a.q->n_shadow, slot, nee);
    }
}

// rays that left the scene (biovolpath.cpp:312-327 with an invalid si): environment radiance + MIS, then the path ends
template <int C, bool TO_FILM>
__global__ void __launch_bounds__(256, LR_WF_OCC) k_bio_escape(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_escape) {
        const bool item = qi < _n; const uint32_t slot = item ? a.q_escape[qi] : 0;
        V3 L = mk3(0.f); uint32_t f = 0; float4 lp = make_float4(0.f, 0.f, 0.f, 0.f);
        if (item) {
            const float4 d4 = p.d(slot), t4 = p.thr(slot), r4 = p.res(slot); lp = p.lane(slot);
            f = __float_as_uint(d4.w); const uint32_t depth = __float_as_uint(t4.w);
            if (wf_medium(f) >= 0) L = mk3(r4.x, r4.y, r4.z);                          // else :304 with mei = zeros: cleared
            if (sc.env >= 0 && !(depth == 0u && sc.hide_emitters)) {
                const V3 dir = mk3(d4.x, d4.y, d4.z), thr = mk3(t4.x, t4.y, t4.z);
                const bool count_direct = depth == 0u || (f & WF_SPEC_CHAIN);
                const V3 emitted = emitter_eval(sc, sc.env, -dir.z, dir);
                if (count_direct) L += thr * emitted;
                else {
                    DirectionSample ds; ds.d = dir; ds.emitter = sc.env; ds.dist = 0.f; ds.n = ds.p = mk3(0.f); ds.uv = mk2(0.f, 0.f); ds.pdf = 0.f; ds.delta = false;
                    L += thr * mis_weight(p.lsp(slot).w, pdf_emitter_direction(sc, ds)) * emitted;
                }
            }
        }
        wv_finish<C, TO_FILM>(sc, a, item, lp, L, f);
        wf_push(a.q_regen, &a.q->n_regen, slot, item && wv_more_passes(a, f));
    }
}

// NEE: the walk along the shadow segment (biovolpath.cpp:444-528) over the slots TRACE_S let through, starting from its closest hit
__global__ void __launch_bounds__(256, LR_WF_OCC) k_bio_nee(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_nee) {
        if (qi < _n) {
            const uint32_t slot = a.q_nee[qi];
            Sampler smp = wv_rng(p, slot);
            const uint32_t channel = __float_as_uint(p.d(slot).w) & WF_CHANNEL_MASK;
            const float4 o4 = p.no(slot), d4 = p.nd(slot), w4 = p.nw(slot);
            Ray ray; ray.o = mk3(o4.x, o4.y, o4.z); ray.d = mk3(d4.x, d4.y, d4.z); ray.maxt = o4.w;
            const Hit h = wf_hit(p.nhit(slot));
            const V3 tr = bio_nee_walk(sc, sc.tris_gid, ray, d4.w, __float_as_int(w4.w), channel, p.adj(slot).y, &h, smp, cnt);
            float4 r4 = p.res(slot);
            r4.x = fmaf(w4.x, tr.x, r4.x); r4.y = fmaf(w4.y, tr.y, r4.y); r4.z = fmaf(w4.z, tr.z, r4.z);
            p.res(slot) = r4;
            wv_rng_store(p, slot, smp);
        }
    }
    flush_counters(cnt, 0, a.stats);
}

// surface: BSDF sampling, next ray, medium transition (biovolpath.cpp:358-386), then the top of the next iteration
template <int C, bool TO_FILM>
__global__ void __launch_bounds__(128, LR_WF_OCC_SURF) k_bio_shade_b(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a, int bucket) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_surf[bucket]) {
        const bool item = qi < _n; const uint32_t slot = item ? a.q_surf[bucket][qi] : 0;
        bool alive = false; uint32_t f = 0; V3 L = mk3(0.f); float4 lp = make_float4(0.f, 0.f, 0.f, 0.f);
        if (item) {
            Sampler smp = wv_rng(p, slot);
            const float4 o4 = p.o(slot), d4 = p.d(slot), t4 = p.thr(slot), r4 = p.res(slot); f = __float_as_uint(d4.w);
            Ray ray; ray.o = mk3(o4.x, o4.y, o4.z); ray.d = mk3(d4.x, d4.y, d4.z); ray.maxt = o4.w;
            const Hit h = wf_hit(p.hit(slot));
            const SurfaceInteraction si = compute_si(sc, ray, h, sc.tris_gid);
            BSDF bsdf(sc, si);
            const float s1 = smp.next_1d(); const V2 s2 = smp.next_2d();
            V3 bsdf_val; const BSDFSample bs = bsdf.sample(si.wi, s1, s2, bsdf_val);
            V3 thr = mk3(t4.x, t4.y, t4.z) * bsdf_val; const float eta = r4.w * bs.eta; uint32_t depth = __float_as_uint(t4.w);
            ray = spawn_ray(si.p, si.n, si.sh.to_world(bs.wo));
            const bool non_null = !(bs.sampled_type & F_NULL);
            if (non_null) { depth += 1u; p.lsp(slot) = make_float4(si.p.x, si.p.y, si.p.z, bs.pdf); f |= WF_VALID_RAY; }
            bool spec = (f & WF_SPEC_CHAIN) != 0;
            spec |= non_null && (bs.sampled_type & F_DELTA); spec &= !(bs.sampled_type & F_SMOOTH);
            f = spec ? (f | WF_SPEC_CHAIN) : (f & ~WF_SPEC_CHAIN);
            if (mesh_is_transition(sc, si.mesh)) f = wf_set_medium(f, target_medium(sc, si, ray.d));
            const float tissue_depth = p.adj(slot).y;
            V3 res = mk3(r4.x, r4.y, r4.z);
            alive = bio_next_iteration(sc, smp, thr, eta, depth, wf_medium(f), ray, h.t, f & WF_CHANNEL_MASK, tissue_depth, res, cnt.collisions);
            wv_rng_store(p, slot, smp);
            if (alive) {
                p.o(slot) = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.maxt);
                p.d(slot) = make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(f));
                p.thr(slot) = make_float4(thr.x, thr.y, thr.z, __uint_as_float(depth));
                if (eta != r4.w) reinterpret_cast<float *>(&p.res(slot))[3] = eta;
                p.adj(slot) = make_float4(h.t, tissue_depth, 0.f, 0.f);               // si.t the next sample_interaction is handed (:222)
            } else { L = res; lp = p.lane(slot); }
        }
        wv_finish<C, TO_FILM>(sc, a, item && !alive, lp, L, f);
        wf_push(a.q_trace, &a.q->n_trace, slot, alive);
        wf_push(a.q_regen, &a.q->n_regen, slot, item && !alive && wv_more_passes(a, f));
    }
    flush_counters(cnt, 0, a.stats);
}

} // namespace lr
