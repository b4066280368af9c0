/*
 * TEST INFRASTRUCTURE ONLY -- CPU oracle, C entry points (`orc_*`).
 *
 * A plain C++ restatement of the reference algorithm for the path / volpath /
 * prbvolpath hot path (global-majorant delta tracking, reference draw order,
 * reference film splat), multi-threaded over lanes.  It consumes the same POD
 * scene description as the product's C ABI (include/lr_api.h) so the parity
 * tests feed both sides identical inputs.  Parity status: pinned against the
 * reference's own known-answer tests and golden EXRs (see tests/ and
 * DESIGN.md); prbvolpath GRADIENTS are unpinned by the reference itself
 *  and are pinned here by finite differences only.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.
 */
#include "kat_fn.h"
#include "orc_integrators.h"
#include <thread>
#include <mutex>
#include <string>

using namespace orc;

struct OrcScene {
    Scene sc;
    std::unique_ptr<Film> film;
    std::atomic<uint64_t> rays{0}, collisions{0}, real_scatters{0}, samples{0};
};

static std::string g_err;

static Texture2D make_texture(const lr_texture &t) {
    Texture2D x; x.w = (int) t.width; x.h = (int) t.height; x.ch = (int) t.channels; x.filter = t.filter; x.wrap = t.wrap;
    memcpy(x.to_uv, t.to_uv, sizeof x.to_uv); x.data.assign(t.data, t.data + (size_t) t.width * t.height * t.channels);
    return x;
}

static void build_scene(OrcScene &o, const lr_scene *s) {
    Scene &sc = o.sc;
    sc.sensor = s->sensor; sc.integ = s->integrator;
    memcpy(sc.cam_to_world.m, s->sensor.to_world, 64); memcpy(sc.sample_to_camera.m, s->sensor.sample_to_camera, 64);
    sc.bsdfs.assign(s->bsdfs, s->bsdfs + s->n_bsdfs);
    for (uint32_t i = 0; i < s->n_textures; ++i) sc.textures.push_back(make_texture(s->textures[i]));
    V3 lo(Inf), hi(-Inf);
    for (uint32_t i = 0; i < s->n_meshes; ++i) {
        const lr_mesh &m = s->meshes[i]; Mesh me;
        me.pos.resize(m.n_vertices);
        for (uint32_t v = 0; v < m.n_vertices; ++v) me.pos[v] = V3(m.positions[3 * v], m.positions[3 * v + 1], m.positions[3 * v + 2]);
        if (m.normals) { me.nrm.resize(m.n_vertices); for (uint32_t v = 0; v < m.n_vertices; ++v) me.nrm[v] = V3(m.normals[3 * v], m.normals[3 * v + 1], m.normals[3 * v + 2]); }
        if (m.texcoords) { me.uv.resize(m.n_vertices); for (uint32_t v = 0; v < m.n_vertices; ++v) me.uv[v] = {m.texcoords[2 * v], m.texcoords[2 * v + 1]}; }
        me.idx.assign(m.indices, m.indices + 3 * (size_t) m.n_faces);
        me.bsdf = m.bsdf; me.interior = m.interior_medium; me.exterior = m.exterior_medium; me.emitter = m.emitter; me.flip = m.flip_normals != 0;
        me.tri_offset = (uint32_t) sc.tris.size(); me.area = 0.f; me.area_cdf.resize(m.n_faces);
        for (uint32_t f = 0; f < m.n_faces; ++f) {
            Tri t; t.p0 = me.pos[me.idx[3 * f]]; t.p1 = me.pos[me.idx[3 * f + 1]]; t.p2 = me.pos[me.idx[3 * f + 2]]; t.mesh = i; t.face = f;
            sc.tris.push_back(t);
            me.area += 0.5f * norm(cross(t.p1 - t.p0, t.p2 - t.p0)); me.area_cdf[f] = me.area;
            for (const V3 *p : {&t.p0, &t.p1, &t.p2}) for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], (*p)[k]); hi[k] = std::max(hi[k], (*p)[k]); }
        }
        sc.meshes.push_back(std::move(me));
    }
    sc.bvh.build(sc.tris);
    /* Scene bbox -> bounding sphere (include/mitsuba/core/bbox.h:343-346) */
    if (!sc.tris.empty()) { sc.bs_center = (lo + hi) * 0.5f; sc.bs_radius = norm(sc.bs_center - hi); }
    for (uint32_t i = 0; i < s->n_grids; ++i) {
        const lr_grid &g = s->grids[i]; Grid gr;
        size_t n = (size_t) g.res[0] * g.res[1] * g.res[2] * g.channels;
        gr.data.assign(g.data, g.data + n); memcpy(gr.res, g.res, 12); gr.ch = g.channels; memcpy(gr.to_local.m, g.to_local, 64);
        gr.bmin = V3(g.bbox_min[0], g.bbox_min[1], g.bbox_min[2]); gr.bmax = V3(g.bbox_max[0], g.bbox_max[1], g.bbox_max[2]);
        gr.filter = g.filter; gr.wrap = g.wrap; gr.differentiable = g.differentiable != 0;
        gr.fixed_max = g.max_value >= 0.f;
        if (gr.fixed_max) gr.maxv = g.max_value; else gr.recompute_max();
        if (gr.differentiable) gr.grad.assign(n, 0.0);
        sc.grids.push_back(std::move(gr));
    }
    for (uint32_t i = 0; i < s->n_media; ++i) {
        Medium m; m.m = s->media[i];
        if (m.m.kind == LR_MEDIUM_HETEROGENEOUS) m.max_density = m.m.scale * sc.grids[m.m.sigma_t_grid].maxv;
        sc.media.push_back(m);
    }
    for (uint32_t i = 0; i < s->n_emitters; ++i) {
        Emitter e; e.e = s->emitters[i];
        if (e.e.kind == LR_EMITTER_ENVMAP) envmap_prepare(e);
        if (e.e.kind != LR_EMITTER_AREA) sc.env = (int) i;
        sc.emitters.push_back(std::move(e));
    }
    if (s->sensor.rfilter == LR_RFILTER_GAUSSIAN) gaussian_coeffs(s->sensor.rfilter_stddev, s->sensor.rfilter_radius, sc.gauss_coeff);
    uint32_t C = s->sensor.has_alpha ? 5 : 4;
    o.film.reset(new Film(s->sensor.film_width, s->sensor.film_height, C));
}

/* lane -> (pixel, stream): SamplingIntegrator::render JIT branch
   (src/render/integrator.cpp:275-338), PCG32Sampler::seed (src/render/sampler.cpp:129-148) */
struct LaneSetup { Sampler smp; V2 pos, sample_pos; Ray ray; };
static LaneSetup lane_setup(const Scene &sc, uint32_t seed, uint32_t spp, uint64_t lane) {
    LaneSetup L; uint32_t W = sc.sensor.film_width;
    if (sc.integ.sampler == LR_SAMPLER_LD) L.smp.seed_ld(sc.integ.base_seed, seed, (uint32_t) lane, spp, spp);
    else L.smp.seed(sc.integ.base_seed + seed, (uint32_t) lane);
    uint32_t pix = (uint32_t) (lane / spp), py = pix / W, px = pix - W * py;
    L.pos = {(float) px, (float) py};
    V2 j = L.smp.next_2d();
    L.sample_pos = {L.pos.x + j.x, L.pos.y + j.y};
    V2 scale = {1.f / (float) W, 1.f / (float) sc.sensor.film_height};            /* 1 / crop_size (:465) */
    float ox = (float) sc.sensor.crop_offset_x, oy = (float) sc.sensor.crop_offset_y; /* pos += crop_offset (:338); offset = -crop_offset * scale (:466) */
    V2 adjusted = {fmaf((L.pos.x + ox) + j.x, scale.x, -(ox * scale.x)), fmaf((L.pos.y + oy) + j.y, scale.y, -(oy * scale.y))};
    L.ray = camera_ray(sc, adjusted);
    return L;
}

template <typename F> static void parallel_lanes(uint64_t begin, uint64_t end, int threads, F f) {
    if (threads <= 0) threads = (int) std::thread::hardware_concurrency();
    if (threads < 1) threads = 1;
    uint64_t n = end - begin; if (n == 0) return;
    uint64_t chunk = std::max<uint64_t>(1024, (n + (uint64_t) threads * 16 - 1) / ((uint64_t) threads * 16));
    std::atomic<uint64_t> next{begin};
    auto worker = [&]() {
        for (;;) {
            uint64_t b = next.fetch_add(chunk); if (b >= end) break;
            uint64_t e = std::min(end, b + chunk);
            f(b, e);
        }
    };
    std::vector<std::thread> ts; for (int i = 1; i < threads; ++i) ts.emplace_back(worker);
    worker(); for (auto &t : ts) t.join();
}

static LaneResult run_lane(Integrator &in, LaneSetup &L) {
    const Scene &sc = in.sc;
    switch (sc.integ.kind) {
        case LR_INTEGRATOR_PATH: return in.sample_path(L.smp, L.ray);
        case LR_INTEGRATOR_VOLPATH: return in.sample_volpath(L.smp, L.ray, sc.sensor.medium);
        case LR_INTEGRATOR_BIOVOLPATH: return in.sample_biovolpath(L.smp, L.ray, sc.sensor.medium);
        case LR_INTEGRATOR_VOLPATHMIS: return in.sample_volpathmis(L.smp, L.ray, sc.sensor.medium);
        default: return in.sample_prb(true, L.smp, L.ray, V3(0.f), V3(0.f));
    }
}

extern "C" {

const char *orc_last_error() { return g_err.c_str(); }

void *orc_scene_create(const lr_scene *s) {
    try { auto *o = new OrcScene(); build_scene(*o, s); return o; }
    catch (std::exception &e) { g_err = e.what(); return nullptr; }
}
void orc_scene_destroy(void *h) { delete (OrcScene *) h; }

/* per-lane radiance: out[(lane-begin)*4] = R,G,B,valid */
int orc_sample_lanes(void *h, uint32_t seed, uint32_t spp, uint64_t begin, uint64_t end, float *out, int threads) {
    OrcScene &o = *(OrcScene *) h;
    parallel_lanes(begin, end, threads, [&](uint64_t b, uint64_t e) {
        Integrator in(o.sc); if (o.sc.integ.kind == LR_INTEGRATOR_PRBVOLPATH) in.prb_prepare();
        for (uint64_t lane = b; lane < e; ++lane) {
            LaneSetup L = lane_setup(o.sc, seed, spp, lane);
            LaneResult r = run_lane(in, L);
            float *p = out + (lane - begin) * 4; p[0] = r.L.x; p[1] = r.L.y; p[2] = r.L.z; p[3] = r.valid ? 1.f : 0.f;
        }
        o.rays += in.cnt.rays; o.collisions += in.cnt.collisions; o.real_scatters += in.cnt.real_scatters; o.samples += e - b;
    });
    return 0;
}

void orc_film_clear(void *h) { ((OrcScene *) h)->film->clear(); }

/* accumulate lanes [begin,end) into the film (render_sample, integrator.cpp:449-521) */
int orc_render_lanes(void *h, uint32_t seed, uint32_t spp, uint64_t begin, uint64_t end, int threads) {
    OrcScene &o = *(OrcScene *) h; const Scene &sc = o.sc;
    bool box = sc.sensor.rfilter == LR_RFILTER_BOX, coalesce = spp >= 4, alpha = sc.sensor.has_alpha != 0;
    parallel_lanes(begin, end, threads, [&](uint64_t b, uint64_t e) {
        Integrator in(sc); if (sc.integ.kind == LR_INTEGRATOR_PRBVOLPATH) in.prb_prepare();
        for (uint64_t lane = b; lane < e; ++lane) {
            LaneSetup L = lane_setup(sc, seed, spp, lane);
            LaneResult r = run_lane(in, L);
            float aovs[5]; aovs[0] = r.L.x; aovs[1] = r.L.y; aovs[2] = r.L.z;
            if (alpha) { aovs[3] = r.valid ? 1.f : 0.f; aovs[4] = 1.f; } else aovs[3] = 1.f;
            film_put(sc, *o.film, box ? L.pos : L.sample_pos, aovs, coalesce);
        }
        o.rays += in.cnt.rays; o.collisions += in.cnt.collisions; o.real_scatters += in.cnt.real_scatters; o.samples += e - b;
    });
    return 0;
}
int orc_film_channels(void *h) { return (int) ((OrcScene *) h)->film->C; }
void orc_film_read(void *h, float *out) {
    Film &f = *((OrcScene *) h)->film; for (size_t i = 0; i < f.acc.size(); ++i) out[i] = (float) f.acc[i].load();
}
void orc_film_develop(void *h, float *out) { film_develop(*((OrcScene *) h)->film, out); }

/* Integrator::render: all W*H*spp lanes, then develop */
int orc_render(void *h, uint32_t seed, uint32_t spp, float *out, int threads) {
    OrcScene &o = *(OrcScene *) h; o.film->clear();
    uint64_t n = (uint64_t) o.sc.sensor.film_width * o.sc.sensor.film_height * spp;
    orc_render_lanes(h, seed, spp, 0, n, threads);
    film_develop(*o.film, out); return 0;
}

/* RBIntegrator.render_backward (ad/integrators/common.py:625-783) for prbvolpath:
   (0) splat weights to get W per pixel, dL_lane = sum_px w(px) * grad_in[px] / W[px];
   (1) primal pass; (2) adjoint replay with the same stream. */
int orc_render_backward(void *h, uint32_t seed, uint32_t spp, const float *grad_in, uint64_t begin, uint64_t end, int threads) {
    OrcScene &o = *(OrcScene *) h; const Scene &sc = o.sc;
    uint32_t W = sc.sensor.film_width, H = sc.sensor.film_height; uint64_t n = (uint64_t) W * H * spp;
    if (end == 0) { begin = 0; end = n; }
    bool box = sc.sensor.rfilter == LR_RFILTER_BOX, coalesce = spp >= 4; uint32_t gc = sc.sensor.has_alpha ? 4 : 3;
    /* pass 0: weight image over ALL lanes (the film normalisation is global) */
    Film wf(W, H, 1);
    parallel_lanes(0, n, threads, [&](uint64_t b, uint64_t e) {
        for (uint64_t lane = b; lane < e; ++lane) { LaneSetup L = lane_setup(sc, seed, spp, lane); float one = 1.f; film_put(sc, wf, box ? L.pos : L.sample_pos, &one, coalesce); }
    });
    parallel_lanes(begin, end, threads, [&](uint64_t b, uint64_t e) {
        Integrator in(sc); in.prb_prepare();
        for (uint64_t lane = b; lane < e; ++lane) {
            LaneSetup L = lane_setup(sc, seed, spp, lane);
            Footprint fp; V3 dL(0.f);
            if (film_footprint(sc, box ? L.pos : L.sample_pos, coalesce, fp))
                for (int j = 0; j < fp.ny; ++j) { int y = fp.y0 + j; if (y < 0 || y >= (int) H) continue;
                    for (int i = 0; i < fp.nx; ++i) { int x = fp.x0 + i; if (x < 0 || x >= (int) W) continue;
                        float w = fp.wy[j] * fp.wx[i]; float wsum = (float) wf.acc[(size_t) y * W + x].load(); float d = wsum == 0.f ? 1.f : wsum;
                        const float *g = grad_in + ((size_t) y * W + x) * gc;
                        dL += V3(g[0], g[1], g[2]) * (w / d); } }
            Sampler s1 = L.smp;                                 /* sampler.clone() for the primal pass */
            LaneResult p = in.sample_prb(true, s1, L.ray, V3(0.f), V3(0.f));
            in.sample_prb(false, L.smp, L.ray, dL, p.L);
        }
        o.rays += in.cnt.rays; o.collisions += in.cnt.collisions; o.real_scatters += in.cnt.real_scatters; o.samples += e - b;
    });
    return 0;
}
size_t orc_grid_size(void *h, uint32_t g) { return ((OrcScene *) h)->sc.grids[g].data.size(); }
void orc_grid_grad_zero(void *h, uint32_t g) { auto &v = ((OrcScene *) h)->sc.grids[g].grad; std::fill(v.begin(), v.end(), 0.0); }
void orc_grid_grad_read(void *h, uint32_t g, float *out) { auto &v = ((OrcScene *) h)->sc.grids[g].grad; for (size_t i = 0; i < v.size(); ++i) out[i] = (float) v[i]; }
void orc_grid_update(void *h, uint32_t g, const float *data) {
    OrcScene &o = *(OrcScene *) h; Grid &gr = o.sc.grids[g];
    std::copy(data, data + gr.data.size(), gr.data.begin()); if (!gr.fixed_max) gr.recompute_max();   /* max_value override survives updates (grid.cpp:314-327) */
    for (Medium &m : o.sc.media) if (m.m.kind == LR_MEDIUM_HETEROGENEOUS && m.m.sigma_t_grid == (int) g) m.max_density = m.m.scale * gr.maxv;
}

/* rays[n][8] -> hits[n][4] = t, prim (bit pattern), u, v */
int orc_ray_intersect(void *h, const float *rays, uint64_t n, float *hits) {
    OrcScene &o = *(OrcScene *) h;
    for (uint64_t i = 0; i < n; ++i) {
        const float *r = rays + i * 8; Ray ray{V3(r[0], r[1], r[2]), V3(r[3], r[4], r[5]), r[6]};
        Hit hh = o.sc.bvh.intersect(ray); float *out = hits + i * 4;
        out[0] = hh.t; uint32_t pr = hh.prim; memcpy(out + 1, &pr, 4); out[2] = hh.u; out[3] = hh.v;
    }
    return 0;
}
void orc_stats(void *h, uint64_t *out4) {
    OrcScene &o = *(OrcScene *) h; out4[0] = o.samples; out4[1] = o.rays; out4[2] = o.collisions; out4[3] = o.real_scatters;
}
void orc_stats_reset(void *h) { OrcScene &o = *(OrcScene *) h; o.samples = 0; o.rays = 0; o.collisions = 0; o.real_scatters = 0; }

/* ---- known-answer probes (checked against the reference's own tests) ------- */
void orc_kat_tea32(uint32_t v0, uint32_t v1, int rounds, uint32_t *out2) { tea32(v0, v1, rounds, out2[0], out2[1]); }
float orc_kat_tea_float32(uint32_t v0, uint32_t v1, int rounds) { return tea_float32(v0, v1, rounds); }
void orc_kat_pcg32_floats(uint64_t initstate, uint64_t initseq, int n, float *out) { PCG32 r; r.seed(initstate, initseq); for (int i = 0; i < n; ++i) out[i] = r.next_f32(); }
void orc_kat_pcg32_uints(uint64_t initstate, uint64_t initseq, int n, uint32_t *out) { PCG32 r; r.seed(initstate, initseq); for (int i = 0; i < n; ++i) out[i] = r.next_u32(); }
/* scalar-variant ldsampler (wavefront size 1, samples_per_wavefront 1): n1 next_1d then n2 next_2d per sample index, `advances` times */
void orc_kat_ldsampler(uint32_t base_seed, uint32_t seed, uint32_t sample_count, int n1, int n2, int advances, float *out) {
    Sampler s; s.seed_ld(base_seed, seed, 0, 1, sample_count);
    for (int a = 0; a <= advances; ++a) {
        if (a) s.advance(1);
        for (int i = 0; i < n1; ++i) *out++ = s.next_1d();
        for (int i = 0; i < n2; ++i) { V2 v = s.next_2d(); *out++ = v.x; *out++ = v.y; }
    }
}
void orc_kat_sampler_floats(uint32_t seed, uint32_t lane, int n, float *out) { Sampler s; s.seed(seed, lane); for (int i = 0; i < n; ++i) out[i] = s.next_1d(); }
void orc_kat_bsdf_eval_pdf(const lr_bsdf *b, const float *wi, const float *wo, float *eval3, float *pdf) {
    BSDF bs(*b); V3 e = bs.eval(V3(wi[0], wi[1], wi[2]), V3(wo[0], wo[1], wo[2])); eval3[0] = e.x; eval3[1] = e.y; eval3[2] = e.z;
    *pdf = bs.pdf(V3(wi[0], wi[1], wi[2]), V3(wo[0], wo[1], wo[2]));
}
/* out[8] = wo(3), pdf, eta, weight(3); returns sampled_type */
uint32_t orc_kat_bsdf_sample(const lr_bsdf *b, const float *wi, float s1, float s2x, float s2y, float *out) {
    BSDF bs(*b); V3 w; BSDFSample r = bs.sample(V3(wi[0], wi[1], wi[2]), s1, {s2x, s2y}, w);
    out[0] = r.wo.x; out[1] = r.wo.y; out[2] = r.wo.z; out[3] = r.pdf; out[4] = r.eta; out[5] = w.x; out[6] = w.y; out[7] = w.z; return r.sampled_type;
}
/* bitmap texture probes: out[6] = eval rgb(3), eval_1, eval_1_grad(2) at each of the n uv pairs */
void orc_kat_texture(const lr_texture *t, int n, const float *uv, float *out) {
    Texture2D x = make_texture(*t);
    for (int i = 0; i < n; ++i) {
        V2 p{uv[2 * i], uv[2 * i + 1]}; V3 c = x.eval(p); V2 g = x.eval_1_grad(p);
        out[6 * i] = c.x; out[6 * i + 1] = c.y; out[6 * i + 2] = c.z; out[6 * i + 3] = x.eval_1(p); out[6 * i + 4] = g.x; out[6 * i + 5] = g.y;
    }
}
/* bumpmap: frame (n, s, t; 9 floats) perturbed by texture `t` for a synthetic interaction:
   si = {n, sh.n, sh.s (t = n x s), dp_du, dp_dv, wi (local), uv}, 20 floats */
void orc_kat_bump_frame(const lr_bsdf *b, const lr_texture *t, const float *sif, float *out9) {
    Texture2D x = make_texture(*t); SurfaceInteraction si;
    si.n = V3(sif[0], sif[1], sif[2]); si.sh.n = V3(sif[3], sif[4], sif[5]); si.sh.s = V3(sif[6], sif[7], sif[8]); si.sh.t = cross(si.sh.n, si.sh.s);
    si.dp_du = V3(sif[9], sif[10], sif[11]); si.dp_dv = V3(sif[12], sif[13], sif[14]); si.wi = V3(sif[15], sif[16], sif[17]); si.uv = {sif[18], sif[19]};
    Frame f = BSDF::bump_frame(x, *b, si);
    out9[0] = f.n.x; out9[1] = f.n.y; out9[2] = f.n.z; out9[3] = f.s.x; out9[4] = f.s.y; out9[5] = f.s.z; out9[6] = f.t.x; out9[7] = f.t.y; out9[8] = f.t.z;
}
/* out[4] = wo(3), pdf */
void orc_kat_phase_sample(uint32_t kind, float g, const float *wi, float u1, float u2, float *out) {
    Phase p{kind, g}; float pdf; V3 wo = p.sample(V3(wi[0], wi[1], wi[2]), {u1, u2}, pdf); out[0] = wo.x; out[1] = wo.y; out[2] = wo.z; out[3] = pdf;
}
float orc_kat_phase_eval(uint32_t kind, float g, const float *wi, const float *wo) { Phase p{kind, g}; return p.eval_pdf(V3(wi[0], wi[1], wi[2]), V3(wo[0], wo[1], wo[2])); }
void orc_kat_grid_eval(void *h, uint32_t g, const float *p, float *out) { ((OrcScene *) h)->sc.grids[g].eval(V3(p[0], p[1], p[2]), out); }
float orc_kat_rfilter_eval(void *h, float x) { return rfilter_eval(((OrcScene *) h)->sc, x); }
void orc_kat_camera_ray(void *h, float x, float y, float *out7) {
    Ray r = camera_ray(((OrcScene *) h)->sc, {x, y}); out7[0] = r.o.x; out7[1] = r.o.y; out7[2] = r.o.z; out7[3] = r.d.x; out7[4] = r.d.y; out7[5] = r.d.z; out7[6] = r.maxt;
}
/* envmap: eval(direction of travel d) -> rgb ; sample(u) -> d(3), pdf, weight(3) ; pdf(d) */
void orc_kat_env_eval(void *h, const float *d, float *rgb) {
    OrcScene &o = *(OrcScene *) h; EmitterOps em(o.sc); SurfaceInteraction si; si.wi = -V3(d[0], d[1], d[2]);
    V3 r = em.eval(o.sc.env, si); rgb[0] = r.x; rgb[1] = r.y; rgb[2] = r.z;
}
void orc_kat_env_sample(void *h, const float *ref_p, float u1, float u2, float *out7) {
    OrcScene &o = *(OrcScene *) h; EmitterOps em(o.sc); Interaction it; it.p = V3(ref_p[0], ref_p[1], ref_p[2]); V3 w;
    DirectionSample ds = em.sample_direction(o.sc.env, it, {u1, u2}, w);
    out7[0] = ds.d.x; out7[1] = ds.d.y; out7[2] = ds.d.z; out7[3] = ds.pdf; out7[4] = w.x; out7[5] = w.y; out7[6] = w.z;
}
float orc_kat_env_pdf(void *h, const float *d) {
    OrcScene &o = *(OrcScene *) h; EmitterOps em(o.sc); Interaction it; DirectionSample ds; ds.d = V3(d[0], d[1], d[2]); ds.emitter = o.sc.env;
    return em.pdf_direction(o.sc.env, it, ds);
}
float orc_kat_erfinv(float x) { return erfinv_(x); }

/* generic probe with the same function ids as oracle/_ref/libref_kat.so (the reference's own headers): kat_fn.h */
/* tissue media probe: in[n][5] = sample, tissue depth, maxt, channel (bits), active (bits);
   out[n][8] = distance, element type (bits), mei.t, transmittance(3), sigma_t[channel], combined_extinction[channel].
   Ray: o = 0, d = +z.  Same layout as the backend's lr_kat_tissue. */
int orc_kat_tissue(const lr_medium *m, const float *in, uint64_t n, float *out) {
    auto bits = [](float f) { uint32_t u; memcpy(&u, &f, 4); return u; };
    auto fbits = [](uint32_t u) { float f; memcpy(&f, &u, 4); return f; };
    Scene sc; Medium md; md.m = *m; sc.media.push_back(md); Integrator integ(sc);
    Ray ray; ray.o = V3(0.f); ray.d = V3(0.f, 0.f, 1.f); ray.maxt = Inf;
    for (uint64_t i = 0; i < n; ++i) {
        const float *q = in + 5 * i; float *o = out + 8 * i; uint32_t ch = bits(q[3]);
        int bt; o[0] = Integrator::bio_compute_distance(*m, ch, q[0], q[1], bt); o[1] = fbits((uint32_t) bt);
        auto mei = integ.bio_sample_interaction(0, ray, q[2], q[0], ch, q[1], bits(q[4]) != 0);
        o[2] = mei.t; o[3] = mei.transmittance.x; o[4] = mei.transmittance.y; o[5] = mei.transmittance.z;
        o[6] = index_spectrum(mei.sigma_t, ch); o[7] = index_spectrum(mei.combined_extinction, ch);
    }
    return 0;
}

int orc_kat_fn(int fn, const float *in, float *out) {
    auto bits = [](float f) { uint32_t u; memcpy(&u, &f, 4); return u; };
    auto fbits = [](uint32_t u) { float f; memcpy(&f, &u, 4); return f; };
    auto v3 = [](const float *p) { return V3(p[0], p[1], p[2]); };
    auto put3 = [](float *o, V3 v) { o[0] = v.x; o[1] = v.y; o[2] = v.z; };
    auto mf = [](const float *i) { return Microfacet(i[0] == 0.f ? LR_MF_BECKMANN : LR_MF_GGX, i[1], i[2], i[3] != 0.f); };
    switch (fn) {
        case KAT_TEA32: { uint32_t a0, a1; tea32(bits(in[0]), bits(in[1]), (int) bits(in[2]), a0, a1); out[0] = fbits(a0); out[1] = fbits(a1); return 2; }
        case KAT_TEA_FLOAT32: out[0] = tea_float32(bits(in[0]), bits(in[1]), (int) bits(in[2])); return 1;
        case KAT_PCG32: {
            PCG32 rng; rng.seed(((uint64_t) bits(in[1]) << 32) | bits(in[0]), ((uint64_t) bits(in[3]) << 32) | bits(in[2]));
            for (int i = 0; i < 4; ++i) out[i] = rng.next_f32();
            for (int i = 0; i < 4; ++i) out[4 + i] = fbits(rng.next_u32());
            return 8; }
        case KAT_WARP_DISK: { V2 p = square_to_uniform_disk_concentric(V2{in[0], in[1]}); out[0] = p.x; out[1] = p.y; return 2; }
        case KAT_WARP_COS_HEMI: { V3 v = square_to_cosine_hemisphere(V2{in[0], in[1]}); put3(out, v); out[3] = square_to_cosine_hemisphere_pdf(v); return 4; }
        case KAT_WARP_SPHERE: put3(out, square_to_uniform_sphere(V2{in[0], in[1]})); return 3;
        case KAT_WARP_TRIANGLE: { V2 p = square_to_uniform_triangle(V2{in[0], in[1]}); out[0] = p.x; out[1] = p.y; return 2; }
        case KAT_FRESNEL: fresnel(in[0], in[1], out[0], out[1], out[2], out[3]); return 4;
        case KAT_REFLECT: put3(out, reflect(v3(in), v3(in + 3))); return 3;
        case KAT_REFRACT: put3(out, refract(v3(in), v3(in + 3), in[6], in[7])); return 3;
        case KAT_MF_EVAL: out[0] = mf(in).eval(v3(in + 4)); return 1;
        case KAT_MF_PDF: out[0] = mf(in).pdf(v3(in + 4), v3(in + 7)); return 1;
        case KAT_MF_SAMPLE: { V3 m; float pdf; mf(in).sample(v3(in + 4), V2{in[7], in[8]}, m, pdf); put3(out, m); out[3] = pdf; return 4; }
        case KAT_MF_G1: out[0] = mf(in).smith_g1(v3(in + 4), v3(in + 7)); return 1;
        case KAT_MF_G: out[0] = mf(in).G(v3(in + 4), v3(in + 7), v3(in + 10)); return 1;
        case KAT_GTR1: { GTR1 g{in[0]}; out[0] = g.eval(v3(in + 1)); out[1] = g.pdf(v3(in + 1)); put3(out + 2, g.sample(V2{in[4], in[5]})); return 5; }
        case KAT_SMITH_GGX1: out[0] = smith_ggx1(v3(in), v3(in + 3), in[6]); out[1] = clearcoat_G(v3(in), v3(in + 7), v3(in + 3), in[6]); return 2;
        case KAT_SCHLICK: out[0] = schlick_weight(in[0]); out[1] = calc_schlick1(in[1], in[0], in[2]); out[2] = schlick_R0_eta(in[2]); return 3;
        case KAT_MAC_MIC: out[0] = mac_mic_compat(v3(in), v3(in + 3), v3(in + 6), in[9], true) ? 1.f : 0.f;
                          out[1] = mac_mic_compat(v3(in), v3(in + 3), v3(in + 6), in[9], false) ? 1.f : 0.f; return 2;
        case KAT_PRINCIPLED_FRESNEL: put3(out, principled_fresnel(in[0], in[1], in[2], v3(in + 3), in[6], in[7], in[8] != 0.f, in[9], in[10], in[11] != 0.f, in[12] != 0.f)); return 3;
        case KAT_DIST_PARAMS: calc_dist_params(in[0], in[1], in[2] != 0.f, out[0], out[1]); return 2;
        case KAT_COORD_SYSTEM: { V3 s_, t_; coordinate_system(v3(in), s_, t_); put3(out, s_); put3(out + 3, t_); return 6; }
        case KAT_ERFINV: out[0] = erfinv_(in[0]); return 1;
    }
    return -1;
}

} // extern "C"
