// lr_kat.inl -- device known-answer probes (included by lr_backend.cu): the shading / lookup arithmetic of the hot path callable
// one query per thread, so that the reference's function-level tests (src/bsdfs/tests/test_principled.py:124-205,
// test_rough_dielectric.py:209-229, test_diffuse.py:14-37, src/volumes/tests/test_grid.py:43-118, src/phase/tests/*.py,
// src/emitters/tests/test_envmap.py) run on the CUDA code itself and not only on the oracle.  Same argument meaning as the
// oracle's orc_kat_* probes (oracle/oracle.cpp), batched.

__global__ void k_kat_bsdf_eval_pdf(const __grid_constant__ DScene sc, const lr_bsdf *b, const float *wi, const float *wo, uint64_t n, float *out) {
    uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; if (i >= n) return;
    BSDF bs(sc, *b, mk2(0.f, 0.f)); float pdf;
    V3 e = bs.eval_pdf_n(mk3(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]), mk3(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]), pdf);
    out[4 * i] = e.x; out[4 * i + 1] = e.y; out[4 * i + 2] = e.z; out[4 * i + 3] = pdf;
}
__global__ void k_kat_bsdf_sample(const __grid_constant__ DScene sc, const lr_bsdf *b, const float *wi, const float *u, uint64_t n, float *out) {
    uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; if (i >= n) return;
    BSDF bs(sc, *b, mk2(0.f, 0.f)); V3 w;
    BSDFSample r = bs.sample_n(mk3(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]), u[3 * i], mk2(u[3 * i + 1], u[3 * i + 2]), w);
    float *o = out + 9 * i; o[0] = r.wo.x; o[1] = r.wo.y; o[2] = r.wo.z; o[3] = r.pdf; o[4] = r.eta; o[5] = w.x; o[6] = w.y; o[7] = w.z; o[8] = __uint_as_float(r.sampled_type);
}
__global__ void k_kat_phase(uint32_t kind, float g, const float *wi, const float *u, uint64_t n, float *out) {
    uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; if (i >= n) return;
    Phase p; p.kind = kind; p.g = g; V3 w = mk3(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]); float pdf;
    V3 wo = p.sample(w, mk2(u[2 * i], u[2 * i + 1]), pdf);
    float *o = out + 5 * i; o[0] = wo.x; o[1] = wo.y; o[2] = wo.z; o[3] = pdf; o[4] = p.eval_pdf(w, wo);
}
__global__ void k_kat_grid_eval(const __grid_constant__ DScene sc, uint32_t grid, const float *p, uint64_t n, float *out, int use_fat) {
    uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; if (i >= n) return;
    DGrid g = sc.grids[grid]; if (!use_fat) g.fat = nullptr;
    V3 v = grid_eval<false>(g, mk3(p[3 * i], p[3 * i + 1], p[3 * i + 2]), nullptr);
    out[3 * i] = v.x; out[3 * i + 1] = v.y; out[3 * i + 2] = v.z;
}
// in[n][5] = ref_p(3), u(2) -> out[n][7] = d(3), pdf, weight(3)  (Scene::sample_emitter_direction on the scene's emitters, no visibility test)
__global__ void k_kat_emitter_sample(const __grid_constant__ DScene sc, const float *in, uint64_t n, float *out) {
    uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; if (i >= n) return;
    V3 w; DirectionSample ds = sample_emitter_direction(sc, mk3(in[5 * i], in[5 * i + 1], in[5 * i + 2]), mk2(in[5 * i + 3], in[5 * i + 4]), w);
    float *o = out + 7 * i; o[0] = ds.d.x; o[1] = ds.d.y; o[2] = ds.d.z; o[3] = ds.pdf; o[4] = w.x; o[5] = w.y; o[6] = w.z;
}
// d[n][3] = direction of travel of an escaping ray -> out[n][4] = environment radiance rgb, pdf_direction of that direction
__global__ void k_kat_env_eval_pdf(const __grid_constant__ DScene sc, const float *d, uint64_t n, float *out) {
    uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; if (i >= n) return;
    V3 dir = mk3(d[3 * i], d[3 * i + 1], d[3 * i + 2]); V3 e = mk3(0.f); float pdf = 0.f;
    if (sc.env >= 0) {
        e = emitter_eval(sc, sc.env, -dir.z, dir);
        DirectionSample ds; ds.d = dir; ds.emitter = sc.env; ds.dist = 0.f; ds.n = ds.p = mk3(0.f); ds.uv = mk2(0.f, 0.f); ds.pdf = 0.f; ds.delta = false;
        pdf = emitter_pdf_direction(sc, sc.env, ds);
    }
    out[4 * i] = e.x; out[4 * i + 1] = e.y; out[4 * i + 2] = e.z; out[4 * i + 3] = pdf;
}

// tissue media probe (lr_biovolpath.cuh): in[n][5] = sample, tissue depth, maxt, channel (bits), active (bits) ->
// out[n][8] = distance, element type (bits), mei.t, transmittance(3), sigma_t[channel], combined_extinction[channel]; ray o = 0, d = +z
__global__ void k_kat_tissue(const lr_medium *m, const float *in, uint64_t n, float *out) {
    uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; if (i >= n) return;
    const float *q = in + 5 * i; float *o = out + 8 * i; const uint32_t ch = __float_as_uint(q[3]);
    Ray ray; ray.o = mk3(0.f); ray.d = mk3(0.f, 0.f, 1.f); ray.maxt = LR_INF;
    int bt; o[0] = bio_compute_distance(*m, ch, q[0], q[1], bt); o[1] = __uint_as_float((uint32_t) bt);
    uint32_t cnt = 0;
    BioInteraction mei = bio_sample_interaction(*m, ray, q[2], q[0], ch, q[1], __float_as_uint(q[4]) != 0u, cnt);
    o[2] = mei.t; o[3] = mei.transmittance.x; o[4] = mei.transmittance.y; o[5] = mei.transmittance.z;
    o[6] = mei.sigma_t_ch; o[7] = idx3(mei.combined_extinction, ch);
}

// host -> device staging of the probe arrays; `outs` floats come back into `out`
struct KatIO {
    lr_ctx *c; std::vector<DevBuf *> bufs;
    ~KatIO() { for (auto *b : bufs) delete b; }
    template <typename T> const T *in(const T *host, size_t n) {
        DevBuf *b = new DevBuf(); bufs.push_back(b);
        if (b->alloc(std::max<size_t>(n, 1) * sizeof(T)) != cudaSuccess) return nullptr;
        if (n && cudaMemcpyAsync(b->p, host, n * sizeof(T), cudaMemcpyHostToDevice, c->stream) != cudaSuccess) return nullptr;
        return b->as<T>();
    }
    float *out(size_t n) { DevBuf *b = new DevBuf(); bufs.push_back(b); return b->alloc(std::max<size_t>(n, 1) * 4) == cudaSuccess ? b->as<float>() : nullptr; }
};
#define LR_KAT_BEGIN(what, need_scene)                                                                           \
    if (!c) return LR_ERR_INVALID_ARG;                                                                            \
    if ((need_scene) && !c->has_scene) LR_FAIL(c, LR_ERR_NO_SCENE, what ": no scene uploaded");                   \
    if (n == 0) return LR_OK;                                                                                     \
    LR_CUDA(c, cudaSetDevice(c->device)); KatIO io{c, {}}; const unsigned blocks = (unsigned) ((n + 127) / 128);
#define LR_KAT_END(dev_out, floats)                                                                               \
    LR_CUDA(c, cudaGetLastError()); return copy_out(c, out, dev_out, (size_t) n * (floats) * 4, 0);

extern "C" {

int lr_kat_bsdf_eval_pdf(lr_ctx *c, const lr_bsdf *b, const float *wi, const float *wo, uint64_t n, float *out) {
    LR_KAT_BEGIN("lr_kat_bsdf_eval_pdf", false)
    if (!b || !wi || !wo || !out) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_kat_bsdf_eval_pdf: NULL argument");
    if (b->tex_kind == 2 || b->bump_tex) LR_FAIL(c, LR_ERR_UNSUPPORTED, "lr_kat_bsdf_eval_pdf: textured BSDFs are probed through a scene");
    const lr_bsdf *db = io.in(b, 1); const float *dwi = io.in(wi, n * 3), *dwo = io.in(wo, n * 3); float *dout = io.out(n * 4);
    if (!db || !dwi || !dwo || !dout) LR_FAIL(c, LR_ERR_OOM, "lr_kat_bsdf_eval_pdf: device allocation failed");
    k_kat_bsdf_eval_pdf<<<blocks, 128, 0, c->stream>>>(c->sc, db, dwi, dwo, n, dout);
    LR_KAT_END(dout, 4)
}
int lr_kat_bsdf_sample(lr_ctx *c, const lr_bsdf *b, const float *wi, const float *u, uint64_t n, float *out) {
    LR_KAT_BEGIN("lr_kat_bsdf_sample", false)
    if (!b || !wi || !u || !out) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_kat_bsdf_sample: NULL argument");
    if (b->tex_kind == 2 || b->bump_tex) LR_FAIL(c, LR_ERR_UNSUPPORTED, "lr_kat_bsdf_sample: textured BSDFs are probed through a scene");
    const lr_bsdf *db = io.in(b, 1); const float *dwi = io.in(wi, n * 3), *du = io.in(u, n * 3); float *dout = io.out(n * 9);
    if (!db || !dwi || !du || !dout) LR_FAIL(c, LR_ERR_OOM, "lr_kat_bsdf_sample: device allocation failed");
    k_kat_bsdf_sample<<<blocks, 128, 0, c->stream>>>(c->sc, db, dwi, du, n, dout);
    LR_KAT_END(dout, 9)
}
int lr_kat_phase(lr_ctx *c, uint32_t kind, float g, const float *wi, const float *u, uint64_t n, float *out) {
    LR_KAT_BEGIN("lr_kat_phase", false)
    if (!wi || !u || !out || kind > LR_PHASE_HG) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_kat_phase: bad argument");
    const float *dwi = io.in(wi, n * 3), *du = io.in(u, n * 2); float *dout = io.out(n * 5);
    if (!dwi || !du || !dout) LR_FAIL(c, LR_ERR_OOM, "lr_kat_phase: device allocation failed");
    k_kat_phase<<<blocks, 128, 0, c->stream>>>(kind, g, dwi, du, n, dout);
    LR_KAT_END(dout, 5)
}
int lr_kat_grid_eval(lr_ctx *c, uint32_t grid, const float *p, uint64_t n, float *out, int use_fat) {
    LR_KAT_BEGIN("lr_kat_grid_eval", true)
    if (grid >= c->h_grids.size() || !p || !out) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_kat_grid_eval: bad argument");
    const float *dp = io.in(p, n * 3); float *dout = io.out(n * 3);
    if (!dp || !dout) LR_FAIL(c, LR_ERR_OOM, "lr_kat_grid_eval: device allocation failed");
    k_kat_grid_eval<<<blocks, 128, 0, c->stream>>>(c->sc, grid, dp, n, dout, use_fat);
    LR_KAT_END(dout, 3)
}
int lr_kat_emitter_sample(lr_ctx *c, const float *ref_p_u, uint64_t n, float *out) {
    LR_KAT_BEGIN("lr_kat_emitter_sample", true)
    if (!ref_p_u || !out) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_kat_emitter_sample: NULL argument");
    const float *din = io.in(ref_p_u, n * 5); float *dout = io.out(n * 7);
    if (!din || !dout) LR_FAIL(c, LR_ERR_OOM, "lr_kat_emitter_sample: device allocation failed");
    k_kat_emitter_sample<<<blocks, 128, 0, c->stream>>>(c->sc, din, n, dout);
    LR_KAT_END(dout, 7)
}
int lr_kat_tissue(lr_ctx *c, const lr_medium *m, const float *in, uint64_t n, float *out) {
    LR_KAT_BEGIN("lr_kat_tissue", false)
    if (!m || !in || !out) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_kat_tissue: NULL argument");
    if (m->kind < LR_MEDIUM_LIVER || m->kind > LR_MEDIUM_PARENCHYMA) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_kat_tissue: not a tissue medium");
    const lr_medium *dm = io.in(m, 1); const float *din = io.in(in, n * 5); float *dout = io.out(n * 8);
    if (!dm || !din || !dout) LR_FAIL(c, LR_ERR_OOM, "lr_kat_tissue: device allocation failed");
    k_kat_tissue<<<blocks, 128, 0, c->stream>>>(dm, din, n, dout);
    LR_KAT_END(dout, 8)
}
int lr_kat_env_eval_pdf(lr_ctx *c, const float *d, uint64_t n, float *out) {
    LR_KAT_BEGIN("lr_kat_env_eval_pdf", true)
    if (!d || !out) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_kat_env_eval_pdf: NULL argument");
    const float *dd = io.in(d, n * 3); float *dout = io.out(n * 4);
    if (!dd || !dout) LR_FAIL(c, LR_ERR_OOM, "lr_kat_env_eval_pdf: device allocation failed");
    k_kat_env_eval_pdf<<<blocks, 128, 0, c->stream>>>(c->sc, dd, n, dout);
    LR_KAT_END(dout, 4)
}

} // extern "C"
