// lr_backend_api.inl -- scene upload and the render / adjoint entry points (included by lr_backend.cu)

// ---------------------------------------------------------------------------------------------
// Sealed media.  In production tracking (LR_TRACKING_DDA) a shadow segment that meets a non-null surface contributes exactly 0
// (eval_null_transmission, volpath.cpp:528-543).  When a medium can only be entered through closed, opaque (non-null BSDF)
// surfaces and no emitter lies inside them, that is the fate of EVERY shadow ray that starts in it, so emitter sampling at its
// scatter vertices -- and the BSDF value / shadow ray of boundary vertices whose light direction points into it -- can be
// skipped without changing a single path.  Conditions checked here, all conservative (a `false` only costs speed):
//   * the medium is the interior of at least one mesh, and every mesh that names it as interior is a closed, consistently
//     oriented manifold (every directed edge has exactly one opposite; vertices welded by position) with a non-null BSDF
//     and without an area emitter
//   * meshes that name it as exterior (nested shells) lie inside the bounding box of one of those enclosing meshes
//   * the sensor is not inside it, and no area emitter's bounding box touches an enclosing mesh's bounding box
// ---------------------------------------------------------------------------------------------
static bool mesh_is_closed(const lr_mesh &m) {
    if (m.n_faces < 4) return false;
    // weld vertices by position bits
    std::vector<uint32_t> order(m.n_vertices), canon(m.n_vertices);
    for (uint32_t i = 0; i < m.n_vertices; ++i) order[i] = i;
    auto key = [&](uint32_t i) { return m.positions + 3 * (size_t) i; };
    std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return memcmp(key(a), key(b), 12) < 0; });
    for (uint32_t k = 0; k < m.n_vertices; ++k) canon[order[k]] = (k && memcmp(key(order[k]), key(order[k - 1]), 12) == 0) ? canon[order[k - 1]] : order[k];
    std::vector<uint64_t> edges; edges.reserve((size_t) m.n_faces * 3);
    for (uint32_t f = 0; f < m.n_faces; ++f) {
        uint32_t v[3] = { canon[m.indices[3 * (size_t) f]], canon[m.indices[3 * (size_t) f + 1]], canon[m.indices[3 * (size_t) f + 2]] };
        if (v[0] == v[1] || v[1] == v[2] || v[0] == v[2]) return false;                       // degenerate face
        for (int k = 0; k < 3; ++k) edges.push_back(((uint64_t) v[k] << 32) | v[(k + 1) % 3]);
    }
    std::sort(edges.begin(), edges.end());
    for (size_t i = 1; i < edges.size(); ++i) if (edges[i] == edges[i - 1]) return false;  // a directed edge used twice
    for (uint64_t e : edges) { uint64_t opp = (e << 32) | (e >> 32); if (!std::binary_search(edges.begin(), edges.end(), opp)) return false; }
    return true;
}
static void mesh_bbox(const lr_mesh &m, float lo[3], float hi[3]) {
    for (int k = 0; k < 3; ++k) { lo[k] = INFINITY; hi[k] = -INFINITY; }
    for (uint32_t i = 0; i < m.n_vertices; ++i) for (int k = 0; k < 3; ++k) { float v = m.positions[3 * (size_t) i + k]; lo[k] = std::min(lo[k], v); hi[k] = std::max(hi[k], v); }
}
static void seal_media(const lr_scene *s, std::vector<DMedium> &media) {
    for (auto &m : media) m.sealed = 0;
    if (getenv("LR_NO_SEAL")) return;
    std::vector<char> closed(s->n_meshes, 0);
    for (uint32_t j = 0; j < s->n_meshes; ++j) {
        const lr_mesh &me = s->meshes[j];
        if (me.interior_medium >= 0 && s->bsdfs[me.bsdf].kind != LR_BSDF_NULL && me.emitter < 0) closed[j] = mesh_is_closed(me) ? 1 : 0;
    }
    for (uint32_t i = 0; i < s->n_media; ++i) {
        if (s->sensor.medium == (int) i) continue;
        bool ok = true, any = false;
        for (uint32_t j = 0; j < s->n_meshes && ok; ++j) if (s->meshes[j].interior_medium == (int) i) { any = true; ok = closed[j] != 0; }
        if (!ok || !any) continue;
        auto inside_an_encloser = [&](const float lo[3], const float hi[3], bool strictly) {
            for (uint32_t j = 0; j < s->n_meshes; ++j) if (s->meshes[j].interior_medium == (int) i) {
                float elo[3], ehi[3]; mesh_bbox(s->meshes[j], elo, ehi); bool in = true, touch = true;
                for (int k = 0; k < 3; ++k) { in = in && lo[k] >= elo[k] && hi[k] <= ehi[k]; touch = touch && lo[k] <= ehi[k] && hi[k] >= elo[k]; }
                if (strictly ? in : touch) return true;
            }
            return false;
        };
        for (uint32_t j = 0; j < s->n_meshes && ok; ++j) {
            const lr_mesh &me = s->meshes[j]; float lo[3], hi[3]; mesh_bbox(me, lo, hi);
            if (me.exterior_medium == (int) i && me.interior_medium != (int) i && !inside_an_encloser(lo, hi, true)) ok = false;
            if (me.emitter >= 0 && inside_an_encloser(lo, hi, false)) ok = false;
        }
        if (ok) media[i].sealed = 1;
    }
}

extern "C" {

int lr_scene_upload(lr_ctx *c, const lr_scene *s) {
    if (!c || !s) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_scene_upload: NULL argument");
    if (s->api_version != LR_API_VERSION) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_scene_upload: api_version %u != %u", s->api_version, LR_API_VERSION);
    if (s->n_bsdfs > LR_MAX_BSDFS || s->n_media > LR_MAX_MEDIA || s->n_grids > LR_MAX_GRIDS || s->n_emitters > LR_MAX_EMITTERS)
        LR_FAIL(c, LR_ERR_UNSUPPORTED, "lr_scene_upload: too many bsdfs/media/grids/emitters");
    if (s->integrator.kind > LR_INTEGRATOR_VOLPATHMIS) LR_FAIL(c, LR_ERR_UNSUPPORTED, "lr_scene_upload: unknown integrator kind %u", s->integrator.kind);
    LR_CUDA(c, cudaSetDevice(c->device));
    c->free_scene();
    DScene &sc = c->sc; memset(&sc, 0, sizeof sc);
    sc.sensor = s->sensor; sc.integ = s->integrator;
    if (sc.integ.kind == LR_INTEGRATOR_BIOVOLPATH || sc.integ.kind == LR_INTEGRATOR_VOLPATHMIS) sc.integ.tracking = LR_TRACKING_REFERENCE;   // no production-tracking variant of these two
    mat_copy(sc.cam_to_world, s->sensor.to_world); mat_copy(sc.sample_to_camera, s->sensor.sample_to_camera);
    sc.max_depth = s->integrator.max_depth < 0 ? 0xffffffffu : (uint32_t) s->integrator.max_depth;
    sc.rr_depth = s->integrator.rr_depth; sc.hide_emitters = s->integrator.hide_emitters;
    if (s->sensor.rfilter == LR_RFILTER_GAUSSIAN) gaussian_coeffs(s->sensor.rfilter_stddev, s->sensor.rfilter_radius, sc.gauss_coeff);
    if (s->sensor.rfilter_radius > 7.5f) LR_FAIL(c, LR_ERR_UNSUPPORTED, "lr_scene_upload: reconstruction filter radius > 7.5 px is outside the hot path");
    if (s->sensor.rfilter_radius > 2.5f && s->sensor.rfilter != LR_RFILTER_BOX && s->integrator.kind == LR_INTEGRATOR_PRBVOLPATH)
        LR_FAIL(c, LR_ERR_UNSUPPORTED, "lr_scene_upload: prbvolpath gathers dL/dI through the 6-tap footprint: reconstruction filter radius > 2.5 px is not supported by the adjoint");
    LR_CUDA(c, c->scratch.alloc(256));

    // ---- meshes: concatenate triangles (gid order), upload per-mesh attribute arrays ----
    std::vector<DTri> tris; std::vector<DMesh> meshes(s->n_meshes); uint32_t surf_mask = 0;
    V3 lo = mk3(INFINITY), hi = mk3(-INFINITY);
    for (uint32_t i = 0; i < s->n_meshes; ++i) {
        const lr_mesh &m = s->meshes[i]; DMesh &d = meshes[i]; memset(&d, 0, sizeof d);
        if (m.bsdf < 0 || (uint32_t) m.bsdf >= s->n_bsdfs) LR_FAIL(c, LR_ERR_INVALID_ARG, "mesh %u: bsdf index out of range", i);
        DevBuf *bp = c->newbuf(), *bi = c->newbuf();
        if (int r = upload(c, bp, m.positions, (size_t) m.n_vertices * 3)) return r;
        if (int r = upload(c, bi, m.indices, (size_t) m.n_faces * 3)) return r;
        d.pos = bp->as<float>(); d.idx = bi->as<uint32_t>();
        if (m.normals) { DevBuf *b = c->newbuf(); if (int r = upload(c, b, m.normals, (size_t) m.n_vertices * 3)) return r; d.nrm = b->as<float>(); }
        if (m.texcoords) { DevBuf *b = c->newbuf(); if (int r = upload(c, b, m.texcoords, (size_t) m.n_vertices * 2)) return r; d.uv = b->as<float>(); }
        d.bsdf = m.bsdf; d.interior = m.interior_medium; d.exterior = m.exterior_medium; d.emitter = m.emitter;
        d.tri_offset = (uint32_t) tris.size(); d.n_faces = m.n_faces;
        d.flags = ((m.interior_medium >= 0 || m.exterior_medium >= 0) ? 1u : 0u) | (m.flip_normals ? 2u : 0u);
        bool is_null = s->bsdfs[m.bsdf].kind == LR_BSDF_NULL; if (is_null) sc.any_null_bsdf = 1;
        uint32_t bucket; { uint32_t k = s->bsdfs[m.bsdf].kind; bucket = k == LR_BSDF_ROUGHDIELECTRIC ? 1u : (k == LR_BSDF_PRINCIPLED ? 2u : 0u); surf_mask |= 1u << bucket; }
        std::vector<float> cdf(m.n_faces); float area = 0.f;
        for (uint32_t f = 0; f < m.n_faces; ++f) {
            const uint32_t *ix = m.indices + 3 * (size_t) f;
            for (int k = 0; k < 3; ++k) if (ix[k] >= m.n_vertices) LR_FAIL(c, LR_ERR_INVALID_ARG, "mesh %u: vertex index out of range", i);
            const float *a = m.positions + 3 * (size_t) ix[0], *b = m.positions + 3 * (size_t) ix[1], *cc = m.positions + 3 * (size_t) ix[2];
            DTri t; uint32_t gid = (uint32_t) tris.size();
            t.v0 = make_float4(a[0], a[1], a[2], 0.f); t.v1 = make_float4(b[0], b[1], b[2], 0.f); t.v2 = make_float4(cc[0], cc[1], cc[2], 0.f);
            memcpy(&t.v0.w, &gid, 4); memcpy(&t.v1.w, &i, 4); uint32_t fl = (is_null ? 1u : 0u) | (bucket << 1); memcpy(&t.v2.w, &fl, 4);
            tris.push_back(t);
            for (const float *p : {a, b, cc}) { lo.x = std::min(lo.x, p[0]); lo.y = std::min(lo.y, p[1]); lo.z = std::min(lo.z, p[2]);
                                                hi.x = std::max(hi.x, p[0]); hi.y = std::max(hi.y, p[1]); hi.z = std::max(hi.z, p[2]); }
            // 0.5 * |e1 x e2| with the device's operation order is not needed bit-exactly: set-up only
            float e1[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, e2[3] = {cc[0] - a[0], cc[1] - a[1], cc[2] - a[2]};
            float cx = e1[1] * e2[2] - e1[2] * e2[1], cy = e1[2] * e2[0] - e1[0] * e2[2], cz = e1[0] * e2[1] - e1[1] * e2[0];
            area += 0.5f * sqrtf(cx * cx + cy * cy + cz * cz); cdf[f] = area;
        }
        d.area = area;
        DevBuf *bc = c->newbuf(); if (int r = upload(c, bc, cdf.data(), cdf.size())) return r; d.area_cdf = bc->as<float>();
    }
    if (!tris.empty()) {                                       // scene bounding sphere (bbox.h:343-346)
        sc.bs_center = mk3((lo.x + hi.x) * 0.5f, (lo.y + hi.y) * 0.5f, (lo.z + hi.z) * 0.5f);
        float dx = sc.bs_center.x - hi.x, dy = sc.bs_center.y - hi.y, dz = sc.bs_center.z - hi.z;
        sc.bs_radius = sqrtf(fmaf(dz, dz, fmaf(dy, dy, dx * dx)));
    }
    { DevBuf *b = c->newbuf(); if (int r = upload(c, b, meshes.data(), meshes.size())) return r; sc.meshes = b->as<DMesh>(); sc.n_meshes = s->n_meshes; }
    { DevBuf *b = c->newbuf(); if (int r = upload(c, b, s->bsdfs, s->n_bsdfs)) return r; sc.bsdfs = b->as<lr_bsdf>(); }
    if (int r = build_bvh(c, tris, lo, hi)) return r;
    sc.tris_gid = c->tris_gid ? c->tris_gid->as<DTri>() : nullptr;
    {   // shared-memory plan of the trace kernels (lr_trace.cuh): staged rays + the top of the tree + the first entries of every lane's stack
        c->bvh4_small = c->bvh4_nodes < 32000 && sc.n_tris <= 32000;
        const size_t esz = c->bvh4_small ? sizeof(short) : sizeof(int), budget = 200 * 1024;
        c->trace_sd = std::min<uint32_t>(std::max<uint32_t>(sc.bvh4_stack, 1u), LR_TRACE_SD);
        const size_t fixed = (size_t) 2 * LR_TRACE_THREADS * 16 + (size_t) c->trace_sd * LR_TRACE_THREADS * esz;
        c->trace_n_smem = (uint32_t) std::min<size_t>(c->bvh4_nodes, (budget - fixed) / sizeof(DNodeQ));
        c->trace_smem = fixed + (size_t) c->trace_n_smem * sizeof(DNodeQ);
        if (getenv("LR_VERBOSE")) fprintf(stderr, "[lr] bvh4: %u nodes (%u in shared memory), worst-case stack %u entries (%u in shared memory), %zu B of shared memory per block\n",
                                          c->bvh4_nodes, c->trace_n_smem, sc.bvh4_stack, c->trace_sd, c->trace_smem);
        // the attribute belongs to the function on this device, not to the context: it is only ever raised, so that an earlier context
        // with a larger tree can still launch after a later context has uploaded a smaller one
        static size_t s_trace_smem_set[64] = {0};
        size_t &cur = s_trace_smem_set[c->device & 63];
        if (c->trace_smem > cur) {
            const int v = (int) c->trace_smem;
            LR_CUDA(c, cudaFuncSetAttribute(k_wv_trace<0, short>, cudaFuncAttributeMaxDynamicSharedMemorySize, v));
            LR_CUDA(c, cudaFuncSetAttribute(k_wv_trace<1, short>, cudaFuncAttributeMaxDynamicSharedMemorySize, v));
            LR_CUDA(c, cudaFuncSetAttribute(k_wv_trace<2, short>, cudaFuncAttributeMaxDynamicSharedMemorySize, v));
            LR_CUDA(c, cudaFuncSetAttribute(k_wv_trace<0, int>, cudaFuncAttributeMaxDynamicSharedMemorySize, v));
            LR_CUDA(c, cudaFuncSetAttribute(k_wv_trace<1, int>, cudaFuncAttributeMaxDynamicSharedMemorySize, v));
            LR_CUDA(c, cudaFuncSetAttribute(k_wv_trace<2, int>, cudaFuncAttributeMaxDynamicSharedMemorySize, v));
            cur = c->trace_smem;
        }
    }

    // ---- bitmap textures (reflectance / bump maps) ----
    {
        if (s->n_textures && !s->textures) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_scene_upload: n_textures = %u but textures is NULL", s->n_textures);
        std::vector<DTex> tex(std::max<uint32_t>(s->n_textures, 1)); memset(tex.data(), 0, tex.size() * sizeof(DTex));
        for (uint32_t i = 0; i < s->n_textures; ++i) {
            const lr_texture &t = s->textures[i]; DTex &d = tex[i];
            if (!t.data || t.width < 2 || t.height < 2) LR_FAIL(c, LR_ERR_INVALID_ARG, "texture %u: needs data of at least 2x2 pixels", i);
            if (t.channels != 1 && t.channels != 3) LR_FAIL(c, LR_ERR_UNSUPPORTED, "texture %u: channels must be 1 or 3", i);
            if (t.filter > LR_TEX_NEAREST || t.wrap > LR_WRAP_CLAMP) LR_FAIL(c, LR_ERR_INVALID_ARG, "texture %u: unknown filter / wrap mode", i);
            DevBuf *b = c->newbuf(); if (int r = upload(c, b, t.data, (size_t) t.width * t.height * t.channels)) return r;
            d.data = b->as<float>(); d.w = (int) t.width; d.h = (int) t.height; d.ch = (int) t.channels; d.filter = t.filter; d.wrap = t.wrap;
            memcpy(d.to_uv, t.to_uv, sizeof d.to_uv);
        }
        DevBuf *b = c->newbuf(); if (int r = upload(c, b, tex.data(), tex.size())) return r; sc.textures = b->as<DTex>();
        for (uint32_t i = 0; i < s->n_bsdfs; ++i) {
            const lr_bsdf &bs = s->bsdfs[i];
            if (bs.bump_tex > s->n_textures || (bs.tex_kind == 2 && (bs.refl_tex == 0 || bs.refl_tex > s->n_textures)))
                LR_FAIL(c, LR_ERR_INVALID_ARG, "bsdf %u: texture slot out of range", i);
            if (bs.tex_kind > 2) LR_FAIL(c, LR_ERR_UNSUPPORTED, "bsdf %u: unknown reflectance texture kind %u", i, bs.tex_kind);
        }
    }

    // ---- grids ----
    c->h_grids.resize(s->n_grids); c->grid_floats.resize(s->n_grids); c->grid_fixed_max.assign(s->n_grids, -1.f);
    for (uint32_t i = 0; i < s->n_grids; ++i) c->grid_fixed_max[i] = s->grids[i].max_value;
    for (uint32_t i = 0; i < s->n_grids; ++i) {
        const lr_grid &g = s->grids[i]; DGrid &d = c->h_grids[i]; memset(&d, 0, sizeof d);
        if (g.channels != 1 && g.channels != 3) LR_FAIL(c, LR_ERR_UNSUPPORTED, "grid %u: channels must be 1 or 3", i);
        size_t n = (size_t) g.res[0] * g.res[1] * g.res[2] * g.channels; c->grid_floats[i] = n;
        DevBuf *b = c->newbuf(); if (int r = upload(c, b, g.data, n)) return r;
        d.data = b->as<float>(); d.rx = g.res[0]; d.ry = g.res[1]; d.rz = g.res[2]; d.ch = g.channels;
        mat_copy(d.to_local, g.to_local); d.bmin = mk3(g.bbox_min[0], g.bbox_min[1], g.bbox_min[2]); d.bmax = mk3(g.bbox_max[0], g.bbox_max[1], g.bbox_max[2]);
        d.filter = g.filter; d.wrap = g.wrap; d.differentiable = g.differentiable;
        d.cell = s->integrator.majorant_cell ? (int) s->integrator.majorant_cell : 8;
        d.mx = (d.rx + d.cell - 1) / d.cell; d.my = (d.ry + d.cell - 1) / d.cell; d.mz = (d.rz + d.cell - 1) / d.cell;
        DevBuf *bm = c->newbuf(); LR_CUDA(c, bm->alloc((((size_t) d.mx * d.my * d.mz + 3) & ~(size_t) 3) * 4)); d.maj = bm->as<float>();   // rounded to 16 B: bulk copies move whole 16-byte units
        if (g.channels == 1 && g.filter == LR_FILTER_TRILINEAR && g.wrap == LR_WRAP_CLAMP && getenv("LR_NO_FAT") == nullptr) {      // fat-voxel copy: 32 B per record
            DevBuf *bf = c->newbuf(); LR_CUDA(c, bf->alloc((size_t) (d.rx + 1) * (d.ry + 1) * (d.rz + 1) * 32)); d.fat = bf->as<float4>();
        }
        if (g.differentiable) { DevBuf *bg = c->newbuf(); LR_CUDA(c, bg->alloc(n * 4)); LR_CUDA(c, cudaMemsetAsync(bg->p, 0, n * 4, c->stream)); d.grad = bg->as<float>(); }
    }
    c->h_media.resize(s->n_media);
    for (uint32_t i = 0; i < s->n_media; ++i) {
        if (s->integrator.kind == LR_INTEGRATOR_PRBVOLPATH && s->media[i].kind == LR_MEDIUM_HETEROGENEOUS && s->media[i].sigma_t_grid >= 0) {
            const lr_grid &sg = s->grids[s->media[i].sigma_t_grid];
            if (sg.differentiable && sg.channels != 1)
                LR_FAIL(c, LR_ERR_UNSUPPORTED, "medium %u: a differentiable sigma_t gridvolume must have 1 channel (the per-channel gradient of an rgb sigma_t grid is not "
                                               "implemented; the adjoint would silently return zeros).  Albedo grids may have 1 or 3 channels.", i);
        }
        c->h_media[i].m = s->media[i]; c->h_media[i].max_density = 0.f;
        const lr_medium &m = s->media[i];
        if (m.kind > LR_MEDIUM_PARENCHYMA) LR_FAIL(c, LR_ERR_UNSUPPORTED, "medium %u: unknown medium kind %u", i, m.kind);
        // the tissue media only implement the five-argument sample_interaction biovolpath calls, and the base class's version of
        // that overload throws (src/render/medium.cpp:84-90): the two always come together
        const bool tissue = m.kind >= LR_MEDIUM_LIVER, bio = s->integrator.kind == LR_INTEGRATOR_BIOVOLPATH;
        if (tissue != bio)
            LR_FAIL(c, LR_ERR_UNSUPPORTED, "medium %u: %s", i, bio ? "biovolpath needs liver / glissonCapsule / parenchyma media (Medium::sample_interaction with a tissue depth is "
                                                                     "not implemented by the other media)"
                                                                   : "liver / glissonCapsule / parenchyma media are only sampled by biovolpath");
        if (tissue && (m.sigma_t_grid >= 0 || m.albedo_grid >= 0)) LR_FAIL(c, LR_ERR_UNSUPPORTED, "medium %u: tissue media take constant sigma_t / albedo only", i);
        if (m.kind == LR_MEDIUM_HETEROGENEOUS && (m.sigma_t_grid < 0 || (uint32_t) m.sigma_t_grid >= s->n_grids))
            LR_FAIL(c, LR_ERR_INVALID_ARG, "medium %u: heterogeneous medium needs a sigma_t grid", i);
        if (m.kind == LR_MEDIUM_HETEROGENEOUS) sc.prb_handle_null = 1; else sc.prb_nee_homogeneous = 1;
    }
    seal_media(s, c->h_media);
    if (getenv("LR_VERBOSE")) for (uint32_t i = 0; i < s->n_media; ++i) fprintf(stderr, "[lr] medium %u: %s\n", i, c->h_media[i].sealed ? "sealed (NEE from inside skipped in DDA mode)" : "open");
    c->d_grids = c->newbuf(); LR_CUDA(c, c->d_grids->alloc(std::max<size_t>(s->n_grids, 1) * sizeof(DGrid)));
    c->d_media = c->newbuf(); LR_CUDA(c, c->d_media->alloc(std::max<size_t>(s->n_media, 1) * sizeof(DMedium)));
    sc.grids = c->d_grids->as<DGrid>(); sc.media = c->d_media->as<DMedium>();
    if (s->n_media) LR_CUDA(c, cudaMemcpyAsync(c->d_media->p, c->h_media.data(), s->n_media * sizeof(DMedium), cudaMemcpyHostToDevice, c->stream));
    for (uint32_t i = 0; i < s->n_grids; ++i) {
        if (int r = refresh_grid_derived(c, i)) return r;
    }

    // ---- emitters ----
    std::vector<DEmitter> ems(s->n_emitters); sc.env = -1; sc.n_emitters = s->n_emitters;
    for (uint32_t i = 0; i < s->n_emitters; ++i) {
        DEmitter &d = ems[i]; memset(&d, 0, sizeof d); d.e = s->emitters[i]; d.e.pixels = nullptr;
        if (d.e.kind != LR_EMITTER_AREA) sc.env = (int) i;
        if (d.e.kind == LR_EMITTER_AREA && !d.e.is_rectangle && (d.e.mesh < 0 || (uint32_t) d.e.mesh >= s->n_meshes))
            LR_FAIL(c, LR_ERR_INVALID_ARG, "emitter %u: area light without a mesh", i);
        if (d.e.kind == LR_EMITTER_ENVMAP) {
            if (!s->emitters[i].pixels || d.e.width < 2 || d.e.height < 2) LR_FAIL(c, LR_ERR_INVALID_ARG, "emitter %u: envmap needs pixels", i);
            std::vector<float> px, warp; build_envmap(s->emitters[i], px, warp, d);
            DevBuf *bp = c->newbuf(), *bw = c->newbuf();
            if (int r = upload(c, bp, px.data(), px.size())) return r; if (int r = upload(c, bw, warp.data(), warp.size())) return r;
            d.pixels = bp->as<float>(); d.warp_data = bw->as<float>();
            mat_copy(d.to_world, d.e.to_world); invert4(d.e.to_world, d.to_local.m);
        }
    }
    { DevBuf *b = c->newbuf(); if (int r = upload(c, b, ems.data(), ems.size())) return r; sc.emitters = b->as<DEmitter>(); }

    // ---- film ----
    c->W = s->sensor.film_width; c->H = s->sensor.film_height; c->film_channels = s->sensor.has_alpha ? 5 : 4;
    LR_CUDA(c, c->film.alloc((size_t) c->W * c->H * c->film_channels * 4));
    LR_CUDA(c, cudaMemsetAsync(c->film.p, 0, c->film.bytes, c->stream));
    LR_CUDA(c, c->weight_img.alloc((size_t) c->W * c->H * 4));
    LR_CUDA(c, cudaStreamSynchronize(c->stream));
    c->surf_bucket_mask = surf_mask;
    c->all_media_dda = s->integrator.tracking == LR_TRACKING_DDA && getenv("LR_TRACK_REFILL_OFF") == nullptr;   // persistent TRACK (k_wf_track_dda)
    for (uint32_t i = 0; i < s->n_media; ++i) if (s->media[i].kind != LR_MEDIUM_HETEROGENEOUS) c->all_media_dda = false;
    // majorant super-grid staged in shared memory by the persistent TRACK (TMA bulk copy at kernel start): the grid of the first medium, if it fits
    c->smaj_grid = -1; c->smaj_floats = 0;
    if (c->all_media_dda && s->n_media && getenv("LR_NO_SMAJ") == nullptr) {
        const int gi = s->media[0].sigma_t_grid; const DGrid &g = c->h_grids[gi];
        const size_t n = ((size_t) g.mx * g.my * g.mz + 3) & ~(size_t) 3;
        if (n * 4 <= 200 * 1024) {
            c->smaj_grid = gi; c->smaj_floats = (uint32_t) n;
            const int lim = 200 * 1024 + 64;
            LR_CUDA(c, cudaFuncSetAttribute(k_wv_track_dda<4, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim));
            LR_CUDA(c, cudaFuncSetAttribute(k_wv_track_dda<4, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim));
            LR_CUDA(c, cudaFuncSetAttribute(k_wv_track_dda<5, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim));
            LR_CUDA(c, cudaFuncSetAttribute(k_wv_track_dda<5, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim));
        }
    }
    c->has_scene = true; memset(&c->st, 0, sizeof c->st);
    return LR_OK;
}

} // extern "C"

// ---------------------------------------------------------------------------------------------
// launches
// ---------------------------------------------------------------------------------------------
struct PassPlan { uint32_t spp_per_pass, n_passes; uint64_t n_lanes; };
// pass splitting when W*H*spp exceeds 2^32-1 lanes (src/render/integrator.cpp:275-293)
// ldsampler: the sample count must be a square power of two (LowDiscrepancySampler::set_sample_count, ldsampler.cpp:80-89;
// the host mirror rounds `sample_count` up the same way before it calls in)
static bool ld_spp_ok(uint32_t spp) { return spp >= 4 && (spp & (spp - 1)) == 0 && (__builtin_ctz(spp) % 2) == 0; }
static PassPlan plan_passes(const lr_ctx *c, uint32_t spp) {
    PassPlan p; p.spp_per_pass = spp; p.n_passes = 1;
    uint64_t wf = (uint64_t) c->W * c->H * spp, limit = 0xffffffffull;
    if (wf > limit) { p.spp_per_pass = spp / (uint32_t) ((wf + limit - 1) / limit); p.n_passes = spp / p.spp_per_pass; }
    p.n_lanes = (uint64_t) c->W * c->H * p.spp_per_pass;
    return p;
}

// per-lane kernels (k_render / k_backward): compiled in lr_lanes.cu, launched through lr_lanes.h
static int launch_render_lanes(lr_ctx *c, const RenderArgs &a, bool to_film) {
    LR_CUDA(c, lanes_launch_render(c->film_channels, to_film, c->n_sm, c->stream, c->sc, a)); return LR_OK;
}

#include "lr_wavefront_host.inl"

static int begin_call(lr_ctx *c, const char *what) {
    if (!c) { g_last_error = std::string(what) + ": NULL context"; return LR_ERR_INVALID_ARG; }
    if (!c->has_scene) LR_FAIL(c, LR_ERR_NO_SCENE, "%s: no scene uploaded", what);
    if (c->nvtx) { nvtxRangePushA(what); c->nvtx_open = true; }
    LR_CUDA(c, cudaSetDevice(c->device));
    *c->cancel_host = 0;
    LR_CUDA(c, cudaMemsetAsync(c->work_counter.p, 0, 8, c->stream));
    LR_CUDA(c, cudaMemsetAsync(c->stats.p, 0, 32, c->stream));
    LR_CUDA(c, cudaEventRecord(c->ev0, c->stream));
    return LR_OK;
}
static int end_call(lr_ctx *c, uint64_t launches, bool tracking) {
    if (c->nvtx_open) { nvtxRangePop(); c->nvtx_open = false; }
    LR_CUDA(c, cudaEventRecord(c->ev1, c->stream));
    unsigned long long h[4]; LR_CUDA(c, cudaMemcpyAsync(h, c->stats.p, 32, cudaMemcpyDeviceToHost, c->stream));
    LR_CUDA(c, cudaStreamSynchronize(c->stream));
    float ms = 0.f; LR_CUDA(c, cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    c->st.samples = h[0]; c->st.rays = h[1]; c->st.collisions = h[2]; c->st.real_scatters = h[3];
    c->st.launches = launches; c->st.device_ms = ms;
    if (tracking) { c->st.tracking_ms = ms; c->st.tracking_launches = 1; }
    if (tracking && c->wf_launches) { c->st.launches = launches - 1 + c->wf_launches; c->st.tracking_ms = c->wf_tracking_ms; c->st.tracking_launches = c->wf_tracking_launches; }
    if (*c->cancel_host) LR_FAIL(c, LR_ERR_CANCELLED, "render cancelled");
    return LR_OK;
}

// lanes of this context's share under lr_set_lane_map (n_lanes when no map is set)
static int lane_map_count(lr_ctx *c, uint64_t n_lanes, uint64_t &n_local) {
    const LaneMap &m = c->lane_map; n_local = n_lanes;
    if (!m.len) return LR_OK;
    if (n_lanes % m.period) LR_FAIL(c, LR_ERR_INVALID_ARG, "lane map: period %llu does not divide the %llu lanes of a pass", (unsigned long long) m.period, (unsigned long long) n_lanes);
    n_local = n_lanes / m.period * m.len; return LR_OK;
}

static int render_lanes_impl(lr_ctx *c, uint32_t seed, uint32_t spp, uint64_t begin, uint64_t end, float *lanes_out) {
    PassPlan p = plan_passes(c, spp);
    if (c->sc.integ.sampler == LR_SAMPLER_LD && !ld_spp_ok(spp)) LR_FAIL(c, LR_ERR_INVALID_ARG, "ldsampler: spp %u is not a square power of two", spp);
    uint64_t n_local = p.n_lanes;
    if (int r = lane_map_count(c, p.n_lanes, n_local)) return r;
    if (end == 0 && begin == 0) end = n_local;
    if (begin > end || end > n_local) LR_FAIL(c, LR_ERR_INVALID_ARG, "lane range [%llu,%llu) outside [0,%llu)", (unsigned long long) begin, (unsigned long long) end, (unsigned long long) n_local);
    RenderArgs a{}; a.map = c->lane_map; a.tris_gid = c->tris_gid->as<DTri>(); a.film = c->film.as<float>(); a.lanes_out = lanes_out;
    a.lane_begin = begin; a.lane_end = end; a.seed = seed; a.spp_per_pass = p.spp_per_pass; a.n_passes = p.n_passes; a.total_spp = spp;
    a.work_counter = c->work_counter.as<unsigned long long>(); a.stats = c->stats.as<unsigned long long>(); a.cancel = c->cancel_dev;
    if (begin == end) return LR_OK;
    c->wf_launches = 0; c->wf_tracking_ms = 0; c->wf_tracking_launches = 0;
    if (c->use_wavefront && c->sc.integ.kind == LR_INTEGRATOR_PRBVOLPATH) {        // primal prbvolpath (INTEGRATOR_MAPPING renders)
        if (int r = wv_prepare(c, (end - begin))) return r;
        if (p.n_passes != 1) LR_FAIL(c, LR_ERR_UNSUPPORTED, "prbvolpath: W*H*spp exceeds 2^32 lanes");
        if (lanes_out) return c->film_channels == 5 ? wavefront_render_t<5, false, 1>(c, a) : wavefront_render_t<4, false, 1>(c, a);
        return c->film_channels == 5 ? wavefront_render_t<5, true, 1>(c, a) : wavefront_render_t<4, true, 1>(c, a);
    }
    if (c->use_wavefront && c->sc.integ.kind == LR_INTEGRATOR_BIOVOLPATH && !getenv("LR_BIO_PER_LANE")) {     // LR_BIO_PER_LANE=1: k_render<biovolpath>, the parity probe
        if (int r = wv_prepare(c, (end - begin))) return r;
        if (lanes_out) return c->film_channels == 5 ? wavefront_render_t<5, false, 3>(c, a) : wavefront_render_t<4, false, 3>(c, a);
        return c->film_channels == 5 ? wavefront_render_t<5, true, 3>(c, a) : wavefront_render_t<4, true, 3>(c, a);
    }
    if (c->use_wavefront && c->sc.integ.kind == LR_INTEGRATOR_VOLPATHMIS && !getenv("LR_MIS_PER_LANE")) {    // LR_MIS_PER_LANE=1: k_render<volpathmis>, the parity probe
        if (int r = wv_prepare(c, (end - begin))) return r;
        const bool s5 = c->film_channels == 5;
        if (c->sc.integ.spectral_mis) {
            if (lanes_out) return s5 ? wavefront_render_t<5, false, 4>(c, a) : wavefront_render_t<4, false, 4>(c, a);
            return s5 ? wavefront_render_t<5, true, 4>(c, a) : wavefront_render_t<4, true, 4>(c, a);
        }
        if (lanes_out) return s5 ? wavefront_render_t<5, false, 5>(c, a) : wavefront_render_t<4, false, 5>(c, a);
        return s5 ? wavefront_render_t<5, true, 5>(c, a) : wavefront_render_t<4, true, 5>(c, a);
    }
    if (c->use_wavefront && c->sc.integ.kind == LR_INTEGRATOR_VOLPATH) {
        if (int r = wv_prepare(c, (end - begin))) return r;
        if (lanes_out) return c->film_channels == 5 ? wavefront_render_t<5, false>(c, a) : wavefront_render_t<4, false>(c, a);
        return c->film_channels == 5 ? wavefront_render_t<5, true>(c, a) : wavefront_render_t<4, true>(c, a);
    }
    return launch_render_lanes(c, a, lanes_out == nullptr);
}

static int copy_out(lr_ctx *c, void *dst, const void *src_dev, size_t bytes, int dst_is_device) {
    LR_CUDA(c, cudaMemcpyAsync(dst, src_dev, bytes, dst_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
    LR_CUDA(c, cudaStreamSynchronize(c->stream)); return LR_OK;
}

extern "C" {

// Scene::sensors()[i]: SamplingIntegrator::render(scene, sensor_index, ...) (src/render/integrator.cpp:33-44) picks one of the scene's
// sensors per call.  Geometry, media, emitters and the acceleration structure stay as uploaded; only the camera, film window, reconstruction
// filter and the sensor's medium change.  The film accumulator is re-created (zeroed) for the new sensor's size.
int lr_sensor_set(lr_ctx *c, const lr_sensor *sn) {
    if (!c || !sn) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_sensor_set: NULL argument");
    if (!c->has_scene) LR_FAIL(c, LR_ERR_NO_SCENE, "lr_sensor_set: no scene uploaded");
    if (sn->film_width == 0 || sn->film_height == 0) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_sensor_set: empty film");
    if (sn->rfilter_radius > 7.5f) LR_FAIL(c, LR_ERR_UNSUPPORTED, "lr_sensor_set: reconstruction filter radius > 7.5 px is outside the hot path");
    if (sn->rfilter_radius > 2.5f && sn->rfilter != LR_RFILTER_BOX && c->sc.integ.kind == LR_INTEGRATOR_PRBVOLPATH)
        LR_FAIL(c, LR_ERR_UNSUPPORTED, "lr_sensor_set: prbvolpath gathers dL/dI through the 6-tap footprint: reconstruction filter radius > 2.5 px is not supported by the adjoint");
    if (sn->medium >= (int) c->h_media.size()) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_sensor_set: sensor medium %d out of range", sn->medium);
    if (sn->medium >= 0 && (c->h_media[sn->medium].m.kind >= LR_MEDIUM_LIVER) != (c->sc.integ.kind == LR_INTEGRATOR_BIOVOLPATH))
        LR_FAIL(c, LR_ERR_UNSUPPORTED, "lr_sensor_set: the sensor's medium does not go with the scene's integrator");
    LR_CUDA(c, cudaSetDevice(c->device));
    LR_CUDA(c, cudaStreamSynchronize(c->stream));
    if (sn->medium >= 0 && c->h_media[sn->medium].sealed) {          // a medium that contains the sensor is not sealed (seal_media)
        c->h_media[sn->medium].sealed = 0;
        LR_CUDA(c, cudaMemcpyAsync(c->d_media->p, c->h_media.data(), c->h_media.size() * sizeof(DMedium), cudaMemcpyHostToDevice, c->stream));
    }
    DScene &sc = c->sc; sc.sensor = *sn;
    mat_copy(sc.cam_to_world, sn->to_world); mat_copy(sc.sample_to_camera, sn->sample_to_camera);
    if (sn->rfilter == LR_RFILTER_GAUSSIAN) gaussian_coeffs(sn->rfilter_stddev, sn->rfilter_radius, sc.gauss_coeff);
    const int ch = sn->has_alpha ? 5 : 4;
    if (c->W != sn->film_width || c->H != sn->film_height || c->film_channels != ch) {
        c->W = sn->film_width; c->H = sn->film_height; c->film_channels = ch;
        LR_CUDA(c, c->film.alloc((size_t) c->W * c->H * c->film_channels * 4));
        LR_CUDA(c, c->weight_img.alloc((size_t) c->W * c->H * 4));
    }
    LR_CUDA(c, cudaMemsetAsync(c->film.p, 0, c->film.bytes, c->stream));
    LR_CUDA(c, cudaStreamSynchronize(c->stream));
    return LR_OK;
}

int lr_film_clear(lr_ctx *c) {
    if (!c || !c->has_scene) return LR_ERR_NO_SCENE;
    LR_CUDA(c, cudaSetDevice(c->device)); LR_CUDA(c, cudaMemsetAsync(c->film.p, 0, c->film.bytes, c->stream)); return LR_OK;
}
int lr_film_channels(const lr_ctx *c) { return c ? c->film_channels : 0; }
void *lr_film_device_ptr(lr_ctx *c) { return c ? c->film.p : nullptr; }

int lr_render_lanes(lr_ctx *c, uint32_t seed, uint32_t spp, uint64_t lane_begin, uint64_t lane_end) {
    if (int r = begin_call(c, "lr_render_lanes")) return r;
    if (spp == 0) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_render_lanes: spp must be > 0");
    if (int r = render_lanes_impl(c, seed, spp, lane_begin, lane_end, nullptr)) return r;
    return end_call(c, 1, true);
}

int lr_film_develop(lr_ctx *c, float *out, int out_is_device) {
    if (!c || !c->has_scene) return LR_ERR_NO_SCENE;
    LR_CUDA(c, cudaSetDevice(c->device));
    uint32_t npix = c->W * c->H; int C = c->film_channels; size_t bytes = (size_t) npix * (C - 1) * 4;
    float *dst = out;
    if (!out_is_device) { if (c->scratch.bytes < bytes) LR_CUDA(c, c->scratch.alloc(bytes)); dst = c->scratch.as<float>(); }
    k_develop<<<(npix + 255) / 256, 256, 0, c->stream>>>(c->film.as<float>(), dst, npix, C);
    LR_CUDA(c, cudaGetLastError());
    if (!out_is_device) return copy_out(c, out, dst, bytes, 0);
    LR_CUDA(c, cudaStreamSynchronize(c->stream)); return LR_OK;
}

int lr_render(lr_ctx *c, uint32_t seed, uint32_t spp, float *out, int out_is_device) {
    if (int r = begin_call(c, "lr_render")) return r;
    if (spp == 0 || !out) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_render: spp must be > 0 and out non-NULL");
    LR_CUDA(c, cudaMemsetAsync(c->film.p, 0, c->film.bytes, c->stream));
    if (int r = render_lanes_impl(c, seed, spp, 0, 0, nullptr)) return r;
    if (int r = end_call(c, 2, true)) return r;
    return lr_film_develop(c, out, out_is_device);
}

int lr_film_read(lr_ctx *c, float *block, int is_device) {
    if (!c || !c->has_scene) return LR_ERR_NO_SCENE;
    LR_CUDA(c, cudaSetDevice(c->device)); return copy_out(c, block, c->film.p, c->film.bytes, is_device);
}
int lr_film_write(lr_ctx *c, const float *block, int is_device) {
    if (!c || !c->has_scene) return LR_ERR_NO_SCENE;
    LR_CUDA(c, cudaSetDevice(c->device));
    LR_CUDA(c, cudaMemcpyAsync(c->film.p, block, c->film.bytes, is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
    LR_CUDA(c, cudaStreamSynchronize(c->stream)); return LR_OK;
}

int lr_sample_lanes(lr_ctx *c, uint32_t seed, uint32_t spp, uint64_t lane_begin, uint64_t lane_end, float *out, int out_is_device) {
    if (int r = begin_call(c, "lr_sample_lanes")) return r;
    if (spp == 0 || !out || lane_end <= lane_begin) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_sample_lanes: bad arguments");
    PassPlan p = plan_passes(c, spp); size_t bytes = (size_t) (lane_end - lane_begin) * p.n_passes * 16;
    float *dst = out; DevBuf tmp;
    if (!out_is_device) { LR_CUDA(c, tmp.alloc(bytes)); dst = tmp.as<float>(); }
    if (int r = render_lanes_impl(c, seed, spp, lane_begin, lane_end, dst)) return r;
    if (int r = end_call(c, 1, true)) return r;
    if (!out_is_device) return copy_out(c, out, dst, bytes, 0);
    return LR_OK;
}

int lr_render_backward(lr_ctx *c, uint32_t seed, uint32_t spp, const float *grad_in, int grad_is_device, uint64_t lane_begin, uint64_t lane_end) {
    if (int r = begin_call(c, "lr_render_backward")) return r;
    if (c->sc.integ.kind != LR_INTEGRATOR_PRBVOLPATH) LR_FAIL(c, LR_ERR_UNSUPPORTED, "lr_render_backward: integrator is not prbvolpath");
    if (spp == 0 || !grad_in) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_render_backward: bad arguments");
    PassPlan p = plan_passes(c, spp);
    if (c->sc.integ.sampler == LR_SAMPLER_LD && !ld_spp_ok(spp)) LR_FAIL(c, LR_ERR_INVALID_ARG, "ldsampler: spp %u is not a square power of two", spp);
    if (p.n_passes != 1) LR_FAIL(c, LR_ERR_UNSUPPORTED, "lr_render_backward: W*H*spp exceeds 2^32 (the reference raises as well, common.py:359-364)");
    uint64_t n_local = p.n_lanes;
    if (int r = lane_map_count(c, p.n_lanes, n_local)) return r;
    if (lane_begin == 0 && lane_end == 0) lane_end = n_local;
    if (lane_begin > lane_end || lane_end > n_local) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_render_backward: lane range out of bounds");
    uint32_t gc = c->film_channels - 1; size_t gbytes = (size_t) c->W * c->H * gc * 4;
    const float *g = grad_in; DevBuf tmp;
    if (!grad_is_device) { LR_CUDA(c, tmp.alloc(gbytes)); LR_CUDA(c, cudaMemcpyAsync(tmp.p, grad_in, gbytes, cudaMemcpyHostToDevice, c->stream)); g = tmp.as<float>(); }
    LR_CUDA(c, cudaMemsetAsync(c->weight_img.p, 0, c->weight_img.bytes, c->stream));
    k_weight_image<<<c->n_sm * 8, 256, 0, c->stream>>>(c->sc, c->weight_img.as<float>(), p.n_lanes, seed, spp);
    LR_CUDA(c, cudaGetLastError());
    RenderArgs a{}; a.tris_gid = c->tris_gid->as<DTri>(); a.lane_begin = lane_begin; a.lane_end = lane_end; a.seed = seed; a.map = c->lane_map;
    a.spp_per_pass = spp; a.n_passes = 1; a.total_spp = spp; a.work_counter = c->work_counter.as<unsigned long long>();
    a.stats = c->stats.as<unsigned long long>(); a.cancel = c->cancel_dev; a.grad_in = g; a.weight_img = c->weight_img.as<float>(); a.grad_channels = gc;
    c->wf_launches = 0;
    if (lane_end > lane_begin) {
        if (c->use_wavefront) { if (int r = wavefront_prb_backward(c, a, g, gc)) return r; }
        else {
            LR_CUDA(c, lanes_launch_backward(c->n_sm, c->stream, c->sc, a));
        }
    }
    return end_call(c, 2, true);
}

int lr_grid_update(lr_ctx *c, uint32_t grid, const float *data, int is_device) {
    if (!c || !c->has_scene) return LR_ERR_NO_SCENE;
    if (grid >= c->h_grids.size() || !data) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_grid_update: bad grid index");
    LR_CUDA(c, cudaSetDevice(c->device));
    LR_CUDA(c, cudaMemcpyAsync((void *) c->h_grids[grid].data, data, c->grid_floats[grid] * 4, is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
    if (int r = refresh_grid_derived(c, grid)) return r;
    LR_CUDA(c, cudaStreamSynchronize(c->stream)); return LR_OK;
}
int lr_grid_grad_zero(lr_ctx *c, uint32_t grid) {
    if (!c || !c->has_scene) return LR_ERR_NO_SCENE;
    if (grid >= c->h_grids.size() || !c->h_grids[grid].grad) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_grid_grad_zero: grid %u is not differentiable", grid);
    LR_CUDA(c, cudaSetDevice(c->device)); LR_CUDA(c, cudaMemsetAsync(c->h_grids[grid].grad, 0, c->grid_floats[grid] * 4, c->stream)); return LR_OK;
}
int lr_grid_grad_read(lr_ctx *c, uint32_t grid, float *out, int is_device) {
    if (!c || !c->has_scene) return LR_ERR_NO_SCENE;
    if (grid >= c->h_grids.size() || !c->h_grids[grid].grad || !out) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_grid_grad_read: grid %u is not differentiable", grid);
    LR_CUDA(c, cudaSetDevice(c->device)); return copy_out(c, out, c->h_grids[grid].grad, c->grid_floats[grid] * 4, is_device);
}
void *lr_grid_grad_device_ptr(lr_ctx *c, uint32_t grid) { return (c && c->has_scene && grid < c->h_grids.size()) ? c->h_grids[grid].grad : nullptr; }
size_t lr_grid_size(const lr_ctx *c, uint32_t grid) { return (c && grid < c->grid_floats.size()) ? c->grid_floats[grid] : 0; }

int lr_ray_intersect(lr_ctx *c, const float *rays, uint64_t n, float *hits) {
    if (!c || !c->has_scene) return LR_ERR_NO_SCENE;
    if (!rays || !hits) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_ray_intersect: NULL buffers");
    if (n == 0) return LR_OK;
    LR_CUDA(c, cudaSetDevice(c->device));
    DevBuf dr, dh; LR_CUDA(c, dr.alloc(n * 32)); LR_CUDA(c, dh.alloc(n * 16));
    LR_CUDA(c, cudaMemcpyAsync(dr.p, rays, n * 32, cudaMemcpyHostToDevice, c->stream));
    k_ray_probe<<<(unsigned) ((n + 127) / 128), 128, 0, c->stream>>>(c->sc, dr.as<float>(), n, dh.as<float>());
    LR_CUDA(c, cudaGetLastError());
    return copy_out(c, hits, dh.p, n * 16, 0);
}

int lr_set_lane_map(lr_ctx *c, uint64_t period, uint64_t offset, uint64_t len) {
    if (!c) return LR_ERR_INVALID_ARG;
    if (len == 0) { c->lane_map = LaneMap{0, 0, 0}; return LR_OK; }
    if (period == 0 || offset + len > period) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_set_lane_map: need 0 < offset + len <= period (got period %llu offset %llu len %llu)",
                                                     (unsigned long long) period, (unsigned long long) offset, (unsigned long long) len);
    c->lane_map = LaneMap{period, offset, len}; return LR_OK;
}

int lr_stats_get(lr_ctx *c, lr_stats *out) { if (!c || !out) return LR_ERR_INVALID_ARG; *out = c->st; return LR_OK; }

int lr_profile_enable(lr_ctx *c, int on) {
    if (!c) return LR_ERR_INVALID_ARG;
    c->profile_on = on != 0; memset(c->prof_ms, 0, sizeof c->prof_ms); memset(c->prof_launches, 0, sizeof c->prof_launches); memset(c->prof_items, 0, sizeof c->prof_items);
    return LR_OK;
}
int lr_profile_get(lr_ctx *c, lr_stage_profile *out) {
    if (!c || !out) return LR_ERR_INVALID_ARG;
    memset(out, 0, sizeof *out); out->n_stages = ST_COUNT;
    for (int k = 0; k < ST_COUNT; ++k) { out->ms[k] = c->prof_ms[k]; out->launches[k] = c->prof_launches[k]; out->items[k] = c->prof_items[k]; }
    return LR_OK;
}
const char *lr_stage_name(uint32_t stage) { return stage < ST_COUNT ? kStageNames[stage] : ""; }

} // extern "C"
