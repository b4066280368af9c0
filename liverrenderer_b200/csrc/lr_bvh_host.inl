// lr_bvh_host.inl -- host-side SAH builder for the two-wide BVH the traversal kernels walk (lr_scene.cuh: bvh_trace).
//
// The reference builds a SAH kd-tree on the host at scene load (include/mitsuba/render/kdtree.h, ShapeKDTree); the scenes of
// the hot path are a few thousand to a few hundred thousand triangles, for which a top-down SAH build takes milliseconds and
// gives markedly cheaper incoherent traversal than the Morton-order LBVH (build_lbvh, kept for very large inputs).  Hits do
// not depend on the tree: every candidate goes through the same Moeller-Trumbore test and ties are broken by the global
// primitive id, so the two builders render bit-identical images (tests/test_gpu_parity.py::test_bvh_builders_agree).
//
// Layout produced: node 0 is the root; nodes in depth-first order; one triangle per leaf, child code ~k = k-th triangle of the
// reordered array.  Leaf boxes are padded by `pad` (absolute) exactly like k_refit does.

namespace sah {

struct Box {
    float lo[3], hi[3];
    void reset() { for (int k = 0; k < 3; ++k) { lo[k] = INFINITY; hi[k] = -INFINITY; } }
    void grow(const Box &o) { for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], o.lo[k]); hi[k] = std::max(hi[k], o.hi[k]); } }
    float half_area() const {
        float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        return (dx < 0.f) ? 0.f : dx * dy + dy * dz + dz * dx;
    }
};

struct Builder {
    static constexpr int kBins = 32, kSweepBelow = 48, kMaxDepth = 40;   // traversal stack holds 48 entries
    const std::vector<DTri> &tris;
    std::vector<Box> box; std::vector<float> cen;                        // per triangle: padded box, centroid (3 floats)
    std::vector<uint32_t> order; std::vector<DNode> nodes;
    std::vector<uint32_t> scratch; std::vector<float> right_area;

    Builder(const std::vector<DTri> &t, float pad) : tris(t) {
        size_t n = t.size(); box.resize(n); cen.resize(3 * n); order.resize(n); nodes.reserve(n);
        for (size_t i = 0; i < n; ++i) {
            const float4 *v = &t[i].v0; Box b; b.reset();
            const float p[3][3] = { { v[0].x, v[0].y, v[0].z }, { v[1].x, v[1].y, v[1].z }, { v[2].x, v[2].y, v[2].z } };
            for (int k = 0; k < 3; ++k) {
                b.lo[k] = std::min(p[0][k], std::min(p[1][k], p[2][k])) - pad; b.hi[k] = std::max(p[0][k], std::max(p[1][k], p[2][k])) + pad;
                cen[3 * i + k] = 0.5f * (b.lo[k] + b.hi[k]);
            }
            box[i] = b; order[i] = (uint32_t) i;
        }
        right_area.resize(n + 1);
    }

    // exact sweep over one axis for small ranges; returns cost, writes the split position (count on the left)
    float sweep(uint32_t b, uint32_t e, int axis, uint32_t &split) {
        std::sort(order.begin() + b, order.begin() + e, [&](uint32_t x, uint32_t y) {
            float cx = cen[3 * x + axis], cy = cen[3 * y + axis]; return cx < cy || (cx == cy && x < y); });
        uint32_t n = e - b; Box acc; acc.reset();
        for (uint32_t i = n; i-- > 1;) { acc.grow(box[order[b + i]]); right_area[i] = acc.half_area(); }
        acc.reset(); float best = INFINITY; split = n / 2;
        for (uint32_t i = 1; i < n; ++i) {
            acc.grow(box[order[b + i - 1]]);
            float cost = acc.half_area() * (float) i + right_area[i] * (float) (n - i);
            if (cost < best) { best = cost; split = i; }
        }
        return best;
    }

    // returns the child code of the subtree over order[b, e) and its bounds
    int build(uint32_t b, uint32_t e, int depth, Box &bounds) {
        if (e - b == 1) { bounds = box[order[b]]; return ~(int) b; }
        int self = (int) nodes.size(); nodes.push_back(DNode{});
        uint32_t n = e - b, mid = b + n / 2; bool have_split = false;
        Box cb; cb.reset();
        for (uint32_t i = b; i < e; ++i) for (int k = 0; k < 3; ++k) {
            float c = cen[3 * order[i] + k]; cb.lo[k] = std::min(cb.lo[k], c); cb.hi[k] = std::max(cb.hi[k], c); }
        if (depth < kMaxDepth) {
            if (n <= (uint32_t) kSweepBelow) {
                float best = INFINITY; int best_axis = -1; uint32_t best_split = 0;
                for (int axis = 0; axis < 3; ++axis) {
                    if (!(cb.hi[axis] > cb.lo[axis])) continue;
                    uint32_t s; float cost = sweep(b, e, axis, s);
                    if (cost < best) { best = cost; best_axis = axis; best_split = s; }
                }
                if (best_axis >= 0) { uint32_t s; sweep(b, e, best_axis, s); mid = b + best_split; have_split = true; }
            } else {
                float best = INFINITY; int best_axis = -1, best_bin = 0;
                for (int axis = 0; axis < 3; ++axis) {
                    float ext = cb.hi[axis] - cb.lo[axis]; if (!(ext > 0.f)) continue;
                    Box bb[kBins]; uint32_t cnt[kBins] = {}; for (auto &x : bb) x.reset();
                    float scale = (float) kBins / ext;
                    for (uint32_t i = b; i < e; ++i) {
                        uint32_t t = order[i]; int k = std::min(kBins - 1, (int) ((cen[3 * t + axis] - cb.lo[axis]) * scale));
                        bb[k].grow(box[t]); ++cnt[k];
                    }
                    float ra[kBins]; uint32_t rc[kBins]; Box acc; acc.reset(); uint32_t c = 0;
                    for (int k = kBins - 1; k > 0; --k) { acc.grow(bb[k]); c += cnt[k]; ra[k] = acc.half_area(); rc[k] = c; }
                    acc.reset(); c = 0;
                    for (int k = 1; k < kBins; ++k) {
                        acc.grow(bb[k - 1]); c += cnt[k - 1];
                        if (c == 0 || rc[k] == 0) continue;
                        float cost = acc.half_area() * (float) c + ra[k] * (float) rc[k];
                        if (cost < best) { best = cost; best_axis = axis; best_bin = k; }
                    }
                }
                if (best_axis >= 0) {
                    float ext = cb.hi[best_axis] - cb.lo[best_axis], scale = (float) kBins / ext, lo = cb.lo[best_axis];
                    auto it = std::partition(order.begin() + b, order.begin() + e, [&](uint32_t t) {
                        return std::min(kBins - 1, (int) ((cen[3 * t + best_axis] - lo) * scale)) < best_bin; });
                    mid = (uint32_t) (it - order.begin()); have_split = mid > b && mid < e;
                }
            }
        }
        if (!have_split) {                                    // coincident centroids or depth cap: median split on the widest axis
            int axis = 0; for (int k = 1; k < 3; ++k) if (cb.hi[k] - cb.lo[k] > cb.hi[axis] - cb.lo[axis]) axis = k;
            mid = b + n / 2;
            std::nth_element(order.begin() + b, order.begin() + mid, order.begin() + e, [&](uint32_t x, uint32_t y) {
                float cx = cen[3 * x + axis], cy = cen[3 * y + axis]; return cx < cy || (cx == cy && x < y); });
        }
        Box l, r; int cl = build(b, mid, depth + 1, l), cr = build(mid, e, depth + 1, r);
        DNode nd;
        nd.a = make_float4(l.lo[0], l.lo[1], l.lo[2], l.hi[0]); nd.b = make_float4(l.hi[1], l.hi[2], r.lo[0], r.lo[1]);
        nd.c = make_float4(r.lo[2], r.hi[0], r.hi[1], r.hi[2]);
        float fl, fr; memcpy(&fl, &cl, 4); memcpy(&fr, &cr, 4); nd.d = make_float4(fl, fr, 0.f, 0.f);
        nodes[self] = nd;
        bounds = l; bounds.grow(r);
        return self;
    }
};

}  // namespace sah

// ---------------------------------------------------------------------------------------------
// 4-wide tree for the wavefront TRACE kernels: the two-wide tree collapsed top-down (the inner child with the largest
// surface area is replaced by its two children until the node holds four), re-ordered breadth-first and quantized to DNodeQ
// (lr_scene.cuh).  Child boxes are the ones the two-wide nodes already carry, rounded outward to the node's 7-bit grid, so
// every box still contains its triangles (hits are decided by the exact triangle test, never by a box).
// Also computes the worst-case number of stack entries of a depth-first traversal that pushes every hit sibling.
// ---------------------------------------------------------------------------------------------
namespace bvh4 {

struct Child { int code; float lo[3], hi[3]; };
struct Wide { Child ch[4]; int n; };                          // one collapsed node; child codes >= 0 index `wide`

static void children_of(const DNode &n, Child &l, Child &r) {
    l.lo[0] = n.a.x; l.lo[1] = n.a.y; l.lo[2] = n.a.z; l.hi[0] = n.a.w; l.hi[1] = n.b.x; l.hi[2] = n.b.y;
    r.lo[0] = n.b.z; r.lo[1] = n.b.w; r.lo[2] = n.c.x; r.hi[0] = n.c.y; r.hi[1] = n.c.z; r.hi[2] = n.c.w;
    memcpy(&l.code, &n.d.x, 4); memcpy(&r.code, &n.d.y, 4);
}
static float half_area(const Child &c) {
    float dx = c.hi[0] - c.lo[0], dy = c.hi[1] - c.lo[1], dz = c.hi[2] - c.lo[2]; return dx * dy + dy * dz + dz * dx;
}

struct Collapser {
    const std::vector<DNode> &bin; std::vector<Wide> wide;
    explicit Collapser(const std::vector<DNode> &b) : bin(b) { wide.reserve(b.size() / 2 + 2); }
    // emits the 4-wide node for two-wide node `node`; returns its index, `need` = stack entries below it
    int emit(int node, uint32_t &need) {
        Wide w; w.n = 2; children_of(bin[node], w.ch[0], w.ch[1]);
        while (w.n < 4) {
            int best = -1; float best_area = -1.f;
            for (int i = 0; i < w.n; ++i) if (w.ch[i].code >= 0) { float a = half_area(w.ch[i]); if (a > best_area) { best_area = a; best = i; } }
            if (best < 0) break;
            Child l, r; children_of(bin[w.ch[best].code], l, r); w.ch[best] = l; w.ch[w.n++] = r;
        }
        int self = (int) wide.size(); wide.push_back(w);
        uint32_t below = 0;
        for (int i = 0; i < w.n; ++i) if (w.ch[i].code >= 0) { uint32_t nb = 0; w.ch[i].code = emit(w.ch[i].code, nb); below = std::max(below, nb); }
        wide[self] = w;
        need = (uint32_t) (w.n - 1) + below;                  // triangles are stack entries too (they are tested when popped)
        return self;
    }
};

// quantize one collapsed node; `remap`: collapsed index -> breadth-first index
static DNodeQ quantize(const Wide &w, const std::vector<int> &remap) {
    float lo[3] = { INFINITY, INFINITY, INFINITY }, hi[3] = { -INFINITY, -INFINITY, -INFINITY };
    for (int i = 0; i < w.n; ++i) for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], w.ch[i].lo[k]); hi[k] = std::max(hi[k], w.ch[i].hi[k]); }
    auto fin = [](float v) { return std::isfinite(v) ? v : (v > 0 ? 1e30f : -1e30f); };
    uint32_t ebits[3]; double cell[3], org[3];
    for (int k = 0; k < 3; ++k) {
        if (!std::isfinite(lo[k]) || !std::isfinite(hi[k])) { lo[k] = -1e30f; hi[k] = 1e30f; }       // the single-triangle node: unbounded box
        // cell = 2^e with 126 cells covering the extent (one cell of slack on either side for the outward rounding)
        double ext = std::max((double) hi[k] - (double) lo[k], 1e-30); int e = (int) std::ceil(std::log2(ext / 125.0));
        e = std::max(-100, std::min(100, e)); cell[k] = std::ldexp(1.0, e); ebits[k] = (uint32_t) (e + 127);
        // origin: an fp32 value at or below lo - one cell's worth of slack is not needed: floor() below rounds outward
        org[k] = (double) (float) lo[k]; if (org[k] > lo[k]) org[k] = (double) std::nextafterf((float) lo[k], -INFINITY);
    }
    DNodeQ q; uint32_t planes[6] = { 0, 0, 0, 0, 0, 0 }; int32_t links[4];
    for (int i = 0; i < 4; ++i) {
        if (i >= w.n) { links[i] = 0x7fffffff; for (int k = 0; k < 3; ++k) { planes[k] |= (0x80u | 127u) << (8 * i); planes[3 + k] |= 0x80u << (8 * i); } continue; }
        links[i] = w.ch[i].code >= 0 ? remap[w.ch[i].code] : w.ch[i].code;
        for (int k = 0; k < 3; ++k) {
            // outward rounding with a margin of 1e-3 cell against the fp32 evaluation order of the device
            int ql = (int) std::floor(((double) fin(w.ch[i].lo[k]) - org[k]) / cell[k] - 1e-3), qh = (int) std::ceil(((double) fin(w.ch[i].hi[k]) - org[k]) / cell[k] + 1e-3);
            ql = std::max(0, std::min(127, ql)); qh = std::max(0, std::min(127, qh));
            planes[k] |= (0x80u | (uint32_t) ql) << (8 * i); planes[3 + k] |= (0x80u | (uint32_t) qh) << (8 * i);
        }
    }
    float of[3] = { (float) org[0], (float) org[1], (float) org[2] }; uint32_t ob[3]; memcpy(ob, of, 12);
    q.c0 = make_uint4(ob[0], ob[1], ob[2], ebits[0] | (ebits[1] << 8) | (ebits[2] << 16));
    q.c1 = make_uint4((uint32_t) links[0], (uint32_t) links[1], (uint32_t) links[2], (uint32_t) links[3]);
    q.c2 = make_uint4(planes[0], planes[1], planes[2], planes[3]);
    q.c3 = make_uint4(planes[4], planes[5], 0u, 0u);
    return q;
}

}  // namespace bvh4

// `bin`: the two-wide nodes (host copy); uploads the collapsed tree and records it in the scene
static int finish_bvh4(lr_ctx *c, const std::vector<DNode> &bin, const std::vector<DTri> &sorted) {
    DScene &sc = c->sc; uint32_t n = (uint32_t) sorted.size();
    sc.nodes4 = nullptr; sc.bvh4_stack = 1; sc.n_nodes4 = 0; c->bvh4_nodes = 0;
    if (n == 0) return LR_OK;
    std::vector<bvh4::Wide> wide; uint32_t need = 0;
    if (n == 1) {                                             // a single triangle: one node with one leaf child, unbounded box
        bvh4::Wide w; w.n = 1; w.ch[0].code = ~0; for (int k = 0; k < 3; ++k) { w.ch[0].lo[k] = -INFINITY; w.ch[0].hi[k] = INFINITY; }
        wide.push_back(w);
    } else {
        bvh4::Collapser col(bin); col.emit(0, need); wide.swap(col.wide);
    }
    // breadth-first order: the top of the tree becomes a prefix of the array
    std::vector<int> order; order.reserve(wide.size()); std::vector<int> remap(wide.size(), -1);
    order.push_back(0);
    for (size_t h = 0; h < order.size(); ++h) { const bvh4::Wide &w = wide[order[h]]; for (int i = 0; i < w.n; ++i) if (w.ch[i].code >= 0) order.push_back(w.ch[i].code); }
    for (size_t i = 0; i < order.size(); ++i) remap[order[i]] = (int) i;
    std::vector<DNodeQ> nodes4(order.size());
    for (size_t i = 0; i < order.size(); ++i) nodes4[i] = bvh4::quantize(wide[order[i]], remap);
    DevBuf *dn = c->newbuf(); if (int r = upload(c, dn, nodes4.data(), nodes4.size())) return r;
    sc.nodes4 = dn->as<DNodeQ>(); sc.bvh4_stack = need + 1; sc.n_nodes4 = (uint32_t) nodes4.size(); c->bvh4_nodes = sc.n_nodes4;
    return LR_OK;
}

static int build_sah_bvh(lr_ctx *c, const std::vector<DTri> &tris, V3 lo, V3 hi) {
    uint32_t n = (uint32_t) tris.size(); DScene &sc = c->sc;
    V3 ext = hi - lo; float pad = 1e-6f * std::max(ext.x, std::max(ext.y, ext.z)) + 1e-30f;
    sah::Builder bld(tris, pad); sah::Box root;
    bld.build(0, n, 0, root);
    std::vector<DTri> sorted(n); for (uint32_t i = 0; i < n; ++i) sorted[i] = tris[bld.order[i]];
    c->tris_gid = c->newbuf(); if (int r = upload(c, c->tris_gid, tris.data(), n)) return r;
    DevBuf *ds = c->newbuf(), *dn = c->newbuf();
    if (int r = upload(c, ds, sorted.data(), n)) return r;
    if (int r = upload(c, dn, bld.nodes.data(), bld.nodes.size())) return r;
    sc.tris = ds->as<DTri>(); sc.nodes = dn->as<DNode>(); sc.n_tris = n;
    return finish_bvh4(c, bld.nodes, sorted);
}
