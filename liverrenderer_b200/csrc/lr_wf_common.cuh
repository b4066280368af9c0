// lr_wf_common.cuh -- pieces shared by the wavefront integrators: path-state flags, warp-ballot queue push, the 256-byte
// AoS slot + queue block of the wavefront `prbvolpath` (lr_wavefront_prb.cuh).  The forward `volpath` wavefront
// (lr_wavefront.cuh) keeps its state in SoA planes instead and only uses the flags / push helpers from here.
#pragma once
#include "lr_integrators.cuh"

// resident 256-thread blocks per SM the stage kernels are compiled for (register cap = 65536 / (256 * LR_WF_OCC))
#ifndef LR_WF_OCC
#define LR_WF_OCC 4
#endif
// resident 128-thread blocks per SM of the two surface-shading kernels (register cap = 65536 / (128 * LR_WF_OCC_SURF))
#ifndef LR_WF_OCC_SURF
#define LR_WF_OCC_SURF (2 * LR_WF_OCC)
#endif

namespace lr {

enum : uint32_t {
    WF_CHANNEL_MASK = 3u, WF_VALID_RAY = 1u << 2, WF_SPEC_CHAIN = 1u << 3, WF_NEE_ACTIVE = 1u << 4, WF_DEAD = 1u << 5,
    WF_NEE_NEEDS_ISECT = 1u << 6, WF_HAS_LANE = 1u << 7,
    WF_MEDIUM_SHIFT = 8, WF_MEDIUM_MASK = 0xffu << 8,          // 0xff: no medium
    WF_PASS_SHIFT = 16                                          // pass index of a multi-pass render (integrator.cpp:279-293)
};
#define WF_N_SURF 3                                             // surface buckets: 0 diffuse/null/dielectric, 1 roughdielectric, 2 principled

// One path slot, 256 B, 16-byte fields: a thread moves its slot with 128-bit loads/stores (full 32 B sectors).
struct __align__(16) WState {
    unsigned long long rng_state, rng_inc;
    float4 o_maxt;          // ray.o, ray.maxt
    float4 d_eta;           // ray.d, eta
    float4 thr_lpdf;        // throughput, last_scatter_direction_pdf
    float4 res_depth;       // result, depth (bits)
    float4 hit;             // t, prim (bits), u, v of the closest hit along the current ray
    float4 lsp_flags;       // last_scatter_p, flags (bits)
    float4 lane_pos;        // lane lo/hi (bits), film sample position
    // ---- NEE walk (volpath.cpp:400-554) ----
    float4 no_maxd;         // shadow ray origin, max_dist
    float4 nd_total;        // ds.d, total_dist
    float4 ntr_mis;         // transmittance, MIS weight
    float4 nw_med;          // throughput * (phase | bsdf) value, walk medium (bits)
    float4 ne_dist;         // emitter_val, ds.dist
    float4 nhit;            // closest hit along the shadow ray
    float4 pad0, pad1;
};
static_assert(sizeof(WState) == 256, "WState layout");

struct WQueues {                       // device-side counters, one 64-byte block
    uint32_t n_active[2]; uint32_t n_scatter; uint32_t n_surf[WF_N_SURF]; uint32_t n_regen; uint32_t live;
    unsigned long long lane_next; uint32_t n_nee; uint32_t cursor; uint32_t pad[4];   // cursor: work cursor of k_wf_track_dda
};

struct WFArgs {
    WState *st; WQueues *q; uint32_t *q_active[2]; uint32_t *q_scatter; uint32_t *q_surf[WF_N_SURF]; uint32_t *q_regen; uint32_t *q_nee;   // q_nee: slots whose NEE walk is live
    uint32_t cap; int cur;            // slots; index of the input active queue
    const DTri *tris_gid; float *film; float *lanes_out;
    uint64_t lane_begin, lane_end; uint32_t seed, spp_per_pass, n_passes;
    unsigned long long *stats; LaneMap map;
};

// ---- queue push: ballot + popc, one atomic per warp; call at a point where the whole warp is converged ----
LR_DEV void wf_push(uint32_t *queue, uint32_t *count, uint32_t slot, bool pred) {
    unsigned m = __ballot_sync(0xffffffffu, pred);
    if (m == 0) return;
    unsigned lane = threadIdx.x & 31; uint32_t base = 0;
    if (lane == (unsigned) (__ffs(m) - 1)) base = atomicAdd(count, (uint32_t) __popc(m));
    base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
    if (pred) queue[base + __popc(m & ((1u << lane) - 1u))] = slot;
}

LR_DEV uint32_t wf_flags(const WState &S) { return __float_as_uint(S.lsp_flags.w); }
LR_DEV int wf_medium(uint32_t f) { uint32_t m = (f & WF_MEDIUM_MASK) >> WF_MEDIUM_SHIFT; return m == 0xffu ? -1 : (int) m; }
LR_DEV uint32_t wf_set_medium(uint32_t f, int m) { return (f & ~WF_MEDIUM_MASK) | (((uint32_t) (m < 0 ? 0xff : m)) << WF_MEDIUM_SHIFT); }
LR_DEV Hit wf_hit(float4 h) { Hit r; r.t = h.x; r.prim = __float_as_uint(h.y); r.u = h.z; r.v = h.w; return r; }
LR_DEV float4 wf_hit4(const Hit &h) { return make_float4(h.t, __uint_as_float(h.prim), h.u, h.v); }
LR_DEV int wf_surf_bucket(const DScene &sc, const DTri *tris_gid, const Hit &h) {
    if (!hit_valid(h)) return 0;
    uint32_t k = sc.bsdfs[sc.meshes[__float_as_uint(tris_gid[h.prim].v1.w)].bsdf].kind;
    return k == LR_BSDF_ROUGHDIELECTRIC ? 1 : (k == LR_BSDF_PRINCIPLED ? 2 : 0);
}

// end of a path: splat (ImageBlock::put) or per-lane output
template <int C, bool TO_FILM>
LR_DEV void wf_finish(const DScene &sc, const WFArgs &a, const WState &S, V3 L, uint32_t flags) {
    bool valid = (flags & WF_VALID_RAY) != 0;
    if (TO_FILM) {
        float aovs[C]; aovs[0] = L.x; aovs[1] = L.y; aovs[2] = L.z;
        if (C == 5) { aovs[3] = valid ? 1.f : 0.f; aovs[4] = 1.f; } else aovs[3] = 1.f;
        film_put<C>(sc, a.film, mk2(S.lane_pos.z, S.lane_pos.w), aovs, a.spp_per_pass >= 4);
    } else {
        uint64_t lane = ((uint64_t) __float_as_uint(S.lane_pos.y) << 32) | __float_as_uint(S.lane_pos.x);
        uint32_t pass = flags >> WF_PASS_SHIFT;
        reinterpret_cast<float4 *>(a.lanes_out)[(lane_local(a.map, lane) - a.lane_begin) * a.n_passes + pass] = make_float4(L.x, L.y, L.z, valid ? 1.f : 0.f);
    }
}

#define WF_FOR_QUEUE(COUNT)                                                                             \
    const uint32_t _n = (COUNT);                                                                        \
    const uint32_t _stride = gridDim.x * blockDim.x;                                                    \
    for (uint32_t _base = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; _base < _n; _base += _stride) \
        if (const uint32_t qi = _base + (threadIdx.x & 31u); true)

__global__ void k_wf_begin_iteration(WQueues *q, int cur) {     // counters of the queues this iteration fills
    q->n_active[cur ^ 1] = 0; q->n_scatter = 0; q->n_regen = 0; q->n_nee = 0; q->cursor = 0;
    for (int b = 0; b < WF_N_SURF; ++b) q->n_surf[b] = 0;
}
__global__ void k_wf_init(WQueues *q, uint32_t *q_regen, uint32_t cap, WState *st) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { q->n_active[0] = q->n_active[1] = 0; q->n_scatter = 0; q->n_regen = cap; q->lane_next = 0; q->n_nee = 0; q->cursor = 0; for (int b = 0; b < WF_N_SURF; ++b) q->n_surf[b] = 0; }
    if (i < cap) { q_regen[i] = i; st[i].lsp_flags.w = __uint_as_float(0u); }
}

} // namespace lr
