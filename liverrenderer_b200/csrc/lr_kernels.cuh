// lr_kernels.cuh -- __global__ entry points: LBVH build, lane rendering, film develop.
#pragma once
#include "lr_integrators.cuh"
#include "lr_prb.cuh"
#include "lr_biovolpath.cuh"
#include "lr_volpathmis.cuh"

namespace lr {

#ifndef LR_LANES_TU      // set-up / utility kernels live in lr_backend.cu only
// ------------------------------------------------------------------------------------------
// LBVH build (Karras 2012): Morton keys -> radix sort (host calls cub) -> hierarchy -> refit.
// ------------------------------------------------------------------------------------------
LR_DEV uint32_t expand_bits10(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu; v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u; v = (v * 0x00000005u) & 0x49249249u; return v;
}
__global__ void k_morton(const DTri *tris_gid, uint32_t n, V3 lo, V3 inv_ext, uint64_t *keys, uint32_t *vals) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; if (i >= n) return;
    const DTri t = tris_gid[i];
    V3 c = (f4xyz(t.v0) + f4xyz(t.v1) + f4xyz(t.v2)) * (1.f / 3.f);
    float x = fminf(fmaxf((c.x - lo.x) * inv_ext.x * 1024.f, 0.f), 1023.f), y = fminf(fmaxf((c.y - lo.y) * inv_ext.y * 1024.f, 0.f), 1023.f),
          z = fminf(fmaxf((c.z - lo.z) * inv_ext.z * 1024.f, 0.f), 1023.f);
    uint32_t m = (expand_bits10((uint32_t) x) << 2) | (expand_bits10((uint32_t) y) << 1) | expand_bits10((uint32_t) z);
    keys[i] = ((uint64_t) m << 32) | i; vals[i] = i;
}
LR_DEV int lbvh_delta(const uint64_t *keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    return __clzll((long long) (keys[i] ^ keys[j]));
}
// children: >= 0 internal node index, < 0 => ~leaf index (leaf k = k-th triangle in Morton order)
__global__ void k_hierarchy(const uint64_t *keys, int n, int2 *children, int *parent_internal, int *parent_leaf) {
    int i = blockIdx.x * blockDim.x + threadIdx.x; if (i >= n - 1) return;
    int d = (lbvh_delta(keys, n, i, i + 1) - lbvh_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = lbvh_delta(keys, n, i, i - d);
    int lmax = 2; while (lbvh_delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0; for (int t = lmax >> 1; t >= 1; t >>= 1) if (lbvh_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = lbvh_delta(keys, n, i, j);
    int s = 0; int t = l;
    do { t = (t + 1) >> 1; if (lbvh_delta(keys, n, i, i + (s + t) * d) > dnode) s += t; } while (t > 1);
    int gamma = i + s * d + min(d, 0);
    int left = (min(i, j) == gamma) ? ~gamma : gamma, right = (max(i, j) == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    children[i] = make_int2(left, right);
    if (left < 0) parent_leaf[~left] = i; else parent_internal[left] = i;
    if (right < 0) parent_leaf[~right] = i; else parent_internal[right] = i;
    if (i == 0) parent_internal[0] = -1;
}
// Bottom-up refit.  Box storage: [0, n-1) internal, [n-1, 2n-1) leaves.  Leaf boxes are padded by
// `pad` (absolute) so the exact Moeller-Trumbore test, not the slab test, decides every hit.
__global__ void k_refit(const DTri *tris_sorted, int n, const int2 *children, const int *parent_internal, const int *parent_leaf,
                        float4 *box_lo, float4 *box_hi, int *flags, float pad) {
    int k = blockIdx.x * blockDim.x + threadIdx.x; if (k >= n) return;
    const DTri t = tris_sorted[k];
    V3 a = f4xyz(t.v0), b = f4xyz(t.v1), c = f4xyz(t.v2);
    V3 lo = mk3(fminf(a.x, fminf(b.x, c.x)) - pad, fminf(a.y, fminf(b.y, c.y)) - pad, fminf(a.z, fminf(b.z, c.z)) - pad);
    V3 hi = mk3(fmaxf(a.x, fmaxf(b.x, c.x)) + pad, fmaxf(a.y, fmaxf(b.y, c.y)) + pad, fmaxf(a.z, fmaxf(b.z, c.z)) + pad);
    box_lo[n - 1 + k] = make_float4(lo.x, lo.y, lo.z, 0.f); box_hi[n - 1 + k] = make_float4(hi.x, hi.y, hi.z, 0.f);
    int node = parent_leaf[k];
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(&flags[node], 1) == 0) return;         // first arrival: the sibling will finish the node
        int2 ch = children[node];
        int il = ch.x < 0 ? n - 1 + (~ch.x) : ch.x, ir = ch.y < 0 ? n - 1 + (~ch.y) : ch.y;
        float4 l0 = box_lo[il], h0 = box_hi[il], l1 = box_lo[ir], h1 = box_hi[ir];
        box_lo[node] = make_float4(fminf(l0.x, l1.x), fminf(l0.y, l1.y), fminf(l0.z, l1.z), 0.f);
        box_hi[node] = make_float4(fmaxf(h0.x, h1.x), fmaxf(h0.y, h1.y), fmaxf(h0.z, h1.z), 0.f);
        node = parent_internal[node];
    }
}
__global__ void k_pack_nodes(int n, const int2 *children, const float4 *box_lo, const float4 *box_hi, DNode *nodes) {
    int i = blockIdx.x * blockDim.x + threadIdx.x; if (i >= n - 1) return;
    int2 ch = children[i];
    int il = ch.x < 0 ? n - 1 + (~ch.x) : ch.x, ir = ch.y < 0 ? n - 1 + (~ch.y) : ch.y;
    float4 l0 = box_lo[il], h0 = box_hi[il], l1 = box_lo[ir], h1 = box_hi[ir];
    DNode nd;
    nd.a = make_float4(l0.x, l0.y, l0.z, h0.x); nd.b = make_float4(h0.y, h0.z, l1.x, l1.y); nd.c = make_float4(l1.z, h1.x, h1.y, h1.z);
    nd.d = make_float4(__int_as_float(ch.x), __int_as_float(ch.y), 0.f, 0.f);
    nodes[i] = nd;
}
__global__ void k_gather_tris(const DTri *tris_gid, const uint32_t *order, uint32_t n, DTri *tris_sorted) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; if (i >= n) return;
    tris_sorted[i] = tris_gid[order[i]];
}

#endif // !LR_LANES_TU

// ------------------------------------------------------------------------------------------
// Lane rendering.  Persistent warps pull 32 consecutive global lanes at a time from a device
// counter (primary rays of a warp stay pixel-coherent); a lane index keys the PCG32 stream, so
// the result does not depend on how lanes are distributed over warps, kernels or GPUs.
// ------------------------------------------------------------------------------------------
// Strided lane sets (multi-GPU split of ONE frame by sample index): with len != 0 a call renders only the lanes l of the whole
// render with (l mod period) in [off, off + len); [lane_begin, lane_end) then counts LOCAL indices i, lane = (i / len) * period +
// off + (i mod len).  period = spp_per_pass, len = spp_per_pass / N gives every rank spp / N samples of every pixel.  Streams stay
// keyed by the GLOBAL lane, so the union of the ranks' films is the single-GPU film.  len == 0: identity.
struct LaneMap { uint64_t period, off, len; };
LR_DEV uint64_t lane_global(const LaneMap &m, uint64_t i) { return m.len ? (i / m.len) * m.period + m.off + (i % m.len) : i; }
LR_DEV uint64_t lane_local(const LaneMap &m, uint64_t lane) { return m.len ? (lane / m.period) * m.len + (lane % m.period - m.off) : lane; }

struct RenderArgs {
    const DTri *tris_gid; float *film; float *lanes_out;
    uint64_t lane_begin, lane_end; uint32_t seed, spp_per_pass, n_passes, total_spp;     // local indices when map.len != 0
    LaneMap map;
    unsigned long long *work_counter, *stats; const volatile int *cancel;
    // adjoint
    const float *grad_in; const float *weight_img; uint32_t grad_channels;
};

struct LaneSetup { Sampler smp; V2 pos, sample_pos; Ray ray; };
// lane -> (pixel, stream): integrator.cpp:321-338 + sampler.cpp:129-148; jitter + camera ray: :465-480
// sampler->seed(seed, wavefront_size) of the render loop (integrator.cpp:303-308) for one lane
LR_DEV void seed_sampler(const DScene &sc, Sampler &smp, uint32_t seed, uint64_t lane, uint32_t spp_per_pass, uint32_t n_passes) {
    if (sc.integ.sampler == LR_SAMPLER_LD) smp.seed_ld(sc.integ.base_seed, seed, (uint32_t) lane, spp_per_pass, spp_per_pass * n_passes);
    else smp.seed(sc.integ.base_seed + seed, (uint32_t) lane);
}
LR_DEV void lane_begin_sample(const DScene &sc, LaneSetup &L, uint64_t lane, uint32_t spp_per_pass) {
    uint32_t W = sc.sensor.film_width;
    uint32_t pix = (uint32_t) (lane / spp_per_pass), py = pix / W, px = pix - W * py;
    L.pos = mk2((float) px, (float) py);
}
LR_DEV void lane_next_ray(const DScene &sc, LaneSetup &L) {
    V2 j = L.smp.next_2d();
    L.sample_pos = mk2(L.pos.x + j.x, L.pos.y + j.y);                       // film position, relative to the crop window
    // integrator.cpp:338, :465-468: pos += crop_offset; scale = 1 / crop_size, offset = -crop_offset * scale;
    // adjusted_pos = fmadd(pos + jitter, scale, offset).  Without a crop this is sample_pos * scale exactly (offset = -0).
    const float sx = 1.f / (float) sc.sensor.film_width, sy = 1.f / (float) sc.sensor.film_height;
    const float ox = (float) sc.sensor.crop_offset_x, oy = (float) sc.sensor.crop_offset_y;
    V2 adjusted = mk2(fmaf((L.pos.x + ox) + j.x, sx, -(ox * sx)), fmaf((L.pos.y + oy) + j.y, sy, -(oy * sy)));
    L.ray = camera_ray(sc, adjusted);
}

#define LR_KIND_VOLPATHMIS_PLAIN 5      // kernel instantiation only: volpathmis with use_spectral_mis = false
template <int KIND>
LR_DEV LaneResult run_lane(const DScene &sc, const DTri *tris_gid, LaneSetup &L, Counters &cnt) {
    if (KIND == LR_INTEGRATOR_PATH) return sample_path(sc, tris_gid, L.smp, L.ray, cnt);
    if (KIND == LR_INTEGRATOR_VOLPATH) return sample_volpath(sc, tris_gid, L.smp, L.ray, sc.sensor.medium, cnt);
    if (KIND == LR_INTEGRATOR_BIOVOLPATH) return sample_biovolpath(sc, tris_gid, L.smp, L.ray, sc.sensor.medium, cnt);
    if (KIND == LR_INTEGRATOR_VOLPATHMIS) return sample_volpathmis<true>(sc, tris_gid, L.smp, L.ray, sc.sensor.medium, cnt);
    if (KIND == LR_KIND_VOLPATHMIS_PLAIN) return sample_volpathmis<false>(sc, tris_gid, L.smp, L.ray, sc.sensor.medium, cnt);
    return sample_prb<true>(sc, tris_gid, L.smp, L.ray, mk3(0.f), mk3(0.f), cnt);
}

LR_DEV void flush_counters(Counters &cnt, unsigned long long nsamples, unsigned long long *stats) {
    unsigned long long r = cnt.rays, c = cnt.collisions, s = cnt.real_scatters, n = nsamples;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        r += __shfl_xor_sync(0xffffffffu, r, o); c += __shfl_xor_sync(0xffffffffu, c, o);
        s += __shfl_xor_sync(0xffffffffu, s, o); n += __shfl_xor_sync(0xffffffffu, n, o);
    }
    if ((threadIdx.x & 31) == 0) { atomicAdd(stats + 0, n); atomicAdd(stats + 1, r); atomicAdd(stats + 2, c); atomicAdd(stats + 3, s); }
}

// KIND: lr_integrator_kind; C: film channels (4 = RGBW, 5 = RGBAW); TO_FILM: splat vs. per-lane output
template <int KIND, int C, bool TO_FILM>
__global__ void __launch_bounds__(128) k_render(const __grid_constant__ DScene sc, const __grid_constant__ RenderArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0; unsigned long long nsamples = 0;
    const uint32_t lane_id = threadIdx.x & 31;
    const bool box = sc.sensor.rfilter == LR_RFILTER_BOX, coalesce = a.spp_per_pass >= 4;
    for (uint32_t it = 0;; ++it) {
        unsigned long long base = 0;
        // the cancel flag lives in mapped host memory: one read costs a PCIe round trip, so a warp looks at it when it starts and
        // then every 64th batch (a read per batch capped the whole kernel at ~1 M batches/s)
        if (lane_id == 0) base = ((it & 63u) == 0u && *a.cancel) ? ~0ull : atomicAdd(a.work_counter, 32ull);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base == ~0ull || a.lane_begin + base >= a.lane_end) break;
        const uint64_t li = a.lane_begin + base + lane_id;
        if (li < a.lane_end) {
            const uint64_t lane = lane_global(a.map, li);
            LaneSetup L; seed_sampler(sc, L.smp, a.seed, lane, a.spp_per_pass, a.n_passes);
            lane_begin_sample(sc, L, lane, a.spp_per_pass);
            for (uint32_t pass = 0; pass < a.n_passes; ++pass) {
                if (pass) L.smp.advance(a.spp_per_pass);
                lane_next_ray(sc, L);
                LaneResult r = run_lane<KIND>(sc, a.tris_gid, L, cnt);
                nsamples++;
                if (TO_FILM) {
                    float aovs[C]; aovs[0] = r.L.x; aovs[1] = r.L.y; aovs[2] = r.L.z;
                    if (C == 5) { aovs[3] = r.valid ? 1.f : 0.f; aovs[4] = 1.f; } else aovs[3] = 1.f;
                    film_put<C>(sc, a.film, box ? L.pos : L.sample_pos, aovs, coalesce);
                } else {
                    float4 o = make_float4(r.L.x, r.L.y, r.L.z, r.valid ? 1.f : 0.f);
                    reinterpret_cast<float4 *>(a.lanes_out)[(li - a.lane_begin) * a.n_passes + pass] = o;
                }
            }
        }
    }
    flush_counters(cnt, nsamples, a.stats);
}

#ifndef LR_LANES_TU
// HDRFilm::develop: out = value / (W == 0 ? 1 : W)
__global__ void k_develop(const float *film, float *out, uint32_t npix, int C) {
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; if (p >= npix) return;
    float w = film[(size_t) p * C + C - 1]; float d = w == 0.f ? 1.f : w;
    for (int k = 0; k < C - 1; ++k) out[(size_t) p * (C - 1) + k] = film[(size_t) p * C + k] / d;
}

// rays[n][8] -> hits[n][4] (parity probe for the LBVH)
__global__ void k_ray_probe(const __grid_constant__ DScene sc, const float *rays, uint64_t n, float *hits) {
    uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; if (i >= n) return;
    const float *r = rays + i * 8; Ray ray; ray.o = mk3(r[0], r[1], r[2]); ray.d = mk3(r[3], r[4], r[5]); ray.maxt = r[6];
    Hit h = bvh_trace<false, false>(sc, ray);
    float *o = hits + i * 4; o[0] = h.t; o[1] = __uint_as_float(h.prim); o[2] = h.u; o[3] = h.v;
}

// max over all channels (VolumeGrid max, src/render/volumegrid.cpp:72-80)
__global__ void k_grid_max(const float *data, size_t n, float *out) {
    float m = 0.f;
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) m = fmaxf(m, data[i]);
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) atomicMax(reinterpret_cast<int *>(out), __float_as_int(m));   // non-negative floats order like ints
}

// ------------------------------------------------------------------------------------------
// Adjoint (RBIntegrator.render_backward, src/python/python/ad/integrators/common.py:625-783)
// ------------------------------------------------------------------------------------------
// pass 0: splat w = 1 for every lane -> per-pixel filter-weight sum W (the develop() normaliser)
__global__ void __launch_bounds__(256) k_weight_image(const __grid_constant__ DScene sc, float *wimg, uint64_t n_lanes, uint32_t seed, uint32_t spp) {
    const bool box = sc.sensor.rfilter == LR_RFILTER_BOX, coalesce = spp >= 4;
    for (uint64_t lane = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; lane < n_lanes; lane += (uint64_t) gridDim.x * blockDim.x) {
        LaneSetup L; seed_sampler(sc, L.smp, seed, lane, spp, 1);
        lane_begin_sample(sc, L, lane, spp);
        V2 j = L.smp.next_2d(); L.sample_pos = mk2(L.pos.x + j.x, L.pos.y + j.y);
        float one = 1.f; film_put<1>(sc, wimg, box ? L.pos : L.sample_pos, &one, coalesce);
    }
}
#endif // !LR_LANES_TU

#ifdef LR_LANES_TU         // the per-lane adjoint kernel is compiled in lr_lanes.cu only
// passes 1+2: primal (cloned sampler) then adjoint replay with dL = sum_px w(px) grad_in[px] / W[px]
__global__ void __launch_bounds__(128) k_backward(const __grid_constant__ DScene sc, const __grid_constant__ RenderArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0; unsigned long long nsamples = 0;
    const uint32_t lane_id = threadIdx.x & 31;
    const bool box = sc.sensor.rfilter == LR_RFILTER_BOX, coalesce = a.spp_per_pass >= 4;
    const int W = (int) sc.sensor.film_width, H = (int) sc.sensor.film_height;
    for (uint32_t it = 0;; ++it) {
        unsigned long long base = 0;
        // the cancel flag lives in mapped host memory: one read costs a PCIe round trip, so a warp looks at it when it starts and
        // then every 64th batch (a read per batch capped the whole kernel at ~1 M batches/s)
        if (lane_id == 0) base = ((it & 63u) == 0u && *a.cancel) ? ~0ull : atomicAdd(a.work_counter, 32ull);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base == ~0ull || a.lane_begin + base >= a.lane_end) break;
        const uint64_t li = a.lane_begin + base + lane_id;
        if (li < a.lane_end) {
            const uint64_t lane = lane_global(a.map, li);
            LaneSetup L; seed_sampler(sc, L.smp, a.seed, lane, a.spp_per_pass, a.n_passes);
            lane_begin_sample(sc, L, lane, a.spp_per_pass);
            lane_next_ray(sc, L);
            Footprint fp; V3 dL = mk3(0.f);
            if (film_footprint(sc, box ? L.pos : L.sample_pos, coalesce, fp)) {
                for (int j = 0; j < fp.ny; ++j) { int y = fp.y0 + j; if (y < 0 || y >= H) continue;
                    for (int i = 0; i < fp.nx; ++i) { int x = fp.x0 + i; if (x < 0 || x >= W) continue;
                        float w = __fmul_rn(fp.wy[j], fp.wx[i]); float ws = a.weight_img[(size_t) y * W + x]; float d = ws == 0.f ? 1.f : ws;
                        const float *g = a.grad_in + ((size_t) y * W + x) * a.grad_channels;
                        float s = w / d; dL.x = fmaf(g[0], s, dL.x); dL.y = fmaf(g[1], s, dL.y); dL.z = fmaf(g[2], s, dL.z); } }
            }
            Sampler s1 = L.smp;
            LaneResult p = sample_prb<true>(sc, a.tris_gid, s1, L.ray, mk3(0.f), mk3(0.f), cnt);
            sample_prb<false>(sc, a.tris_gid, L.smp, L.ray, dL, p.L, cnt);
            nsamples++;
        }
    }
    flush_counters(cnt, nsamples, a.stats);
}

#endif // LR_LANES_TU

} // namespace lr

#ifndef LR_LANES_TU
#include "lr_wavefront.cuh"
#include "lr_wavefront_prb.cuh"
#include "lr_wavefront_bio.cuh"
#include "lr_wavefront_mis.cuh"
#endif
