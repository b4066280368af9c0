// lr_backend.cu -- host side of the C ABI (include/lr_api.h): context, scene upload
// (flatten -> HBM, LBVH build, envmap warp, majorants), kernel launches, film / gradient I/O.
// sm_100a only; no CPU path: every entry point fails with LR_ERR_NO_DEVICE without a GPU.
#include "lr_kernels.cuh"
#include "lr_lanes.h"
#include <cub/device/device_radix_sort.cuh>
#include <nvtx3/nvToolsExt.h>          // header-only: ranges are no-ops unless a tool (nsys / ncu --nvtx) is attached
#include <vector>
#include <string>
#include <algorithm>
#include <cstring>
#include <cstdio>
#include <cmath>
#include <cstdlib>

using namespace lr;

static thread_local std::string g_last_error;

struct DevBuf {
    void *p = nullptr; size_t bytes = 0;
    ~DevBuf() { if (p) cudaFree(p); }
    DevBuf() {}
    DevBuf(const DevBuf &) = delete; DevBuf &operator=(const DevBuf &) = delete;
    cudaError_t alloc(size_t n) { if (p) { cudaFree(p); p = nullptr; } bytes = n; return n ? cudaMalloc(&p, n) : cudaSuccess; }
    template <typename T> T *as() const { return (T *) p; }
};

// forward wavefront pool (lr_wavefront.cuh): SoA state planes + index queues + counters, and the mapped pinned ring the
// iterations write their live counts to
struct WVRing { uint32_t live; uint32_t pad; unsigned long long lane_next; };
struct WVPool {
    DevBuf planes, queues, counters, trace_ovf, mis; uint32_t cap = 0;      // mis: 128 B per slot of p / f state, volpathmis scenes only
    WVRing *ring_host = nullptr, *ring_dev = nullptr; std::vector<cudaEvent_t> ring_ev;
    ~WVPool() { if (ring_host) cudaFreeHost(ring_host); for (auto e : ring_ev) cudaEventDestroy(e); }
};

struct lr_ctx {
    int device = 0; cudaStream_t stream = nullptr; bool own_stream = false;
    int smaj_grid = -1; uint32_t smaj_floats = 0;         // majorant super-grid staged in shared memory by k_wv_track_dda (0: none)
    bool nvtx_open = false; bool nvtx = getenv("LR_NVTX") != nullptr;      // per-stage / per-call NVTX ranges (include/mitsuba/core/profiler.h:20-49 has ITT / NVTX scopes per phase)
    LaneMap lane_map{0, 0, 0};                      // lr_set_lane_map: strided lane set of the following render / backward calls
    void *noalloc_ptr = nullptr;                    // device address of g_noalloc_scratch on `device` (lr_noalloc.inl)
    std::string err; bool has_scene = false;
    DScene sc{};
    // device storage
    std::vector<DevBuf *> pool;                     // everything referenced by sc
    DevBuf film, weight_img, work_counter, stats, scratch, prb_L;
    DevBuf *tris_gid = nullptr;
    std::vector<DGrid> h_grids; DevBuf *d_grids = nullptr;
    std::vector<DMedium> h_media; DevBuf *d_media = nullptr;
    std::vector<size_t> grid_floats;
    std::vector<float> grid_fixed_max;              // `max_value` override (grid.cpp:303-306, 324), < 0: none
    int film_channels = 4; uint32_t W = 0, H = 0;
    int *cancel_host = nullptr; int *cancel_dev = nullptr;      // mapped pinned flag polled by the kernels
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    lr_stats st{};
    int blocks_per_sm[8] = {0}; int n_sm = 148;
    // wavefront volpath (lr_wavefront.cuh / lr_wavefront_host.inl)
    WVPool *wv[4] = {nullptr, nullptr, nullptr, nullptr}; cudaStream_t stream_more[3] = {nullptr, nullptr, nullptr};   // forward pools: batches in flight
    uint32_t bvh4_nodes = 0; bool bvh4_small = false; uint32_t trace_n_smem = 0, trace_sd = 1; size_t trace_smem = 0;   // trace-kernel shared-memory plan
    DevBuf lane_cursor;                                                // device-side lane cursor shared by the batches of one render
    bool profile_on = false; std::vector<cudaEvent_t> prof_ev; double prof_ms[16] = {0}; uint64_t prof_launches[16] = {0}, prof_items[16] = {0};
    cudaEvent_t fork_ev = nullptr; int wf_runs = 3, wf_grid_mult = LR_WF_OCC, wf_grid_mult128 = LR_WF_OCC_SURF;   // one resident wave of CTAs per stage launch (profiles/r1_ab_register_budgets.txt)
    uint64_t wf_pool_slots = 1ull << 25; uint32_t surf_bucket_mask = 0; bool use_wavefront = true, all_media_dda = false;
    uint64_t wf_launches = 0, wf_tracking_launches = 0; double wf_tracking_ms = 0, wf_track_ms = 0, wf_nee_ms = 0;
    DevBuf *newbuf() { pool.push_back(new DevBuf()); return pool.back(); }
    void free_scene() { for (auto *b : pool) delete b; pool.clear(); h_grids.clear(); h_media.clear(); grid_floats.clear(); grid_fixed_max.clear(); has_scene = false; }
};

#define LR_FAIL(ctx, code, ...) do { char _b[512]; snprintf(_b, sizeof _b, __VA_ARGS__); if (ctx) (ctx)->err = _b; g_last_error = _b; return (code); } while (0)
#define LR_CUDA(ctx, call) do { cudaError_t _e = (call); if (_e != cudaSuccess) { \
    LR_FAIL(ctx, _e == cudaErrorMemoryAllocation ? LR_ERR_OOM : LR_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__, __LINE__); } } while (0)

template <typename T> static int upload(lr_ctx *c, DevBuf *b, const T *src, size_t n) {
    LR_CUDA(c, b->alloc(std::max<size_t>(n, 1) * sizeof(T)));
    if (n) LR_CUDA(c, cudaMemcpyAsync(b->p, src, n * sizeof(T), cudaMemcpyHostToDevice, c->stream));
    return LR_OK;
}

extern "C" {

int lr_api_version(void) { return LR_API_VERSION; }

const char *lr_last_error(const lr_ctx *ctx) { return ctx ? ctx->err.c_str() : g_last_error.c_str(); }

int lr_create(int device, lr_ctx **out) {
    if (!out) LR_FAIL((lr_ctx *) nullptr, LR_ERR_INVALID_ARG, "lr_create: out is NULL");
    *out = nullptr;
    int n = 0; cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        LR_FAIL((lr_ctx *) nullptr, LR_ERR_NO_DEVICE, "lr_create: no CUDA device available (%s); this backend has no CPU path", cudaGetErrorString(e));
    if (device < 0 || device >= n) LR_FAIL((lr_ctx *) nullptr, LR_ERR_INVALID_ARG, "lr_create: device %d out of range (0..%d)", device, n - 1);
    lr_ctx *c = new lr_ctx(); c->device = device;
    if (const char *e = getenv("LR_WF_RUNS")) { int v = atoi(e); if (v >= 1 && v <= 4) c->wf_runs = v; }
    if (const char *e = getenv("LR_WF_GRID")) { int v = atoi(e); if (v >= 1 && v <= 64) c->wf_grid_mult = v; }
    if (const char *e = getenv("LR_WF_GRID128")) { int v = atoi(e); if (v >= 1 && v <= 64) c->wf_grid_mult128 = v; }
    if (const char *e = getenv("LR_WF_SLOTS")) { long long v = atoll(e); if (v >= 1024) c->wf_pool_slots = (uint64_t) v; }
    if (const char *e = getenv("LR_MEGAKERNEL")) c->use_wavefront = atoi(e) == 0;
    LR_CUDA(c, cudaSetDevice(device));
    cudaDeviceProp prop; LR_CUDA(c, cudaGetDeviceProperties(&prop, device));
    c->n_sm = prop.multiProcessorCount;
    LR_CUDA(c, cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)); c->own_stream = true;
    LR_CUDA(c, cudaEventCreate(&c->ev0)); LR_CUDA(c, cudaEventCreate(&c->ev1));
    LR_CUDA(c, cudaHostAlloc((void **) &c->cancel_host, sizeof(int), cudaHostAllocMapped));
    *c->cancel_host = 0;
    LR_CUDA(c, cudaHostGetDevicePointer((void **) &c->cancel_dev, c->cancel_host, 0));
    LR_CUDA(c, c->work_counter.alloc(sizeof(unsigned long long))); LR_CUDA(c, c->stats.alloc(4 * sizeof(unsigned long long)));
    *out = c; return LR_OK;
}

void lr_destroy(lr_ctx *c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    c->free_scene();
    for (int k = 0; k < 4; ++k) { delete c->wv[k]; c->wv[k] = nullptr; }
    for (int k = 0; k < 3; ++k) if (c->stream_more[k]) { cudaStreamDestroy(c->stream_more[k]); c->stream_more[k] = nullptr; }
    for (auto e : c->prof_ev) cudaEventDestroy(e);
    c->prof_ev.clear();
    if (c->fork_ev) { cudaEventDestroy(c->fork_ev); c->fork_ev = nullptr; }
    if (c->cancel_host) cudaFreeHost(c->cancel_host);
    if (c->ev0) cudaEventDestroy(c->ev0); if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int lr_set_stream(lr_ctx *c, void *s) {
    if (!c) return LR_ERR_INVALID_ARG;
    if (c->own_stream && c->stream) { cudaStreamSynchronize(c->stream); cudaStreamDestroy(c->stream); }
    c->stream = (cudaStream_t) s; c->own_stream = false; return LR_OK;
}

int lr_cancel(lr_ctx *c) { if (!c || !c->cancel_host) return LR_ERR_INVALID_ARG; *(volatile int *) c->cancel_host = 1; return LR_OK; }

} // extern "C"

// ---------------------------------------------------------------------------------------------
// scene upload helpers
// ---------------------------------------------------------------------------------------------
static void mat_copy(M4 &d, const float *s) { memcpy(d.m, s, 64); }

static void invert4(const float *A, float *R) {                 // general 4x4 inverse, double precision (set-up only)
    double m[16], inv[16]; for (int i = 0; i < 16; ++i) m[i] = A[i];
    auto cof = [&](int r, int c) {
        double s[9]; int k = 0;
        for (int i = 0; i < 4; ++i) if (i != r) for (int j = 0; j < 4; ++j) if (j != c) s[k++] = m[i * 4 + j];
        double d = s[0] * (s[4] * s[8] - s[5] * s[7]) - s[1] * (s[3] * s[8] - s[5] * s[6]) + s[2] * (s[3] * s[7] - s[4] * s[6]);
        return ((r + c) & 1) ? -d : d;
    };
    double det = 0; for (int j = 0; j < 4; ++j) det += m[j] * cof(0, j);
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) inv[j * 4 + i] = cof(i, j) / det;
    for (int i = 0; i < 16; ++i) R[i] = (float) inv[i];
}

// GaussianFilter ctor, non-CUDA branch (src/rfilters/gaussian.cpp:60-96)
static void gaussian_coeffs(float stddev, float radius, float *out) {
    static const double remez[10] = {9.992604880e-1, -4.977025247e-1, 1.222248550e-1, -1.932406282e-2, 2.136713061e-3,
                                     -1.679873860e-4, 9.202145248e-6, -3.329417433e-7, 7.128382794e-9, -6.821193280e-11};
    double scale = 1.0, s2 = (double) stddev * (double) stddev;
    for (int i = 0; i < 10; ++i) { out[i] = (float) (remez[i] * scale); scale /= s2; }
    float x = radius * radius, x2 = x * x, x4 = x2 * x2, x8 = x4 * x4;   // Estrin, fp32, fused
    float p01 = fmaf(out[1], x, out[0]), p23 = fmaf(out[3], x, out[2]), p45 = fmaf(out[5], x, out[4]), p67 = fmaf(out[7], x, out[6]), p89 = fmaf(out[9], x, out[8]);
    out[0] -= fmaf(p89, x8, fmaf(fmaf(p67, x2, p45), x4, fmaf(p23, x2, p01)));
}

// envmap set-up (src/emitters/envmap.cpp:141-232 + distr_2d.h:385-503): periodic column, lum*sin(theta)
// importance map, MIP hierarchy with 2x2-blocked levels.  Returns packed warp levels.
static void build_envmap(const lr_emitter &e, std::vector<float> &pixels, std::vector<float> &warp, DEmitter &de) {
    uint32_t W = e.width, H = e.height, W1 = W + 1;
    pixels.assign((size_t) W1 * H * 3, 0.f); std::vector<float> lum((size_t) W1 * H);
    float theta_scale = 1.f / (float) (H - 1) * kPi;
    for (uint32_t y = 0; y < H; ++y) {
        float st = sinf((float) y * theta_scale);
        for (uint32_t x = 0; x < W; ++x) {
            const float *px = e.pixels + ((size_t) y * W + x) * 3;
            float l = px[0] * 0.212671f + px[1] * 0.715160f + px[2] * 0.072169f; l = std::max(l, 0.f);
            lum[(size_t) y * W1 + x] = l * st;
            memcpy(&pixels[((size_t) y * W1 + x) * 3], px, 12);
        }
        lum[(size_t) y * W1 + W] = lum[(size_t) y * W1];
        memcpy(&pixels[((size_t) y * W1 + W) * 3], &pixels[((size_t) y * W1) * 3], 12);
    }
    uint32_t npx = W1 - 1, npy = H - 1, mx = std::max(npx, npy), max_level = 0; while ((1u << max_level) < mx) ++max_level;
    struct Lv { uint32_t w, h; size_t off; };
    std::vector<Lv> lv; size_t off = 0;
    auto push = [&](uint32_t w, uint32_t h) { lv.push_back({w, h, off}); off += ((size_t) w * h + 3) & ~(size_t) 3; };
    push(W1, H);
    uint32_t lx = npx, ly = npy;
    for (int level = (int) max_level; level >= 0; --level) { lx += lx & 1u; ly += ly & 1u; push(lx, ly); lx >>= 1; ly >>= 1; }
    warp.assign(off, 0.f);
    auto index = [](uint32_t x, uint32_t y, uint32_t width) { return ((x & 1u) | (((x & ~1u) | (y & 1u)) << 1)) + ((y & ~1u) * width); };
    double sum = 0.0; const float *in = lum.data();
    for (uint32_t y = 0; y < npy; ++y) { for (uint32_t x = 0; x < npx; ++x) {
            float avg = .25f * (in[0] + in[1] + in[W1] + in[W1 + 1]); sum += (double) avg;
            warp[lv[1].off + index(x, y, lv[1].w)] = avg; ++in; } ++in; }
    float scale = (float) ((double) (npx * npy) / sum);
    for (size_t i = 0; i < (size_t) W1 * H; ++i) warp[lv[0].off + i] = lum[i] * scale;
    for (size_t i = 0; i < (size_t) lv[1].w * lv[1].h; ++i) warp[lv[1].off + i] *= scale;
    lx = npx; ly = npy;
    for (uint32_t level = 2; level <= max_level + 1; ++level) {
        lx = (lx + 1u) >> 1; ly = (ly + 1u) >> 1;
        for (uint32_t y = 0; y < ly; ++y) for (uint32_t x = 0; x < lx; ++x) {
            const float *d0 = &warp[lv[level - 1].off + index(x * 2, y * 2, lv[level - 1].w)];
            warp[lv[level].off + index(x, y, lv[level].w)] = d0[0] + d0[1] + d0[2] + d0[3];
        }
    }
    de.n_levels = (int) lv.size();
    for (size_t i = 0; i < lv.size() && i < 24; ++i) { de.level_off[i] = (int) lv[i].off; de.level_w[i] = (int) lv[i].w; }
}

// majorant super-grid: max of the raw grid over each cell's trilinear support (cell voxels +1 apron)
__global__ void k_majorant_grid(const float *data, int rx, int ry, int rz, int ch, int cell, int mx, int my, int mz, int wrap, float *maj) {
    int c = blockIdx.x * blockDim.x + threadIdx.x; if (c >= mx * my * mz) return;
    int cx = c % mx, cy = (c / mx) % my, cz = c / (mx * my);
    // a lookup at local coordinate u in [cx*cell, (cx+1)*cell) voxels touches voxel indices floor(u-.5) .. floor(u-.5)+1: the cell's
    // voxels plus a one-voxel apron, whose indices leave the grid at its faces and are brought back by the grid's wrap mode exactly as
    // the lookup does it (grid_wrap, lr_medium.cuh): clamp / mirror stay at the face, REPEAT reaches the voxels of the opposite face
    auto wr = [wrap](int i, int n) { return wrap == LR_WRAP_REPEAT ? ((i % n) + n) % n : min(max(i, 0), n - 1); };
    const int x0 = cx * cell - 1, x1 = min((cx + 1) * cell, rx), y0 = cy * cell - 1, y1 = min((cy + 1) * cell, ry), z0 = cz * cell - 1, z1 = min((cz + 1) * cell, rz);
    float m = 0.f;
    for (int z = z0; z <= z1; ++z) for (int y = y0; y <= y1; ++y) for (int x = x0; x <= x1; ++x) {
        const size_t v = ((size_t) wr(z, rz) * ry + wr(y, ry)) * rx + wr(x, rx);
        for (int k = 0; k < ch; ++k) m = fmaxf(m, data[v * ch + k]);
    }
    maj[c] = m;
}

// fat-voxel copy (DGrid::fat): one thread per record
__global__ void k_grid_fatten(const float *data, int rx, int ry, int rz, float4 *fat) {
    const size_t n = (size_t) (rx + 1) * (ry + 1) * (rz + 1);
    for (size_t r = (size_t) blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (size_t) gridDim.x * blockDim.x) {
        const int rix = (int) (r % (rx + 1)), riy = (int) ((r / (rx + 1)) % (ry + 1)), riz = (int) (r / ((size_t) (rx + 1) * (ry + 1)));
        const int x0 = max(rix - 1, 0), x1 = min(rix, rx - 1), y0 = max(riy - 1, 0), y1 = min(riy, ry - 1), z0 = max(riz - 1, 0), z1 = min(riz, rz - 1);
        auto at = [&](int x, int y, int z) { return data[((size_t) z * ry + y) * rx + x]; };
        fat[2 * r] = make_float4(at(x0, y0, z0), at(x1, y0, z0), at(x0, y1, z0), at(x1, y1, z0));
        fat[2 * r + 1] = make_float4(at(x0, y0, z1), at(x1, y0, z1), at(x0, y1, z1), at(x1, y1, z1));
    }
}

static int refresh_grid_derived(lr_ctx *c, uint32_t gi) {
    DGrid &g = c->h_grids[gi]; size_t n = c->grid_floats[gi];
    // max over all channels
    LR_CUDA(c, cudaMemsetAsync(c->scratch.p, 0, 4, c->stream));
    k_grid_max<<<std::min<size_t>((n + 255) / 256, 1184), 256, 0, c->stream>>>(g.data, n, c->scratch.as<float>());
    float mx = 0.f; LR_CUDA(c, cudaMemcpyAsync(&mx, c->scratch.p, 4, cudaMemcpyDeviceToHost, c->stream));
    LR_CUDA(c, cudaStreamSynchronize(c->stream));
    g.maxv = (gi < c->grid_fixed_max.size() && c->grid_fixed_max[gi] >= 0.f) ? c->grid_fixed_max[gi] : mx;
    int nc = g.mx * g.my * g.mz;
    k_majorant_grid<<<(nc + 127) / 128, 128, 0, c->stream>>>(g.data, g.rx, g.ry, g.rz, g.ch, g.cell, g.mx, g.my, g.mz, (int) g.wrap, (float *) g.maj);
    if (g.fat) k_grid_fatten<<<c->n_sm * 8, 256, 0, c->stream>>>(g.data, g.rx, g.ry, g.rz, (float4 *) g.fat);
    LR_CUDA(c, cudaGetLastError());
    for (auto &m : c->h_media) if (m.m.kind == LR_MEDIUM_HETEROGENEOUS && m.m.sigma_t_grid == (int) gi) m.max_density = m.m.scale * g.maxv;
    LR_CUDA(c, cudaMemcpyAsync(c->d_grids->p, c->h_grids.data(), c->h_grids.size() * sizeof(DGrid), cudaMemcpyHostToDevice, c->stream));
    LR_CUDA(c, cudaMemcpyAsync(c->d_media->p, c->h_media.data(), std::max<size_t>(c->h_media.size(), 1) * sizeof(DMedium), cudaMemcpyHostToDevice, c->stream));
    return LR_OK;
}

static int finish_bvh4(lr_ctx *c, const std::vector<DNode> &bin, const std::vector<DTri> &sorted);

static int build_lbvh(lr_ctx *c, const std::vector<DTri> &tris, V3 lo, V3 hi) {
    uint32_t n = (uint32_t) tris.size(); DScene &sc = c->sc;
    c->tris_gid = c->newbuf(); if (int r = upload(c, c->tris_gid, tris.data(), n)) return r;
    DevBuf *sorted = c->newbuf(), *nodes = c->newbuf();
    LR_CUDA(c, sorted->alloc(std::max<uint32_t>(n, 1) * sizeof(DTri))); LR_CUDA(c, nodes->alloc(std::max<uint32_t>(n, 2) * sizeof(DNode)));
    sc.tris = sorted->as<DTri>(); sc.nodes = nodes->as<DNode>(); sc.n_tris = n;
    sc.nodes4 = nullptr; sc.bvh4_stack = 1;
    if (n == 0) return LR_OK;
    DevBuf keys, keys2, vals, vals2, children, par_i, par_l, blo, bhi, flags, tmp;
    LR_CUDA(c, keys.alloc(n * 8)); LR_CUDA(c, keys2.alloc(n * 8)); LR_CUDA(c, vals.alloc(n * 4)); LR_CUDA(c, vals2.alloc(n * 4));
    V3 ext = hi - lo; V3 inv_ext = mk3(ext.x > 0 ? 1.f / ext.x : 0.f, ext.y > 0 ? 1.f / ext.y : 0.f, ext.z > 0 ? 1.f / ext.z : 0.f);
    k_morton<<<(n + 255) / 256, 256, 0, c->stream>>>(c->tris_gid->as<DTri>(), n, lo, inv_ext, keys.as<uint64_t>(), vals.as<uint32_t>());
    size_t tb = 0; cub::DeviceRadixSort::SortPairs(nullptr, tb, keys.as<uint64_t>(), keys2.as<uint64_t>(), vals.as<uint32_t>(), vals2.as<uint32_t>(), (int) n, 0, 64, c->stream);
    LR_CUDA(c, tmp.alloc(tb));
    LR_CUDA(c, cub::DeviceRadixSort::SortPairs(tmp.p, tb, keys.as<uint64_t>(), keys2.as<uint64_t>(), vals.as<uint32_t>(), vals2.as<uint32_t>(), (int) n, 0, 64, c->stream));
    k_gather_tris<<<(n + 255) / 256, 256, 0, c->stream>>>(c->tris_gid->as<DTri>(), vals2.as<uint32_t>(), n, sorted->as<DTri>());
    if (n >= 2) {
        LR_CUDA(c, children.alloc((n - 1) * sizeof(int2))); LR_CUDA(c, par_i.alloc(n * 4)); LR_CUDA(c, par_l.alloc(n * 4));
        LR_CUDA(c, blo.alloc((2 * n - 1) * 16)); LR_CUDA(c, bhi.alloc((2 * n - 1) * 16)); LR_CUDA(c, flags.alloc(n * 4));
        LR_CUDA(c, cudaMemsetAsync(flags.p, 0, n * 4, c->stream));
        k_hierarchy<<<(n + 255) / 256, 256, 0, c->stream>>>(keys2.as<uint64_t>(), (int) n, children.as<int2>(), par_i.as<int>(), par_l.as<int>());
        float pad = 1e-6f * std::max(ext.x, std::max(ext.y, ext.z)) + 1e-30f;
        k_refit<<<(n + 255) / 256, 256, 0, c->stream>>>(sorted->as<DTri>(), (int) n, children.as<int2>(), par_i.as<int>(), par_l.as<int>(),
                                                        blo.as<float4>(), bhi.as<float4>(), flags.as<int>(), pad);
        k_pack_nodes<<<(n + 255) / 256, 256, 0, c->stream>>>((int) n, children.as<int2>(), blo.as<float4>(), bhi.as<float4>(), nodes->as<DNode>());
    }
    LR_CUDA(c, cudaGetLastError());
    LR_CUDA(c, cudaStreamSynchronize(c->stream));
    // the 4-wide tree of the TRACE kernels is collapsed on the host from the device-built two-wide tree
    std::vector<DNode> bin(n >= 2 ? n - 1 : 0); std::vector<DTri> srt(n);
    if (n >= 2) LR_CUDA(c, cudaMemcpy(bin.data(), nodes->p, (size_t) (n - 1) * sizeof(DNode), cudaMemcpyDeviceToHost));
    LR_CUDA(c, cudaMemcpy(srt.data(), sorted->p, (size_t) n * sizeof(DTri), cudaMemcpyDeviceToHost));
    return finish_bvh4(c, bin, srt);
}

#include "lr_bvh_host.inl"

// Builder choice: SAH on the host up to 4 Mi triangles (scene-load cost of milliseconds to seconds, cheapest traversal),
// Morton LBVH on the device above that.  LR_BVH=lbvh|sah overrides.
static int build_bvh(lr_ctx *c, const std::vector<DTri> &tris, V3 lo, V3 hi) {
    const char *e = getenv("LR_BVH");
    bool lbvh = e ? strcmp(e, "lbvh") == 0 : tris.size() > (size_t(1) << 22);
    if (tris.size() < 2 || lbvh) return build_lbvh(c, tris, lo, hi);
    return build_sah_bvh(c, tris, lo, hi);
}

#include "lr_backend_api.inl"
#include "lr_kat.inl"
#include "lr_noalloc.inl"
