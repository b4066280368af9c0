// lr_noalloc.inl -- lr_ray_intersect_noalloc (included by lr_backend.cu, so the array below, the kernel that uses it and the
// host code that takes its address are one translation unit).
//
// NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).  It answers the same query as lr_ray_intersect
// (Scene::ray_intersect, src/render/scene.cpp:204-226) and allocates nothing: no cudaMalloc / cudaFree / DevBuf / torch.
//
// What lr_ray_intersect allocates per call is scratch -- the device staging of the caller's host buffers -- and the wide-tree
// walk below needs more of it: a per-ray traversal stack in global memory.  All of that lives in ONE file-scope device array
// of fixed size, held by every GPU that loads the library until the process exits:
//     [ rays  : LR_NOALLOC_MAX_RAYS x 32 B ][ hits : LR_NOALLOC_MAX_RAYS x 16 B ][ stack : LR_NOALLOC_MAX_STACK x LR_NOALLOC_MAX_RAYS x 4 B ]
//     = 8192 x (32 + 16 + 96 x 4) B = 3.375 MiB.
// In host code `g_noalloc_scratch` is only a stand-in for the symbol, never a device address: its address is looked up with
// cudaGetSymbolAddress once per context (a context is bound to one device, which is made current first) and that pointer is
// what the kernel receives.  Every call shares the array and finds the previous call's contents in it; nothing in it has to
// start at zero (the staged rays and hits of the first n entries are overwritten before they are read, a stack entry is
// written by its push before any pop reads it), so nothing is cleared.  Calls that overlap on the GPU (two streams, two
// contexts on one device, two host threads) RACE on the array: serialise them.
#pragma once

#define LR_NOALLOC_MAX_RAYS  8192u          // rays per call
#define LR_NOALLOC_MAX_STACK 96u            // traversal-stack entries per ray (worst case of the scene's 4-wide tree, DScene::bvh4_stack)
#define LR_NOALLOC_RAYS_OFF  0u
#define LR_NOALLOC_HITS_OFF  (LR_NOALLOC_MAX_RAYS * 32u)
#define LR_NOALLOC_STACK_OFF (LR_NOALLOC_MAX_RAYS * 48u)
#define LR_NOALLOC_BYTES     (LR_NOALLOC_MAX_RAYS * (48u + LR_NOALLOC_MAX_STACK * 4u))

static __device__ __align__(256) unsigned char g_noalloc_scratch[LR_NOALLOC_BYTES];

namespace lr {

// One ray per thread through the 4-wide quantized tree of the wavefront TRACE stage (trace_visit / trace_leaf, lr_trace.cuh)
// with every stack entry in the scratch column of the thread: stack[depth * stride + i].  Same Moeller-Trumbore test and
// tie rule as every other traversal here, so (t, prim, u, v) equal k_ray_probe's bit for bit.
__global__ void k_ray_probe_scratch(const __grid_constant__ DScene sc, const float *rays, uint32_t n, float *hits, int *stack, uint32_t stride) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; if (i >= n) return;
    const float4 r0 = reinterpret_cast<const float4 *>(rays)[2 * i], r1 = reinterpret_cast<const float4 *>(rays)[2 * i + 1];
    TraceStack<int> st; st.stack = nullptr; st.stride = 0; st.sd = 0; st.ovf = stack + i; st.ovf_stride = (int) stride;
    TraceRay r; trace_begin(sc, r, mk3(r0.x, r0.y, r0.z), mk3(r0.w, r1.x, r1.y), r1.z);
    while (r.node != TR_DONE) {
        if (r.node < 0) trace_leaf<false, false, int>(sc, r, st);
        else trace_visit<int>(sc, r, nullptr, 0u, st);
    }
    reinterpret_cast<float4 *>(hits)[i] = make_float4(r.h.t, __uint_as_float(r.h.prim), r.h.u, r.h.v);
}

} // namespace lr

// address of the scratch array on the context's device (looked up once per context)
static int noalloc_scratch(lr_ctx *c, unsigned char **out) {
    LR_CUDA(c, cudaSetDevice(c->device));
    if (!c->noalloc_ptr) {
        void *p = nullptr;
        LR_CUDA(c, cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) LR_FAIL(c, LR_ERR_CUDA, "lr_ray_intersect_noalloc: cudaGetSymbolAddress returned NULL");
        c->noalloc_ptr = p;
    }
    *out = static_cast<unsigned char *>(c->noalloc_ptr); return LR_OK;
}

extern "C" {

uint64_t lr_noalloc_max_rays(void) { return LR_NOALLOC_MAX_RAYS; }

int lr_ray_intersect_noalloc(lr_ctx *c, const float *rays, uint64_t n, float *hits, int is_device) {
    if (!c || !c->has_scene) return LR_ERR_NO_SCENE;
    if (!rays || !hits) LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_ray_intersect_noalloc: NULL buffers");
    if (n > LR_NOALLOC_MAX_RAYS)
        LR_FAIL(c, LR_ERR_INVALID_ARG, "lr_ray_intersect_noalloc: %llu rays exceed the static scratch (at most %u per call; split the batch or use lr_ray_intersect)",
                (unsigned long long) n, LR_NOALLOC_MAX_RAYS);
    if (c->sc.bvh4_stack > LR_NOALLOC_MAX_STACK)
        LR_FAIL(c, LR_ERR_UNSUPPORTED, "lr_ray_intersect_noalloc: the scene's tree needs a %u-entry traversal stack, the static scratch holds %u per ray (use lr_ray_intersect)",
                c->sc.bvh4_stack, LR_NOALLOC_MAX_STACK);
    if (n == 0) return LR_OK;
    unsigned char *s = nullptr; if (int r = noalloc_scratch(c, &s)) return r;
    const float *d_rays = rays; float *d_hits = hits;
    if (!is_device) {                                           // host buffers: staged through the scratch array
        d_rays = reinterpret_cast<const float *>(s + LR_NOALLOC_RAYS_OFF); d_hits = reinterpret_cast<float *>(s + LR_NOALLOC_HITS_OFF);
        LR_CUDA(c, cudaMemcpyAsync(s + LR_NOALLOC_RAYS_OFF, rays, n * 32, cudaMemcpyHostToDevice, c->stream));
    }
    lr::k_ray_probe_scratch<<<(unsigned) ((n + 127) / 128), 128, 0, c->stream>>>(c->sc, d_rays, (uint32_t) n, d_hits,
                                                                                 reinterpret_cast<int *>(s + LR_NOALLOC_STACK_OFF), LR_NOALLOC_MAX_RAYS);
    LR_CUDA(c, cudaGetLastError());
    if (is_device) return LR_OK;                                // stream-ordered: the caller synchronises (or captures)
    return copy_out(c, hits, d_hits, n * 16, 0);
}

} // extern "C"
