/*
 * lr_api.h -- C ABI of the B200-native volumetric path-tracing backend.
 *
 * This is the drop-in boundary for ONE hot path of the LiverRenderer fork of
 * Mitsuba 3: everything below `SamplingIntegrator::render`
 * (reference: src/render/integrator.cpp:151-397) for the `path` / `volpath`
 * integrators and `RBIntegrator.render_backward`
 * (reference: src/python/python/ad/integrators/common.py:625-783) for
 * `prbvolpath`.  The host (Mitsuba's plugin / scene API, or the Python mirror
 * in liverrenderer_b200/) flattens its scene graph into the POD `lr_scene`
 * below and calls these entry points; no C++ types, no exceptions and no
 * torch types cross the boundary.
 *
 * Conventions
 *   - matrices are row-major float[16], points are transformed as M*(x,y,z,1)
 *   - colours are linear RGB float[3] (Mitsuba `*_rgb` variants)
 *   - every function returns LR_OK (0) or a negative lr_status; the message is
 *     available from lr_last_error()
 *   - host pointers passed in are only read during the call; the context owns
 *     all device memory
 */
#ifndef LR_API_H
#define LR_API_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LR_API_VERSION 4

typedef enum lr_status {
    LR_OK                = 0,
    LR_ERR_INVALID_ARG   = -1,
    LR_ERR_NO_DEVICE     = -2,   /* no CUDA device / driver: there is NO CPU fallback */
    LR_ERR_CUDA          = -3,
    LR_ERR_UNSUPPORTED   = -4,   /* plugin / option outside the hot path */
    LR_ERR_NO_SCENE      = -5,
    LR_ERR_CANCELLED     = -6,
    LR_ERR_OOM           = -7
} lr_status;

/* ---- plugin kinds (names follow the reference's plugin names) ---------- */
typedef enum lr_integrator_kind {
    LR_INTEGRATOR_PATH       = 0,  /* src/integrators/path.cpp:95-349            */
    LR_INTEGRATOR_VOLPATH    = 1,  /* src/integrators/volpath.cpp:93-554         */
    LR_INTEGRATOR_PRBVOLPATH = 2,  /* src/python/python/ad/integrators/prbvolpath.py:96-444 */
    LR_INTEGRATOR_VOLPATHMIS = 4,  /* src/integrators/volpathmis.cpp:97-700 (lr_integrator.spectral_mis = use_spectral_mis) */
    LR_INTEGRATOR_BIOVOLPATH = 3   /* the fork's src/integrators/biovolpath.cpp:95-551: volpath with the tissue media below
                                      (accumulated tissue depth, one-hot channel transmittance); needs every medium of the
                                      scene to be one of the tissue kinds, as the base Medium's five-argument
                                      sample_interaction throws (src/render/medium.cpp:84-90) */
} lr_integrator_kind;

typedef enum lr_tracking_mode {
    /* bit-compatible draw order with the reference: one global majorant,
       plain delta tracking (src/render/medium.cpp:42-82)                      */
    LR_TRACKING_REFERENCE = 0,
    /* production mode: majorant super-grid + DDA, opaque-occluder pre-test for
       NEE.  Same expectation, different RNG consumption (statistical parity). */
    LR_TRACKING_DDA       = 1
} lr_tracking_mode;

typedef enum lr_bsdf_kind {
    LR_BSDF_DIFFUSE         = 0,  /* src/bsdfs/diffuse.cpp:100-178             */
    LR_BSDF_ROUGHDIELECTRIC = 1,  /* src/bsdfs/roughdielectric.cpp:242-499     */
    LR_BSDF_PRINCIPLED      = 2,  /* src/bsdfs/principled.cpp:333-822          */
    LR_BSDF_NULL            = 3,  /* src/bsdfs/null.cpp (index-matched boundary) */
    LR_BSDF_DIELECTRIC      = 4   /* src/bsdfs/dielectric.cpp (smooth)         */
} lr_bsdf_kind;
/* `bumpmap` (src/bsdfs/bumpmap.cpp:96-281) is not a kind: it wraps any of the above through
   lr_bsdf.bump_tex (the nested BSDF is the block itself).                                    */

typedef enum lr_microfacet_kind { LR_MF_BECKMANN = 0, LR_MF_GGX = 1 } lr_microfacet_kind;

typedef enum lr_phase_kind {
    LR_PHASE_ISOTROPIC = 0,       /* src/phase/isotropic.cpp:39-58             */
    LR_PHASE_HG        = 1        /* src/phase/hg.cpp:64-99                    */
} lr_phase_kind;

typedef enum lr_sampler_kind {
    LR_SAMPLER_INDEPENDENT = 0,   /* src/samplers/independent.cpp (PCG32, src/render/sampler.cpp:129-148) */
    LR_SAMPLER_LD          = 1    /* src/samplers/ldsampler.cpp:107-141: (0,2)-sequence, per-pixel scramble + per-dimension
                                     permutation; spp must be a square power of two (set_sample_count, :80-89) */
} lr_sampler_kind;

typedef enum lr_medium_kind {
    LR_MEDIUM_HOMOGENEOUS   = 0,  /* src/media/homogeneous.cpp:137-181         */
    LR_MEDIUM_HETEROGENEOUS = 1,  /* src/media/heterogeneous.cpp:155-201       */
    /* the fork's tissue media (biovolpath only): per-element exponential free paths drawn from an inner PCG32
       seeded with the bit pattern of the integrator's sample */
    LR_MEDIUM_LIVER         = 2,  /* src/media/liver.cpp:133-574: capsule layers by tissue depth, parenchyma below   */
    LR_MEDIUM_GLISSON       = 3,  /* src/media/glissonCapsule.cpp:134-414: capsule layers only                      */
    LR_MEDIUM_PARENCHYMA    = 4   /* src/media/parenchyma.cpp:134-393: parenchyma elements only, hard-coded sigma_t  */
} lr_medium_kind;

typedef enum lr_emitter_kind {
    LR_EMITTER_AREA     = 0,      /* src/emitters/area.cpp:83-197              */
    LR_EMITTER_CONSTANT = 1,      /* src/emitters/constant.cpp:92-163          */
    LR_EMITTER_ENVMAP   = 2       /* src/emitters/envmap.cpp:353-476           */
} lr_emitter_kind;

typedef enum lr_rfilter_kind {
    LR_RFILTER_BOX      = 0,      /* src/rfilters/box.cpp:33-41                */
    LR_RFILTER_GAUSSIAN = 1,      /* src/rfilters/gaussian.cpp:48-105 (LLVM polynomial form) */
    LR_RFILTER_TENT     = 2       /* src/rfilters/tent.cpp:45-53               */
} lr_rfilter_kind;

typedef enum lr_grid_filter { LR_FILTER_NEAREST = 0, LR_FILTER_TRILINEAR = 1 } lr_grid_filter;
typedef enum lr_grid_wrap   { LR_WRAP_REPEAT = 0, LR_WRAP_MIRROR = 1, LR_WRAP_CLAMP = 2 } lr_grid_wrap;

/* ---- POD scene description -------------------------------------------- */

/* Triangle mesh (reference: src/render/mesh.cpp, include/mitsuba/render/mesh.h).
   Vertices are already in world space.  `normals` / `texcoords` may be NULL.   */
typedef struct lr_mesh {
    const float    *positions;      /* [n_vertices*3]                           */
    const float    *normals;        /* [n_vertices*3] or NULL (=> face normals) */
    const float    *texcoords;      /* [n_vertices*2] or NULL                   */
    const uint32_t *indices;        /* [n_faces*3]                              */
    uint32_t n_vertices;
    uint32_t n_faces;
    int32_t  bsdf;                  /* index into lr_scene.bsdfs                */
    int32_t  interior_medium;       /* index into lr_scene.media or -1          */
    int32_t  exterior_medium;       /* index into lr_scene.media or -1          */
    int32_t  emitter;               /* index into lr_scene.emitters or -1       */
    uint32_t flip_normals;
    uint32_t reserved;
} lr_mesh;

/* `bitmap` texture (src/textures/bitmap.cpp:395-930) evaluated the way the LLVM variants do:
   dr::Texture<Float,2>::eval_nonaccel / eval_fetch_nonaccel
   (ext/drjit/include/drjit/texture.h:484-556,704-735), wrap (:1369-1396).
   `data` is what the reference holds after load: linear values (the host converts 8-bit sRGB
   files unless `raw`), row-major [height][width][channels].                                  */
typedef enum lr_tex_filter { LR_TEX_BILINEAR = 0, LR_TEX_NEAREST = 1 } lr_tex_filter;
typedef struct lr_texture {
    const float *data;              /* [height*width*channels]                  */
    uint32_t width, height;
    uint32_t channels;              /* 1 or 3                                   */
    uint32_t filter;                /* lr_tex_filter (`filter_type`)            */
    uint32_t wrap;                  /* lr_grid_wrap (`wrap_mode`, default repeat) */
    uint32_t reserved;
    float    to_uv[6];              /* 2x3 affine, row-major: uv' = to_uv * (u, v, 1) */
} lr_texture;

/* BSDF parameter block.  Texture slots hold 0 for "none" or 1 + index into lr_scene.textures. */
typedef struct lr_bsdf {
    uint32_t kind;                  /* lr_bsdf_kind                             */
    uint32_t mf_kind;               /* lr_microfacet_kind (roughdielectric)     */
    uint32_t sample_visible;        /* roughdielectric, default 1               */
    uint32_t flags;                 /* principled: bit i set => optional lobe i present (see LR_PR_*) */
    float reflectance[3];           /* diffuse reflectance / principled base_color / spec. reflectance */
    float transmittance[3];         /* roughdielectric specular_transmittance   */
    float eta;                      /* int_ior / ext_ior                        */
    float alpha_u, alpha_v;         /* roughdielectric                          */
    /* principled (src/bsdfs/principled.cpp:190-238)                            */
    float roughness, anisotropic, metallic, spec_trans, specular, spec_tint;
    float sheen, sheen_tint, flatness, clearcoat, clearcoat_gloss;
    float diff_refl_srate, main_spec_srate, clearcoat_srate;
    /* `reflectance` texture: 0 = constant, 1 = checkerboard (src/textures/checkerboard.cpp:70-89):
       color0 = reflectance, color1 = tex_color1, uv' = tex_to_uv (2x3 affine, row-major) * (u, v, 1);
       2 = bitmap: Texture::eval of slot `refl_tex` (bitmap.cpp:443-485, 1-channel => grey)     */
    uint32_t tex_kind;
    float tex_color1[3];
    float tex_to_uv[6];
    uint32_t refl_tex;              /* texture slot for tex_kind == 2           */
    /* `bumpmap` wrapper (src/bsdfs/bumpmap.cpp): slot != 0 => sample/eval/pdf run in the frame
       perturbed by `bump_scale * eval_1_grad(texture)` (:248-276)                            */
    uint32_t bump_tex;              /* texture slot, 0 = no bump map            */
    float    bump_scale;            /* `scale`, default 1                       */
    uint32_t bump_flags;            /* LR_BUMP_*                                */
} lr_bsdf;

#define LR_BUMP_FLIP_INVALID_NORMALS (1u << 0)   /* `flip_invalid_normals`, default on  (:265-269) */
#define LR_BUMP_SHADOWING_FUNCTION   (1u << 1)   /* `use_shadowing_function`, default on (normalmap_helpers.h:20-25) */

#define LR_PR_HAS_ANISOTROPIC (1u << 0)
#define LR_PR_HAS_METALLIC    (1u << 1)
#define LR_PR_HAS_SPEC_TRANS  (1u << 2)
#define LR_PR_HAS_SPEC_TINT   (1u << 3)
#define LR_PR_HAS_SHEEN       (1u << 4)
#define LR_PR_HAS_SHEEN_TINT  (1u << 5)
#define LR_PR_HAS_FLATNESS    (1u << 6)
#define LR_PR_HAS_CLEARCOAT   (1u << 7)
#define LR_PR_ETA_SPECULAR    (1u << 8)   /* `eta` given instead of `specular` */

/* `gridvolume` (reference: src/volumes/grid.cpp:164-366, 553-583).            */
typedef struct lr_grid {
    const float *data;              /* [res_z][res_y][res_x][channels], fp32    */
    uint32_t res[3];                /* x, y, z                                  */
    uint32_t channels;              /* 1 or 3                                   */
    float to_local[16];             /* world -> unit cube                       */
    float bbox_min[3], bbox_max[3]; /* world AABB of the unit cube (Volume::update_bbox, src/render/volume.cpp:13-18) */
    uint32_t filter;                /* lr_grid_filter                           */
    uint32_t wrap;                  /* lr_grid_wrap                             */
    float max_value;                /* <0 => computed by the backend (max over all channels) */
    uint32_t differentiable;        /* allocate a gradient buffer (prbvolpath)  */
} lr_grid;

typedef struct lr_medium {
    uint32_t kind;                  /* lr_medium_kind                           */
    uint32_t phase;                 /* lr_phase_kind                            */
    float    phase_g;               /* hg                                       */
    float    scale;
    int32_t  sigma_t_grid;          /* index into lr_scene.grids or -1 => sigma_t_const */
    int32_t  albedo_grid;           /* index into lr_scene.grids or -1 => albedo_const  */
    float    sigma_t_const[3];
    float    albedo_const[3];
    uint32_t sample_emitters;       /* default 1 (src/render/medium.cpp:31)     */
    uint32_t has_spectral_extinction; /* default 1 (heterogeneous.cpp:161)      */
    /* tissue media only (ignored for the other kinds).  Values AS THE PLUGIN STORES THEM: the constructors read the
       "_B" property into the G slot and "_G" into B for every collagen layer and for elastin layers 1-2
       (liver.cpp:149-150,...,174-175; glissonCapsule.cpp the same) -- the host loader reproduces that swap.
       sigma_t_const / albedo_const / scale above keep their meaning (constant volumes only; parenchyma ignores
       them, parenchyma.cpp:164,182-183). */
    float    layer_limit[4];        /* layer1Limit..layer4Limit (liver.cpp:144-147)                      */
    float    sigma_collagen[4][3];  /* m_sigma_collagen_layer1..4                                        */
    float    sigma_elastin[4][3];   /* m_sigma_elastin_layer1..4                                         */
    float    sigma_blood[3], sigma_bile[3], sigma_lipid_water[3];   /* constant volumes (liver.cpp:189-193) */
    float    sigma_hepatocity;
} lr_medium;

typedef struct lr_emitter {
    uint32_t kind;                  /* lr_emitter_kind                          */
    int32_t  mesh;                  /* area: index of the mesh it is attached to */
    float    radiance[3];           /* area / constant; envmap: unused          */
    float    scale;                 /* envmap                                   */
    /* area light on a `rectangle`: p = origin + (2u-1)*edge_u + (2v-1)*edge_v
       (src/shapes/rectangle.cpp:181-199); is_rectangle=0 => generic mesh sampling */
    uint32_t is_rectangle;
    float    rect_origin[3], rect_edge_u[3], rect_edge_v[3], rect_normal[3];
    float    inv_area;
    /* envmap: pixels are linear RGB, row-major [height][width][3], WITHOUT the
       duplicated periodic column (the backend appends it, envmap.cpp:141-226) */
    const float *pixels;
    uint32_t width, height;
    float    to_world[16];          /* envmap rotation                          */
    float    reserved[2];
} lr_emitter;

typedef struct lr_sensor {
    float    to_world[16];          /* camera -> world (no scale)               */
    float    sample_to_camera[16];  /* inverse of perspective_projection (include/mitsuba/render/sensor.h:234-269) */
    float    near_clip, far_clip;
    uint32_t film_width, film_height;       /* size of the rendered window = the film's CROP size (Film::crop_size, integrator.cpp:163);
                                               sample_to_camera is built for that window (perspective_projection, sensor.h:234-269) */
    uint32_t rfilter;               /* lr_rfilter_kind                          */
    float    rfilter_radius;        /* box .5, tent 1, gaussian 4*stddev        */
    float    rfilter_stddev;        /* gaussian                                 */
    uint32_t has_alpha;             /* pixel_format rgba => 1                   */
    int32_t  medium;                /* medium the sensor sits in, or -1         */
    uint32_t crop_offset_x, crop_offset_y;  /* Film::crop_offset (hdrfilm crop_offset_x / _y): added to the pixel position before the
                                               camera sample position is formed (integrator.cpp:338, :465-468)                       */
} lr_sensor;

typedef struct lr_integrator {
    uint32_t kind;                  /* lr_integrator_kind                       */
    int32_t  max_depth;             /* -1 => unbounded (2^32-1)                 */
    int32_t  rr_depth;
    uint32_t hide_emitters;
    uint32_t tracking;              /* lr_tracking_mode                         */
    uint32_t base_seed;             /* independent sampler `seed`               */
    uint32_t majorant_cell;         /* DDA super-voxel edge in voxels (0 => default 8) */
    uint32_t sampler;               /* lr_sampler_kind: the sensor's sampler plugin */
    uint32_t spectral_mis;          /* volpathmis: use_spectral_mis (default 1, volpathmis.cpp:79) */
} lr_integrator;

typedef struct lr_scene {
    uint32_t api_version;           /* LR_API_VERSION                           */
    uint32_t n_meshes, n_bsdfs, n_media, n_grids, n_emitters;
    uint32_t n_textures;
    const lr_texture *textures;
    const lr_mesh    *meshes;
    const lr_bsdf    *bsdfs;
    const lr_medium  *media;
    const lr_grid    *grids;
    const lr_emitter *emitters;
    lr_sensor     sensor;
    lr_integrator integrator;
} lr_scene;

typedef struct lr_stats {
    uint64_t samples;               /* camera paths started                     */
    uint64_t rays;                  /* closest-hit + any-hit queries            */
    uint64_t collisions;            /* tentative free-flight events (= sigma_t lookups) */
    uint64_t real_scatters;
    uint64_t launches;              /* kernels launched by the last call        */
    double   device_ms;             /* CUDA-event time of the last render call  */
    double   tracking_ms;           /* time of the dominant (tracking) kernel launches in the last call */
    uint64_t tracking_launches;
} lr_stats;

/* Per-stage device time of the wavefront `volpath` (lr_profile_enable): CUDA events around every stage launch of a render
   that keeps ONE batch in flight, so the intervals do not overlap (reference counterpart: the ProfilerPhase scopes of
   include/mitsuba/core/profiler.h:20-49).  Stage names: lr_stage_name(i), i < n_stages.                                   */
#define LR_MAX_STAGES 16
typedef struct lr_stage_profile {
    uint32_t n_stages, reserved;
    double   ms[LR_MAX_STAGES];
    uint64_t launches[LR_MAX_STAGES];
    uint64_t items[LR_MAX_STAGES];       /* queue entries the stage consumed (rays for the trace stages, lanes for gen); counted in every mode */
} lr_stage_profile;

typedef struct lr_ctx lr_ctx;

/* ---- lifecycle ---------------------------------------------------------- */
int  lr_api_version(void);
/* Create a context on CUDA device `device`.  Fails with LR_ERR_NO_DEVICE when
   there is no usable GPU: this backend has no CPU path.                       */
int  lr_create(int device, lr_ctx **out);
void lr_destroy(lr_ctx *ctx);
const char *lr_last_error(const lr_ctx *ctx);   /* ctx may be NULL: last global error */

/* Flatten-and-upload: replaces Scene construction + accel build
   (reference: src/render/scene.cpp:23-115).  Builds the LBVH, the envmap
   hierarchical warp, the majorant grids.                                       */
int  lr_scene_upload(lr_ctx *ctx, const lr_scene *scene);

/* ---- forward: Integrator::render (include/mitsuba/render/integrator.h:25-30) */
/* Renders `spp` samples per pixel and develops the film into
   out[H][W][3 or 4].  `out_is_device` != 0 means `out` is a CUDA device pointer. */
int  lr_render(lr_ctx *ctx, uint32_t seed, uint32_t spp, float *out, int out_is_device);

/* Multi-GPU / partial renders: accumulate global lanes [lane_begin, lane_end)
   of the W*H*spp wavefront (streams keyed by the GLOBAL lane index,
   integrator.cpp:321-338) into the context's film accumulator.                 */
int  lr_film_clear(lr_ctx *ctx);
int  lr_render_lanes(lr_ctx *ctx, uint32_t seed, uint32_t spp,
                     uint64_t lane_begin, uint64_t lane_end);
/* raw accumulator [H][W][C] with C = 4 (RGBW) or 5 (RGBAW) */
/* Split of ONE frame over devices by sample index (the reference has no multi-GPU code; its render loop's lane <-> (pixel,
   sample) map is src/render/integrator.cpp:321-338): after lr_set_lane_map(ctx, period, offset, len) the lane arguments of
   lr_render_lanes / lr_sample_lanes / lr_render_backward count LOCAL indices i of the context's share,
   lane = (i / len) * period + offset + (i mod len); begin = end = 0 renders the whole share.  period = spp per pass, len =
   spp / N, offset = rank * len gives each of N devices spp / N samples of every pixel (perfectly balanced); PCG32 streams
   stay keyed by the global lane, so the sum of the devices' films is the single-device film.  len = 0 removes the map.   */
int  lr_set_lane_map(lr_ctx *ctx, uint64_t period, uint64_t offset, uint64_t len);
/* Render the uploaded scene through another of its sensors: the `sensor_index` argument of SamplingIntegrator::render
   (src/render/integrator.cpp:33-44, include/mitsuba/render/integrator.h:89).  Replaces camera, film window, reconstruction filter and the
   sensor's medium; everything else of the upload stays.  The film accumulator is re-created and zeroed.                                  */
int  lr_sensor_set(lr_ctx *ctx, const lr_sensor *sensor);
int  lr_film_channels(const lr_ctx *ctx);
int  lr_film_read(lr_ctx *ctx, float *block, int is_device);
int  lr_film_write(lr_ctx *ctx, const float *block, int is_device);
int  lr_film_develop(lr_ctx *ctx, float *out, int out_is_device);
void *lr_film_device_ptr(lr_ctx *ctx);          /* for NCCL reductions done by the host */

/* Per-lane radiance (no film): out[(lane-lane_begin)*4 + {0,1,2}] = RGB,
   [3] = valid.  Used by the per-path parity tests.                             */
int  lr_sample_lanes(lr_ctx *ctx, uint32_t seed, uint32_t spp,
                     uint64_t lane_begin, uint64_t lane_end, float *out, int out_is_device);

/* ---- adjoint: render_backward (include/mitsuba/render/integrator.h:73-78) */
/* grad_in[H][W][3 or 4] = dLoss/dImage.  Accumulates into the gradient buffers
   of every grid uploaded with differentiable=1.  Lane range as above
   (0,0 => all lanes).                                                          */
int  lr_render_backward(lr_ctx *ctx, uint32_t seed, uint32_t spp,
                        const float *grad_in, int grad_is_device,
                        uint64_t lane_begin, uint64_t lane_end);

/* scene-parameter path used by mi.traverse()/params.update()
   (src/volumes/grid.cpp:309-327, src/media/heterogeneous.cpp:173-175):
   recomputes the majorant and the majorant super-grid.                         */
int  lr_grid_update(lr_ctx *ctx, uint32_t grid, const float *data, int is_device);
int  lr_grid_grad_zero(lr_ctx *ctx, uint32_t grid);
int  lr_grid_grad_read(lr_ctx *ctx, uint32_t grid, float *out, int is_device);
void *lr_grid_grad_device_ptr(lr_ctx *ctx, uint32_t grid);   /* for ncclAllReduce */
size_t lr_grid_size(const lr_ctx *ctx, uint32_t grid);       /* number of floats */

/* ---- ray-cast probe (parity of the LBVH against the oracle) ------------- */
/* rays[n][8] = ox,oy,oz,dx,dy,dz,maxt,unused ; hits[n][4] = t, prim(as float bits), u, v */
int  lr_ray_intersect(lr_ctx *ctx, const float *rays, uint64_t n, float *hits);
/* The same query without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
   callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops).  Its scratch -- staging of host buffers and
   the per-ray traversal stack of the 4-wide tree -- is one file-scope 3.375 MiB device array held by every GPU that loads the
   library (liverrenderer_b200/csrc/lr_noalloc.inl).  Limits: n <= lr_noalloc_max_rays() (8192) per call, else
   LR_ERR_INVALID_ARG; scenes whose tree needs more than 96 stack entries per ray return LR_ERR_UNSUPPORTED.  is_device = 0:
   HOST buffers, the call returns with hits filled; is_device = 1: 16-byte-aligned DEVICE buffers, the kernel is only
   enqueued on the context's stream.  Every call shares the array (nothing in it needs clearing between calls): calls that
   overlap on the GPU race -- serialise them.  Results equal lr_ray_intersect's bit for bit.                                */
int  lr_ray_intersect_noalloc(lr_ctx *ctx, const float *rays, uint64_t n, float *hits, int is_device);
uint64_t lr_noalloc_max_rays(void);

/* ---- device known-answer probes ------------------------------------------ */
/* The shading / lookup arithmetic of the hot path, one query per CUDA thread, so the reference's function-level tests run on
   the device code: src/bsdfs/tests/test_principled.py:124-205, test_rough_dielectric.py:209-229, test_diffuse.py:14-37,
   src/phase/tests/test_hg.py, test_isotropic.py, src/volumes/tests/test_grid.py:43-118, src/emitters/tests/test_envmap.py.
   All buffers are HOST pointers; directions are in the local shading frame (BSDF) or world space (phase, emitters).       */
/* BSDF::eval + BSDF::pdf: wi[n][3], wo[n][3] -> out[n][4] = value rgb (cosine included), pdf.  No textured parameter blocks. */
int  lr_kat_bsdf_eval_pdf(lr_ctx *ctx, const lr_bsdf *bsdf, const float *wi, const float *wo, uint64_t n, float *out);
/* BSDF::sample: wi[n][3], u[n][3] = sample1, sample2.x, sample2.y -> out[n][9] = wo(3), pdf, eta, weight(3), sampled_type (bits) */
int  lr_kat_bsdf_sample(lr_ctx *ctx, const lr_bsdf *bsdf, const float *wi, const float *u, uint64_t n, float *out);
/* PhaseFunction::sample + eval_pdf: wi[n][3], u[n][2] -> out[n][5] = wo(3), sampled pdf, eval_pdf(wi, wo) */
int  lr_kat_phase(lr_ctx *ctx, uint32_t kind, float g, const float *wi, const float *u, uint64_t n, float *out);
/* Volume::eval of grid `grid` of the uploaded scene at world points p[n][3] -> out[n][3]; use_fat = 0 forces the reference
   layout path, 1 lets the lookup use the fat-voxel copy where the grid has one (identical results)                        */
int  lr_kat_grid_eval(lr_ctx *ctx, uint32_t grid, const float *p, uint64_t n, float *out, int use_fat);
/* Scene::sample_emitter_direction (no visibility test): in[n][5] = ref_p(3), u(2) -> out[n][7] = d(3), pdf, weight rgb */
int  lr_kat_emitter_sample(lr_ctx *ctx, const float *ref_p_u, uint64_t n, float *out);
/* environment emitter of the scene: d[n][3] = direction of travel -> out[n][4] = radiance rgb, pdf_direction(d) */
int  lr_kat_env_eval_pdf(lr_ctx *ctx, const float *d, uint64_t n, float *out);
/* tissue media (liver.cpp:229-547 computeDistance + sample_interaction along the ray o = 0, d = +z):
   in[n][5] = sample, tissue depth, maxt, channel (uint32 bits), active mask (uint32 bits) ->
   out[n][8] = sampled distance, element type (bits: 0 absorber, 1 attenuator, 2 both), mei.t, mei.transmittance rgb,
               sigma_t[channel], combined_extinction[channel]                                                        */
int  lr_kat_tissue(lr_ctx *ctx, const lr_medium *medium, const float *in, uint64_t n, float *out);

/* ---- control ------------------------------------------------------------ */
int  lr_cancel(lr_ctx *ctx);                    /* async-signal-safe             */
int  lr_stats_get(lr_ctx *ctx, lr_stats *out);
int  lr_set_stream(lr_ctx *ctx, void *cuda_stream);  /* run on the caller's stream (torch current stream) */
/* profile mode (off by default): the next renders time every stage launch; lr_profile_get returns the sums since enable */
int  lr_profile_enable(lr_ctx *ctx, int on);
int  lr_profile_get(lr_ctx *ctx, lr_stage_profile *out);
const char *lr_stage_name(uint32_t stage);

#ifdef __cplusplus
}
#endif
#endif /* LR_API_H */
