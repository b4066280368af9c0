// src/integrators/b200_volpath.cpp   (variant: cuda_ad_rgb or llvm_ad_rgb host, fp32 RGB)
#include <mitsuba/render/integrator.h>
#include <mitsuba/render/scene.h>
#include "lr_api.h"

NAMESPACE_BEGIN(mitsuba)
template <typename Float, typename Spectrum>
class B200Volpath final : public MonteCarloIntegrator<Float, Spectrum> {
public:
    MI_IMPORT_BASE(MonteCarloIntegrator, m_max_depth, m_rr_depth, m_hide_emitters)
    MI_IMPORT_TYPES(Scene, Sensor, Film, Mesh, Medium, BSDF, Emitter)

    B200Volpath(const Properties &props) : Base(props) {
        m_kind = props.get<std::string>("algorithm", "volpath") == "path" ? LR_INTEGRATOR_PATH : LR_INTEGRATOR_VOLPATH;
        m_tracking = props.get<bool>("reference_order", false) ? LR_TRACKING_REFERENCE : LR_TRACKING_DDA;
        if (lr_create(props.get<int>("device", 0), &m_ctx) != LR_OK) Throw("%s", lr_last_error(nullptr));
    }
    ~B200Volpath() { lr_destroy(m_ctx); }

    TensorXf render(Scene *scene, Sensor *sensor, UInt32 seed, uint32_t spp, bool develop, bool /*evaluate*/) override {
        if (!develop) Throw("b200_volpath: develop=false is outside the hot path");
        upload_if_dirty(scene, sensor);                          // flatten Scene -> lr_scene (see below)
        ScalarVector2u size = sensor->film()->crop_size();
        size_t ch = has_flag(sensor->film()->flags(), FilmFlags::Alpha) ? 4 : 3;
        std::vector<float> img((size_t) size.x() * size.y() * ch);
        if (lr_render(m_ctx, dr::slice(seed), spp ? spp : sensor->sampler()->sample_count(), img.data(), 0) != LR_OK)
            Throw("%s", lr_last_error(m_ctx));
        size_t shape[3] = { size.y(), size.x(), ch };
        return TensorXf(dr::load<DynamicBuffer<Float>>(img.data(), img.size()), 3, shape);
    }
    void cancel() override { lr_cancel(m_ctx); }
    // render_backward: same pattern -> lr_render_backward(m_ctx, seed, spp, grad_in.data(), 0, 0, 0), then
    // dr::accum_grad(params["…sigma_t.data"], lr_grid_grad_read(...))
    MI_DECLARE_CLASS(B200Volpath)
private:
    void upload_if_dirty(Scene *scene, Sensor *sensor);          // meshes: Mesh::vertex_positions_buffer()/faces_buffer() -> lr_mesh;
                                                                 // bsdf->id()/class name + traverse() values -> lr_bsdf; medium ->
                                                                 // lr_medium + GridVolume data tensor -> lr_grid; emitters; sensor -> lr_sensor;
                                                                 // BumpMap: traverse() gives "nested_bsdf", "nested_texture", "scale" ->
                                                                 // the nested BSDF's lr_bsdf with bump_tex/bump_scale/bump_flags set;
                                                                 // BitmapTexture "data" tensor + "to_uv" -> lr_texture (bitmap.cpp:430-437)
    lr_ctx *m_ctx = nullptr; uint32_t m_kind, m_tracking;
};
MI_EXPORT_PLUGIN(B200Volpath)                                    // generates extern "C" init_plugin (object.h:345-349)
NAMESPACE_END(mitsuba)
