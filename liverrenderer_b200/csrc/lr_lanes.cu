// lr_lanes.cu -- the one-lane-per-thread kernels (k_render: 5 integrators x 2 film layouts x {film, per-lane output};
// k_backward: per-lane prbvolpath adjoint) in their own translation unit.  They are the parity probes
// (lr_sample_lanes) and the LR_MEGAKERNEL=1 A/B path, and take half of the build time; the production wavefront
// kernels live in lr_backend.cu.  No relocatable device code: each TU links its own device image.
#define LR_LANES_TU 1
#include "lr_kernels.cuh"
#include "lr_lanes.h"
#include <algorithm>
#include <cstdlib>

namespace lr {

template <typename K> static cudaError_t persistent_grid(K kernel, int block, int n_sm, int &grid) {
    int per_sm = 0; cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, block, 0);
    grid = std::max(per_sm, 1) * n_sm; return e;                    // persistent: a whole number of CTAs per SM
}
template <int KIND, int C, bool TO_FILM> static cudaError_t launch_render_t(int n_sm, cudaStream_t s, const DScene &sc, const RenderArgs &a) {
    int grid; if (cudaError_t e = persistent_grid(k_render<KIND, C, TO_FILM>, 128, n_sm, grid)) return e;
    uint64_t warps_needed = (a.lane_end - a.lane_begin + 31) / 32;
    grid = (int) std::min<uint64_t>((uint64_t) grid, (warps_needed + 3) / 4); if (grid < 1) grid = 1;
    k_render<KIND, C, TO_FILM><<<grid, 128, 0, s>>>(sc, a);
    return cudaGetLastError();
}
template <int C, bool TO_FILM> static cudaError_t launch_render_k(int n_sm, cudaStream_t s, const DScene &sc, const RenderArgs &a) {
    switch (sc.integ.kind) {
        case LR_INTEGRATOR_PATH: return launch_render_t<LR_INTEGRATOR_PATH, C, TO_FILM>(n_sm, s, sc, a);
        case LR_INTEGRATOR_VOLPATH: return launch_render_t<LR_INTEGRATOR_VOLPATH, C, TO_FILM>(n_sm, s, sc, a);
        case LR_INTEGRATOR_BIOVOLPATH: return launch_render_t<LR_INTEGRATOR_BIOVOLPATH, C, TO_FILM>(n_sm, s, sc, a);
        case LR_INTEGRATOR_VOLPATHMIS: return sc.integ.spectral_mis ? launch_render_t<LR_INTEGRATOR_VOLPATHMIS, C, TO_FILM>(n_sm, s, sc, a)
                                                                     : launch_render_t<LR_KIND_VOLPATHMIS_PLAIN, C, TO_FILM>(n_sm, s, sc, a);
        default: return launch_render_t<LR_INTEGRATOR_PRBVOLPATH, C, TO_FILM>(n_sm, s, sc, a);
    }
}
cudaError_t lanes_launch_render(int channels, bool to_film, int n_sm, cudaStream_t s, const DScene &sc, const RenderArgs &a) {
    if (channels == 4) return to_film ? launch_render_k<4, true>(n_sm, s, sc, a) : launch_render_k<4, false>(n_sm, s, sc, a);
    return to_film ? launch_render_k<5, true>(n_sm, s, sc, a) : launch_render_k<5, false>(n_sm, s, sc, a);
}
cudaError_t lanes_launch_backward(int n_sm, cudaStream_t s, const DScene &sc, const RenderArgs &a) {
    int grid; if (cudaError_t e = persistent_grid(k_backward, 128, n_sm, grid)) return e;
    k_backward<<<grid, 128, 0, s>>>(sc, a);
    return cudaGetLastError();
}

} // namespace lr
