// lr_lanes.h -- host entry points of lr_lanes.cu (per-lane kernels), called from lr_backend.cu
#pragma once
#include <cuda_runtime.h>

namespace lr {
struct DScene; struct RenderArgs;
// k_render<kind, channels, to_film> on a persistent grid (kind = sc.integ.kind; channels 4 = RGBW, 5 = RGBAW)
cudaError_t lanes_launch_render(int channels, bool to_film, int n_sm, cudaStream_t s, const DScene &sc, const RenderArgs &a);
// k_backward: per-lane prbvolpath primal + adjoint replay
cudaError_t lanes_launch_backward(int n_sm, cudaStream_t s, const DScene &sc, const RenderArgs &a);
}
