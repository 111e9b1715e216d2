// kernels.hpp — host-callable launchers of the CUDA kernels (kernels.cu, wavefront.cu).
#pragma once
#include <cuda_runtime.h>

#include "../../include/rtw_b200.h"
#include "device_scene.hpp"

namespace rtw {

uint32_t pick_variant(uint32_t features);

// Megakernel: one launch renders samples [spp_begin, spp_begin+spp_count) of every pixel.
cudaError_t launch_render(const DevScene& sc, const DevRenderParams& rp, float4* accum, unsigned long long* ray_counter,
                          cudaStream_t st);

cudaError_t launch_intersect(const DevScene& sc, const rtw_ray* rays, rtw_hit* hits, long long n, float tmin, float tmax,
                             cudaStream_t st);

}  // namespace rtw
