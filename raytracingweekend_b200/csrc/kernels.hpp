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

// Megakernel with material-sorted shading (sorted.cu): same contract; `sorted_supports` = its packing limits.
bool sorted_supports(const DevRenderParams& rp);
cudaError_t launch_render_sorted(const DevScene& sc, const DevRenderParams& rp, float4* accum, unsigned long long* ray_counter,
                                 cudaStream_t st);

cudaError_t launch_intersect(const DevScene& sc, const rtw_ray* rays, rtw_hit* hits, long long n, float tmin, float tmax,
                             cudaStream_t st);

// Wavefront pipeline (wavefront.cu): generate / extend / shade kernels over a pool of `n_slots` path slots in `scratch`
// (wavefront_scratch_bytes(n_slots) bytes of device memory).  Synchronises the stream between chunks of iterations to
// poll the dead-slot counter through `host_flag` (pinned host memory, >= 8 bytes).
size_t wavefront_scratch_bytes(uint32_t n_slots, bool nee = false);  // nee: + the shadow-ray queue of RTW_FLAG_NEE
cudaError_t launch_wavefront(const DevScene& sc, const DevRenderParams& rp, float4* accum, unsigned long long* ray_counter,
                             void* scratch, uint32_t n_slots, unsigned long long* host_flag, int* launches_out, cudaStream_t st);

// The wavefront pipeline on the library's static device scratch (wavefront.cu): allocates nothing, synchronous,
// n_slots <= wavefront_noalloc_max_slots(); `rays_out` is a HOST pointer.
uint32_t wavefront_noalloc_max_slots();
cudaError_t launch_wavefront_noalloc(const DevScene& sc, const DevRenderParams& rp, float4* accum, uint32_t n_slots,
                                     unsigned long long* rays_out, int* launches_out, cudaStream_t st);

// Known-answer hooks (kernels.cu): Philox4x32-10 on given counters; the Philox-driven samplers (0 ball, 1 lens disk, 2 cosine).
cudaError_t launch_philox(const uint4* counters, uint4* out, long long n, const DevRenderParams& rp);
cudaError_t launch_sampler(int which, float* out, long long n, const DevRenderParams& rp);

// FFMA throughput of the current device in TFLOP/s (roofline denominator, measured not assumed).
cudaError_t measure_fp32_peak(double* tflops);

}  // namespace rtw
