// kernels.cu — the megakernel path tracer and intersect_batch (sm_100a).
//
// Megakernel: one thread owns one pixel and runs its whole sample range.  Paths are regenerated in place: as
// soon as a lane's path ends it starts the pixel's next sample in the same loop iteration, so a warp never
// waits for its longest path (only for its most expensive pixel).  Radiance sums stay in registers; HBM sees a
// single 16-byte store per pixel per launch.  Warps cover 8x4 pixel tiles so primary rays and
// the first bounce stay coherent.
#include "kernels.hpp"
#include "rt_path.cuh"
#include "rt_stage.cuh"

namespace rtw {

constexpr int kBlock = 128;   // 4 warps
#ifndef RTW_MIN_BLOCKS
#define RTW_MIN_BLOCKS 6  // resident blocks per SM the register allocator must allow (tuned from ncu)
#endif
constexpr int kTileW = 16;    // block tile: 16 x 8 pixels, warp tile 8 x 4
constexpr int kTileH = 8;

template <uint32_t F, int SM>
__global__ void __launch_bounds__(kBlock, RTW_MIN_BLOCKS) render_megakernel(const __grid_constant__ DevScene sc,
                                                           const __grid_constant__ DevRenderParams rp,
                                                           float4* __restrict__ accum,
                                                           unsigned long long* __restrict__ ray_counter) {
    extern __shared__ uint4 smem[];
    const SceneView sv = stage_scene<SM>(sc, smem);
    int* stack = stack_base(sc, SM, smem) + threadIdx.x;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tiles_x = (rp.nx + kTileW - 1) / kTileW;
    const int tx = blockIdx.x % tiles_x, ty = blockIdx.x / tiles_x;
    const int i = tx * kTileW + (warp & 1) * 8 + (lane & 7);
    const int j = ty * kTileH + (warp >> 1) * 4 + (lane >> 3);
    const bool inside = i < rp.nx && j < rp.ny;

    unsigned long long rays = 0;
    if (inside) {
        int gi = i, gj = j;
        if (rp.flags & RTW_FLAG_DEBUG_RAY) {  // RayTracingWeekend.cpp:221-225
            gi = rp.nx / 2;
            gj = rp.ny / 2;
        }
        Rng rng;
        rng.pixel = (uint32_t)(j * rp.nx + i);
        rng.seed_lo = rp.seed_lo;
        rng.seed_hi = rp.seed_hi;
        rng.sample = 0;

        V3 sum = mk(0.0f, 0.0f, 0.0f);
        V3 beta = mk(1.0f, 1.0f, 1.0f), L = mk(0.0f, 0.0f, 0.0f);
        Ray ray;
        ray.o = ray.d = mk(0.0f, 0.0f, 0.0f);
        ray.time = 0.0f;
        int s = 0;
        uint32_t bounce = 0;
        bool alive = false;
        for (;;) {
            if (!alive) {
                if (s >= rp.spp_count) break;
                rng.sample = (uint32_t)(rp.spp_begin + s);  // global sample index: shards are disjoint streams
                s++;
                ray = camera_ray(sc, rp, rng, gi, gj);
                beta = mk(1.0f, 1.0f, 1.0f);
                L = mk(0.0f, 0.0f, 0.0f);
                bounce = 0;
                alive = true;
            }
            bounce++;
            rays++;
            alive = path_step<F, SM>(sc, sv, rng, bounce, ray, beta, L, stack, kBlock);
            if (alive && (int)bounce >= rp.max_depth) alive = false;  // :47-48: the next color() returns 0
            if (!alive) sum = sum + L;                                // :235-239
        }
        size_t idx = (size_t)j * rp.nx + i;  // canvas[j*nx+i] (:247)
        float4 out = make_float4(sum.x, sum.y, sum.z, (float)rp.spp_count);
        if (rp.flags & RTW_FLAG_ACCUMULATE) {
            float4 prev = accum[idx];
            out.x += prev.x, out.y += prev.y, out.z += prev.z, out.w += prev.w;
        }
        accum[idx] = out;
    }
    // one atomic per warp for the device-side ray counter
    for (int o = 16; o > 0; o >>= 1) rays += __shfl_down_sync(0xffffffffu, rays, o);
    if (lane == 0 && rays) atomicAdd(ray_counter, rays);
}

// world.hit() on a fixed ray batch (parity tiers 1-2): one thread per ray, media skipped.
template <uint32_t F, int SM>
__global__ void __launch_bounds__(kBlock) intersect_kernel(const __grid_constant__ DevScene sc,
                                                          const rtw_ray* __restrict__ rays, rtw_hit* __restrict__ hits,
                                                          long long n, float tmin, float tmax) {
    extern __shared__ uint4 smem[];
    const SceneView sv = stage_scene<SM>(sc, smem);
    int* stack = stack_base(sc, SM, smem) + threadIdx.x;
    long long idx = (long long)blockIdx.x * kBlock + threadIdx.x;
    if (idx >= n) return;
    rtw_ray r = rays[idx];
    V3 o = mk(r.ox, r.oy, r.oz), d = mk(r.dx, r.dy, r.dz);
    HitId h = closest_hit<F, SM, true>(sc, sv, o, d, r.time, tmin, tmax, nullptr, 0, stack, kBlock);
    rtw_hit out;
    if (h.prim < 0) {
        out.prim = -1, out.material = -1;
        out.t = out.u = out.v = out.nx = out.ny = out.nz = 0.0f;
    } else {
        uint32_t pm;
        (void)hit_material<SM>(sc, sv, h, pm);
        HitRec rec = complete_hit<F, SM>(sc, sv, h, pm, o, d, r.time, true);
        out.prim = h.prim, out.material = rec.mat;
        out.t = h.t, out.u = rec.u, out.v = rec.v;
        out.nx = rec.n.x, out.ny = rec.n.y, out.nz = rec.n.z;
    }
    hits[idx] = out;
}

// ------------------------------------------------------------------------------------------------ launchers
uint32_t pick_variant(uint32_t features) { return pick_variant_impl(features); }

namespace {
template <uint32_t F, int SM>
cudaError_t launch_render_t(const DevScene& sc, const DevRenderParams& rp, float4* accum, unsigned long long* counter,
                            size_t smem, cudaStream_t st) {
    int tiles_x = (rp.nx + kTileW - 1) / kTileW, tiles_y = (rp.ny + kTileH - 1) / kTileH;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(render_megakernel<F, SM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    render_megakernel<F, SM><<<tiles_x * tiles_y, kBlock, smem, st>>>(sc, rp, accum, counter);
    return cudaGetLastError();
}
template <uint32_t F, int SM>
cudaError_t launch_intersect_t(const DevScene& sc, const rtw_ray* rays, rtw_hit* hits, long long n, float tmin, float tmax,
                               size_t smem, cudaStream_t st) {
    long long blocks = (n + kBlock - 1) / kBlock;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(intersect_kernel<F, SM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    intersect_kernel<F, SM><<<(unsigned)blocks, kBlock, smem, st>>>(sc, rays, hits, n, tmin, tmax);
    return cudaGetLastError();
}
}  // namespace

cudaError_t launch_render(const DevScene& sc, const DevRenderParams& rp, float4* accum, unsigned long long* counter,
                          cudaStream_t st) {
    RTW_DISPATCH(kBlock, launch_render_t, sc, rp, accum, counter);
}

cudaError_t launch_intersect(const DevScene& sc, const rtw_ray* rays, rtw_hit* hits, long long n, float tmin, float tmax,
                             cudaStream_t st) {
    RTW_DISPATCH(kBlock, launch_intersect_t, sc, rays, hits, n, tmin, tmax);
}

}  // namespace rtw

// ------------------------------------------------------------------------------------------------ FP32 peak probe
// Roofline denominator: dependent-chain-free FFMA throughput measured in the same session.
namespace rtw {
__global__ void __launch_bounds__(256) fma_peak_kernel(float* out, int iters) {
    float a[16];
#pragma unroll
    for (int k = 0; k < 16; k++) a[k] = 1.0f + 1e-3f * (float)(threadIdx.x + k);
    const float m = 0.9999f, c = 1e-4f;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 16; k++) a[k] = fmaf(a[k], m, c);
    }
    float s = 0.0f;
#pragma unroll
    for (int k = 0; k < 16; k++) s += a[k];
    if (s == 123.456f) out[0] = s;  // keep the chains alive
}

cudaError_t measure_fp32_peak(double* tflops) {
    float* d = nullptr;
    cudaError_t e = cudaMalloc(&d, sizeof(float));
    if (e != cudaSuccess) return e;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    int sms = 0;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int iters = 1 << 14, blocks = sms * 8, threads = 256;
    double best = 0.0;
    for (int rep = 0; rep < 4; rep++) {
        cudaEventRecord(e0);
        fma_peak_kernel<<<blocks, threads>>>(d, iters);
        cudaEventRecord(e1);
        if ((e = cudaEventSynchronize(e1)) != cudaSuccess) break;
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        double fl = 2.0 * 16.0 * iters * (double)blocks * threads;
        if (rep > 0 && ms > 0) best = fmax(best, fl / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    *tflops = best;
    return e;
}
}  // namespace rtw
