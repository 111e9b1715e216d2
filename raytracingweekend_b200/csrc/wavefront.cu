// wavefront.cu — the same hot path as a wavefront pipeline (north_star: "generate / extend / shade kernels with
// warp-ballot stream compaction"), built from the SAME device functions as the megakernel (closest_hit, shade_hit,
// camera_ray), so both designs trace bit-identical paths for the same Philox keys; only the order in which a pixel's
// samples are summed differs (atomic adds).
//
// A pool of N path slots lives in global memory as three 16-byte SoA arrays plus a hit array (64 B per slot):
//   A = (ox, oy, oz, time)      B = (dx, dy, dz, pixel)      C = (beta.x, beta.y, beta.z, sample << 8 | bounce)
//   D = (t, prim, item, -)      written by extend, read by shade
// The pool is sized to stay L2-resident (default 1 Mi slots = 64 MiB of the 126 MB L2), so queue traffic is L2
// traffic, not HBM traffic.
//
//   wf_generate : every slot draws its first path (work item -> pixel, sample; camera ray).
//   wf_extend   : world.hit() for every live slot (all lanes converged on the traversal), classifies the hit by material
//                 type and appends the slot index to that class's queue — one ballot + one atomicAdd per warp per class.
//   wf_shade    : walks the class queues back to back; every warp sees one material class, so the material switch is
//                 converged.  Terminated paths add beta*e to the pixel accumulator (one float4 atomic) and the slot is
//                 regenerated in place (next work item), so the pool stays full until the work runs out.
//
// Work items are numbered sample-major over 8x4 pixel tiles: 32 consecutive items are 32 neighbouring pixels of one
// sample, so regenerated primary rays are coherent.
#include "kernels.hpp"
#include "rt_path.cuh"
#include "rt_stage.cuh"

namespace rtw {

constexpr int kWfBlock = 256;
constexpr int kWfClasses = 6;  // RTW_MAT_* (0..4) + miss/terminal (5)
constexpr uint32_t kDeadPixel = 0xffffffffu;
constexpr uint32_t kNoAllocMaxSlots = 65536;  // RTW_NOALLOC_MAX_POOL_SLOTS (include/rtw_b200.h): the static scratch holds this many path slots

struct WfParams {
    float4 *A, *B, *C, *D;
    uint32_t* queue;        // kWfClasses arrays of n_slots indices
    uint32_t* counts;       // 2 x kWfClasses (double buffered by iteration parity)
    unsigned long long* next_item;  // work-item allocator
    unsigned long long* dead;       // slots that found no more work
    unsigned long long total_items;  // tiles * 32 * spp_count
    uint32_t n_slots;
    uint32_t tiles_x, n_tile_pixels;  // padded pixel count = tiles * 32
    // RTW_FLAG_NEE: shadow rays of this iteration (o.xyz, time) (d.xyz, pixel) (weight.rgb, bounce | sample << 8); null = off
    float4 *S0, *S1, *S2;
    uint32_t* n_shadow;
};
constexpr uint32_t kSkipEmit = 0x80000000u;  // sign bit of the slot's ray time (>= 0): the next emitter hit was already sampled (NEE)

// Work item -> (pixel, sample) -> camera ray; writes the slot.  Lanes that find no work mark the slot dead.
// Called with any subset of lanes: the allocation is warp-aggregated over the calling lanes.
template <uint32_t F, int SM>
RTW_DEV void wf_start_path(const DevScene& sc, const DevRenderParams& rp, const WfParams& wp, uint32_t slot) {
    uint32_t pixel = kDeadPixel, sample = 0;
    int gi = 0, gj = 0;
    // until a work item lands inside the image or the items run out (the tiles are padded past the image: on a 1-pixel-wide
    // frame 7 of 8 items are empty, and a slot must not be declared dead while items remain)
    for (;;) {
        const unsigned mask = __activemask();
        const int lane = threadIdx.x & 31;
        const int leader = __ffs(mask) - 1;
        unsigned long long base = 0;
        if (lane == leader) base = atomicAdd(wp.next_item, (unsigned long long)__popc(mask));
        base = __shfl_sync(mask, base, leader);
        const unsigned long long w = base + __popc(mask & ((1u << lane) - 1u));
        if (w >= wp.total_items) break;  // no work left
        const uint32_t s = (uint32_t)(w / wp.n_tile_pixels), q = (uint32_t)(w % wp.n_tile_pixels);
        const uint32_t tile = q >> 5, l = q & 31;
        const int i = (int)((tile % wp.tiles_x) * 8 + (l & 7)), j = (int)((tile / wp.tiles_x) * 4 + (l >> 3));
        if (i < rp.nx && j < rp.ny) {
            pixel = (uint32_t)(j * rp.nx + i);
            sample = (uint32_t)rp.spp_begin + s;
            gi = i, gj = j;
            break;
        }
    }
    if (pixel == kDeadPixel) {
        wp.B[slot] = make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(kDeadPixel));
        atomicAdd(wp.dead, 1ull);
        return;
    }
    if (rp.flags & RTW_FLAG_DEBUG_RAY) gi = rp.nx / 2, gj = rp.ny / 2;  // RayTracingWeekend.cpp:221-225
    Rng rng;
    rng.pixel = pixel, rng.sample = sample, rng.rk = rp.rk;
    const Ray ray = camera_ray(sc, rp, rng, gi, gj);
    wp.A[slot] = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.time);
    wp.B[slot] = make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(pixel));
    wp.C[slot] = make_float4(1.0f, 1.0f, 1.0f, __uint_as_float(sample << 8));
}

template <uint32_t F, int SM>
__global__ void __launch_bounds__(kWfBlock) wf_generate(const __grid_constant__ DevScene sc, const __grid_constant__ DevRenderParams rp,
                                                        const __grid_constant__ WfParams wp) {
    for (uint32_t slot = blockIdx.x * kWfBlock + threadIdx.x; slot < wp.n_slots; slot += gridDim.x * kWfBlock)
        wf_start_path<F, SM>(sc, rp, wp, slot);
}

template <uint32_t F, int SM>
__global__ void __launch_bounds__(kWfBlock) wf_extend(const __grid_constant__ DevScene sc, const __grid_constant__ DevRenderParams rp,
                                                      const __grid_constant__ WfParams wp, int parity,
                                                      unsigned long long* __restrict__ ray_counter) {
    extern __shared__ uint4 smem[];
    const SceneView sv = stage_scene<SM>(sc, smem);
    int* stack = stack_base(sc, SM, smem) + threadIdx.x;
    uint32_t* counts = wp.counts + parity * kWfClasses;
    if (wp.n_shadow && blockIdx.x == 0 && threadIdx.x == 0) *wp.n_shadow = 0;  // consumed by the previous iteration's wf_shadow
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __shared__ uint32_t s_warp_count[kWfClasses][kWfBlock / 32];  // per class: lanes of each warp in this round
    __shared__ uint32_t s_base[kWfClasses][kWfBlock / 32];        // per class: queue position of each warp's first entry
    unsigned long long rays = 0;
    const uint32_t n_round = (wp.n_slots + kWfBlock - 1u) / kWfBlock * kWfBlock;  // block-uniform trip count
    for (uint32_t slot = blockIdx.x * kWfBlock + threadIdx.x; slot < n_round; slot += gridDim.x * kWfBlock) {
        int cls = -1;
        if (slot < wp.n_slots) {
            const float4 b = wp.B[slot];
            const uint32_t pixel = __float_as_uint(b.w);
            if (pixel != kDeadPixel) {
                const float4 a = wp.A[slot];
                const uint32_t sb = __float_as_uint(wp.C[slot].w);
                Rng rng;
                rng.pixel = pixel, rng.sample = sb >> 8, rng.rk = rp.rk;
                const uint32_t bounce = (sb & 0xffu) + 1u;
                // :52 world.hit(r, 0.001f, DBL_MAX, rec)
                const HitId h = closest_hit<F, SM, false>(sc, sv, mk(a.x, a.y, a.z), mk(b.x, b.y, b.z), fabsf(a.w), 0.001f, FLT_MAX, &rng,
                                                          bounce, stack, kWfBlock);
                wp.D[slot] = make_float4(h.t, __int_as_float(h.prim), __int_as_float(h.item), 0.0f);
                rays++;
                cls = kWfClasses - 1;
                if (h.prim != -1 && sc.render_type != RTW_RENDER_NORMAL) {
                    uint32_t pm;
                    const int mat = hit_material<SM>(sc, sv, h, pm);
                    cls = (int)cold1<SM>(sv, sc.off[SEC_MATS] + (uint32_t)mat * 32u + 16u);  // DevMat.type
                }
            }
        }
        // Stream compaction: warp ballots give every lane its rank inside the warp; the block then claims ONE queue range
        // per class with a single global atomic (a per-warp atomic on six shared counters serialises in L2).
        unsigned my_mask = 0;
#pragma unroll
        for (int c = 0; c < kWfClasses; c++) {
            const unsigned m = __ballot_sync(0xffffffffu, cls == c);
            if (lane == 0) s_warp_count[c][warp] = (uint32_t)__popc(m);
            if (cls == c) my_mask = m;
        }
        __syncthreads();
        if (threadIdx.x < kWfClasses) {
            const int c = threadIdx.x;
            uint32_t total = 0;
#pragma unroll
            for (int w = 0; w < kWfBlock / 32; w++) total += s_warp_count[c][w];
            uint32_t base = total ? atomicAdd(&counts[c], total) : 0u;
#pragma unroll
            for (int w = 0; w < kWfBlock / 32; w++) {
                s_base[c][w] = base;
                base += s_warp_count[c][w];
            }
        }
        __syncthreads();
        if (cls >= 0) wp.queue[(size_t)cls * wp.n_slots + s_base[cls][warp] + __popc(my_mask & ((1u << lane) - 1u))] = slot;
    }
    for (int o = 16; o > 0; o >>= 1) rays += __shfl_down_sync(0xffffffffu, rays, o);
    if (lane == 0 && rays) atomicAdd(ray_counter, rays);
}

template <uint32_t F, int SM>
__global__ void __launch_bounds__(kWfBlock) wf_shade(const __grid_constant__ DevScene sc, const __grid_constant__ DevRenderParams rp,
                                                     const __grid_constant__ WfParams wp, int parity, float4* __restrict__ accum) {
    extern __shared__ uint4 smem[];
    const SceneView sv = stage_scene<SM>(sc, smem);
    const uint32_t* counts = wp.counts + parity * kWfClasses;
    // the other parity's counters are zeroed here for the next extend (which runs after this kernel in stream order)
    if (blockIdx.x == 0 && threadIdx.x < kWfClasses) wp.counts[(parity ^ 1) * kWfClasses + threadIdx.x] = 0;
    // class ranges in the virtual index space, each padded to a warp multiple so no warp straddles two classes
    uint32_t start[kWfClasses + 1];
    start[0] = 0;
#pragma unroll
    for (int c = 0; c < kWfClasses; c++) start[c + 1] = start[c] + ((counts[c] + 31u) & ~31u);
    for (uint32_t e = blockIdx.x * kWfBlock + threadIdx.x; e < start[kWfClasses]; e += gridDim.x * kWfBlock) {
        int c = 0;
#pragma unroll
        for (int k = 1; k < kWfClasses; k++) c += (e >= start[k]) ? 1 : 0;
        const uint32_t idx = e - start[c];
        if (idx >= counts[c]) continue;
        const uint32_t slot = wp.queue[(size_t)c * wp.n_slots + idx];
        const float4 a = wp.A[slot], b = wp.B[slot], cc = wp.C[slot], d = wp.D[slot];
        const uint32_t pixel = __float_as_uint(b.w), sb = __float_as_uint(cc.w);
        Rng rng;
        rng.pixel = pixel, rng.sample = sb >> 8, rng.rk = rp.rk;
        const uint32_t bounce = (sb & 0xffu) + 1u;
        Ray ray;
        ray.o = mk(a.x, a.y, a.z), ray.d = mk(b.x, b.y, b.z), ray.time = fabsf(a.w);
        V3 beta = mk(cc.x, cc.y, cc.z), L = mk(0.0f, 0.0f, 0.0f);
        HitId h;
        h.t = d.x, h.prim = __float_as_int(d.y), h.item = __float_as_int(d.z);
        bool alive, skip_next = false;
        const bool nee = wp.n_shadow != nullptr && sc.n_nee_lights > 0;
        if (nee && c == RTW_MAT_LAMBERTIAN) {  // warp-uniform (c is the queue's class): rt_path.cuh shade_lambert_nee
            ShadowRay sh;
            shade_lambert_nee<F, SM>(sc, sv, rng, bounce, h, ray, beta, sh);
            if (sh.cast) {
                const uint32_t q = atomicAdd(wp.n_shadow, 1u);
                wp.S0[q] = make_float4(sh.o.x, sh.o.y, sh.o.z, ray.time);
                wp.S1[q] = make_float4(sh.d.x, sh.d.y, sh.d.z, b.w);
                wp.S2[q] = make_float4(sh.w.x, sh.w.y, sh.w.z, __uint_as_float((sb & ~0xffu) | bounce));
            }
            alive = skip_next = true;
        } else if (nee && c == RTW_MAT_DIFFUSE_LIGHT && (__float_as_uint(a.w) & kSkipEmit) &&
                   nee_samples_prim<SM>(sc, sv, h.prim, cold1<SM>(sv, sc.off[SEC_PM] + (uint32_t)max(h.prim, 0) * 4u))) {
            alive = false;  // this light's emission was counted by the shadow ray of the previous (lambertian) bounce
        } else {
            alive = shade_hit<F, SM>(sc, sv, rng, bounce, h, ray, beta, L);
        }
        if (alive && (int)bounce >= rp.max_depth) alive = false;  // :47-48: the next color() returns 0
        if (alive) {
            if (skip_next) ray.time = __uint_as_float(__float_as_uint(ray.time) | kSkipEmit);
            wp.A[slot] = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.time);
            wp.B[slot] = make_float4(ray.d.x, ray.d.y, ray.d.z, b.w);
            wp.C[slot] = make_float4(beta.x, beta.y, beta.z, __uint_as_float((sb & ~0xffu) | bounce));
        } else {
            // :235-239: radiance enters only at terminal events, so the sample's value is L
            if (L.x != 0.0f || L.y != 0.0f || L.z != 0.0f) atomicAdd(&accum[pixel], make_float4(L.x, L.y, L.z, 0.0f));
            wf_start_path<F, SM>(sc, rp, wp, slot);
        }
    }
}

// RTW_FLAG_NEE: the shadow rays of this iteration.  A shadow ray is an ordinary world.hit() (media on its way scatter it with
// their usual probability, from Philox counters of their own: bounce + 0x8000); it delivers  weight * Le  when the first thing
// it meets is one of the sampled lights, seen from its emitting side (material.h:238-244).
template <uint32_t F, int SM>
__global__ void __launch_bounds__(kWfBlock) wf_shadow(const __grid_constant__ DevScene sc, const __grid_constant__ DevRenderParams rp,
                                                      const __grid_constant__ WfParams wp, float4* __restrict__ accum) {
    extern __shared__ uint4 smem[];
    const SceneView sv = stage_scene<SM>(sc, smem);
    int* stack = stack_base(sc, SM, smem) + threadIdx.x;
    const uint32_t n = *wp.n_shadow;
    for (uint32_t q = blockIdx.x * kWfBlock + threadIdx.x; q < n; q += gridDim.x * kWfBlock) {
        const float4 s0 = wp.S0[q], s1 = wp.S1[q], s2 = wp.S2[q];
        const uint32_t pixel = __float_as_uint(s1.w), sb = __float_as_uint(s2.w);
        Rng rng;
        rng.pixel = pixel, rng.sample = sb >> 8, rng.rk = rp.rk;
        const V3 o = mk(s0.x, s0.y, s0.z), dv = mk(s1.x, s1.y, s1.z);
        const HitId h = closest_hit<F, SM, false>(sc, sv, o, dv, s0.w, 0.001f, FLT_MAX, &rng, (sb & 0xffu) + 0x8000u, stack, kWfBlock);
        if (h.prim < 0) continue;
        uint32_t pm;
        const int mat_idx = hit_material<SM>(sc, sv, h, pm);
        const DevMat m = ld_mat<SM>(sv, sc.off[SEC_MATS] + (uint32_t)mat_idx * 32u);
        if (m.type != RTW_MAT_DIFFUSE_LIGHT || !nee_samples_prim<SM>(sc, sv, h.prim, pm)) continue;
        const bool need_uv = (F & F_TEXTURES) && !m.tex_is_constant;
        const HitRec rec = complete_hit<F, SM>(sc, sv, h, pm, o, dv, s0.w, need_uv);
        if (!(dot(rec.n, dv) > 0.0f)) continue;  // one-sided emission (Q23)
        const V3 le = material_texture<F, SM>(sc, sv, m, rec.u, rec.v, rec.p);
        atomicAdd(&accum[pixel], make_float4(s2.x * le.x, s2.y * le.y, s2.z * le.z, 0.0f));
    }
}

// accum.w = number of samples (RayTracingWeekend.cpp:241): added once at the end
__global__ void wf_finalize(float4* accum, int n, float spp) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) accum[i].w += spp;
}

namespace {
template <uint32_t F, int SM>
cudaError_t wf_run_t(const DevScene& sc, const DevRenderParams& rp, const WfParams& wp, float4* accum, unsigned long long* ray_counter,
                     int grid, int iterations, int first_iteration, bool generate, size_t smem, cudaStream_t st) {
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(wf_extend<F, SM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(wf_shade<F, SM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(wf_shadow<F, SM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    if (generate) wf_generate<F, SM><<<grid, kWfBlock, 0, st>>>(sc, rp, wp);
    for (int it = 0; it < iterations; it++) {
        const int parity = (first_iteration + it) & 1;
        wf_extend<F, SM><<<grid, kWfBlock, smem, st>>>(sc, rp, wp, parity, ray_counter);
        wf_shade<F, SM><<<grid, kWfBlock, smem, st>>>(sc, rp, wp, parity, accum);
        if (wp.n_shadow) wf_shadow<F, SM><<<grid, kWfBlock, smem, st>>>(sc, rp, wp, accum);
    }
    return cudaGetLastError();
}
cudaError_t wf_run(const DevScene& sc, const DevRenderParams& rp, const WfParams& wp, float4* accum, unsigned long long* ray_counter,
                   int grid, int iterations, int first_iteration, bool generate, cudaStream_t st) {
    RTW_DISPATCH(kWfBlock, wf_run_t, sc, rp, wp, accum, ray_counter, grid, iterations, first_iteration, generate);
}
}  // namespace

size_t wavefront_scratch_bytes(uint32_t n_slots, bool nee) {
    return (size_t)n_slots * ((nee ? 7 : 4) * sizeof(float4) + kWfClasses * sizeof(uint32_t)) + 256;
}

// Renders the whole sample range.  `scratch` holds wavefront_scratch_bytes(n_slots) bytes; `host_flag` is pinned host
// memory (2 x u64) used to poll the dead-slot counter between chunks of iterations.
cudaError_t launch_wavefront(const DevScene& sc, const DevRenderParams& rp, float4* accum, unsigned long long* ray_counter,
                             void* scratch, uint32_t n_slots, unsigned long long* host_flag, int* launches_out, cudaStream_t st) {
    WfParams wp{};
    uint8_t* p = (uint8_t*)scratch;
    wp.A = (float4*)p, p += (size_t)n_slots * sizeof(float4);
    wp.B = (float4*)p, p += (size_t)n_slots * sizeof(float4);
    wp.C = (float4*)p, p += (size_t)n_slots * sizeof(float4);
    wp.D = (float4*)p, p += (size_t)n_slots * sizeof(float4);
    wp.queue = (uint32_t*)p, p += (size_t)n_slots * kWfClasses * sizeof(uint32_t);
    wp.counts = (uint32_t*)p, p += 2 * kWfClasses * sizeof(uint32_t) + 16;
    p = (uint8_t*)(((uintptr_t)p + 15) & ~(uintptr_t)15);
    wp.next_item = (unsigned long long*)p, p += 8;
    wp.dead = (unsigned long long*)p, p += 8;
    if (rp.flags & RTW_FLAG_NEE) {  // scratch was sized with nee = true
        wp.n_shadow = (uint32_t*)p, p += 16;
        p = (uint8_t*)(((uintptr_t)p + 15) & ~(uintptr_t)15);
        wp.S0 = (float4*)p, p += (size_t)n_slots * sizeof(float4);
        wp.S1 = (float4*)p, p += (size_t)n_slots * sizeof(float4);
        wp.S2 = (float4*)p, p += (size_t)n_slots * sizeof(float4);
    }
    wp.n_slots = n_slots;
    wp.tiles_x = (uint32_t)((rp.nx + 7) / 8);
    const uint32_t tiles_y = (uint32_t)((rp.ny + 3) / 4);
    wp.n_tile_pixels = wp.tiles_x * tiles_y * 32u;
    wp.total_items = (unsigned long long)wp.n_tile_pixels * (unsigned long long)rp.spp_count;

    cudaError_t e;
    const size_t npix = (size_t)rp.nx * rp.ny;
    if (!(rp.flags & RTW_FLAG_ACCUMULATE))
        if ((e = cudaMemsetAsync(accum, 0, npix * sizeof(float4), st)) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(wp.counts, 0, 2 * kWfClasses * sizeof(uint32_t) + 16 + 32, st)) != cudaSuccess) return e;

    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int grid = (int)((n_slots + kWfBlock - 1) / kWfBlock);
    if (grid > sms * 8) grid = sms * 8;

    int launches = 1, iteration = 0;
    const int chunk = 32;  // iterations between polls of the dead-slot counter
    for (;;) {
        if ((e = wf_run(sc, rp, wp, accum, ray_counter, grid, chunk, iteration, iteration == 0, st)) != cudaSuccess) return e;
        iteration += chunk;
        launches += (wp.n_shadow ? 3 : 2) * chunk;
        if ((e = cudaMemcpyAsync(host_flag, wp.dead, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
        if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return e;
        if (*host_flag >= (unsigned long long)n_slots) break;
        if (iteration > (1 << 24)) return cudaErrorUnknown;  // cannot happen: every path ends within max_depth bounces
    }
    wf_finalize<<<(unsigned)((npix + 255) / 256), 256, 0, st>>>(accum, (int)npix, (float)rp.spp_count);
    launches++;
    if (launches_out) *launches_out = launches;
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ no-allocation variant
// The same pipeline with its scratch (path-slot pool, class queues, counters, ray counter) in a file-scope device array
// instead of memory from the pool allocator: for callers that may not allocate (CUDA graph capture around the call site,
// tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Every GPU that loads the library holds the array (kNoAllocBytes, ~5.8 MB) until the process exits.
// ALL calls on a device share it and it keeps the last call's contents: what must start at zero (queue counters, work-item
// and dead-slot counters, the ray counter) is cleared at the start of every call, the slot pool is fully rewritten by
// wf_generate, queue entries are written before they are read.  Two calls that overlap on the GPU race on it.
constexpr size_t kNoAllocRayCounterOff = ((size_t)kNoAllocMaxSlots * (4 * sizeof(float4) + kWfClasses * sizeof(uint32_t)) + 256 + 255) & ~(size_t)255;
constexpr size_t kNoAllocBytes = kNoAllocRayCounterOff + 256;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address is resolved once per device,
// with that device current, and only that pointer is handed to the kernels.
static cudaError_t noalloc_scratch_ptr(uint8_t** out) {
    static uint8_t* ptrs[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (!ptrs[dev]) {
        void* p = nullptr;
        if ((e = cudaGetSymbolAddress(&p, g_noalloc_scratch)) != cudaSuccess) return e;
        ptrs[dev] = (uint8_t*)p;
    }
    *out = ptrs[dev];
    return cudaSuccess;
}

uint32_t wavefront_noalloc_max_slots() { return kNoAllocMaxSlots; }

// Synchronous like launch_wavefront (the host polls the dead-slot counter).  Calls no cudaMalloc / cudaFree / cudaMallocHost:
// the polls land in a stack variable.  n_slots <= kNoAllocMaxSlots (checked by the caller, re-checked here).
cudaError_t launch_wavefront_noalloc(const DevScene& sc, const DevRenderParams& rp, float4* accum, uint32_t n_slots,
                                     unsigned long long* rays_out, int* launches_out, cudaStream_t st) {
    if (n_slots == 0 || n_slots > kNoAllocMaxSlots || wavefront_scratch_bytes(n_slots, false) > kNoAllocRayCounterOff || (rp.flags & RTW_FLAG_NEE)) return cudaErrorInvalidValue;
    uint8_t* base = nullptr;
    cudaError_t e = noalloc_scratch_ptr(&base);
    if (e != cudaSuccess) return e;
    unsigned long long* ray_counter = (unsigned long long*)(base + kNoAllocRayCounterOff);
    if ((e = cudaMemsetAsync(ray_counter, 0, sizeof(unsigned long long), st)) != cudaSuccess) return e;
    unsigned long long host_flag[2] = {0, 0};  // pageable: the copies are staged by the driver, nothing is allocated here
    if ((e = launch_wavefront(sc, rp, accum, ray_counter, base, n_slots, host_flag, launches_out, st)) != cudaSuccess) return e;
    unsigned long long rays = 0;
    if ((e = cudaMemcpyAsync(&rays, ray_counter, sizeof rays, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
    if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return e;
    if (rays_out) *rays_out = rays;
    return cudaSuccess;
}

}  // namespace rtw
