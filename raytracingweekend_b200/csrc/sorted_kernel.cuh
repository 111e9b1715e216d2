// sorted.cu — the megakernel with material-sorted shading (sm_100a).
//
// The lane-owns-a-path megakernel (kernels.cu) traverses at ~31 of 32 lanes but shades at 3-15: after the hit a warp
// holds a mix of lambertian, dielectric, metal and terminal events and runs every branch.  Here the path state lives in
// shared memory instead of registers, and the warps of a block exchange paths through per-class queues so that a
// warp always shades 32 paths of ONE material class and then traces their 32 new rays:
//
//   block  = 256, 512 or 768 threads (per variant class, launch_sorted_t), one 16x16 pixel tile, one pool of (pixel, sample) items like the
//            warp pool of kernels.cu
//   slots  = kSPool path slots in shared memory, 64 B each: (o, t) (d, time) (beta, prim) (item, pixel, sample, bounce)
//   queues = one ring of slot ids per class in shared memory {lambertian, dielectric, metal, isotropic, terminal
//            (miss / emitter), needs-a-camera-ray}; producers reserve with one atomicAdd per (warp, class) found by
//            match_any, consumers claim 32 entries with a CAS on the head.  No block-wide barrier after start-up.
//   unit   = claim 32 entries of one class -> load state -> shade_hit (coherent) -> new ray -> closest_hit (full
//            lanes) -> classify -> store state -> push.  Terminal units add their radiance and start 32 new camera
//            rays on the spot, so camera maths also runs at full lanes.
//
// Rings: capacity >= number of slots, 16-bit entries = slot id | lap of the ring index (4 bits).  A consumer takes (and clears) its entries right after
// the CAS that claimed them, waiting for an entry of ITS lap where the producer has not written yet; a producer waits until its
// target position is empty before it writes, so a delayed consumer is never overwritten (ring_take / ring_put below).
// Per-pixel sums are 64-bit fixed point (2^-24) in shared memory (kept as two 32-bit words, see the adds): integer adds commute, so the image stays
// bit-reproducible although units run in a timing-dependent order.  Samples that are not finite or >= 2^30 go to a
// float side accumulator (NaN / Inf propagate like in the reference).  HBM traffic is unchanged: one 16-byte
// store per pixel per launch.
// This file is the kernel TEMPLATE and its launcher; it is included by sorted.cu (all variants that fit the SM instruction
// cache) and by sorted_big.cu (the all-features unstaged variants, compiled in compact mode: RTW_COMPACT).
#pragma once
#include <cstdio>
#include <cstdlib>

#include "kernels.hpp"
#include "rt_path.cuh"
#include "rt_stage.cuh"

namespace rtw {

constexpr int kSBlock = 256;
constexpr int kSTile = 16;  // 16 x 16 pixels per block
constexpr int kSPix = kSTile * kSTile;
// BS = threads per block, PT = path slots per 256 threads, MB = resident blocks per SM (launch bounds -> registers).
// Measured choices (profiles/r1_ab_*.txt), set in launch_sorted_t.  The common thread: ONE big pool per SM shared by many
// warps keeps the class queues full (whole 32-path units, few waits), as long as the registers suffice:
//   Cornell-class variant (F_XF only):        512 threads, 1024 slots, 2 blocks/SM: 64 registers, issue-bound, wants 32
//                                             warps per SM: 6.40 -> 6.53 Gpaths/s over 256 / 512 / 4 blocks; 1024 threads
//                                             with 2048 slots in one block loses again (6.33);
//   texture variant (Perlin, image, checker): 768 threads, 1536 slots, 1 block/SM, 85 registers: Perlin 7.41 -> 8.88,
//                                             checker spheres 5.43 -> 7.53, light_sample 6.78 -> 10.2 over 256 threads /
//                                             768 slots / 2 blocks (896 threads the same, 512 and 1024 threads less);
//   BVH variants on unstaged scenes:          256 threads, 512 slots, 2 blocks/SM: 128 registers, no spills, and two 64 KB
//                                             blocks leave ~100 KB of L1 for nodes and records (Book-2 final 0.485 -> 0.59;
//                                             bigger pools or blocks lose: they eat the L1);
//   all-features variant on staged scenes:    256 threads, 512 slots, 3 blocks/SM (Cornell smoke 4.14; 768 threads in one
//                                             block the same, 512 threads 3.37, 2 blocks of 256 3.32).
__host__ __device__ constexpr int pool_slots(int pt) { return 256 * pt; }
#ifndef RTW_QCAP_MUL
#define RTW_QCAP_MUL 1  // A/B: ring capacity = this many times the smallest power of two >= slots
#endif
__host__ __device__ constexpr int queue_cap(int pt) {  // ring capacity per class (power of two >= slots)
    return RTW_QCAP_MUL * (pool_slots(pt) <= 512 ? 512 : (pool_slots(pt) <= 1024 ? 1024 : 2048));
}
__host__ __device__ constexpr int min_blocks(int pt) { return pt <= 2 ? 4 : (pt == 3 ? 3 : 2); }

// C_TRACE exists in the PHASED kernels only (PH below): rays that wait for their closest_hit.
enum : int { C_LAMB = 0, C_DIEL, C_METAL, C_ISO, C_TERM, C_REGEN, C_TRACE, C_DEAD };
__host__ __device__ constexpr int n_queues(bool ph) { return ph ? C_TRACE + 1 : C_TRACE; }
constexpr uint16_t kQEmpty = 0xffffu;
// Phased kernels (template parameter PH; instantiated where RTW_PHASED is defined — sorted_big.cu, the unified binary-tree variants —
// with 512 threads, 1024 slots, one block per SM).  The all-features kernel touches 44.6 KB of code on the Book-2 final scene and
// the SM's instruction cache holds 32 KB: with every warp somewhere else in the unit, 13 % of its fetches miss, and the GPC-level
// instruction cache behind it serves ~7.4 G requests/s for the whole chip whatever the kernel does (profiles/r2_icache_wall.txt)
// — that, not issue slots or lanes, is what this kernel's time follows.  Here a unit is EITHER "shade 32 paths of one class and
// make their rays" (queued on C_TRACE) OR "closest_hit for 32 rays of C_TRACE, then the class queue of what they hit", and the
// block has a MODE: trace units while C_TRACE holds a whole one, then shade units while a class does, and so on.  No barrier: a
// warp that finds its mode's queues short takes the other kind and turns the mode (with a __syncthreads per phase the cache hit
// rate is 97 % but the warps idle at the barrier: 0.42 against 0.60 Gpaths/s; with the mode 92 % and 0.64, 0.71 at 1024 spp).
// Random numbers, hits and the integer pixel sums do not depend on the order: images are bit-identical to the plain kernel's.
#if defined(RTW_PHASED_BVH)  // A/B: every variant with trees
__host__ __device__ constexpr bool phased_variant(uint32_t f, int) { return (f & F_BVH) != 0; }
#elif defined(RTW_PHASED)
__host__ __device__ constexpr bool phased_variant(uint32_t f, int sm) { return unified_variant(f, sm); }
#else
__host__ __device__ constexpr bool phased_variant(uint32_t, int) { return false; }
#endif

#ifdef RTW_SORTED_STATS
__device__ int g_stats[8][2];
#endif
// The two spin loops of the rings.  Two properties, both found with the phased kernels, whose rings run nearly FULL (the plain
// kernels' rings run nearly empty and never showed either; tools: -DRTW_SORTED_CHECK, profiles/r2_ab_phased.txt):
//  * written in PTX so that each LANE acts in the very iteration in which its own condition holds.  From C++
//    (`while (*e == kQEmpty) {}  *e = kQEmpty;`) the store lands behind the loop's reconvergence point (BSYNC, then STS): a lane
//    that has its entry clears its position only when the slowest lane of its warp has one too, and a lane whose position is
//    free writes only when all are — "claimed but not cleared" then lasts as long as the slowest producer, not a few cycles;
//  * entries carry the LAP of their ring index in bits 12-15 ((index / capacity) & 15; slot ids are < 2^11, so no entry equals
//    the empty mark 0xffff), and a consumer only accepts an entry of its own lap.  Seen without it (256 threads x 512 slots,
//    capacity 512: 3 of 11 renders wrong, then 3 of 41 with the PTX loops, 0 of 41 with one lap bit, which the four replaced):
//    a producer held up between its reservation and its write while the ring went round once; the consumers of index i and of
//    index i + capacity both spin on the position and both take the entry the moment it lands — one slot owned twice,
//    another orphaned for a lap.  With four lap bits it takes a hold-up of SIXTEEN laps.
// ring_take: wait until the position holds an entry of lap `tag` (lap << 12), take it and leave the position empty.
__device__ __forceinline__ int ring_take(uint32_t saddr, uint32_t tag) {
    uint32_t v;
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        ".reg .b16 t, e;\n\t"
        ".reg .b32 w, x;\n\t"
        "mov.b16 e, 0xffff;\n"
        "RTW_RING_TAKE:\n\t"
        "ld.volatile.shared.u16 t, [%1];\n\t"
        "cvt.u32.u16 w, t;\n\t"
        "and.b32 x, w, 0xf000;\n\t"
        "setp.eq.u32 p, x, %2;\n\t"
        "setp.ne.and.u32 p, w, 0xffff, p;\n\t"
        "@p st.volatile.shared.u16 [%1], e;\n\t"
        "@!p bra RTW_RING_TAKE;\n\t"
        "and.b32 %0, w, 0x0fff;\n\t"
        "}"
        : "=r"(v)
        : "r"(saddr), "r"(tag)
        : "memory");
    return (int)v;
}
// ring_put: wait until the position is empty (its consumer of one lap ago has read it), then write the entry (slot | lap << 12).
__device__ __forceinline__ void ring_put(uint32_t saddr, uint16_t entry) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        ".reg .b16 t, e;\n\t"
        "mov.b16 e, 0xffff;\n"
        "RTW_RING_PUT:\n\t"
        "ld.volatile.shared.u16 t, [%0];\n\t"
        "setp.eq.u16 p, t, e;\n\t"
        "@p st.volatile.shared.u16 [%0], %1;\n\t"
        "@!p bra RTW_RING_PUT;\n\t"
        "}"
        :
        : "r"(saddr), "h"(entry)
        : "memory");
}

struct SortedCtl {
    unsigned int head[8], tail[8];  // per class ring: [head, tail) are queued entries (tail = reserved)
    unsigned int item_next;         // the block's (pixel, sample) pool
    int live;                       // slots that still carry or may get a path
    int mode;                       // phased kernels: 0 = trace units preferred, 1 = shade units
};

#ifdef RTW_SORTED_CHECK
#define RTW_CHECK_BYTES(pt) ((size_t)pool_slots(pt) * 16)
__device__ int g_check_first = 0;
#else
#define RTW_CHECK_BYTES(pt) 0
#endif
constexpr size_t sorted_pool_bytes(int pt, bool ph) {
    return RTW_CHECK_BYTES(pt) + (size_t)pool_slots(pt) * 64 + (size_t)n_queues(ph) * queue_cap(pt) * 2 + (size_t)kSPix * (3 * (8 + 4) + 4) + sizeof(SortedCtl) + 64;
}

// BS = threads per block (a multiple of 32; the pixel tile and the pool do not depend on it)
template <uint32_t F, int SM, int PT, int MB, int BS, bool PH = false>
__global__ void __launch_bounds__(BS, MB) render_sorted(const __grid_constant__ DevScene sc,
                                                                              const __grid_constant__ DevRenderParams rp,
                                                                              float4* __restrict__ accum,
                                                                              unsigned long long* __restrict__ ray_counter,
                                                                              uint32_t pool_off) {
    constexpr int kSPool = pool_slots(PT), kQCap = queue_cap(PT), C_QUEUES = n_queues(PH);
    extern __shared__ uint4 smem[];
    const SceneView sv = stage_scene<SM>(sc, smem);
    int* stack = stack_base(sc, SM, smem) + threadIdx.x;
    // pool carve-up (pool_off = staged bytes + stacks, 16-byte aligned by the host)
    uint8_t* base = reinterpret_cast<uint8_t*>(smem) + pool_off;
    float4* stA = reinterpret_cast<float4*>(base);  // o.xyz, t
    float4* stB = stA + kSPool;                     // d.xyz, time
    float4* stC = stB + kSPool;                     // beta.xyz, prim
    uint4* stD = reinterpret_cast<uint4*>(stC + kSPool);  // item, global pixel index, sample, bounce | pixel slot << 16
    unsigned int* fxlo = reinterpret_cast<unsigned int*>(stD + kSPool);  // [3][kSPix] fixed-point sums, low words
    unsigned int* fxhi = fxlo + 3 * kSPix;                               //                                high words
    float* fl = reinterpret_cast<float*>(fxhi + 3 * kSPix);              // [3][kSPix] side sums
    unsigned int* prays = reinterpret_cast<unsigned int*>(fl + 3 * kSPix);          // [kSPix] color() invocations per pixel
    volatile uint16_t* queue = reinterpret_cast<volatile uint16_t*>(prays + kSPix);  // [C_QUEUES][kQCap]
    SortedCtl* ctl = reinterpret_cast<SortedCtl*>(const_cast<uint16_t*>(queue) + C_QUEUES * kQCap);
#ifdef RTW_SORTED_CHECK
    int* own = reinterpret_cast<int*>(reinterpret_cast<uint8_t*>(ctl) + sizeof(SortedCtl) + 64);  // [kSPool] 1 while a warp holds the slot
    int* push_idx = own + kSPool;   // ring index the slot was last queued at
    int* push_cls = push_idx + kSPool;  // ... and its class | pushing warp << 8
    int* take_idx = push_cls + kSPool;  // ring index it was last taken from
    int anomalies = 0;
#endif
    volatile unsigned int* vhead = ctl->head;
    volatile unsigned int* vtail = ctl->tail;

    const int tid = threadIdx.x, lane = tid & 31;
    const int tiles_x = (rp.nx + kSTile - 1) / kSTile;
    const int x0 = (blockIdx.x % tiles_x) * kSTile, y0 = (blockIdx.x / tiles_x) * kSTile;
    const int w_in = min(kSTile, rp.nx - x0), h_in = min(kSTile, rp.ny - y0);
    const int n_pix = w_in * h_in;  // > 0 for every launched block
    const bool full_tile = n_pix == kSPix;
    const uint32_t total = (uint32_t)n_pix * (uint32_t)rp.spp_count;  // < 2^31 (host check: spp_count <= 2^23)

    // pixel slot -> pixel: full tiles use an 8x4-blocked order so that 32 consecutive items are a compact tile
    auto slot_pixel = [&](uint32_t sl, int& pi, int& pj) {
        if (full_tile) {
            pi = x0 + (int)((sl & 7u) | ((sl >> 2) & 8u));
            pj = y0 + (int)(((sl >> 3) & 3u) | ((sl >> 4) & 12u));
        } else {
            const int row = (int)sl / w_in;
            pj = y0 + row, pi = x0 + (int)sl - row * w_in;
        }
    };

    for (int k = tid; k < 3 * kSPix; k += BS) fxlo[k] = 0u, fxhi[k] = 0u, fl[k] = 0.0f;
    for (int k = tid; k < kSPix; k += BS) prays[k] = 0u;
#ifdef RTW_SORTED_CHECK
    for (int k = tid; k < kSPool; k += BS) own[k] = 0, push_idx[k] = k, push_cls[k] = C_REGEN | (99 << 8), take_idx[k] = -1;
#endif
    for (int k = tid; k < C_QUEUES * kQCap; k += BS) queue[k] = kQEmpty;
    __syncthreads();
    for (int s = tid; s < kSPool; s += BS) queue[C_REGEN * kQCap + s] = (uint16_t)s;  // every slot needs a camera ray
    if (tid < 8) ctl->head[tid] = 0, ctl->tail[tid] = (tid == C_REGEN) ? kSPool : 0;
    if (tid == 8) ctl->item_next = 0, ctl->live = kSPool, ctl->mode = 1;
    __syncthreads();

    Rng rng;
    rng.rk = rp.rk;
    const int pref = (tid >> 5) % C_TERM;  // the class this warp looks at first: one of the four scatter classes
    // entry `pos` of class ring `cls`, claimed by this warp: wait for its producer, take it, leave the position empty
    auto take = [&](int cls, unsigned int pos) -> int {
        const int v = ring_take((uint32_t)__cvta_generic_to_shared(const_cast<uint16_t*>(queue + cls * kQCap + (pos & (kQCap - 1)))),
                                ((pos / (unsigned int)kQCap) & 15u) << 12);
#ifdef RTW_SORTED_CHECK
        {
            const int prev = atomicExch(&own[v], 1 + (tid >> 5));
            if (prev != 0 && atomicAdd(&g_check_first, 1) < 40)
                printf("block %d warp %d: slot %d taken at class %d index %u while held by warp %d; last queued at class %d index %d by warp %d, last taken at index %d; "
                       "ring head %u tail %u\n",
                       blockIdx.x, tid >> 5, v, cls, pos, prev - 1, push_cls[v] & 255, push_idx[v], push_cls[v] >> 8, take_idx[v], ctl->head[cls],
                       ctl->tail[cls]);
            take_idx[v] = (int)pos;
        }
#endif
        return v;
    };
    // the lanes of `mask` (all of them call) queue their slot on class `ncls`, one reservation per class present (C_DEAD: the slot retires)
    auto push = [&](unsigned mask, int ncls, int slot) {
#ifdef RTW_SORTED_CHECK
        {
            const int prev = atomicExch(&own[slot], 0);
            if (prev != 1 + (tid >> 5) && atomicAdd(&g_check_first, 1) < 40)
                printf("block %d warp %d: slot %d queued on class %d by a warp that does not hold it (holder %d)\n", blockIdx.x, tid >> 5, slot, ncls, prev - 1);
        }
#endif
        const unsigned same = __match_any_sync(mask, ncls);
        const int leader = __ffs(same) - 1;
        unsigned int pos = 0;
        if (lane == leader) {
            if (ncls != C_DEAD) pos = atomicAdd(&ctl->tail[ncls], (unsigned int)__popc(same));
            else atomicSub(&ctl->live, __popc(same));
        }
        pos = __shfl_sync(mask, pos, leader);
#ifdef RTW_SORTED_CHECK
        if (ncls != C_DEAD) push_idx[slot] = (int)(pos + __popc(same & ((1u << lane) - 1u))), push_cls[slot] = ncls | ((tid >> 5) << 8);
#endif
        if (ncls != C_DEAD) {
            const unsigned int idx = pos + (unsigned int)__popc(same & ((1u << lane) - 1u));
            ring_put((uint32_t)__cvta_generic_to_shared(const_cast<uint16_t*>(queue + ncls * kQCap + (idx & (kQCap - 1)))),
                     (uint16_t)((unsigned int)slot | (((idx / (unsigned int)kQCap) & 15u) << 12)));
        }
    };
    // class queue of a hit: the material's class, or terminal for a miss (and for every hit of a normal-render)
    auto classify = [&](const HitId& hit) -> int {
        int ncls = C_TERM;
        if (hit.prim != -1 && sc.render_type != RTW_RENDER_NORMAL) {
            uint32_t pm;
            const int mat_idx = hit_material<SM>(sc, sv, hit, pm);
            const int mt = ld_mat<SM>(sv, sc.off[SEC_MATS] + (uint32_t)mat_idx * 32u).type;
            // class of material type `mt` from a packed table (one shift instead of a jump table with four divergent
            // arms): nibble mt of kClassOfMat; any other value shades as lambertian, like the switch in shade_hit
            constexpr uint32_t kClassOfMat = (uint32_t)C_LAMB << (4 * RTW_MAT_LAMBERTIAN) | (uint32_t)C_METAL << (4 * RTW_MAT_METAL) |
                                             (uint32_t)C_DIEL << (4 * RTW_MAT_DIELECTRIC) | (uint32_t)C_TERM << (4 * RTW_MAT_DIFFUSE_LIGHT) |
                                             (uint32_t)C_ISO << (4 * RTW_MAT_ISOTROPIC);
            static_assert(C_LAMB == 0, "nibbles of unknown material types must read as lambertian");
            ncls = (int)((kClassOfMat >> (min((uint32_t)mt, 7u) * 4u)) & 7u);
        }
        return ncls;
    };
    for (;;) {
        // ---- claim a unit: 32 entries of a class that has them (each warp starts looking at a different class, which
        // spreads the CAS traffic), else - after a short wait for the warps that are mid-unit - the longest partial run.
        // Lanes 0..5 each look at one queue.
        int c = -1, n = 0;
        unsigned int h = 0;
        int idle_polls = 0;
        for (int spins = 0;;) {
            unsigned int hk = 0;
            int avail = 0;
            if (lane < C_QUEUES) hk = vhead[lane], avail = (int)(vtail[lane] - hk);
            unsigned full = __ballot_sync(0xffffffffu, avail >= 32);
            int k;
            if (full) {
                if constexpr (PH) {
                    // the block's mode: trace units while C_TRACE holds a whole one, then shade units while a class does, and so
                    // on — no barrier: a warp that finds its mode's queues short takes the other kind and turns the mode
                    const int mode = *(volatile int*)&ctl->mode;
                    const unsigned full_t = full & (1u << C_TRACE), full_s = full & ((1u << C_TRACE) - 1u);
                    const bool trace = mode == 0 ? full_t != 0 : full_s == 0;
                    if (trace != (mode == 0) && lane == 0) *(volatile int*)&ctl->mode = trace ? 0 : 1;
                    full = trace ? full_t : full_s;
                }
                const unsigned from_pref = full & ~((1u << pref) - 1u);
                k = __ffs(from_pref ? from_pref : full) - 1;
                n = 32;
            } else {
                const unsigned best = __reduce_max_sync(0xffffffffu, (unsigned)(max(avail, 0) << 3) | (unsigned)(lane & 7));
                if ((best >> 3) == 0) {  // nothing queued anywhere
                    if (*(volatile int*)&ctl->live <= 0) break;
                    // watchdog: paths in flight but nothing to do for > 1.6 s (a tile's tail is milliseconds) means a slot was
                    // lost; leave with an error flag (ray_counter[1], read by rtw_render_device_finish) instead of polling forever
                    if (++idle_polls > (1 << 23)) {
                        if (lane == 0) atomicExch(ray_counter + 1, 1ull);
                        break;
                    }
                    __nanosleep(200);  // (doubling the wait up to 1.6 or 6.4 us while nothing turns up changes nothing: r2_ab_phased.txt)
                    continue;
                }
                if (spins < 2) {
                    spins++;
                    __nanosleep(300);
                    continue;
                }
                k = (int)(best & 7u), n = (int)(best >> 3);
            }
            h = __shfl_sync(0xffffffffu, hk, k);
            int ok = 0;
            if (lane == 0) ok = atomicCAS(&ctl->head[k], h, h + (unsigned int)n) == h;
            if (__shfl_sync(0xffffffffu, ok, 0)) {
                c = k;
                break;
            }
        }
        if (c < 0) break;
        const bool valid = lane < n;
        const unsigned vmask = __ballot_sync(0xffffffffu, valid);
        if constexpr (PH) {
            if (c == C_TRACE) {  // ---- trace unit: closest_hit for the claimed rays, then the class queue of what they hit
                int slot = 0, ncls = C_TERM;
                if (valid) {
                    slot = take(C_TRACE, h + (unsigned int)lane);
                    __threadfence_block();
                    const float4 a = stA[slot], b = stB[slot];
                    const uint4 d = stD[slot];
                    rng.pixel = d.y, rng.sample = d.z;
                    const HitId hit = closest_hit<F, SM, false>(sc, sv, mk(a.x, a.y, a.z), mk(b.x, b.y, b.z), b.w, 0.001f, FLT_MAX, &rng, d.w & 0xffffu,
                                                                stack, BS);
                    ncls = classify(hit);
                    reinterpret_cast<volatile float*>(stA + slot)[3] = hit.t;
                    reinterpret_cast<volatile int*>(stC + slot)[3] = hit.prim;
                    reinterpret_cast<volatile unsigned int*>(stD + slot)[0] = (unsigned int)hit.item;
                    __threadfence_block();  // state before the queue entry
                    push(vmask, ncls, slot);
                }
                __syncwarp();  // the lanes without a ray wait here, not inside the next claim (where they would read the ring counters early)
                continue;
            }
        }
#ifdef RTW_SORTED_STATS
        if (lane == 0 && blockIdx.x == 1000) atomicAdd(&g_stats[c][0], 1), atomicAdd(&g_stats[c][1], n);
#endif
        int slot = 0;
        Ray ray;
        ray.o = ray.d = mk(0.0f, 0.0f, 0.0f);
        ray.time = 0.0f;
        V3 beta = mk(1.0f, 1.0f, 1.0f);
        uint32_t pix = 0, bounce = 0;
        bool alive = false;
        if (valid) {
            slot = take(c, h + (unsigned int)lane);
            __threadfence_block();  // the entry was published after the slot's state
            if (c <= C_TERM) {      // warp-uniform
                const float4 a = stA[slot], b = stB[slot], cc = stC[slot];
                const uint4 d = stD[slot];
                ray.o = mk(a.x, a.y, a.z), ray.d = mk(b.x, b.y, b.z), ray.time = b.w;
                beta = mk(cc.x, cc.y, cc.z);
                HitId hit;
                hit.t = a.w, hit.prim = fi(cc.w), hit.item = (int)d.x;
                rng.pixel = d.y, rng.sample = d.z;
                bounce = d.w & 0xffffu, pix = d.w >> 16;
                V3 L = mk(0.0f, 0.0f, 0.0f);
                alive = shade_hit<F, SM>(sc, sv, rng, bounce, hit, ray, beta, L, c == C_LAMB || c == C_DIEL);
                if (alive && (int)bounce >= rp.max_depth) alive = false;  // :47-48: the next color() returns 0
                if (!alive) {                                             // :235-239
                    // the path's color() invocations = its last bounce index; counted per pixel at the path's end (a per-lane
                    // counter live across the whole loop was the one value this kernel spilled to local memory)
                    atomicAdd(prays + pix, bounce);
                    const float v[3] = {L.x, L.y, L.z};
#pragma unroll
                    for (int ch = 0; ch < 3; ch++) {
                        if (v[ch] != 0.0f) {
                            // below 8192 a pixel's 2^23 samples cannot overflow the 64-bit sum; fireflies up to 2^30 take the
                            // same (exact, commutative) integer path — it would take > 448 of them at 1e9 in ONE pixel of one
                            // launch to wrap; only samples >= 2^30 or not finite go to the float side sum
                            if (fabsf(v[ch]) < 1073741824.0f) {
                                // two native 32-bit ATOMS.ADD instead of a 64-bit CAS loop: the carry out of the low word (and a
                                // sample >= 256) goes to the high word; integer adds commute, so the sum stays exact and order-free
                                const long long fx = __float2ll_rn(v[ch] * 16777216.0f);
                                const unsigned int lo = (unsigned int)fx;
                                unsigned int hi = (unsigned int)((unsigned long long)fx >> 32);
                                const unsigned int old = atomicAdd(fxlo + ch * kSPix + pix, lo);
                                hi += (old + lo < old) ? 1u : 0u;
                                if (hi) atomicAdd(fxhi + ch * kSPix + pix, hi);
                            } else {
                                atomicAdd(fl + ch * kSPix + pix, v[ch]);
                            }
                        }
                    }
                }
            }
        }
        __syncwarp();       // reconverge the lanes that added radiance with those that did not
        if (c >= C_TERM) {  // warp-uniform: every valid lane needs a new path
            unsigned int item = 0;
            if (lane == 0) item = atomicAdd(&ctl->item_next, (unsigned int)n);
            item = __shfl_sync(0xffffffffu, item, 0) + lane;
            if (valid && item < total) {
                uint32_t sm;
                if (full_tile) sm = item >> 8, pix = item & 255u;
                else sm = item / (uint32_t)n_pix, pix = item - sm * (uint32_t)n_pix;
                int pi, pj;
                slot_pixel(pix, pi, pj);
                rng.pixel = (uint32_t)(pj * rp.nx + pi);
                rng.sample = (uint32_t)rp.spp_begin + sm;  // global sample index: shards are disjoint streams
                if (rp.flags & RTW_FLAG_DEBUG_RAY) pi = rp.nx / 2, pj = rp.ny / 2;  // RayTracingWeekend.cpp:221-225
                ray = camera_ray(sc, rp, rng, pi, pj);
                beta = mk(1.0f, 1.0f, 1.0f);
                bounce = 0;
                alive = true;
            }
        }
        __syncwarp();
        int ncls = (c >= C_TERM) ? C_DEAD : C_REGEN;  // a terminal/regen lane only stays dead when the pool is empty
        if (alive) {
            bounce++;
            HitId hit;
            hit.t = 0.0f, hit.prim = -1, hit.item = -1;
            if constexpr (PH) {
                ncls = C_TRACE;  // a trace unit finds the hit in the next phase
            } else {
                hit = closest_hit<F, SM, false>(sc, sv, ray.o, ray.d, ray.time, 0.001f, FLT_MAX, &rng, bounce, stack, BS);
                ncls = classify(hit);
            }
            stA[slot] = make_float4(ray.o.x, ray.o.y, ray.o.z, hit.t);
            stB[slot] = make_float4(ray.d.x, ray.d.y, ray.d.z, ray.time);
            stC[slot] = make_float4(beta.x, beta.y, beta.z, __int_as_float(hit.prim));
            stD[slot] = make_uint4((uint32_t)hit.item, rng.pixel, rng.sample, bounce | (pix << 16));
            __threadfence_block();  // state before the queue entry
        }
        if (valid) push(vmask, ncls, slot);
    }
#ifdef RTW_SORTED_CHECK
    if (lane == 0 && *(volatile int*)&ctl->live != 0 && atomicAdd(&g_check_first, 1) < 40) printf("block %d warp %d leaves with live = %d\n", blockIdx.x, tid >> 5, *(volatile int*)&ctl->live);
#endif
    __syncthreads();
#ifdef RTW_SORTED_CHECK
    if (tid == 0) {
        int left = 0;
        for (int k = 0; k < C_QUEUES; k++) left += (int)(ctl->tail[k] - ctl->head[k]);
        if ((left != 0 || ctl->item_next < total) && atomicAdd(&g_check_first, 1) < 40) printf("block %d ends with %d queued entries, item_next %u of %u\n", blockIdx.x, left, ctl->item_next, total);
    }
#endif
#ifdef RTW_SORTED_STATS
    if (tid == 0 && blockIdx.x == 1000)
        for (int k = 0; k < C_QUEUES; k++) printf("class %d: units %d lanes %d\n", k, g_stats[k][0], g_stats[k][1]);
#endif

    // one 16-byte store per pixel (canvas[j*nx+i], RayTracingWeekend.cpp:247)
    for (int p = tid; p < n_pix; p += BS) {
        int pi, pj;
        slot_pixel((uint32_t)p, pi, pj);
        const size_t idx = (size_t)pj * rp.nx + pi;
        float4 out;
        out.x = (float)((double)(long long)(((unsigned long long)fxhi[p] << 32) | fxlo[p]) * (1.0 / 16777216.0)) + fl[p];
        out.y = (float)((double)(long long)(((unsigned long long)fxhi[kSPix + p] << 32) | fxlo[kSPix + p]) * (1.0 / 16777216.0)) + fl[kSPix + p];
        out.z = (float)((double)(long long)(((unsigned long long)fxhi[2 * kSPix + p] << 32) | fxlo[2 * kSPix + p]) * (1.0 / 16777216.0)) + fl[2 * kSPix + p];
        out.w = (float)rp.spp_count;
        if (rp.flags & RTW_FLAG_ACCUMULATE) {
            const float4 prev = accum[idx];
            out.x += prev.x, out.y += prev.y, out.z += prev.z, out.w += prev.w;
        }
        accum[idx] = out;
    }
    unsigned long long warp_rays = 0;
    for (int p = tid; p < kSPix; p += BS) warp_rays += prays[p];
    for (int o = 16; o > 0; o >>= 1) warp_rays += __shfl_down_sync(0xffffffffu, warp_rays, o);
    if (lane == 0 && warp_rays) atomicAdd(ray_counter, warp_rays);
}

namespace {
template <uint32_t F, int SM>
cudaError_t launch_sorted_t(const DevScene& sc, const DevRenderParams& rp, float4* accum, unsigned long long* counter,
                            size_t smem, cudaStream_t st) {
    // threads per block, pool slots per 256 threads, resident blocks per SM — one measured triple per variant class (header
    // comment), written BS * 1000 + PT * 10 + MB so that -DRTW_CFG_x=... can override a class for A/B builds (tools/ab_libs.sh;
    // nvcc splits -D values at commas)
#ifndef RTW_CFG_CORNELL
#define RTW_CFG_CORNELL 512042
#endif
#ifndef RTW_CFG_TEX
#define RTW_CFG_TEX 768061
#endif
#ifndef RTW_CFG_BVH_UNSTAGED
#define RTW_CFG_BVH_UNSTAGED 256022
#endif
#ifndef RTW_CFG_BVH_STAGED
#define RTW_CFG_BVH_STAGED 256023
#endif
#ifndef RTW_CFG_SMOKE
#define RTW_CFG_SMOKE RTW_CFG_CORNELL
#endif
    constexpr int cfg = (F == kVariants[0]) ? RTW_CFG_CORNELL : (F == (F_XF | F_MEDIA)) ? RTW_CFG_SMOKE : ((F & F_BVH) ? (SM < 2 ? RTW_CFG_BVH_UNSTAGED : RTW_CFG_BVH_STAGED) : RTW_CFG_TEX);
    constexpr int BS = cfg / 1000, PT = cfg / 10 % 100, MB = cfg % 10;
    static_assert(BS % 32 == 0 && BS <= 1024 && pool_slots(PT) <= 2048, "block size / pool out of range (ring entries hold 11-bit slot ids)");
    // `smem` (staged bytes + traversal stacks) was planned for kSBlock threads: the stacks scale with the block size
    const size_t staged = SM == 2 ? sc.all_bytes : (SM == 1 ? sc.hot_bytes : 0);
    const int tiles_x = (rp.nx + kSTile - 1) / kSTile, tiles_y = (rp.ny + kSTile - 1) / kSTile;
    static const bool debug = getenv("RTW_DEBUG_TREES") != nullptr;  // read once per process
    auto launch = [&](auto kernel, int bs, int pt, int mb, bool ph) -> cudaError_t {
        // `smem` (staged bytes + traversal stacks) was planned for kSBlock threads: the stacks scale with the block size
        const size_t sm_bytes = staged + (smem - staged) / kSBlock * bs;
        const uint32_t pool_off = (uint32_t)((sm_bytes + 15) & ~(size_t)15);
        const size_t bytes = pool_off + sorted_pool_bytes(pt, ph);
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
        if (e != cudaSuccess) return e;
        if (debug) {
            int nb = 0;
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, bs, bytes);
            fprintf(stderr, "rtw render_sorted<%u,%d,%d,%d,%d,%d>: %zu bytes of shared memory per block (stack %zu), %d blocks/SM\n", F, SM, pt,
                    mb, bs, (int)ph, bytes, sm_bytes - staged, nb);
        }
        kernel<<<tiles_x * tiles_y, bs, bytes, st>>>(sc, rp, accum, counter, pool_off);
        return cudaGetLastError();
    };
    if constexpr (phased_variant(F, SM)) {
#ifndef RTW_CFG_PHASED
#define RTW_CFG_PHASED 512041
#endif
        constexpr int PBS = RTW_CFG_PHASED / 1000, PPT = RTW_CFG_PHASED / 10 % 100, PMB = RTW_CFG_PHASED % 10;
        static const bool phased = [] { const char* v = getenv("RTW_PHASED"); return !v || atoi(v) != 0; }();  // RTW_PHASED=0: the plain kernel (A/B)
        // its 512 stack columns and 1024 slots must fit the block's shared memory (deep trees: the plain kernel has half the columns)
        static const size_t smem_max = [] {
            int dev = 0, v = 0;
            cudaGetDevice(&dev);
            cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
            return (size_t)v;
        }();
        const size_t need = ((staged + (smem - staged) / kSBlock * PBS + 15) & ~(size_t)15) + sorted_pool_bytes(PPT, true);
        if (phased && need <= smem_max) return launch(render_sorted<F, SM, PPT, PMB, PBS, true>, PBS, PPT, PMB, true);
    }
    return launch(render_sorted<F, SM, PT, MB, BS, false>, BS, PT, MB, false);
}
}  // namespace


}  // namespace rtw
