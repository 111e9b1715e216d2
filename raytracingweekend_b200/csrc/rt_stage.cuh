// rt_stage.cuh — shared-memory staging of the scene blob and the kernel-variant dispatch shared by kernels.cu and
// wavefront.cu.
#pragma once
#include <cstdlib>
#include "rt_path.cuh"

namespace rtw {

// Stage a prefix of the scene blob into shared memory (16-byte vectors, whole block) and return the view.
// `sbase` is laundered through an empty volatile asm after the barrier: every staged load depends on it, so none can
// be scheduled above the barrier, and the compiler keeps it in a register instead of re-deriving the shared window
// address at each use.
template <int SM>
RTW_DEV SceneView stage_scene(const DevScene& sc, uint4* smem) {
    const uint32_t staged = SM == 2 ? sc.all_bytes : (SM == 1 ? sc.hot_bytes : 0u);
    if (SM > 0) {
        const uint4* src = reinterpret_cast<const uint4*>(sc.blob);
        for (uint32_t i = threadIdx.x; i < staged / 16; i += blockDim.x) smem[i] = __ldg(src + i);
        __syncthreads();
    }
    SceneView v;
    v.sbase = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("" : "+r"(v.sbase)::"memory");
    v.g = sc.blob;
    v.texels = sc.texels;
    return v;
}
// dynamic shared memory: [staged blob prefix][BVH traversal stacks]
RTW_DEV int* stack_base(const DevScene& sc, int sm_level, uint4* smem) {
    uint32_t staged = sm_level == 2 ? sc.all_bytes : (sm_level == 1 ? sc.hot_bytes : 0u);
    return reinterpret_cast<int*>(reinterpret_cast<uint8_t*>(smem) + staged);
}

}  // namespace rtw

namespace rtw {
// Kernel specialisations: the smallest feature superset is chosen per scene (registers and code size shrink
// a lot when BVH / media / textures are compiled out, e.g. for the Cornell box), and the staging level from the
// blob size: everything in shared memory (<= kStageMax), the hot sections only, or nothing.
constexpr uint32_t kStageMax = 24 * 1024;  // per block; keeps >= 8 blocks of 128 threads resident per SM

constexpr uint32_t kVariants[] = {
    F_XF,                                       // Cornell box, dielectric_scene
    F_XF | F_MOVING | F_PRECISE | F_BVH,        // random_balls_scene (485-sphere list under a tree built at upload)
    F_XF | F_TEXTURES | F_PRECISE,              // light_sample, two_perlin_spheres_earth
    F_ALL};

inline uint32_t pick_variant_impl(uint32_t features) {
    for (uint32_t v : kVariants)
        if ((features & ~v) == 0) return v;
    return F_ALL;
}
inline int pick_stage(const DevScene& sc) {
    if (sc.all_bytes <= kStageMax) return 2;
    if (sc.hot_bytes <= kStageMax) return 1;
    return 0;
}

struct Launch {
    uint32_t variant;
    int stage;
    size_t smem;
};
// `block` = threads per block of the kernel being launched (sizes the per-thread BVH stack columns); `exact` = the kernel
// walks the reference-topology binary trees (intersect_batch), else the render trees.
inline Launch plan(const DevScene& sc, int block, bool exact = false) {
    Launch l;
    l.stage = pick_stage(sc);
    l.variant = pick_variant_impl(sc.features);
    // instantiated combinations: every variant at stage 2; {moving, all} at stages 1 and 0
    if (l.stage < 2 && l.variant != kVariants[1]) l.variant = F_ALL;
    size_t staged = l.stage == 2 ? sc.all_bytes : (l.stage == 1 ? sc.hot_bytes : 0);
    // exact needs computed at upload (api.cu push_need), + 1 entry of slack
    int entries = 1 + (exact ? sc.stack_exact : (bvh4_variant(l.variant) ? sc.bvh4_stack : sc.stack_render));
    static const int extra = [] { const char* v = getenv("RTW_STACK_EXTRA"); return v ? atoi(v) : 0; }();  // diagnostics: cost of a deeper stack
    entries += extra;
    size_t stack = (l.variant & F_BVH) ? (size_t)entries * block * sizeof(int) : 0;
    l.smem = staged + stack;
    return l;
}

}  // namespace rtw

#define RTW_DISPATCH(BLOCK, FN, ...) RTW_DISPATCH_X(false, BLOCK, FN, __VA_ARGS__)
#define RTW_DISPATCH_X(EXACT, BLOCK, FN, ...)                                                      \
    do {                                                                                           \
        Launch l = plan(sc, BLOCK, EXACT);                                                         \
        if (l.stage == 2) {                                                                        \
            if (l.variant == kVariants[0]) return FN<kVariants[0], 2>(__VA_ARGS__, l.smem, st);    \
            if (l.variant == kVariants[1]) return FN<kVariants[1], 2>(__VA_ARGS__, l.smem, st);    \
            if (l.variant == kVariants[2]) return FN<kVariants[2], 2>(__VA_ARGS__, l.smem, st);    \
            return FN<F_ALL, 2>(__VA_ARGS__, l.smem, st);                                          \
        }                                                                                          \
        if (l.stage == 1) {                                                                        \
            if (l.variant == kVariants[1]) return FN<kVariants[1], 1>(__VA_ARGS__, l.smem, st);    \
            return FN<F_ALL, 1>(__VA_ARGS__, l.smem, st);                                          \
        }                                                                                          \
        if (l.variant == kVariants[1]) return FN<kVariants[1], 0>(__VA_ARGS__, l.smem, st);        \
        return FN<F_ALL, 0>(__VA_ARGS__, l.smem, st);                                              \
    } while (0)
