// sorted.cu — instantiations and dispatch of the material-sorted megakernel (sorted_kernel.cuh).
#include "sorted_kernel.cuh"

namespace rtw {

// spp * max_depth < 2^32: the per-pixel count of color() invocations is 32 bits
bool sorted_supports(const DevRenderParams& rp) {
    return rp.spp_count <= (1 << 23) && rp.max_depth < (1 << 16) && (unsigned long long)rp.spp_count * (unsigned long long)rp.max_depth < (1ull << 32);
}

#ifdef RTW_COUNT_EVENTS
// diagnostics build: read and clear the traversal event counters of the render kernel (tools/count_tree_events.py)
extern "C" int rtw_debug_events(unsigned long long* out, int n, int clear) {
    unsigned long long h[8] = {};
    if (cudaMemcpyFromSymbol(h, g_evc, sizeof h) != cudaSuccess) return -1;
    for (int k = 0; k < n && k < 8; k++) out[k] = h[k];
    if (clear) {
        unsigned long long z[8] = {};
        if (cudaMemcpyToSymbol(g_evc, z, sizeof z) != cudaSuccess) return -1;
    }
    return EVC_N;
}
#endif

// The all-features variants on unstaged scenes (Book-2 final scene) are bound by the SM instruction cache: they
// are instantiated in sorted_big.cu, which compiles the same template in compact mode (RTW_COMPACT: one out-of-line copy of the
// Philox rounds instead of one per draw site — +6 % there, -9...-15 % on the variants that fit the cache, which stay here).
cudaError_t launch_render_sorted_big(const DevScene& sc, const DevRenderParams& rp, float4* accum, unsigned long long* counter, int stage,
                                     size_t smem, cudaStream_t st);

cudaError_t launch_render_sorted(const DevScene& sc, const DevRenderParams& rp, float4* accum, unsigned long long* counter,
                                 cudaStream_t st) {
#ifdef RTW_COUNT_EVENTS
    RTW_DISPATCH(kSBlock, launch_sorted_t, sc, rp, accum, counter);  // diagnostics build: one translation unit, one set of counters
#else
    Launch l = plan(sc, kSBlock);
    // Staged scenes with media but without trees, textures, moving or huge spheres (Cornell smoke, hittable.h:430-479 over box
    // boundaries): a variant of their own, for this kernel only.  The all-features kernel they would otherwise run touches far
    // more code than the SM instruction cache holds (76.9 % hit rate).
    constexpr uint32_t kSmoke = F_XF | F_MEDIA;
    static const bool smoke_variant = [] { const char* v = getenv("RTW_SMOKE_VARIANT"); return !v || atoi(v) != 0; }();  // =0: A/B
    if (smoke_variant && l.stage == 2 && (sc.features & F_MEDIA) && (sc.features & ~kSmoke) == 0)
        return launch_sorted_t<kSmoke, 2>(sc, rp, accum, counter, sc.all_bytes, st);
    if (l.stage == 2) {
        if (l.variant == kVariants[0]) return launch_sorted_t<kVariants[0], 2>(sc, rp, accum, counter, l.smem, st);
        if (l.variant == kVariants[1]) return launch_sorted_t<kVariants[1], 2>(sc, rp, accum, counter, l.smem, st);
        if (l.variant == kVariants[2]) return launch_sorted_t<kVariants[2], 2>(sc, rp, accum, counter, l.smem, st);
        return launch_sorted_t<F_ALL, 2>(sc, rp, accum, counter, l.smem, st);
    }
    if (l.variant == kVariants[1]) {
        if (l.stage == 1) return launch_sorted_t<kVariants[1], 1>(sc, rp, accum, counter, l.smem, st);
        return launch_sorted_t<kVariants[1], 0>(sc, rp, accum, counter, l.smem, st);
    }
    return launch_render_sorted_big(sc, rp, accum, counter, l.stage, l.smem, st);
#endif
}

}  // namespace rtw
