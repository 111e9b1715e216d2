// sorted_big.cu — the all-features variants of the material-sorted megakernel on unstaged scenes (<F_ALL, 0> and <F_ALL, 1>:
// the Book-2 final scene), compiled in COMPACT mode.  These kernels are ~3 700 SASS instructions and run at a 91 % SM
// instruction-cache hit rate with `no_inst` as a top stall; every instruction of code they lose is worth more than
// the call it costs: with ONE out-of-line copy of the ten Philox rounds (six inlined draw sites before) the Book-2 final scene
// renders 6 % faster (profiles/r2_ab_philox_ool.txt), while the variants that fit the cache lose 3-15 % to the same change —
// hence a translation unit of their own.
#ifndef RTW_COUNT_EVENTS
#define RTW_COMPACT 1
#define RTW_PHILOX_OOL 1
#define RTW_PHASED 1
#include "sorted_kernel.cuh"

namespace rtw {

cudaError_t launch_render_sorted_big(const DevScene& sc, const DevRenderParams& rp, float4* accum, unsigned long long* counter, int stage,
                                     size_t smem, cudaStream_t st) {
    if (stage == 1) return launch_sorted_t<F_ALL, 1>(sc, rp, accum, counter, smem, st);
    return launch_sorted_t<F_ALL, 0>(sc, rp, accum, counter, smem, st);
}

}  // namespace rtw
#endif
