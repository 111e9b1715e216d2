// rt_trace.cuh — world.hit(): list walk, bvh_node traversal, constant_medium, hit completion.
#pragma once
#include "rt_device.cuh"

namespace rtw {

#ifndef RTW_BVH4
#define RTW_BVH4 1  // render kernels walk the 4-wide tree where it was measured faster (A/B switch: 0 = never, 2 = always)
#endif
#ifndef RTW_UNIFIED_BVH
#define RTW_UNIFIED_BVH 1  // binary-tree render variants walk all bvh_node items of the world in one loop (A/B switch)
#endif
#ifndef RTW_BVH4_OCTANT
#define RTW_BVH4_OCTANT 1
#endif
#ifndef RTW_BVH4_SORT
#define RTW_BVH4_SORT 0  // 1: full distance sort of the hit children; 0: nearest first, the others in slot order (+2.5 %)
#endif
#ifndef RTW_FAST_BOX
#define RTW_FAST_BOX 0  // min/max slab test on the render path: measured 1.2-1.6x SLOWER than aabb_hit (A/B switch, off)
#endif

// bvh_node::hit (hittable.h:97-129, as intended — SURVEY F1).  A node is entered iff its box passes aabb::hit with the
// caller's [tmin, tmax]; both children are then examined; the smaller t wins and on equal t the right (later) subtree
// wins (:107).  Leaves are lists (a box is six rects) evaluated with list semantics against the caller's tmax.
//
// Device form (DevNode, device_scene.hpp): both child boxes are tested at the parent; the nearer child is descended
// first and the farther one pushed; the interval used for box culling shrinks to the best hit found so far, widened by
// a relative margin that covers the 1/d-multiply vs divide rounding gap, so no candidate with t <= best is ever culled.
// Because the visiting order is no longer left-to-right, the reference's tie rule is applied explicitly: among equal t
// the leaf further right in the tree (larger leaf offset) wins.
// The loop is "while-while": every lane first descends inner nodes until it holds a leaf, then all lanes of the warp
// that hold one test their leaves together, so node steps and primitive tests do not serialise against each other.
// Stack: shared memory, one column per thread (`stack[level * stride]`).
constexpr int kBvhDone = (int)0x80000000;  // not a valid reference: ~0x7fffffff is never a leaf offset
template <uint32_t F, int SM, bool SHRINK, bool FAST_BOX, bool BOXREC>
RTW_DEV void bvh_hit(const SceneView& sv, int root, const RayK& rk, float time, float tmin, float tmax, int* stack,
                     int stride, float& t_out, int& prim_out) {
    const V3 o = rk.o, inv = rk.inv;
    const V3 oi = mk(o.x * inv.x, o.y * inv.y, o.z * inv.z);  // FAST_BOX: (b - o) * inv as fma(b, inv, -o*inv)
    float best = tmax;
    int best_prim = -1, best_leaf = 0;
    int sp = 0;
    int ref = root;
    for (;;) {
        while (ref >= 0) {  // inner node: test both child boxes
            const float4 llo = hot4<SM>(sv, (uint32_t)ref), lhi = hot4<SM>(sv, (uint32_t)ref + 16);
            const float4 rlo = hot4<SM>(sv, (uint32_t)ref + 32), rhi = hot4<SM>(sv, (uint32_t)ref + 48);
            const float cull = (SHRINK && best_prim >= 0) ? fmaf(best, 4e-6f, best) : tmax;
            RTW_EV(EVC_AABB, 2);
            float tl, tr;
            const bool hl = FAST_BOX ? slab_hit(llo, lhi, inv, oi, tmin, cull, tl) : aabb_hit(llo, lhi, o, inv, tmin, cull, tl);
            const bool hr = FAST_BOX ? slab_hit(rlo, rhi, inv, oi, tmin, cull, tr) : aabb_hit(rlo, rhi, o, inv, tmin, cull, tr);
            const int l = fi(llo.w), r = fi(lhi.w);
            if (hl && hr) {
                const bool left_first = !(tr < tl);
                stack[sp * stride] = left_first ? r : l;
                sp++;
                ref = left_first ? l : r;
            } else if (hl || hr) {
                ref = hl ? l : r;
            } else if (sp > 0) {
                sp--;
                ref = stack[sp * stride];
            } else {
                ref = kBvhDone;
            }
        }
        if (ref == kBvhDone) break;
        {  // leaf: list semantics against the caller's tmax
            const uint2 lf = hot2<SM>(sv, (uint32_t)~ref);
            float t_leaf = tmax;
            int p_leaf = -1;
            test_groups<F, SM, BOXREC>(sv, lf.x, lf.y, rk, time, tmin, t_leaf, p_leaf);
            const int leaf = ~ref;
            if (p_leaf >= 0 && (best_prim < 0 || t_leaf < best || (t_leaf == best && leaf >= best_leaf))) {
                best = t_leaf;
                best_prim = p_leaf;
                best_leaf = leaf;
            }
            if (sp > 0) {
                sp--;
                ref = stack[sp * stride];
            } else {
                ref = kBvhDone;
            }
        }
        if (ref == kBvhDone) break;
    }
    if (best_prim >= 0) {
        t_out = best;
        prim_out = best_prim;
    }
}

// Which kernel variants walk the 4-wide tree.  Measured on B200 (tools/ab_libs.sh): random_balls_scene 2.27 -> 2.49
// Gpaths/s (node visits x0.57), but the all-features kernels lose 4-7 % (Book-2 final 0.452 -> 0.422, and Cornell smoke,
// which has no tree at all, 3.96 -> 3.80: those kernels are bound by instruction fetch and by the longest lane of a warp,
// not by the number of node fetches), so they keep the binary tree.
__host__ __device__ constexpr bool bvh4_variant(uint32_t f) { return RTW_BVH4 == 2 || (RTW_BVH4 == 1 && (f & F_BVH) && !(f & F_MEDIA)); }

// The same search over the 4-wide render tree (DevNode4, device_scene.hpp): one 128-byte node decides four children.
// The near and far plane of each slab are picked by the sign of the ray direction (three per-ray address offsets), so a
// child costs 6 FFMA + 4 min/max + 1 compare; fminf/fmaxf drop a NaN plane (0 * inf, inf - inf), which only widens the
// interval — the boxes are padded bounds of what the leaves test exactly, so widening never changes a result.  The hit
// children are ordered by entry distance (comparator network, RTW_BVH4_SORT); the nearest is descended, the others pushed
// farthest first.
template <uint32_t F, int SM, bool BOXREC>
RTW_DEV void bvh4_hit(const SceneView& sv, int root, const RayK& rk, float time, float tmin, float tmax, int* stack, int stride,
                      float& t_out, int& prim_out) {
    const V3 inv = rk.inv;
    const V3 oi = mk(rk.o.x * inv.x, rk.o.y * inv.y, rk.o.z * inv.z);
#if RTW_BVH4_OCTANT
    const uint32_t nx = inv.x < 0.0f ? 16u : 0u, ny = inv.y < 0.0f ? 48u : 32u, nz = inv.z < 0.0f ? 80u : 64u;  // near planes
#endif
    float best = tmax;
    int best_prim = -1, best_leaf = 0;
    int sp = 0;
    int ref = root;
    for (;;) {
        while (ref >= 0) {
            const uint32_t a = (uint32_t)ref;
#if RTW_BVH4_OCTANT
            const float4 nxv = hot4<SM>(sv, a + nx), fxv = hot4<SM>(sv, a + (nx ^ 16u));
            const float4 nyv = hot4<SM>(sv, a + ny), fyv = hot4<SM>(sv, a + (ny ^ 16u));
            const float4 nzv = hot4<SM>(sv, a + nz), fzv = hot4<SM>(sv, a + (nz ^ 16u));
#else
            const float4 nxv = hot4<SM>(sv, a), fxv = hot4<SM>(sv, a + 16u);
            const float4 nyv = hot4<SM>(sv, a + 32u), fyv = hot4<SM>(sv, a + 48u);
            const float4 nzv = hot4<SM>(sv, a + 64u), fzv = hot4<SM>(sv, a + 80u);
#endif
            const float4 rf = hot4<SM>(sv, a + 96u);
            const float cull = best_prim >= 0 ? fmaf(best, 4e-6f, best) : tmax;
            RTW_EV(EVC_AABB, 4);
            float t0, t1, t2, t3;
#if RTW_BVH4_OCTANT
#define RTW_SLAB4(T, C)                                                                                                       \
    {                                                                                                                         \
        const float tn = fmaxf(fmaxf(fmaf(nxv.C, inv.x, -oi.x), fmaf(nyv.C, inv.y, -oi.y)), fmaxf(fmaf(nzv.C, inv.z, -oi.z), tmin)); \
        const float tf = fminf(fminf(fmaf(fxv.C, inv.x, -oi.x), fmaf(fyv.C, inv.y, -oi.y)), fminf(fmaf(fzv.C, inv.z, -oi.z), cull)); \
        T = tn <= tf ? tn : INFINITY;                                                                                         \
    }
#else
#define RTW_SLAB4(T, C)                                                                                                       \
    {                                                                                                                         \
        const float x0 = fmaf(nxv.C, inv.x, -oi.x), x1 = fmaf(fxv.C, inv.x, -oi.x);                                           \
        const float y0 = fmaf(nyv.C, inv.y, -oi.y), y1 = fmaf(fyv.C, inv.y, -oi.y);                                           \
        const float z0 = fmaf(nzv.C, inv.z, -oi.z), z1 = fmaf(fzv.C, inv.z, -oi.z);                                           \
        const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), tmin));                              \
        const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), cull));                              \
        T = tn <= tf ? tn : INFINITY;                                                                                         \
    }
#endif
            RTW_SLAB4(t0, x)
            RTW_SLAB4(t1, y)
            RTW_SLAB4(t2, z)
            RTW_SLAB4(t3, w)
#undef RTW_SLAB4
            int r0 = fi(rf.x), r1 = fi(rf.y), r2 = fi(rf.z), r3 = fi(rf.w);
#define RTW_CSWAP(TA, RA, TB, RB)            \
    {                                        \
        const bool s = TB < TA;              \
        const float tt = s ? TB : TA;        \
        TB = s ? TA : TB, TA = tt;           \
        const int rr = s ? RB : RA;          \
        RB = s ? RA : RB, RA = rr;           \
    }
            RTW_CSWAP(t0, r0, t1, r1)
            RTW_CSWAP(t2, r2, t3, r3)
            RTW_CSWAP(t0, r0, t2, r2)
#if RTW_BVH4_SORT
            RTW_CSWAP(t1, r1, t3, r3)
            RTW_CSWAP(t1, r1, t2, r2)
#endif
#undef RTW_CSWAP
            if (t0 < INFINITY) {
                if (t3 < INFINITY) stack[sp * stride] = r3, sp++;
                if (t2 < INFINITY) stack[sp * stride] = r2, sp++;
                if (t1 < INFINITY) stack[sp * stride] = r1, sp++;
                ref = r0;
            } else if (sp > 0) {
                sp--;
                ref = stack[sp * stride];
            } else {
                ref = kBvhDone;
            }
        }
        if (ref == kBvhDone) break;
        {  // leaf: list semantics against the caller's tmax, as in bvh_hit
            const uint2 lf = hot2<SM>(sv, (uint32_t)~ref);
            float t_leaf = tmax;
            int p_leaf = -1;
            test_groups<F, SM, BOXREC>(sv, lf.x, lf.y, rk, time, tmin, t_leaf, p_leaf);
            const int leaf = ~ref;
            if (p_leaf >= 0 && (best_prim < 0 || t_leaf < best || (t_leaf == best && leaf >= best_leaf))) {
                best = t_leaf;
                best_prim = p_leaf;
                best_leaf = leaf;
            }
            if (sp > 0) {
                sp--;
                ref = stack[sp * stride];
            } else {
                ref = kBvhDone;
            }
        }
        if (ref == kBvhDone) break;
    }
    if (best_prim >= 0) {
        t_out = best;
        prim_out = best_prim;
    }
}

// All bvh_node items of the world in ONE traversal loop (render kernels of the binary-tree variants).  closest_hit defers
// the tree items: it pushes an ENTER marker per item and this loop pops them like any other reference — a lane that
// finishes one tree enters the next (ray into that item's frame) while its neighbours are still inside the previous
// one.  A warp then waits for its longest lane ONCE, max_i(steps_A(i) + steps_B(i)), instead of once per tree,
// max_i steps_A(i) + max_i steps_B(i): on the Book-2 final scene four warps out of five hold a lane or two that enter the
// small sphere-cluster tree, and walked it at 1-2 active lanes after everyone had waited for the ground tree.
// Same boxes, same leaves, same tie rule as bvh_hit; the items see the closest hit of the lists and of each other as
// their cull distance, which can only shorten the walk.  `h` carries the closest hit so far in and the result out.
constexpr int kBvhEnter = (int)0x80000001;  // ENTER marker of item i = kBvhEnter + i; markers and kBvhDone are < -2^30
#ifndef RTW_UNIFIED_INLINE
#define RTW_UNIFIED_INLINE 1  // A/B switch (0 = one out-of-line copy)
#endif
#if RTW_UNIFIED_INLINE
#define RTW_UNIFIED_FN RTW_DEV
#else
#define RTW_UNIFIED_FN __device__ __noinline__
#endif
template <uint32_t F, int SM>
RTW_UNIFIED_FN HitId bvh_hit_items(uint32_t items0, SceneView sv, RayK rk, float time, float tmin, HitId h, int* stack, int stride, int sp) {
    const float tmax = h.t;
    float best = tmax;
    int best_prim = -1, best_leaf = 0, best_item = -1;
    RayK lk = rk;
    int cur_item = -1;
    sp--;
    int ref = stack[sp * stride];
    for (;;) {
        while (ref >= 0) {  // inner node: test both child boxes (as in bvh_hit)
            const float4 llo = hot4<SM>(sv, (uint32_t)ref), lhi = hot4<SM>(sv, (uint32_t)ref + 16);
            const float4 rlo = hot4<SM>(sv, (uint32_t)ref + 32), rhi = hot4<SM>(sv, (uint32_t)ref + 48);
            const float cull = best_prim >= 0 ? fmaf(best, 4e-6f, best) : tmax;
            RTW_EV(EVC_AABB, 2);
            float tl, tr;
            const bool hl = aabb_hit(llo, lhi, lk.o, lk.inv, tmin, cull, tl);
            const bool hr = aabb_hit(rlo, rhi, lk.o, lk.inv, tmin, cull, tr);
            const int l = fi(llo.w), r = fi(lhi.w);
            if (hl && hr) {
                const bool left_first = !(tr < tl);
                stack[sp * stride] = left_first ? r : l;
                sp++;
                ref = left_first ? l : r;
            } else if (hl || hr) {
                ref = hl ? l : r;
            } else if (sp > 0) {
                sp--;
                ref = stack[sp * stride];
            } else {
                ref = kBvhDone;
            }
        }
        if (ref < -0x40000000) {
            if (ref == kBvhDone) break;
            // ENTER: the next tree item
            cur_item = ref - kBvhEnter;
            const float4 i0 = hot4<SM>(sv, items0 + (uint32_t)cur_item * 32u);
            const int ixf = fi(i0.y);
            lk = rk;
            if (ixf >= 0) lk = xf_rayk(ld_xf<SM>(sv, (uint32_t)ixf), rk);
            ref = fi(i0.w);  // DevItem.end: the render tree
            continue;
        }
        {  // leaf: list semantics against the closest hit of the non-tree items
            const uint2 lf = hot2<SM>(sv, (uint32_t)~ref);
            float t_leaf = tmax;
            int p_leaf = -1;
            test_groups<F, SM, true>(sv, lf.x, lf.y, lk, time, tmin, t_leaf, p_leaf);
            const int leaf = ~ref;
            if (p_leaf >= 0 && (best_prim < 0 || t_leaf < best || (t_leaf == best && leaf >= best_leaf))) {
                best = t_leaf;
                best_prim = p_leaf;
                best_leaf = leaf;
                best_item = cur_item;
            }
            if (sp > 0) {
                sp--;
                ref = stack[sp * stride];
            } else {
                ref = kBvhDone;
            }
        }
        if (ref == kBvhDone) break;
    }
    if (best_prim >= 0) {
        h.t = best;
        h.prim = best_prim;
        h.item = best_item;
    }
    return h;
}

// world.hit(r, tmin, tmax) (RayTracingWeekend.cpp:52 -> hittable_list.h:11-37) over the flattened items.
// `rng` == nullptr skips constant_medium items (intersect_batch: media are stochastic).
// EXACT_TIES replays the second pass of hittable_list::hit (hittable_list.h:26-34) for scenes with bvh_node items:
// with tmax == closest the node boxes are re-tested against the final t, which can change which of several
// primitives at exactly that t is reported (coplanar faces of adjacent boxes, seen from inside a box).  Rays of a
// render never start inside closed boxes, so the megakernel skips it; intersect_batch (parity tier 1) runs it.
// which variants walk all tree items of the world in one loop (kUnified in closest_hit)
__host__ __device__ constexpr bool unified_variant(uint32_t f, int sm) { return RTW_UNIFIED_BVH && (f & F_BVH) != 0 && !bvh4_variant(f) && sm < 2; }
template <uint32_t F, int SM, bool EXACT_TIES>
RTW_DEV HitId closest_hit(const DevScene& sc, const SceneView& sv, V3 o, V3 d, float time, float tmin, float tmax,
                          const Rng* rng, uint32_t bounce, int* stack, int stride) {
    HitId h;
    h.t = tmax;
    h.prim = -1;
    h.item = -1;
    int medium_idx = 0;
    const RayK rk = make_rayk(o, d);
    const uint32_t items0 = sc.off[SEC_ITEMS];
    constexpr bool kOnlyLists = (F & (F_BVH | F_MEDIA)) == 0;
    // big (unstaged) scenes only: a small staged scene compiled from the same variant (Cornell smoke: no trees at all) loses
    // 4 % to the longer hot path, and calling the loop out of line gives the gain back on the Book-2 final scene
    // (0.484 inline / 0.463 out of line / 0.471 one tree after the other, Gpaths/s)
    constexpr bool kUnified = RTW_UNIFIED_BVH && !EXACT_TIES && (F & F_BVH) != 0 && !bvh4_variant(F) && SM < 2;
    int deferred = 0;  // kUnified: tree items wait on the traversal stack (bvh_hit_items)
#pragma unroll 1
    for (int it = 0; it < sc.n_world_items; it++) {
        const uint32_t ia = items0 + (uint32_t)it * 32u;
        float4 i0 = hot4<SM>(sv, ia);
        const int type = fi(i0.x), ixf = fi(i0.y), first = fi(i0.z), end = fi(i0.w);
        if (kOnlyLists || type == RTW_ITEM_PRIMS) {
            int before = h.prim;
            test_groups<F, SM, !EXACT_TIES>(sv, (uint32_t)first, (uint32_t)end, rk, time, tmin, h.t, h.prim);
            if (h.prim != before) h.item = it;
        } else if (kUnified && type == RTW_ITEM_BVH) {
            stack[deferred * stride] = kBvhEnter + it;
            deferred++;
        } else if (!kUnified && (F & F_BVH) && type == RTW_ITEM_BVH) {
            RayK lk = rk;
            if (ixf >= 0) lk = xf_rayk(ld_xf<SM>(sv, (uint32_t)ixf), rk);
            int p = -1;
            float t = h.t;
            // intersect_batch walks the reference's own topology (`first`); rendering walks the SAH tree over the same leaves
            // (`end`) or, in the variants bvh4_variant() names, its 4-wide collapse (`root4`)
            if constexpr (EXACT_TIES || !bvh4_variant(F)) {
                bvh_hit<F, SM, true, RTW_FAST_BOX && !EXACT_TIES, !EXACT_TIES>(sv, EXACT_TIES ? first : end, lk, time, tmin, h.t, stack, stride, t, p);
            } else {
                bvh4_hit<F, SM, true>(sv, (int)hot1<SM>(sv, ia + 28u), lk, time, tmin, h.t, stack, stride, t, p);
            }
            if (p >= 0) {
                h.t = t;
                h.prim = p;
                h.item = it;
            }
        } else if ((F & F_MEDIA) && type == RTW_ITEM_MEDIUM) {
            // constant_medium::hit (hittable.h:430-479)
            int mi = medium_idx++;
            if (!rng) continue;
            RayK lk = rk;
            if (ixf >= 0) lk = xf_rayk(ld_xf<SM>(sv, (uint32_t)ixf), rk);
            float t1 = FLT_MAX, t2 = FLT_MAX;
            // the two boundary searches (:438-442, :445-449) share one call site (code size): pass 1 starts at
            // `rec1.t + 0.0001f` (:447); in FP32 that step can vanish for |t| > ~800, so advance at least one ulp
            bool both = false;
            const int one_sphere = (int)hot1<SM>(sv, ia + 28u);  // DevItem.root4 of a medium: its boundary is ONE FP32 sphere record
            if (one_sphere >= 0) {
                // one quadratic gives both boundary hits, with the very numbers of the two searches below: the first search
                // (from -inf) returns the near root, the second (from just behind it) can only return the far one
                const float4 ra = hot4<SM>(sv, (uint32_t)one_sphere);
                float r0, r1;
                if (sphere_roots(lk.o - mk(ra.x, ra.y, ra.z), lk.d, lk.a, lk.inv_a, ra.w, r0, r1)) {
                    const float from = fmaxf(r0 + 0.0001f, nextafterf(r0, FLT_MAX));
                    both = (r0 > -FLT_MAX) & (r0 < FLT_MAX) & (r1 > from) & (r1 < FLT_MAX);
                    t1 = r0, t2 = r1;
                }
            }
#pragma unroll 1
            for (int pass = 0; pass < 2 && one_sphere < 0; pass++) {
                float tb = FLT_MAX;
                int pb = -1;
                const float from = pass == 0 ? -FLT_MAX : fmaxf(t1 + 0.0001f, nextafterf(t1, FLT_MAX));
                test_groups<F, SM, !EXACT_TIES>(sv, (uint32_t)first, (uint32_t)end, lk, time, from, tb, pb);
                if (pb < 0) break;
                if (pass == 0) t1 = tb;
                else t2 = tb, both = true;
            }
            if (!both) continue;
            if (t1 < tmin) t1 = tmin;  // :452-457
            if (t2 > h.t) t2 = h.t;
            if (t1 >= t2) continue;
            if (t1 < 0.0f) t1 = 0.0f;  // :460-461
            float len = sqrtf(lk.a);
            float inside = (t2 - t1) * len;  // :463-464
            float u = rng->draw(bounce, SITE_MEDIUM0 + (uint32_t)mi).x;
            float density_eff = __uint_as_float(hot1<SM>(sv, ia + 24));
            float hit_distance = -(1.0f / density_eff) * logf(u);  // :465 with density*multiplicity (Q5)
            if (hit_distance < inside) {
                h.t = t1 + hit_distance / len;  // :469
                h.prim = -2 - it;
                h.item = it;
            }
        }
    }
    if constexpr (kUnified) {
        if (deferred > 0) h = bvh_hit_items<F, SM>(items0, sv, rk, time, tmin, h, stack, stride, deferred);
    }
    if (EXACT_TIES && (F & F_BVH) && h.prim >= 0) {
#pragma unroll 1
        for (int it = 0; it < sc.n_world_items; it++) {
            float4 i0 = hot4<SM>(sv, items0 + (uint32_t)it * 32u);
            const int type = fi(i0.x), ixf = fi(i0.y), first = fi(i0.z), end = fi(i0.w);
            if (type == RTW_ITEM_PRIMS) {
                int before = h.prim;
                test_groups<F, SM, false>(sv, (uint32_t)first, (uint32_t)end, rk, time, tmin, h.t, h.prim);
                if (h.prim != before) h.item = it;
            } else if (type == RTW_ITEM_BVH) {
                RayK lk = rk;
                if (ixf >= 0) lk = xf_rayk(ld_xf<SM>(sv, (uint32_t)ixf), rk);
                int p = -1;
                float t = h.t;
                bvh_hit<F, SM, false, false, false>(sv, first, lk, time, tmin, h.t, stack, stride, t, p);
                if (p >= 0) {
                    h.t = t;
                    h.prim = p;
                    h.item = it;
                }
            }
        }
    }
    return h;
}

// hit_record (hittable.h:16-29) rebuilt from the winning id.
struct HitRec {
    V3 p, n;
    float u, v;
    int mat;
};

RTW_DEV void sphere_uv(V3 n, float& u, float& v) {  // sphere.h:115-122 / utility.h:6-12
    const float PI = 3.14159265358979323846f;
    float phi = atan2f(n.z, n.x);
    float theta = asinf(n.y);
    u = 1.0f - (phi + PI) / (2.0f * PI);
    v = (theta + PI / 2.0f) / PI;
}

// Material index of the hit (needed before complete_hit: whether u,v are wanted depends on the material's texture).
template <int SM>
RTW_DEV int hit_material(const DevScene& sc, const SceneView& sv, const HitId& h, uint32_t& pm_out) {
    if (h.prim < 0) {
        pm_out = 0;
        return (int)hot1<SM>(sv, sc.off[SEC_ITEMS] + (uint32_t)(-2 - h.prim) * 32u + 16u);
    }
    pm_out = cold1<SM>(sv, sc.off[SEC_PM] + (uint32_t)h.prim * 4u);
    return (int)(pm_out >> PM_MAT_SHIFT);
}

template <uint32_t F, int SM>
RTW_DEV HitRec complete_hit(const DevScene& sc, const SceneView& sv, const HitId& h, uint32_t m, V3 o, V3 d, float time,
                            bool need_uv) {
    HitRec r;
    r.u = r.v = 0.0f;
    if ((F & F_MEDIA) && h.prim < 0) {  // constant_medium (hittable.h:469-472): arbitrary normal, u/v never written (Q13)
        const uint32_t ia = sc.off[SEC_ITEMS] + (uint32_t)(-2 - h.prim) * 32u;
        uint2 mf = hot2<SM>(sv, ia + 16);
        r.p = madd(o, h.t, d);
        r.n = mk(mf.y ? -1.0f : 1.0f, 0.0f, 0.0f);
        r.mat = (int)mf.x;
        return r;
    }
    const int type = m & PM_TYPE_MASK;
    r.mat = (int)(m >> PM_MAT_SHIFT);
    // object-space ray: item transform, then the primitive's own chain
    V3 ol = o, dl = d;
    int ixf = -1, pxf = 0;
    DevXf xi, xp;
    if (F & F_XF) {
        if (F & (F_BVH | F_MEDIA)) {  // only BVH / medium items carry an item-level transform
            ixf = (int)hot1<SM>(sv, sc.off[SEC_ITEMS] + (uint32_t)h.item * 32u + 4u);
            if (ixf >= 0) {
                xi = ld_xf<SM>(sv, (uint32_t)ixf);
                xf_ray(xi, o, d, ol, dl);
            }
        }
        pxf = (m >> PM_XF_SHIFT) & PM_XF_MASK;
        if (pxf) {
            xp = ld_xf<SM>(sv, sc.off[SEC_XFS] + (uint32_t)(pxf - 1) * 32u);
            V3 o2, d2;
            xf_ray(xp, ol, dl, o2, d2);
            ol = o2, dl = d2;
        }
    }
    float4 pa = cold4<SM>(sv, sc.off[SEC_PA] + (uint32_t)h.prim * 16u);
    V3 pl = madd(ol, h.t, dl);  // point_at_parameter in object space
    V3 nl;
    if (type <= RTW_PRIM_MOVING_SPHERE) {
        V3 c = mk(pa.x, pa.y, pa.z);
        float radius = pa.w;
        if ((F & F_MOVING) && type == RTW_PRIM_MOVING_SPHERE) {
            float4 pb = cold4<SM>(sv, sc.off[SEC_PB] + (uint32_t)h.prim * 16u);
            c = madd(c, time, mk(pb.x, pb.y, pb.z));
        }
        float inv_r = rcp_fast(radius);
        if ((F & F_PRECISE) && (m & PM_PRECISE)) {
            float4 pb = cold4<SM>(sv, sc.off[SEC_PB] + (uint32_t)h.prim * 16u);
            const uint32_t pd = (uint32_t)fi(pb.w);
            double cx = cold_f64<SM>(sv, pd), cy = cold_f64<SM>(sv, pd + 8), cz = cold_f64<SM>(sv, pd + 16);
            double rad = cold_f64<SM>(sv, pd + 24);
            // p - centre in FP64: at r = 1000 the FP32 difference would quantise the normal to ~1e-4
            double t = h.t;
            nl = mk((float)(((double)ol.x + t * dl.x - cx) / rad), (float)(((double)ol.y + t * dl.y - cy) / rad),
                    (float)(((double)ol.z + t * dl.z - cz) / rad));
        } else {
            nl = (pl - c) * inv_r;  // sphere.h:62 — negative radius flips the normal (Q8)
        }
        if (need_uv) sphere_uv(nl, r.u, r.v);
    } else {
        float k = __uint_as_float(cold1<SM>(sv, sc.off[SEC_PB] + (uint32_t)h.prim * 16u));
        float a, b;
        // snap the plane coordinate: in exact arithmetic p lies on the rect (FP32 hygiene against self-hits)
        if (type == RTW_PRIM_XY_RECT) {
            pl.z = k; a = pl.x; b = pl.y; nl = mk(0.0f, 0.0f, 1.0f);
        } else if (type == RTW_PRIM_XZ_RECT) {
            pl.y = k; a = pl.x; b = pl.z; nl = mk(0.0f, 1.0f, 0.0f);
        } else {
            pl.x = k; a = pl.y; b = pl.z; nl = mk(1.0f, 0.0f, 0.0f);
        }
        if (need_uv) {  // hittable.h:158-159
            r.u = (a - pa.x) / (pa.y - pa.x);
            r.v = (b - pa.z) / (pa.w - pa.z);
        }
    }
    if (m & PM_FLIP) nl = -nl;  // hittable.h:277
    if (F & F_XF) {
        if (pxf) {
            nl = xf_vec_out(xp, nl);
            pl = xf_point_out(xp, pl);
        }
        if ((F & (F_BVH | F_MEDIA)) && ixf >= 0) {
            nl = xf_vec_out(xi, nl);
            pl = xf_point_out(xi, pl);
        }
    }
    r.p = pl;
    r.n = nl;
    return r;
}

}  // namespace rtw
