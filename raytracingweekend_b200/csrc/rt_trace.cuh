// rt_trace.cuh — world.hit(): list walk, bvh_node traversal, constant_medium, hit completion.
#pragma once
#include "rt_device.cuh"

namespace rtw {

constexpr int kStackDepth = 32;  // shared-memory traversal stack entries per thread

// bvh_node::hit (hittable.h:97-129, as intended — SURVEY F1).  A node is entered iff its box passes
// aabb::hit with the caller's [tmin, tmax]; both children are then examined; the smaller t wins and on equal t
// the right (later) subtree wins (:107).  Leaves are lists (a box is six rects) evaluated with list semantics
// against the caller's tmax.  The interval used for box culling shrinks to the best hit found so far, widened
// by a relative margin that covers the 1/d-multiply vs divide rounding gap, so no candidate with
// t <= best is ever culled and the winner is the reference's.
// Stack: shared memory, one column per thread (`stack[level * stride]`), depth-first, left first.
template <uint32_t F, int SM, bool SHRINK>
RTW_DEV void bvh_hit(const SceneView& sv, int root, const RayK& rk, float time, float tmin, float tmax, int* stack,
                     int stride, float& t_out, int& prim_out) {
    const V3 o = rk.o, inv = rk.inv;
    float best = tmax;
    int best_prim = -1;
    int sp = 0;
    int ref = root;
    for (;;) {
        if (ref >= 0) {
            float4 lo = ld_hot<SM>(&sv.node_lo[ref]), hi = ld_hot<SM>(&sv.node_hi[ref]);
            float cull = (SHRINK && best_prim >= 0) ? fmaf(best, 4e-6f, best) : tmax;
            if (aabb_hit(lo, hi, o, inv, tmin, cull)) {
                int l = __float_as_int(lo.w), r = __float_as_int(hi.w);
                if (r != l) {  // n == 1 node has left == right (hittable.h:68-72): one visit is enough
                    stack[sp * stride] = r;
                    sp++;
                }
                ref = l;
                continue;
            }
        } else {
            int2 lf = ld_hot<SM>(&sv.leaves[~ref]);
            float t_leaf = tmax;
            int p_leaf = -1;
            test_groups<F, SM>(sv, lf.x, lf.y, rk, time, tmin, t_leaf, p_leaf);
            if (p_leaf >= 0 && (best_prim < 0 || t_leaf <= best)) {
                best = t_leaf;
                best_prim = p_leaf;
            }
        }
        if (sp == 0) break;
        sp--;
        ref = stack[sp * stride];
    }
    if (best_prim >= 0) {
        t_out = best;
        prim_out = best_prim;
    }
}

// world.hit(r, tmin, tmax) (RayTracingWeekend.cpp:52 -> hittable_list.h:11-37) over the flattened items.
// `rng` == nullptr skips constant_medium items (intersect_batch: media are stochastic).
// EXACT_TIES replays the second pass of hittable_list::hit (hittable_list.h:26-34) for scenes with bvh_node items:
// with tmax == closest the node boxes are re-tested against the final t, which can change which of several
// primitives at exactly that t is reported (coplanar faces of adjacent boxes, seen from inside a box).  Rays of a
// render never start inside closed boxes, so the megakernel skips it; intersect_batch (parity tier 1) runs it.
template <uint32_t F, int SM, bool EXACT_TIES>
RTW_DEV HitId closest_hit(const DevScene& sc, const SceneView& sv, V3 o, V3 d, float time, float tmin, float tmax,
                          const Rng* rng, uint32_t bounce, int* stack, int stride) {
    HitId h;
    h.t = tmax;
    h.prim = -1;
    h.item = -1;
    int medium_idx = 0;
    const RayK rk = make_rayk(o, d);
    for (int it = 0; it < sc.n_world_items; it++) {
        DevItem item = ld_hot<SM>(&sv.items[it]);
        if (item.type == RTW_ITEM_PRIMS) {
            int before = h.prim;
            test_groups<F, SM>(sv, item.first, item.count, rk, time, tmin, h.t, h.prim);
            if (h.prim != before) h.item = it;
        } else if ((F & F_BVH) && item.type == RTW_ITEM_BVH) {
            RayK lk = rk;
            if (item.xf >= 0) {
                V3 ol, dl;
                xf_ray(ld_hot<SM>(&sv.xfs[item.xf]), o, d, ol, dl);
                lk = make_rayk(ol, dl);
            }
            int p = -1;
            float t = h.t;
            bvh_hit<F, SM, true>(sv, item.first, lk, time, tmin, h.t, stack, stride, t, p);
            if (p >= 0) {
                h.t = t;
                h.prim = p;
                h.item = it;
            }
        } else if ((F & F_MEDIA) && item.type == RTW_ITEM_MEDIUM) {
            // constant_medium::hit (hittable.h:430-479)
            int mi = medium_idx++;
            if (!rng) continue;
            RayK lk = rk;
            if (item.xf >= 0) {
                V3 ol, dl;
                xf_ray(ld_hot<SM>(&sv.xfs[item.xf]), o, d, ol, dl);
                lk = make_rayk(ol, dl);
            }
            float t1 = FLT_MAX, t2 = FLT_MAX;
            int p1 = -1, p2 = -1;
            test_groups<F, SM>(sv, item.first, item.count, lk, time, -FLT_MAX, t1, p1);  // :438-442
            if (p1 < 0) continue;
            // :447 `rec1.t + 0.0001f`; in FP32 the step can vanish for |t| > ~800, so advance at least one ulp
            float step = fmaxf(t1 + 0.0001f, nextafterf(t1, FLT_MAX));
            test_groups<F, SM>(sv, item.first, item.count, lk, time, step, t2, p2);  // :445-449
            if (p2 < 0) continue;
            if (t1 < tmin) t1 = tmin;  // :452-457
            if (t2 > h.t) t2 = h.t;
            if (t1 >= t2) continue;
            if (t1 < 0.0f) t1 = 0.0f;  // :460-461
            float len = sqrtf(lk.a);
            float inside = (t2 - t1) * len;  // :463-464
            float u = rng->draw(bounce, SITE_MEDIUM0 + (uint32_t)mi).x;
            float hit_distance = -(1.0f / item.density_eff) * logf(u);  // :465 with density*multiplicity (Q5)
            if (hit_distance < inside) {
                h.t = t1 + hit_distance / len;  // :469
                h.prim = -2 - it;
                h.item = it;
            }
        }
    }
    if (EXACT_TIES && (F & F_BVH) && h.prim >= 0) {
        for (int it = 0; it < sc.n_world_items; it++) {
            DevItem item = ld_hot<SM>(&sv.items[it]);
            if (item.type == RTW_ITEM_PRIMS) {
                int before = h.prim;
                test_groups<F, SM>(sv, item.first, item.count, rk, time, tmin, h.t, h.prim);
                if (h.prim != before) h.item = it;
            } else if (item.type == RTW_ITEM_BVH) {
                RayK lk = rk;
                if (item.xf >= 0) {
                    V3 ol, dl;
                    xf_ray(ld_hot<SM>(&sv.xfs[item.xf]), o, d, ol, dl);
                    lk = make_rayk(ol, dl);
                }
                int p = -1;
                float t = h.t;
                bvh_hit<F, SM, false>(sv, item.first, lk, time, tmin, h.t, stack, stride, t, p);
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

template <uint32_t F, int SM>
RTW_DEV HitRec complete_hit(const SceneView& sv, const HitId& h, V3 o, V3 d, float time, bool need_uv) {
    HitRec r;
    r.u = r.v = 0.0f;
    if (h.prim < 0) {  // constant_medium (hittable.h:469-472): arbitrary normal, u/v never written (Q13)
        DevItem item = ld_hot<SM>(&sv.items[-2 - h.prim]);
        r.p = madd(o, h.t, d);
        r.n = mk(item.flip ? -1.0f : 1.0f, 0.0f, 0.0f);
        r.mat = item.material;
        return r;
    }
    uint32_t m = ld_cold<SM>(&sv.pm[h.prim]);
    int type = m & PM_TYPE_MASK;
    r.mat = (int)(m >> PM_MAT_SHIFT);
    // object-space ray: item transform, then the primitive's own chain
    V3 ol = o, dl = d;
    int ixf = -1, pxf = 0;
    if (F & F_XF) {
        ixf = ld_hot<SM>(&sv.items[h.item].xf);
        if (ixf >= 0) xf_ray(ld_hot<SM>(&sv.xfs[ixf]), o, d, ol, dl);
        pxf = (m >> PM_XF_SHIFT) & PM_XF_MASK;
        if (pxf) {
            V3 o2, d2;
            xf_ray(ld_hot<SM>(&sv.xfs[pxf - 1]), ol, dl, o2, d2);
            ol = o2, dl = d2;
        }
    }
    float4 pa = ld_cold<SM>(&sv.pa[h.prim]);
    V3 pl = madd(ol, h.t, dl);  // point_at_parameter in object space
    V3 nl;
    if (type <= RTW_PRIM_MOVING_SPHERE) {
        V3 c = mk(pa.x, pa.y, pa.z);
        float radius = pa.w;
        if ((F & F_MOVING) && type == RTW_PRIM_MOVING_SPHERE) {
            float4 pb = ld_cold<SM>(&sv.pb[h.prim]);
            c = madd(c, time, mk(pb.x, pb.y, pb.z));
        }
        float inv_r = 1.0f / radius;
        if ((F & F_PRECISE) && (m & PM_PRECISE)) {
            float4 pb = ld_cold<SM>(&sv.pb[h.prim]);
            const double* pd = sv.pd + 4 * __float_as_int(pb.w);
            // p - centre in FP64: at r = 1000 the FP32 difference would quantise the normal to ~1e-4
            double t = h.t;
            nl = mk((float)(((double)ol.x + t * dl.x - pd[0]) / pd[3]), (float)(((double)ol.y + t * dl.y - pd[1]) / pd[3]),
                    (float)(((double)ol.z + t * dl.z - pd[2]) / pd[3]));
        } else {
            nl = (pl - c) * inv_r;  // sphere.h:62 — negative radius flips the normal (Q8)
        }
        if (need_uv) sphere_uv(nl, r.u, r.v);
    } else {
        float4 pb = ld_cold<SM>(&sv.pb[h.prim]);
        float a, b;
        // snap the plane coordinate: in exact arithmetic p lies on the rect (FP32 hygiene against self-hits)
        if (type == RTW_PRIM_XY_RECT) {
            pl.z = pb.x; a = pl.x; b = pl.y; nl = mk(0.0f, 0.0f, 1.0f);
        } else if (type == RTW_PRIM_XZ_RECT) {
            pl.y = pb.x; a = pl.x; b = pl.z; nl = mk(0.0f, 1.0f, 0.0f);
        } else {
            pl.x = pb.x; a = pl.y; b = pl.z; nl = mk(1.0f, 0.0f, 0.0f);
        }
        if (need_uv) {  // hittable.h:158-159
            r.u = (a - pa.x) / (pa.y - pa.x);
            r.v = (b - pa.z) / (pa.w - pa.z);
        }
    }
    if (m & PM_FLIP) nl = -nl;  // hittable.h:277
    if (F & F_XF) {
        if (pxf) {
            DevXf x = ld_hot<SM>(&sv.xfs[pxf - 1]);
            nl = xf_vec_out(x, nl);
            pl = xf_point_out(x, pl);
        }
        if (ixf >= 0) {
            DevXf x = ld_hot<SM>(&sv.xfs[ixf]);
            nl = xf_vec_out(x, nl);
            pl = xf_point_out(x, pl);
        }
    }
    r.p = pl;
    r.n = nl;
    return r;
}

}  // namespace rtw
