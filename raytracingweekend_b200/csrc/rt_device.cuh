// rt_device.cuh — device functions of the hot path, shared by the megakernel, the wavefront pipeline and
// intersect_batch.  Every function cites the reference code it evaluates (paths relative to
// /root/reference/RayTracingWeekend/).  Arithmetic is FP32 (FP64 only for PM_PRECISE spheres).
#pragma once
#include <cfloat>
#include <cstdint>
#include <cuda_runtime.h>

#include "device_scene.hpp"

namespace rtw {

#define RTW_DEV __device__ __forceinline__

// ------------------------------------------------------------------------------------------------ vec3
struct V3 {
    float x, y, z;
};
RTW_DEV V3 mk(float x, float y, float z) { return V3{x, y, z}; }
RTW_DEV V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
RTW_DEV V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
RTW_DEV V3 operator*(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
RTW_DEV V3 operator*(float s, V3 a) { return mk(s * a.x, s * a.y, s * a.z); }
RTW_DEV V3 operator*(V3 a, float s) { return mk(s * a.x, s * a.y, s * a.z); }
RTW_DEV V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
RTW_DEV float dot(V3 a, V3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
RTW_DEV V3 cross(V3 a, V3 b) {  // vec3.h:54-59
    return mk(a.y * b.z - a.z * b.y, -(a.x * b.z - a.z * b.x), a.x * b.y - a.y * b.x);
}
RTW_DEV V3 normalize(V3 a) {  // vec3.h:61-67
    float inv = rsqrtf(dot(a, a));
    return a * inv;
}
RTW_DEV V3 madd(V3 o, float t, V3 d) { return mk(fmaf(t, d.x, o.x), fmaf(t, d.y, o.y), fmaf(t, d.z, o.z)); }

struct Ray {  // ray.h:17-20 — direction is NOT normalised
    V3 o, d;
    float time;
};

// ------------------------------------------------------------------------------------------------ RNG
// Philox4x32-10 (Salmon et al. 2011), counter = (pixel, sample, bounce, site), key = seed.
// One call yields four U[0,1) values with 24 random bits each, exactly representable in FP32 and FP64,
// so the CPU oracle in philox mode consumes bit-identical uniforms.
struct U4 {
    float x, y, z, w;
};
RTW_DEV uint4 philox4x32_10(uint4 c, uint2 k) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int i = 0; i < 10; i++) {
        uint32_t hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
        uint32_t hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += W0;
        k.y += W1;
    }
    return c;
}
RTW_DEV float u01(uint32_t x) { return (float)(x >> 8) * 5.9604644775390625e-8f; }  // 2^-24
struct Rng {
    uint32_t pixel, sample, seed_lo, seed_hi;
    RTW_DEV U4 draw(uint32_t bounce, uint32_t site) const {
        uint4 r = philox4x32_10(make_uint4(pixel, sample, bounce, site), make_uint2(seed_lo, seed_hi));
        return U4{u01(r.x), u01(r.y), u01(r.z), u01(r.w)};
    }
};
enum : uint32_t { SITE_MAIN = 0, SITE_BALL = 1, SITE_LENS = 1, SITE_MEDIUM0 = 2 };

// ------------------------------------------------------------------------------------------------ loads
// SM = how much of the scene blob the block staged into shared memory: 0 nothing, 1 the hot sections, 2 all.
// Staged sections are read with plain loads (ld.shared after address-space inference), the others through the
// read-only global path in 16-byte vectors.
template <bool STAGED, typename T>
RTW_DEV T ld_pod(const T* p) {
    static_assert(sizeof(T) % 4 == 0, "pod size");
    if constexpr (STAGED) {
        return *p;
    } else if constexpr (sizeof(T) % 16 == 0) {
        union {
            T v;
            float4 q[sizeof(T) / 16];
        } u;
#pragma unroll
        for (int i = 0; i < (int)(sizeof(T) / 16); i++) u.q[i] = __ldg(reinterpret_cast<const float4*>(p) + i);
        return u.v;
    } else if constexpr (sizeof(T) == 8) {
        union {
            T v;
            int2 q;
        } u;
        u.q = __ldg(reinterpret_cast<const int2*>(p));
        return u.v;
    } else {
        union {
            T v;
            uint32_t q[sizeof(T) / 4];
        } u;
#pragma unroll
        for (int i = 0; i < (int)(sizeof(T) / 4); i++) u.q[i] = __ldg(reinterpret_cast<const uint32_t*>(p) + i);
        return u.v;
    }
}
template <int SM, typename T>
RTW_DEV T ld_hot(const T* p) { return ld_pod<(SM >= 1), T>(p); }
template <int SM, typename T>
RTW_DEV T ld_cold(const T* p) { return ld_pod<(SM >= 2), T>(p); }
template <int SM>
RTW_DEV int ld_u8(const uint8_t* p) {  // cold section bytes (Perlin permutations)
    if constexpr (SM >= 2) return *p;
    else return __ldg(p);
}

// ------------------------------------------------------------------------------------------------ transforms
// translate/rotate_y chain composed on the host (hittable.h:299-311, 373-404): ray into object space.
RTW_DEV void xf_ray(const DevXf& x, V3 o, V3 d, V3& ol, V3& dl) {
    ol = mk(fmaf(x.c, o.x, -x.s * o.z) + x.tx, o.y + x.ty, fmaf(x.s, o.x, x.c * o.z) + x.tz);
    dl = mk(fmaf(x.c, d.x, -x.s * d.z), d.y, fmaf(x.s, d.x, x.c * d.z));
}
RTW_DEV V3 xf_vec_out(const DevXf& x, V3 n) {  // object -> world for directions (hittable.h:393-394)
    return mk(fmaf(x.c, n.x, x.s * n.z), n.y, fmaf(-x.s, n.x, x.c * n.z));
}
RTW_DEV V3 xf_point_out(const DevXf& x, V3 p) {  // object -> world for points (hittable.h:390-391, :304)
    V3 q = mk(p.x - x.tx, p.y - x.ty, p.z - x.tz);
    return xf_vec_out(x, q);
}

// ------------------------------------------------------------------------------------------------ primitives
// sphere_base::hit (sphere.h:46-81): near root, then far root, each with STRICT t_min < t < t_max (Q7).
// FP32 form: the far-from-cancellation root comes from q = -b -/+ sqrt(disc), the other one from c/q, and the
// discriminant from the perpendicular distance (|oc - (b/a) d|^2), which is what keeps r = 1000 usable.
// `a` = dot(d, d) and inv_a = 1/a are per-ray constants hoisted out of the primitive loop.
RTW_DEV bool sphere_roots(V3 oc, V3 d, float a, float inv_a, float radius, float& t0, float& t1) {
    float b = dot(oc, d);
    float ba = b * inv_a;
    V3 l = mk(fmaf(-ba, d.x, oc.x), fmaf(-ba, d.y, oc.y), fmaf(-ba, d.z, oc.z));
    float r2 = radius * radius;
    float disc = a * (r2 - dot(l, l));  // == b*b - a*c (sphere.h:53)
    if (!(disc > 0.0f)) return false;
    float sq = sqrtf(disc);
    float c = dot(oc, oc) - r2;
    if (b < 0.0f) {
        float q = sq - b;
        t1 = q * inv_a;
        t0 = c / q;
    } else {
        float q = -b - sq;
        t0 = q * inv_a;
        t1 = c / q;
    }
    return true;
}
RTW_DEV bool sphere_roots_f64(V3 o, V3 d, const double* pd, float& t0, float& t1) {
    double ocx = (double)o.x - pd[0], ocy = (double)o.y - pd[1], ocz = (double)o.z - pd[2];
    double dx = d.x, dy = d.y, dz = d.z;
    double a = dx * dx + dy * dy + dz * dz;
    double b = ocx * dx + ocy * dy + ocz * dz;
    double c = ocx * ocx + ocy * ocy + ocz * ocz - pd[3] * pd[3];
    double disc = b * b - a * c;
    if (!(disc > 0.0)) return false;
    double sq = sqrt(disc);
    t0 = (float)((-b - sq) / a);
    t1 = (float)((-b + sq) / a);
    return true;
}
// Pick the root the reference would return; false if neither lies in (tmin, tmax).
RTW_DEV bool sphere_pick(float t0, float t1, float tmin, float tmax, float& t) {
    if (t0 < tmax && t0 > tmin) {
        t = t0;
        return true;
    }
    if (t1 < tmax && t1 > tmin) {
        t = t1;
        return true;
    }
    return false;
}

// x?_rect::hit (hittable.h:149-165, 184-200, 241-257): INCLUSIVE bounds; a NaN t passes like the reference (Q39).
// (on, dn) are the ray's origin/direction along the plane normal, (oa, da), (ob, db) along the two in-plane axes;
// inv_n = 1/dn is shared by every rect of that orientation in the group.  (k - on) * inv_n is exactly 0 for a ray
// that starts on the plane and +-inf / NaN for dn == 0 exactly as the division is.
RTW_DEV bool rect_test(float4 ra, float k, float on, float inv_n, float oa, float da, float ob, float db, float tmin,
                       float tmax, float& t_out) {
    float t = (k - on) * inv_n;
    float a = fmaf(t, da, oa), b = fmaf(t, db, ob);
    bool miss = (t < tmin) | (t > tmax) | (a < ra.x) | (a > ra.y) | (b < ra.z) | (b > ra.w);
    t_out = t;
    return !miss;
}

// aabb::hit (aabb.h:17-47) with std::max/std::min written out so NaN behaves as in the reference (Q32).
RTW_DEV bool aabb_hit(float4 lo, float4 hi, V3 o, V3 inv, float tmin, float tmax) {
    {
        float t0 = (lo.x - o.x) * inv.x, t1 = (hi.x - o.x) * inv.x;
        if (inv.x < 0.0f) { float s = t0; t0 = t1; t1 = s; }
        tmin = (t0 < tmin) ? tmin : t0;
        tmax = (tmax < t1) ? tmax : t1;
        if (tmax <= tmin) return false;
    }
    {
        float t0 = (lo.y - o.y) * inv.y, t1 = (hi.y - o.y) * inv.y;
        if (inv.y < 0.0f) { float s = t0; t0 = t1; t1 = s; }
        tmin = (t0 < tmin) ? tmin : t0;
        tmax = (tmax < t1) ? tmax : t1;
        if (tmax <= tmin) return false;
    }
    {
        float t0 = (lo.z - o.z) * inv.z, t1 = (hi.z - o.z) * inv.z;
        if (inv.z < 0.0f) { float s = t0; t0 = t1; t1 = s; }
        tmin = (t0 < tmin) ? tmin : t0;
        tmax = (tmax < t1) ? tmax : t1;
        if (tmax <= tmin) return false;
    }
    return true;
}

// ------------------------------------------------------------------------------------------------ closest hit
// Result of world.hit(): winning primitive (or medium item) and t.  Normal/uv/p are reconstructed afterwards
// from the id (complete_hit), so the search loop carries two registers.
struct HitId {
    float t;
    int prim;  // >= 0 primitive; -1 miss; <= -2: medium item ~prim
    int item;  // item that owns the hit (its xf applies)
};

// Per-ray constants of the primitive tests: reciprocal direction (rect planes, aabb slabs), a = d.d and 1/a (spheres).
struct RayK {
    V3 o, d, inv;
    float a, inv_a;
};
RTW_DEV RayK make_rayk(V3 o, V3 d) {
    RayK k;
    k.o = o, k.d = d;
    k.inv = mk(__frcp_rn(d.x), __frcp_rn(d.y), __frcp_rn(d.z));  // aabb.h:22 (1.0 / direction)
    k.a = dot(d, d);
    k.inv_a = __frcp_rn(k.a);
    return k;
}

// hittable_list::hit (hittable_list.h:16-24) over groups [first, first+count): tmax shrinks to the running closest
// hit — rects accept t == closest (a later object wins a tie), spheres need t < closest (Q38).  Inside a group the
// records are ordered spheres, moving spheres, xy, xz, yz rects, each in list order, which evaluates the same
// winner as the reference's interleaved order (see device_scene.hpp).  The second pass of hittable_list::hit
// (:26-34) re-selects the same winner for deterministic children.
// All loop bounds and record addresses are warp-uniform: one broadcast load per record.
template <uint32_t F, int SM>
RTW_DEV void test_groups(const SceneView& sv, int first, int count, const RayK& rk, float time, float tmin, float& t_best,
                         int& prim_best) {
    for (int gi = first; gi < first + count; gi++) {
        DevGroup g = ld_hot<SM>(&sv.groups[gi]);
        RayK k = rk;
        if ((F & F_XF) && g.xf >= 0) {
            DevXf x = ld_hot<SM>(&sv.xfs[g.xf]);
            V3 ol, dl;
            xf_ray(x, rk.o, rk.d, ol, dl);
            k = make_rayk(ol, dl);
        }
        for (int r = g.sph0; r < g.sph1; r++) {
            DevRec rec = ld_hot<SM>(&sv.recs[r]);
            int id = __float_as_int(rec.b.w);
            float t0, t1, t;
            bool roots;
            if ((F & F_PRECISE) && (id & REC_PRECISE)) {
                roots = sphere_roots_f64(k.o, k.d, sv.pd + 4 * __float_as_int(rec.b.x), t0, t1);
                id &= ~REC_PRECISE;
            } else {
                roots = sphere_roots(k.o - mk(rec.a.x, rec.a.y, rec.a.z), k.d, k.a, k.inv_a, rec.a.w, t0, t1);
            }
            if (roots && sphere_pick(t0, t1, tmin, t_best, t)) {
                t_best = t;
                prim_best = id;
            }
        }
        if (F & F_MOVING) {
            for (int r = g.sph1; r < g.msph1; r++) {
                DevRec rec = ld_hot<SM>(&sv.recs[r]);
                // sphere.h:22-25 folded to c' + time*delta on the host
                V3 c = madd(mk(rec.a.x, rec.a.y, rec.a.z), time, mk(rec.b.x, rec.b.y, rec.b.z));
                float t0, t1, t;
                if (sphere_roots(k.o - c, k.d, k.a, k.inv_a, rec.a.w, t0, t1) && sphere_pick(t0, t1, tmin, t_best, t)) {
                    t_best = t;
                    prim_best = __float_as_int(rec.b.w);
                }
            }
        }
        for (int r = g.msph1; r < g.xy1; r++) {  // xy_rect: plane z = k
            DevRec rec = ld_hot<SM>(&sv.recs[r]);
            float t;
            if (rect_test(rec.a, rec.b.x, k.o.z, k.inv.z, k.o.x, k.d.x, k.o.y, k.d.y, tmin, t_best, t)) {
                t_best = t;
                prim_best = __float_as_int(rec.b.w);
            }
        }
        for (int r = g.xy1; r < g.xz1; r++) {  // xz_rect: plane y = k
            DevRec rec = ld_hot<SM>(&sv.recs[r]);
            float t;
            if (rect_test(rec.a, rec.b.x, k.o.y, k.inv.y, k.o.x, k.d.x, k.o.z, k.d.z, tmin, t_best, t)) {
                t_best = t;
                prim_best = __float_as_int(rec.b.w);
            }
        }
        for (int r = g.xz1; r < g.yz1; r++) {  // yz_rect: plane x = k
            DevRec rec = ld_hot<SM>(&sv.recs[r]);
            float t;
            if (rect_test(rec.a, rec.b.x, k.o.x, k.inv.x, k.o.y, k.d.y, k.o.z, k.d.z, tmin, t_best, t)) {
                t_best = t;
                prim_best = __float_as_int(rec.b.w);
            }
        }
    }
}

}  // namespace rtw
