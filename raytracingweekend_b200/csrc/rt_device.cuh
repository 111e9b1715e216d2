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

// Diagnostics build only (-DRTW_COUNT_EVENTS, librtw_b200_count.so): counts the traversal events the render kernel's own
// acceleration structures cost, for the algorithmic-flop figure bench.py reports beside the reference-math count.
#ifdef RTW_COUNT_EVENTS
enum { EVC_AABB = 0, EVC_SPHERE_TEST, EVC_SPHERE_ROOT, EVC_MOVING, EVC_RECT, EVC_BOX, EVC_XF, EVC_N };
static __device__ unsigned long long g_evc[8];
#define RTW_EV(k, n) atomicAdd(&g_evc[k], (unsigned long long)(n))
#else
#define RTW_EV(k, n) ((void)0)
#endif

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
RTW_DEV V3 sel(bool c, V3 a, V3 b) { return mk(c ? a.x : b.x, c ? a.y : b.y, c ? a.z : b.z); }

// 1/x as a single MUFU.RCP (1 ulp; 1/0 = inf, 1/denormal = inf under ftz — the slab and plane tests only need the
// IEEE special values, not correct rounding; tier-2 tolerance is 1e-5 relative).
RTW_DEV float rcp_fast(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

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
// `rk` = the ten round keys (seed + i * Weyl constants), computed once per launch on the host (DevRenderParams::rk):
// they sit in the kernel-parameter constant bank and feed the round's LOP3 directly, which takes the 18 key-schedule
// adds out of every draw.
RTW_DEV uint4 philox4x32_10(uint4 c, const uint32_t* __restrict__ rk) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
#pragma unroll
    for (int i = 0; i < 10; i++) {
        uint32_t hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
        uint32_t hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
        c = make_uint4(hi1 ^ c.y ^ rk[2 * i], lo1, hi0 ^ c.w ^ rk[2 * i + 1], lo0);
    }
    return c;
}
RTW_DEV float u01(uint32_t x) { return (float)(x >> 8) * 5.9604644775390625e-8f; }  // 2^-24
// RTW_PHILOX_OOL (per translation unit): ONE out-of-line copy of the ten rounds instead of one inlined copy per draw site (six
// sites x 57 instructions in the all-features kernels, which are bound by the SM instruction cache: DESIGN.md 4.6).
#ifdef RTW_PHILOX_OOL
static __device__ __noinline__ uint4 philox4x32_10_call(uint4 c, const uint32_t* __restrict__ rk) { return philox4x32_10(c, rk); }
#endif
struct Rng {
    uint32_t pixel, sample;
    const uint32_t* rk;  // DevRenderParams::rk of this launch
    RTW_DEV U4 draw(uint32_t bounce, uint32_t site) const {
#ifdef RTW_PHILOX_OOL
        uint4 r = philox4x32_10_call(make_uint4(pixel, sample, bounce, site), rk);
#else
        uint4 r = philox4x32_10(make_uint4(pixel, sample, bounce, site), rk);
#endif
        return U4{u01(r.x), u01(r.y), u01(r.z), u01(r.w)};
    }
};
enum : uint32_t { SITE_MAIN = 0, SITE_BALL = 1, SITE_LENS = 1, SITE_MEDIUM0 = 2 };

// ------------------------------------------------------------------------------------------------ loads
// SM = how much of the scene blob the block staged into shared memory: 0 nothing, 1 the hot sections, 2 all.
// `off` is a byte offset from the blob start (device_scene.hpp).  Staged data is read with explicit ld.shared on a
// 32-bit address (one LDS per 16 bytes, no generic-address arithmetic), the rest through the read-only global path.
// The asm is not volatile on purpose: the blob is immutable once staged, so loads may be scheduled freely; they
// depend on `sbase`, which is produced after the staging barrier (stage_scene), and cannot move above it.
template <bool STAGED>
RTW_DEV float4 ldb4(const SceneView& sv, uint32_t off) {
    float4 v;
    if constexpr (STAGED) {
        asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(sv.sbase + off));
    } else {
        v = __ldg(reinterpret_cast<const float4*>(sv.g + off));
    }
    return v;
}
template <bool STAGED>
RTW_DEV uint32_t ldb1(const SceneView& sv, uint32_t off) {
    uint32_t v;
    if constexpr (STAGED) {
        asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(sv.sbase + off));
    } else {
        v = __ldg(reinterpret_cast<const uint32_t*>(sv.g + off));
    }
    return v;
}
template <bool STAGED>
RTW_DEV uint2 ldb2(const SceneView& sv, uint32_t off) {
    uint2 v;
    if constexpr (STAGED) {
        asm("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(sv.sbase + off));
    } else {
        v = __ldg(reinterpret_cast<const uint2*>(sv.g + off));
    }
    return v;
}
template <bool STAGED>
RTW_DEV uint32_t ldb_u8(const SceneView& sv, uint32_t off) {
    uint32_t v;
    if constexpr (STAGED) {
        asm("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(sv.sbase + off));
    } else {
        v = __ldg(sv.g + off);
    }
    return v;
}
template <bool STAGED>
RTW_DEV double ldb_f64(const SceneView& sv, uint32_t off) {
    uint2 v = ldb2<STAGED>(sv, off);
    return __hiloint2double((int)v.y, (int)v.x);
}
template <int SM> RTW_DEV float4 hot4(const SceneView& sv, uint32_t off) { return ldb4<(SM >= 1)>(sv, off); }
template <int SM> RTW_DEV uint32_t hot1(const SceneView& sv, uint32_t off) { return ldb1<(SM >= 1)>(sv, off); }
template <int SM> RTW_DEV uint2 hot2(const SceneView& sv, uint32_t off) { return ldb2<(SM >= 1)>(sv, off); }
template <int SM> RTW_DEV float4 cold4(const SceneView& sv, uint32_t off) { return ldb4<(SM >= 2)>(sv, off); }
template <int SM> RTW_DEV uint32_t cold1(const SceneView& sv, uint32_t off) { return ldb1<(SM >= 2)>(sv, off); }
template <int SM> RTW_DEV uint32_t cold_u8(const SceneView& sv, uint32_t off) { return ldb_u8<(SM >= 2)>(sv, off); }
template <int SM> RTW_DEV double cold_f64(const SceneView& sv, uint32_t off) { return ldb_f64<(SM >= 2)>(sv, off); }
RTW_DEV int fi(float f) { return __float_as_int(f); }

template <int SM>
RTW_DEV DevXf ld_xf(const SceneView& sv, uint32_t off) {
    float4 a = hot4<SM>(sv, off);
    DevXf x;
    x.c = a.x, x.s = a.y, x.tx = a.z, x.ty = a.w;
    x.tz = __uint_as_float(hot1<SM>(sv, off + 16));
    return x;
}
template <int SM>
RTW_DEV DevItem ld_item(const SceneView& sv, uint32_t off) {
    float4 a = hot4<SM>(sv, off), b = hot4<SM>(sv, off + 16);
    DevItem it;
    it.type = fi(a.x), it.xf = fi(a.y), it.first = fi(a.z), it.end = fi(a.w);
    it.material = fi(b.x), it.flip = fi(b.y), it.density_eff = b.z, it.root4 = fi(b.w);
    return it;
}
template <int SM>
RTW_DEV DevLight ld_light(const SceneView& sv, uint32_t off) {
    float4 a = hot4<SM>(sv, off), b = hot4<SM>(sv, off + 16);
    DevLight l;
    l.a0 = a.x, l.a1 = a.y, l.a2 = a.z, l.a3 = a.w;
    l.k = b.x, l.area = b.y, l.type = fi(b.z), l.pad = fi(b.w);
    return l;
}
template <int SM>
RTW_DEV DevMat ld_mat(const SceneView& sv, uint32_t off) {
    float4 a = cold4<SM>(sv, off), b = cold4<SM>(sv, off + 16);
    DevMat m;
    m.r = a.x, m.g = a.y, m.b = a.z, m.param = a.w;
    m.type = fi(b.x), m.texture = fi(b.y), m.tex_is_constant = fi(b.z), m.pad = 0;
    return m;
}
template <int SM>
RTW_DEV DevTex ld_tex(const SceneView& sv, uint32_t off) {
    float4 a = cold4<SM>(sv, off), b = cold4<SM>(sv, off + 16);
    uint2 c = ldb2<(SM >= 2)>(sv, off + 32);
    DevTex t;
    t.r = a.x, t.g = a.y, t.b = a.z, t.scale = a.w;
    t.type = fi(b.x), t.even = fi(b.y), t.odd = fi(b.z), t.image_offset = fi(b.w);
    t.nx = (int)c.x, t.ny = (int)c.y, t.pad0 = t.pad1 = 0;
    return t;
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
    RTW_EV(EVC_SPHERE_TEST, 1);
    if (!(disc > 0.0f)) return false;
    RTW_EV(EVC_SPHERE_ROOT, 1);
    float sq = sqrtf(disc);
    float c = dot(oc, oc) - r2;
    // q = -b - sign(b) sqrt(disc) never cancels; the roots are q/a and c/q
    float q = -(b + copysignf(sq, b));
    float ta = q * inv_a, tb = c * rcp_fast(q);
    bool neg = b < 0.0f;  // q > 0: q/a is the far root
    t0 = neg ? tb : ta;
    t1 = neg ? ta : tb;
    return true;
}
template <int SM>
RTW_DEV bool sphere_roots_f64(const SceneView& sv, uint32_t pd_off, V3 o, V3 d, float inv_a, float& t0, float& t1) {
    double cx = cold_f64<SM>(sv, pd_off), cy = cold_f64<SM>(sv, pd_off + 8), cz = cold_f64<SM>(sv, pd_off + 16);
    double rad = cold_f64<SM>(sv, pd_off + 24);
    double ocx = (double)o.x - cx, ocy = (double)o.y - cy, ocz = (double)o.z - cz;
    double dx = d.x, dy = d.y, dz = d.z;
    double a = dx * dx + dy * dy + dz * dz;
    double b = ocx * dx + ocy * dy + ocz * dz;
    double c = ocx * ocx + ocy * ocy + ocz * ocz - rad * rad;
    double disc = b * b - a * c;
    RTW_EV(EVC_SPHERE_TEST, 1);
    if (!(disc > 0.0)) return false;
    RTW_EV(EVC_SPHERE_ROOT, 1);
    // FP64 only where FP32 cancels (b*b - a*c and -b -+ sq for |r| >= 100); the division by a = d.d is the ray's FP32
    // reciprocal (2 ulp), which keeps two inlined FP64 divisions out of every copy of the sphere loop
    double sq = sqrt(disc);
    t0 = (float)(-b - sq) * inv_a;
    t1 = (float)(-b + sq) * inv_a;
    return true;
}
// Pick the root the reference would return; false if neither lies in (tmin, tmax).
RTW_DEV bool sphere_pick(float t0, float t1, float tmin, float tmax, float& t) {
    bool ok0 = (t0 < tmax) & (t0 > tmin);
    bool ok1 = (t1 < tmax) & (t1 > tmin);
    t = ok0 ? t0 : t1;
    return ok0 | ok1;
}

// x?_rect::hit (hittable.h:149-165, 184-200, 241-257): INCLUSIVE bounds; a NaN t passes like the reference (Q39).
// (on, dn) are the ray's origin/direction along the plane normal, (oa, da), (ob, db) along the two in-plane axes;
// inv_n = 1/dn is shared by every rect of that orientation in the group.  (k - on) * inv_n is exactly 0 for a ray
// that starts on the plane and +-inf / NaN for dn == 0 exactly as the division is.
// The six ordered comparisons are chained into one predicate (FSETP with a predicate operand) and the running
// closest hit is updated under it: 4 FP32 ops + 6 FSETP + 2 predicated moves per rect.
RTW_DEV void rect_update(float4 ra, float k, int id, float on, float inv_n, float oa, float da, float ob, float db, float tmin,
                         float& t_best, int& prim_best) {
    float t = (k - on) * inv_n;
    float a = fmaf(t, da, oa), b = fmaf(t, db, ob);
    RTW_EV(EVC_RECT, 1);
    // two independent 3-deep predicate chains (t range + first axis | second axis) joined at the end: one more
    // instruction than a single 6-deep chain, half the dependent latency
    asm("{\n\t.reg .pred p, q;\n\t"
        "setp.lt.ftz.f32 p, %2, %4;\n\t"
        "setp.lt.ftz.f32 q, %5, %7;\n\t"
        "setp.gt.or.ftz.f32 p, %2, %0, p;\n\t"
        "setp.gt.or.ftz.f32 q, %5, %8, q;\n\t"
        "setp.lt.or.ftz.f32 p, %6, %9, p;\n\t"
        "setp.gt.or.ftz.f32 q, %6, %10, q;\n\t"
        "or.pred p, p, q;\n\t"
        "@!p mov.f32 %0, %2;\n\t"
        "@!p mov.b32 %1, %3;\n\t}"
        : "+f"(t_best), "+r"(prim_best)
        : "f"(t), "r"(id), "f"(tmin), "f"(a), "f"(b), "f"(ra.x), "f"(ra.y), "f"(ra.z), "f"(ra.w));
}

// aabb::hit (aabb.h:17-47) with std::max/std::min written out so NaN behaves as in the reference (Q32).
// `t_entry` = the interval's lower end after the three slabs (orders the two children of a node).
RTW_DEV bool aabb_hit(float4 lo, float4 hi, V3 o, V3 inv, float tmin, float tmax, float& t_entry) {
    bool ok;
    {
        float t0 = (lo.x - o.x) * inv.x, t1 = (hi.x - o.x) * inv.x;
        if (inv.x < 0.0f) { float s = t0; t0 = t1; t1 = s; }
        tmin = (t0 < tmin) ? tmin : t0;
        tmax = (tmax < t1) ? tmax : t1;
        ok = !(tmax <= tmin);
    }
    {
        float t0 = (lo.y - o.y) * inv.y, t1 = (hi.y - o.y) * inv.y;
        if (inv.y < 0.0f) { float s = t0; t0 = t1; t1 = s; }
        tmin = (t0 < tmin) ? tmin : t0;
        tmax = (tmax < t1) ? tmax : t1;
        ok = ok & !(tmax <= tmin);
    }
    {
        float t0 = (lo.z - o.z) * inv.z, t1 = (hi.z - o.z) * inv.z;
        if (inv.z < 0.0f) { float s = t0; t0 = t1; t1 = s; }
        tmin = (t0 < tmin) ? tmin : t0;
        tmax = (tmax < t1) ? tmax : t1;
        ok = ok & !(tmax <= tmin);
    }
    t_entry = tmin;
    return ok;
}

// Conservative slab test for the trees this backend builds itself (padded boxes, rounded outwards): min/max form, the
// reference's NaN conventions are not needed — fminf/fmaxf drop a NaN slab (0 * inf), which only widens the interval.
RTW_DEV bool slab_hit(float4 lo, float4 hi, V3 inv, V3 oi, float tmin, float tmax, float& t_entry) {
    const float x0 = fmaf(lo.x, inv.x, -oi.x), x1 = fmaf(hi.x, inv.x, -oi.x);
    const float y0 = fmaf(lo.y, inv.y, -oi.y), y1 = fmaf(hi.y, inv.y, -oi.y);
    const float z0 = fmaf(lo.z, inv.z, -oi.z), z1 = fmaf(hi.z, inv.z, -oi.z);
    const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), tmin));
    const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), tmax));
    t_entry = tn;
    return tn <= tf;
}

// ------------------------------------------------------------------------------------------------ closest hit
// Result of world.hit(): winning primitive (or medium item) and t.  Normal/uv/p are reconstructed afterwards
// from the id (complete_hit), so the search loop carries two registers.
struct HitId {
    float t;
    int prim;  // >= 0 primitive; -1 miss; <= -2: medium item -2 - prim
    int item;  // index of the item that owns the hit (its xf applies)
};

// Per-ray constants of the primitive tests: reciprocal direction (rect planes, aabb slabs), a = d.d and 1/a (spheres).
struct RayK {
    V3 o, d, inv;
    float a, inv_a;
};
RTW_DEV RayK make_rayk(V3 o, V3 d) {
    RayK k;
    k.o = o, k.d = d;
    k.inv = mk(rcp_fast(d.x), rcp_fast(d.y), rcp_fast(d.z));  // aabb.h:22 (1.0 / direction)
    k.a = dot(d, d);
    k.inv_a = rcp_fast(k.a);
    return k;
}
// Ray constants in the object space of a translate/rotate_y chain: y is untouched by rotate_y, so 1/d.y carries over.
RTW_DEV RayK xf_rayk(const DevXf& x, const RayK& w) {
    RayK k;
    RTW_EV(EVC_XF, 1);
    xf_ray(x, w.o, w.d, k.o, k.d);
    k.inv = mk(rcp_fast(k.d.x), w.inv.y, rcp_fast(k.d.z));
    k.a = dot(k.d, k.d);
    k.inv_a = rcp_fast(k.a);
    return k;
}

// hittable_list::hit (hittable_list.h:16-24) over the groups at byte offsets [gfirst, gend): tmax shrinks to the running
// closest hit — rects accept t == closest (a later object wins a tie), spheres need t < closest (Q38).  Inside a group
// the records are ordered spheres, moving spheres, xy, xz, yz rects, each in list order, which evaluates the same
// winner as the reference's interleaved order (see device_scene.hpp).  The second pass of hittable_list::hit
// (:26-34) re-selects the same winner for deterministic children.
// All loop bounds and record addresses are warp-uniform: one broadcast load per 16 bytes.  Loops are kept rolled
// (`unroll 1`): trip counts are tiny (Cornell: 1-3) and unrolled remainders cost more branches than they save.
// Loads at an ABSOLUTE address: for staged sections the 32-bit shared address (sbase already added, so the inner loops
// advance one register and every LDS is [reg + imm]); for unstaged ones the blob offset.
template <int SM>
RTW_DEV uint32_t hot_abs(const SceneView& sv, uint32_t off) { return (SM >= 1) ? sv.sbase + off : off; }
template <int SM>
RTW_DEV float4 hot4_abs(const SceneView& sv, uint32_t addr) {
    float4 v;
    if constexpr (SM >= 1) {
        asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    } else {
        v = __ldg(reinterpret_cast<const float4*>(sv.g + addr));
    }
    return v;
}

// One slab test for the six rects of a reference `box` (hittable_list.h:69-104; record layout in device_scene.hpp).
// The nearest of the six rect hits with t in [tmin, t_best] is the entry face if the ray starts before it, else the
// exit face; t = (k - o_n) * inv_n is the expression of rect_update, so the winner's t is bit-identical.  The rect ids are
// id0 + {0: z = hi, 1: z = lo, 2: y = hi, 3: y = lo, 4: x = hi, 5: x = lo} (builder.cpp K_BOX).
RTW_DEV void box_update(float4 lo, float4 hi, const RayK& k, float tmin, float& t_best, int& prim_best) {
    RTW_EV(EVC_BOX, 1);
    const float tx0 = (lo.x - k.o.x) * k.inv.x, tx1 = (hi.x - k.o.x) * k.inv.x;
    const float ty0 = (lo.y - k.o.y) * k.inv.y, ty1 = (hi.y - k.o.y) * k.inv.y;
    const float tz0 = (lo.z - k.o.z) * k.inv.z, tz1 = (hi.z - k.o.z) * k.inv.z;
    const float nx = fminf(tx0, tx1), fx = fmaxf(tx0, tx1);
    const float ny = fminf(ty0, ty1), fy = fmaxf(ty0, ty1);
    const float nz = fminf(tz0, tz1), fz = fmaxf(tz0, tz1);
    const float te = fmaxf(fmaxf(nx, ny), nz), tf = fminf(fminf(fx, fy), fz);
    const bool entry = te >= tmin;
    const float t = entry ? te : tf;
    if (te <= tf && t >= tmin && t <= t_best) {
        // the face: the axis that produced t; on that axis the lo plane is the near one iff t?0 <= t?1
        const float ax = entry ? nx : fx, ay = entry ? ny : fy;
        const bool on_x = ax == t, on_y = ay == t;
        const bool lo_near = on_x ? (tx0 <= tx1) : (on_y ? (ty0 <= ty1) : (tz0 <= tz1));
        const int axis_base = on_x ? 4 : (on_y ? 2 : 0);
        t_best = t;
        prim_best = fi(hi.w) + axis_base + ((lo_near == entry) ? 1 : 0);
    }
}

// One slab test for the walls of a "room": >= 3 rects that are faces of one box, at most one per face (api.cu find_room; the
// Cornell room is five of six).  Record = box record with a.w != 0, followed by the table of the six face ids in box face
// order {z = hi, z = lo, y = hi, y = lo, x = hi, x = lo}, -1 = no such wall.  The nearest wall hit with t in [tmin, t_best]
// is the ENTRY face if the ray starts before it and that wall exists, else the EXIT face (the box is convex: no other face
// can lie between); t is the expression of rect_update, so it is bit-identical.  At an edge (two planes at the same t) the
// wall that is later in the list wins, like the rect loop's `t <= t_best` in list order.
template <int SM>
RTW_DEV void room_update(const SceneView& sv, uint32_t ids_addr, float4 lo, float4 hi, const RayK& k, float tmin, float& t_best,
                         int& prim_best) {
    RTW_EV(EVC_BOX, 1);
    const float tx0 = (lo.x - k.o.x) * k.inv.x, tx1 = (hi.x - k.o.x) * k.inv.x;
    const float ty0 = (lo.y - k.o.y) * k.inv.y, ty1 = (hi.y - k.o.y) * k.inv.y;
    const float tz0 = (lo.z - k.o.z) * k.inv.z, tz1 = (hi.z - k.o.z) * k.inv.z;
    const float nx = fminf(tx0, tx1), fx = fmaxf(tx0, tx1);
    const float ny = fminf(ty0, ty1), fy = fmaxf(ty0, ty1);
    const float nz = fminf(tz0, tz1), fz = fmaxf(tz0, tz1);
    const float te = fmaxf(fmaxf(nx, ny), nz), tf = fminf(fminf(fx, fy), fz);
    if (!(te <= tf) || !(tf >= tmin) || !(te <= t_best)) return;  // past the box, behind the ray, or beyond the closest hit
    const float4 ia = hot4_abs<SM>(sv, ids_addr), ib = hot4_abs<SM>(sv, ids_addr + 16);
    const bool xl = tx0 <= tx1, yl = ty0 <= ty1, zl = tz0 <= tz1;  // the lo plane is the near one on this axis
    const int x_near = fi(xl ? ib.y : ib.x), x_far = fi(xl ? ib.x : ib.y);
    const int y_near = fi(yl ? ia.w : ia.z), y_far = fi(yl ? ia.z : ia.w);
    const int z_near = fi(zl ? ia.y : ia.x), z_far = fi(zl ? ia.x : ia.y);
    int id = -1;
    float t = te;
    if (te >= tmin) id = max(max(nx == te ? x_near : -1, ny == te ? y_near : -1), nz == te ? z_near : -1);
    if (id < 0) {  // started inside, or entered through an open face
        t = tf;
        id = max(max(fx == tf ? x_far : -1, fy == tf ? y_far : -1), fz == tf ? z_far : -1);
    }
    if (id >= 0 && t <= t_best) {
        t_best = t;
        prim_best = id;
    }
}

// The five kind loops of one group, for ray constants `k` already in the group's frame.  BOXREC (render kernels): a group
// that carries a box record is tested with it instead of its six rects.
template <uint32_t F, int SM, bool BOXREC>
RTW_DEV void test_group_kinds(const SceneView& sv, float4 g0, float4 g1, const RayK& k, float time, float tmin, float& t_best,
                              int& prim_best) {
    const uint32_t sph0 = hot_abs<SM>(sv, (uint32_t)fi(g0.y)), sph1 = hot_abs<SM>(sv, (uint32_t)fi(g0.z));
    const uint32_t msph1 = hot_abs<SM>(sv, (uint32_t)fi(g0.w));
    const uint32_t xy1 = hot_abs<SM>(sv, (uint32_t)fi(g1.x)), xz1 = hot_abs<SM>(sv, (uint32_t)fi(g1.y));
    const uint32_t yz1 = hot_abs<SM>(sv, (uint32_t)fi(g1.z));
    if (BOXREC) {
        const uint32_t box1 = hot_abs<SM>(sv, (uint32_t)fi(g1.w));
        if (box1 != yz1) {  // warp-uniform
#pragma unroll 1
            for (uint32_t r = yz1; r < box1; r += 32) {
                const float4 lo = hot4_abs<SM>(sv, r), hi = hot4_abs<SM>(sv, r + 16);
                if (fi(lo.w) != 0) {  // warp-uniform: a room record, its face-id table in the next slot
                    room_update<SM>(sv, r + 32, lo, hi, k, tmin, t_best, prim_best);
                    r += 32;
                } else {
                    box_update(lo, hi, k, tmin, t_best, prim_best);
                }
            }
            return;
        }
    }
#pragma unroll 1
    for (uint32_t r = sph0; r < sph1; r += 32) {
        float4 ra = hot4_abs<SM>(sv, r), rb = hot4_abs<SM>(sv, r + 16);
        int id = fi(rb.w);
        float t0, t1, t;
        bool roots;
        if ((F & F_PRECISE) && (id & REC_PRECISE)) {
            roots = sphere_roots_f64<SM>(sv, (uint32_t)fi(rb.x), k.o, k.d, k.inv_a, t0, t1);
            id &= ~REC_PRECISE;
        } else {
            roots = sphere_roots(k.o - mk(ra.x, ra.y, ra.z), k.d, k.a, k.inv_a, ra.w, t0, t1);
        }
        if (roots && sphere_pick(t0, t1, tmin, t_best, t)) {
            t_best = t;
            prim_best = id;
        }
    }
    if (F & F_MOVING) {
#pragma unroll 1
        for (uint32_t r = sph1; r < msph1; r += 32) {
            float4 ra = hot4_abs<SM>(sv, r), rb = hot4_abs<SM>(sv, r + 16);
            // sphere.h:22-25 folded to c' + time*delta on the host
            V3 c = madd(mk(ra.x, ra.y, ra.z), time, mk(rb.x, rb.y, rb.z));
            RTW_EV(EVC_MOVING, 1);
            float t0, t1, t;
            if (sphere_roots(k.o - c, k.d, k.a, k.inv_a, ra.w, t0, t1) && sphere_pick(t0, t1, tmin, t_best, t)) {
                t_best = t;
                prim_best = fi(rb.w);
            }
        }
    }
#pragma unroll 1
    for (uint32_t r = msph1; r < xy1; r += 32) {  // xy_rect: plane z = k
        float4 ra = hot4_abs<SM>(sv, r), rb = hot4_abs<SM>(sv, r + 16);
        rect_update(ra, rb.x, fi(rb.w), k.o.z, k.inv.z, k.o.x, k.d.x, k.o.y, k.d.y, tmin, t_best, prim_best);
    }
#pragma unroll 1
    for (uint32_t r = xy1; r < xz1; r += 32) {  // xz_rect: plane y = k
        float4 ra = hot4_abs<SM>(sv, r), rb = hot4_abs<SM>(sv, r + 16);
        rect_update(ra, rb.x, fi(rb.w), k.o.y, k.inv.y, k.o.x, k.d.x, k.o.z, k.d.z, tmin, t_best, prim_best);
    }
#pragma unroll 1
    for (uint32_t r = xz1; r < yz1; r += 32) {  // yz_rect: plane x = k
        float4 ra = hot4_abs<SM>(sv, r), rb = hot4_abs<SM>(sv, r + 16);
        rect_update(ra, rb.x, fi(rb.w), k.o.x, k.inv.x, k.o.y, k.d.y, k.o.z, k.d.z, tmin, t_best, prim_best);
    }
}

template <uint32_t F, int SM, bool BOXREC>
RTW_DEV void test_groups_body(const SceneView& sv, uint32_t gfirst, uint32_t gend, const RayK& rk, float time, float tmin,
                              float& t_best, int& prim_best) {
    const uint32_t ga0 = hot_abs<SM>(sv, gfirst), ga1 = hot_abs<SM>(sv, gend);
#pragma unroll 1
    for (uint32_t ga = ga0; ga < ga1; ga += 32) {
        const float4 g0 = hot4_abs<SM>(sv, ga), g1 = hot4_abs<SM>(sv, ga + 16);
        // two instantiations of the kind loops instead of `k = xf ? xf_rayk(rk) : rk`: no 11-register copy per group.
        // The all-features variant keeps ONE copy: its kernels are > 100 KB of SASS and instruction-cache bound.
        if constexpr ((F & F_MEDIA) != 0) {
            RayK k = rk;
            if ((F & F_XF) && fi(g0.x) >= 0) k = xf_rayk(ld_xf<SM>(sv, (uint32_t)fi(g0.x)), rk);
            test_group_kinds<F, SM, BOXREC>(sv, g0, g1, k, time, tmin, t_best, prim_best);
        } else if ((F & F_XF) && fi(g0.x) >= 0) {
            const RayK k = xf_rayk(ld_xf<SM>(sv, (uint32_t)fi(g0.x)), rk);
            test_group_kinds<F, SM, BOXREC>(sv, g0, g1, k, time, tmin, t_best, prim_best);
        } else {
            test_group_kinds<F, SM, BOXREC>(sv, g0, g1, rk, time, tmin, t_best, prim_best);
        }
    }
}

// The all-features kernels call ONE out-of-line copy of the primitive loops from their three call sites (world lists,
// BVH leaves, medium boundaries): they are instruction-fetch bound (ncu: 42 % of stall samples `no_instruction` on the
// Book-2 final scene with the loops inlined three times); the small variants keep them inline.
struct GroupHit {
    float t;
    int prim;
};
template <uint32_t F, int SM, bool BOXREC>
__device__ __noinline__ GroupHit test_groups_call(SceneView sv, uint32_t gfirst, uint32_t gend, RayK rk, float time, float tmin,
                                                  float t_best, int prim_best) {
    test_groups_body<F, SM, BOXREC>(sv, gfirst, gend, rk, time, tmin, t_best, prim_best);
    return GroupHit{t_best, prim_best};
}
template <uint32_t F, int SM, bool BOXREC>
RTW_DEV void test_groups(const SceneView& sv, uint32_t gfirst, uint32_t gend, const RayK& rk, float time, float tmin,
                         float& t_best, int& prim_best) {
    if constexpr ((F & F_MEDIA) != 0 && SM < 2) {  // big scenes only: a small staged scene (Cornell smoke) loses 19 % to the calls
        const GroupHit g = test_groups_call<F, SM, BOXREC>(sv, gfirst, gend, rk, time, tmin, t_best, prim_best);
        t_best = g.t, prim_best = g.prim;
    } else {
        test_groups_body<F, SM, BOXREC>(sv, gfirst, gend, rk, time, tmin, t_best, prim_best);
    }
}

}  // namespace rtw
