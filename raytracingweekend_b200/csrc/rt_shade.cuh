// rt_shade.cuh — textures, sampling utilities, pdfs, materials, camera, background (FP32 device forms).
#pragma once
#include "rt_trace.cuh"

namespace rtw {

constexpr float kPi = 3.14159265358979323846f;

// ------------------------------------------------------------------------------------------------ Perlin
// perlin::noise (noise.h:89-151, InterpolationType::PERLIN) + perlin_interp (noise.h:40-60) with the Hermite
// cubic `smooth` (noise.h:9-12, :19).
template <int SM>
RTW_DEV float perlin_noise(const DevScene& sc, const SceneView& sv, V3 p) {
    float fx = floorf(p.x), fy = floorf(p.y), fz = floorf(p.z);
    float u = p.x - fx, v = p.y - fy, w = p.z - fz;
    int i = (int)fx, j = (int)fy, k = (int)fz;
    float uu = u * u * (3.0f - 2.0f * u), vv = v * v * (3.0f - 2.0f * v), ww = w * w * (3.0f - 2.0f * w);
    const uint32_t px = sc.off[SEC_PERM], py = px + 256, pz = px + 512, rv = sc.off[SEC_RANVEC];
    float accum = 0.0f;
#pragma unroll
    for (int di = 0; di < 2; di++) {
        int hx = (int)cold_u8<SM>(sv, px + (uint32_t)((i + di) & 255));
        float wi = di ? uu : 1.0f - uu;
#pragma unroll
        for (int dj = 0; dj < 2; dj++) {
            int hy = (int)cold_u8<SM>(sv, py + (uint32_t)((j + dj) & 255));
            float wj = dj ? vv : 1.0f - vv;
#pragma unroll
            for (int dk = 0; dk < 2; dk++) {
                int hz = (int)cold_u8<SM>(sv, pz + (uint32_t)((k + dk) & 255));
                float wk = dk ? ww : 1.0f - ww;
                float4 c = cold4<SM>(sv, rv + (uint32_t)(hx ^ hy ^ hz) * 16u);
                float dp = fmaf(c.x, u - di, fmaf(c.y, v - dj, c.z * (w - dk)));
                accum = fmaf(wi * wj * wk, dp, accum);
            }
        }
    }
    return accum;
}
template <int SM>
RTW_DEV float perlin_turb(const DevScene& sc, const SceneView& sv, V3 p) {  // noise.h:74-86, depth 7
    float accum = 0.0f, weight = 1.0f;
#pragma unroll 1
    for (int i = 0; i < 7; i++) {  // rolled: seven unrolled octaves are 750 SASS instructions at every call site
        accum = fmaf(weight, perlin_noise<SM>(sc, sv, p), accum);
        weight *= 0.5f;
        p = p * 2.0f;
    }
    return fabsf(accum);
}

// sin(x) for texture arguments (|x| up to ~1e4): two-constant Cody-Waite reduction to [-pi, pi], then MUFU.SIN (2^-21
// absolute there).  libdevice sinf is ~40 instructions plus a slow path per call site, and texture_value is inlined at
// three call sites of shade_hit.
RTW_DEV float sin_reduced(float x) {
    const float k = rintf(x * 0.15915494309189535f);
    float r = fmaf(-k, 6.28318548202514648f, x);   // 2*pi rounded to FP32
    r = fmaf(-k, -1.74845553e-7f, r);              // 2*pi - fl(2*pi)
    return __sinf(r);
}

// texture::value(u, v, p) (texture.h:21-24, 35-46, 57-68, 81-94)
template <int SM>
RTW_DEV V3 texture_value(const DevScene& sc, const SceneView& sv, int tex, float u, float v, V3 p) {
#pragma unroll 1
    for (int guard = 0; guard < 16; guard++) {  // checker textures chain to their even/odd child
        DevTex t = ld_tex<SM>(sv, sc.off[SEC_TEXS] + (uint32_t)tex * 48u);
        if (t.type == RTW_TEX_CONSTANT) return mk(t.r, t.g, t.b);
        if (t.type == RTW_TEX_CHECKER) {
            float sines = sin_reduced(10.0f * p.x) * sin_reduced(10.0f * p.y) * sin_reduced(10.0f * p.z);
            tex = (sines < 0.0f) ? t.odd : t.even;
            continue;
        }
        if (t.type == RTW_TEX_NOISE) {
            float g = 0.5f * (1.0f + sin_reduced(fmaf(t.scale, p.z, 10.0f * perlin_turb<SM>(sc, sv, p))));
            return mk(g, g, g);
        }
        // image_texture (texture.h:81-94): nearest texel, v flipped
        int i = (int)(u * t.nx);
        int j = (int)((1.0f - v) * t.ny - 0.001f);
        i = max(min(i, t.nx - 1), 0);
        j = max(min(j, t.ny - 1), 0);
        const uint8_t* px = sv.texels + t.image_offset + 3 * i + 3 * t.nx * j;
        return mk(__ldg(px) / 255.0f, __ldg(px + 1) / 255.0f, __ldg(px + 2) / 255.0f);
    }
    return mk(0.0f, 0.0f, 0.0f);
}

// out-of-line copy for the big (unstaged) scenes, whose kernels are instruction-fetch bound: shade_hit evaluates textures
// at three call sites (emitter, isotropic, lambertian)
template <int SM>
__device__ __noinline__ V3 texture_value_call(const DevScene* sc, SceneView sv, int tex, float u, float v, V3 p) {
    return texture_value<SM>(*sc, sv, tex, u, v, p);
}
template <uint32_t F, int SM>
RTW_DEV V3 material_texture(const DevScene& sc, const SceneView& sv, const DevMat& m, float u, float v, V3 p) {
    if (!(F & F_TEXTURES) || m.tex_is_constant) return mk(m.r, m.g, m.b);
    if constexpr (SM < 2) return texture_value_call<SM>(&sc, sv, m.texture, u, v, p);
    return texture_value<SM>(sc, sv, m.texture, u, v, p);
}

// ------------------------------------------------------------------------------------------------ sampling
// sin and cos of 2*pi*u for u in [0, 1): the angle is shifted into [-pi, pi), where MUFU.SIN/COS are accurate to 2^-21
// absolute (sin(x + pi) = -sin x, cos(x + pi) = -cos x) — two MUFU ops instead of the ~30-instruction sincospif.
RTW_DEV void sincos_2pi(float u, float& s, float& c) {
    const float a = (u - 0.5f) * 6.283185307179586f;
    s = -__sinf(a);
    c = -__cosf(a);
}

struct Onb {  // onb.h:32-38
    V3 u, v, w;
};
// onb::build_from_w with the two possible helper axes folded in: cross(w, x^) = (0, w.z, -w.y), cross(w, y^) = (-w.z, 0, w.x).
RTW_DEV Onb onb_from_w(V3 n) {
    Onb b;
    b.w = normalize(n);
    const bool big = fabsf(b.w.x) > 0.9f;
    V3 c = big ? mk(-b.w.z, 0.0f, b.w.x) : mk(0.0f, b.w.z, -b.w.y);
    b.v = normalize(c);
    b.u = cross(b.w, b.v);
    return b;
}
RTW_DEV V3 onb_local(const Onb& b, V3 a) { return a.x * b.u + a.y * b.v + a.z * b.w; }  // onb.h:21-24

// random_in_unit_sphere (utility.h:27-35) without the rejection loop: uniform direction * cbrt(U) is the same
// distribution (uniform in the unit ball).
RTW_DEV V3 random_in_unit_ball(float u0, float u1, float u2) {
    float z = 1.0f - 2.0f * u0;
    float r = sqrtf(fmaxf(0.0f, 1.0f - z * z));
    float s, c;
    sincos_2pi(u1, s, c);
    float rad = cbrtf(u2);
    return mk(rad * r * c, rad * r * s, rad * z);
}

// ------------------------------------------------------------------------------------------------ light pdfs
// xz_rect::pdf_value (hittable.h:208-222): self-hit test with t in [0.001, +inf], no occlusion.
// `len2` = dot(v, v), `len` = sqrt(len2): distance^2 / (cosine * area) = t^2 len2 len / (|v.y| area).
RTW_DEV float light_rect_pdf(const DevLight& l, V3 o, V3 v, float len2, float len) {
    float t = (l.k - o.y) * rcp_fast(v.y);
    float x = fmaf(t, v.x, o.x), z = fmaf(t, v.z, o.z);
    bool miss = (t < 0.001f) | (t > INFINITY) | (x < l.a0) | (x > l.a1) | (z < l.a2) | (z > l.a3);
    float pdf = (t * t) * (len2 * len) * rcp_fast(fabsf(v.y) * l.area);
    return miss ? 0.0f : pdf;
}
// sphere_base::pdf_value (sphere.h:88-99)
RTW_DEV float light_sphere_pdf(const DevLight& l, V3 o, V3 v, float len2, float inv_len2) {
    V3 co = mk(l.a0 - o.x, l.a1 - o.y, l.a2 - o.z);
    float t0, t1, t;
    bool hit = sphere_roots(-co, v, len2, inv_len2, l.a3, t0, t1) && sphere_pick(t0, t1, 0.001f, INFINITY, t);
    float q = l.a3 * l.a3 * rcp_fast(dot(co, co));
    float one_minus_cmax = q * rcp_fast(1.0f + sqrtf(1.0f - q));  // NaN when the origin is inside, like the reference (Q17)
    float pdf = rcp_fast(2.0f * kPi * one_minus_cmax);
    return hit ? pdf : 0.0f;
}
// hittable_list::pdf_value (hittable_list.h:44-53)
template <int SM>
RTW_DEV float lights_pdf(const DevScene& sc, const SceneView& sv, V3 o, V3 v) {
    const float weight = 1.0f / (float)sc.n_lights;
    const float len2 = dot(v, v);
    const float inv_len2 = rcp_fast(len2), len = sqrtf(len2);
    float sum = 0.0f;
    const uint32_t l0 = sc.off[SEC_LIGHTS];
#pragma unroll 1
    for (int i = 0; i < sc.n_lights; i++) {
        DevLight l = ld_light<SM>(sv, l0 + (uint32_t)i * 32u);
        float pv = 0.0f;  // hittable.h:36 default
        if (l.type == RTW_LIGHT_XZ_RECT) pv = light_rect_pdf(l, o, v, len2, len);
        else if (l.type == RTW_LIGHT_SPHERE) pv = light_sphere_pdf(l, o, v, len2, inv_len2);
        sum = fmaf(weight, pv, sum);
    }
    return sum;
}

// mixture_pdf::generate (pdf.h:69-75) over cosine_pdf (pdf.h:26-29 -> utility.h:54-67) and hittable_pdf on the lights
// list (hittable_list.h:55-59 -> xz_rect::random hittable.h:224-228 | sphere_base::random sphere.h:101-108 ->
// utility.h:69-82 | hittable::random default hittable.h:37).
// The cosine lobe and the cone towards a sphere light are the same computation — a polar direction (z, sqrt(1-z^2))
// with azimuth 2*pi*r1 in an orthonormal frame about an axis — with different axis and z, so they run as ONE
// instruction stream with selected operands (one frame, one sincospi) instead of two divergent branches:
//   cosine:  axis = n,            z = sqrt(1 - r2),                  sin^2 = r2
//   sphere:  axis = centre - o,   z = 1 - r2 (1 - cos_max),          sin^2 = (1 - z)(1 + z)
// with 1 - cos_max = q / (1 + sqrt(1 - q)), q = radius^2 / |centre - o|^2 (no cancellation for far spheres).
// u: .x mixture choice, .y light pick (utility.h:22-25 with a = 0, b = n-1), .z = r1, .w = r2.
template <int SM>
RTW_DEV V3 mixture_generate(const DevScene& sc, const SceneView& sv, V3 o, V3 n, U4 u) {
    int ltype = -1;  // -1: cosine lobe
    DevLight l;
    l.a0 = l.a1 = l.a2 = l.a3 = l.k = 0.0f;
    if (sc.n_lights > 0 && !(u.x < 0.5f)) {
        int idx = min(sc.n_lights - 1, (int)((float)sc.n_lights * u.y));
        l = ld_light<SM>(sv, sc.off[SEC_LIGHTS] + (uint32_t)idx * 32u);
        ltype = l.type;
    }
    if (ltype == RTW_LIGHT_XZ_RECT) return mk(fmaf(l.a1 - l.a0, u.z, l.a0) - o.x, l.k - o.y, fmaf(l.a3 - l.a2, u.w, l.a2) - o.z);
    if (ltype == RTW_LIGHT_DEFAULT) return mk(1.0f, 0.0f, 0.0f);
    const bool sph = ltype == RTW_LIGHT_SPHERE;
    V3 axis = sph ? mk(l.a0 - o.x, l.a1 - o.y, l.a2 - o.z) : n;
    float q = sph ? l.a3 * l.a3 * rcp_fast(dot(axis, axis)) : u.w;
    float s1 = sqrtf(1.0f - q);
    float omz = u.w * q * rcp_fast(1.0f + s1);
    float z = sph ? 1.0f - omz : s1;
    float st = sqrtf(sph ? omz * (1.0f + z) : u.w);
    float sn, cs;
    sincos_2pi(u.z, sn, cs);
    Onb uvw = onb_from_w(axis);
    return onb_local(uvw, mk(cs * st, sn * st, z));
}

// random_in_unit_disk (camera.h:61-69) without the rejection loop: polar (sqrt(u), 2*pi*v) is uniform in the unit disk.
RTW_DEV void lens_disk(float u, float v, float& x, float& y) {
    const float rad = sqrtf(u);
    float sn, cs;
    sincos_2pi(v, sn, cs);
    x = rad * cs, y = rad * sn;
}

// ------------------------------------------------------------------------------------------------ camera
// RayTracingWeekend.cpp:227-231 + camera::get_ray (camera.h:36-50).  The lens disk is sampled directly
// (sqrt(u), 2*pi*v) instead of by rejection (camera.h:61-69) — same distribution — and skipped when aperture = 0.
RTW_DEV Ray camera_ray(const DevScene& sc, const DevRenderParams& rp, const Rng& rng, int i, int j) {
    const DevCamera& c = sc.cam;
    U4 r = rng.draw(0, SITE_MAIN);
    float s = ((float)i + r.x) / (float)rp.nx;
    float t = ((float)j + r.y) / (float)rp.ny;
    V3 origin = mk(c.ox, c.oy, c.oz);
    V3 offset = mk(0.0f, 0.0f, 0.0f);
    if (c.lens_radius > 0.0f) {
        U4 q = rng.draw(0, SITE_LENS);
        float dx, dy;
        lens_disk(q.x, q.y, dx, dy);
        offset = (c.lens_radius * dx) * mk(c.ux, c.uy, c.uz) + (c.lens_radius * dy) * mk(c.wx, c.wy, c.wz);
    }
    Ray ray;
    ray.time = fmaf(r.z, c.time1 - c.time0, c.time0);
    V3 dir = mk(c.llx, c.lly, c.llz) + s * mk(c.hx, c.hy, c.hz) + t * mk(c.vx, c.vy, c.vz) - origin - offset;
    ray.o = origin + offset;
    ray.d = normalize(dir);
    return ray;
}

// RayTracingWeekend.cpp:143-158
RTW_DEV V3 background(const DevScene& sc, V3 d) {
    if (sc.background_type == RTW_BG_GRADIENT) {
        V3 unit = normalize(d);
        float t = 0.5f * (unit.y + 1.0f);
        // lerp(from, to, t) = (1-t)*to + t*from (vec3.h:84-87): white at t = 0, (0.5, 0.7, 1) at t = 1
        return (1.0f - t) * mk(1.0f, 1.0f, 1.0f) + t * mk(0.5f, 0.7f, 1.0f);
    }
    return mk(0.0f, 0.0f, 0.0f);
}

}  // namespace rtw
