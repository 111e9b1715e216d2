// rt_path.cuh — one color() invocation (RayTracingWeekend.cpp:45-160) as an iterative path step.
//
// The reference recursion  L = e1 + w1*(e2 + w2*(...))  is unrolled with a throughput register:
// radiance only enters at terminal events (an emitter, the background, the Normal debug view), so
// `L += beta * e` at those events and `beta *= w` at every scatter is algebraically the same value.
#pragma once
#include "rt_shade.cuh"

namespace rtw {

// Returns true when the path continues with the new `ray`/`beta`; false when it ended (L updated).
// `bounce` is the 1-based index of this color() invocation (Philox counter word).
// `shade_hit` is everything of color() after world.hit() (:54-159); `path_step` = world.hit() + shade_hit.  The
// wavefront pipeline runs the two halves in different kernels.
template <uint32_t F, int SM>
RTW_DEV bool shade_hit(const DevScene& sc, const SceneView& sv, const Rng& rng, uint32_t bounce, const HitId& h, Ray& ray,
                       V3& beta, V3& L, bool need_main = true) {
    // The bounce's main uniforms are drawn here, before control flow diverges by material: every lane of the
    // warp runs the ten Philox rounds together instead of once per material branch.
    // `need_main` = false (warp-uniform) when the caller already knows that no lane consumes them (sorted.cu: units of
    // metal, isotropic or terminal hits).
    U4 rmain = U4{0.0f, 0.0f, 0.0f, 0.0f};
    if (need_main) rmain = rng.draw(bounce, SITE_MAIN);
    if (h.prim == -1) {  // :141-159
        L = L + beta * background(sc, ray.d);
        return false;
    }
    // material first: whether u,v are needed depends on its texture
    uint32_t pm;
    const int mat_idx = hit_material<SM>(sc, sv, h, pm);
    const DevMat m = ld_mat<SM>(sv, sc.off[SEC_MATS] + (uint32_t)mat_idx * 32u);
    const bool need_uv = (F & F_TEXTURES) && !m.tex_is_constant &&
                         (m.type == RTW_MAT_LAMBERTIAN || m.type == RTW_MAT_DIFFUSE_LIGHT || m.type == RTW_MAT_ISOTROPIC);
    const HitRec rec = complete_hit<F, SM>(sc, sv, h, pm, ray.o, ray.d, ray.time, need_uv);

    if (sc.render_type == RTW_RENDER_NORMAL) {  // :135-136
        L = L + beta * (0.5f * (rec.n + mk(1.0f, 1.0f, 1.0f)));
        return false;
    }

    switch (m.type) {
        case RTW_MAT_DIFFUSE_LIGHT: {  // material.h:232-244: no scatter; one-sided emission (Q23)
            if (dot(rec.n, ray.d) > 0.0f) L = L + beta * material_texture<F, SM>(sc, sv, m, rec.u, rec.v, rec.p);
            return false;  // :62-63
        }
        case RTW_MAT_METAL: {  // material.h:128-136 -> specular branch :114-115
            V3 unit = normalize(ray.d);
            V3 reflected = unit - (2.0f * dot(unit, rec.n)) * rec.n;
            if (m.param != 0.0f) {  // the ball draw is skipped when fuzz == 0 (Q24)
                U4 b = rng.draw(bounce, SITE_BALL);
                reflected = reflected + m.param * random_in_unit_ball(b.x, b.y, b.z);
            }
            beta = beta * mk(m.r, m.g, m.b);
            ray.o = rec.p;
            ray.d = reflected;
            return true;
        }
        case RTW_MAT_DIELECTRIC: {  // material.h:146-222
            float ref_idx = m.param;
            float ddn = dot(ray.d, rec.n);
            float len = sqrtf(dot(ray.d, ray.d));
            V3 outward;
            float ni_over_nt, cosine;
            if (ddn > 0.0f) {  // inside -> outside (:153-184)
                outward = -rec.n;
                ni_over_nt = ref_idx;
                cosine = ddn / len;
                cosine = sqrtf(1.0f - ref_idx * ref_idx * (1.0f - cosine * cosine));
            } else {  // :185-192
                outward = rec.n;
                ni_over_nt = 1.0f / ref_idx;
                cosine = -ddn / len;
            }
            V3 reflected = ray.d - (2.0f * ddn) * rec.n;  // :194 reflect() of the UN-normalised direction (Q25)
            // refract (material.h:17-39)
            V3 uv = ray.d * (1.0f / len);
            float dt = dot(uv, outward);
            float disc = 1.0f - ni_over_nt * ni_over_nt * (1.0f - dt * dt);
            float reflect_prob = 1.0f;
            V3 refracted = reflected;
            if (disc > 0.0f) {
                refracted = ni_over_nt * (uv - outward * dt) - outward * sqrtf(disc);
                float r0 = (1.0f - ref_idx) / (1.0f + ref_idx);  // schlick (material.h:44-49)
                r0 = r0 * r0;
                float om = 1.0f - cosine;
                float om2 = om * om;
                reflect_prob = r0 + (1.0f - r0) * (om2 * om2 * om);
            }
            float u = rmain.x;  // :209-212 (drawn even on total internal reflection)
            ray.o = rec.p;
            ray.d = (u < reflect_prob) ? reflected : refracted;
            return true;  // attenuation (1,1,1)
        }
        case RTW_MAT_ISOTROPIC: {  // material.h:257-262
            U4 b = rng.draw(bounce, SITE_BALL);
            beta = beta * material_texture<F, SM>(sc, sv, m, rec.u, rec.v, rec.p);
            ray.o = rec.p;
            ray.d = random_in_unit_ball(b.x, b.y, b.z);
            return true;
        }
        default: {  // lambertian: material.h:81-119 + mixture importance sampling :117-132
            const V3 att = material_texture<F, SM>(sc, sv, m, rec.u, rec.v, rec.p);
            const V3 dir = mixture_generate<SM>(sc, sv, rec.p, rec.n, rmain);  // :120-122
            const V3 nd = normalize(dir);
            const float cos_w = dot(nd, normalize(rec.n));                // cosine_pdf: uvw.w = normalize(n) (pdf.h:18)
            const float pdf_cos = (cos_w <= 0.0f) ? 0.0f : cos_w / kPi;    // cosine_pdf::value (pdf.h:20-24)
            float pdf_val = pdf_cos;
            if (sc.n_lights > 0) pdf_val = 0.5f * pdf_cos + 0.5f * lights_pdf<SM>(sc, sv, rec.p, dir);  // pdf.h:64-67
            if (pdf_val <= 0.0f) return false;  // :126-127 (emitted is zero for a lambertian)
            const float cos_n = dot(rec.n, nd);
            const float spdf = (cos_n < 0.0f) ? 0.0f : cos_n / kPi;  // lambertian::scattering_pdf (material.h:115-119)
            beta = beta * att * (spdf * rcp_fast(pdf_val));      // :129-132
            ray.o = rec.p;
            ray.d = dir;  // NOT normalised (:123)
            return true;
        }
    }
}

// ------------------------------------------------------------------------------------------------ next-event estimation
// OPT-IN estimator (RTW_FLAG_NEE, wavefront pipeline only): changes RayTracingWeekend.cpp:117-132 — the
// reference draws ONE direction from the 50/50 mixture of the cosine lobe and the lights list and never casts an occlusion
// ray.  Here a lambertian hit (a) casts a SHADOW ray at a uniformly chosen point of the emitting xz_rect lights of the lights
// list and, if it arrives, adds  beta * albedo * (cos/pi) * Le / pdf_light  to the pixel, and (b) continues with a
// cosine-sampled direction (weight = albedo); a path that then runs into one of those lights by itself must not count its
// emission again (`skip_emit`), while emission reached through a specular bounce, from emitters outside the lights list, or
// seen by the camera is counted as before.  Same expectation as the reference's estimator, different variance.
constexpr uint32_t SITE_NEE = 0x100u;  // Philox draw site of the light sample (the medium sites are SITE_MEDIUM0 + index)

struct ShadowRay {
    V3 o, d, w;  // origin, direction to the light sample (not normalised), weight = beta * albedo * cos / (pi * pdf)
    bool cast;
};

// Emitting rect lights of the lights list = the set NEE samples.  Index of `prim` in it, or -1 (world-space xz_rects only).
template <int SM>
RTW_DEV bool nee_samples_prim(const DevScene& sc, const SceneView& sv, int prim, uint32_t pm) {
    if (prim < 0 || (int)(pm & PM_TYPE_MASK) != RTW_PRIM_XZ_RECT || ((pm >> PM_XF_SHIFT) & PM_XF_MASK)) return false;
    const float4 pa = cold4<SM>(sv, sc.off[SEC_PA] + (uint32_t)prim * 16u);
    const float k = __uint_as_float(cold1<SM>(sv, sc.off[SEC_PB] + (uint32_t)prim * 16u));
    for (int i = 0; i < sc.n_lights; i++) {
        const DevLight l = ld_light<SM>(sv, sc.off[SEC_LIGHTS] + (uint32_t)i * 32u);
        if (l.type == RTW_LIGHT_XZ_RECT && l.a0 == pa.x && l.a1 == pa.y && l.a2 == pa.z && l.a3 == pa.w && l.k == k) return true;
    }
    return false;
}
RTW_DEV int nee_light_count(const DevScene& sc) { return sc.n_nee_lights; }

// Lambertian hit under NEE: fills the shadow ray, replaces `ray` by the cosine-sampled continuation and updates beta.
template <uint32_t F, int SM>
RTW_DEV void shade_lambert_nee(const DevScene& sc, const SceneView& sv, const Rng& rng, uint32_t bounce, const HitId& h, Ray& ray, V3& beta,
                               ShadowRay& sh) {
    const U4 rmain = rng.draw(bounce, SITE_MAIN), rl = rng.draw(bounce, SITE_NEE);
    uint32_t pm;
    const int mat_idx = hit_material<SM>(sc, sv, h, pm);
    const DevMat m = ld_mat<SM>(sv, sc.off[SEC_MATS] + (uint32_t)mat_idx * 32u);
    const bool need_uv = (F & F_TEXTURES) && !m.tex_is_constant;
    const HitRec rec = complete_hit<F, SM>(sc, sv, h, pm, ray.o, ray.d, ray.time, need_uv);
    const V3 att = material_texture<F, SM>(sc, sv, m, rec.u, rec.v, rec.p);
    // (a) the light sample: the pick-th emitting rect of the lights list, a uniform point on it (hittable.h:224-228)
    sh.cast = false;
    const int nl = sc.n_nee_lights;
    int pick = min(nl - 1, (int)((float)nl * rl.x));
    float pdf = 0.0f;
    V3 dir = mk(0.0f, 0.0f, 0.0f);
    const uint32_t l0 = sc.off[SEC_LIGHTS];
    for (int i = 0, seen = 0; i < sc.n_lights; i++) {  // the pick
        const DevLight l = ld_light<SM>(sv, l0 + (uint32_t)i * 32u);
        if (l.type != RTW_LIGHT_XZ_RECT || !l.pad) continue;  // pad = 1: the rect's material emits (set at upload)
        if (seen++ == pick) dir = mk(fmaf(l.a1 - l.a0, rl.y, l.a0) - rec.p.x, l.k - rec.p.y, fmaf(l.a3 - l.a2, rl.z, l.a2) - rec.p.z);
    }
    const float len2 = dot(dir, dir), len = sqrtf(len2);
    for (int i = 0; i < sc.n_lights; i++) {  // its density under "a uniformly picked emitting rect, a uniform point on it"
        const DevLight l = ld_light<SM>(sv, l0 + (uint32_t)i * 32u);
        if (l.type == RTW_LIGHT_XZ_RECT && l.pad) pdf += light_rect_pdf(l, rec.p, dir, len2, len);
    }
    pdf *= 1.0f / (float)nl;
    const float cos_x = dot(rec.n, dir) * rcp_fast(len);
    if (pdf > 0.0f && cos_x > 0.0f) {
        sh.cast = true;
        sh.o = rec.p, sh.d = dir;
        sh.w = beta * att * (cos_x * rcp_fast(kPi * pdf));
    }
    // (b) the continuation: cosine lobe (pdf.h:26-29); scattering_pdf / pdf = 1, so the weight is the albedo
    const V3 out = mixture_generate<SM>(sc, sv, rec.p, rec.n, U4{0.0f, rmain.y, rmain.z, rmain.w});
    beta = beta * att;
    ray.o = rec.p;
    ray.d = out;
}

template <uint32_t F, int SM>
RTW_DEV bool path_step(const DevScene& sc, const SceneView& sv, const Rng& rng, uint32_t bounce, Ray& ray, V3& beta, V3& L,
                       int* stack, int stride) {
    // :52 world.hit(r, 0.001f, DBL_MAX, rec)
    const HitId h = closest_hit<F, SM, false>(sc, sv, ray.o, ray.d, ray.time, 0.001f, FLT_MAX, &rng, bounce, stack, stride);
    return shade_hit<F, SM>(sc, sv, rng, bounce, h, ray, beta, L);
}

}  // namespace rtw
