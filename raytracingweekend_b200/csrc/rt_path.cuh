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

template <uint32_t F, int SM>
RTW_DEV bool path_step(const DevScene& sc, const SceneView& sv, const Rng& rng, uint32_t bounce, Ray& ray, V3& beta, V3& L,
                       int* stack, int stride) {
    // :52 world.hit(r, 0.001f, DBL_MAX, rec)
    const HitId h = closest_hit<F, SM, false>(sc, sv, ray.o, ray.d, ray.time, 0.001f, FLT_MAX, &rng, bounce, stack, stride);
    return shade_hit<F, SM>(sc, sv, rng, bounce, h, ray, beta, L);
}

}  // namespace rtw
