// scenes.cpp — the reference's Scene/ functions composed on the builder C API.
//
// Each function mirrors one class of /root/reference/RayTracingWeekend/Scene/scene.h (cited), or one of
// the two scenes BASELINE.json names that the reference does not ship (cfg2, cfg3),
// composed from the same constructors.  Float literals keep their `f` suffix where the reference has
// one: the builder stores doubles and (double)0.1f != 0.1.
//
// Construction-time randomness is the reference's: a local default-seeded std::minstd_rand +
// std::uniform_real_distribution<double> (scene.h:103-104) from the same libstdc++, so the scenes are
// object-for-object those the reference builds in this image (tests/test_host_parity.py compares the
// flattened arrays byte for byte with the adapter-flattened reference objects).
#include <cmath>
#include <cstring>
#include <random>
#include <string>
#include <vector>

#include "flat.hpp"

namespace {

struct V3 {
    double x, y, z;
};

struct SceneCtx {
    rtw_builder* b;
    std::vector<rtw_id> objects;
    std::vector<rtw_id> lights;
    void add(rtw_id h) { objects.push_back(h); }
    bool finish(const V3& lookfrom, const V3& lookat, double vfov, double aspect, double aperture, double dist_to_focus,
                int background) {
        for (rtw_id o : objects)
            if (o < 0) return false;
        rtw_id w = rtw_hittable_list(b, objects.data(), (int)objects.size());
        rtw_id l = rtw_hittable_list(b, lights.data(), (int)lights.size());  // scene.h:35: always a list, maybe empty
        if (w < 0 || l < 0) return false;
        double lf[3] = {lookfrom.x, lookfrom.y, lookfrom.z}, la[3] = {lookat.x, lookat.y, lookat.z}, up[3] = {0.0, 1.0, 0.0};
        return rtw_set_world(b, w) == RTW_OK && rtw_set_lights(b, l) == RTW_OK &&
               rtw_set_camera(b, lf, la, up, vfov, aspect, aperture, dist_to_focus, 0.0, 1.0) == RTW_OK &&
               rtw_set_background_type(b, background) == RTW_OK && rtw_set_render_type(b, RTW_RENDER_SHADED) == RTW_OK;
    }
};

rtw_id lambert_const(rtw_builder* b, double r, double g, double bl) {
    return rtw_lambertian(b, rtw_constant_texture(b, r, g, bl));
}

// Scene/scene.h:42-70
bool light_sample(rtw_builder* b, double aspect) {
    SceneCtx s{b};
    rtw_id pertext = rtw_noise_texture(b, 4.0);
    rtw_id four = rtw_constant_texture(b, 4, 4, 4);
    s.add(rtw_sphere(b, 0, -1000, 0, 1000.0, rtw_lambertian(b, pertext)));
    s.add(rtw_sphere(b, 0, 2, 0, 2.0, rtw_lambertian(b, pertext)));
    s.add(rtw_sphere(b, 0, 7, 0, 2.0, rtw_diffuse_light(b, four)));
    s.add(rtw_xy_rect(b, 3.0, 5.0, 1.0, 3.0, -2.0, rtw_diffuse_light(b, four)));
    V3 lookfrom{24, 5, 5}, lookat{0, 3, 0};
    double dx = lookfrom.x - lookat.x, dy = lookfrom.y - lookat.y, dz = lookfrom.z - lookat.z;
    double dist_to_focus = std::sqrt(dx * dx + dy * dy + dz * dz);
    float aperture = 0.2f;
    return s.finish(lookfrom, lookat, 20.0, aspect, aperture, dist_to_focus, RTW_BG_GRADIENT);
}

// Scene/scene.h:72-96
bool dielectric_scene(rtw_builder* b, double aspect) {
    SceneCtx s{b};
    s.add(rtw_sphere(b, 0, 0, -1, 0.5f, lambert_const(b, 0.1f, 0.2f, 0.5f)));
    s.add(rtw_sphere(b, 0, -100.5f, -1, 100.0, lambert_const(b, 0.8f, 0.8f, 0.0)));
    s.add(rtw_sphere(b, 1, 0, -1, 0.5f, rtw_metal(b, 0.8f, 0.6f, 0.2f, 0.0)));
    s.add(rtw_sphere(b, -1, 0, -1, 0.5f, rtw_dielectric(b, 1.5f)));
    s.add(rtw_sphere(b, -1, 0, -1, -0.45f, rtw_dielectric(b, 1.5f)));
    return s.finish({0, 0, 0}, {0, 0, -1}, 120.0, aspect, 0.0, 10.0, RTW_BG_GRADIENT);
}

// Scene/scene.h:98-174
bool random_balls_scene(rtw_builder* b, double aspect) {
    SceneCtx s{b};
    std::uniform_real_distribution<double> uniform;
    std::minstd_rand engine;
    s.add(rtw_sphere(b, 0, -1000, 0, 1000.0, lambert_const(b, 0.5f, 0.5f, 0.5f)));
    for (int a = -11; a < 11; a++)
        for (int bb = -11; bb < 11; bb++) {
            double choose_mat = uniform(engine);
            // scene.h:116 draws twice inside one constructor call; g++ 13 evaluates the arguments of
            // this non-aggregate constructor right to left: the z draw happens first (pinned by the
            // byte-for-byte comparison in tests/test_host_parity.py).
            double zr = uniform(engine);
            double xr = uniform(engine);
            V3 center{a + 0.9f * xr, 0.2f, bb + 0.9f * zr};
            double ex = center.x - 4.0, ey = center.y - (double)0.2f, ez = center.z - 0.0;
            if (std::sqrt(ex * ex + ey * ey + ez * ez) > 0.9f) {
                if (choose_mat < 0.8f) {  // diffuse, always a moving_sphere (scene.h:126)
                    double r1 = uniform(engine), r2 = uniform(engine);
                    double cr = r1 * r2;
                    double g1 = uniform(engine), g2 = uniform(engine);
                    double cg = g1 * g2;
                    double b1 = uniform(engine), b2 = uniform(engine);
                    double cb = b1 * b2;
                    rtw_id mat = lambert_const(b, cr, cg, cb);
                    double c1y = center.y + 0.5f * uniform(engine);
                    s.add(rtw_moving_sphere(b, center.x, center.y, center.z, 0.2f, mat, center.x + 0.0, c1y, center.z + 0.0, 0.0, 1.0));
                } else if (choose_mat < 0.95) {  // metal
                    double cr = 0.5f * (1 + uniform(engine));
                    double cg = 0.5f * (1 + uniform(engine));
                    double cb = 0.5f * (1 + uniform(engine));
                    double fuzz = 0.5f * uniform(engine);
                    s.add(rtw_sphere(b, center.x, center.y, center.z, 0.2f, rtw_metal(b, cr, cg, cb, fuzz)));
                } else {  // glass
                    s.add(rtw_sphere(b, center.x, center.y, center.z, 0.2f, rtw_dielectric(b, 1.5f)));
                }
            }
        }
    s.add(rtw_sphere(b, 0, 1, 0, 1.0, rtw_dielectric(b, 1.5f)));
    s.add(rtw_sphere(b, -4, 1, 0, 1.0, lambert_const(b, 0.4f, 0.2f, 0.1f)));
    s.add(rtw_sphere(b, 4, 1, 0, 1.0, rtw_metal(b, 0.7f, 0.6f, 0.5f, 0.0)));
    return s.finish({13, 2, 3}, {0, 0, 0}, 20.0, aspect, 0.0, 10.0, RTW_BG_GRADIENT);
}

// Scene/scene.h:176-249; `aluminum` enables the #if at :228-231 (BASELINE.json cfg4/5); `glass` = false switches the #if at
// :219-225 off: the short box (:212-217) stays and nothing but the area light is in `lights` ("cornell_two_boxes";
// with `aluminum` it is the scene of the reference's Sampling/12.2.png, "cornell_aluminum_box").
bool cornell_box_scene(rtw_builder* b, double aspect, bool aluminum, bool glass = true) {
    SceneCtx s{b};
    rtw_id red = lambert_const(b, 0.65f, 0.05f, 0.05f);
    rtw_id white = lambert_const(b, 0.73f, 0.73f, 0.73f);
    rtw_id green = lambert_const(b, 0.12f, 0.45f, 0.15f);
    rtw_id light = rtw_diffuse_light(b, rtw_constant_texture(b, 15.0, 15.0, 15.0));

    s.add(rtw_xz_rect(b, 213.0, 343.0, 227.0, 332.0, 554.0, light));
    s.lights.push_back(s.objects.back());
    s.add(rtw_flip_normals(b, rtw_yz_rect(b, 0.0, 555.0, 0.0, 555.0, 555.0, green)));
    s.add(rtw_yz_rect(b, 0.0, 555.0, 0.0, 555.0, 0.0, red));
    s.add(rtw_flip_normals(b, rtw_xz_rect(b, 0.0, 555.0, 0.0, 555.0, 555.0, white)));
    s.add(rtw_xz_rect(b, 0.0, 555.0, 0.0, 555.0, 0.0, white));
    s.add(rtw_flip_normals(b, rtw_xy_rect(b, 0.0, 555.0, 0.0, 555.0, 555.0, white)));
    if (glass) {
        // the short box is pushed and popped again (scene.h:212-220); the glass sphere replaces it
        s.add(rtw_sphere(b, 190, 90, 190, 90, rtw_dielectric(b, 1.5)));
        s.lights.push_back(s.objects.back());
    } else {
        double q0[3] = {0.0, 0.0, 0.0}, q1[3] = {165.0, 165.0, 165.0};
        s.add(rtw_translate(b, rtw_rotate_y(b, rtw_box(b, q0, q1, white), -18.0), 130.0, 0.0, 65.0));  // scene.h:212-217
    }
    rtw_id tall_mat = aluminum ? rtw_metal(b, 0.8, 0.85, 0.88, 0.0) : white;
    double p0[3] = {0.0, 0.0, 0.0}, p1[3] = {165.0, 330.0, 165.0};
    s.add(rtw_translate(b, rtw_rotate_y(b, rtw_box(b, p0, p1, tall_mat), 15.0), 265.0, 0.0, 295.0));
    return s.finish({278.0, 278.0, -800.0}, {278.0, 278.0, 0.0}, 40.0, aspect, 0.0, 10.0, RTW_BG_BLACK);
}

// BASELINE.json cfg2: Book-2 Perlin spheres + an earth image_texture sphere.
// Sphere placement and camera framing are this project's choice (nothing in the reference fixes them).
bool two_perlin_spheres_earth(rtw_builder* b, double aspect, const uint8_t* earth, int enx, int eny) {
    if (!earth) {
        rtw::set_error("two_perlin_spheres_earth needs the earth RGB8 texels");
        return false;
    }
    SceneCtx s{b};
    rtw_id pertext = rtw_noise_texture(b, 4.0);
    s.add(rtw_sphere(b, 0, -1000, 0, 1000.0, rtw_lambertian(b, pertext)));
    s.add(rtw_sphere(b, 0, 2, 2.2, 2.0, rtw_lambertian(b, pertext)));
    s.add(rtw_sphere(b, 0, 2, -2.2, 2.0, rtw_lambertian(b, rtw_image_texture(b, earth, enx, eny))));
    return s.finish({13, 2, 3}, {0, 2, 0}, 40.0, aspect, 0.0, 10.0, RTW_BG_GRADIENT);
}

// BASELINE.json cfg3: Book-2 final scene through the reference's constructors.
bool book2_final_scene(rtw_builder* b, double aspect, const uint8_t* earth, int enx, int eny) {
    if (!earth) {
        rtw::set_error("book2_final_scene needs the earth RGB8 texels");
        return false;
    }
    SceneCtx s{b};
    std::uniform_real_distribution<double> uniform;
    std::minstd_rand engine;

    rtw_id ground = lambert_const(b, 0.48, 0.83, 0.53);
    std::vector<rtw_id> boxes;
    for (int i = 0; i < 20; i++)
        for (int j = 0; j < 20; j++) {
            double w = 100.0;
            double x0 = -1000.0 + i * w, z0 = -1000.0 + j * w, y0 = 0.0;
            double x1 = x0 + w, y1 = 100.0 * (uniform(engine) + 0.01), z1 = z0 + w;
            double p0[3] = {x0, y0, z0}, p1[3] = {x1, y1, z1};
            boxes.push_back(rtw_box(b, p0, p1, ground));
        }
    s.add(rtw_bvh_node(b, boxes.data(), (int)boxes.size(), 0.0, 1.0));

    rtw_id light = rtw_diffuse_light(b, rtw_constant_texture(b, 7, 7, 7));
    s.add(rtw_xz_rect(b, 123, 423, 147, 412, 554, light));
    s.lights.push_back(s.objects.back());

    s.add(rtw_moving_sphere(b, 400, 400, 200, 50, lambert_const(b, 0.7, 0.3, 0.1), 430, 400, 200, 0.0, 1.0));
    s.add(rtw_sphere(b, 260, 150, 45, 50, rtw_dielectric(b, 1.5)));
    // the book passes fuzz 10 and clamps to 1; the reference's metal does not clamp (material.h:127)
    s.add(rtw_sphere(b, 0, 150, 145, 50, rtw_metal(b, 0.8, 0.8, 0.9, 1.0)));

    rtw_id boundary = rtw_sphere(b, 360, 150, 145, 70, rtw_dielectric(b, 1.5));
    s.add(boundary);
    s.add(rtw_constant_medium(b, boundary, 0.2, rtw_isotropic(b, rtw_constant_texture(b, 0.2, 0.4, 0.9))));
    rtw_id fog_boundary = rtw_sphere(b, 0, 0, 0, 5000, rtw_dielectric(b, 1.5));
    s.add(rtw_constant_medium(b, fog_boundary, 0.0001, rtw_isotropic(b, rtw_constant_texture(b, 1, 1, 1))));

    s.add(rtw_sphere(b, 400, 200, 400, 100, rtw_lambertian(b, rtw_image_texture(b, earth, enx, eny))));
    s.add(rtw_sphere(b, 220, 280, 300, 80, rtw_lambertian(b, rtw_noise_texture(b, 0.1))));

    rtw_id white = lambert_const(b, 0.73, 0.73, 0.73);
    std::vector<rtw_id> spheres;
    for (int j = 0; j < 1000; j++) {
        double x = 165 * uniform(engine);
        double y = 165 * uniform(engine);
        double z = 165 * uniform(engine);
        spheres.push_back(rtw_sphere(b, x, y, z, 10, white));
    }
    s.add(rtw_translate(b, rtw_rotate_y(b, rtw_bvh_node(b, spheres.data(), (int)spheres.size(), 0.0, 1.0), 15.0), -100, 270, 395));

    return s.finish({478, 278, -600}, {278, 278, 0}, 40.0, aspect, 0.0, 10.0, RTW_BG_BLACK);
}

// Book 2 "Cornell smoke" (the reference's Volume.png, README figure): the two boxes become constant_medium boundaries
// (hittable.h:420-489) — white and black smoke.  Composed from the reference's constructors; the light is the wider,
// dimmer panel of the book.  Exercises constant_medium over translate(rotate_y(box)) boundaries.
bool cornell_smoke_scene(rtw_builder* b, double aspect) {
    SceneCtx s{b};
    rtw_id red = lambert_const(b, 0.65f, 0.05f, 0.05f);
    rtw_id white = lambert_const(b, 0.73f, 0.73f, 0.73f);
    rtw_id green = lambert_const(b, 0.12f, 0.45f, 0.15f);
    rtw_id light = rtw_diffuse_light(b, rtw_constant_texture(b, 7.0, 7.0, 7.0));
    s.add(rtw_xz_rect(b, 113.0, 443.0, 127.0, 432.0, 554.0, light));
    s.lights.push_back(s.objects.back());
    s.add(rtw_flip_normals(b, rtw_yz_rect(b, 0.0, 555.0, 0.0, 555.0, 555.0, green)));
    s.add(rtw_yz_rect(b, 0.0, 555.0, 0.0, 555.0, 0.0, red));
    s.add(rtw_flip_normals(b, rtw_xz_rect(b, 0.0, 555.0, 0.0, 555.0, 555.0, white)));
    s.add(rtw_xz_rect(b, 0.0, 555.0, 0.0, 555.0, 0.0, white));
    s.add(rtw_flip_normals(b, rtw_xy_rect(b, 0.0, 555.0, 0.0, 555.0, 555.0, white)));
    double p0[3] = {0.0, 0.0, 0.0}, p1[3] = {165.0, 165.0, 165.0}, p2[3] = {165.0, 330.0, 165.0};
    rtw_id b1 = rtw_translate(b, rtw_rotate_y(b, rtw_box(b, p0, p1, white), -18.0), 130.0, 0.0, 65.0);
    rtw_id b2 = rtw_translate(b, rtw_rotate_y(b, rtw_box(b, p0, p2, white), 15.0), 265.0, 0.0, 295.0);
    s.add(rtw_constant_medium(b, b1, 0.01, rtw_isotropic(b, rtw_constant_texture(b, 1.0, 1.0, 1.0))));
    s.add(rtw_constant_medium(b, b2, 0.01, rtw_isotropic(b, rtw_constant_texture(b, 0.0, 0.0, 0.0))));
    return s.finish({278.0, 278.0, -800.0}, {278.0, 278.0, 0.0}, 40.0, aspect, 0.0, 10.0, RTW_BG_BLACK);
}

// Book 2 "two checker spheres": checker_texture (texture.h:29-50) on two large spheres under the gradient sky.
bool checker_spheres_scene(rtw_builder* b, double aspect) {
    SceneCtx s{b};
    rtw_id checker = rtw_checker_texture(b, rtw_constant_texture(b, 0.2, 0.3, 0.1), rtw_constant_texture(b, 0.9, 0.9, 0.9));
    s.add(rtw_sphere(b, 0, -10, 0, 10.0, rtw_lambertian(b, checker)));
    s.add(rtw_sphere(b, 0, 10, 0, 10.0, rtw_lambertian(b, checker)));
    return s.finish({13, 2, 3}, {0, 0, 0}, 20.0, aspect, 0.0, 10.0, RTW_BG_GRADIENT);
}

}  // namespace

extern "C" int rtw_build_named_scene(rtw_builder* b, const char* name, double aspect, const uint8_t* earth, int enx,
                                     int eny) {
    if (!b || !name) {
        rtw::set_error("build_named_scene: null argument");
        return RTW_ERR_INVALID;
    }
    std::string n(name);
    bool ok;
    if (n == "light_sample") ok = light_sample(b, aspect);
    else if (n == "dielectric_scene") ok = dielectric_scene(b, aspect);
    else if (n == "random_balls_scene") ok = random_balls_scene(b, aspect);
    else if (n == "cornell_box_scene") ok = cornell_box_scene(b, aspect, false);
    else if (n == "cornell_box_aluminum") ok = cornell_box_scene(b, aspect, true);
    else if (n == "cornell_two_boxes") ok = cornell_box_scene(b, aspect, false, false);
    else if (n == "cornell_aluminum_box") ok = cornell_box_scene(b, aspect, true, false);
    else if (n == "two_perlin_spheres_earth") ok = two_perlin_spheres_earth(b, aspect, earth, enx, eny);
    else if (n == "book2_final_scene") ok = book2_final_scene(b, aspect, earth, enx, eny);
    else if (n == "cornell_smoke_scene") ok = cornell_smoke_scene(b, aspect);
    else if (n == "checker_spheres_scene") ok = checker_spheres_scene(b, aspect);
    else {
        rtw::set_error("unknown scene name: " + n);
        return RTW_ERR_INVALID;
    }
    if (!ok) return RTW_ERR_INVALID;
    return RTW_OK;
}
