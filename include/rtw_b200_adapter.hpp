// rtw_b200_adapter.hpp — C++ adapter between the reference's scene classes and the rtw_b200 C ABI.
//
// Usage inside the reference's own translation unit (RayTracingWeekend/RayTracingWeekend.cpp), AFTER its
// headers have been included (vec3.h, hittable.h, hittable_list.h, sphere.h, material.h, texture.h,
// camera.h, Scene/scene.h):
//
//     #include "rtw_b200_adapter.hpp"
//     ...
//     scene_type scene(nx * 1.0 / ny);
//     std::vector<vec3> canvas(nx * ny);
//     rtw_b200::render(scene, nx, ny, subPixelCount, max_depth, canvas);   // replaces :211-250
//     // the PPM writer at :252-282 stays untouched
//
// The adapter only READS the object graph (every data member it needs is public in the reference — SURVEY.md
// §8b); it never includes, edits or copies a reference header.  `perlin`'s tables are private statics
// (noise.h:215-225): they are reached with the explicit-instantiation access idiom below, after forcing the
// lazy initialisation (noise.h:91-94) with one noise() call.
#pragma once

#include <cstdint>
#include <cstdio>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <typeinfo>
#include <vector>

#include "rtw_b200.h"

namespace rtw_b200 {

namespace detail {

// --- private static member access without touching noise.h ---------------------------------------
template <typename Tag, typename Tag::type M>
struct rob {
    friend typename Tag::type get(Tag) { return M; }
};
struct perlin_ranvec_tag {
    typedef vec3 (*type)[256];
    friend type get(perlin_ranvec_tag);
};
struct perlin_perm_x_tag {
    typedef int (*type)[256];
    friend type get(perlin_perm_x_tag);
};
struct perlin_perm_y_tag {
    typedef int (*type)[256];
    friend type get(perlin_perm_y_tag);
};
struct perlin_perm_z_tag {
    typedef int (*type)[256];
    friend type get(perlin_perm_z_tag);
};
template struct rob<perlin_ranvec_tag, &perlin::ranvec>;
template struct rob<perlin_perm_x_tag, &perlin::perm_x>;
template struct rob<perlin_perm_y_tag, &perlin::perm_y>;
template struct rob<perlin_perm_z_tag, &perlin::perm_z>;

struct walker {
    rtw_builder* b;
    std::map<const hittable*, rtw_id> hmap;
    std::map<const material*, rtw_id> mmap;
    std::map<const texture*, rtw_id> tmap;
    bool saw_noise = false;

    explicit walker(rtw_builder* bb) : b(bb) {}

    [[noreturn]] static void fail(const std::string& what) {
        throw std::runtime_error("rtw_b200 adapter: " + what + " (" + rtw_last_error() + ")");
    }
    static rtw_id chk(rtw_id id, const char* what) {
        if (id < 0) fail(what);
        return id;
    }

    rtw_id tex(const texture* t) {
        if (!t) fail("null texture");
        auto it = tmap.find(t);
        if (it != tmap.end()) return it->second;
        rtw_id id;
        if (auto c = dynamic_cast<const constant_texture*>(t)) {
            id = rtw_constant_texture(b, c->color.x, c->color.y, c->color.z);
        } else if (auto c = dynamic_cast<const checker_texture*>(t)) {
            id = rtw_checker_texture(b, tex(c->even.get()), tex(c->odd.get()));
        } else if (auto c = dynamic_cast<const noise_texture*>(t)) {
            saw_noise = true;
            id = rtw_noise_texture(b, c->scale);
        } else if (auto c = dynamic_cast<const image_texture*>(t)) {
            if (!c->data) fail("image_texture without data");
            id = rtw_image_texture(b, c->data->data(), c->nx, c->ny);
        } else {
            fail(std::string("unsupported texture class ") + typeid(*t).name());
        }
        return tmap[t] = chk(id, "texture");
    }

    rtw_id mat(const material* m) {
        if (!m) fail("null material");
        auto it = mmap.find(m);
        if (it != mmap.end()) return it->second;
        rtw_id id;
        if (auto c = dynamic_cast<const lambertian*>(m)) {
            id = rtw_lambertian(b, tex(c->albedo.get()));
        } else if (auto c = dynamic_cast<const metal*>(m)) {
            id = rtw_metal(b, c->albedo.x, c->albedo.y, c->albedo.z, c->fuzz);
        } else if (auto c = dynamic_cast<const dielectric*>(m)) {
            id = rtw_dielectric(b, c->ref_idx);
        } else if (auto c = dynamic_cast<const diffuse_light*>(m)) {
            id = rtw_diffuse_light(b, tex(c->emit.get()));
        } else if (auto c = dynamic_cast<const isotropic*>(m)) {
            id = rtw_isotropic(b, tex(c->albedo.get()));
        } else {
            fail(std::string("unsupported material class ") + typeid(*m).name());
        }
        return mmap[m] = chk(id, "material");
    }

    rtw_id list(const hittable_list& l) {
        std::vector<rtw_id> ids;
        for (const auto& o : l.objects) ids.push_back(hit(o.get()));
        return chk(rtw_hittable_list(b, ids.data(), (int)ids.size()), "hittable_list");
    }

    rtw_id hit(const hittable* h) {
        if (!h) fail("null hittable");
        auto it = hmap.find(h);
        if (it != hmap.end()) return it->second;
        rtw_id id;
        if (auto c = dynamic_cast<const sphere*>(h)) {
            id = rtw_sphere(b, c->center.x, c->center.y, c->center.z, c->radius, mat(c->mat.get()));
        } else if (auto c = dynamic_cast<const moving_sphere*>(h)) {
            id = rtw_moving_sphere(b, c->center.x, c->center.y, c->center.z, c->radius, mat(c->mat.get()),
                                   c->movement.center1.x, c->movement.center1.y, c->movement.center1.z,
                                   c->movement.time0, c->movement.time1);
        } else if (auto c = dynamic_cast<const xy_rect*>(h)) {
            id = rtw_xy_rect(b, c->x0, c->x1, c->y0, c->y1, c->k, mat(c->mp.get()));
        } else if (auto c = dynamic_cast<const xz_rect*>(h)) {
            id = rtw_xz_rect(b, c->x0, c->x1, c->z0, c->z1, c->k, mat(c->mp.get()));
        } else if (auto c = dynamic_cast<const yz_rect*>(h)) {
            id = rtw_yz_rect(b, c->y0, c->y1, c->z0, c->z1, c->k, mat(c->mp.get()));
        } else if (auto c = dynamic_cast<const flip_normals*>(h)) {
            id = rtw_flip_normals(b, hit(c->ptr.get()));
        } else if (auto c = dynamic_cast<const translate*>(h)) {
            id = rtw_translate(b, hit(c->ptr.get()), c->offset.x, c->offset.y, c->offset.z);
        } else if (auto c = dynamic_cast<const rotate_y*>(h)) {
            id = rtw_rotate_y_sincos(b, hit(c->ptr.get()), c->sin_theta, c->cos_theta);
        } else if (auto c = dynamic_cast<const constant_medium*>(h)) {
            id = rtw_constant_medium(b, hit(c->boundary.get()), c->density, mat(c->mp.get()));
        } else if (auto c = dynamic_cast<const box*>(h)) {
            // box keeps its 6 rects in list_ptr (hittable_list.h:112-113); the material is the rects'
            const material* m = nullptr;
            if (!c->list_ptr.objects.empty())
                if (auto r0 = dynamic_cast<const xy_rect*>(c->list_ptr.objects[0].get())) m = r0->mp.get();
            if (!m) fail("box without its xy_rect face");
            double p0[3] = {c->pmin.x, c->pmin.y, c->pmin.z}, p1[3] = {c->pmax.x, c->pmax.y, c->pmax.z};
            id = rtw_box(b, p0, p1, mat(m));
        } else if (auto c = dynamic_cast<const hittable_list*>(h)) {
            id = list(*c);
        } else if (auto c = dynamic_cast<const bvh_node*>(h)) {
            rtw_id l = hit(c->left), r = hit(c->right);
            vec3 mn = c->box.min(), mx = c->box.max();
            double a[3] = {mn.x, mn.y, mn.z}, d[3] = {mx.x, mx.y, mx.z};
            id = rtw_bvh_node_import(b, l, r, a, d);
        } else {
            fail(std::string("unsupported hittable class ") + typeid(*h).name());
        }
        return hmap[h] = chk(id, "hittable");
    }

    void perlin_tables() {
        perlin().noise(vec3(0, 0, 0));  // force the lazy initialisation (noise.h:91-94)
        vec3(*rv)[256] = get(perlin_ranvec_tag());
        int(*px)[256] = get(perlin_perm_x_tag());
        int(*py)[256] = get(perlin_perm_y_tag());
        int(*pz)[256] = get(perlin_perm_z_tag());
        std::vector<double> ranvec(768);
        std::vector<int32_t> x(256), y(256), z(256);
        for (int i = 0; i < 256; i++) {
            ranvec[3 * i] = (*rv)[i].x, ranvec[3 * i + 1] = (*rv)[i].y, ranvec[3 * i + 2] = (*rv)[i].z;
            x[i] = (*px)[i], y[i] = (*py)[i], z[i] = (*pz)[i];
        }
        if (rtw_set_perlin_tables(b, ranvec.data(), x.data(), y.data(), z.data()) != RTW_OK) fail("perlin tables");
    }
};

}  // namespace detail

// Walk `s` (Scene/scene.h:18-40 accessors) into `b`.  Throws std::runtime_error on unsupported graphs.
inline void describe(rtw_builder* b, scene& s) {
    detail::walker w(b);
    rtw_id world = w.list(s.GetWorld());
    if (rtw_set_world(b, world) != RTW_OK) detail::walker::fail("set_world");
    auto lights = s.GetLights();  // may be null or empty (RayTracingWeekend.cpp:118)
    rtw_id l = -1;
    if (lights) l = w.list(*lights);
    if (rtw_set_lights(b, l) != RTW_OK) detail::walker::fail("set_lights");
    rtw_set_render_type(b, s.GetRenderType() == RenderType::Normal ? RTW_RENDER_NORMAL : RTW_RENDER_SHADED);
    rtw_set_background_type(b, s.GetBackgroundType() == BackgroundType::Gradient ? RTW_BG_GRADIENT : RTW_BG_BLACK);
    camera& c = s.GetCamera();
    rtw_flat_camera fc{};
    auto put = [](double* d, const vec3& v) { d[0] = v.x, d[1] = v.y, d[2] = v.z; };
    put(fc.origin, c.origin), put(fc.lower_left_corner, c.lower_left_corner), put(fc.horizontal, c.horizontal);
    put(fc.vertical, c.vertical), put(fc.u, c.u), put(fc.v, c.v), put(fc.w, c.w);
    fc.time0 = c.time0, fc.time1 = c.time1, fc.lens_radius = c.lens_radius;
    if (rtw_set_camera_members(b, &fc) != RTW_OK) detail::walker::fail("set_camera");
    if (w.saw_noise) w.perlin_tables();
}

struct render_options {
    int device = 0;
    int spp_begin = 0;
    uint64_t seed = 0;
    uint32_t flags = 0;
    rtw_render_stats* stats = nullptr;
    // multi-GPU (one process or thread per GPU, include/rtw_b200.h section 4): every rank calls render() with the same
    // scene and `spp` = ALL samples; rank r renders its share, the sums are reduced onto `root` (ncclReduce), and only the
    // root's canvas is filled — the others come back empty.  nullptr = single GPU.
    rtw_comm* comm = nullptr;
    int root = 0;
};

// Drop-in for RayTracingWeekend.cpp:211-250: fills `canvas[j*nx+i]` with the gamma-2, clamped colour.
inline void render(scene& s, int nx, int ny, int spp, int max_depth, std::vector<vec3>& canvas,
                   const render_options& opt = render_options()) {
    struct guard {
        rtw_builder* b = nullptr;
        rtw_scene* sc = nullptr;
        ~guard() {
            if (sc) rtw_scene_destroy(sc);
            if (b) rtw_builder_destroy(b);
        }
    } g;
    g.b = rtw_builder_create();
    if (!g.b) detail::walker::fail("builder_create");
    describe(g.b, s);
    const rtw_flat_scene* fs = rtw_builder_flatten(g.b);
    if (!fs) detail::walker::fail("flatten");
    g.sc = rtw_scene_create(fs, opt.device);
    if (!g.sc) detail::walker::fail("scene_create");
    rtw_render_params p{};
    p.nx = nx, p.ny = ny, p.spp_begin = opt.spp_begin, p.spp_count = spp, p.max_depth = max_depth;
    p.flags = opt.flags, p.seed = opt.seed;
    const bool is_root = !opt.comm || rtw_comm_rank(opt.comm) == opt.root;
    std::vector<float> accum(is_root ? (size_t)4 * nx * ny : 0);
    rtw_render_stats st{};
    if (opt.comm) {
        if (rtw_render_sharded(g.sc, opt.comm, &p, opt.root, is_root ? accum.data() : nullptr, &st) != RTW_OK) detail::walker::fail("render_sharded");
    } else if (rtw_render(g.sc, &p, accum.data(), &st) != RTW_OK) {
        detail::walker::fail("render");
    }
    if (opt.stats) *opt.stats = st;
    if (!is_root) {  // :241-250 happen on the root only
        canvas.clear();
        return;
    }
    std::vector<double> rgb((size_t)3 * nx * ny);
    if (rtw_tonemap(accum.data(), nx, ny, rgb.data()) != RTW_OK) detail::walker::fail("tonemap");
    canvas.resize((size_t)nx * ny);
    for (size_t i = 0; i < (size_t)nx * ny; i++) canvas[i] = vec3(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]);
}

}  // namespace rtw_b200
