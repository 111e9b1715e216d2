// ref_harness.cpp — TEST INFRASTRUCTURE.  Drives the UNMODIFIED reference sources
// (/root/reference/RayTracingWeekend/*.h, RayTracingWeekend.cpp, noise.cpp) as the parity oracle and as
// the timed CPU baseline.  Built by oracle/Makefile into oracle/_ref/ref_harness (git-ignored); never
// linked into, imported by or executed from the product path.
//
// The reference translation unit is included verbatim (its main() renamed by a macro) so that color()
// (RayTracingWeekend.cpp:45-160), camera::get_ray, every hit()/scatter()/pdf and the RNG engines are the
// reference's own object code.  The only restated piece is the image loop of :214-250 with nx/ny/spp
// turned into run-time arguments (the reference fixes them at compile time, :32-36).
//
// One invocation = one fresh process = fresh `static std::minstd_rand` engines (utility.h:16-17 etc.),
// which is what makes serial runs (RTW_PPL_THREADS=1) bit-reproducible.
//
// Sub-commands (binary I/O is little-endian float64 unless noted):
//   flat   <scene> <aspect> <earth.rgb8|-> <out>                         adapter -> rtw_flat_serialize bytes
//   hits   <scene> <aspect> <earth|-> <rays> <out> <tmin> <tmax|inf>     world.hit() per ray
//   render <scene> <aspect> <earth|-> <W> <H> <spp> <depth> <out>        linear radiance sums + sums of squares
//   perlin <out>                                                         noise.h tables
//   kat                                                                  unit tests of CppTest/unittest1.cpp as JSON
//   funcs  <out>                                                         per-function sequences (fresh RNG state)
//   main                                                                 the reference's own main(), unmodified
#include <cfloat>
#include <climits>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <vector>
#define __int64 long long

#define main reference_main
#include "RayTracingWeekend.cpp"  // resolved through -I to the (overlaid) reference directory
#undef main

#include "ref_scenes.h"
#include "rtw_b200_adapter.hpp"

extern "C" int32_t rtw_builder_flat_material_index(const rtw_builder* b, rtw_id material);

namespace {

std::shared_ptr<image_texture::byte_array> g_earth;
int g_enx = 0, g_eny = 0;

bool load_earth(const char* path) {
    if (!strcmp(path, "-")) return true;
    FILE* f = fopen(path, "rb");
    if (!f) return false;
    g_earth = std::make_shared<image_texture::byte_array>();
    unsigned char buf[65536];
    size_t n;
    while ((n = fread(buf, 1, sizeof buf, f)) > 0) g_earth->insert(g_earth->end(), buf, buf + n);
    fclose(f);
    // square RGB8
    size_t px = g_earth->size() / 3;
    int side = 1;
    while ((size_t)side * side < px) side++;
    g_enx = g_eny = side;
    return (size_t)side * side * 3 == g_earth->size();
}

std::unique_ptr<scene> make_scene(const std::string& name, double aspect) {
    if (name == "light_sample") return std::unique_ptr<scene>(new light_sample(aspect));
    if (name == "dielectric_scene") return std::unique_ptr<scene>(new dielectric_scene(aspect));
    if (name == "random_balls_scene") return std::unique_ptr<scene>(new random_balls_scene(aspect));
    if (name == "cornell_box_scene") return std::unique_ptr<scene>(new cornell_box_scene(aspect));
    if (name == "cornell_box_aluminum") return std::unique_ptr<scene>(new cornell_box_aluminum_scene(aspect));
    if (name == "cornell_two_boxes") return std::unique_ptr<scene>(new cornell_no_glass_scene(aspect, false));
    if (name == "cornell_aluminum_box") return std::unique_ptr<scene>(new cornell_no_glass_scene(aspect, true));
    if (name == "cornell_smoke_scene") return std::unique_ptr<scene>(new cornell_smoke_scene(aspect));
    if (name == "checker_spheres_scene") return std::unique_ptr<scene>(new checker_spheres_scene(aspect));
    if (name == "two_perlin_spheres_earth" && g_earth)
        return std::unique_ptr<scene>(new two_perlin_spheres_earth_scene(aspect, g_earth, g_enx, g_eny));
    if (name == "book2_final_scene" && g_earth)
        return std::unique_ptr<scene>(new book2_final_scene(aspect, g_earth, g_enx, g_eny));
    fprintf(stderr, "unknown scene '%s' (or missing earth texels)\n", name.c_str());
    return nullptr;
}

bool write_file(const char* path, const void* data, size_t bytes) {
    FILE* f = fopen(path, "wb");
    if (!f) return false;
    bool ok = fwrite(data, 1, bytes, f) == bytes;
    fclose(f);
    return ok;
}

std::vector<uint8_t> read_file(const char* path) {
    std::vector<uint8_t> v;
    FILE* f = fopen(path, "rb");
    if (!f) return v;
    unsigned char buf[65536];
    size_t n;
    while ((n = fread(buf, 1, sizeof buf, f)) > 0) v.insert(v.end(), buf, buf + n);
    fclose(f);
    return v;
}

int cmd_flat(int argc, char** argv) {
    if (argc < 6 || !load_earth(argv[4])) return 2;
    auto s = make_scene(argv[2], atof(argv[3]));
    if (!s) return 2;
    rtw_builder* b = rtw_builder_create();
    rtw_b200::describe(b, *s);
    const rtw_flat_scene* fs = rtw_builder_flatten(b);
    if (!fs) {
        fprintf(stderr, "flatten failed: %s\n", rtw_last_error());
        return 1;
    }
    size_t n = rtw_flat_serialize(fs, nullptr, 0);
    std::vector<uint8_t> buf(n);
    rtw_flat_serialize(fs, buf.data(), n);
    bool ok = write_file(argv[5], buf.data(), n);
    rtw_builder_destroy(b);
    return ok ? 0 : 1;
}

int cmd_hits(int argc, char** argv) {
    if (argc < 9 || !load_earth(argv[4])) return 2;
    auto s = make_scene(argv[2], atof(argv[3]));
    if (!s) return 2;
    double tmin = atof(argv[7]);
    double tmax = !strcmp(argv[8], "inf") ? std::numeric_limits<double>::max() : atof(argv[8]);
    // material pointer -> flat material index through the adapter + builder
    rtw_builder* b = rtw_builder_create();
    rtw_b200::detail::walker w(b);
    rtw_id world = w.list(s->GetWorld());
    rtw_set_world(b, world);
    rtw_flat_camera fc{};
    rtw_set_camera_members(b, &fc);
    if (!rtw_builder_flatten(b)) {
        fprintf(stderr, "flatten failed: %s\n", rtw_last_error());
        return 1;
    }
    std::vector<uint8_t> raw = read_file(argv[5]);
    size_t n = raw.size() / (7 * sizeof(double));
    const double* r = reinterpret_cast<const double*>(raw.data());
    std::vector<double> out(n * 11, 0.0);
    for (size_t i = 0; i < n; i++) {
        ray rr(vec3(r[7 * i], r[7 * i + 1], r[7 * i + 2]), vec3(r[7 * i + 3], r[7 * i + 4], r[7 * i + 5]), r[7 * i + 6]);
        hit_record rec;
        rec.u = rec.v = 0;  // constant_medium leaves them unwritten (hittable.h:468-473)
        bool h = s->GetWorld().hit(rr, tmin, tmax, rec);
        double* o = &out[11 * i];
        o[0] = h ? 1.0 : 0.0;
        if (h) {
            o[1] = rec.t;
            o[2] = rec.p.x, o[3] = rec.p.y, o[4] = rec.p.z;
            o[5] = rec.normal.x, o[6] = rec.normal.y, o[7] = rec.normal.z;
            o[8] = rec.u, o[9] = rec.v;
            auto it = w.mmap.find(rec.mat_ptr);
            o[10] = it == w.mmap.end() ? -1.0 : (double)rtw_builder_flat_material_index(b, it->second);
        }
    }
    bool ok = write_file(argv[6], out.data(), out.size() * sizeof(double));
    rtw_builder_destroy(b);
    return ok ? 0 : 1;
}

// The image loop of RayTracingWeekend.cpp:204-250 with W/H/spp/depth as arguments.  Everything it calls is
// the reference's own code.  Output is pre-gamma: per pixel sum(rgb) and sum(rgb^2) over the samples.
int cmd_render(int argc, char** argv) {
    if (argc < 10 || !load_earth(argv[4])) return 2;
    const int W = atoi(argv[5]), H = atoi(argv[6]), spp = atoi(argv[7]), depth = atoi(argv[8]);
    auto sp = make_scene(argv[2], atof(argv[3]) > 0 ? atof(argv[3]) : W * 1.0 / H);
    if (!sp) return 2;
    scene& sc = *sp;
    auto& cam = sc.GetCamera();  // :205
    std::uniform_real_distribution<double> uniform;  // :207
    std::minstd_rand engine;                         // :208
    std::vector<double> out((size_t)W * H * 6, 0.0);
    __int64 elapsedTrace = time_call([&] {
        _for(0, H, 1, [&](int j) {       // :214
            _for(0, W, 1, [&](int i) {   // :216
                double* o = &out[(size_t)(j * W + i) * 6];
                std::vector<vec3> subPixels(spp);  // :218 (stack array there)
                _for(0, spp, 1, [&](int s) {       // :219
                    double u = double(i + uniform(engine)) / double(W);  // :227
                    double v = double(j + uniform(engine)) / double(H);  // :228
                    ray r = cam.get_ray(u, v);                           // :231
                    subPixels[s] = color(r, &sc, depth);                 // :232
                });
                vec3 sum(0, 0, 0), sq(0, 0, 0);
                for (vec3& c : subPixels) {  // :236-239
                    sum += c;
                    sq += c * c;
                }
                o[0] = sum.x, o[1] = sum.y, o[2] = sum.z, o[3] = sq.x, o[4] = sq.y, o[5] = sq.z;
            });
        });
    });
    double paths = (double)W * H * spp;
    printf("{\"scene\": \"%s\", \"nx\": %d, \"ny\": %d, \"spp\": %d, \"max_depth\": %d, \"threads\": %d, "
           "\"trace_ms\": %lld, \"paths\": %.0f, \"paths_per_s\": %.1f}\n",
           argv[2], W, H, spp, depth, concurrency::shim_thread_count(), (long long)elapsedTrace, paths,
           elapsedTrace > 0 ? paths / (elapsedTrace * 1e-3) : 0.0);
    return write_file(argv[9], out.data(), out.size() * sizeof(double)) ? 0 : 1;
}

int cmd_perlin(int argc, char** argv) {
    if (argc < 3) return 2;
    perlin().noise(vec3(0, 0, 0));
    using namespace rtw_b200::detail;
    vec3(*rv)[256] = get(perlin_ranvec_tag());
    int(*p[3])[256] = {get(perlin_perm_x_tag()), get(perlin_perm_y_tag()), get(perlin_perm_z_tag())};
    std::vector<double> out;
    for (int i = 0; i < 256; i++) out.push_back((*rv)[i].x), out.push_back((*rv)[i].y), out.push_back((*rv)[i].z);
    for (int t = 0; t < 3; t++)
        for (int i = 0; i < 256; i++) out.push_back((double)(*p[t])[i]);
    return write_file(argv[2], out.data(), out.size() * sizeof(double)) ? 0 : 1;
}

// CppTest/unittest1.cpp:20-40, 72-109 evaluated with the reference's own vec3/ray/aabb.
int cmd_kat() {
    auto d = dot(vec3(1, 1, 0), vec3(1, 1, 0));
    auto c = cross(vec3(1, 0, 0), vec3(0, 1, 0));
    auto r = ray(vec3(1, 1, 1), vec3(2, 2, 2), 0.0);
    aabb bx(vec3(2, 2, 2), vec3(4, 4, 4));
    bool h[5] = {bx.hit(ray(vec3(0, 0, 0), vec3(1, 1, 1), 0.0), 0, FLT_MAX),
                 bx.hit(ray(vec3(0, 0, 0), -vec3(1, 1, 1), 0.0), 0, FLT_MAX),
                 bx.hit(ray(vec3(3, 3, 3), vec3(0, 1, 0), 0.0), 0, FLT_MAX),
                 bx.hit(ray(vec3(0, 3, 0), vec3(1, 0, 1), 0.0), 0, FLT_MAX),
                 bx.hit(ray(vec3(0, 5, 0), vec3(1, 0, 1), 0.0), 0, FLT_MAX)};
    aabb box0(vec3(0, 0, 0), vec3(1, 1, 1)), box1(vec3(3, 3, 3), vec3(4, 4, 4));
    auto s = aabb::surrounding(box0, box1);
    printf("{\"dot\": %.17g, \"cross\": [%.17g, %.17g, %.17g], \"point_at_parameter_3\": [%.17g, %.17g, %.17g], "
           "\"aabb_hit\": [%d, %d, %d, %d, %d], \"surrounding_min\": [%.17g, %.17g, %.17g], "
           "\"surrounding_max\": [%.17g, %.17g, %.17g], \"sizeof_vec3\": %zu}\n",
           d, c.x, c.y, c.z, r.point_at_parameter(3).x, r.point_at_parameter(3).y, r.point_at_parameter(3).z, h[0],
           h[1], h[2], h[3], h[4], s.min()[0], s.min()[1], s.min()[2], s.max()[0], s.max()[1], s.max()[2],
           sizeof(vec3));
    return 0;
}

// Per-function sequences from FRESH static engines, in a fixed call order; the CPU restatement
// (oracle/rtw_oracle.cpp, minstd mode) must reproduce every number bit for bit.
int cmd_funcs(int argc, char** argv) {
    if (argc < 3) return 2;
    std::vector<double> o;
    auto put3 = [&](const vec3& v) { o.push_back(v.x), o.push_back(v.y), o.push_back(v.z); };
    for (int i = 0; i < 8; i++) o.push_back(random_double());              // utility.h:14-20
    for (int i = 0; i < 8; i++) o.push_back((double)random_int(0, 1));     // utility.h:22-25
    for (int i = 0; i < 8; i++) put3(random_in_unit_sphere());             // utility.h:27-35
    for (int i = 0; i < 8; i++) put3(random_cosine_direction());           // utility.h:54-67
    for (int i = 0; i < 8; i++) put3(random_to_sphere(90.0, 250000.0));    // utility.h:69-82
    {   // onb.h:32-38 on two normals (both branches of the 0.9 test)
        onb a, b;
        a.build_from_w(vec3(0.3, -0.5, 0.8));
        b.build_from_w(vec3(-2.0, 0.1, 0.2));
        put3(a.u()), put3(a.v()), put3(a.w()), put3(b.u()), put3(b.v()), put3(b.w());
        put3(a.local(vec3(0.1, 0.2, 0.3)));
    }
    {   // pdfs on a Cornell-like configuration
        auto light = std::make_shared<xz_rect>(213.0, 343.0, 227.0, 332.0, 554.0, nullptr);
        auto ball = std::make_shared<sphere>(vec3(190, 90, 190), 90, nullptr);
        auto lst = std::make_shared<hittable_list>();
        lst->objects.push_back(light);
        lst->objects.push_back(ball);
        vec3 org(100.0, 0.0, 300.0);
        for (int i = 0; i < 6; i++) {
            auto cp = std::make_shared<cosine_pdf>(vec3(0, 1, 0));
            mixture_pdf mix(cp, std::make_shared<hittable_pdf>(lst, org));
            vec3 dir = mix.generate();
            put3(dir);
            o.push_back(mix.value(dir));
            o.push_back(light->pdf_value(org, dir));
            o.push_back(ball->pdf_value(org, dir));
        }
    }
    {   // materials: scatter on fixed records (dielectric draws from its own static engine)
        hit_record rec;
        rec.t = 1.0;
        rec.p = vec3(1, 2, 3);
        rec.normal = normalize(vec3(0.2, 1.0, -0.1));
        rec.u = 0.25, rec.v = 0.75;
        ray rin(vec3(0, 5, 0), vec3(0.3, -1.2, 0.4), 0.5);
        ray rout(vec3(1, 1, 4), vec3(-0.3, 1.2, -0.4), 0.5);
        dielectric di(1.5);
        metal me(vec3(0.8, 0.85, 0.88), 0.3);
        isotropic iso(std::make_shared<constant_texture>(vec3(0.2, 0.4, 0.9)));
        for (int i = 0; i < 6; i++) {
            scatter_record sr;
            di.scatter(i % 2 ? rout : rin, rec, sr);
            put3(sr.scattered_ray_without_pdf.direction());
        }
        for (int i = 0; i < 4; i++) {
            scatter_record sr;
            me.scatter(rin, rec, sr);
            put3(sr.scattered_ray_without_pdf.direction());
        }
        for (int i = 0; i < 4; i++) {
            scatter_record sr;
            iso.scatter(rin, rec, sr);
            put3(sr.scattered_ray_without_pdf.direction());
        }
        lambertian la(std::make_shared<constant_texture>(vec3(0.5, 0.5, 0.5)));
        o.push_back(la.scattering_pdf(rin, rec, rout));
        o.push_back(schlick(0.3, 1.5));
        vec3 refr;
        o.push_back(refract(vec3(0.3, -1.2, 0.4), rec.normal, 1.0 / 1.5, refr) ? 1.0 : 0.0);
        put3(refr);
        put3(reflect(vec3(0.3, -1.2, 0.4), rec.normal));
    }
    {   // textures
        noise_texture nt(4.0);
        std::shared_ptr<texture> t0 = std::make_shared<constant_texture>(vec3(0.2, 0.3, 0.1));
        std::shared_ptr<texture> t1 = std::make_shared<constant_texture>(vec3(0.9, 0.9, 0.9));
        checker_texture ck(t0, t1);
        vec3 pts[4] = {vec3(0.1, 0.2, 0.3), vec3(-3.7, 2.2, 9.1), vec3(100.5, -0.25, 0.001), vec3(-0.5, -0.5, -0.5)};
        perlin pn;
        for (auto& p : pts) {
            put3(nt.value(0, 0, p));
            put3(ck.value(0, 0, p));
            o.push_back(pn.noise(p));
            o.push_back(pn.turb(p));
        }
        auto bytes = std::make_shared<image_texture::byte_array>();
        for (int i = 0; i < 4 * 4 * 3; i++) bytes->push_back((unsigned char)(i * 5));
        image_texture it(bytes, 4, 4);
        double uv[5][2] = {{0, 0}, {1, 1}, {0.5, 0.5}, {0.99, 0.01}, {0.26, 0.74}};
        for (auto& q : uv) put3(it.value(q[0], q[1], vec3(0, 0, 0)));
    }
    {   // camera (thin lens; private engines)
        camera cam(vec3(24, 5, 5), vec3(0, 3, 0), vec3(0, 1, 0), 20.0, 1.5, 0.2f, 10.0, 0.0, 1.0);
        for (int i = 0; i < 4; i++) {
            ray r = cam.get_ray(0.25 * i, 0.1 + 0.2 * i);
            put3(r.origin()), put3(r.direction());
            o.push_back(r.time());
        }
    }
    {   // constant_medium called directly and through a one-element list
        auto bnd = std::make_shared<sphere>(vec3(0, 0, 0), 50, nullptr);
        auto med = std::make_shared<constant_medium>(bnd, 0.02, nullptr);
        hittable_list one;
        one.objects.push_back(med);
        ray r(vec3(0, 0, -200), vec3(0, 0, 1), 0.0);
        for (int i = 0; i < 6; i++) {
            hit_record rec;
            bool h = med->hit(r, 0.001f, std::numeric_limits<double>::max(), rec);
            o.push_back(h ? rec.t : -1.0);
        }
        for (int i = 0; i < 6; i++) {
            hit_record rec;
            bool h = one.hit(r, 0.001f, std::numeric_limits<double>::max(), rec);
            o.push_back(h ? rec.t : -1.0);
        }
    }
    return write_file(argv[2], o.data(), o.size() * sizeof(double)) ? 0 : 1;
}

}  // namespace

int main(int argc, char** argv) {
    if (argc < 2) {
        fprintf(stderr, "usage: ref_harness flat|hits|render|perlin|kat|funcs|main ...\n");
        return 2;
    }
    std::string c = argv[1];
    try {
        if (c == "flat") return cmd_flat(argc, argv);
        if (c == "hits") return cmd_hits(argc, argv);
        if (c == "render") return cmd_render(argc, argv);
        if (c == "perlin") return cmd_perlin(argc, argv);
        if (c == "kat") return cmd_kat();
        if (c == "funcs") return cmd_funcs(argc, argv);
        if (c == "main") return reference_main(argc, argv);
    } catch (const std::exception& e) {
        fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
    fprintf(stderr, "unknown sub-command %s\n", c.c_str());
    return 2;
}
