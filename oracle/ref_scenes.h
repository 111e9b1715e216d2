// ref_scenes.h — TEST INFRASTRUCTURE (oracle side).  Scene functions for the BASELINE.json configs the
// reference does not ship, written against the REFERENCE's own API (Scene/scene.h `class scene`), so the
// reference's unmodified color()/hit() code renders them.  Included after the reference headers.
// Parameters: SURVEY.md Appendix C.  The builder-side twins live in raytracingweekend_b200/csrc/scenes.cpp;
// tests/test_host_parity.py compares the two flattened forms byte for byte.
#pragma once

// cfg4/5: Scene/scene.h:176-249 with the `#if 0` at :228-231 switched on (aluminium tall box).
class cornell_box_aluminum_scene : public scene {
public:
    cornell_box_aluminum_scene(double aspect) : scene() {
        std::shared_ptr<texture> red_tex = std::make_shared<constant_texture>(vec3(0.65f, 0.05f, 0.05f));
        auto red = std::make_shared<lambertian>(red_tex);
        std::shared_ptr<texture> white_tex = std::make_shared<constant_texture>(vec3(0.73f, 0.73f, 0.73f));
        auto white = std::make_shared<lambertian>(white_tex);
        std::shared_ptr<texture> green_tex = std::make_shared<constant_texture>(vec3(0.12f, 0.45f, 0.15f));
        auto green = std::make_shared<lambertian>(green_tex);
        std::shared_ptr<texture> light_tex = std::make_shared<constant_texture>(vec3(15.0, 15.0, 15.0));
        auto light = std::make_shared<diffuse_light>(light_tex);
        std::vector<std::shared_ptr<hittable>> objects;
        objects.push_back(std::make_shared<xz_rect>(213.0, 343.0, 227.0, 332.0, 554.0, light));
        lights->objects.push_back(objects.back());
        objects.push_back(std::make_shared<flip_normals>(std::make_shared<yz_rect>(0.0, 555.0, 0.0, 555.0, 555.0, green)));
        objects.push_back(std::make_shared<yz_rect>(0.0, 555.0, 0.0, 555.0, 0.0, red));
        objects.push_back(std::make_shared<flip_normals>(std::make_shared<xz_rect>(0.0, 555.0, 0.0, 555.0, 555.0, white)));
        objects.push_back(std::make_shared<xz_rect>(0.0, 555.0, 0.0, 555.0, 0.0, white));
        objects.push_back(std::make_shared<flip_normals>(std::make_shared<xy_rect>(0.0, 555.0, 0.0, 555.0, 555.0, white)));
        objects.push_back(std::make_shared<sphere>(vec3(190, 90, 190), 90, std::make_shared<dielectric>(1.5)));
        lights->objects.push_back(objects.back());
        std::shared_ptr<material> aluminum = std::make_shared<metal>(vec3(0.8, 0.85, 0.88), 0.0);
        objects.push_back(std::make_shared<translate>(
            std::make_shared<rotate_y>(std::make_shared<box>(vec3(0.0, 0.0, 0.0), vec3(165.0, 330.0, 165.0), aluminum), 15.0),
            vec3(265.0, 0.0, 295.0)));
        this->world = hittable_list(objects);
        this->cam = camera(vec3(278.0, 278.0, -800.0), vec3(278.0, 278.0, 0.0), vec3(0.0, 1.0, 0.0), 40.0, aspect, 0.0, 10.0, 0.0, 1.0);
        this->background_type = BackgroundType::Black;
    }
};

// Scene/scene.h:176-249 with the `#if 1` at :219-225 switched OFF: the short box (:212-217) stays, no glass sphere, and the
// area light alone is in `lights` ("cornell_two_boxes", the reference's CornellBox.png strip); with `aluminum` also the
// `#if 0` at :228-231 ON: the scene of Sampling/12.2.png ("cornell_aluminum_box").  Twin of scenes.cpp cornell_box_scene.
class cornell_no_glass_scene : public scene {
public:
    cornell_no_glass_scene(double aspect, bool aluminum_tall_box) : scene() {
        std::shared_ptr<texture> red_tex = std::make_shared<constant_texture>(vec3(0.65f, 0.05f, 0.05f));
        auto red = std::make_shared<lambertian>(red_tex);
        std::shared_ptr<texture> white_tex = std::make_shared<constant_texture>(vec3(0.73f, 0.73f, 0.73f));
        auto white = std::make_shared<lambertian>(white_tex);
        std::shared_ptr<texture> green_tex = std::make_shared<constant_texture>(vec3(0.12f, 0.45f, 0.15f));
        auto green = std::make_shared<lambertian>(green_tex);
        std::shared_ptr<texture> light_tex = std::make_shared<constant_texture>(vec3(15.0, 15.0, 15.0));
        auto light = std::make_shared<diffuse_light>(light_tex);
        std::vector<std::shared_ptr<hittable>> objects;
        objects.push_back(std::make_shared<xz_rect>(213.0, 343.0, 227.0, 332.0, 554.0, light));
        lights->objects.push_back(objects.back());
        objects.push_back(std::make_shared<flip_normals>(std::make_shared<yz_rect>(0.0, 555.0, 0.0, 555.0, 555.0, green)));
        objects.push_back(std::make_shared<yz_rect>(0.0, 555.0, 0.0, 555.0, 0.0, red));
        objects.push_back(std::make_shared<flip_normals>(std::make_shared<xz_rect>(0.0, 555.0, 0.0, 555.0, 555.0, white)));
        objects.push_back(std::make_shared<xz_rect>(0.0, 555.0, 0.0, 555.0, 0.0, white));
        objects.push_back(std::make_shared<flip_normals>(std::make_shared<xy_rect>(0.0, 555.0, 0.0, 555.0, 555.0, white)));
        objects.push_back(std::make_shared<translate>(
            std::make_shared<rotate_y>(std::make_shared<box>(vec3(0.0, 0.0, 0.0), vec3(165.0, 165.0, 165.0), white), -18.0),
            vec3(130.0, 0.0, 65.0)));
        std::shared_ptr<material> tall = white;
        if (aluminum_tall_box) tall = std::make_shared<metal>(vec3(0.8, 0.85, 0.88), 0.0);
        objects.push_back(std::make_shared<translate>(
            std::make_shared<rotate_y>(std::make_shared<box>(vec3(0.0, 0.0, 0.0), vec3(165.0, 330.0, 165.0), tall), 15.0),
            vec3(265.0, 0.0, 295.0)));
        this->world = hittable_list(objects);
        this->cam = camera(vec3(278.0, 278.0, -800.0), vec3(278.0, 278.0, 0.0), vec3(0.0, 1.0, 0.0), 40.0, aspect, 0.0, 10.0, 0.0, 1.0);
        this->background_type = BackgroundType::Black;
    }
};

// cfg2: Book-2 Perlin spheres + earth image_texture sphere.
class two_perlin_spheres_earth_scene : public scene {
public:
    two_perlin_spheres_earth_scene(double aspect, std::shared_ptr<image_texture::byte_array> earth, int enx, int eny) : scene() {
        std::shared_ptr<texture> pertext = std::make_shared<noise_texture>(4.0);
        std::shared_ptr<texture> earthtex = std::make_shared<image_texture>(earth, enx, eny);
        Add(std::make_shared<sphere>(vec3(0, -1000, 0), 1000.0, std::make_shared<lambertian>(pertext)));
        Add(std::make_shared<sphere>(vec3(0, 2, 2.2), 2.0, std::make_shared<lambertian>(pertext)));
        Add(std::make_shared<sphere>(vec3(0, 2, -2.2), 2.0, std::make_shared<lambertian>(earthtex)));
        this->cam = camera(vec3(13, 2, 3), vec3(0, 2, 0), vec3(0.0, 1.0, 0.0), 40.0, aspect, 0.0, 10.0, 0.0, 1.0);
    }
};

// cfg3: Book-2 final scene.  bvh_node takes a raw hittable** (hittable.h:45): the owning shared_ptrs are
// kept alive in members.  Uses the two-token bvh_node overlay of oracle/Makefile.
class book2_final_scene : public scene {
public:
    std::vector<std::shared_ptr<hittable>> keep;
    std::vector<hittable*> box_ptrs, sphere_ptrs;

    book2_final_scene(double aspect, std::shared_ptr<image_texture::byte_array> earth, int enx, int eny) : scene() {
        std::uniform_real_distribution<double> uniform;
        std::minstd_rand engine;

        std::shared_ptr<texture> ground_tex = std::make_shared<constant_texture>(vec3(0.48, 0.83, 0.53));
        auto ground = std::make_shared<lambertian>(ground_tex);
        for (int i = 0; i < 20; i++)
            for (int j = 0; j < 20; j++) {
                double w = 100.0;
                double x0 = -1000.0 + i * w, z0 = -1000.0 + j * w, y0 = 0.0;
                double x1 = x0 + w, y1 = 100.0 * (uniform(engine) + 0.01), z1 = z0 + w;
                keep.push_back(std::make_shared<box>(vec3(x0, y0, z0), vec3(x1, y1, z1), ground));
                box_ptrs.push_back(keep.back().get());
            }
        Add(std::shared_ptr<hittable>(new bvh_node(box_ptrs.data(), (int)box_ptrs.size(), 0.0, 1.0)));

        std::shared_ptr<texture> light_tex = std::make_shared<constant_texture>(vec3(7, 7, 7));
        auto light = std::make_shared<diffuse_light>(light_tex);
        Add(std::make_shared<xz_rect>(123, 423, 147, 412, 554, light));
        lights->objects.push_back(world.objects.back());

        std::shared_ptr<texture> ms_tex = std::make_shared<constant_texture>(vec3(0.7, 0.3, 0.1));
        auto ms = std::make_shared<moving_sphere>(vec3(400, 400, 200), 50, std::make_shared<lambertian>(ms_tex));
        movement_linear m;
        m.center1 = vec3(430, 400, 200);
        m.time0 = 0.0;
        m.time1 = 1.0;
        ms->set_movement(m);
        Add(ms);
        Add(std::make_shared<sphere>(vec3(260, 150, 45), 50, std::make_shared<dielectric>(1.5)));
        Add(std::make_shared<sphere>(vec3(0, 150, 145), 50, std::make_shared<metal>(vec3(0.8, 0.8, 0.9), 1.0)));

        auto boundary = std::make_shared<sphere>(vec3(360, 150, 145), 70, std::make_shared<dielectric>(1.5));
        Add(boundary);
        std::shared_ptr<texture> blue = std::make_shared<constant_texture>(vec3(0.2, 0.4, 0.9));
        Add(std::make_shared<constant_medium>(boundary, 0.2, std::make_shared<isotropic>(blue)));
        auto fog_boundary = std::make_shared<sphere>(vec3(0, 0, 0), 5000, std::make_shared<dielectric>(1.5));
        std::shared_ptr<texture> fog = std::make_shared<constant_texture>(vec3(1, 1, 1));
        Add(std::make_shared<constant_medium>(fog_boundary, 0.0001, std::make_shared<isotropic>(fog)));

        std::shared_ptr<texture> earthtex = std::make_shared<image_texture>(earth, enx, eny);
        Add(std::make_shared<sphere>(vec3(400, 200, 400), 100, std::make_shared<lambertian>(earthtex)));
        std::shared_ptr<texture> pertext = std::make_shared<noise_texture>(0.1);
        Add(std::make_shared<sphere>(vec3(220, 280, 300), 80, std::make_shared<lambertian>(pertext)));

        std::shared_ptr<texture> white_tex = std::make_shared<constant_texture>(vec3(0.73, 0.73, 0.73));
        auto white = std::make_shared<lambertian>(white_tex);
        for (int j = 0; j < 1000; j++) {
            double x = 165 * uniform(engine);
            double y = 165 * uniform(engine);
            double z = 165 * uniform(engine);
            keep.push_back(std::make_shared<sphere>(vec3(x, y, z), 10, white));
            sphere_ptrs.push_back(keep.back().get());
        }
        std::shared_ptr<hittable> sphere_bvh(new bvh_node(sphere_ptrs.data(), (int)sphere_ptrs.size(), 0.0, 1.0));
        Add(std::make_shared<translate>(std::make_shared<rotate_y>(sphere_bvh, 15.0), vec3(-100, 270, 395)));

        this->cam = camera(vec3(478, 278, -600), vec3(278, 278, 0), vec3(0.0, 1.0, 0.0), 40.0, aspect, 0.0, 10.0, 0.0, 1.0);
        this->background_type = BackgroundType::Black;
    }
};

// Book 2 "Cornell smoke" (Volume.png): both boxes as constant_medium boundaries.  Twin of scenes.cpp cornell_smoke_scene.
class cornell_smoke_scene : public scene {
public:
    cornell_smoke_scene(double aspect) : scene() {
        std::shared_ptr<texture> red_tex = std::make_shared<constant_texture>(vec3(0.65f, 0.05f, 0.05f));
        auto red = std::make_shared<lambertian>(red_tex);
        std::shared_ptr<texture> white_tex = std::make_shared<constant_texture>(vec3(0.73f, 0.73f, 0.73f));
        auto white = std::make_shared<lambertian>(white_tex);
        std::shared_ptr<texture> green_tex = std::make_shared<constant_texture>(vec3(0.12f, 0.45f, 0.15f));
        auto green = std::make_shared<lambertian>(green_tex);
        std::shared_ptr<texture> light_tex = std::make_shared<constant_texture>(vec3(7.0, 7.0, 7.0));
        auto light = std::make_shared<diffuse_light>(light_tex);
        std::vector<std::shared_ptr<hittable>> objects;
        objects.push_back(std::make_shared<xz_rect>(113.0, 443.0, 127.0, 432.0, 554.0, light));
        lights->objects.push_back(objects.back());
        objects.push_back(std::make_shared<flip_normals>(std::make_shared<yz_rect>(0.0, 555.0, 0.0, 555.0, 555.0, green)));
        objects.push_back(std::make_shared<yz_rect>(0.0, 555.0, 0.0, 555.0, 0.0, red));
        objects.push_back(std::make_shared<flip_normals>(std::make_shared<xz_rect>(0.0, 555.0, 0.0, 555.0, 555.0, white)));
        objects.push_back(std::make_shared<xz_rect>(0.0, 555.0, 0.0, 555.0, 0.0, white));
        objects.push_back(std::make_shared<flip_normals>(std::make_shared<xy_rect>(0.0, 555.0, 0.0, 555.0, 555.0, white)));
        std::shared_ptr<hittable> b1 = std::make_shared<translate>(
            std::make_shared<rotate_y>(std::make_shared<box>(vec3(0.0, 0.0, 0.0), vec3(165.0, 165.0, 165.0), white), -18.0),
            vec3(130.0, 0.0, 65.0));
        std::shared_ptr<hittable> b2 = std::make_shared<translate>(
            std::make_shared<rotate_y>(std::make_shared<box>(vec3(0.0, 0.0, 0.0), vec3(165.0, 330.0, 165.0), white), 15.0),
            vec3(265.0, 0.0, 295.0));
        std::shared_ptr<texture> smoke_white = std::make_shared<constant_texture>(vec3(1.0, 1.0, 1.0));
        std::shared_ptr<texture> smoke_black = std::make_shared<constant_texture>(vec3(0.0, 0.0, 0.0));
        objects.push_back(std::make_shared<constant_medium>(b1, 0.01, std::make_shared<isotropic>(smoke_white)));
        objects.push_back(std::make_shared<constant_medium>(b2, 0.01, std::make_shared<isotropic>(smoke_black)));
        this->world = hittable_list(objects);
        this->cam = camera(vec3(278.0, 278.0, -800.0), vec3(278.0, 278.0, 0.0), vec3(0.0, 1.0, 0.0), 40.0, aspect, 0.0, 10.0, 0.0, 1.0);
        this->background_type = BackgroundType::Black;
    }
};

// Book 2 "two checker spheres".  Twin of scenes.cpp checker_spheres_scene.
class checker_spheres_scene : public scene {
public:
    checker_spheres_scene(double aspect) : scene() {
        std::shared_ptr<texture> even = std::make_shared<constant_texture>(vec3(0.2, 0.3, 0.1));
        std::shared_ptr<texture> odd = std::make_shared<constant_texture>(vec3(0.9, 0.9, 0.9));
        std::shared_ptr<texture> checker = std::make_shared<checker_texture>(even, odd);
        Add(std::make_shared<sphere>(vec3(0, -10, 0), 10.0, std::make_shared<lambertian>(checker)));
        Add(std::make_shared<sphere>(vec3(0, 10, 0), 10.0, std::make_shared<lambertian>(checker)));
        this->cam = camera(vec3(13, 2, 3), vec3(0, 0, 0), vec3(0.0, 1.0, 0.0), 20.0, aspect, 0.0, 10.0, 0.0, 1.0);
    }
};
