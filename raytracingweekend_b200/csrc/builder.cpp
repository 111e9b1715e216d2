// builder.cpp — host-side scene graph + flattener behind the rtw_* builder calls.
//
// The builder mirrors the reference's scene-construction API one call per constructor
// (see include/rtw_b200.h for the file:line of each) and flattens the resulting graph into
// the POD arrays of `rtw_flat_scene`.  Semantics the flattener preserves:
//   * list order (tie-breaking at equal t, Q38) — nested hittable_lists are inlined in order;
//   * box = 6 rects in the order of hittable_list.h:76-95, min-side ones flipped;
//   * flip_normals folds to a parity bit, translate/rotate_y to a transform chain (Q12);
//   * bvh_node trees keep their exact topology (imported node by node, or built here with the
//     reference's algorithm on the same libstdc++: random axis, std::sort(&l[0], &l[n-1]), Q3);
//   * constant_medium records how many independent free-flight draws one color() call makes
//     (2 per enclosing hittable_list pass pair, Q5).
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <map>
#include <memory>
#include <random>
#include <string>
#include <vector>

#include "flat.hpp"

namespace rtw {

static thread_local std::string g_err;
void set_error(const std::string& msg) { g_err = msg; }
void clear_error() { g_err.clear(); }

void FlatStorage::refresh_view() {
    view.n_world_items = n_world_items;
    view.n_items = (int32_t)items.size();
    view.n_prims = (int32_t)prims.size();
    view.n_xfs = (int32_t)xfs.size();
    view.n_nodes = (int32_t)nodes.size();
    view.n_leaves = (int32_t)leaves.size();
    view.n_materials = (int32_t)materials.size();
    view.n_textures = (int32_t)textures.size();
    view.n_lights = (int32_t)lights.size();
    view.n_texel_bytes = (int32_t)texels.size();
    view.render_type = render_type;
    view.background_type = background_type;
    view.items = items.data();
    view.prims = prims.data();
    view.xfs = xfs.data();
    view.nodes = nodes.data();
    view.leaves = leaves.data();
    view.materials = materials.data();
    view.textures = textures.data();
    view.lights = lights.data();
    view.texels = texels.empty() ? nullptr : texels.data();
    view.perlin_ranvec = perlin_ranvec.empty() ? nullptr : perlin_ranvec.data();
    view.perlin_perm = perlin_perm.empty() ? nullptr : perlin_perm.data();
    view.camera = camera;
}

enum Kind {
    K_TEX_CONSTANT,
    K_TEX_CHECKER,
    K_TEX_NOISE,
    K_TEX_IMAGE,
    K_MAT_LAMBERTIAN,
    K_MAT_METAL,
    K_MAT_DIELECTRIC,
    K_MAT_DIFFUSE_LIGHT,
    K_MAT_ISOTROPIC,
    K_SPHERE,
    K_MOVING_SPHERE,
    K_XY_RECT,
    K_XZ_RECT,
    K_YZ_RECT,
    K_BOX,
    K_FLIP,
    K_TRANSLATE,
    K_ROTATE_Y,
    K_MEDIUM,
    K_LIST,
    K_BVH
};

static bool is_texture(int k) { return k >= K_TEX_CONSTANT && k <= K_TEX_IMAGE; }
static bool is_material(int k) { return k >= K_MAT_LAMBERTIAN && k <= K_MAT_ISOTROPIC; }
static bool is_hittable(int k) { return k >= K_SPHERE && k <= K_BVH; }

struct Aabb {
    double mn[3] = {0, 0, 0}, mx[3] = {0, 0, 0};  // aabb() leaves vec3() zeros (aabb.h:11, vec3.h:12)
};

struct Node {
    int kind = 0;
    std::vector<rtw_id> ch;  // children / referenced ids
    double v[12] = {0};
    int img_nx = 0, img_ny = 0;
    std::shared_ptr<std::vector<uint8_t>> img;
    // rotate_y keeps the box computed at construction (hittable.h:334-372)
    bool hasbox = false;
    Aabb bbox;
};

}  // namespace rtw

using namespace rtw;

struct rtw_builder {
    std::vector<Node> nodes;
    rtw_id world = -1;
    rtw_id lights = -1;
    int render_type = RTW_RENDER_SHADED;
    int background_type = RTW_BG_GRADIENT;  // Scene/scene.h:38-39 defaults
    bool have_camera = false;
    rtw_flat_camera cam{};
    bool have_perlin = false;
    std::vector<double> ranvec;
    std::vector<int32_t> perm;
    // hittable.h:63-64: function-local statics in the reference; one process = one builder here.
    std::uniform_int_distribution<int> bvh_uniform{0, 2};
    std::minstd_rand bvh_engine;
    FlatStorage flat;
    std::map<rtw_id, int> flat_material_index;  // filled by rtw_builder_flatten

    bool valid(rtw_id id) const { return id >= 0 && (size_t)id < nodes.size(); }
    rtw_id add(Node&& n) {
        nodes.push_back(std::move(n));
        return (rtw_id)nodes.size() - 1;
    }
};

// ------------------------------------------------------------------------------------------------
// bounding boxes, reference semantics
// ------------------------------------------------------------------------------------------------
namespace {

const double kPad = (double)0.0001f;  // hittable.h:169, 204, 261 (float literal)

Aabb surrounding(const Aabb& a, const Aabb& b) {  // aabb.h:49-62
    Aabb r;
    for (int i = 0; i < 3; i++) {
        r.mn[i] = fmin(a.mn[i], b.mn[i]);
        r.mx[i] = fmax(a.mx[i], b.mx[i]);
    }
    return r;
}

// hittable::bounding_box(t0, t1, box): `box` is an in/out parameter exactly like the reference
// (hittable_list::bounding_box returns true without writing it, hittable_list.h:39-42).
bool bounding_box(const rtw_builder* b, rtw_id id, double t0, double t1, Aabb& box) {
    const Node& n = b->nodes[id];
    switch (n.kind) {
        case K_SPHERE: {  // sphere.h:13-17
            double r = n.v[3];
            for (int i = 0; i < 3; i++) {
                box.mn[i] = n.v[i] - r;
                box.mx[i] = n.v[i] + r;
            }
            return true;
        }
        case K_MOVING_SPHERE: {  // sphere.h:27-33 (ignores t0,t1 — Q36)
            double r = n.v[3];
            Aabb b0, b1;
            for (int i = 0; i < 3; i++) {
                b0.mn[i] = n.v[i] - r;
                b0.mx[i] = n.v[i] + r;
                b1.mn[i] = n.v[4 + i] - r;
                b1.mx[i] = n.v[4 + i] + r;
            }
            box = surrounding(b0, b1);
            return true;
        }
        case K_XY_RECT:  // hittable.h:167-171
            box.mn[0] = n.v[0], box.mn[1] = n.v[2], box.mn[2] = n.v[4] - kPad;
            box.mx[0] = n.v[1], box.mx[1] = n.v[3], box.mx[2] = n.v[4] + kPad;
            return true;
        case K_XZ_RECT:  // hittable.h:202-206
            box.mn[0] = n.v[0], box.mn[1] = n.v[4] - kPad, box.mn[2] = n.v[2];
            box.mx[0] = n.v[1], box.mx[1] = n.v[4] + kPad, box.mx[2] = n.v[3];
            return true;
        case K_YZ_RECT:  // hittable.h:259-263
            box.mn[0] = n.v[4] - kPad, box.mn[1] = n.v[0], box.mn[2] = n.v[2];
            box.mx[0] = n.v[4] + kPad, box.mx[1] = n.v[1], box.mx[2] = n.v[3];
            return true;
        case K_BOX:  // hittable_list.h:106-110
            for (int i = 0; i < 3; i++) {
                box.mn[i] = n.v[i];
                box.mx[i] = n.v[3 + i];
            }
            return true;
        case K_FLIP:  // hittable.h:286-289
            return bounding_box(b, n.ch[0], t0, t1, box);
        case K_TRANSLATE:  // hittable.h:313-324
            if (bounding_box(b, n.ch[0], t0, t1, box)) {
                for (int i = 0; i < 3; i++) {
                    box.mn[i] += n.v[i];
                    box.mx[i] += n.v[i];
                }
                return true;
            }
            return false;
        case K_ROTATE_Y:  // hittable.h:405-409
            box = n.bbox;
            return n.hasbox;
        case K_MEDIUM:  // hittable.h:481-484
            return bounding_box(b, n.ch[0], t0, t1, box);
        case K_LIST:  // hittable_list.h:39-42 — true, box untouched (Q6)
            return true;
        case K_BVH:  // hittable.h:131-135
            for (int i = 0; i < 3; i++) {
                box.mn[i] = n.v[i];
                box.mx[i] = n.v[3 + i];
            }
            return true;
        default:
            return false;
    }
}

int fail(int code, const std::string& msg) {
    set_error(msg);
    return code;
}
rtw_id fail_id(const std::string& msg) {
    set_error(msg);
    return -1;
}

bool check_kind(const rtw_builder* b, rtw_id id, bool (*pred)(int), const char* what) {
    if (!b || !b->valid(id) || !pred(b->nodes[id].kind)) {
        set_error(std::string("invalid ") + what + " id");
        return false;
    }
    return true;
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// C ABI: construction
// ------------------------------------------------------------------------------------------------
extern "C" {

const char* rtw_last_error(void) { return g_err.c_str(); }
int rtw_abi_version(void) { return RTW_B200_ABI_VERSION; }

rtw_builder* rtw_builder_create(void) {
    clear_error();
    return new (std::nothrow) rtw_builder();
}
void rtw_builder_destroy(rtw_builder* b) { delete b; }

rtw_id rtw_constant_texture(rtw_builder* b, double r, double g, double bl) {
    if (!b) return fail_id("null builder");
    Node n;
    n.kind = K_TEX_CONSTANT;
    n.v[0] = r, n.v[1] = g, n.v[2] = bl;
    return b->add(std::move(n));
}
rtw_id rtw_checker_texture(rtw_builder* b, rtw_id t0, rtw_id t1) {
    if (!check_kind(b, t0, is_texture, "texture") || !check_kind(b, t1, is_texture, "texture")) return -1;
    Node n;
    n.kind = K_TEX_CHECKER;
    n.ch = {t0, t1};  // even = t0, odd = t1 (texture.h:33)
    return b->add(std::move(n));
}
rtw_id rtw_noise_texture(rtw_builder* b, double scale) {
    if (!b) return fail_id("null builder");
    Node n;
    n.kind = K_TEX_NOISE;
    n.v[0] = scale;
    return b->add(std::move(n));
}
rtw_id rtw_image_texture(rtw_builder* b, const uint8_t* rgb8, int nx, int ny) {
    if (!b || !rgb8 || nx <= 0 || ny <= 0) return fail_id("image_texture: bad arguments");
    Node n;
    n.kind = K_TEX_IMAGE;
    n.img_nx = nx, n.img_ny = ny;
    n.img = std::make_shared<std::vector<uint8_t>>(rgb8, rgb8 + (size_t)3 * nx * ny);
    return b->add(std::move(n));
}

static rtw_id textured_material(rtw_builder* b, int kind, rtw_id tex) {
    if (!check_kind(b, tex, is_texture, "texture")) return -1;
    Node n;
    n.kind = kind;
    n.ch = {tex};
    return b->add(std::move(n));
}
rtw_id rtw_lambertian(rtw_builder* b, rtw_id tex) { return textured_material(b, K_MAT_LAMBERTIAN, tex); }
rtw_id rtw_diffuse_light(rtw_builder* b, rtw_id tex) { return textured_material(b, K_MAT_DIFFUSE_LIGHT, tex); }
rtw_id rtw_isotropic(rtw_builder* b, rtw_id tex) { return textured_material(b, K_MAT_ISOTROPIC, tex); }
rtw_id rtw_metal(rtw_builder* b, double r, double g, double bl, double fuzz) {
    if (!b) return fail_id("null builder");
    Node n;
    n.kind = K_MAT_METAL;
    n.v[0] = r, n.v[1] = g, n.v[2] = bl, n.v[3] = fuzz;
    return b->add(std::move(n));
}
rtw_id rtw_dielectric(rtw_builder* b, double ref_idx) {
    if (!b) return fail_id("null builder");
    Node n;
    n.kind = K_MAT_DIELECTRIC;
    n.v[0] = ref_idx;
    return b->add(std::move(n));
}

rtw_id rtw_sphere(rtw_builder* b, double cx, double cy, double cz, double radius, rtw_id material) {
    if (!check_kind(b, material, is_material, "material")) return -1;
    Node n;
    n.kind = K_SPHERE;
    n.v[0] = cx, n.v[1] = cy, n.v[2] = cz, n.v[3] = radius;
    n.ch = {material};
    return b->add(std::move(n));
}
rtw_id rtw_moving_sphere(rtw_builder* b, double cx, double cy, double cz, double radius, rtw_id material, double c1x,
                         double c1y, double c1z, double time0, double time1) {
    if (!check_kind(b, material, is_material, "material")) return -1;
    Node n;
    n.kind = K_MOVING_SPHERE;
    n.v[0] = cx, n.v[1] = cy, n.v[2] = cz, n.v[3] = radius;
    n.v[4] = c1x, n.v[5] = c1y, n.v[6] = c1z, n.v[7] = time0, n.v[8] = time1;
    n.ch = {material};
    return b->add(std::move(n));
}
static rtw_id rect(rtw_builder* b, int kind, double a0, double a1, double b0, double b1, double k, rtw_id material) {
    if (!check_kind(b, material, is_material, "material")) return -1;
    Node n;
    n.kind = kind;
    n.v[0] = a0, n.v[1] = a1, n.v[2] = b0, n.v[3] = b1, n.v[4] = k;
    n.ch = {material};
    return b->add(std::move(n));
}
rtw_id rtw_xy_rect(rtw_builder* b, double x0, double x1, double y0, double y1, double k, rtw_id m) {
    return rect(b, K_XY_RECT, x0, x1, y0, y1, k, m);
}
rtw_id rtw_xz_rect(rtw_builder* b, double x0, double x1, double z0, double z1, double k, rtw_id m) {
    return rect(b, K_XZ_RECT, x0, x1, z0, z1, k, m);
}
rtw_id rtw_yz_rect(rtw_builder* b, double y0, double y1, double z0, double z1, double k, rtw_id m) {
    return rect(b, K_YZ_RECT, y0, y1, z0, z1, k, m);
}
rtw_id rtw_box(rtw_builder* b, const double p0[3], const double p1[3], rtw_id material) {
    if (!p0 || !p1) return fail_id("box: null corner");
    if (!check_kind(b, material, is_material, "material")) return -1;
    Node n;
    n.kind = K_BOX;
    for (int i = 0; i < 3; i++) n.v[i] = p0[i], n.v[3 + i] = p1[i];
    n.ch = {material};
    return b->add(std::move(n));
}
rtw_id rtw_flip_normals(rtw_builder* b, rtw_id child) {
    if (!check_kind(b, child, is_hittable, "hittable")) return -1;
    Node n;
    n.kind = K_FLIP;
    n.ch = {child};
    return b->add(std::move(n));
}
rtw_id rtw_translate(rtw_builder* b, rtw_id child, double ox, double oy, double oz) {
    if (!check_kind(b, child, is_hittable, "hittable")) return -1;
    Node n;
    n.kind = K_TRANSLATE;
    n.v[0] = ox, n.v[1] = oy, n.v[2] = oz;
    n.ch = {child};
    return b->add(std::move(n));
}
rtw_id rtw_rotate_y_sincos(rtw_builder* b, rtw_id child, double sin_theta, double cos_theta) {
    if (!check_kind(b, child, is_hittable, "hittable")) return -1;
    Node n;
    n.kind = K_ROTATE_Y;
    n.v[0] = sin_theta, n.v[1] = cos_theta;
    n.ch = {child};
    // hittable.h:340-371: box of the 8 rotated corners of bounding_box(0, 1)
    Aabb bb;
    n.hasbox = bounding_box(b, child, 0, 1, bb);
    double mn[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, mx[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
    for (int i = 0; i < 2; i++)
        for (int j = 0; j < 2; j++)
            for (int k = 0; k < 2; k++) {
                // NB the reference indexes all three axes with `i` (hittable.h:350-352); kept.
                double x = i * bb.mx[0] + (1 - i) * bb.mn[0];
                double y = i * bb.mx[1] + (1 - i) * bb.mn[1];
                double z = i * bb.mx[2] + (1 - i) * bb.mn[2];
                double newx = cos_theta * x + sin_theta * z;
                double newz = -sin_theta * x + cos_theta * z;
                double t[3] = {newx, y, newz};
                for (int c = 0; c < 3; c++) {
                    if (t[c] > mx[c]) mx[c] = t[c];
                    if (t[c] < mn[c]) mn[c] = t[c];
                }
            }
    for (int c = 0; c < 3; c++) n.bbox.mn[c] = mn[c], n.bbox.mx[c] = mx[c];
    return b->add(std::move(n));
}
rtw_id rtw_rotate_y(rtw_builder* b, rtw_id child, double angle) {
    double radians = ((double)M_PI / 180.0) * angle;  // hittable.h:337-339
    return rtw_rotate_y_sincos(b, child, sin(radians), cos(radians));
}
rtw_id rtw_constant_medium(rtw_builder* b, rtw_id boundary, double density, rtw_id material) {
    if (!check_kind(b, boundary, is_hittable, "hittable") || !check_kind(b, material, is_material, "material")) return -1;
    Node n;
    n.kind = K_MEDIUM;
    n.v[0] = density;
    n.ch = {boundary, material};
    return b->add(std::move(n));
}
rtw_id rtw_hittable_list(rtw_builder* b, const rtw_id* children, int nch) {
    if (!b || nch < 0 || (nch > 0 && !children)) return fail_id("hittable_list: bad arguments");
    Node n;
    n.kind = K_LIST;
    for (int i = 0; i < nch; i++) {
        if (!check_kind(b, children[i], is_hittable, "hittable")) return -1;
        n.ch.push_back(children[i]);
    }
    return b->add(std::move(n));
}

rtw_id rtw_bvh_node_import(rtw_builder* b, rtw_id left, rtw_id right, const double mn[3], const double mx[3]) {
    if (!mn || !mx) return fail_id("bvh_node_import: null box");
    if (!check_kind(b, left, is_hittable, "hittable") || !check_kind(b, right, is_hittable, "hittable")) return -1;
    Node n;
    n.kind = K_BVH;
    n.ch = {left, right};
    for (int i = 0; i < 3; i++) n.v[i] = mn[i], n.v[3 + i] = mx[i];
    return b->add(std::move(n));
}

rtw_id rtw_bvh_node(rtw_builder* b, rtw_id* l, int n, double time0, double time1) {
    if (!b || !l || n < 1) return fail_id("bvh_node: need n >= 1 children");
    for (int i = 0; i < n; i++)
        if (!check_kind(b, l[i], is_hittable, "hittable")) return -1;
    // hittable.h:63-66
    int axis = b->bvh_uniform(b->bvh_engine);
    std::sort(&l[0], &l[n - 1], [b, axis](rtw_id lhs, rtw_id rhs) {
        Aabb bl, br;  // hittable.h:51-58
        bounding_box(b, lhs, 0, 0, bl);
        bounding_box(b, rhs, 0, 0, br);
        return bl.mn[axis] < br.mn[axis];
    });
    rtw_id left, right;
    if (n == 1) {
        left = right = l[0];
    } else if (n == 2) {
        left = l[0];
        right = l[1];
    } else {  // as intended: the second assignment targets `right`
        left = rtw_bvh_node(b, l, n / 2, time0, time1);
        if (left < 0) return -1;
        right = rtw_bvh_node(b, l + n / 2, n - n / 2, time0, time1);
        if (right < 0) return -1;
    }
    Aabb bl, br;  // hittable.h:87-94
    bounding_box(b, left, time0, time1, bl);
    bounding_box(b, right, time0, time1, br);
    Aabb box = surrounding(bl, br);
    return rtw_bvh_node_import(b, left, right, box.mn, box.mx);
}

int rtw_bounding_box(const rtw_builder* b, rtw_id node, double t0, double t1, double mn[3], double mx[3]) {
    if (!check_kind(b, node, is_hittable, "hittable") || !mn || !mx) return -1;
    Aabb bb;
    bool r = bounding_box(b, node, t0, t1, bb);
    for (int i = 0; i < 3; i++) mn[i] = bb.mn[i], mx[i] = bb.mx[i];
    return r ? 1 : 0;
}

int rtw_set_world(rtw_builder* b, rtw_id list) {
    if (!b || !b->valid(list) || b->nodes[list].kind != K_LIST) return fail(RTW_ERR_INVALID, "set_world: need a hittable_list id");
    b->world = list;
    return RTW_OK;
}
int rtw_set_lights(rtw_builder* b, rtw_id list) {
    if (!b) return fail(RTW_ERR_INVALID, "null builder");
    if (list >= 0 && (!b->valid(list) || b->nodes[list].kind != K_LIST)) return fail(RTW_ERR_INVALID, "set_lights: need a hittable_list id or -1");
    b->lights = list < 0 ? -1 : list;
    return RTW_OK;
}
int rtw_set_render_type(rtw_builder* b, int t) {
    if (!b || (t != RTW_RENDER_SHADED && t != RTW_RENDER_NORMAL)) return fail(RTW_ERR_INVALID, "bad render type");
    b->render_type = t;
    return RTW_OK;
}
int rtw_set_background_type(rtw_builder* b, int t) {
    if (!b || (t != RTW_BG_BLACK && t != RTW_BG_GRADIENT)) return fail(RTW_ERR_INVALID, "bad background type");
    b->background_type = t;
    return RTW_OK;
}

int rtw_set_camera_members(rtw_builder* b, const rtw_flat_camera* cam) {
    if (!b || !cam) return fail(RTW_ERR_INVALID, "set_camera_members: null");
    b->cam = *cam;
    b->have_camera = true;
    return RTW_OK;
}

int rtw_set_camera(rtw_builder* b, const double lookfrom[3], const double lookat[3], const double vup[3], double vfov,
                   double aspect, double aperture, double focus_dist, double t0, double t1) {
    if (!b || !lookfrom || !lookat || !vup) return fail(RTW_ERR_INVALID, "set_camera: null");
    // camera.h:13-34, same operation order (vec3 ops are component-wise doubles)
    rtw_flat_camera c{};
    c.time0 = t0;
    c.time1 = t1;
    c.lens_radius = aperture / 2;
    double theta = vfov * static_cast<double>(M_PI) / 180.0;
    double half_height = tan(theta / 2);
    double half_width = aspect * half_height;
    auto norm = [](double* v) {
        double len = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
        v[0] /= len, v[1] /= len, v[2] /= len;
    };
    auto cross = [](const double* a, const double* bb, double* o) {  // vec3.h:54-59
        o[0] = (a[1] * bb[2] - a[2] * bb[1]);
        o[1] = (-(a[0] * bb[2] - a[2] * bb[0]));
        o[2] = (a[0] * bb[1] - a[1] * bb[0]);
    };
    for (int i = 0; i < 3; i++) c.origin[i] = lookfrom[i];
    for (int i = 0; i < 3; i++) c.w[i] = lookfrom[i] - lookat[i];
    norm(c.w);
    cross(vup, c.w, c.u);
    norm(c.u);
    cross(c.w, c.u, c.v);
    for (int i = 0; i < 3; i++) {
        // origin - half_width*focus_dist*u - half_height*focus_dist*v - focus_dist*w  (camera.h:30)
        // operator precedence: ((half_width*focus_dist)*u) etc. — scalar products first (double*double),
        // then the implicit double->vec3 conversion (vec3.h:15) and component-wise multiply.
        c.lower_left_corner[i] = ((c.origin[i] - (half_width * focus_dist) * c.u[i]) - (half_height * focus_dist) * c.v[i]) -
                                 focus_dist * c.w[i];
        c.horizontal[i] = ((2.0 * half_width) * focus_dist) * c.u[i];  // camera.h:32
        c.vertical[i] = ((2.0 * half_height) * focus_dist) * c.v[i];   // camera.h:33
    }
    b->cam = c;
    b->have_camera = true;
    return RTW_OK;
}

int rtw_set_perlin_tables(rtw_builder* b, const double ranvec[768], const int32_t px[256], const int32_t py[256],
                          const int32_t pz[256]) {
    if (!b || !ranvec || !px || !py || !pz) return fail(RTW_ERR_INVALID, "set_perlin_tables: null");
    b->ranvec.assign(ranvec, ranvec + 768);
    b->perm.resize(768);
    memcpy(b->perm.data(), px, 256 * sizeof(int32_t));
    memcpy(b->perm.data() + 256, py, 256 * sizeof(int32_t));
    memcpy(b->perm.data() + 512, pz, 256 * sizeof(int32_t));
    b->have_perlin = true;
    return RTW_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// Perlin tables exactly as noise.h:166-213 builds them (each generator owns a fresh default-seeded
// std::minstd_rand, so the three permutations are identical — SURVEY F6).
// ------------------------------------------------------------------------------------------------
namespace rtw {
void default_perlin_tables(std::vector<double>& ranvec, std::vector<int32_t>& perm) {
    ranvec.resize(768);
    {
        std::uniform_real_distribution<double> uniform;
        std::minstd_rand engine;
        for (int i = 0; i < 256; i++) {
            // noise.h:207-210: three draws inside one constructor call; g++ evaluates them in the
            // order pinned by tests/test_host_parity.py against the reference's own tables.
            double a = -1.0 + 2.0 * uniform(engine);
            double bb = -1.0 + 2.0 * uniform(engine);
            double c = -1.0 + 2.0 * uniform(engine);
            double x = c, y = bb, z = a;  // g++ evaluates the three constructor arguments right to left
            double len = std::sqrt(x * x + y * y + z * z);
            ranvec[3 * i + 0] = x / len;
            ranvec[3 * i + 1] = y / len;
            ranvec[3 * i + 2] = z / len;
        }
    }
    perm.resize(768);
    for (int t = 0; t < 3; t++) {
        int32_t* p = perm.data() + 256 * t;
        for (int i = 0; i < 256; i++) p[i] = i;
        std::uniform_real_distribution<double> uniform;
        std::minstd_rand engine;
        for (int i = 255; i > 0; i--) {  // noise.h:171-177
            int target = int(uniform(engine) * (i + 1));
            int tmp = p[i];
            p[i] = p[target];
            p[target] = tmp;
        }
    }
}
}  // namespace rtw

// ------------------------------------------------------------------------------------------------
// Flattener
// ------------------------------------------------------------------------------------------------
namespace {

struct Link {
    int kind;
    double a, b, c;
    bool operator<(const Link& o) const {
        if (kind != o.kind) return kind < o.kind;
        if (a != o.a) return a < o.a;
        if (b != o.b) return b < o.b;
        return c < o.c;
    }
};
typedef std::vector<Link> Chain;  // outermost first

struct Ctx {
    Chain chain;
    int flip = 0;
    int mult = 1;
};

struct Flattener {
    rtw_builder* b;
    FlatStorage& fs;
    std::map<rtw_id, int> mat_map, tex_map;
    std::map<Chain, int> chain_map;
    bool uses_noise = false;
    bool ok = true;
    std::string err;
    bool run_open = false;  // last world item is a PRIMS run that may be extended

    Flattener(rtw_builder* bb, FlatStorage& f) : b(bb), fs(f) {}

    void error(const std::string& m) {
        if (ok) err = m;
        ok = false;
    }

    int intern_chain(const Chain& c) {
        if (c.empty()) return -1;
        auto it = chain_map.find(c);
        if (it != chain_map.end()) return it->second;
        int head = (int)fs.xfs.size();
        for (size_t i = 0; i < c.size(); i++) {
            rtw_flat_xf x{};
            x.kind = c[i].kind;
            x.next = (i + 1 < c.size()) ? head + (int)i + 1 : -1;
            x.a = c[i].a, x.b = c[i].b, x.c = c[i].c;
            fs.xfs.push_back(x);
        }
        chain_map[c] = head;
        return head;
    }

    int texture(rtw_id id) {
        auto it = tex_map.find(id);
        if (it != tex_map.end()) return it->second;
        const Node& n = b->nodes[id];
        rtw_flat_texture t{};
        t.even = t.odd = -1;
        int idx = (int)fs.textures.size();
        fs.textures.push_back(t);  // reserve the slot first: children get larger indices
        tex_map[id] = idx;
        switch (n.kind) {
            case K_TEX_CONSTANT:
                t.type = RTW_TEX_CONSTANT;
                t.color[0] = n.v[0], t.color[1] = n.v[1], t.color[2] = n.v[2];
                break;
            case K_TEX_CHECKER:
                t.type = RTW_TEX_CHECKER;
                t.even = texture(n.ch[0]);
                t.odd = texture(n.ch[1]);
                break;
            case K_TEX_NOISE:
                t.type = RTW_TEX_NOISE;
                t.scale = n.v[0];
                uses_noise = true;
                break;
            case K_TEX_IMAGE:
                t.type = RTW_TEX_IMAGE;
                t.image_offset = (int32_t)fs.texels.size();
                t.nx = n.img_nx, t.ny = n.img_ny;
                fs.texels.insert(fs.texels.end(), n.img->begin(), n.img->end());
                break;
        }
        fs.textures[idx] = t;
        return idx;
    }

    int material(rtw_id id) {
        auto it = mat_map.find(id);
        if (it != mat_map.end()) return it->second;
        const Node& n = b->nodes[id];
        rtw_flat_material m{};
        m.texture = -1;
        switch (n.kind) {
            case K_MAT_LAMBERTIAN: m.type = RTW_MAT_LAMBERTIAN; m.texture = texture(n.ch[0]); break;
            case K_MAT_DIFFUSE_LIGHT: m.type = RTW_MAT_DIFFUSE_LIGHT; m.texture = texture(n.ch[0]); break;
            case K_MAT_ISOTROPIC: m.type = RTW_MAT_ISOTROPIC; m.texture = texture(n.ch[0]); break;
            case K_MAT_METAL:
                m.type = RTW_MAT_METAL;
                m.albedo[0] = n.v[0], m.albedo[1] = n.v[1], m.albedo[2] = n.v[2];
                m.fuzz = n.v[3];
                break;
            case K_MAT_DIELECTRIC: m.type = RTW_MAT_DIELECTRIC; m.ref_idx = n.v[0]; break;
        }
        int idx = (int)fs.materials.size();
        fs.materials.push_back(m);
        mat_map[id] = idx;
        return idx;
    }

    void push_prim(int type, const Node& n, int mat, const Ctx& c, const double* v, int nv) {
        rtw_flat_prim p{};
        p.type = type;
        p.material = mat;
        p.flip = c.flip & 1;
        p.xf = intern_chain(c.chain);
        for (int i = 0; i < nv; i++) p.v[i] = v[i];
        (void)n;
        fs.prims.push_back(p);
    }

    // Append the primitives of `id` (prim / box / wrappers / nested list of those) to fs.prims.
    // Returns false if the subtree holds something that is not a pure primitive run.
    bool emit_prims(rtw_id id, Ctx c) {
        const Node& n = b->nodes[id];
        switch (n.kind) {
            case K_SPHERE: push_prim(RTW_PRIM_SPHERE, n, material(n.ch[0]), c, n.v, 4); return true;
            case K_MOVING_SPHERE: push_prim(RTW_PRIM_MOVING_SPHERE, n, material(n.ch[0]), c, n.v, 9); return true;
            case K_XY_RECT: push_prim(RTW_PRIM_XY_RECT, n, material(n.ch[0]), c, n.v, 5); return true;
            case K_XZ_RECT: push_prim(RTW_PRIM_XZ_RECT, n, material(n.ch[0]), c, n.v, 5); return true;
            case K_YZ_RECT: push_prim(RTW_PRIM_YZ_RECT, n, material(n.ch[0]), c, n.v, 5); return true;
            case K_BOX: {  // hittable_list.h:76-95
                const double *p0 = n.v, *p1 = n.v + 3;
                int m = material(n.ch[0]);
                Ctx f = c;
                f.flip ^= 1;
                double r[5];
                r[0] = p0[0], r[1] = p1[0], r[2] = p0[1], r[3] = p1[1];
                r[4] = p1[2]; push_prim(RTW_PRIM_XY_RECT, n, m, c, r, 5);
                r[4] = p0[2]; push_prim(RTW_PRIM_XY_RECT, n, m, f, r, 5);
                r[0] = p0[0], r[1] = p1[0], r[2] = p0[2], r[3] = p1[2];
                r[4] = p1[1]; push_prim(RTW_PRIM_XZ_RECT, n, m, c, r, 5);
                r[4] = p0[1]; push_prim(RTW_PRIM_XZ_RECT, n, m, f, r, 5);
                r[0] = p0[1], r[1] = p1[1], r[2] = p0[2], r[3] = p1[2];
                r[4] = p1[0]; push_prim(RTW_PRIM_YZ_RECT, n, m, c, r, 5);
                r[4] = p0[0]; push_prim(RTW_PRIM_YZ_RECT, n, m, f, r, 5);
                return true;
            }
            case K_FLIP: c.flip ^= 1; return emit_prims(n.ch[0], c);
            case K_TRANSLATE: c.chain.push_back({RTW_XF_TRANSLATE, n.v[0], n.v[1], n.v[2]}); return emit_prims(n.ch[0], c);
            case K_ROTATE_Y: c.chain.push_back({RTW_XF_ROTATE_Y, n.v[0], n.v[1], 0.0}); return emit_prims(n.ch[0], c);
            case K_LIST:
                for (rtw_id ch : n.ch)
                    if (!emit_prims(ch, c)) return false;
                return true;
            default: return false;
        }
    }

    static bool pure_prims(const rtw_builder* b, rtw_id id) {
        const Node& n = b->nodes[id];
        switch (n.kind) {
            case K_SPHERE: case K_MOVING_SPHERE: case K_XY_RECT: case K_XZ_RECT: case K_YZ_RECT: case K_BOX: return true;
            case K_FLIP: case K_TRANSLATE: case K_ROTATE_Y: return pure_prims(b, n.ch[0]);
            case K_LIST:
                for (rtw_id ch : n.ch)
                    if (!pure_prims(b, ch)) return false;
                return true;
            default: return false;
        }
    }

    // bvh_node subtree -> fs.nodes (pre-order). Returns child reference (>=0 node, <0 ~leaf).
    int emit_bvh(rtw_id id, const Ctx& leaf_ctx, std::map<rtw_id, int>& leaf_of) {
        const Node& n = b->nodes[id];
        if (n.kind == K_BVH) {
            int idx = (int)fs.nodes.size();
            fs.nodes.push_back(rtw_flat_node{});
            rtw_flat_node fn{};
            for (int i = 0; i < 3; i++) fn.bmin[i] = n.v[i], fn.bmax[i] = n.v[3 + i];
            fn.left = emit_bvh(n.ch[0], leaf_ctx, leaf_of);
            // n == 1 leaf: left == right (hittable.h:68-72) -> same reference twice
            fn.right = (n.ch[1] == n.ch[0]) ? fn.left : emit_bvh(n.ch[1], leaf_ctx, leaf_of);
            fs.nodes[idx] = fn;
            return idx;
        }
        auto it = leaf_of.find(id);
        if (it != leaf_of.end()) return ~it->second;
        if (!pure_prims(b, id)) {
            error("bvh_node leaf must be a primitive, box, or flip/translate/rotate_y/list of those");
            return ~0;
        }
        rtw_flat_leaf lf{};
        lf.first_prim = (int32_t)fs.prims.size();
        emit_prims(id, leaf_ctx);
        lf.prim_count = (int32_t)fs.prims.size() - lf.first_prim;
        int li = (int)fs.leaves.size();
        fs.leaves.push_back(lf);
        leaf_of[id] = li;
        return ~li;
    }

    // One object of the world list (or of a list nested in it).
    void emit_object(rtw_id id, Ctx c) {
        if (!ok) return;
        const Node& n = b->nodes[id];
        switch (n.kind) {
            case K_FLIP: c.flip ^= 1; emit_object(n.ch[0], c); return;
            case K_TRANSLATE: c.chain.push_back({RTW_XF_TRANSLATE, n.v[0], n.v[1], n.v[2]}); emit_object(n.ch[0], c); return;
            case K_ROTATE_Y: c.chain.push_back({RTW_XF_ROTATE_Y, n.v[0], n.v[1], 0.0}); emit_object(n.ch[0], c); return;
            case K_LIST: {  // nested hittable_list: inline in order; two passes per level (hittable_list.h:16-34)
                c.mult *= 2;
                for (rtw_id ch : n.ch) emit_object(ch, c);
                return;
            }
            case K_BVH: {
                run_open = false;
                rtw_flat_item it{};
                it.type = RTW_ITEM_BVH;
                it.xf = intern_chain(c.chain);
                it.material = -1;
                Ctx leaf_ctx;
                leaf_ctx.flip = c.flip;
                leaf_ctx.mult = c.mult;
                int first_node = (int)fs.nodes.size();
                std::map<rtw_id, int> leaf_of;
                int root = emit_bvh(id, leaf_ctx, leaf_of);
                it.first = root;
                it.count = (int)fs.nodes.size() - first_node;
                fs.items.push_back(it);
                return;
            }
            case K_MEDIUM: {
                run_open = false;
                if (!pure_prims(b, n.ch[0])) {
                    error("constant_medium boundary must be a primitive, box, or flip/translate/rotate_y/list of those");
                    return;
                }
                rtw_flat_item it{};
                it.type = RTW_ITEM_MEDIUM;
                it.xf = intern_chain(c.chain);
                it.first = (int32_t)fs.prims.size();
                Ctx bc;  // boundary prims are relative to the item's chain
                emit_prims(n.ch[0], bc);
                it.count = (int32_t)fs.prims.size() - it.first;
                it.material = material(n.ch[1]);
                it.flip = c.flip & 1;
                it.multiplicity = c.mult;
                it.density = n.v[0];
                fs.items.push_back(it);
                return;
            }
            default: {  // primitive or box: extend / open a PRIMS run
                int first = (int)fs.prims.size();
                emit_prims(id, c);
                int cnt = (int)fs.prims.size() - first;
                if (run_open) {
                    fs.items.back().count += cnt;
                } else {
                    rtw_flat_item it{};
                    it.type = RTW_ITEM_PRIMS;
                    it.xf = -1;
                    it.first = first;
                    it.count = cnt;
                    it.material = -1;
                    fs.items.push_back(it);
                    run_open = true;
                }
                return;
            }
        }
    }

    void emit_lights(rtw_id list) {
        const Node& l = b->nodes[list];
        for (rtw_id id : l.ch) {
            const Node& n = b->nodes[id];
            rtw_flat_light lt{};
            if (n.kind == K_XZ_RECT) {
                lt.type = RTW_LIGHT_XZ_RECT;
                for (int i = 0; i < 5; i++) lt.v[i] = n.v[i];
            } else if (n.kind == K_SPHERE) {
                lt.type = RTW_LIGHT_SPHERE;
                for (int i = 0; i < 4; i++) lt.v[i] = n.v[i];
            } else if (n.kind == K_MOVING_SPHERE) {
                // sphere.h:92 evaluates the centre at time FLT_MAX -> garbage (Q15)
                error("moving_sphere in the lights list is not supported (reference evaluates it at time FLT_MAX)");
            } else if (n.kind == K_LIST) {
                error("nested hittable_list in the lights list is not supported");
            } else {
                lt.type = RTW_LIGHT_DEFAULT;  // hittable.h:36-37
            }
            fs.lights.push_back(lt);
        }
    }

    bool run() {
        fs = FlatStorage();
        if (b->world < 0) {
            error("no world set");
            return false;
        }
        if (!b->have_camera) {
            error("no camera set");
            return false;
        }
        Ctx c;
        c.mult = 2;  // scene.world is itself a hittable_list
        for (rtw_id ch : b->nodes[b->world].ch) emit_object(ch, c);
        fs.n_world_items = (int32_t)fs.items.size();
        if (b->lights >= 0) emit_lights(b->lights);
        fs.render_type = b->render_type;
        fs.background_type = b->background_type;
        fs.camera = b->cam;
        if (uses_noise) {
            if (b->have_perlin) {
                fs.perlin_ranvec = b->ranvec;
                fs.perlin_perm = b->perm;
            } else {
                default_perlin_tables(fs.perlin_ranvec, fs.perlin_perm);
            }
        }
        fs.refresh_view();
        return ok;
    }
};

template <typename T>
void put(std::vector<uint8_t>& out, const T& v) {
    const uint8_t* p = reinterpret_cast<const uint8_t*>(&v);
    out.insert(out.end(), p, p + sizeof(T));
}

}  // namespace

extern "C" {

const rtw_flat_scene* rtw_builder_flatten(rtw_builder* b) {
    if (!b) {
        set_error("null builder");
        return nullptr;
    }
    Flattener f(b, b->flat);
    if (!f.run()) {
        set_error("flatten: " + f.err);
        return nullptr;
    }
    b->flat_material_index = f.mat_map;
    return &b->flat.view;
}

int32_t rtw_builder_flat_material_index(const rtw_builder* b, rtw_id material) {
    if (!b) return -1;
    auto it = b->flat_material_index.find(material);
    return it == b->flat_material_index.end() ? -1 : it->second;
}

size_t rtw_flat_serialize(const rtw_flat_scene* fs, void* buf, size_t cap) {
    if (!fs) return 0;
    std::vector<uint8_t> out;
    const char magic[8] = {'R', 'T', 'W', 'F', 'L', 'A', 'T', '1'};
    out.insert(out.end(), magic, magic + 8);
    int32_t hdr[12] = {fs->n_world_items, fs->n_items, fs->n_prims, fs->n_xfs, fs->n_nodes, fs->n_leaves,
                       fs->n_materials, fs->n_textures, fs->n_lights, fs->n_texel_bytes, fs->render_type,
                       fs->background_type};
    for (int32_t h : hdr) put(out, h);
    // field-by-field (no struct padding in the stream)
    for (int i = 0; i < fs->n_items; i++) {
        const rtw_flat_item& x = fs->items[i];
        put(out, x.type), put(out, x.xf), put(out, x.first), put(out, x.count), put(out, x.material), put(out, x.flip),
            put(out, x.multiplicity), put(out, x.density);
    }
    for (int i = 0; i < fs->n_prims; i++) {
        const rtw_flat_prim& x = fs->prims[i];
        put(out, x.type), put(out, x.material), put(out, x.flip), put(out, x.xf);
        for (double d : x.v) put(out, d);
    }
    for (int i = 0; i < fs->n_xfs; i++) {
        const rtw_flat_xf& x = fs->xfs[i];
        put(out, x.kind), put(out, x.next), put(out, x.a), put(out, x.b), put(out, x.c);
    }
    for (int i = 0; i < fs->n_nodes; i++) {
        const rtw_flat_node& x = fs->nodes[i];
        for (double d : x.bmin) put(out, d);
        for (double d : x.bmax) put(out, d);
        put(out, x.left), put(out, x.right);
    }
    for (int i = 0; i < fs->n_leaves; i++) put(out, fs->leaves[i].first_prim), put(out, fs->leaves[i].prim_count);
    for (int i = 0; i < fs->n_materials; i++) {
        const rtw_flat_material& x = fs->materials[i];
        put(out, x.type), put(out, x.texture);
        for (double d : x.albedo) put(out, d);
        put(out, x.fuzz), put(out, x.ref_idx);
    }
    for (int i = 0; i < fs->n_textures; i++) {
        const rtw_flat_texture& x = fs->textures[i];
        put(out, x.type), put(out, x.even), put(out, x.odd), put(out, x.image_offset), put(out, x.nx), put(out, x.ny);
        for (double d : x.color) put(out, d);
        put(out, x.scale);
    }
    for (int i = 0; i < fs->n_lights; i++) {
        put(out, fs->lights[i].type);
        for (double d : fs->lights[i].v) put(out, d);
    }
    if (fs->n_texel_bytes) out.insert(out.end(), fs->texels, fs->texels + fs->n_texel_bytes);
    int32_t has_perlin = fs->perlin_ranvec ? 1 : 0;
    put(out, has_perlin);
    if (has_perlin) {
        for (int i = 0; i < 768; i++) put(out, fs->perlin_ranvec[i]);
        for (int i = 0; i < 768; i++) put(out, fs->perlin_perm[i]);
    }
    const double* cam = reinterpret_cast<const double*>(&fs->camera);
    for (size_t i = 0; i < sizeof(rtw_flat_camera) / sizeof(double); i++) put(out, cam[i]);
    if (buf && cap) memcpy(buf, out.data(), std::min(cap, out.size()));
    return out.size();
}

}  // extern "C"
