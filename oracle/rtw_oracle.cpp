// rtw_oracle.cpp — TEST INFRASTRUCTURE: CPU restatement (FP64) of the reference's hot path, evaluated on the
// flattened scene (`rtw_flat_scene`, include/rtw_b200.h).  Only tests/, __graft_entry__.smoke() and the
// cpu_baseline leg of bench.py may load this; the product path never does.
//
// PARITY PIN: in `minstd` mode this file owns the same set of default-seeded std::minstd_rand engines as the
// reference (utility.h:16-17, material.h:209-210, hittable.h:432-433, camera.h:72-73,
// RayTracingWeekend.cpp:207-208) and draws from them in the reference's order, so a serial render and every
// per-function sequence is compared BIT FOR BIT (float64) with the reference's own object code run through
// oracle/_ref/ref_harness (tests/test_oracle_vs_reference.py).  hit() needs no RNG and is compared bit for bit
// on fixed ray batches for all seven scenes.
// In `philox` mode the same color() logic consumes the Philox4x32-10 uniforms the CUDA kernels consume
// (same counters, same 24-bit values), with the bounded samplers the device uses, which lets tests compare
// GPU and oracle path by path.
//
// Each function cites the reference file:line it restates (paths relative to /root/reference/RayTracingWeekend/).
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <random>
#include <thread>
#include <vector>

#include "rtw_b200.h"

namespace {

// ------------------------------------------------------------------------------------------------ vec3.h
struct V {
    double x, y, z;
};
inline V mk(double x, double y, double z) { return V{x, y, z}; }
inline V operator+(V a, V b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V operator-(V a, V b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V operator*(V a, V b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V operator/(V a, V b) { return mk(a.x / b.x, a.y / b.y, a.z / b.z); }
inline V operator*(double s, V a) { return mk(s * a.x, s * a.y, s * a.z); }  // vec3(double) * vec3 (vec3.h:15, :49)
inline V operator*(V a, double s) { return mk(a.x * s, a.y * s, a.z * s); }
inline V operator/(V a, double s) { return mk(a.x / s, a.y / s, a.z / s); }
inline V operator-(V a) { return mk(-a.x, -a.y, -a.z); }
inline double dot(V a, V b) { return a.x * b.x + a.y * b.y + a.z * b.z; }  // vec3.h:52
inline V cross(V a, V b) {                                                  // vec3.h:54-59
    return mk((a.y * b.z - a.z * b.y), (-(a.x * b.z - a.z * b.x)), (a.x * b.y - a.y * b.x));
}
inline double length_squared(V a) { return a.x * a.x + a.y * a.y + a.z * a.z; }  // vec3.h:31
inline double length(V a) { return std::sqrt(length_squared(a)); }
inline V normalize(V a) {  // vec3.h:61-67: divide by the length, component-wise
    double len = length(a);
    return mk(a.x / len, a.y / len, a.z / len);
}
inline double comp(V a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

struct Ray {  // ray.h
    V o, d;
    double time;
    V at(double t) const { return o + d * t; }  // ray.h:11
};

// ------------------------------------------------------------------------------------------------ RNG
uint32_t mulhi(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) >> 32); }
void philox4x32_10(const uint32_t c_in[4], const uint32_t k_in[2], uint32_t out[4]) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    uint32_t c[4] = {c_in[0], c_in[1], c_in[2], c_in[3]}, k[2] = {k_in[0], k_in[1]};
    for (int i = 0; i < 10; i++) {
        uint32_t hi0 = mulhi(M0, c[0]), lo0 = M0 * c[0], hi1 = mulhi(M1, c[2]), lo1 = M1 * c[2];
        uint32_t n[4] = {hi1 ^ c[1] ^ k[0], lo1, hi0 ^ c[3] ^ k[1], lo0};
        memcpy(c, n, sizeof n);
        k[0] += W0;
        k[1] += W1;
    }
    memcpy(out, c, 4 * sizeof(uint32_t));
}

enum { MODE_MINSTD = 0, MODE_PHILOX = 1 };
double g_jitter_eps = 0.0;   // oracle_set_jitter: relative size of the per-ray perturbation (0 = off)
uint64_t g_jitter_seed = 0;  // which perturbation
enum { SITE_MAIN = 0, SITE_BALL = 1, SITE_LENS = 1, SITE_MEDIUM0 = 2 };

struct Rng {
    int mode = MODE_MINSTD;
    // the reference's engines (all default-seeded)
    std::minstd_rand util, dielectric, medium, cam_ray, cam_time, image;
    std::uniform_real_distribution<double> uniform;
    // philox stream identity
    uint32_t pixel = 0, sample = 0, seed_lo = 0, seed_hi = 0;

    double random_double() { return 0.0 + (1.0 - 0.0) * uniform(util); }  // utility.h:14-20
    double random_double(double a, double b) { return a + (b - a) * uniform(util); }
    void philox(uint32_t bounce, uint32_t site, double u[4]) const {
        uint32_t c[4] = {pixel, sample, bounce, site}, k[2] = {seed_lo, seed_hi}, r[4];
        philox4x32_10(c, k, r);
        for (int i = 0; i < 4; i++) u[i] = (double)(r[i] >> 8) * 5.9604644775390625e-8;
    }
};

// ------------------------------------------------------------------------------------------------ hit_record
struct Rec {  // hittable.h:16-29 + ids
    double t = 0;
    V p{0, 0, 0}, normal{0, 0, 0};
    double u = 0, v = 0;
    int mat = -1;
    int prim = -1;  // flat primitive index; -2-item for a constant_medium
};


// ------------------------------------------------------------------------------------------------ event counts
// Algorithmic-work accounting for the roofline: every event of the reference's math is
// counted once per occurrence on a SINGLE list pass (the second pass of hittable_list::hit repeats work without
// changing the result) and with one self-hit test per light pdf evaluation folded into that pdf's cost.
enum Ev {
    EV_CAMERA, EV_LENS, EV_RECT_TEST, EV_RECT_ACCEPT, EV_SPHERE_TEST, EV_SPHERE_ROOT, EV_SPHERE_ACCEPT, EV_MOVING_CENTER,
    EV_AABB, EV_TRANSLATE, EV_ROTATE, EV_FLIP, EV_MEDIUM, EV_BACKGROUND, EV_EMITTED, EV_ONB, EV_COSINE_GEN, EV_COSINE_VALUE,
    EV_SCATTERING_PDF, EV_LIGHT_PICK, EV_RECT_RANDOM, EV_RECT_PDF, EV_SPHERE_RANDOM, EV_SPHERE_PDF, EV_MIXTURE, EV_LAMBERT_UPDATE,
    EV_SPECULAR_UPDATE, EV_METAL, EV_DIELECTRIC, EV_ISOTROPIC, EV_CHECKER, EV_IMAGE_TEX, EV_PERLIN_NOISE, EV_NOISE_TEX, EV_RAY,
    EV_COUNT
};
thread_local uint64_t* g_ev = nullptr;  // null = counting off (also while inside a light pdf's self-hit test)
inline void ev(int e, uint64_t n = 1) {
    if (g_ev) g_ev[e] += n;
}
struct EvSuspend {
    uint64_t* saved;
    EvSuspend() : saved(g_ev) { g_ev = nullptr; }
    ~EvSuspend() { g_ev = saved; }
};

// sphere_base::hit (sphere.h:46-81)
bool sphere_hit(const rtw_flat_prim& s, const Ray& r, double t_min, double t_max, Rec& rec) {
    V center = mk(s.v[0], s.v[1], s.v[2]);
    double radius = s.v[3];
    V cur = center;
    ev(EV_SPHERE_TEST);
    if (s.type == RTW_PRIM_MOVING_SPHERE) {  // sphere.h:22-25
        ev(EV_MOVING_CENTER);
        V c1 = mk(s.v[4], s.v[5], s.v[6]);
        cur = center + ((r.time - s.v[7]) / (s.v[8] - s.v[7])) * (c1 - center);
    }
    V oc = r.o - cur;
    double a = dot(r.d, r.d);
    double b = dot(oc, r.d);
    double c = dot(oc, oc) - radius * radius;
    double discriminant = b * b - a * c;
    if (discriminant > 0) {
        double temp = (-b - std::sqrt(discriminant)) / a;
        for (int root = 0; root < 2; root++) {
            ev(EV_SPHERE_ROOT);
            if (temp < t_max && temp > t_min) {
                ev(EV_SPHERE_ACCEPT);
                rec.t = temp;
                rec.p = r.at(rec.t);
                rec.normal = (rec.p - cur) / radius;
                double phi = atan2(rec.normal.z, rec.normal.x);  // sphere.h:115-122
                double theta = asin(rec.normal.y);
                rec.u = double(1 - (phi + M_PI) / (2 * M_PI));
                rec.v = double((theta + M_PI / 2) / M_PI);
                rec.mat = s.material;
                return true;
            }
            temp = (-b + std::sqrt(discriminant)) / a;
        }
    }
    return false;
}

// xy_rect / xz_rect / yz_rect ::hit (hittable.h:149-165, 184-200, 241-257)
bool rect_hit(const rtw_flat_prim& q, const Ray& r, double t0, double t1, Rec& rec) {
    int n, a, b;  // plane axis, first and second in-plane axes
    if (q.type == RTW_PRIM_XY_RECT) n = 2, a = 0, b = 1;
    else if (q.type == RTW_PRIM_XZ_RECT) n = 1, a = 0, b = 2;
    else n = 0, a = 1, b = 2;
    double k = q.v[4];
    ev(EV_RECT_TEST);
    double t = (k - comp(r.o, n)) / comp(r.d, n);
    if (t < t0 || t > t1) return false;
    double ca = comp(r.o, a) + t * comp(r.d, a);
    double cb = comp(r.o, b) + t * comp(r.d, b);
    if (ca < q.v[0] || ca > q.v[1] || cb < q.v[2] || cb > q.v[3]) return false;
    ev(EV_RECT_ACCEPT);
    rec.u = (ca - q.v[0]) / (q.v[1] - q.v[0]);
    rec.v = (cb - q.v[2]) / (q.v[3] - q.v[2]);
    rec.t = t;
    rec.mat = q.material;
    rec.p = r.at(t);
    rec.normal = mk(n == 0 ? 1 : 0, n == 1 ? 1 : 0, n == 2 ? 1 : 0);
    return true;
}

// aabb::hit (aabb.h:17-47)
bool aabb_hit(const double* mn, const double* mx, const Ray& r, double tmin, double tmax) {
    ev(EV_AABB);
    for (int axis = 0; axis < 3; axis++) {
        double invD = 1.0 / comp(r.d, axis);
        double t0 = (mn[axis] - comp(r.o, axis)) * invD;
        double t1 = (mx[axis] - comp(r.o, axis)) * invD;
        if (invD < 0.0) std::swap(t0, t1);
        tmin = std::max(t0, tmin);
        tmax = std::min(t1, tmax);
        if (tmax <= tmin) return false;
    }
    return true;
}

// translate / rotate_y chain (hittable.h:299-311, 373-404) around one primitive test; flip_normals (:273-284)
// is applied innermost-first like the wrappers unwind.
bool prim_hit(const rtw_flat_scene* fs, int pi, int chain, const Ray& r, double t_min, double t_max, Rec& rec) {
    if (chain >= 0) {
        const rtw_flat_xf& x = fs->xfs[chain];
        ev(x.kind == RTW_XF_TRANSLATE ? EV_TRANSLATE : EV_ROTATE);
        if (x.kind == RTW_XF_TRANSLATE) {
            V off = mk(x.a, x.b, x.c);
            Ray moved{r.o - off, r.d, r.time};
            if (prim_hit(fs, pi, x.next, moved, t_min, t_max, rec)) {
                rec.p = rec.p + off;
                return true;
            }
            return false;
        }
        double sin_theta = x.a, cos_theta = x.b;
        V origin = r.o, direction = r.d;
        origin.x = cos_theta * r.o.x - sin_theta * r.o.z;
        origin.z = sin_theta * r.o.x + cos_theta * r.o.z;
        direction.x = cos_theta * r.d.x - sin_theta * r.d.z;
        direction.z = sin_theta * r.d.x + cos_theta * r.d.z;
        Ray rotated{origin, direction, r.time};
        if (prim_hit(fs, pi, x.next, rotated, t_min, t_max, rec)) {
            V p = rec.p, normal = rec.normal;
            p.x = cos_theta * rec.p.x + sin_theta * rec.p.z;
            p.z = -sin_theta * rec.p.x + cos_theta * rec.p.z;
            normal.x = cos_theta * rec.normal.x + sin_theta * rec.normal.z;
            normal.z = -sin_theta * rec.normal.x + cos_theta * rec.normal.z;
            rec.p = p;
            rec.normal = normal;
            return true;
        }
        return false;
    }
    const rtw_flat_prim& q = fs->prims[pi];
    bool h = q.type <= RTW_PRIM_MOVING_SPHERE ? sphere_hit(q, r, t_min, t_max, rec) : rect_hit(q, r, t_min, t_max, rec);
    if (h) {
        rec.prim = pi;
        if (q.flip) {
            ev(EV_FLIP);
            rec.normal = -rec.normal;
        }
    }
    return h;
}

}  // namespace

namespace {

// ------------------------------------------------------------------------------------------------ world.hit
struct World {
    const rtw_flat_scene* fs;
    Rng* rng;          // may be null: constant_medium items are then skipped (they are stochastic)
    uint32_t bounce;   // philox counter word for medium draws
};

// hittable_list::hit over a run of primitives (hittable_list.h:11-37), both passes.
bool prims_list_hit(const rtw_flat_scene* fs, int first, int count, const Ray& r, double t_min, double t_max, Rec& rec) {
    Rec temp;
    bool hit_anything = false;
    double closest = t_max;
    uint64_t* saved = g_ev;
    for (int pass = 0; pass < 2; pass++) {
        if (pass == 1) g_ev = nullptr;  // the second pass repeats the first (event accounting: single pass)
        for (int i = first; i < first + count; i++)
            if (prim_hit(fs, i, fs->prims[i].xf, r, t_min, closest, temp)) {
                hit_anything = true;
                closest = temp.t;
                rec = temp;
            }
    }
    g_ev = saved;
    return hit_anything;
}

// bvh_node::hit (hittable.h:97-129) as intended: box test, both children with the caller's
// interval, smaller t wins, right wins ties.
bool bvh_hit(const rtw_flat_scene* fs, int ref, const Ray& r, double tmin, double tmax, Rec& rec) {
    if (ref < 0) {
        const rtw_flat_leaf& lf = fs->leaves[~ref];
        if (lf.prim_count == 1) return prim_hit(fs, lf.first_prim, fs->prims[lf.first_prim].xf, r, tmin, tmax, rec);
        return prims_list_hit(fs, lf.first_prim, lf.prim_count, r, tmin, tmax, rec);  // box::hit (hittable_list.h:100-104)
    }
    const rtw_flat_node& n = fs->nodes[ref];
    if (!aabb_hit(n.bmin, n.bmax, r, tmin, tmax)) return false;
    Rec left_rec, right_rec;
    bool hit_left = bvh_hit(fs, n.left, r, tmin, tmax, left_rec);
    bool hit_right = bvh_hit(fs, n.right, r, tmin, tmax, right_rec);
    if (hit_left && hit_right) {
        rec = left_rec.t < right_rec.t ? left_rec : right_rec;
        return true;
    }
    if (hit_left) {
        rec = left_rec;
        return true;
    }
    if (hit_right) {
        rec = right_rec;
        return true;
    }
    return false;
}

// Apply an item-level translate/rotate_y chain around `inner` (same arithmetic as prim_hit's wrappers).
template <typename Inner>
bool with_chain(const rtw_flat_scene* fs, int chain, const Ray& r, Rec& rec, Inner inner) {
    if (chain < 0) return inner(r, rec);
    const rtw_flat_xf& x = fs->xfs[chain];
    ev(x.kind == RTW_XF_TRANSLATE ? EV_TRANSLATE : EV_ROTATE);
    if (x.kind == RTW_XF_TRANSLATE) {
        V off = mk(x.a, x.b, x.c);
        Ray moved{r.o - off, r.d, r.time};
        if (with_chain(fs, x.next, moved, rec, inner)) {
            rec.p = rec.p + off;
            return true;
        }
        return false;
    }
    double s = x.a, c = x.b;
    V origin = r.o, direction = r.d;
    origin.x = c * r.o.x - s * r.o.z;
    origin.z = s * r.o.x + c * r.o.z;
    direction.x = c * r.d.x - s * r.d.z;
    direction.z = s * r.d.x + c * r.d.z;
    Ray rotated{origin, direction, r.time};
    if (with_chain(fs, x.next, rotated, rec, inner)) {
        V p = rec.p, normal = rec.normal;
        p.x = c * rec.p.x + s * rec.p.z;
        p.z = -s * rec.p.x + c * rec.p.z;
        normal.x = c * rec.normal.x + s * rec.normal.z;
        normal.z = -s * rec.normal.x + c * rec.normal.z;
        rec.p = p;
        rec.normal = normal;
        return true;
    }
    return false;
}

// constant_medium::hit (hittable.h:430-479).  `medium_index` numbers the media in list order (philox site).
bool medium_hit(const World& w, int item_index, int medium_index, const Ray& r0, double t_min, double t_max, Rec& out) {
    const rtw_flat_scene* fs = w.fs;
    const rtw_flat_item& it = fs->items[item_index];
    return with_chain(fs, it.xf, r0, out, [&](const Ray& r, Rec& rec) {
        auto boundary = [&](double a, double b, Rec& q) {
            if (it.count == 1) return prim_hit(fs, it.first, fs->prims[it.first].xf, r, a, b, q);
            return prims_list_hit(fs, it.first, it.count, r, a, b, q);
        };
        Rec rec1, rec2;
        if (boundary(-std::numeric_limits<double>::max(), std::numeric_limits<double>::max(), rec1)) {
            if (boundary(rec1.t + 0.0001f, std::numeric_limits<double>::max(), rec2)) {
                if (rec1.t < t_min) rec1.t = t_min;
                if (rec2.t > t_max) rec2.t = t_max;
                if (rec1.t >= rec2.t) return false;
                if (rec1.t < 0) rec1.t = 0;
                ev(EV_MEDIUM);
                double distance_inside_boundary = (rec2.t - rec1.t) * length(r.d);
                double hit_distance;
                if (w.rng->mode == MODE_MINSTD) {
                    hit_distance = -(1 / it.density) * log(w.rng->uniform(w.rng->medium));
                } else {  // device form: one draw at density*multiplicity (min of iid exponentials)
                    double u[4];
                    w.rng->philox(w.bounce, SITE_MEDIUM0 + (uint32_t)medium_index, u);
                    hit_distance = -(1 / (it.density * it.multiplicity)) * log(u[0]);
                }
                if (hit_distance < distance_inside_boundary) {
                    rec.t = rec1.t + hit_distance / length(r.d);
                    rec.p = r.at(rec.t);
                    rec.normal = it.flip ? mk(-1, 0, 0) : mk(1, 0, 0);
                    rec.mat = it.material;
                    rec.prim = -2 - item_index;
                    rec.u = rec.v = 0;  // never written by the reference (Q13); defined as 0 here and on device
                    return true;
                }
            }
        }
        return false;
    });
}

// scene.world.hit (hittable_list.h:11-37) over the flattened items.
bool world_hit(const World& w, const Ray& r, double t_min, double t_max, Rec& rec) {
    const rtw_flat_scene* fs = w.fs;
    Rec temp;
    bool hit_anything = false;
    double closest = t_max;
    int passes = 2;
    uint64_t* saved_ev = g_ev;
    for (int pass = 0; pass < passes; pass++) {
        if (pass == 1) g_ev = nullptr;
        int medium_index = 0;
        int prev_xf = -1;
        for (int ii = 0; ii < fs->n_world_items; ii++) {
            const rtw_flat_item& it = fs->items[ii];
            if (it.type == RTW_ITEM_PRIMS) {
                for (int i = it.first; i < it.first + it.count; i++) {
                    // a wrapped object (e.g. translate(rotate_y(box))) transforms the ray once for all its rects
                    int xf = fs->prims[i].xf;
                    if (xf != prev_xf && xf >= 0)
                        for (int l = xf; l >= 0; l = fs->xfs[l].next) ev(fs->xfs[l].kind == RTW_XF_TRANSLATE ? EV_TRANSLATE : EV_ROTATE);
                    prev_xf = xf;
                    uint64_t* keep = g_ev;
                    uint64_t tr0 = keep ? keep[EV_TRANSLATE] : 0, ro0 = keep ? keep[EV_ROTATE] : 0;
                    bool h = prim_hit(fs, i, xf, r, t_min, closest, temp);
                    if (keep) keep[EV_TRANSLATE] = tr0, keep[EV_ROTATE] = ro0;  // per-primitive chain walks are not reference work
                    if (h) {
                        hit_anything = true;
                        closest = temp.t;
                        rec = temp;
                    }
                }
            } else if (it.type == RTW_ITEM_BVH) {
                bool h = with_chain(fs, it.xf, r, temp, [&](const Ray& rl, Rec& q) { return bvh_hit(fs, it.first, rl, t_min, closest, q); });
                if (h) {
                    hit_anything = true;
                    closest = temp.t;
                    rec = temp;
                }
            } else {
                int mi = medium_index++;
                if (!w.rng) continue;
                if (w.rng->mode == MODE_PHILOX && pass == 1) continue;  // device: single draw at density*multiplicity
                int calls = w.rng->mode == MODE_MINSTD ? std::max(1, it.multiplicity / 2) : 1;
                for (int c = 0; c < calls; c++)
                    if (medium_hit(w, ii, mi, r, t_min, closest, temp)) {
                        hit_anything = true;
                        closest = temp.t;
                        rec = temp;
                    }
            }
        }
    }
    g_ev = saved_ev;
    if (hit_anything && rec.prim >= 0 && fs->prims[rec.prim].xf >= 0)  // p / normal transformed back once for the winner
        for (int l = fs->prims[rec.prim].xf; l >= 0; l = fs->xfs[l].next) ev(fs->xfs[l].kind == RTW_XF_TRANSLATE ? EV_TRANSLATE : EV_ROTATE);
    return hit_anything;
}

}  // namespace

namespace {

// ------------------------------------------------------------------------------------------------ textures
// perlin::noise / perlin_interp / turb (noise.h:40-60, 74-86, 89-151), Hermite-cubic smoothing (noise.h:9-12,19)
double perlin_noise(const rtw_flat_scene* fs, V p) {
    ev(EV_PERLIN_NOISE);
    double u = p.x - floor(p.x), v = p.y - floor(p.y), w = p.z - floor(p.z);
    int i = (int)floor(p.x), j = (int)floor(p.y), k = (int)floor(p.z);
    const int32_t *px = fs->perlin_perm, *py = fs->perlin_perm + 256, *pz = fs->perlin_perm + 512;
    V c[2][2][2];
    for (int di = 0; di < 2; di++)
        for (int dj = 0; dj < 2; dj++)
            for (int dk = 0; dk < 2; dk++) {
                int idx = px[(i + di) & 255] ^ py[(j + dj) & 255] ^ pz[(k + dk) & 255];
                c[di][dj][dk] = mk(fs->perlin_ranvec[3 * idx], fs->perlin_ranvec[3 * idx + 1], fs->perlin_ranvec[3 * idx + 2]);
            }
    double uu = u * u * (3 - 2 * u), vv = v * v * (3 - 2 * v), ww = w * w * (3 - 2 * w);
    double accum = 0;
    for (int a = 0; a < 2; a++)
        for (int b = 0; b < 2; b++)
            for (int cc = 0; cc < 2; cc++) {
                V weight_v = mk(u - a, v - b, w - cc);
                accum += (a * uu + (1 - a) * (1 - uu)) * (b * vv + (1 - b) * (1 - vv)) * (cc * ww + (1 - cc) * (1 - ww)) *
                         dot(c[a][b][cc], weight_v);
            }
    return accum;
}
double perlin_turb(const rtw_flat_scene* fs, V p) {
    double accum = 0, weight = 1.0;
    V temp_p = p;
    for (int i = 0; i < 7; i++) {
        accum += weight * perlin_noise(fs, temp_p);
        weight *= 0.5f;
        temp_p = temp_p * 2.0;
    }
    return fabs(accum);
}

// texture::value (texture.h:21-24, 35-46, 57-68, 81-94)
V texture_value(const rtw_flat_scene* fs, int ti, double u, double v, V p) {
    const rtw_flat_texture& t = fs->textures[ti];
    switch (t.type) {
        case RTW_TEX_CONSTANT: return mk(t.color[0], t.color[1], t.color[2]);
        case RTW_TEX_CHECKER: {
            ev(EV_CHECKER);
            double sines = std::sin(10.0 * p.x) * std::sin(10.0 * p.y) * std::sin(10.0 * p.z);
            return texture_value(fs, sines < 0 ? t.odd : t.even, u, v, p);
        }
        case RTW_TEX_NOISE:
            ev(EV_NOISE_TEX);
            return mk(1, 1, 1) * 0.5f * (1 + sin(t.scale * p.z + 10 * perlin_turb(fs, p)));
        default: {
            ev(EV_IMAGE_TEX);
            int i = (int)(u * t.nx);
            int j = (int)((1 - v) * t.ny - 0.001f);
            i = std::max(std::min(i, t.nx - 1), 0);
            j = std::max(std::min(j, t.ny - 1), 0);
            const uint8_t* d = fs->texels + t.image_offset;
            double r = int(d[3 * i + 3 * t.nx * j]) / 255.0;
            double g = int(d[3 * i + 3 * t.nx * j + 1]) / 255.0;
            double b = int(d[3 * i + 3 * t.nx * j + 2]) / 255.0;
            return mk(r, g, b);
        }
    }
}

// ------------------------------------------------------------------------------------------------ sampling
struct Onb {  // onb.h
    V u, v, w;
    V local(V a) const { return a.x * u + a.y * v + a.z * w; }  // onb.h:21-24
};
Onb onb_from_w(V n) {  // onb.h:32-38
    ev(EV_ONB);
    Onb b;
    b.w = normalize(n);
    V a = (fabs(b.w.x) > 0.9) ? mk(0, 1, 0) : mk(1, 0, 0);
    b.v = normalize(cross(b.w, a));
    b.u = cross(b.w, b.v);
    return b;
}

// random_in_unit_sphere (utility.h:27-35).  minstd: the rejection loop, constructor arguments drawn right to
// left as g++ evaluates them; philox: direction * cbrt(U) from one draw (same distribution, no loop).
// philox mode: uniform direction * cbrt(U) — the same distribution as the rejection loop (uniform in the unit ball)
V ball_from_uniforms(double u0, double u1, double u2) {
    double z = 1.0 - 2.0 * u0, r = std::sqrt(std::max(0.0, 1.0 - z * z)), phi = 2 * M_PI * u1, rad = cbrt(u2);
    return mk(rad * r * cos(phi), rad * r * sin(phi), rad * z);
}
// philox mode of random_in_unit_disk (camera.h:61-69): polar (sqrt(U), 2 pi V) — uniform in the unit disk
V disk_from_uniforms(double q0, double q1) {
    double rad = sqrt(q0);
    return mk(rad * cos(2 * M_PI * q1), rad * sin(2 * M_PI * q1), 0);
}
V random_in_unit_sphere(Rng& g, uint32_t bounce) {
    if (g.mode == MODE_PHILOX) {
        double u[4];
        g.philox(bounce, SITE_BALL, u);
        return ball_from_uniforms(u[0], u[1], u[2]);
    }
    V p;
    do {
        double c = g.random_double(), b = g.random_double(), a = g.random_double();
        p = 2.0 * mk(a, b, c) - mk(1, 1, 1);
    } while (dot(p, p) >= 1.0);
    return p;
}
V random_cosine_direction(double r1, double r2) {  // utility.h:54-67
    double z = sqrt(1 - r2);
    double phi = 2 * M_PI * r1;
    double x = cos(phi) * sqrt(r2);
    double y = sin(phi) * sqrt(r2);
    return mk(x, y, z);
}
V random_to_sphere(double radius, double distance_squared, double r1, double r2) {  // utility.h:69-82
    double z = 1 + r2 * (sqrt(1 - radius * radius / distance_squared) - 1);
    double phi = 2 * M_PI * r1;
    double x = cos(phi) * sqrt(1 - z * z);
    double y = sin(phi) * sqrt(1 - z * z);
    return mk(x, y, z);
}

// ------------------------------------------------------------------------------------------------ light pdfs
rtw_flat_prim light_as_prim(const rtw_flat_light& l) {
    rtw_flat_prim q{};
    q.type = l.type == RTW_LIGHT_XZ_RECT ? RTW_PRIM_XZ_RECT : RTW_PRIM_SPHERE;
    q.material = -1, q.xf = -1;
    for (int i = 0; i < 5; i++) q.v[i] = l.v[i];
    return q;
}
double light_pdf_value(const rtw_flat_light& l, V o, V v) {
    ev(l.type == RTW_LIGHT_XZ_RECT ? EV_RECT_PDF : (l.type == RTW_LIGHT_SPHERE ? EV_SPHERE_PDF : EV_COUNT - 1), l.type ? 1 : 0);
    EvSuspend quiet;  // the pdf's own hit test is part of its cost (Appendix D)
    Rec rec;
    Ray r{o, v, (double)FLT_MAX};
    if (l.type == RTW_LIGHT_XZ_RECT) {  // hittable.h:208-222
        if (!rect_hit(light_as_prim(l), r, 0.001, std::numeric_limits<double>::infinity(), rec)) return 0;
        double area = (l.v[1] - l.v[0]) * (l.v[3] - l.v[2]);
        double distance_squared = rec.t * rec.t * length_squared(v);
        double cosine = fabs(dot(v, rec.normal) / length(v));
        return distance_squared / (cosine * area);
    }
    if (l.type == RTW_LIGHT_SPHERE) {  // sphere.h:88-99
        if (!sphere_hit(light_as_prim(l), r, 0.001, std::numeric_limits<double>::infinity(), rec)) return 0.0;
        V center = mk(l.v[0], l.v[1], l.v[2]);
        double radius = l.v[3];
        double cos_theta_max = sqrt(1 - radius * radius / length_squared(center - o));
        double solid_angle = 2.0 * M_PI * (1.0 - cos_theta_max);
        return 1.0 / solid_angle;
    }
    return 0.0;  // hittable.h:36
}
double lights_pdf_value(const rtw_flat_scene* fs, V o, V v) {  // hittable_list.h:44-53
    double weight = 1.0 / fs->n_lights;
    double sum = 0.0;
    for (int i = 0; i < fs->n_lights; i++) sum += weight * light_pdf_value(fs->lights[i], o, v);
    return sum;
}
V light_random(const rtw_flat_light& l, V o, double r1, double r2) {
    ev(EV_LIGHT_PICK);
    if (l.type == RTW_LIGHT_XZ_RECT) ev(EV_RECT_RANDOM);
    if (l.type == RTW_LIGHT_SPHERE) ev(EV_SPHERE_RANDOM);
    EvSuspend quiet;  // ONB build is inside the sphere-random cost
    if (l.type == RTW_LIGHT_XZ_RECT) {  // hittable.h:224-228 (r1 -> x, r2 -> z)
        V random = mk(l.v[0] + (l.v[1] - l.v[0]) * r1, l.v[4], l.v[2] + (l.v[3] - l.v[2]) * r2);
        return random - o;
    }
    if (l.type == RTW_LIGHT_SPHERE) {  // sphere.h:101-108
        V direction = mk(l.v[0], l.v[1], l.v[2]) - o;
        double distance_squared = length_squared(direction);
        Onb uvw = onb_from_w(direction);
        return uvw.local(random_to_sphere(l.v[3], distance_squared, r1, r2));
    }
    return mk(1, 0, 0);  // hittable.h:37
}

}  // namespace

namespace {

// ------------------------------------------------------------------------------------------------ color()
struct Tracer {
    const rtw_flat_scene* fs;
    Rng rng;
    uint64_t rays = 0;

    V reflect(V v, V n) { return v - 2.0 * dot(v, n) * n; }  // material.h:10-13
    bool refract(V v, V n, double ni_over_nt, V& refracted) {  // material.h:17-39
        V uv = normalize(v);
        double dt = dot(uv, n);
        double discriminant = 1.0 - ni_over_nt * ni_over_nt * (1 - dt * dt);
        if (discriminant > 0) {
            refracted = ni_over_nt * (uv - n * dt) - n * sqrt(discriminant);
            return true;
        }
        return false;
    }
    double schlick(double cosine, double ref_idx) {  // material.h:44-49
        double r0 = (1 - ref_idx) / (1 + ref_idx);
        r0 = r0 * r0;
        return r0 + (1 - r0) * pow((1 - cosine), 5);
    }

    // dielectric::scatter (material.h:146-222); returns the scattered direction
    V dielectric_scatter(const rtw_flat_material& m, const Ray& r_in, const Rec& rec, uint32_t bounce) {
        double ref_idx = m.ref_idx;
        V outward_normal;
        double ni_over_nt, cosine;
        if (dot(r_in.d, rec.normal) > 0) {
            outward_normal = -rec.normal;
            ni_over_nt = ref_idx;
            cosine = dot(r_in.d, rec.normal) / length(r_in.d);
            cosine = sqrt(1 - ref_idx * ref_idx * (1 - cosine * cosine));
        } else {
            outward_normal = rec.normal;
            ni_over_nt = 1.0 / ref_idx;
            cosine = -dot(r_in.d, rec.normal) / length(r_in.d);
        }
        V reflected = reflect(r_in.d, rec.normal);
        V refracted = mk(0, 0, 0);
        double reflect_prob;
        if (refract(r_in.d, outward_normal, ni_over_nt, refracted)) reflect_prob = schlick(cosine, ref_idx);
        else reflect_prob = 1.0;
        double rand;
        if (rng.mode == MODE_MINSTD) rand = rng.uniform(rng.dielectric);  // material.h:209-211
        else {
            double u[4];
            rng.philox(bounce, SITE_MAIN, u);
            rand = u[0];
        }
        return rand < reflect_prob ? reflected : refracted;
    }

    V background(const Ray& r) {  // RayTracingWeekend.cpp:143-158
        if (fs->background_type == RTW_BG_GRADIENT) {
            ev(EV_BACKGROUND);
            V unit_direction = normalize(r.d);
            double t = 0.5f * (unit_direction.y + 1.0);
            V from = mk(0.5f, 0.7f, 1.0), to = mk(1.0, 1.0, 1.0);
            return (1.0 - t) * to + t * from;  // lerp, vec3.h:84-87
        }
        return mk(0, 0, 0);
    }

    // vec3 color(const ray&, const scene*, int depth)  (RayTracingWeekend.cpp:45-160); `bounce` = 1-based call index
    V color(const Ray& r_in, int depth, uint32_t bounce) {
        if (depth <= 0) return mk(0, 0, 0);
        Ray r = r_in;
        if (g_jitter_eps > 0 && rng.mode == MODE_PHILOX) {
            // sensitivity probe (oracle_set_jitter): every ray entering color() is perturbed component-wise by a relative
            // amount of FP32-rounding size, from Philox draws no sampler uses — a path whose result changes under this is one
            // whose branch decisions FP32 arithmetic can legitimately flip
            double a[4], b[4];
            rng.philox(bounce, 0x7000u + 2u * (uint32_t)g_jitter_seed, a);
            rng.philox(bounce, 0x7001u + 2u * (uint32_t)g_jitter_seed, b);
            const double e = g_jitter_eps;
            r.o = mk(r.o.x * (1 + e * (2 * a[0] - 1)), r.o.y * (1 + e * (2 * a[1] - 1)), r.o.z * (1 + e * (2 * a[2] - 1)));
            r.d = mk(r.d.x * (1 + e * (2 * b[0] - 1)), r.d.y * (1 + e * (2 * b[1] - 1)), r.d.z * (1 + e * (2 * b[2] - 1)));
        }
        rays++;
        ev(EV_RAY);
        Rec rec;
        World w{fs, &rng, bounce};
        if (!world_hit(w, r, 0.001f, std::numeric_limits<double>::max(), rec)) return background(r);
        if (fs->render_type == RTW_RENDER_NORMAL) return 0.5f * (rec.normal + mk(1, 1, 1));  // :135-136
        const rtw_flat_material& m = fs->materials[rec.mat];
        switch (m.type) {
            case RTW_MAT_DIFFUSE_LIGHT:  // material.h:232-244: emitted one-sided, no scatter (:62-63)
                ev(EV_EMITTED);
                if (dot(rec.normal, r.d) > 0) return texture_value(fs, m.texture, rec.u, rec.v, rec.p);
                return mk(0, 0, 0);
            case RTW_MAT_METAL: {  // material.h:128-136, then :114-115
                ev(EV_METAL), ev(EV_SPECULAR_UPDATE);
                V reflected = reflect(normalize(r.d), rec.normal);
                V dir;
                if (rng.mode == MODE_PHILOX && m.fuzz == 0.0) dir = reflected;  // device skips the unused draw (Q24)
                else dir = reflected + m.fuzz * random_in_unit_sphere(rng, bounce);
                V att = mk(m.albedo[0], m.albedo[1], m.albedo[2]);
                return att * color(Ray{rec.p, dir, r.time}, depth - 1, bounce + 1);
            }
            case RTW_MAT_DIELECTRIC: {
                ev(EV_DIELECTRIC), ev(EV_SPECULAR_UPDATE);
                V dir = dielectric_scatter(m, r, rec, bounce);
                return mk(1.0, 1.0, 1.0) * color(Ray{rec.p, dir, r.time}, depth - 1, bounce + 1);
            }
            case RTW_MAT_ISOTROPIC: {  // material.h:257-262
                ev(EV_ISOTROPIC), ev(EV_SPECULAR_UPDATE);
                V dir = random_in_unit_sphere(rng, bounce);
                V att = texture_value(fs, m.texture, rec.u, rec.v, rec.p);
                return att * color(Ray{rec.p, dir, r.time}, depth - 1, bounce + 1);
            }
            default: break;
        }
        // lambertian (material.h:81-119) + mixture pdf (:117-132)
        V emitted = mk(0, 0, 0);
        V attenuation = texture_value(fs, m.texture, rec.u, rec.v, rec.p);
        Onb uvw = onb_from_w(rec.normal);  // cosine_pdf ctor, pdf.h:18
        bool have_lights = fs->n_lights > 0;
        double pu[4] = {0, 0, 0, 0};
        if (rng.mode == MODE_PHILOX) rng.philox(bounce, SITE_MAIN, pu);
        V dir;
        bool pick_cosine = true;
        if (have_lights) pick_cosine = (rng.mode == MODE_MINSTD ? rng.random_double() : pu[0]) < 0.5;  // pdf.h:71
        if (pick_cosine) {
            double r1, r2;
            if (rng.mode == MODE_MINSTD) r1 = rng.random_double(), r2 = rng.random_double();
            else r1 = pu[2], r2 = pu[3];
            ev(EV_COSINE_GEN);
            dir = uvw.local(random_cosine_direction(r1, r2));  // pdf.h:26-29
        } else {
            int n = fs->n_lights;
            double up = rng.mode == MODE_MINSTD ? rng.random_double() : pu[1];
            int idx = 0 + std::min(n - 1 - 0, (int)((n - 1 - 0 + 1) * up));  // utility.h:22-25
            const rtw_flat_light& l = fs->lights[idx];
            double r1 = 0, r2 = 0;
            if (rng.mode == MODE_PHILOX) r1 = pu[2], r2 = pu[3];
            else if (l.type == RTW_LIGHT_XZ_RECT) {
                // hittable.h:226: vec3(random_double(x0,x1), k, random_double(z0,z1)) — g++ draws z first
                r2 = rng.uniform(rng.util);
                r1 = rng.uniform(rng.util);
            } else if (l.type == RTW_LIGHT_SPHERE) {
                r1 = rng.random_double(), r2 = rng.random_double();  // utility.h:73-74
            }
            dir = light_random(l, rec.p, r1, r2);
        }
        Ray scattered{rec.p, dir, r.time};
        ev(EV_COSINE_VALUE), ev(EV_SCATTERING_PDF), ev(EV_LAMBERT_UPDATE);
        if (have_lights) ev(EV_MIXTURE);
        double cosine = dot(normalize(dir), uvw.w);  // cosine_pdf::value (pdf.h:20-24)
        double pdf_cos = (cosine <= 0) ? 0 : cosine / M_PI;
        double pdf_val = pdf_cos;
        if (have_lights) pdf_val = 0.5 * pdf_cos + 0.5 * lights_pdf_value(fs, rec.p, dir);  // pdf.h:64-67
        if (pdf_val <= 0.0) return emitted;  // :126-127
        double c2 = dot(rec.normal, normalize(scattered.d));  // material.h:115-119
        double spdf = c2 < 0 ? 0 : c2 / M_PI;
        return emitted + attenuation * spdf * color(scattered, depth - 1, bounce + 1) / pdf_val;  // :129-132
    }

    // camera::get_ray (camera.h:36-50)
    Ray get_ray(double s, double t) {
        const rtw_flat_camera& c = fs->camera;
        V origin = mk(c.origin[0], c.origin[1], c.origin[2]), llc = mk(c.lower_left_corner[0], c.lower_left_corner[1], c.lower_left_corner[2]);
        V horizontal = mk(c.horizontal[0], c.horizontal[1], c.horizontal[2]), vertical = mk(c.vertical[0], c.vertical[1], c.vertical[2]);
        V cu = mk(c.u[0], c.u[1], c.u[2]), cv = mk(c.v[0], c.v[1], c.v[2]);
        V rd;
        double time;
        ev(EV_CAMERA);
        if (c.lens_radius > 0) ev(EV_LENS);
        if (rng.mode == MODE_MINSTD) {
            V p;
            do {  // camera.h:61-69; the two constructor-argument draws happen right to left (y first)
                double y = rng.uniform(rng.cam_ray), x = rng.uniform(rng.cam_ray);
                p = 2.0 * mk(x, y, 0) - mk(1, 1, 0);
            } while (dot(p, p) >= 1.0);
            rd = c.lens_radius * p;
            time = c.time0 + rng.uniform(rng.cam_time) * (c.time1 - c.time0);
        } else {
            double u[4], q[4] = {0, 0, 0, 0};
            rng.philox(0, SITE_MAIN, u);
            rd = mk(0, 0, 0);
            if (c.lens_radius > 0) {
                rng.philox(0, SITE_LENS, q);
                rd = c.lens_radius * disk_from_uniforms(q[0], q[1]);
            }
            time = c.time0 + u[2] * (c.time1 - c.time0);
        }
        V offset = cu * rd.x + cv * rd.y;
        V dir = llc + s * horizontal + t * vertical - origin - offset;
        return Ray{origin + offset, normalize(dir), time};
    }
};

}  // namespace

// ------------------------------------------------------------------------------------------------ C API (ctypes)
namespace {
bool g_events_on = false;
uint64_t g_event_totals[EV_COUNT] = {0};
// Appendix D cost table C_e (FMA = 2; add/mul/div/sqrt/transcendental = 1; compares, selects, integer = 0)
const double kEventCost[EV_COUNT] = {
    34, 25, 6, 12, 23, 4, 19, 12, 21, 3, 12, 3, 20, 21, 5, 36, 25, 15, 15, 3, 9, 43, 76, 60, 3, 12, 3, 48, 75, 21, 8, 7, 140,
    7 * 6 + 1 + 5 /* noise_texture::value on top of its 7 perlin::noise calls */, 0};
}  // namespace

extern "C" {

// Event accounting for the roofline: enable, render, then read counts and the flop total.
void oracle_events_enable(int on) {
    g_events_on = on != 0;
    for (auto& t : g_event_totals) t = 0;
}
// out_counts[EV_COUNT] (may be null); returns sum_e N_e * C_e (algorithmic FP32 flops of the reference's math).
double oracle_events_read(uint64_t* out_counts, int cap) {
    double flops = 0;
    for (int k = 0; k < EV_COUNT; k++) {
        if (out_counts && k < cap) out_counts[k] = g_event_totals[k];
        flops += (double)g_event_totals[k] * kEventCost[k];
    }
    return flops;
}
int oracle_event_count() { return EV_COUNT; }

// world.hit() per ray, media skipped.  rays: n x 7 doubles (o, d, time).  out: n x 12 doubles
// [hit, t, px, py, pz, nx, ny, nz, u, v, material, prim].
static void hit_range(const rtw_flat_scene* fs, const double* rays, int64_t lo, int64_t hi, double tmin, double tmax, double* out) {
    World w{fs, nullptr, 0};
    for (int64_t i = lo; i < hi; i++) {
        const double* q = rays + 7 * i;
        Ray r{mk(q[0], q[1], q[2]), mk(q[3], q[4], q[5]), q[6]};
        Rec rec;
        bool h = world_hit(w, r, tmin, tmax, rec);
        double* o = out + 12 * i;
        for (int k = 0; k < 12; k++) o[k] = 0;
        o[10] = o[11] = -1;
        if (h) {
            o[0] = 1, o[1] = rec.t, o[2] = rec.p.x, o[3] = rec.p.y, o[4] = rec.p.z;
            o[5] = rec.normal.x, o[6] = rec.normal.y, o[7] = rec.normal.z, o[8] = rec.u, o[9] = rec.v;
            o[10] = rec.mat, o[11] = rec.prim;
        }
    }
}
void oracle_hit_batch(const rtw_flat_scene* fs, const double* rays, int64_t n, double tmin, double tmax, double* out) {
    // rays are independent and world_hit keeps no state between them: big batches are split over the host threads
    const int threads = n >= (1 << 16) ? (int)std::max(1u, std::thread::hardware_concurrency()) : 1;
    if (threads == 1) {
        hit_range(fs, rays, 0, n, tmin, tmax, out);
        return;
    }
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; t++) pool.emplace_back(hit_range, fs, rays, n * t / threads, n * (t + 1) / threads, tmin, tmax, out);
    for (auto& t : pool) t.join();
}

// Sensitivity probe for the path-by-path parity tests: with eps > 0 a philox-mode render perturbs every ray by a relative
// eps per component (tests use 2^-22, the size of accumulated FP32 rounding) — the pixels that change are the ones where
// an FP32 kernel may legitimately leave the FP64 path.  eps = 0 switches it off.
void oracle_set_jitter(double eps, uint64_t seed) { g_jitter_eps = eps, g_jitter_seed = seed; }

// Known-answer hooks for the RNG and the philox-mode samplers (tests/test_rng_kat.py).
void oracle_philox(const uint32_t* counter, const uint32_t* key, uint32_t* out) { philox4x32_10(counter, key, out); }
// which: 0 unit ball (SITE_BALL draw), 1 lens disk (SITE_LENS draw), 2 cosine_pdf::generate about +z (SITE_MAIN draw, r1 = u[2],
// r2 = u[3] as in the lambertian branch).  Sample i uses the counter (pixel = i, sample = 0, bounce = 1, site).  out: n x 3.
void oracle_sampler_batch(int which, int64_t n, uint64_t seed, double* out) {
    Rng g;
    g.mode = MODE_PHILOX;
    g.seed_lo = (uint32_t)seed, g.seed_hi = (uint32_t)(seed >> 32);
    for (int64_t i = 0; i < n; i++) {
        g.pixel = (uint32_t)i, g.sample = 0;
        double u[4];
        V v;
        if (which == 0) g.philox(1, SITE_BALL, u), v = ball_from_uniforms(u[0], u[1], u[2]);
        else if (which == 1) g.philox(1, SITE_LENS, u), v = disk_from_uniforms(u[0], u[1]);
        else g.philox(1, SITE_MAIN, u), v = onb_from_w(mk(0, 0, 1)).local(random_cosine_direction(u[2], u[3]));  // cosine_pdf about +z (pdf.h:18, :26-29)
        out[3 * i] = v.x, out[3 * i + 1] = v.y, out[3 * i + 2] = v.z;
    }
}

// The image loop of RayTracingWeekend.cpp:214-250.  mode 0 (minstd): serial, rows outer, draws in the
// reference's order — bit-comparable with `ref_harness render` under RTW_PPL_THREADS=1.  mode 1 (philox):
// the device's streams; rows are split over `threads`.  Output: per pixel sum(rgb) and sum(rgb^2), pre-gamma.
void oracle_render(const rtw_flat_scene* fs, int nx, int ny, int spp_begin, int spp, int max_depth, int mode, uint64_t seed,
                   int threads, double* out_sum, double* out_sq, uint64_t* out_rays) {
    if (mode == MODE_MINSTD) threads = 1;
    if (threads < 1) threads = 1;
    std::vector<uint64_t> rays(threads, 0);
    std::vector<std::vector<uint64_t>> evs(threads, std::vector<uint64_t>(EV_COUNT, 0));
    auto work = [&](int tid) {
        g_ev = g_events_on ? evs[tid].data() : nullptr;
        Tracer tr{fs};
        tr.rng.mode = mode;
        tr.rng.seed_lo = (uint32_t)(seed & 0xffffffffu);
        tr.rng.seed_hi = (uint32_t)(seed >> 32);
        for (int j = tid; j < ny; j += threads)
            for (int i = 0; i < nx; i++) {
                V sum = mk(0, 0, 0), sq = mk(0, 0, 0);
                for (int s = 0; s < spp; s++) {
                    double ju, jv;
                    if (mode == MODE_MINSTD) {
                        ju = tr.rng.uniform(tr.rng.image);  // :227
                        jv = tr.rng.uniform(tr.rng.image);  // :228
                    } else {
                        tr.rng.pixel = (uint32_t)(j * nx + i);
                        tr.rng.sample = (uint32_t)(spp_begin + s);
                        double u[4];
                        tr.rng.philox(0, SITE_MAIN, u);
                        ju = u[0], jv = u[1];
                    }
                    double u = double(i + ju) / double(nx);
                    double v = double(j + jv) / double(ny);
                    Ray r = tr.get_ray(u, v);
                    V c = tr.color(r, max_depth, 1);
                    sum = sum + c;
                    sq = sq + c * c;
                }
                double* o = out_sum + 3 * ((size_t)j * nx + i);
                o[0] = sum.x, o[1] = sum.y, o[2] = sum.z;
                if (out_sq) {
                    double* q = out_sq + 3 * ((size_t)j * nx + i);
                    q[0] = sq.x, q[1] = sq.y, q[2] = sq.z;
                }
            }
        rays[tid] = tr.rays;
        g_ev = nullptr;
    };
    if (threads == 1) work(0);
    else {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; t++) pool.emplace_back(work, t);
        for (auto& t : pool) t.join();
    }
    uint64_t total = 0;
    for (uint64_t r : rays) total += r;
    if (out_rays) *out_rays = total;
    for (auto& e : evs)
        for (int k = 0; k < EV_COUNT; k++) g_event_totals[k] += e[k];
}

int oracle_hardware_threads() { return (int)std::thread::hardware_concurrency(); }

}  // extern "C"

// ------------------------------------------------------------------------------------------------ per-function KAT
// Same call sequence as `ref_harness funcs` (oracle/ref_harness.cpp cmd_funcs) from fresh engines; every number
// must equal the reference's bit for bit.  `tables`: 768 ranvec doubles + 768 perm values as doubles (as written
// by `ref_harness perlin`).  Returns the count written.
extern "C" int oracle_funcs(const double* tables, double* out, int cap) {
    std::vector<double> o;
    auto put3 = [&](V v) { o.push_back(v.x), o.push_back(v.y), o.push_back(v.z); };
    Tracer tr{nullptr};
    Rng& g = tr.rng;
    g.mode = MODE_MINSTD;
    for (int i = 0; i < 8; i++) o.push_back(g.random_double());
    for (int i = 0; i < 8; i++) o.push_back((double)(0 + std::min(1 - 0, (int)((1 - 0 + 1) * g.random_double()))));
    for (int i = 0; i < 8; i++) put3(random_in_unit_sphere(g, 0));
    for (int i = 0; i < 8; i++) {
        double r1 = g.random_double(), r2 = g.random_double();
        put3(random_cosine_direction(r1, r2));
    }
    for (int i = 0; i < 8; i++) {
        double r1 = g.random_double(), r2 = g.random_double();
        put3(random_to_sphere(90.0, 250000.0, r1, r2));
    }
    {
        Onb a = onb_from_w(mk(0.3, -0.5, 0.8)), b = onb_from_w(mk(-2.0, 0.1, 0.2));
        put3(a.u), put3(a.v), put3(a.w), put3(b.u), put3(b.v), put3(b.w);
        put3(a.local(mk(0.1, 0.2, 0.3)));
    }
    // flat scene with the two Cornell lights only
    rtw_flat_light lights[2] = {};
    lights[0].type = RTW_LIGHT_XZ_RECT;
    double lr[5] = {213.0, 343.0, 227.0, 332.0, 554.0}, ls[4] = {190, 90, 190, 90};
    memcpy(lights[0].v, lr, sizeof lr);
    lights[1].type = RTW_LIGHT_SPHERE;
    memcpy(lights[1].v, ls, sizeof ls);
    rtw_flat_scene fs{};
    fs.n_lights = 2;
    fs.lights = lights;
    std::vector<int32_t> perm(768);
    if (tables) {
        for (int i = 0; i < 768; i++) perm[i] = (int32_t)tables[768 + i];
        fs.perlin_ranvec = tables;
        fs.perlin_perm = perm.data();
    }
    {
        V org = mk(100.0, 0.0, 300.0);
        Onb uvw = onb_from_w(mk(0, 1, 0));
        for (int i = 0; i < 6; i++) {
            V dir;
            if (g.random_double() < 0.5) {
                double r1 = g.random_double(), r2 = g.random_double();
                dir = uvw.local(random_cosine_direction(r1, r2));
            } else {
                int idx = 0 + std::min(1, (int)(2 * g.random_double()));
                double r1, r2;
                if (idx == 0) r2 = g.uniform(g.util), r1 = g.uniform(g.util);
                else r1 = g.random_double(), r2 = g.random_double();
                dir = light_random(lights[idx], org, r1, r2);
            }
            put3(dir);
            double cosine = dot(normalize(dir), uvw.w);
            double pc = (cosine <= 0) ? 0 : cosine / M_PI;
            o.push_back(0.5 * pc + 0.5 * lights_pdf_value(&fs, org, dir));
            o.push_back(light_pdf_value(lights[0], org, dir));
            o.push_back(light_pdf_value(lights[1], org, dir));
        }
    }
    {
        Rec rec;
        rec.t = 1.0;
        rec.p = mk(1, 2, 3);
        rec.normal = normalize(mk(0.2, 1.0, -0.1));
        rec.u = 0.25, rec.v = 0.75;
        Ray rin{mk(0, 5, 0), mk(0.3, -1.2, 0.4), 0.5}, rout{mk(1, 1, 4), mk(-0.3, 1.2, -0.4), 0.5};
        rtw_flat_material di{};
        di.type = RTW_MAT_DIELECTRIC, di.ref_idx = 1.5;
        for (int i = 0; i < 6; i++) put3(tr.dielectric_scatter(di, i % 2 ? rout : rin, rec, 0));
        for (int i = 0; i < 4; i++) put3(tr.reflect(normalize(rin.d), rec.normal) + 0.3 * random_in_unit_sphere(g, 0));
        for (int i = 0; i < 4; i++) put3(random_in_unit_sphere(g, 0));
        double c2 = dot(rec.normal, normalize(rout.d));
        o.push_back(c2 < 0 ? 0 : c2 / M_PI);
        o.push_back(tr.schlick(0.3, 1.5));
        V refr = mk(0, 0, 0);
        o.push_back(tr.refract(mk(0.3, -1.2, 0.4), rec.normal, 1.0 / 1.5, refr) ? 1.0 : 0.0);
        put3(refr);
        put3(tr.reflect(mk(0.3, -1.2, 0.4), rec.normal));
    }
    if (tables) {
        rtw_flat_texture tx[4] = {};
        tx[0].type = RTW_TEX_NOISE, tx[0].scale = 4.0;
        tx[1].type = RTW_TEX_CONSTANT, tx[1].color[0] = 0.2, tx[1].color[1] = 0.3, tx[1].color[2] = 0.1;
        tx[2].type = RTW_TEX_CONSTANT, tx[2].color[0] = tx[2].color[1] = tx[2].color[2] = 0.9;
        tx[3].type = RTW_TEX_CHECKER, tx[3].even = 1, tx[3].odd = 2;
        fs.textures = tx;
        V pts[4] = {mk(0.1, 0.2, 0.3), mk(-3.7, 2.2, 9.1), mk(100.5, -0.25, 0.001), mk(-0.5, -0.5, -0.5)};
        for (V p : pts) {
            put3(texture_value(&fs, 0, 0, 0, p));
            put3(texture_value(&fs, 3, 0, 0, p));
            o.push_back(perlin_noise(&fs, p));
            o.push_back(perlin_turb(&fs, p));
        }
        uint8_t bytes[48];
        for (int i = 0; i < 48; i++) bytes[i] = (uint8_t)(i * 5);
        rtw_flat_texture im{};
        im.type = RTW_TEX_IMAGE, im.nx = 4, im.ny = 4, im.image_offset = 0;
        fs.textures = &im;
        fs.texels = bytes;
        double uv[5][2] = {{0, 0}, {1, 1}, {0.5, 0.5}, {0.99, 0.01}, {0.26, 0.74}};
        for (auto& q : uv) put3(texture_value(&fs, 0, q[0], q[1], mk(0, 0, 0)));
    }
    {
        // camera(vec3(24,5,5), vec3(0,3,0), vup, 20, 1.5, 0.2f, 10, 0, 1) — members computed like camera.h:13-34
        rtw_flat_camera& c = fs.camera;
        double theta = 20.0 * static_cast<double>(M_PI) / 180.0, half_height = tan(theta / 2), half_width = 1.5 * half_height;
        V origin = mk(24, 5, 5), w = normalize(origin - mk(0, 3, 0)), u = normalize(cross(mk(0, 1, 0), w)), v = cross(w, u);
        double fd = 10.0;
        V llc = origin - half_width * fd * u - half_height * fd * v - fd * w;
        V hor = 2.0 * half_width * fd * u, ver = 2.0 * half_height * fd * v;
        auto st = [](double* d, V a) { d[0] = a.x, d[1] = a.y, d[2] = a.z; };
        st(c.origin, origin), st(c.lower_left_corner, llc), st(c.horizontal, hor), st(c.vertical, ver), st(c.u, u), st(c.v, v), st(c.w, w);
        c.time0 = 0, c.time1 = 1, c.lens_radius = 0.2f / 2;
        tr.fs = &fs;
        for (int i = 0; i < 4; i++) {
            Ray r = tr.get_ray(0.25 * i, 0.1 + 0.2 * i);
            put3(r.o), put3(r.d);
            o.push_back(r.time);
        }
    }
    {
        rtw_flat_prim bnd{};
        bnd.type = RTW_PRIM_SPHERE, bnd.xf = -1, bnd.material = -1, bnd.v[3] = 50;
        rtw_flat_item it{};
        it.type = RTW_ITEM_MEDIUM, it.xf = -1, it.first = 0, it.count = 1, it.material = -1, it.multiplicity = 2, it.density = 0.02;
        fs.prims = &bnd, fs.n_prims = 1, fs.items = &it, fs.n_items = 1, fs.n_world_items = 1;
        World w{&fs, &g, 0};
        Ray r{mk(0, 0, -200), mk(0, 0, 1), 0.0};
        for (int i = 0; i < 6; i++) {
            Rec rec;
            bool h = medium_hit(w, 0, 0, r, 0.001f, std::numeric_limits<double>::max(), rec);
            o.push_back(h ? rec.t : -1.0);
        }
        for (int i = 0; i < 6; i++) {
            Rec rec;
            bool h = world_hit(w, r, 0.001f, std::numeric_limits<double>::max(), rec);
            o.push_back(h ? rec.t : -1.0);
        }
    }
    int n = (int)std::min<size_t>(o.size(), (size_t)cap);
    memcpy(out, o.data(), n * sizeof(double));
    return (int)o.size();
}

// CppTest/unittest1.cpp:20-40, 72-109 restated on this file's vec3/ray/aabb forms.  out[22]:
// dot, cross(3), point_at_parameter(3), aabb hits(5), surrounding min(3), max(3), sizeof(vec3), reserved(3)
extern "C" void oracle_kat(double* out) {
    int k = 0;
    out[k++] = dot(mk(1, 1, 0), mk(1, 1, 0));
    V c = cross(mk(1, 0, 0), mk(0, 1, 0));
    out[k++] = c.x, out[k++] = c.y, out[k++] = c.z;
    Ray r{mk(1, 1, 1), mk(2, 2, 2), 0.0};
    V p = r.at(3);
    out[k++] = p.x, out[k++] = p.y, out[k++] = p.z;
    double mn[3] = {2, 2, 2}, mx[3] = {4, 4, 4};
    Ray cases[5] = {{mk(0, 0, 0), mk(1, 1, 1), 0}, {mk(0, 0, 0), -mk(1, 1, 1), 0}, {mk(3, 3, 3), mk(0, 1, 0), 0},
                    {mk(0, 3, 0), mk(1, 0, 1), 0}, {mk(0, 5, 0), mk(1, 0, 1), 0}};
    for (auto& q : cases) out[k++] = aabb_hit(mn, mx, q, 0, FLT_MAX) ? 1 : 0;
    double a0[3] = {0, 0, 0}, a1[3] = {1, 1, 1}, b0[3] = {3, 3, 3}, b1[3] = {4, 4, 4};
    for (int i = 0; i < 3; i++) out[k++] = fmin(a0[i], b0[i]);  // aabb.h:49-62
    for (int i = 0; i < 3; i++) out[k++] = fmax(a1[i], b1[i]);
    out[k++] = (double)sizeof(V);
}
