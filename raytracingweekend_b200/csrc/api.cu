// api.cu — device scene upload + the hot-path entry points of include/rtw_b200.h.
// No CPU fallback: every function here fails with RTW_ERR_CUDA when the CUDA runtime or an sm_100 device
// is missing.
#include <algorithm>
#include <cmath>
#include <functional>
#include <limits>
#include <mutex>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <dlfcn.h>
#include <nccl.h>  // types only: the library is bound with dlopen (multi-GPU section at the end)

#include "flat.hpp"
#include "kernels.hpp"
#include "lbvh.hpp"

using namespace rtw;

namespace {

// Device memory comes from a PRIVATE stream-ordered pool per device (created once, release threshold unlimited): a host
// that creates, renders and destroys a scene per frame (bench.py's end-to-end step, the adapter's render()) gets the same
// memory back in microseconds instead of paying cudaMalloc / cudaFree (~1 ms each, device-wide synchronisation) every
// time, and the host application's default pool keeps its own attributes.  The pool keeps what was freed (at most the peak
// use) until the process ends.  Allocation is followed by a synchronisation of the legacy stream, so the pointer is valid
// on every stream, as with cudaMalloc.
cudaError_t dev_alloc(void** p, size_t bytes) {
    static std::mutex mu;
    static cudaMemPool_t pools[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    cudaMemPool_t pool;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!pools[dev]) {
            cudaMemPoolProps props{};
            props.allocType = cudaMemAllocationTypePinned;
            props.handleTypes = cudaMemHandleTypeNone;
            props.location.type = cudaMemLocationTypeDevice;
            props.location.id = dev;
            if ((e = cudaMemPoolCreate(&pools[dev], &props)) != cudaSuccess) {
                pools[dev] = nullptr;
                return e;
            }
            unsigned long long keep = ~0ull;
            cudaMemPoolSetAttribute(pools[dev], cudaMemPoolAttrReleaseThreshold, &keep);
        }
        pool = pools[dev];
    }
    if ((e = cudaMallocFromPoolAsync(p, bytes, pool, 0)) != cudaSuccess) return e;
    return cudaStreamSynchronize(0);
}
// Stream-ordered free on the legacy stream.  The CALLER guarantees that no work on another stream still uses `p`:
// rtw_scene::release waits for the scene's last render (one event, not a device-wide synchronisation per buffer), the
// host-buffer entry points are synchronous.
void dev_free(void* p) {
    if (!p) return;
    cudaFreeAsync(p, 0);
}

struct DevBuf {
    void* p = nullptr;
    template <typename T>
    cudaError_t upload(const std::vector<T>& v) {
        if (v.empty()) return cudaSuccess;
        cudaError_t e = dev_alloc(&p, v.size() * sizeof(T));
        if (e != cudaSuccess) return e;
        return cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice);
    }
    void release() {
        dev_free(p);
        p = nullptr;
    }
};

int cuda_fail(cudaError_t e, const char* what) {
    set_error(std::string(what) + ": " + cudaGetErrorString(e));
    return RTW_ERR_CUDA;
}

// Compose a translate/rotate_y chain (outermost first) into local = Ry(o) + t, in FP64.
void compose_chain(const rtw_flat_scene* fs, int head, DevXf& out) {
    double C = 1, S = 0, t[3] = {0, 0, 0};
    for (int i = head; i >= 0; i = fs->xfs[i].next) {
        const rtw_flat_xf& x = fs->xfs[i];
        if (x.kind == RTW_XF_TRANSLATE) {  // hittable.h:301: origin - offset
            t[0] -= x.a, t[1] -= x.b, t[2] -= x.c;
        } else {  // hittable.h:378-382: x' = cos*x - sin*z, z' = sin*x + cos*z
            double s = x.a, c = x.b;
            double nC = c * C - s * S, nS = s * C + c * S;
            double tx = c * t[0] - s * t[2], tz = s * t[0] + c * t[2];
            C = nC, S = nS, t[0] = tx, t[2] = tz;
        }
    }
    out = DevXf{(float)C, (float)S, (float)t[0], (float)t[1], (float)t[2], 0, 0, 0};
}

// Composition in FP64 (bounding boxes): local = Ry(o) + t with Ry: x' = C x - S z, z' = S x + C z.
void compose_chain_f64(const rtw_flat_scene* fs, int head, double& C, double& S, double t[3]) {
    C = 1, S = 0, t[0] = t[1] = t[2] = 0;
    for (int i = head; i >= 0; i = fs->xfs[i].next) {
        const rtw_flat_xf& x = fs->xfs[i];
        if (x.kind == RTW_XF_TRANSLATE) {
            t[0] -= x.a, t[1] -= x.b, t[2] -= x.c;
        } else {
            double s = x.a, c = x.b;
            double nC = c * C - s * S, nS = s * C + c * S;
            double tx = c * t[0] - s * t[2], tz = s * t[0] + c * t[2];
            C = nC, S = nS, t[0] = tx, t[2] = tz;
        }
    }
}

struct Box {
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    void grow_axis(int k, double a, double b) {
        lo[k] = std::min(lo[k], std::min(a, b));
        hi[k] = std::max(hi[k], std::max(a, b));
    }
    void grow_point(double x, double y, double z) { grow_axis(0, x, x), grow_axis(1, y, y), grow_axis(2, z, z); }
    void merge(const Box& o) {
        for (int k = 0; k < 3; k++) lo[k] = std::min(lo[k], o.lo[k]), hi[k] = std::max(hi[k], o.hi[k]);
    }
    double centre(int k) const { return 0.5 * (lo[k] + hi[k]); }
    double area() const {
        if (hi[0] < lo[0]) return 0.0;
        double x = hi[0] - lo[0], y = hi[1] - lo[1], z = hi[2] - lo[2];
        return 2.0 * (x * y + y * z + z * x);
    }
};

inline float __int_as_float_host(int v) {
    float f;
    memcpy(&f, &v, sizeof f);
    return f;
}

int bvh_depth(const rtw_flat_scene* fs, int ref, int depth) {
    if (ref < 0) return depth;
    if (depth > 4096) return depth;
    const rtw_flat_node& n = fs->nodes[ref];
    int l = bvh_depth(fs, n.left, depth + 1);
    int r = n.right == n.left ? l : bvh_depth(fs, n.right, depth + 1);
    return l > r ? l : r;
}

}  // namespace

struct rtw_scene {
    int device = 0;
    DevScene dev{};
    DevBuf blob, texels;
    unsigned long long* d_counter = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // pending asynchronous render
    bool pending = false;
    rtw_render_params pending_params{};
    int pending_launches = 0;
    float* scratch_accum = nullptr;  // device buffer behind rtw_render (host API)
    size_t scratch_pixels = 0;
    bool prefer_wavefront = false;  // RTW_FLAG_AUTO: set at upload from what the scene's shading costs
    void* wf_scratch = nullptr;  // wavefront path pool + queues (lazy)
    uint32_t wf_slots = 0;
    size_t wf_bytes = 0;
    unsigned long long* wf_host_flag = nullptr;  // pinned
    bool rendered = false;  // ev1 was recorded at least once
    void release() {
        // one wait for the scene's last render instead of a device-wide synchronisation per buffer: every other use of
        // these buffers (upload, intersect_batch, host-buffer renders) is synchronous
        if (rendered && ev1) cudaEventSynchronize(ev1);
        blob.release();
        texels.release();
        dev_free(d_counter);
        dev_free(scratch_accum);
        dev_free(wf_scratch);
        if (wf_host_flag) cudaFreeHost(wf_host_flag);
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
    }
};

extern "C" int rtw_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    int ok = 0;
    for (int i = 0; i < n; i++) {
        cudaDeviceProp p;
        if (cudaGetDeviceProperties(&p, i) == cudaSuccess && p.major == 10) ok++;
    }
    return ok;
}

// ------------------------------------------------------------------------------------------------ upload
extern "C" rtw_scene* rtw_scene_create(const rtw_flat_scene* fs, int device) {
    if (!fs || fs->n_world_items < 0) {
        set_error("scene_create: null flat scene");
        return nullptr;
    }
    // RTW_DRY_RUN=1 (diagnostics only): build the device-side tables without a device, print the RTW_DEBUG_TREES line, fail
    const bool dry_run = getenv("RTW_DRY_RUN") != nullptr;
    int ndev = 0;
    cudaError_t e = dry_run ? cudaSuccess : cudaGetDeviceCount(&ndev);
    if (dry_run) ndev = device + 1;
    if (e != cudaSuccess || device < 0 || device >= ndev) {
        set_error(std::string("scene_create: no usable CUDA device (") + (e != cudaSuccess ? cudaGetErrorString(e) : "bad index") +
                  "); this backend has no CPU fallback");
        cudaGetLastError();
        return nullptr;
    }
    int cc_major = 0;  // one attribute, not cudaGetDeviceProperties (which takes ~1 ms and this runs once per scene)
    if (!dry_run && ((e = cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, device)) != cudaSuccess || cc_major != 10)) {
        set_error("scene_create: device is not compute capability 10.x (kernels are built for sm_100a only)");
        return nullptr;
    }
    if (!dry_run && (e = cudaSetDevice(device)) != cudaSuccess) {
        cuda_fail(e, "cudaSetDevice");
        return nullptr;
    }
    if (fs->n_xfs > (int)PM_XF_MASK - 1 || fs->n_materials >= (1 << 15)) {
        set_error("scene_create: too many transforms/materials for the packed primitive word");
        return nullptr;
    }

    uint32_t feat = 0;
    std::vector<DevXf> xfs(fs->n_xfs);
    for (int i = 0; i < fs->n_xfs; i++) compose_chain(fs, i, xfs[i]);  // entry i = chain starting at link i
    if (fs->n_xfs) feat |= F_XF;

    // Spheres of |radius| >= this are tested in FP64 (ground spheres, where FP32 cancels in b*b - a*c).  RTW_PRECISE_MIN_RADIUS
    // overrides it for A/B runs.  Measured at 500 (profiles/r2_ab_precise_radius.txt): the Book-2 final scene gains 2.7 % (its
    // r = 100 earth sphere leaves the FP64 square root) and stays inside the tier-2 gates, but the r = 100 GROUND sphere of the
    // book-1 scenes, seen at grazing angles, lands 5.4 coordinate-granularity units from the oracle (gate: 2) — so 100 stays.
    static const double precise_min_radius = [] { const char* v = getenv("RTW_PRECISE_MIN_RADIUS"); return v ? atof(v) : 100.0; }();
    std::vector<float4> pa(fs->n_prims), pb(fs->n_prims);
    std::vector<uint32_t> pm(fs->n_prims);
    std::vector<double> pd;
    for (int i = 0; i < fs->n_prims; i++) {
        const rtw_flat_prim& p = fs->prims[i];
        uint32_t m = (uint32_t)p.type | (p.flip ? PM_FLIP : 0u) | ((uint32_t)(p.xf + 1) << PM_XF_SHIFT) |
                     ((uint32_t)p.material << PM_MAT_SHIFT);
        pb[i] = make_float4(0, 0, 0, 0);
        if (p.type == RTW_PRIM_SPHERE) {
            pa[i] = make_float4((float)p.v[0], (float)p.v[1], (float)p.v[2], (float)p.v[3]);
            if (std::fabs(p.v[3]) >= precise_min_radius) {
                m |= PM_PRECISE;
                feat |= F_PRECISE;
                pb[i].w = __int_as_float_host((int)(pd.size() / 4));
                for (int k = 0; k < 4; k++) pd.push_back(p.v[k]);
            }
        } else if (p.type == RTW_PRIM_MOVING_SPHERE) {
            // sphere.h:22-25: center0 + ((time - time0)/(time1 - time0))*(center1 - center0) = c' + time*delta'
            double t0 = p.v[7], t1 = p.v[8], inv = 1.0 / (t1 - t0);
            double d[3], c[3];
            for (int k = 0; k < 3; k++) {
                d[k] = (p.v[4 + k] - p.v[k]) * inv;
                c[k] = p.v[k] - t0 * d[k];
            }
            pa[i] = make_float4((float)c[0], (float)c[1], (float)c[2], (float)p.v[3]);
            pb[i] = make_float4((float)d[0], (float)d[1], (float)d[2], 0.0f);
            feat |= F_MOVING;
        } else {
            pa[i] = make_float4((float)p.v[0], (float)p.v[1], (float)p.v[2], (float)p.v[3]);
            pb[i].x = (float)p.v[4];
        }
        pm[i] = m;
    }

    // primitive-test groups and records (device_scene.hpp): split a set of primitives (ascending list order) into maximal
    // runs that share a transform chain, and order each run's records spheres, moving spheres, xy, xz, yz rects.
    // Built with INDICES first; converted to blob byte offsets once the section offsets are known (below).
    std::vector<DevGroup> groups;
    std::vector<DevRec> recs;
    // The six rects emitted for a reference `box` (builder.cpp K_BOX: xy z1, xy z0, xz y1, xz y0, yz x1, yz x0 over the
    // same extents): detected here so that they get a group of their own with a box record (device_scene.hpp).
    static const bool use_boxes = [] { const char* v = getenv("RTW_BOX_PRIMS"); return !v || atoi(v) != 0; }();
    auto is_box_at = [&](const std::vector<int>& ids, size_t q, size_t end) {
        if (!use_boxes || q + 6 > end) return false;
        const rtw_flat_prim* P[6];
        for (int k = 0; k < 6; k++) {
            if (ids[q + k] != ids[q] + k) return false;
            P[k] = &fs->prims[ids[q + k]];
        }
        const int want[6] = {RTW_PRIM_XY_RECT, RTW_PRIM_XY_RECT, RTW_PRIM_XZ_RECT, RTW_PRIM_XZ_RECT, RTW_PRIM_YZ_RECT, RTW_PRIM_YZ_RECT};
        for (int k = 0; k < 6; k++)
            if (P[k]->type != want[k]) return false;
        const double x0 = P[0]->v[0], x1 = P[0]->v[1], y0 = P[0]->v[2], y1 = P[0]->v[3], z1 = P[0]->v[4], z0 = P[1]->v[4];
        if (!(x0 < x1 && y0 < y1 && z0 < z1)) return false;
        auto same = [](const rtw_flat_prim* a, double a0, double a1, double b0, double b1, double k) {
            return a->v[0] == a0 && a->v[1] == a1 && a->v[2] == b0 && a->v[3] == b1 && a->v[4] == k;
        };
        return same(P[1], x0, x1, y0, y1, z0) && same(P[2], x0, x1, z0, z1, y1) && same(P[3], x0, x1, z0, z1, y0) &&
               same(P[4], y0, y1, z0, z1, x1) && same(P[5], y0, y1, z0, z1, x0);
    };
    // Room records: a set of >= 3 axis-aligned rects of one run that are faces of ONE box (the five walls of the Cornell
    // room, Scene/scene.h:197-210: at most one rect per face, extents exactly the box's) is tested by the render kernels with
    // one slab test instead of one plane test per wall: the nearest rect hit is the entry face when the ray starts outside and
    // that face exists, else the exit face (the box is convex, so no other face can be nearer); t is the very expression of
    // the rect test.  Same record as a box plus a table of the six face ids (-1 = no such wall).  Only formed when the run's
    // other rects all come before (or all after) the room's in list order, so that the "later rect wins an exact tie" rule
    //  is kept by the order of the two groups; intersect_batch keeps the rect-by-rect order either way.
    // Measured on B200 (profiles/r2_ab_rooms.txt): NEUTRAL — Cornell + glass + aluminium 6.60 Gpaths/s without, 6.56 with: the
    // slab test saves ~60 of the five plane tests' ~110 instructions per ray, and splitting the run into "room" and "the rest"
    // adds a group (~30 instructions of header loads and kind-loop bounds) — so it is opt-in (RTW_ROOM_PRIMS=1), off by default.
    static const bool use_rooms = [] { const char* v = getenv("RTW_ROOM_PRIMS"); return v && atoi(v) != 0; }();
    struct Room {
        double lo[3], hi[3];
        int face[6];  // box face order (builder.cpp K_BOX): z = hi, z = lo, y = hi, y = lo, x = hi, x = lo
        int n = 0;
    };
    auto rect_axes = [&](int p, int ax[3]) {  // in-plane axes a, b and the normal axis
        const int t = fs->prims[p].type;
        if (t == RTW_PRIM_XY_RECT) ax[0] = 0, ax[1] = 1, ax[2] = 2;
        else if (t == RTW_PRIM_XZ_RECT) ax[0] = 0, ax[1] = 2, ax[2] = 1;
        else ax[0] = 1, ax[1] = 2, ax[2] = 0;
    };
    auto find_room = [&](const std::vector<int>& members, Room& best) {
        best.n = 0;
        std::vector<int> rects;
        for (int p : members)
            if (fs->prims[p].type >= RTW_PRIM_XY_RECT && fs->prims[p].type <= RTW_PRIM_YZ_RECT) rects.push_back(p);
        if (!use_rooms || rects.size() < 3 || rects.size() > 64) return;
        for (int seed : rects) {
            // the seed fixes two ranges; the third comes from another rect's in-plane range that ends on the seed's plane
            Room r;
            bool known[3] = {false, false, false};
            for (int k = 0; k < 6; k++) r.face[k] = -1;
            int ax[3];
            rect_axes(seed, ax);
            const double* v = fs->prims[seed].v;
            if (!(v[0] < v[1] && v[2] < v[3])) continue;
            r.lo[ax[0]] = v[0], r.hi[ax[0]] = v[1], r.lo[ax[1]] = v[2], r.hi[ax[1]] = v[3];
            known[ax[0]] = known[ax[1]] = true;
            for (int q : rects) {
                int bx[3];
                rect_axes(q, bx);
                const double* w = fs->prims[q].v;
                for (int e = 0; e < 2 && !known[ax[2]]; e++)
                    if (bx[e] == ax[2] && w[2 * e] < w[2 * e + 1] && (v[4] == w[2 * e] || v[4] == w[2 * e + 1]))
                        r.lo[ax[2]] = w[2 * e], r.hi[ax[2]] = w[2 * e + 1], known[ax[2]] = true;
            }
            if (!known[ax[2]]) continue;
            for (int q : rects) {
                int bx[3];
                rect_axes(q, bx);
                const double* w = fs->prims[q].v;
                if (w[0] != r.lo[bx[0]] || w[1] != r.hi[bx[0]] || w[2] != r.lo[bx[1]] || w[3] != r.hi[bx[1]]) continue;
                const bool at_hi = w[4] == r.hi[bx[2]], at_lo = w[4] == r.lo[bx[2]];
                if (!at_hi && !at_lo) continue;
                const int f = (bx[2] == 2 ? 0 : (bx[2] == 1 ? 2 : 4)) + (at_hi ? 0 : 1);
                if (r.face[f] >= 0) {  // two rects on one face: not a room
                    r.n = -1000;
                    break;
                }
                r.face[f] = q, r.n++;
            }
            if (r.n > best.n) best = r;
        }
        if (best.n < 3) {
            best.n = 0;
            return;
        }
        // a rect of the run that is not a face but lies in a face's plane could tie with it exactly: leave such runs alone
        int lo_idx = 1 << 30, hi_idx = -1;
        for (int k = 0; k < 6; k++)
            if (best.face[k] >= 0) lo_idx = std::min(lo_idx, best.face[k]), hi_idx = std::max(hi_idx, best.face[k]);
        bool before = true, after = true;
        for (int q : rects) {
            bool is_face = false;
            for (int k = 0; k < 6; k++) is_face |= best.face[k] == q;
            if (is_face) continue;
            int bx[3];
            rect_axes(q, bx);
            const double k4 = fs->prims[q].v[4];
            if (k4 == best.lo[bx[2]] || k4 == best.hi[bx[2]]) {
                best.n = 0;
                return;
            }
            before &= q < lo_idx, after &= q > hi_idx;
        }
        if (!before && !after) best.n = 0;
        else if (!before) best.n = -best.n;  // n < 0: the run's other members go AFTER the room
    };
    // one group of records for `members` (ascending list order, one transform chain); `box_first` >= 0: also the box record of
    // the six rects starting there; `room`: also a room record
    auto emit_group = [&](const std::vector<int>& members, int box_first, const Room* room) {
        DevGroup g{};
        g.xf = fs->prims[members[0]].xf;
        g.sph0 = (int)recs.size();
        const int kinds[5] = {RTW_PRIM_SPHERE, RTW_PRIM_MOVING_SPHERE, RTW_PRIM_XY_RECT, RTW_PRIM_XZ_RECT, RTW_PRIM_YZ_RECT};
        int32_t* ends[5] = {&g.sph1, &g.msph1, &g.xy1, &g.xz1, &g.yz1};
        for (int k = 0; k < 5; k++) {
            for (int p : members) {
                if (fs->prims[p].type != kinds[k]) continue;
                DevRec r{};
                r.a = pa[p];
                if (k == 0) {
                    int id = p;
                    if (pm[p] & PM_PRECISE) {
                        id |= REC_PRECISE;
                        r.b.x = pb[p].w;  // pd index bits (-> byte offset below)
                    }
                    r.b.w = __int_as_float_host(id);
                } else if (k == 1) {
                    r.b = make_float4(pb[p].x, pb[p].y, pb[p].z, __int_as_float_host(p));
                } else {
                    r.b = make_float4(pb[p].x, 0.0f, 0.0f, __int_as_float_host(p));
                }
                recs.push_back(r);
            }
            *ends[k] = (int)recs.size();
        }
        if (box_first >= 0) {
            const rtw_flat_prim& xy1r = fs->prims[box_first];      // x0 x1 y0 y1 | z1
            const rtw_flat_prim& xy0r = fs->prims[box_first + 1];  // z0
            DevRec r{};
            // the same FP32 constants as the rect records (round to nearest), so t is bit-identical to the rect test
            r.a = make_float4((float)xy1r.v[0], (float)xy1r.v[2], (float)xy0r.v[4], 0.0f);
            r.b = make_float4((float)xy1r.v[1], (float)xy1r.v[3], (float)xy1r.v[4], __int_as_float_host(box_first));
            recs.push_back(r);
        }
        if (room) {
            DevRec r{}, ids{};
            r.a = make_float4((float)room->lo[0], (float)room->lo[1], (float)room->lo[2], __int_as_float_host(1));  // a.w != 0: room
            r.b = make_float4((float)room->hi[0], (float)room->hi[1], (float)room->hi[2], 0.0f);
            ids.a = make_float4(__int_as_float_host(room->face[0]), __int_as_float_host(room->face[1]), __int_as_float_host(room->face[2]),
                                __int_as_float_host(room->face[3]));
            ids.b = make_float4(__int_as_float_host(room->face[4]), __int_as_float_host(room->face[5]), 0.0f, 0.0f);
            recs.push_back(r);
            recs.push_back(ids);  // the face-id table rides in the next record slot (the box loop skips it)
        }
        g.box1 = (int)recs.size();
        groups.push_back(g);
    };
    auto add_groups_of = [&](const std::vector<int>& ids, int& first_group, int& end_group) {
        first_group = (int)groups.size();
        size_t i = 0;
        while (i < ids.size()) {
            // a run = consecutive primitives with one transform chain, cut so that a box is a run of its own
            size_t j = i;
            const bool box = is_box_at(ids, i, ids.size());
            if (box) {
                j = i + 6;
            } else {
                while (j < ids.size() && fs->prims[ids[j]].xf == fs->prims[ids[i]].xf && !is_box_at(ids, j, ids.size())) j++;
            }
            std::vector<int> members(ids.begin() + i, ids.begin() + j);
            Room room;
            room.n = 0;
            if (!box) find_room(members, room);
            if (room.n != 0) {
                std::vector<int> faces, others;
                for (int p : members) {
                    bool is_face = false;
                    for (int k = 0; k < 6; k++) is_face |= room.face[k] == p;
                    (is_face ? faces : others).push_back(p);
                }
                if (room.n > 0 && !others.empty()) emit_group(others, -1, nullptr);
                emit_group(faces, -1, &room);
                if (room.n < 0 && !others.empty()) emit_group(others, -1, nullptr);
            } else {
                emit_group(members, box ? ids[i] : -1, nullptr);
            }
            i = j;
        }
        end_group = (int)groups.size();
    };
    auto add_groups = [&](int first, int count, int& first_group, int& end_group) {
        std::vector<int> ids(count);
        for (int k = 0; k < count; k++) ids[k] = first + k;
        add_groups_of(ids, first_group, end_group);
    };

    // ---- bounding boxes (FP64, in the frame the primitive's own chain starts from), padded for FP32 slab tests -----------
    const double cam_t0 = fs->camera.time0, cam_t1 = fs->camera.time1;
    auto prim_box = [&](int p) {
        const rtw_flat_prim& pr = fs->prims[p];
        Box b;
        const double* v = pr.v;
        if (pr.type == RTW_PRIM_SPHERE) {
            double r = std::fabs(v[3]);
            for (int k = 0; k < 3; k++) b.grow_axis(k, v[k] - r, v[k] + r);
        } else if (pr.type == RTW_PRIM_MOVING_SPHERE) {  // sphere.h:22-25 at both ends of the shutter interval
            double r = std::fabs(v[3]), mt0 = v[7], mt1 = v[8];
            for (double tm : {cam_t0, cam_t1}) {
                double f = (tm - mt0) / (mt1 - mt0);
                for (int k = 0; k < 3; k++) {
                    double c = v[k] + f * (v[4 + k] - v[k]);
                    b.grow_axis(k, c - r, c + r);
                }
            }
        } else {
            const double e = 0.0001;  // hittable.h:169
            int ax[3];  // in-plane axes a, b and the normal axis
            if (pr.type == RTW_PRIM_XY_RECT) ax[0] = 0, ax[1] = 1, ax[2] = 2;
            else if (pr.type == RTW_PRIM_XZ_RECT) ax[0] = 0, ax[1] = 2, ax[2] = 1;
            else ax[0] = 1, ax[1] = 2, ax[2] = 0;
            b.grow_axis(ax[0], v[0], v[1]), b.grow_axis(ax[1], v[2], v[3]), b.grow_axis(ax[2], v[4] - e, v[4] + e);
        }
        if (pr.xf >= 0) {  // local = Ry(outer) + t  ->  outer = Ry^-1(local - t)
            double C, S, t[3];
            compose_chain_f64(fs, pr.xf, C, S, t);
            Box w;
            for (int c = 0; c < 8; c++) {
                double lx = ((c & 1) ? b.hi[0] : b.lo[0]) - t[0], ly = ((c & 2) ? b.hi[1] : b.lo[1]) - t[1],
                       lz = ((c & 4) ? b.hi[2] : b.lo[2]) - t[2];
                w.grow_point(C * lx + S * lz, ly, -S * lx + C * lz);
            }
            b = w;
        }
        return b;
    };
    auto padded = [](Box b) {
        for (int k = 0; k < 3; k++) {
            double m = std::max(std::fabs(b.lo[k]), std::fabs(b.hi[k]));
            double pad = 1e-3 + 2e-5 * m;
            b.lo[k] -= pad, b.hi[k] += pad;
        }
        return b;
    };

    // ---- traversal nodes ---------------------------------------------------------------------------------------------------
    std::vector<DevNode> nodes;
    std::vector<int2> leaves;
    std::vector<Box> leaf_box;
    const float kInf = std::numeric_limits<float>::infinity();
    auto set_child = [&](DevNode& n, bool right, const Box* b, int ref) {
        float4 lo = make_float4(kInf, kInf, kInf, 0.0f), hi = make_float4(-kInf, -kInf, -kInf, 0.0f);  // empty: never hit
        if (b) {
            // round outwards so the FP32 box contains the FP64 one
            lo = make_float4(std::nextafterf((float)b->lo[0], -kInf), std::nextafterf((float)b->lo[1], -kInf),
                             std::nextafterf((float)b->lo[2], -kInf), 0.0f);
            hi = make_float4(std::nextafterf((float)b->hi[0], kInf), std::nextafterf((float)b->hi[1], kInf),
                             std::nextafterf((float)b->hi[2], kInf), 0.0f);
        }
        if (right) {
            n.rlo = lo, n.rhi = hi;
            n.lhi.w = __int_as_float_host(ref);
        } else {
            n.llo.x = lo.x, n.llo.y = lo.y, n.llo.z = lo.z, n.llo.w = __int_as_float_host(ref);
            n.lhi.x = hi.x, n.lhi.y = hi.y, n.lhi.z = hi.z;
        }
    };
    auto exact_child = [&](DevNode& n, bool right, const double* bmin, const double* bmax, int ref) {
        // a reference bvh_node's own box, converted like every other scene constant (round to nearest)
        float4 lo = make_float4((float)bmin[0], (float)bmin[1], (float)bmin[2], 0.0f);
        float4 hi = make_float4((float)bmax[0], (float)bmax[1], (float)bmax[2], 0.0f);
        if (right) {
            n.rlo = lo, n.rhi = hi;
            n.lhi.w = __int_as_float_host(ref);
        } else {
            n.llo.x = lo.x, n.llo.y = lo.y, n.llo.z = lo.z, n.llo.w = __int_as_float_host(ref);
            n.lhi.x = hi.x, n.lhi.y = hi.y, n.lhi.z = hi.z;
        }
    };

    // (1) the reference's bvh_node trees, same topology: node i of the flat scene is device node i
    nodes.resize(fs->n_nodes);
    leaves.resize(fs->n_leaves);
    leaf_box.resize(fs->n_leaves);
    for (int i = 0; i < fs->n_leaves; i++) {
        add_groups(fs->leaves[i].first_prim, fs->leaves[i].prim_count, leaves[i].x, leaves[i].y);
        Box b;
        for (int p = 0; p < fs->leaves[i].prim_count; p++) b.merge(prim_box(fs->leaves[i].first_prim + p));
        leaf_box[i] = padded(b);
    }
    for (int i = 0; i < fs->n_nodes; i++) {
        const rtw_flat_node& n = fs->nodes[i];
        DevNode d{};
        for (int side = 0; side < 2; side++) {
            int ref = side ? n.right : n.left;
            if (side && n.right == n.left) {  // n == 1 (hittable.h:68-72): one visit is enough
                set_child(d, true, nullptr, ref);
            } else if (ref >= 0) {
                exact_child(d, side, fs->nodes[ref].bmin, fs->nodes[ref].bmax, ref);
            } else {
                set_child(d, side, &leaf_box[~ref], ref);
            }
        }
        nodes[i] = d;
    }

    // (2) trees built here (binned SAH).  Two uses, same builder over "objects" = (box, what the leaf holds):
    //   * a long hittable_list (hittable_list.h:11-37 walks it linearly): objects are its primitives, leaves hold up to 4;
    //     the closest hit is the same, the list's order-dependent tie rules only matter for bit-equal t of different prims;
    //   * the leaves of a reference bvh_node tree, for the RENDER path: the reference topology (random split axis) is kept in
    //     the flat scene and walked by intersect_batch; rendering walks a SAH tree over the very same leaf objects — one
    //     reference leaf per tree leaf, same leaf offsets — so the winner, including the right-most-leaf tie rule, is the same.
    struct Obj {
        Box box;   // unpadded
        int id;    // primitive index, or reference leaf index
    };
    int list_bvh_min = 32;
    if (const char* env = getenv("RTW_LIST_BVH_MIN")) list_bvh_min = atoi(env);
    bool rebuild_ref = true;
    if (const char* env = getenv("RTW_RENDER_REF_TOPOLOGY")) rebuild_ref = atoi(env) == 0;
    int max_depth = 0;
    auto node_box = [&](int ref) {
        if (ref < 0) return leaf_box[~ref];
        Box b;  // union of the node's two child boxes (already padded and rounded outwards)
        const DevNode& c = nodes[ref];
        b.grow_point(c.llo.x, c.llo.y, c.llo.z), b.grow_point(c.lhi.x, c.lhi.y, c.lhi.z);
        if (c.rlo.x <= c.rhi.x) b.grow_point(c.rlo.x, c.rlo.y, c.rlo.z), b.grow_point(c.rhi.x, c.rhi.y, c.rhi.z);
        return b;
    };
    std::function<int(std::vector<Obj>&, int, bool)> build_tree = [&](std::vector<Obj>& objs, int depth, bool prims) -> int {
        if (depth > max_depth) max_depth = depth;
        auto make_leaf = [&](std::vector<Obj>& lo) {
            if (!prims) return ~lo[0].id;  // an existing reference leaf
            std::vector<int> ids;
            Box b;
            for (auto& o : lo) ids.push_back(o.id), b.merge(o.box);
            std::sort(ids.begin(), ids.end());  // list order inside the leaf
            int2 lf;
            add_groups_of(ids, lf.x, lf.y);
            leaves.push_back(lf);
            leaf_box.push_back(padded(b));
            return ~(int)(leaves.size() - 1);
        };
        if (objs.size() == 1 || (prims && objs.size() <= 2)) return make_leaf(objs);
        // binned SAH on centroids (16 bins, 3 axes)
        Box cb;
        for (auto& o : objs) cb.grow_point(o.box.centre(0), o.box.centre(1), o.box.centre(2));
        double best_cost = 1e300;
        int best_axis = -1, best_bin = 0;
        const int NB = 16;
        for (int ax = 0; ax < 3; ax++) {
            double lo = cb.lo[ax], ext = cb.hi[ax] - cb.lo[ax];
            if (!(ext > 1e-12)) continue;
            Box bb[NB];
            int cnt[NB] = {};
            for (auto& o : objs) {
                int bi = std::min(NB - 1, (int)((o.box.centre(ax) - lo) / ext * NB));
                bb[bi].merge(o.box);
                cnt[bi]++;
            }
            Box lb[NB], rb[NB];
            int lc[NB], rc[NB];
            Box acc;
            int c = 0;
            for (int k = 0; k < NB; k++) acc.merge(bb[k]), c += cnt[k], lb[k] = acc, lc[k] = c;
            acc = Box(), c = 0;
            for (int k = NB - 1; k >= 0; k--) acc.merge(bb[k]), c += cnt[k], rb[k] = acc, rc[k] = c;
            for (int k = 0; k + 1 < NB; k++) {
                if (!lc[k] || !rc[k + 1]) continue;
                double cost = lb[k].area() * lc[k] + rb[k + 1].area() * rc[k + 1];
                if (cost < best_cost) best_cost = cost, best_axis = ax, best_bin = k;
            }
        }
        std::vector<Obj> L, R;
        if (best_axis >= 0) {
            double lo = cb.lo[best_axis], ext = cb.hi[best_axis] - cb.lo[best_axis];
            for (auto& o : objs) {
                int bi = std::min(NB - 1, (int)((o.box.centre(best_axis) - lo) / ext * NB));
                (bi <= best_bin ? L : R).push_back(o);
            }
        }
        if (L.empty() || R.empty()) {  // coincident centroids: split in order
            L.assign(objs.begin(), objs.begin() + objs.size() / 2);
            R.assign(objs.begin() + objs.size() / 2, objs.end());
        }
        // leaves of up to 4 primitives when splitting further does not pay (SAH with node cost ~ 2 primitive tests)
        if (prims && objs.size() <= 4) {
            Box all, lbx, rbx;
            for (auto& o : objs) all.merge(o.box);
            for (auto& o : L) lbx.merge(o.box);
            for (auto& o : R) rbx.merge(o.box);
            double split_cost = 2.0 + (lbx.area() * L.size() + rbx.area() * R.size()) / std::max(all.area(), 1e-300);
            if (split_cost >= (double)objs.size()) return make_leaf(objs);
        }
        int idx = (int)nodes.size();
        nodes.push_back(DevNode{});
        int lref = build_tree(L, depth + 1, prims), rref = build_tree(R, depth + 1, prims);
        DevNode d{};
        Box lb = node_box(lref), rb = node_box(rref);
        set_child(d, false, &lb, lref);
        set_child(d, true, &rb, rref);
        nodes[idx] = d;
        return idx;
    };
    // Device-side alternative (RTW_DEVICE_BVH=1, lbvh.cu): Morton codes, sort, Karras' radix tree and the box fit run on the GPU;
    // the nodes come back as (left, right, box) and are written as DevNode records here, so padding, traversal, stack needs and
    // the 4-wide collapse are those of the SAH build.  Leaves hold ONE object (a primitive, or a reference leaf).
    static const bool device_bvh = [] { const char* v = getenv("RTW_DEVICE_BVH"); return v && atoi(v) != 0; }();
    bool device_build_failed = false;
    std::function<int(std::vector<Obj>&, bool)> build_tree_device = [&](std::vector<Obj>& objs, bool prims) -> int {
        const int n = (int)objs.size();
        std::vector<int> leaf_ref(n);
        for (int i = 0; i < n; i++) {  // leaves first, in object (= list) order: the tie rule orders equal hits by leaf offset
            if (!prims) {
                leaf_ref[i] = ~objs[i].id;
                continue;
            }
            std::vector<int> ids{objs[i].id};
            int2 lf;
            add_groups_of(ids, lf.x, lf.y);
            leaves.push_back(lf);
            leaf_box.push_back(padded(objs[i].box));
            leaf_ref[i] = ~(int)(leaves.size() - 1);
        }
        if (n == 1) return leaf_ref[0];
        std::vector<float> boxes((size_t)n * 6);
        for (int i = 0; i < n; i++)
            for (int k = 0; k < 3; k++) boxes[6 * i + k] = (float)objs[i].box.lo[k], boxes[6 * i + 3 + k] = (float)objs[i].box.hi[k];
        std::vector<int> L, R;
        std::vector<float> nb;
        if (lbvh_build(boxes, L, R, nb) != cudaSuccess || (int)L.size() != n - 1) {
            device_build_failed = true;
            cudaGetLastError();
            return leaf_ref[0];
        }
        const int base = (int)nodes.size();
        nodes.resize(base + n - 1);
        std::function<int(int, int)> depth_of = [&](int ref, int d) -> int {
            if (ref < 0) return d;
            return std::max(depth_of(L[ref], d + 1), depth_of(R[ref], d + 1));
        };
        max_depth = std::max(max_depth, depth_of(0, 1));
        // boxes of the children live in the parent (DevNode); node boxes are unions of the padded leaf boxes, as in build_tree
        std::function<Box(int)> box_of = [&](int ref) -> Box {
            if (ref < 0) return leaf_box[~leaf_ref[~ref]];
            Box b = box_of(L[ref]);
            b.merge(box_of(R[ref]));
            return b;
        };
        for (int i = 0; i < n - 1; i++) {
            DevNode d{};
            const int lref = L[i] < 0 ? leaf_ref[~L[i]] : base + L[i], rref = R[i] < 0 ? leaf_ref[~R[i]] : base + R[i];
            Box lb = prims || L[i] >= 0 ? box_of(L[i]) : leaf_box[~leaf_ref[~L[i]]];
            Box rb = prims || R[i] >= 0 ? box_of(R[i]) : leaf_box[~leaf_ref[~R[i]]];
            set_child(d, false, &lb, lref);
            set_child(d, true, &rb, rref);
            nodes[base + i] = d;
        }
        (void)nb;  // the device's own boxes bound the UNPADDED objects; the render tree needs the padded leaf boxes (FP32 slab test)
        return base;
    };
    auto pseudo_root = [&](int root) {  // left child = the tree, right child empty
        DevNode pr{};
        Box rb = node_box(root);
        set_child(pr, false, &rb, root);
        set_child(pr, true, nullptr, root);
        nodes.push_back(pr);
        return (int)nodes.size() - 1;
    };
    std::function<void(int, std::vector<int>&)> collect_leaves = [&](int ref, std::vector<int>& out) {
        if (ref < 0) {
            out.push_back(~ref);
            return;
        }
        collect_leaves(fs->nodes[ref].left, out);
        if (fs->nodes[ref].right != fs->nodes[ref].left) collect_leaves(fs->nodes[ref].right, out);
    };

    std::vector<DevItem> items(fs->n_items);
    for (int i = 0; i < fs->n_items; i++) {
        const rtw_flat_item& it = fs->items[i];
        DevItem d{};
        d.type = it.type, d.xf = it.xf, d.first = it.first, d.end = it.first + it.count;
        d.material = it.material, d.flip = it.flip;
        d.density_eff = (float)(it.density * (it.multiplicity > 0 ? it.multiplicity : 1));
        if (it.type == RTW_ITEM_BVH) {
            feat |= F_BVH;
            int dep = bvh_depth(fs, it.first, 0);
            if (dep > max_depth) max_depth = dep;
            DevNode pr{};  // exact path: the reference's own root box
            exact_child(pr, false, fs->nodes[it.first].bmin, fs->nodes[it.first].bmax, it.first);
            set_child(pr, true, nullptr, it.first);
            d.first = (int)nodes.size();
            nodes.push_back(pr);
            d.end = d.first;
            if (rebuild_ref) {
                std::vector<int> lv;
                collect_leaves(it.first, lv);
                std::sort(lv.begin(), lv.end());
                lv.erase(std::unique(lv.begin(), lv.end()), lv.end());
                std::vector<Obj> objs;
                for (int l : lv) {
                    Obj o;
                    o.id = l;
                    for (int p = 0; p < fs->leaves[l].prim_count; p++) o.box.merge(prim_box(fs->leaves[l].first_prim + p));
                    objs.push_back(o);
                }
                d.end = pseudo_root(device_bvh ? build_tree_device(objs, false) : build_tree(objs, 1, false));
            }
        } else if (it.type == RTW_ITEM_PRIMS && list_bvh_min > 0 && it.count >= list_bvh_min) {
            feat |= F_BVH;
            std::vector<Obj> objs(it.count);
            for (int k = 0; k < it.count; k++) objs[k].id = it.first + k, objs[k].box = prim_box(it.first + k);
            d.type = RTW_ITEM_BVH;
            d.first = d.end = pseudo_root(device_bvh ? build_tree_device(objs, true) : build_tree(objs, 1, true));
        } else {
            const size_t rec0 = recs.size();
            add_groups(it.first, it.count, d.first, d.end);
            // constant_medium boundaries (hittable.h:438-449) take the FP32 sphere test whatever their radius: the two boundary
            // roots only clip the free-flight interval (an error of a few 1e-4 on a chord of 1e4 moves the scatter decision with
            // probability ~1e-7), the boundary's normal is never used (:469-472), and rtw_intersect_batch skips media — while
            // the FP64 test of the r = 5000 fog sphere ran twice per ray of the Book-2 final scene (442 M FP64 instructions per
            // 32-spp frame, profiles/r1_sorted_v9_book2_final_ncu_details.txt).  RTW_MEDIUM_FP64=1 restores it (A/B).
            static const bool medium_fp64 = [] { const char* v = getenv("RTW_MEDIUM_FP64"); return v && atoi(v) != 0; }();
            // a medium whose boundary is exactly one (non-moving) sphere: remember the record, closest_hit takes both boundary
            // roots from one quadratic (RTW_MEDIUM_ONE_SPHERE=0 switches this off: A/B)
            static const bool one_sphere_on = [] { const char* v = getenv("RTW_MEDIUM_ONE_SPHERE"); return !v || atoi(v) != 0; }();
            d.root4 = -1;
            if (it.type == RTW_ITEM_MEDIUM && !medium_fp64 && one_sphere_on && it.count == 1 && fs->prims[it.first].type == RTW_PRIM_SPHERE &&
                fs->prims[it.first].xf < 0 && recs.size() == rec0 + 1)
                d.root4 = -2 - (int)rec0;  // record index, turned into a byte offset with the others below
            if (it.type == RTW_ITEM_MEDIUM && !medium_fp64)
                for (size_t r = rec0; r < recs.size(); r++) {
                    int id;
                    memcpy(&id, &recs[r].b.w, 4);
                    if (id & REC_PRECISE) {
                        id &= ~REC_PRECISE;
                        memcpy(&recs[r].b.w, &id, 4);
                        recs[r].b.x = 0.0f;
                    }
                }
        }
        if (it.type == RTW_ITEM_MEDIUM) feat |= F_MEDIA;
        items[i] = d;
    }
    max_depth += 1;  // pseudo-root

    // (3) the 4-wide render trees (DevNode4): every tree the render kernels walk (`DevItem.end`) is collapsed — a child that is an
    // inner node is replaced by its own two children, largest box first, until the node holds four — so a ray fetches about
    // half as many nodes, each one cache line.  Leaves (and with them the right-most-leaf tie rule) are untouched.
    std::vector<DevNode4> nodes4;
    int stack4_need = 0;
    {
        struct Child {
            int ref;
            float lo[3], hi[3];
        };
        auto children_of = [&](int b, std::vector<Child>& out, size_t at) {
            const DevNode& n = nodes[b];
            int l, r;
            memcpy(&l, &n.llo.w, 4), memcpy(&r, &n.lhi.w, 4);
            std::vector<Child> cs;
            cs.push_back(Child{l, {n.llo.x, n.llo.y, n.llo.z}, {n.lhi.x, n.lhi.y, n.lhi.z}});
            if (n.rlo.x <= n.rhi.x) cs.push_back(Child{r, {n.rlo.x, n.rlo.y, n.rlo.z}, {n.rhi.x, n.rhi.y, n.rhi.z}});
            for (Child& c : cs)
                if (b < fs->n_nodes && c.ref >= 0)  // a reference bvh_node's own (unpadded) box: pad it like the boxes built here
                    for (int k = 0; k < 3; k++) {
                        float m = std::max(std::fabs(c.lo[k]), std::fabs(c.hi[k])), pad = 1e-3f + 2e-5f * m;
                        c.lo[k] = std::nextafterf(c.lo[k] - pad, -kInf), c.hi[k] = std::nextafterf(c.hi[k] + pad, kInf);
                    }
            out.insert(out.begin() + at, cs.begin(), cs.end());
        };
        const int empty_leaf = (int)leaves.size();  // first == end: no groups
        if (feat & F_BVH) leaves.push_back(int2{0, 0});
        std::function<int(int, int&)> collapse = [&](int b, int& need) -> int {
            std::vector<Child> cs;
            children_of(b, cs, 0);
            while (cs.size() < 4) {
                int pick = -1;
                double pick_area = -1.0;
                for (size_t i = 0; i < cs.size(); i++) {
                    if (cs[i].ref < 0) continue;
                    double x = (double)cs[i].hi[0] - cs[i].lo[0], y = (double)cs[i].hi[1] - cs[i].lo[1], z = (double)cs[i].hi[2] - cs[i].lo[2];
                    double a = x * y + y * z + z * x;
                    if (a > pick_area) pick_area = a, pick = (int)i;
                }
                if (pick < 0) break;
                int inner = cs[pick].ref;
                cs.erase(cs.begin() + pick);
                children_of(inner, cs, (size_t)pick);
            }
            const int idx = (int)nodes4.size();
            nodes4.push_back(DevNode4{});
            DevNode4 d{};
            float* lo[3] = {&d.lox.x, &d.loy.x, &d.loz.x};
            float* hi[3] = {&d.hix.x, &d.hiy.x, &d.hiz.x};
            int deepest = 0;
            for (int i = 0; i < 4; i++) {
                if (i < (int)cs.size()) {
                    for (int k = 0; k < 3; k++) lo[k][i] = cs[i].lo[k], hi[k][i] = cs[i].hi[k];
                    if (cs[i].ref >= 0) {
                        int sub = 0;
                        d.ref[i] = collapse(cs[i].ref, sub);
                        deepest = std::max(deepest, sub);
                    } else {
                        d.ref[i] = cs[i].ref;
                    }
                } else {
                    for (int k = 0; k < 3; k++) lo[k][i] = hi[k][i] = 3.0e38f;
                    d.ref[i] = ~empty_leaf;
                }
            }
            need = (int)cs.size() - 1 + deepest;  // siblings pushed here + the deepest subtree's own need
            nodes4[idx] = d;
            return idx;
        };
        for (DevItem& d : items)
            if (d.type == RTW_ITEM_BVH) {
                int need = 0;
                d.root4 = collapse(d.end, need);
                stack4_need = std::max(stack4_need, need);
            }
    }
    // Exact traversal-stack needs of the binary trees: a node pushes one reference when both of its children are hit, so a
    // tree needs as many entries as the most two-child nodes on any root-to-leaf path.  The render kernels of the binary
    // variants walk all tree items in one loop (bvh_hit_items): the k-th tree item of the world is walked with k-1 ENTER
    // markers still below it.  Shared memory per block decides whether three blocks fit the 196 KB carve-out (60 KB of L1)
    // or need the 228 KB one (28 KB): on the Book-2 final scene that is 0.473 vs 0.434 Gpaths/s, so no slack entries.
    std::function<int(int)> push_need = [&](int ref) -> int {
        if (ref < 0) return 0;
        const DevNode& n = nodes[ref];
        int l, r;
        memcpy(&l, &n.llo.w, 4), memcpy(&r, &n.lhi.w, 4);
        const bool two = n.rlo.x <= n.rhi.x;
        const int nl = push_need(l);
        return two ? 1 + std::max(nl, push_need(r)) : nl;
    };
    int stack_exact = 0, stack_render = 0;
    {
        int k = 0;
        for (int i = 0; i < fs->n_world_items; i++)
            if (items[i].type == RTW_ITEM_BVH) {
                stack_exact = std::max(stack_exact, push_need(items[i].first));
                stack_render = std::max(stack_render, push_need(items[i].end) + k);
                k++;
            }
        stack_render = std::max(stack_render, k);  // all markers are pushed before the first is popped
    }
    if (device_build_failed) {
        set_error("scene_create: the device-side tree build (RTW_DEVICE_BVH=1) failed");
        return nullptr;
    }
    if (getenv("RTW_DEBUG_TREES"))
        fprintf(stderr, "rtw trees: %zu binary nodes (depth %d; stack need: exact %d, render %d), %zu 4-wide nodes (stack need %d), %zu leaves\n",
                nodes.size(), max_depth, stack_exact, stack_render, nodes4.size(), stack4_need, leaves.size());
    if (dry_run) {
        set_error("scene_create: RTW_DRY_RUN is set (diagnostics only, nothing was uploaded)");
        return nullptr;
    }
    if (max_depth > 60 || stack4_need > 60) {
        set_error("scene_create: bvh deeper than the 62-entry traversal stack");
        return nullptr;
    }

    std::vector<DevTex> texs(fs->n_textures);
    for (int i = 0; i < fs->n_textures; i++) {
        const rtw_flat_texture& t = fs->textures[i];
        DevTex d{};
        d.r = (float)t.color[0], d.g = (float)t.color[1], d.b = (float)t.color[2], d.scale = (float)t.scale;
        d.type = t.type, d.even = t.even, d.odd = t.odd, d.image_offset = t.image_offset, d.nx = t.nx, d.ny = t.ny;
        texs[i] = d;
    }
    std::vector<DevMat> mats(fs->n_materials);
    for (int i = 0; i < fs->n_materials; i++) {
        const rtw_flat_material& m = fs->materials[i];
        DevMat d{};
        d.type = m.type, d.texture = m.texture;
        if (m.type == RTW_MAT_METAL) {
            d.r = (float)m.albedo[0], d.g = (float)m.albedo[1], d.b = (float)m.albedo[2], d.param = (float)m.fuzz;
            d.tex_is_constant = 1;
        } else if (m.type == RTW_MAT_DIELECTRIC) {
            d.param = (float)m.ref_idx;
            d.tex_is_constant = 1;
        } else if (m.texture >= 0 && fs->textures[m.texture].type == RTW_TEX_CONSTANT) {
            const rtw_flat_texture& t = fs->textures[m.texture];
            d.r = (float)t.color[0], d.g = (float)t.color[1], d.b = (float)t.color[2];
            d.tex_is_constant = 1;
        } else {
            feat |= F_TEXTURES;
        }
        mats[i] = d;
    }
    std::vector<DevLight> lights(fs->n_lights);
    int n_nee_lights = 0;
    for (int i = 0; i < fs->n_lights; i++) {
        const rtw_flat_light& l = fs->lights[i];
        DevLight d{};
        d.type = l.type;
        d.a0 = (float)l.v[0], d.a1 = (float)l.v[1], d.a2 = (float)l.v[2], d.a3 = (float)l.v[3];
        if (l.type == RTW_LIGHT_XZ_RECT) {
            d.k = (float)l.v[4];
            d.area = (float)((l.v[1] - l.v[0]) * (l.v[3] - l.v[2]));  // hittable.h:217
            // the NEE set (RTW_FLAG_NEE): lights whose shape is a world-space xz_rect of the scene with an emitting material
            for (int p = 0; p < fs->n_prims && !d.pad; p++) {
                const rtw_flat_prim& q = fs->prims[p];
                if (q.type == RTW_PRIM_XZ_RECT && q.xf < 0 && q.v[0] == l.v[0] && q.v[1] == l.v[1] && q.v[2] == l.v[2] && q.v[3] == l.v[3] &&
                    q.v[4] == l.v[4] && q.material >= 0 && fs->materials[q.material].type == RTW_MAT_DIFFUSE_LIGHT)
                    d.pad = 1;
            }
            n_nee_lights += d.pad;
        }
        lights[i] = d;
    }
    std::vector<float4> ranvec;
    std::vector<uint8_t> perm;
    if (fs->perlin_ranvec) {
        for (int i = 0; i < 256; i++)
            ranvec.push_back(make_float4((float)fs->perlin_ranvec[3 * i], (float)fs->perlin_ranvec[3 * i + 1],
                                         (float)fs->perlin_ranvec[3 * i + 2], 0.0f));
        for (int i = 0; i < 768; i++) perm.push_back((uint8_t)fs->perlin_perm[i]);
    }
    std::vector<uint8_t> texels;
    if (fs->n_texel_bytes) texels.assign(fs->texels, fs->texels + fs->n_texel_bytes);

    // blob layout: 16-byte aligned sections, hot first; offsets first (the records hold byte offsets), then the bytes
    const size_t sec_bytes[SEC_COUNT] = {
        items.size() * sizeof(DevItem), groups.size() * sizeof(DevGroup), recs.size() * sizeof(DevRec), xfs.size() * sizeof(DevXf),
        lights.size() * sizeof(DevLight), nodes.size() * sizeof(DevNode), leaves.size() * sizeof(int2), nodes4.size() * sizeof(DevNode4),
        pm.size() * sizeof(uint32_t), pa.size() * sizeof(float4), pb.size() * sizeof(float4), pd.size() * sizeof(double),
        mats.size() * sizeof(DevMat), texs.size() * sizeof(DevTex), ranvec.size() * sizeof(float4), perm.size()};
    uint32_t off[SEC_COUNT] = {};
    uint32_t hot_bytes = 0;
    size_t total = 0;
    for (int k = 0; k < SEC_COUNT; k++) {
        if (k == SEC_NODES4) total = (total + 127) & ~(size_t)127;  // a DevNode4 is one 128-byte line
        off[k] = (uint32_t)total;
        total = (total + sec_bytes[k] + 15) & ~(size_t)15;
        if (k == SEC_NODES4) hot_bytes = (uint32_t)total;
    }
    if (total == 0) total = 16;
    if (total > 0x3fffffffu) {  // leaf references ~offset stay above -2^30; below that are the traversal's markers
        set_error("scene_create: scene too large for 30-bit blob offsets");
        return nullptr;
    }
    // indices -> byte offsets
    auto xf_off = [&](int idx) { return idx < 0 ? -1 : (int32_t)(off[SEC_XFS] + (uint32_t)idx * sizeof(DevXf)); };
    auto rec_off = [&](int idx) { return (int32_t)(off[SEC_RECS] + (uint32_t)idx * sizeof(DevRec)); };
    auto group_off = [&](int idx) { return (int32_t)(off[SEC_GROUPS] + (uint32_t)idx * sizeof(DevGroup)); };
    auto ref_off = [&](int ref) {
        return ref >= 0 ? (int32_t)(off[SEC_NODES] + (uint32_t)ref * sizeof(DevNode)) : ~(int32_t)(off[SEC_LEAVES] + (uint32_t)(~ref) * sizeof(int2));
    };
    for (DevGroup& g : groups) {
        g.xf = xf_off(g.xf);
        g.sph0 = rec_off(g.sph0), g.sph1 = rec_off(g.sph1), g.msph1 = rec_off(g.msph1);
        g.xy1 = rec_off(g.xy1), g.xz1 = rec_off(g.xz1), g.yz1 = rec_off(g.yz1), g.box1 = rec_off(g.box1);
    }
    for (DevItem& d : items) {
        d.xf = xf_off(d.xf);
        if (d.type == RTW_ITEM_BVH) {
            d.first = ref_off(d.first), d.end = ref_off(d.end), d.root4 = (int32_t)(off[SEC_NODES4] + (uint32_t)d.root4 * sizeof(DevNode4));
        } else {
            d.first = group_off(d.first), d.end = group_off(d.end);
            d.root4 = (d.type == RTW_ITEM_MEDIUM && d.root4 <= -2) ? rec_off(-2 - d.root4) : -1;
        }
    }
    for (DevNode& n : nodes) {
        int l, r;
        memcpy(&l, &n.llo.w, 4), memcpy(&r, &n.lhi.w, 4);
        n.llo.w = __int_as_float_host(ref_off(l)), n.lhi.w = __int_as_float_host(ref_off(r));
    }
    for (int2& lf : leaves) lf.x = group_off(lf.x), lf.y = group_off(lf.y);
    for (DevNode4& n : nodes4)
        for (int i = 0; i < 4; i++)
            n.ref[i] = n.ref[i] >= 0 ? (int32_t)(off[SEC_NODES4] + (uint32_t)n.ref[i] * sizeof(DevNode4)) : ref_off(n.ref[i]);
    auto pd_off = [&](float idx_bits) {
        int idx;
        memcpy(&idx, &idx_bits, 4);
        return __int_as_float_host((int32_t)(off[SEC_PD] + (uint32_t)idx * 4u * sizeof(double)));
    };
    for (int i = 0; i < fs->n_prims; i++)
        if (pm[i] & PM_PRECISE) pb[i].w = pd_off(pb[i].w);
    for (DevRec& r : recs) {
        int id;
        memcpy(&id, &r.b.w, 4);
        if (id & REC_PRECISE) r.b.x = pd_off(r.b.x);
    }

    std::vector<uint8_t> blob(total, 0);
    const void* sec_data[SEC_COUNT] = {items.data(), groups.data(), recs.data(), xfs.data(), lights.data(), nodes.data(), leaves.data(), nodes4.data(),
                                       pm.data(), pa.data(), pb.data(), pd.data(), mats.data(), texs.data(), ranvec.data(), perm.data()};
    for (int k = 0; k < SEC_COUNT; k++)
        if (sec_bytes[k]) memcpy(blob.data() + off[k], sec_data[k], sec_bytes[k]);

    rtw_scene* s = new (std::nothrow) rtw_scene();
    if (!s) return nullptr;
    s->device = device;
    cudaError_t err[] = {s->blob.upload(blob), s->texels.upload(texels), dev_alloc((void**)&s->d_counter, 2 * sizeof(unsigned long long)),  // [0] rays, [1] render_sorted's watchdog flag
                         cudaEventCreate(&s->ev0), cudaEventCreate(&s->ev1)};
    for (cudaError_t x : err)
        if (x != cudaSuccess) {
            cuda_fail(x, "scene upload");
            s->release();
            delete s;
            return nullptr;
        }
    DevScene& d = s->dev;
    d.blob = (const uint8_t*)s->blob.p, d.texels = (const uint8_t*)s->texels.p;
    for (int k = 0; k < SEC_COUNT; k++) d.off[k] = off[k];
    d.hot_bytes = hot_bytes, d.all_bytes = (uint32_t)blob.size();
    d.n_world_items = fs->n_world_items, d.n_items = fs->n_items, d.n_prims = fs->n_prims, d.n_lights = fs->n_lights;
    d.render_type = fs->render_type, d.background_type = fs->background_type;
    d.features = feat, d.stack_exact = stack_exact, d.stack_render = stack_render, d.bvh4_stack = stack4_need;
    d.n_nee_lights = n_nee_lights;
    d.n_bvh_items = 0;
    for (int i = 0; i < fs->n_world_items; i++) d.n_bvh_items += items[i].type == RTW_ITEM_BVH;
    // RTW_FLAG_AUTO: the multi-kernel wavefront pipeline beat the lane-owns-a-path megakernel on noise-textured and on
    // media + BVH scenes; the material-sorted megakernel is faster than both on every measured class, so
    // AUTO only falls back to the wavefront rule when the sorted kernel is switched off.
    {
        const char* v = getenv("RTW_MEGA_SORTED");
        if (v && atoi(v) == 0) {
            for (int i = 0; i < fs->n_textures; i++)
                if (fs->textures[i].type == RTW_TEX_NOISE) s->prefer_wavefront = true;
            if ((feat & F_MEDIA) && fs->n_nodes > 0) s->prefer_wavefront = true;
        }
    }
    const rtw_flat_camera& c = fs->camera;
    DevCamera& k = d.cam;
    k.ox = (float)c.origin[0], k.oy = (float)c.origin[1], k.oz = (float)c.origin[2];
    k.llx = (float)c.lower_left_corner[0], k.lly = (float)c.lower_left_corner[1], k.llz = (float)c.lower_left_corner[2];
    k.hx = (float)c.horizontal[0], k.hy = (float)c.horizontal[1], k.hz = (float)c.horizontal[2];
    k.vx = (float)c.vertical[0], k.vy = (float)c.vertical[1], k.vz = (float)c.vertical[2];
    k.ux = (float)c.u[0], k.uy = (float)c.u[1], k.uz = (float)c.u[2];
    k.wx = (float)c.v[0], k.wy = (float)c.v[1], k.wz = (float)c.v[2];
    k.time0 = (float)c.time0, k.time1 = (float)c.time1, k.lens_radius = (float)c.lens_radius;
    return s;
}

extern "C" void rtw_scene_destroy(rtw_scene* s) {
    if (!s) return;
    cudaSetDevice(s->device);
    s->release();
    delete s;
}

// ------------------------------------------------------------------------------------------------ hot path
namespace {

int check_params(const rtw_scene* s, const rtw_render_params* p) {
    if (!s || !p) {
        set_error("render: null scene or params");
        return RTW_ERR_INVALID;
    }
    if (p->nx <= 0 || p->ny <= 0 || p->spp_count <= 0 || p->spp_begin < 0 || p->max_depth < 0 ||
        (long long)p->nx * p->ny > 0x7fffffffLL || p->spp_count > (1 << 26)) {  // a warp's pool holds 32*spp items
        set_error("render: bad nx/ny/spp/max_depth (spp_count per launch <= 2^26)");
        return RTW_ERR_INVALID;
    }
    return RTW_OK;
}

}  // namespace

extern "C" int rtw_render_device(rtw_scene* s, const rtw_render_params* p, float* accum_dev, void* cuda_stream) {
    int rc = check_params(s, p);
    if (rc != RTW_OK) return rc;
    if (!accum_dev) {
        set_error("render: null accumulator");
        return RTW_ERR_INVALID;
    }
    if (s->pending) {  // the ray counter and the two timing events live in the scene: one render in flight per rtw_scene
        set_error("render: the previous rtw_render_device of this scene has not been finished (rtw_render_device_finish)");
        return RTW_ERR_INVALID;
    }
    cudaStream_t st = (cudaStream_t)cuda_stream;
    cudaError_t e = cudaSetDevice(s->device);
    if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    DevRenderParams rp{};
    rp.nx = p->nx, rp.ny = p->ny, rp.spp_begin = p->spp_begin, rp.spp_count = p->spp_count, rp.max_depth = p->max_depth;
    rp.flags = p->flags;
    rp.seed_lo = (uint32_t)(p->seed & 0xffffffffu), rp.seed_hi = (uint32_t)(p->seed >> 32);
    philox_round_keys(rp);
    if ((e = cudaMemsetAsync(s->d_counter, 0, 2 * sizeof(unsigned long long), st)) != cudaSuccess) return cuda_fail(e, "memset");
    if ((e = cudaEventRecord(s->ev0, st)) != cudaSuccess) return cuda_fail(e, "event");
    int launches = 1;
    const bool nee = (p->flags & RTW_FLAG_NEE) != 0;  // the shadow stage lives in the wavefront pipeline
    bool wavefront = (p->flags & RTW_FLAG_WAVEFRONT) != 0 || nee;
    if ((p->flags & RTW_FLAG_AUTO) && !wavefront) {
        // measured on B200 (profiles/README.md): the sorted shade kernel pays off when part of each warp would evaluate
        // 7-octave Perlin turbulence, or walk big trees and media together; Cornell-class scenes (constant or checker
        // textures, smoke in small boxes, long sphere lists) run faster in the megakernel
        wavefront = s->prefer_wavefront && p->max_depth <= 255 && (long long)p->spp_begin + p->spp_count < (1 << 24);
    }
    if (wavefront) {
        // pool of path slots: L2-resident by default (1 Mi slots = 64 MiB), never larger than the work
        if (p->max_depth > 255 || (long long)p->spp_begin + p->spp_count >= (1 << 24)) {
            set_error("render: the wavefront pipeline packs (sample, bounce) into 24+8 bits: max_depth <= 255, samples < 2^24");
            return RTW_ERR_INVALID;
        }
        unsigned long long want = 1ull << 20;
        if (const char* env = getenv("RTW_WF_SLOTS")) want = strtoull(env, nullptr, 10);
        unsigned long long paths = (unsigned long long)p->nx * p->ny * p->spp_count;
        if (want > paths) want = paths;
        want = (want + 255) & ~255ull;
        if (want < 256) want = 256;
        const size_t want_bytes = wavefront_scratch_bytes((uint32_t)want, nee);
        if ((uint32_t)want != s->wf_slots || want_bytes > s->wf_bytes) {
            if (s->wf_scratch) cudaStreamSynchronize(st);  // a previous render on this stream may still use it
            dev_free(s->wf_scratch);
            s->wf_scratch = nullptr, s->wf_slots = 0, s->wf_bytes = 0;
            if ((e = dev_alloc((void**)&s->wf_scratch, want_bytes)) != cudaSuccess) return cuda_fail(e, "cudaMalloc wavefront pool");
            s->wf_slots = (uint32_t)want, s->wf_bytes = want_bytes;
        }
        if (!s->wf_host_flag && (e = cudaMallocHost((void**)&s->wf_host_flag, 16)) != cudaSuccess) return cuda_fail(e, "cudaMallocHost");
        if ((e = launch_wavefront(s->dev, rp, (float4*)accum_dev, s->d_counter, s->wf_scratch, s->wf_slots, s->wf_host_flag, &launches, st)) !=
            cudaSuccess)
            return cuda_fail(e, "wavefront launch");
    } else {
        // the material-sorted megakernel (sorted.cu) is the default; RTW_MEGA_SORTED=0 selects the warp-pool kernel
        static const bool use_sorted = [] { const char* v = getenv("RTW_MEGA_SORTED"); return !v || atoi(v) != 0; }();
        if (use_sorted && sorted_supports(rp)) e = launch_render_sorted(s->dev, rp, (float4*)accum_dev, s->d_counter, st);
        else e = launch_render(s->dev, rp, (float4*)accum_dev, s->d_counter, st);
        if (e != cudaSuccess) return cuda_fail(e, "render launch");
    }
    if ((e = cudaEventRecord(s->ev1, st)) != cudaSuccess) return cuda_fail(e, "event");
    s->rendered = true;
    s->pending = true;
    s->pending_params = *p;
    s->pending_launches = launches;
    return RTW_OK;
}

extern "C" int rtw_render_device_finish(rtw_scene* s, rtw_render_stats* stats) {
    if (!s || !s->pending) {
        set_error("render_device_finish: no pending render");
        return RTW_ERR_INVALID;
    }
    cudaError_t e = cudaSetDevice(s->device);
    if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    if ((e = cudaEventSynchronize(s->ev1)) != cudaSuccess) return cuda_fail(e, "render kernel");
    s->pending = false;
    unsigned long long counters[2] = {0, 0};  // [0] color() invocations, [1] set by render_sorted's watchdog (sorted_kernel.cuh)
    if ((e = cudaMemcpy(counters, s->d_counter, sizeof counters, cudaMemcpyDeviceToHost)) != cudaSuccess) return cuda_fail(e, "counter");
    if (counters[1] != 0) {
        set_error("render: a block of the sorted kernel found no work for over a second with paths still in flight and gave up (path slot lost); the image is incomplete");
        return RTW_ERR_CUDA;
    }
    if (stats) {
        const unsigned long long rays = counters[0];
        float ms = 0;
        cudaEventElapsedTime(&ms, s->ev0, s->ev1);
        const rtw_render_params& p = s->pending_params;
        stats->paths = (uint64_t)p.nx * p.ny * p.spp_count;
        stats->rays = rays;
        stats->kernel_ms = ms;
        stats->kernel_launches = s->pending_launches;
        stats->pad_ = 0;
    }
    return RTW_OK;
}

extern "C" int rtw_render(rtw_scene* s, const rtw_render_params* p, float* accum_host, rtw_render_stats* stats) {
    int rc = check_params(s, p);
    if (rc != RTW_OK) return rc;
    if (!accum_host) {
        set_error("render: null accumulator");
        return RTW_ERR_INVALID;
    }
    cudaError_t e = cudaSetDevice(s->device);
    if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    size_t pixels = (size_t)p->nx * p->ny;
    if (pixels > s->scratch_pixels) {
        dev_free(s->scratch_accum);
        s->scratch_accum = nullptr;
        s->scratch_pixels = 0;
        if ((e = dev_alloc((void**)&s->scratch_accum, pixels * 4 * sizeof(float))) != cudaSuccess) return cuda_fail(e, "cudaMalloc accum");
        s->scratch_pixels = pixels;
    }
    if (p->flags & RTW_FLAG_ACCUMULATE)
        if ((e = cudaMemcpy(s->scratch_accum, accum_host, pixels * 16, cudaMemcpyHostToDevice)) != cudaSuccess) return cuda_fail(e, "H2D accum");
    rc = rtw_render_device(s, p, s->scratch_accum, nullptr);
    if (rc != RTW_OK) return rc;
    rc = rtw_render_device_finish(s, stats);
    if (rc != RTW_OK) return rc;
    if ((e = cudaMemcpy(accum_host, s->scratch_accum, pixels * 16, cudaMemcpyDeviceToHost)) != cudaSuccess) return cuda_fail(e, "D2H accum");
    return RTW_OK;
}

// No-allocation variant of the wavefront render (no counterpart in the reference):
// scratch = the library's static device array (wavefront.cu), nothing is allocated or freed by this call.
extern "C" int rtw_render_device_noalloc(rtw_scene* s, const rtw_render_params* p, float* accum_dev, uint32_t pool_slots, void* cuda_stream,
                                         rtw_render_stats* stats) {
    int rc = check_params(s, p);
    if (rc != RTW_OK) return rc;
    if (!accum_dev) {
        set_error("render_device_noalloc: null accumulator");
        return RTW_ERR_INVALID;
    }
    if (s->pending) {
        set_error("render_device_noalloc: the previous rtw_render_device of this scene has not been finished");
        return RTW_ERR_INVALID;
    }
    const uint32_t max_slots = wavefront_noalloc_max_slots();
    if (pool_slots > max_slots) {
        set_error("render_device_noalloc: pool_slots = " + std::to_string(pool_slots) + " exceeds RTW_NOALLOC_MAX_POOL_SLOTS = " +
                  std::to_string(max_slots) + " (the static device scratch holds no more; use rtw_render_device for bigger pools)");
        return RTW_ERR_INVALID;
    }
    if (p->flags & RTW_FLAG_NEE) {
        set_error("render_device_noalloc: RTW_FLAG_NEE needs the shadow-ray queue, which the static scratch does not hold");
        return RTW_ERR_INVALID;
    }
    if (p->max_depth > 255 || (long long)p->spp_begin + p->spp_count >= (1 << 24)) {
        set_error("render_device_noalloc: the wavefront pipeline packs (sample, bounce) into 24+8 bits: max_depth <= 255, samples < 2^24");
        return RTW_ERR_INVALID;
    }
    cudaStream_t st = (cudaStream_t)cuda_stream;
    cudaError_t e = cudaSetDevice(s->device);
    if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    DevRenderParams rp{};
    rp.nx = p->nx, rp.ny = p->ny, rp.spp_begin = p->spp_begin, rp.spp_count = p->spp_count, rp.max_depth = p->max_depth;
    rp.flags = p->flags;
    rp.seed_lo = (uint32_t)(p->seed & 0xffffffffu), rp.seed_hi = (uint32_t)(p->seed >> 32);
    philox_round_keys(rp);
    unsigned long long want = pool_slots ? pool_slots : max_slots;
    const unsigned long long paths = (unsigned long long)p->nx * p->ny * p->spp_count;
    if (want > paths) want = paths;        // never more slots than paths
    want = (want + 255) & ~255ull;         // whole blocks, like rtw_render_device
    if (want > max_slots) want = max_slots;
    unsigned long long rays = 0;
    int launches = 0;
    if ((e = cudaEventRecord(s->ev0, st)) != cudaSuccess) return cuda_fail(e, "event");
    if ((e = launch_wavefront_noalloc(s->dev, rp, (float4*)accum_dev, (uint32_t)want, &rays, &launches, st)) != cudaSuccess)
        return cuda_fail(e, "wavefront (static scratch) launch");
    if ((e = cudaEventRecord(s->ev1, st)) != cudaSuccess) return cuda_fail(e, "event");
    if ((e = cudaEventSynchronize(s->ev1)) != cudaSuccess) return cuda_fail(e, "render kernels");
    s->rendered = true;
    if (stats) {
        float ms = 0;
        cudaEventElapsedTime(&ms, s->ev0, s->ev1);
        stats->paths = paths, stats->rays = rays, stats->kernel_ms = ms, stats->kernel_launches = launches, stats->pad_ = 0;
    }
    return RTW_OK;
}

extern "C" int rtw_intersect_batch_device(rtw_scene* s, const rtw_ray* rays, rtw_hit* hits, int64_t n, float tmin, float tmax,
                                          void* cuda_stream) {
    if (!s || n < 0 || (n > 0 && (!rays || !hits))) {
        set_error("intersect_batch: bad arguments");
        return RTW_ERR_INVALID;
    }
    if (n == 0) return RTW_OK;
    cudaError_t e = cudaSetDevice(s->device);
    if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    if ((e = launch_intersect(s->dev, rays, hits, n, tmin, tmax, (cudaStream_t)cuda_stream)) != cudaSuccess)
        return cuda_fail(e, "intersect launch");
    return RTW_OK;
}

extern "C" int rtw_intersect_batch(rtw_scene* s, const rtw_ray* rays, rtw_hit* hits, int64_t n, float tmin, float tmax) {
    if (!s || n < 0 || (n > 0 && (!rays || !hits))) {
        set_error("intersect_batch: bad arguments");
        return RTW_ERR_INVALID;
    }
    if (n == 0) return RTW_OK;
    cudaError_t e = cudaSetDevice(s->device);
    if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    rtw_ray* d_rays = nullptr;
    rtw_hit* d_hits = nullptr;
    int rc = RTW_OK;
    if ((e = dev_alloc((void**)&d_rays, n * sizeof(rtw_ray))) != cudaSuccess || (e = dev_alloc((void**)&d_hits, n * sizeof(rtw_hit))) != cudaSuccess ||
        (e = cudaMemcpy(d_rays, rays, n * sizeof(rtw_ray), cudaMemcpyHostToDevice)) != cudaSuccess) {
        rc = cuda_fail(e, "intersect_batch buffers");
    } else if ((rc = rtw_intersect_batch_device(s, d_rays, d_hits, n, tmin, tmax, nullptr)) == RTW_OK) {
        if ((e = cudaMemcpy(hits, d_hits, n * sizeof(rtw_hit), cudaMemcpyDeviceToHost)) != cudaSuccess) rc = cuda_fail(e, "intersect_batch D2H");
    }
    dev_free(d_rays);
    dev_free(d_hits);
    return rc;
}

extern "C" int rtw_measure_fp32_peak(int device, double* tflops) {
    if (!tflops) {
        set_error("measure_fp32_peak: null");
        return RTW_ERR_INVALID;
    }
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    if ((e = measure_fp32_peak(tflops)) != cudaSuccess) return cuda_fail(e, "fp32 peak probe");
    return RTW_OK;
}

// ------------------------------------------------------------------------------------------------ known-answer hooks
extern "C" int rtw_philox4x32_10_batch(int device, const uint32_t* counters, int64_t n, uint64_t key, uint32_t* out) {
    if (n < 0 || (n > 0 && (!counters || !out))) {
        set_error("philox batch: bad arguments");
        return RTW_ERR_INVALID;
    }
    if (n == 0) return RTW_OK;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    DevRenderParams rp{};
    rp.seed_lo = (uint32_t)key, rp.seed_hi = (uint32_t)(key >> 32);
    philox_round_keys(rp);
    uint4 *dc = nullptr, *dout = nullptr;
    int rc = RTW_OK;
    if ((e = dev_alloc((void**)&dc, n * 16)) != cudaSuccess || (e = dev_alloc((void**)&dout, n * 16)) != cudaSuccess ||
        (e = cudaMemcpy(dc, counters, n * 16, cudaMemcpyHostToDevice)) != cudaSuccess || (e = launch_philox(dc, dout, n, rp)) != cudaSuccess ||
        (e = cudaMemcpy(out, dout, n * 16, cudaMemcpyDeviceToHost)) != cudaSuccess)
        rc = cuda_fail(e, "philox batch");
    dev_free(dc);
    dev_free(dout);
    return rc;
}

extern "C" int rtw_sampler_batch(int device, int which, int64_t n, uint64_t seed, float* out_xyz) {
    if (n < 0 || which < 0 || which > 2 || (n > 0 && !out_xyz)) {
        set_error("sampler batch: bad arguments");
        return RTW_ERR_INVALID;
    }
    if (n == 0) return RTW_OK;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    DevRenderParams rp{};
    rp.seed_lo = (uint32_t)seed, rp.seed_hi = (uint32_t)(seed >> 32);
    philox_round_keys(rp);
    float* d = nullptr;
    int rc = RTW_OK;
    if ((e = dev_alloc((void**)&d, n * 12)) != cudaSuccess || (e = launch_sampler(which, d, n, rp)) != cudaSuccess ||
        (e = cudaMemcpy(out_xyz, d, n * 12, cudaMemcpyDeviceToHost)) != cudaSuccess)
        rc = cuda_fail(e, "sampler batch");
    dev_free(d);
    return rc;
}

// ------------------------------------------------------------------------------------------------ multi-GPU
// Samples-per-pixel sharding + ONE sum-reduce of the float4 accumulators onto the root.  The seam is still
// RayTracingWeekend.cpp:211-250: the root ends up with the radiance sums of ALL samples and alone applies :241-244.
// NCCL is bound at run time (dlopen of libnccl.so.2: the copy the host process already loaded, e.g. torch's, else the
// system one), so the library carries no link-time dependency and single-GPU users never touch it.
namespace {

struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
};

const NcclApi* nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* names[] = {getenv("RTW_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
        for (const char* n : names) {
            if (!n) continue;
            if ((api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL)) != nullptr) break;
        }
        if (!api.handle) return;
        api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(api.handle, "ncclGetUniqueId");
        api.CommInitRank = (decltype(api.CommInitRank))dlsym(api.handle, "ncclCommInitRank");
        api.CommDestroy = (decltype(api.CommDestroy))dlsym(api.handle, "ncclCommDestroy");
        api.Reduce = (decltype(api.Reduce))dlsym(api.handle, "ncclReduce");
        api.GetErrorString = (decltype(api.GetErrorString))dlsym(api.handle, "ncclGetErrorString");
        api.GetVersion = (decltype(api.GetVersion))dlsym(api.handle, "ncclGetVersion");
        if (!api.GetUniqueId || !api.CommInitRank || !api.CommDestroy || !api.Reduce || !api.GetErrorString) api.handle = nullptr;
    });
    if (!api.handle) {
        set_error("multi-GPU: libnccl.so.2 not found (set RTW_NCCL_LIB to its path)");
        return nullptr;
    }
    return &api;
}

int nccl_fail(const NcclApi* a, ncclResult_t r, const char* what) {
    set_error(std::string(what) + ": " + a->GetErrorString(r));
    return RTW_ERR_CUDA;
}

}  // namespace

struct rtw_comm {
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1, device = 0;
    bool owned = true;
};

static_assert(sizeof(ncclUniqueId) == RTW_COMM_ID_BYTES, "rtw_comm id size");

extern "C" int rtw_comm_unique_id(uint8_t* id) {
    const NcclApi* a = nccl_api();
    if (!a) return RTW_ERR_CUDA;
    if (!id) {
        set_error("comm_unique_id: null");
        return RTW_ERR_INVALID;
    }
    ncclUniqueId u;
    ncclResult_t r = a->GetUniqueId(&u);
    if (r != ncclSuccess) return nccl_fail(a, r, "ncclGetUniqueId");
    memcpy(id, &u, sizeof u);
    return RTW_OK;
}

extern "C" rtw_comm* rtw_comm_create(const uint8_t* id, int rank, int world, int device) {
    const NcclApi* a = nccl_api();
    if (!a) return nullptr;
    if (!id || world < 1 || rank < 0 || rank >= world) {
        set_error("comm_create: bad id / rank / world");
        return nullptr;
    }
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) {
        cuda_fail(e, "cudaSetDevice");
        return nullptr;
    }
    ncclUniqueId u;
    memcpy(&u, id, sizeof u);
    rtw_comm* c = new rtw_comm();
    c->rank = rank, c->world = world, c->device = device;
    ncclResult_t r = a->CommInitRank(&c->comm, world, u, rank);
    if (r != ncclSuccess) {
        nccl_fail(a, r, "ncclCommInitRank");
        delete c;
        return nullptr;
    }
    return c;
}

extern "C" rtw_comm* rtw_comm_adopt(void* nccl_comm, int rank, int world, int device) {
    if (!nccl_api()) return nullptr;
    if (!nccl_comm || world < 1 || rank < 0 || rank >= world) {
        set_error("comm_adopt: bad communicator / rank / world");
        return nullptr;
    }
    rtw_comm* c = new rtw_comm();
    c->comm = (ncclComm_t)nccl_comm, c->rank = rank, c->world = world, c->device = device, c->owned = false;
    return c;
}

extern "C" void rtw_comm_destroy(rtw_comm* c) {
    if (!c) return;
    const NcclApi* a = nccl_api();
    if (a && c->owned && c->comm) {
        cudaSetDevice(c->device);
        a->CommDestroy(c->comm);
    }
    delete c;
}

extern "C" int rtw_comm_rank(const rtw_comm* c) { return c ? c->rank : 0; }
extern "C" int rtw_comm_world(const rtw_comm* c) { return c ? c->world : 1; }

// [rank*spp/world, (rank+1)*spp/world): the ranges tile [0, spp) exactly; Philox streams are keyed by the GLOBAL sample index
extern "C" int rtw_shard_range(int spp, int rank, int world, int* spp_begin, int* spp_count) {
    if (spp < 0 || world < 1 || rank < 0 || rank >= world || !spp_begin || !spp_count) {
        set_error("shard_range: bad arguments");
        return RTW_ERR_INVALID;
    }
    const long long lo = (long long)rank * spp / world, hi = (long long)(rank + 1) * spp / world;
    *spp_begin = (int)lo, *spp_count = (int)(hi - lo);
    return RTW_OK;
}

extern "C" int rtw_reduce_accumulators(rtw_comm* c, float* accum_dev, int nx, int ny, int root, void* cuda_stream) {
    if (!c || c->world == 1) return RTW_OK;  // one rank: the accumulator is already the sum
    const NcclApi* a = nccl_api();
    if (!a) return RTW_ERR_CUDA;
    if (!accum_dev || nx <= 0 || ny <= 0 || root < 0 || root >= c->world) {
        set_error("reduce_accumulators: bad arguments");
        return RTW_ERR_INVALID;
    }
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    // in place: the w channel carries sample counts, so the reduced buffer is again (sum r, sum g, sum b, samples)
    ncclResult_t r = a->Reduce(accum_dev, accum_dev, (size_t)4 * nx * ny, ncclFloat32, ncclSum, root, c->comm, (cudaStream_t)cuda_stream);
    if (r != ncclSuccess) return nccl_fail(a, r, "ncclReduce");
    return RTW_OK;
}

// The whole multi-GPU step behind one call: `whole` describes ALL samples [spp_begin, spp_begin + spp_count); this rank
// renders its share of them, the accumulators are summed onto `root`, and the root alone receives the image
// (accum_host_root may be NULL elsewhere).  stats = this rank's shard.
extern "C" int rtw_render_sharded(rtw_scene* s, rtw_comm* c, const rtw_render_params* whole, int root, float* accum_host_root,
                                  rtw_render_stats* stats) {
    int rc = check_params(s, whole);
    if (rc != RTW_OK) return rc;
    const int rank = c ? c->rank : 0, world = c ? c->world : 1;
    if (root < 0 || root >= world || (rank == root && !accum_host_root)) {
        set_error("render_sharded: bad root / null image on the root");
        return RTW_ERR_INVALID;
    }
    if (whole->flags & RTW_FLAG_ACCUMULATE) {
        set_error("render_sharded: RTW_FLAG_ACCUMULATE is not supported here (accumulate on the root after the call)");
        return RTW_ERR_INVALID;
    }
    cudaError_t e = cudaSetDevice(s->device);
    if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    const size_t pixels = (size_t)whole->nx * whole->ny;
    if (pixels > s->scratch_pixels) {
        dev_free(s->scratch_accum);
        s->scratch_accum = nullptr, s->scratch_pixels = 0;
        if ((e = dev_alloc((void**)&s->scratch_accum, pixels * 16)) != cudaSuccess) return cuda_fail(e, "cudaMalloc accum");
        s->scratch_pixels = pixels;
    }
    rtw_render_params p = *whole;
    int b = 0, n = 0;
    rtw_shard_range(whole->spp_count, rank, world, &b, &n);
    p.spp_begin = whole->spp_begin + b, p.spp_count = n;
    if (stats) memset(stats, 0, sizeof *stats);
    if (n > 0) {
        if ((rc = rtw_render_device(s, &p, s->scratch_accum, nullptr)) != RTW_OK) return rc;
    } else if ((e = cudaMemsetAsync(s->scratch_accum, 0, pixels * 16, nullptr)) != cudaSuccess) {  // more ranks than samples
        return cuda_fail(e, "memset");
    }
    // same stream: the reduce starts when the kernel ends; the stats are read after both are queued
    if ((rc = rtw_reduce_accumulators(c, s->scratch_accum, whole->nx, whole->ny, root, nullptr)) != RTW_OK) return rc;
    if (n > 0 && (rc = rtw_render_device_finish(s, stats)) != RTW_OK) return rc;
    if (rank == root) {
        if ((e = cudaMemcpy(accum_host_root, s->scratch_accum, pixels * 16, cudaMemcpyDeviceToHost)) != cudaSuccess) return cuda_fail(e, "D2H accum");
    } else if ((e = cudaStreamSynchronize(nullptr)) != cudaSuccess) {
        return cuda_fail(e, "reduce");
    }
    return RTW_OK;
}
