/*
 * rtw_b200.h — C ABI of the B200-native path-tracing backend.
 *
 * Drop-in boundary for ONE hot path of silvesthu/RayTracingWeekend: the body of
 * `time_call([&]{ ... })` at RayTracingWeekend/RayTracingWeekend.cpp:211-250, i.e. the
 * rows x columns x samples nest that calls `vec3 color(const ray&, const scene*, int)`
 * (RayTracingWeekend.cpp:45-160).  The reference has no FFI of its own;
 * the seam is its call graph.  A reference user keeps building scenes with the
 * reference's own classes (Scene/scene.h), includes `rtw_b200_adapter.hpp` (which
 * walks that object graph and drives the builder calls below) and replaces the
 * `_for` nest by `rtw_b200::render(scene, nx, ny, spp, max_depth, canvas)`.
 *
 * Everything here is plain C: opaque handles, plain pointers, sizes, integer status
 * codes; no exceptions cross the boundary.  There is NO CPU fallback: every compute
 * entry point fails with RTW_ERR_CUDA when no sm_100 device is usable.
 *
 * Three groups of entry points:
 *   1. builder   — one call per reference constructor (file:line cited at each);
 *                  flattens the scene graph into SoA/POD arrays (`rtw_flat_scene`),
 *                  keeping the exact topology of the reference's bvh_node build.
 *   2. scene     — upload the flat arrays once to a device; immutable afterwards.
 *   3. hot path  — rtw_render*, rtw_intersect_batch*; plus tonemap/PPM on the host
 *                  (RayTracingWeekend.cpp:241-244, 252-282).
 */
#ifndef RTW_B200_H
#define RTW_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTW_B200_ABI_VERSION 1

/* ------------------------------------------------------------------ status codes */
enum {
    RTW_OK = 0,
    RTW_ERR_INVALID = 1,     /* bad handle / id / argument                                  */
    RTW_ERR_UNSUPPORTED = 2, /* scene graph shape the flattener cannot express               */
    RTW_ERR_CUDA = 3,        /* CUDA runtime error or no usable sm_100 device (no fallback)  */
    RTW_ERR_NOMEM = 4,
    RTW_ERR_IO = 5
};

typedef struct rtw_builder rtw_builder; /* host-side scene graph under construction */
typedef struct rtw_scene rtw_scene;     /* device-resident immutable flattened scene */
typedef int32_t rtw_id;                 /* node id inside one builder; < 0 = error    */

/* Thread-local message of the last failing call on this thread ("" if none). */
const char* rtw_last_error(void);
int rtw_abi_version(void);

/* ------------------------------------------------------------------ enums of the flat format */
enum { /* rtw_flat_prim.type */
    RTW_PRIM_SPHERE = 0,        /* sphere.h:33-130 sphere_base<movement_none>   */
    RTW_PRIM_MOVING_SPHERE = 1, /* sphere.h:20-38 movement_linear               */
    RTW_PRIM_XY_RECT = 2,       /* hittable.h:142-175 */
    RTW_PRIM_XZ_RECT = 3,       /* hittable.h:177-232 */
    RTW_PRIM_YZ_RECT = 4        /* hittable.h:234-267 */
};
enum { /* rtw_flat_item.type */
    RTW_ITEM_PRIMS = 0, /* run of primitives tested in list order (hittable_list.h:11-37) */
    RTW_ITEM_BVH = 1,   /* bvh_node tree (hittable.h:41-140, as intended — SURVEY F1)      */
    RTW_ITEM_MEDIUM = 2 /* constant_medium (hittable.h:420-489)                             */
};
enum { /* rtw_flat_material.type */
    RTW_MAT_LAMBERTIAN = 0,    /* material.h:77-122  */
    RTW_MAT_METAL = 1,         /* material.h:124-140 */
    RTW_MAT_DIELECTRIC = 2,    /* material.h:142-225 */
    RTW_MAT_DIFFUSE_LIGHT = 3, /* material.h:227-247 */
    RTW_MAT_ISOTROPIC = 4      /* material.h:252-265 */
};
enum { /* rtw_flat_texture.type */
    RTW_TEX_CONSTANT = 0, /* texture.h:16-27 */
    RTW_TEX_CHECKER = 1,  /* texture.h:29-50 */
    RTW_TEX_NOISE = 2,    /* texture.h:52-71 */
    RTW_TEX_IMAGE = 3     /* texture.h:73-99 */
};
enum { /* rtw_flat_light.type */
    RTW_LIGHT_DEFAULT = 0, /* hittable.h:36-37 defaults: pdf_value 0, random (1,0,0) */
    RTW_LIGHT_XZ_RECT = 1, /* hittable.h:208-228 */
    RTW_LIGHT_SPHERE = 2   /* sphere.h:88-108    */
};
enum { RTW_XF_TRANSLATE = 1, RTW_XF_ROTATE_Y = 2 };  /* hittable.h:295-328, 331-416 */
enum { RTW_RENDER_SHADED = 0, RTW_RENDER_NORMAL = 1 }; /* Scene/scene.h:6-10  */
enum { RTW_BG_BLACK = 0, RTW_BG_GRADIENT = 1 };        /* Scene/scene.h:12-16 */

/* ------------------------------------------------------------------ flat scene (host, FP64 POD)
 * Host arrays keep the reference's doubles so the CPU oracle can re-evaluate them
 * bit-exactly against the reference object graph; the device copy is FP32.          */
typedef struct rtw_flat_prim {
    int32_t type;     /* RTW_PRIM_*                                                   */
    int32_t material; /* index into materials                                         */
    int32_t flip;     /* parity of flip_normals wrappers (hittable.h:269-292)         */
    int32_t xf;       /* transform chain head relative to the owning item, or -1      */
    /* sphere: cx cy cz radius [c1x c1y c1z time0 time1];  rect: a0 a1 b0 b1 k        */
    double v[9];
} rtw_flat_prim;

typedef struct rtw_flat_xf { /* one link of a translate/rotate_y chain, outermost first */
    int32_t kind;            /* RTW_XF_*                                                 */
    int32_t next;            /* next (inner) link or -1                                  */
    double a, b, c;          /* translate: offset xyz; rotate_y: sin_theta, cos_theta, 0 */
} rtw_flat_xf;

typedef struct rtw_flat_node { /* bvh_node (hittable.h:137-139)                         */
    double bmin[3], bmax[3];
    int32_t left, right;       /* >= 0: node index; < 0: ~leaf index                     */
} rtw_flat_node;

typedef struct rtw_flat_leaf { /* BVH leaf object = run of primitives in list order      */
    int32_t first_prim, prim_count;
} rtw_flat_leaf;

typedef struct rtw_flat_item {
    int32_t type;         /* RTW_ITEM_*                                                  */
    int32_t xf;           /* transform chain applied to the ray before the item, or -1   */
    int32_t first;        /* PRIMS: first prim; BVH: root node; MEDIUM: first boundary prim */
    int32_t count;        /* PRIMS/MEDIUM: prim count; BVH: node count of this tree       */
    int32_t material;     /* MEDIUM: phase material                                       */
    int32_t flip;         /* MEDIUM: parity of flip_normals wrappers                      */
    int32_t multiplicity; /* MEDIUM: independent free-flight draws per color() call
                             (2 per enclosing hittable_list — hittable_list.h:16-34, Q5) */
    int32_t pad_;
    double density;       /* MEDIUM: nominal density as passed to the constructor         */
} rtw_flat_item;

typedef struct rtw_flat_material {
    int32_t type;    /* RTW_MAT_*                                  */
    int32_t texture; /* lambertian/diffuse_light/isotropic: texture */
    double albedo[3];/* metal                                       */
    double fuzz;     /* metal                                       */
    double ref_idx;  /* dielectric                                  */
} rtw_flat_material;

typedef struct rtw_flat_texture {
    int32_t type;         /* RTW_TEX_*                                   */
    int32_t even, odd;    /* checker children (texture.h:33: even=t0, odd=t1) */
    int32_t image_offset; /* byte offset into texels                     */
    int32_t nx, ny;       /* image size                                  */
    double color[3];      /* constant                                    */
    double scale;         /* noise                                       */
} rtw_flat_texture;

typedef struct rtw_flat_light { /* entry of scene::lights (Scene/scene.h:35), unwrapped shape */
    int32_t type;               /* RTW_LIGHT_*                                                 */
    int32_t pad_;
    double v[5];                /* xz_rect: x0 x1 z0 z1 k;  sphere: cx cy cz radius            */
} rtw_flat_light;

typedef struct rtw_flat_camera { /* camera.h:52-58 public members */
    double origin[3], lower_left_corner[3], horizontal[3], vertical[3], u[3], v[3], w[3];
    double time0, time1, lens_radius;
} rtw_flat_camera;

typedef struct rtw_flat_scene {
    int32_t n_world_items; /* items[0..n_world_items) are scene.world in list order           */
    int32_t n_items, n_prims, n_xfs, n_nodes, n_leaves, n_materials, n_textures, n_lights;
    int32_t n_texel_bytes;
    int32_t render_type;     /* RTW_RENDER_* */
    int32_t background_type; /* RTW_BG_*     */
    const rtw_flat_item* items;
    const rtw_flat_prim* prims;
    const rtw_flat_xf* xfs;
    const rtw_flat_node* nodes;
    const rtw_flat_leaf* leaves;
    const rtw_flat_material* materials;
    const rtw_flat_texture* textures;
    const rtw_flat_light* lights;
    const uint8_t* texels;     /* RGB8 rows, top row first (texture.h:84) */
    const double* perlin_ranvec; /* 256*3 (noise.h:216) or NULL when no noise texture          */
    const int32_t* perlin_perm;  /* 3*256: perm_x, perm_y, perm_z (noise.h:218-220) or NULL    */
    rtw_flat_camera camera;
} rtw_flat_scene;

/* ------------------------------------------------------------------ 1. builder
 * One function per reference constructor.  Ids are only meaningful inside their builder. */
rtw_builder* rtw_builder_create(void);
void rtw_builder_destroy(rtw_builder* b);

/* texture.h:20 / :33 / :56 / :80 */
rtw_id rtw_constant_texture(rtw_builder* b, double r, double g, double bl);
rtw_id rtw_checker_texture(rtw_builder* b, rtw_id t0_even, rtw_id t1_odd);
rtw_id rtw_noise_texture(rtw_builder* b, double scale);
rtw_id rtw_image_texture(rtw_builder* b, const uint8_t* rgb8, int nx, int ny); /* bytes are copied */

/* material.h:80 / :127 / :145 / :230 / :255 */
rtw_id rtw_lambertian(rtw_builder* b, rtw_id albedo_texture);
rtw_id rtw_metal(rtw_builder* b, double r, double g, double bl, double fuzz);
rtw_id rtw_dielectric(rtw_builder* b, double ref_idx);
rtw_id rtw_diffuse_light(rtw_builder* b, rtw_id emit_texture);
rtw_id rtw_isotropic(rtw_builder* b, rtw_id albedo_texture);

/* sphere.h:45, :20-38 + :110-113 (set_movement) */
rtw_id rtw_sphere(rtw_builder* b, double cx, double cy, double cz, double radius, rtw_id material);
rtw_id rtw_moving_sphere(rtw_builder* b, double cx, double cy, double cz, double radius, rtw_id material,
                         double c1x, double c1y, double c1z, double time0, double time1);
/* hittable.h:146-148 / :181-183 / :238-240 */
rtw_id rtw_xy_rect(rtw_builder* b, double x0, double x1, double y0, double y1, double k, rtw_id material);
rtw_id rtw_xz_rect(rtw_builder* b, double x0, double x1, double z0, double z1, double k, rtw_id material);
rtw_id rtw_yz_rect(rtw_builder* b, double y0, double y1, double z0, double z1, double k, rtw_id material);
/* hittable_list.h:69 — six rects, min-side ones flipped, in the reference's order */
rtw_id rtw_box(rtw_builder* b, const double p0[3], const double p1[3], rtw_id material);
/* hittable.h:272 / :298 / :334 */
rtw_id rtw_flip_normals(rtw_builder* b, rtw_id child);
rtw_id rtw_translate(rtw_builder* b, rtw_id child, double ox, double oy, double oz);
rtw_id rtw_rotate_y(rtw_builder* b, rtw_id child, double angle_degrees);
/* same, from the already computed members rotate_y::sin_theta/cos_theta (hittable.h:412-413) */
rtw_id rtw_rotate_y_sincos(rtw_builder* b, rtw_id child, double sin_theta, double cos_theta);
/* hittable.h:423-428 — third argument is a *material* (isotropic), as in the reference */
rtw_id rtw_constant_medium(rtw_builder* b, rtw_id boundary, double density, rtw_id material);
/* hittable_list.h:9 */
rtw_id rtw_hittable_list(rtw_builder* b, const rtw_id* children, int n);
/* hittable.h:45-95, as intended (`right` assigned, `right->hit` tested):
 * random axis from a builder-local default-seeded std::minstd_rand +
 * uniform_int_distribution<int>(0,2), std::sort(&l[0], &l[n-1]) on bounding_box(0,0).min()[axis],
 * median split, pre-order recursion.  Children array is permuted in place like `l`. */
rtw_id rtw_bvh_node(rtw_builder* b, rtw_id* children, int n, double time0, double time1);
/* Import one node of a tree the reference already built in this process (adapter path):
 * public members left/right/box (hittable.h:137-139). */
rtw_id rtw_bvh_node_import(rtw_builder* b, rtw_id left, rtw_id right, const double box_min[3],
                           const double box_max[3]);
/* hittable::bounding_box(t0,t1) of any builder node, reference semantics. Returns 1/0 like the
 * reference's bool, < 0 on error. */
int rtw_bounding_box(const rtw_builder* b, rtw_id node, double t0, double t1, double out_min[3],
                     double out_max[3]);

/* Scene/scene.h:18-40 */
int rtw_set_world(rtw_builder* b, rtw_id hittable_list);
int rtw_set_lights(rtw_builder* b, rtw_id hittable_list_or_minus1);
int rtw_set_render_type(rtw_builder* b, int render_type);
int rtw_set_background_type(rtw_builder* b, int background_type);
/* camera.h:13-34 (same argument order) */
int rtw_set_camera(rtw_builder* b, const double lookfrom[3], const double lookat[3], const double vup[3],
                   double vfov, double aspect, double aperture, double focus_dist, double t0, double t1);
/* camera.h:52-58, already derived members */
int rtw_set_camera_members(rtw_builder* b, const rtw_flat_camera* cam);
/* noise.h:215-220 tables read from the reference (adapter path). Without this call the builder
 * generates them exactly as noise.h:166-213 does (default-seeded std::minstd_rand). */
int rtw_set_perlin_tables(rtw_builder* b, const double ranvec[768], const int32_t perm_x[256],
                          const int32_t perm_y[256], const int32_t perm_z[256]);

/* Named scene functions composed on the builder (mirror Scene/scene.h:42-249 and SURVEY App. C).
 * names: "light_sample", "dielectric_scene", "random_balls_scene", "cornell_box_scene",
 *        "cornell_box_aluminum" (scene.h:228 enabled), "two_perlin_spheres_earth" (cfg2),
 *        "book2_final_scene" (cfg3).  earth_rgb8 may be NULL for scenes without image_texture. */
int rtw_build_named_scene(rtw_builder* b, const char* name, double aspect, const uint8_t* earth_rgb8,
                          int earth_nx, int earth_ny);

/* Flatten.  The returned view is owned by the builder and valid until the next mutating call or
 * rtw_builder_destroy.  NULL on error (RTW_ERR_UNSUPPORTED: see rtw_last_error). */
const rtw_flat_scene* rtw_builder_flatten(rtw_builder* b);
/* Index in rtw_flat_scene.materials of a builder material id after rtw_builder_flatten (-1: unused). */
int32_t rtw_builder_flat_material_index(const rtw_builder* b, rtw_id material);
/* Canonical little-endian serialisation of the flat scene (fixtures, hashing). Returns bytes needed;
 * writes at most cap bytes. */
size_t rtw_flat_serialize(const rtw_flat_scene* fs, void* buf, size_t cap);

/* ------------------------------------------------------------------ 2. device scene */
rtw_scene* rtw_scene_create(const rtw_flat_scene* fs, int device);
void rtw_scene_destroy(rtw_scene* s);

/* ------------------------------------------------------------------ 3. hot path */
enum { /* rtw_render_params.flags */
    RTW_FLAG_WAVEFRONT = 1u << 0,   /* use the wavefront pipeline instead of the megakernel      */
    RTW_FLAG_ACCUMULATE = 1u << 1,  /* add into accum instead of overwriting (progressive/resume) */
    RTW_FLAG_DEBUG_RAY = 1u << 2,   /* RayTracingWeekend.cpp:221-225: aim every sample at the centre pixel */
    RTW_FLAG_AUTO = 1u << 3         /* let the backend pick megakernel or wavefront from the scene's features: the wavefront
                                       pipeline where shading is heavy and divergent (Perlin noise textures; bvh_node trees
                                       together with media) — the ncu evidence is in profiles/ — the megakernel otherwise.  Sums are then only
                                       reproducible up to the order of float adds. */,
    RTW_FLAG_NEE = 1u << 4          /* OPT-IN, changes the estimator of RayTracingWeekend.cpp:117-132 (the reference casts no occlusion
                                       rays): next-event estimation — every lambertian hit casts a shadow ray at a point of the
                                       emitting xz_rect lights of the lights list and continues with a cosine-sampled direction.
                                       Same expected image, different (on Cornell-like scenes much lower) variance; runs in the
                                       wavefront pipeline (generate / extend / shade / shadow kernels), implies RTW_FLAG_WAVEFRONT.
                                       stats.rays still counts color() invocations only, not shadow rays. */
};

typedef struct rtw_render_params {
    int32_t nx, ny;        /* RayTracingWeekend.cpp:35-36                                      */
    int32_t spp_begin;     /* first global sample index of this shard                          */
    int32_t spp_count;     /* samples rendered by this call (subPixelCount, :33)               */
    int32_t max_depth;     /* :42 (100)                                                        */
    uint32_t flags;
    uint64_t seed;         /* Philox key; streams are (seed, global sample index)              */
} rtw_render_params;

typedef struct rtw_render_stats {
    uint64_t paths;   /* camera samples = nx*ny*spp_count                          */
    uint64_t rays;    /* color() invocations (primary + scattered), device-counted */
    double kernel_ms; /* CUDA-event time of the render kernels on their stream     */
    int32_t kernel_launches;
    int32_t pad_;
} rtw_render_stats;

/* Linear radiance SUMS (not means, not gamma-corrected): accum[j*nx+i] = (sum r, sum g, sum b, samples),
 * j = 0 is the BOTTOM row exactly like `canvas` (RayTracingWeekend.cpp:247, 261).
 * rtw_render:        host buffer, synchronous (device->host copy included).
 * rtw_render_device: device buffer on the scene's device, asynchronous on `cuda_stream`
 *                    (a cudaStream_t passed as void*; NULL = legacy default stream);
 *                    stats->rays/kernel_ms are valid after rtw_render_device_finish.  ONE render in flight per
 *                    rtw_scene (its ray counter and timing events live in the scene): a second rtw_render_device
 *                    before rtw_render_device_finish returns RTW_ERR_INVALID.  With RTW_FLAG_WAVEFRONT the call is
 *                    synchronous (the host polls the pool between chunks of iterations).
 *                    rtw_render / rtw_render_device_finish return RTW_ERR_CUDA (and the image must be discarded) if the
 *                    render kernel's watchdog fired: a block found no work for over a second with paths still in flight. */
int rtw_render(rtw_scene* s, const rtw_render_params* p, float* accum_host_rgba, rtw_render_stats* stats);
int rtw_render_device(rtw_scene* s, const rtw_render_params* p, float* accum_dev_rgba, void* cuda_stream);
int rtw_render_device_finish(rtw_scene* s, rtw_render_stats* stats);

/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The wavefront render
 * (rtw_render_device with RTW_FLAG_WAVEFRONT) with its scratch memory - the pool of path slots, the class queues and the
 * counters - in a static device array of the library instead of allocated memory: the call allocates and frees NOTHING
 * (no cudaMalloc / cudaFree / cudaMallocHost, no allocating library routine).  Limit: the array holds at most
 * RTW_NOALLOC_MAX_POOL_SLOTS path slots (64 Ki slots, about 5.8 MB that every GPU which loads the library holds until the
 * process exits); pool_slots = 0 picks the largest useful pool, a larger request fails with RTW_ERR_INVALID.  Any image size
 * works (slots are refilled as paths end), a small pool only costs throughput.  The caller passes the scene (uploaded
 * beforehand: set-up may allocate) and the device accumulator; results are those of rtw_render_device with
 * RTW_FLAG_WAVEFRONT: identical ray counts, sums equal up to the order of the float adds.  Synchronous.
 * ALL calls on one GPU share the array and it keeps the previous call's contents (the call clears what must start at
 * zero): calls that overlap on the GPU - two host threads, two streams - race on it. */
#define RTW_NOALLOC_MAX_POOL_SLOTS 65536u
int rtw_render_device_noalloc(rtw_scene* s, const rtw_render_params* p, float* accum_dev_rgba, uint32_t pool_slots,
                              void* cuda_stream, rtw_render_stats* stats);

typedef struct rtw_ray { /* ray.h:17-20, FP32 */
    float ox, oy, oz, dx, dy, dz, time, pad_;
} rtw_ray;
typedef struct rtw_hit { /* hit_record (hittable.h:16-29) + stable ids */
    int32_t prim;     /* index into rtw_flat_scene.prims, -1 = miss; media report ~item index - see docs */
    int32_t material; /* index into materials, -1 = miss                                              */
    float t, u, v;
    float nx, ny, nz;
} rtw_hit;
/* Closest hit of world.hit(r, t_min, t_max) for n rays (RayTracingWeekend.cpp:52 uses 0.001f, DBL_MAX).
 * constant_medium items are skipped (they are stochastic). Host buffers, synchronous. */
int rtw_intersect_batch(rtw_scene* s, const rtw_ray* rays_host, rtw_hit* hits_host, int64_t n, float t_min,
                        float t_max);
int rtw_intersect_batch_device(rtw_scene* s, const rtw_ray* rays_dev, rtw_hit* hits_dev, int64_t n, float t_min,
                               float t_max, void* cuda_stream);

/* ------------------------------------------------------------------ 4. multi-GPU
 * The reference has ONE process; its seam RayTracingWeekend.cpp:211-250 produces all samples of every pixel and :241-244
 * tonemaps them.  Here GPU g of G renders the global sample indices [g*spp/G, (g+1)*spp/G) of every pixel (Philox
 * streams are keyed by the global sample index, so shards are disjoint and their union is the 1-GPU render), the float4
 * accumulators are summed onto the root with one ncclReduce over NVLink, and the root alone tonemaps and writes.
 * One process (or thread) per GPU.  NCCL is bound at run time (dlopen libnccl.so.2; RTW_NCCL_LIB overrides). */
#define RTW_COMM_ID_BYTES 128
typedef struct rtw_comm rtw_comm;
/* ncclGetUniqueId: call on ONE rank and hand the 128 bytes to the others (file, socket, MPI, torch.distributed ...). */
int rtw_comm_unique_id(uint8_t id[RTW_COMM_ID_BYTES]);
/* ncclCommInitRank on `device`; collective over all ranks.  NULL on error. */
rtw_comm* rtw_comm_create(const uint8_t id[RTW_COMM_ID_BYTES], int rank, int world, int device);
/* Wrap an ncclComm_t the host application already owns (not destroyed by rtw_comm_destroy). */
rtw_comm* rtw_comm_adopt(void* nccl_comm, int rank, int world, int device);
void rtw_comm_destroy(rtw_comm* c);
int rtw_comm_rank(const rtw_comm* c);
int rtw_comm_world(const rtw_comm* c);
/* Sample range of `rank`: [rank*spp/world, (rank+1)*spp/world); the ranges tile [0, spp) exactly. */
int rtw_shard_range(int spp, int rank, int world, int* spp_begin, int* spp_count);
/* In-place sum of the per-rank accumulators (nx*ny float4, device memory) onto `root`, asynchronous on `cuda_stream`;
 * the w channel carries sample counts, so the root's buffer is again (sum r, sum g, sum b, samples).  c == NULL or one
 * rank: no-op. */
int rtw_reduce_accumulators(rtw_comm* c, float* accum_dev_rgba, int nx, int ny, int root, void* cuda_stream);
/* The whole multi-GPU step: `whole` names ALL samples; this rank renders its shard, the sums are reduced onto `root`,
 * which alone receives the image in host memory (accum_host_root may be NULL on the other ranks).  Synchronous.
 * stats = this rank's shard. */
int rtw_render_sharded(rtw_scene* s, rtw_comm* c, const rtw_render_params* whole, int root, float* accum_host_root,
                       rtw_render_stats* stats);

/* ------------------------------------------------------------------ host-side output (rank 0)
 * RayTracingWeekend.cpp:241-244: mean, sqrt gamma, min(.,1) -> canvas (3 doubles per pixel). */
int rtw_tonemap(const float* accum_rgba, int nx, int ny, double* canvas_rgb);
/* RayTracingWeekend.cpp:252-275: "P3", top row first, int(255.99f*c), one pixel per line. */
int rtw_write_ppm(const char* path, const double* canvas_rgb, int nx, int ny);

/* "P6" binary twin of rtw_write_ppm: same rows, same int(255.99f*c) quantisation (RayTracingWeekend.cpp:252-275). */
int rtw_write_ppm_binary(const char* path, const double* canvas_rgb, int nx, int ny);
/* 8-bit RGB PNG of the same pixels — replaces `system("magick 1.ppm 1.png")` (RayTracingWeekend.cpp:285). */
int rtw_write_png(const char* path, const double* canvas_rgb, int nx, int ny);
/* Texture ingest for image_texture (texture.h:73-99 takes caller-decoded bytes; the reference ships earth.jpg but no
 * loader): baseline JPEG -> RGB8 rows, top row first.  rgb8 == NULL only reports the size.  RTW_ERR_UNSUPPORTED for
 * progressive / arithmetic / >2x subsampled files. */
int rtw_decode_jpeg(const uint8_t* data, size_t size, uint8_t* rgb8, size_t rgb8_capacity, int* out_nx, int* out_ny);

/* Known-answer hooks: the device functions of the render kernels, evaluated on caller-chosen inputs.
 * rtw_philox4x32_10_batch: Philox4x32-10 (Salmon et al., SC'11; replaces the engines of utility.h:14-20) of n counters
 *   (4 x uint32 each) under `key` -> n x 4 uint32; checked against the published Random123 vectors.
 * rtw_sampler_batch: n samples of the bounded samplers that replace the reference's rejection loops, sample i drawn from
 *   the Philox counter (pixel = i, sample = 0, bounce = 1, site): which = 0 uniform unit ball (utility.h:27-35), 1 unit
 *   lens disk (camera.h:61-69), 2 cosine-weighted direction about +z (utility.h:54-67) -> n x 3 floats (host). */
int rtw_philox4x32_10_batch(int device, const uint32_t* counters, int64_t n, uint64_t key, uint32_t* out);
int rtw_sampler_batch(int device, int which, int64_t n, uint64_t seed, float* out_xyz);

/* Measured FP32 FFMA throughput of `device` in TFLOP/s (roofline denominator). */
int rtw_measure_fp32_peak(int device, double* tflops);
/* Device utility: number of CUDA devices with compute capability 10.x (0 = none / no driver). */
int rtw_device_count(void);

#ifdef __cplusplus
}
#endif
#endif /* RTW_B200_H */
