// device_scene.hpp — FP32 device-resident form of `rtw_flat_scene` (HBM layout).
//
// The whole scene except image texels is ONE packed blob of 16-byte aligned sections, ordered hot-first:
//
//   [ items | groups | recs | xfs | lights | nodes | leaves | nodes4 ]      <- "hot": touched per primitive test
//   [ pm | pa | pb | pd | mats | texs | perlin_ranvec | perlin_perm ]       <- "cold": touched once per ray
//
// A render block stages a prefix of the blob into shared memory (everything, the hot part, or nothing,
// depending on its size) and walks it with warp-uniform addresses, so every primitive-test load is a shared
// memory broadcast; larger scenes read the same sections from global memory through L1 (ld.global.nc).
//
// Everything inside the blob is addressed by BYTE OFFSET FROM THE BLOB START (uint32): the same offset is
// valid against the shared-memory copy (32-bit shared address = staged base + offset, `ld.shared.v4`) and against
// the global copy (`ld.global.nc.v4`), and the inner loops advance one address register by 32 bytes per record.
//
// Primitive tests are driven by `DevGroup`s: a group is a run of primitives of one item that share a
// transform chain, re-ordered by kind (spheres, moving spheres, xy/xz/yz rects) so the inner loops are
// branch-free per kind.  The reference's list-order tie rules survive the re-ordering because they
// can be stated order-independently: among equal t a rect beats a sphere, the later rect beats the earlier
// rect, the earlier sphere beats the later sphere — and spheres-then-rects in index order evaluates exactly that.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace rtw {

// prim meta word: type[0:3) | flip[3] | precise[4] | xf+1 [5:17) | material [17:32)
enum : uint32_t {
    PM_TYPE_MASK = 0x7u,
    PM_FLIP = 1u << 3,
    PM_PRECISE = 1u << 4,  // sphere with |radius| >= 100 (api.cu precise_min_radius): quadratic evaluated in FP64 (ground / fog spheres)
    PM_XF_SHIFT = 5,
    PM_XF_MASK = 0xFFFu,
    PM_MAT_SHIFT = 17
};

// composed translate/rotate_y chain: local = Ry(o) + t with Ry: x' = c*x - s*z, z' = s*x + c*z
struct DevXf {
    float c, s, tx, ty, tz, pad0, pad1, pad2;
};

struct DevItem {
    int32_t type;     // RTW_ITEM_*
    int32_t xf;       // byte offset of the item's DevXf, or -1
    int32_t first;    // PRIMS/MEDIUM: byte offset of the first DevGroup; BVH: root reference (see DevNode)
    int32_t end;      // PRIMS/MEDIUM: byte offset one past the last DevGroup
    int32_t material, flip;
    float density_eff;  // density * multiplicity
    int32_t root4;      // BVH: byte offset of the root DevNode4 of the render tree
};

// Run of primitive-test records sharing one transform chain.  All members are byte offsets into the blob:
// [sph0, sph1) spheres; [sph1, msph1) moving spheres; [msph1, xy1) xy_rects; [xy1, xz1) xz_rects; [xz1, yz1) yz_rects;
// [yz1, box1) box records.  The six rects of a reference `box` (hittable_list.h:69-104) form a group of their own
// that ALSO carries one box record: the render kernels run one slab test instead of six rect tests (same t expression,
// same winning face = same primitive id), intersect_batch keeps the rect-by-rect order of the reference.
struct DevGroup {
    int32_t xf;  // byte offset of the DevXf, or -1
    int32_t sph0, sph1, msph1;
    int32_t xy1, xz1, yz1, box1;
};

// One primitive-test record (32 bytes = two 16-byte loads).
//   box:           a = (lo.x, lo.y, lo.z, -)       b = (hi.x, hi.y, hi.z, id of the first of its six rect prims)
//   sphere:        a = (cx, cy, cz, radius)        b = (pd index bits, -, -, prim id | REC_PRECISE)
//   moving sphere: a = (c'x, c'y, c'z, radius)     b = (dx, dy, dz, prim id)     centre(time) = c' + time*d
//   rect:          a = (a0, a1, b0, b1)            b = (k, -, -, prim id)
struct DevRec {
    float4 a, b;
};
constexpr int32_t REC_PRECISE = 1 << 30;

// One traversal step of a bounding-volume tree, 64 bytes: the boxes of BOTH children live in the parent, so one visit
// decides both (bvh_node::hit, hittable.h:97-129, enters a child iff the child's own box is hit — the same boxes, tested
// one level up).  Child references: >= 0 byte offset of a DevNode; < 0: ~(byte offset of the leaf's int2 {first group
// offset, end group offset}).  Leaves are numbered left to right, so the byte offset orders them as the tree does.
//   * a reference bvh_node child keeps the reference's box (hittable.h:84-94);
//   * a LEAF child (a list/box/primitive — the reference calls its hit() without any box test) carries a padded
//     bounding box of its primitives, so most leaves are skipped without touching their records: a ray that hits a
//     primitive passes through its padded box, so no hit is ever lost;
//   * a node with left == right (n == 1, hittable.h:68-72) has an empty right box: one visit;
//   * every tree gets a pseudo-root whose left child is the real root (with the root's own box) and whose right box is empty.
// The same structure backs the trees this backend builds itself over long hittable_lists (api.cu build_list_bvh).
struct DevNode {
    float4 llo;  // left child box min.xyz, left reference (int bits)
    float4 lhi;  // left child box max.xyz, right reference (int bits)
    float4 rlo;  // right child box min.xyz
    float4 rhi;  // right child box max.xyz
};

// One step of the 4-wide tree the RENDER kernels walk (128 bytes = one cache line): the binary tree built at upload,
// collapsed so that a node holds the boxes of up to four descendants in SoA form — half the dependent node fetches per
// ray of the binary tree, which is what bounds the big scenes (node fetches are L1/L2 latency, not instruction issue).
// Child references as in DevNode (>= 0: byte offset of a DevNode4, < 0: ~leaf offset).  An unused slot has a far-away
// degenerate box that no finite ray reaches and references an empty leaf (a NaN ray, whose slabs all drop out, lands there).
struct DevNode4 {
    float4 lox, hix, loy, hiy, loz, hiz;  // component i = child i
    int32_t ref[4];
    int32_t pad[4];
};

struct DevMat {
    float r, g, b, param;  // metal: albedo + fuzz; dielectric: param = ref_idx; textured: resolved constant colour
    int32_t type, texture, tex_is_constant, pad;
};

struct DevTex {
    float r, g, b, scale;
    int32_t type, even, odd, image_offset;
    int32_t nx, ny, pad0, pad1;
};

struct DevLight {
    float a0, a1, a2, a3;  // xz_rect: x0 x1 z0 z1; sphere: cx cy cz r
    float k, area;         // xz_rect
    int32_t type, pad;     // pad = 1: an xz_rect light whose material is a diffuse_light (the NEE set)
};

struct DevCamera {
    float ox, oy, oz, llx, lly, llz, hx, hy, hz, vx, vy, vz, ux, uy, uz, wx, wy, wz;  // w* = camera.v (named to avoid clash)
    float time0, time1, lens_radius, pad;
};

enum : uint32_t {  // feature mask -> kernel specialisation
    F_MOVING = 1u << 0,
    F_TEXTURES = 1u << 1,  // any non-constant texture
    F_BVH = 1u << 2,
    F_MEDIA = 1u << 3,
    F_XF = 1u << 4,
    F_PRECISE = 1u << 5,
    F_ALL = 0x3Fu
};

enum {  // blob sections, in blob order
    SEC_ITEMS = 0, SEC_GROUPS, SEC_RECS, SEC_XFS, SEC_LIGHTS, SEC_NODES, SEC_LEAVES, SEC_NODES4,  // hot
    SEC_PM, SEC_PA, SEC_PB, SEC_PD, SEC_MATS, SEC_TEXS, SEC_RANVEC, SEC_PERM,        // cold
    SEC_COUNT
};

struct DevScene {
    const uint8_t* blob;   // global memory, 16-byte aligned
    const uint8_t* texels;  // RGB8 image texels (global only)
    uint32_t off[SEC_COUNT];  // byte offset of each section inside the blob
    uint32_t hot_bytes, all_bytes;  // blob prefix sizes (multiples of 16)
    int32_t n_world_items, n_items, n_prims, n_lights;
    int32_t render_type, background_type;
    uint32_t features;
    int32_t stack_exact;    // binary trees, reference topology (intersect_batch): most references ever pushed at once
    int32_t stack_render;   // binary render trees walked in one loop, ENTER markers included (bvh_hit_items)
    int32_t bvh4_stack;     // 4-wide render trees: most references ever pushed at once
    int32_t n_bvh_items;    // tree items among the world items
    int32_t n_nee_lights;   // emitting xz_rect lights of the lights list (DevLight.pad = 1): what RTW_FLAG_NEE samples
    DevCamera cam;
};

struct DevRenderParams {
    int32_t nx, ny, spp_begin, spp_count, max_depth;
    uint32_t flags;
    uint32_t seed_lo, seed_hi;
    uint32_t rk[20];  // Philox4x32-10 round keys of the seed: rk[2i] = seed_lo + i*0x9E3779B9, rk[2i+1] = seed_hi + i*0xBB67AE85
};
inline void philox_round_keys(DevRenderParams& rp) {
    for (uint32_t i = 0; i < 10; i++) {
        rp.rk[2 * i] = rp.seed_lo + i * 0x9E3779B9u;
        rp.rk[2 * i + 1] = rp.seed_hi + i * 0xBB67AE85u;
    }
}

// Where the blob lives for this block: `sbase` = 32-bit shared address of the staged prefix, `g` = global copy.
struct SceneView {
    uint32_t sbase;
    const uint8_t* g;
    const uint8_t* texels;
};

}  // namespace rtw
