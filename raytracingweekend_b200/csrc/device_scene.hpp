// device_scene.hpp — FP32 device-resident form of `rtw_flat_scene` (HBM layout).
//
// All arrays live in global memory and are read through the read-only path (ld.global.nc, 16-byte
// vectors).  In the linear-list loops every lane of a warp reads the SAME element, so each load is one
// broadcast transaction served by L1; BVH nodes are the only divergent reads.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace rtw {

// prim meta word: type[0:3) | flip[3] | precise[4] | xf+1 [5:17) | material [17:32)
enum : uint32_t {
    PM_TYPE_MASK = 0x7u,
    PM_FLIP = 1u << 3,
    PM_PRECISE = 1u << 4,  // sphere with |radius| >= 100: quadratic evaluated in FP64 (ground / fog spheres)
    PM_XF_SHIFT = 5,
    PM_XF_MASK = 0xFFFu,
    PM_MAT_SHIFT = 17
};

// composed translate/rotate_y chain: local = Ry(o) + t with Ry: x' = c*x - s*z, z' = s*x + c*z
struct DevXf {
    float c, s, tx, ty, tz, pad0, pad1, pad2;
};

struct DevItem {
    int32_t type, xf, first, count;  // xf: index into xfs or -1
    int32_t material, flip;
    float density_eff;  // density * multiplicity
    float pad;
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
    int32_t type, pad;
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

struct DevScene {
    const DevItem* items;
    const float4* pa;
    const float4* pb;
    const uint32_t* pm;
    const double* pd;  // 4 doubles per precise sphere (cx cy cz r); index in pb.w bits
    const DevXf* xfs;
    const float4* node_lo;  // bmin.xyz, left (int bits)
    const float4* node_hi;  // bmax.xyz, right (int bits)
    const int2* leaves;
    const DevMat* mats;
    const DevTex* texs;
    const uint8_t* texels;
    const float4* perlin_ranvec;
    const uint8_t* perlin_perm;  // 3*256
    const DevLight* lights;
    int32_t n_world_items, n_items, n_prims, n_lights;
    int32_t render_type, background_type;
    uint32_t features;
    int32_t max_bvh_depth;
    DevCamera cam;
};

struct DevRenderParams {
    int32_t nx, ny, spp_begin, spp_count, max_depth;
    uint32_t flags;
    uint32_t seed_lo, seed_hi;
};

}  // namespace rtw
