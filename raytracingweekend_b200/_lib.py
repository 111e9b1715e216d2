"""ctypes binding of include/rtw_b200.h.  Fails loudly when the native library is missing: there is no
Python or CPU fallback for the hot path."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# RTW_B200_LIB: alternative build of the same library (kernel tuning experiments); never a CPU fallback
LIB_PATH = os.environ.get("RTW_B200_LIB") or os.path.join(HERE, "lib", "librtw_b200.so")

RTW_OK = 0
RTW_FLAG_WAVEFRONT = 1
RTW_FLAG_AUTO = 8
RTW_FLAG_ACCUMULATE = 2
RTW_FLAG_DEBUG_RAY = 4
RTW_FLAG_NEE = 16
RTW_NOALLOC_MAX_POOL_SLOTS = 65536
RTW_COMM_ID_BYTES = 128


class FlatCamera(C.Structure):
    _fields_ = [(n, C.c_double * 3) for n in ("origin", "lower_left_corner", "horizontal", "vertical", "u", "v", "w")] + [
        ("time0", C.c_double), ("time1", C.c_double), ("lens_radius", C.c_double)]


class FlatPrim(C.Structure):
    _fields_ = [("type", C.c_int32), ("material", C.c_int32), ("flip", C.c_int32), ("xf", C.c_int32), ("v", C.c_double * 9)]


class FlatXf(C.Structure):
    _fields_ = [("kind", C.c_int32), ("next", C.c_int32), ("a", C.c_double), ("b", C.c_double), ("c", C.c_double)]


class FlatNode(C.Structure):
    _fields_ = [("bmin", C.c_double * 3), ("bmax", C.c_double * 3), ("left", C.c_int32), ("right", C.c_int32)]


class FlatLeaf(C.Structure):
    _fields_ = [("first_prim", C.c_int32), ("prim_count", C.c_int32)]


class FlatItem(C.Structure):
    _fields_ = [("type", C.c_int32), ("xf", C.c_int32), ("first", C.c_int32), ("count", C.c_int32), ("material", C.c_int32),
                ("flip", C.c_int32), ("multiplicity", C.c_int32), ("pad_", C.c_int32), ("density", C.c_double)]


class FlatMaterial(C.Structure):
    _fields_ = [("type", C.c_int32), ("texture", C.c_int32), ("albedo", C.c_double * 3), ("fuzz", C.c_double), ("ref_idx", C.c_double)]


class FlatTexture(C.Structure):
    _fields_ = [("type", C.c_int32), ("even", C.c_int32), ("odd", C.c_int32), ("image_offset", C.c_int32), ("nx", C.c_int32),
                ("ny", C.c_int32), ("color", C.c_double * 3), ("scale", C.c_double)]


class FlatLight(C.Structure):
    _fields_ = [("type", C.c_int32), ("pad_", C.c_int32), ("v", C.c_double * 5)]


class FlatScene(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("n_world_items", "n_items", "n_prims", "n_xfs", "n_nodes", "n_leaves", "n_materials",
                                         "n_textures", "n_lights", "n_texel_bytes", "render_type", "background_type")] + [
        ("items", C.POINTER(FlatItem)), ("prims", C.POINTER(FlatPrim)), ("xfs", C.POINTER(FlatXf)), ("nodes", C.POINTER(FlatNode)),
        ("leaves", C.POINTER(FlatLeaf)), ("materials", C.POINTER(FlatMaterial)), ("textures", C.POINTER(FlatTexture)),
        ("lights", C.POINTER(FlatLight)), ("texels", C.POINTER(C.c_uint8)), ("perlin_ranvec", C.POINTER(C.c_double)),
        ("perlin_perm", C.POINTER(C.c_int32)), ("camera", FlatCamera)]


class RenderParams(C.Structure):
    _fields_ = [("nx", C.c_int32), ("ny", C.c_int32), ("spp_begin", C.c_int32), ("spp_count", C.c_int32), ("max_depth", C.c_int32),
                ("flags", C.c_uint32), ("seed", C.c_uint64)]


class RenderStats(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("rays", C.c_uint64), ("kernel_ms", C.c_double), ("kernel_launches", C.c_int32), ("pad_", C.c_int32)]


class Ray(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("ox", "oy", "oz", "dx", "dy", "dz", "time", "pad_")]


class Hit(C.Structure):
    _fields_ = [("prim", C.c_int32), ("material", C.c_int32), ("t", C.c_float), ("u", C.c_float), ("v", C.c_float),
                ("nx", C.c_float), ("ny", C.c_float), ("nz", C.c_float)]


_D, _I, _P = C.c_double, C.c_int32, C.c_void_p
_D3 = C.POINTER(C.c_double)

# name -> (restype, argtypes); every symbol declared in include/rtw_b200.h
SIGNATURES = {
    "rtw_last_error": (C.c_char_p, []),
    "rtw_abi_version": (C.c_int, []),
    "rtw_builder_create": (_P, []),
    "rtw_builder_destroy": (None, [_P]),
    "rtw_constant_texture": (_I, [_P, _D, _D, _D]),
    "rtw_checker_texture": (_I, [_P, _I, _I]),
    "rtw_noise_texture": (_I, [_P, _D]),
    "rtw_image_texture": (_I, [_P, C.c_char_p, C.c_int, C.c_int]),
    "rtw_lambertian": (_I, [_P, _I]),
    "rtw_metal": (_I, [_P, _D, _D, _D, _D]),
    "rtw_dielectric": (_I, [_P, _D]),
    "rtw_diffuse_light": (_I, [_P, _I]),
    "rtw_isotropic": (_I, [_P, _I]),
    "rtw_sphere": (_I, [_P, _D, _D, _D, _D, _I]),
    "rtw_moving_sphere": (_I, [_P, _D, _D, _D, _D, _I, _D, _D, _D, _D, _D]),
    "rtw_xy_rect": (_I, [_P, _D, _D, _D, _D, _D, _I]),
    "rtw_xz_rect": (_I, [_P, _D, _D, _D, _D, _D, _I]),
    "rtw_yz_rect": (_I, [_P, _D, _D, _D, _D, _D, _I]),
    "rtw_box": (_I, [_P, _D3, _D3, _I]),
    "rtw_flip_normals": (_I, [_P, _I]),
    "rtw_translate": (_I, [_P, _I, _D, _D, _D]),
    "rtw_rotate_y": (_I, [_P, _I, _D]),
    "rtw_rotate_y_sincos": (_I, [_P, _I, _D, _D]),
    "rtw_constant_medium": (_I, [_P, _I, _D, _I]),
    "rtw_hittable_list": (_I, [_P, C.POINTER(_I), C.c_int]),
    "rtw_bvh_node": (_I, [_P, C.POINTER(_I), C.c_int, _D, _D]),
    "rtw_bvh_node_import": (_I, [_P, _I, _I, _D3, _D3]),
    "rtw_bounding_box": (C.c_int, [_P, _I, _D, _D, _D3, _D3]),
    "rtw_set_world": (C.c_int, [_P, _I]),
    "rtw_set_lights": (C.c_int, [_P, _I]),
    "rtw_set_render_type": (C.c_int, [_P, C.c_int]),
    "rtw_set_background_type": (C.c_int, [_P, C.c_int]),
    "rtw_set_camera": (C.c_int, [_P, _D3, _D3, _D3, _D, _D, _D, _D, _D, _D]),
    "rtw_set_camera_members": (C.c_int, [_P, C.POINTER(FlatCamera)]),
    "rtw_set_perlin_tables": (C.c_int, [_P, _D3, C.POINTER(_I), C.POINTER(_I), C.POINTER(_I)]),
    "rtw_build_named_scene": (C.c_int, [_P, C.c_char_p, _D, C.c_char_p, C.c_int, C.c_int]),
    "rtw_builder_flatten": (C.POINTER(FlatScene), [_P]),
    "rtw_builder_flat_material_index": (_I, [_P, _I]),
    "rtw_flat_serialize": (C.c_size_t, [C.POINTER(FlatScene), _P, C.c_size_t]),
    "rtw_scene_create": (_P, [C.POINTER(FlatScene), C.c_int]),
    "rtw_scene_destroy": (None, [_P]),
    "rtw_render": (C.c_int, [_P, C.POINTER(RenderParams), _P, C.POINTER(RenderStats)]),
    "rtw_render_device": (C.c_int, [_P, C.POINTER(RenderParams), _P, _P]),
    "rtw_render_device_finish": (C.c_int, [_P, C.POINTER(RenderStats)]),
    "rtw_render_device_noalloc": (C.c_int, [_P, C.POINTER(RenderParams), _P, C.c_uint32, _P, C.POINTER(RenderStats)]),
    "rtw_intersect_batch": (C.c_int, [_P, _P, _P, C.c_int64, C.c_float, C.c_float]),
    "rtw_intersect_batch_device": (C.c_int, [_P, _P, _P, C.c_int64, C.c_float, C.c_float, _P]),
    "rtw_tonemap": (C.c_int, [_P, C.c_int, C.c_int, _P]),
    "rtw_write_ppm": (C.c_int, [C.c_char_p, _P, C.c_int, C.c_int]),
    "rtw_write_ppm_binary": (C.c_int, [C.c_char_p, _P, C.c_int, C.c_int]),
    "rtw_write_png": (C.c_int, [C.c_char_p, _P, C.c_int, C.c_int]),
    "rtw_decode_jpeg": (C.c_int, [_P, C.c_size_t, _P, C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "rtw_comm_unique_id": (C.c_int, [_P]),
    "rtw_comm_create": (_P, [_P, C.c_int, C.c_int, C.c_int]),
    "rtw_comm_adopt": (_P, [_P, C.c_int, C.c_int, C.c_int]),
    "rtw_comm_destroy": (None, [_P]),
    "rtw_comm_rank": (C.c_int, [_P]),
    "rtw_comm_world": (C.c_int, [_P]),
    "rtw_shard_range": (C.c_int, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "rtw_reduce_accumulators": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, _P]),
    "rtw_render_sharded": (C.c_int, [_P, _P, C.POINTER(RenderParams), C.c_int, _P, C.POINTER(RenderStats)]),
    "rtw_philox4x32_10_batch": (C.c_int, [C.c_int, _P, C.c_int64, C.c_uint64, _P]),
    "rtw_sampler_batch": (C.c_int, [C.c_int, C.c_int, C.c_int64, C.c_uint64, _P]),
    "rtw_device_count": (C.c_int, []),
    "rtw_measure_fp32_peak": (C.c_int, [C.c_int, C.POINTER(C.c_double)]),
}

_lib = None


def load() -> C.CDLL:
    """dlopen librtw_b200.so (built in-tree by raytracingweekend_b200/build.py) and type every entry point."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m raytracingweekend_b200.build` "
            "(or __graft_entry__.build()). There is no CPU fallback for the render path.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        if os.environ.get("RTW_B200_LIB") and not hasattr(lib, name):
            continue  # an experimental / older build named explicitly (tools/ab_libs.sh): bind what it has
        fn = getattr(lib, name)  # AttributeError = ABI mismatch: fail loudly
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class RtwError(RuntimeError):
    pass


def check(rc: int, what: str) -> None:
    if rc != RTW_OK:
        raise RtwError(f"{what} failed (status {rc}): {load().rtw_last_error().decode()}")
