"""Host-side mirror of the reference's scene-construction interface over the C ABI.

`Builder` has one method per reference constructor (same names and argument order as
RayTracingWeekend/*.h — see include/rtw_b200.h for the file:line of each); `Scene` is the device-resident
flattened scene; `Scene.render*` replace the `_for` nest around color() (RayTracingWeekend.cpp:211-250).
Python is plumbing only: every call lands in librtw_b200.so.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import numpy as np

from . import _lib
from ._lib import RenderParams, RenderStats, check

EARTH_TEXELS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "earth_256x256_rgb8.bin")

SCENE_NAMES = ("light_sample", "dielectric_scene", "random_balls_scene", "cornell_box_scene", "cornell_box_aluminum",
               "two_perlin_spheres_earth", "book2_final_scene", "cornell_smoke_scene", "checker_spheres_scene", "cornell_two_boxes",
               "cornell_aluminum_box")


def _d3(v: Sequence[float]):
    return (C.c_double * 3)(*[float(x) for x in v])


class Builder:
    """Scene graph under construction (rtw_builder)."""

    def __init__(self):
        self._l = _lib.load()
        self._b = self._l.rtw_builder_create()
        if not self._b:
            raise _lib.RtwError("rtw_builder_create failed")
        self._keep = []

    def close(self):
        if self._b:
            self._l.rtw_builder_destroy(self._b)
            self._b = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _id(self, v: int, what: str) -> int:
        if v < 0:
            raise _lib.RtwError(f"{what}: {self._l.rtw_last_error().decode()}")
        return v

    # textures (texture.h)
    def constant_texture(self, r, g, b): return self._id(self._l.rtw_constant_texture(self._b, r, g, b), "constant_texture")
    def checker_texture(self, t0, t1): return self._id(self._l.rtw_checker_texture(self._b, t0, t1), "checker_texture")
    def noise_texture(self, scale=5.0): return self._id(self._l.rtw_noise_texture(self._b, scale), "noise_texture")

    def image_texture(self, rgb8: np.ndarray, nx: int, ny: int):
        buf = np.ascontiguousarray(rgb8, dtype=np.uint8).tobytes()
        if len(buf) != 3 * nx * ny:
            raise ValueError("image_texture: need nx*ny*3 bytes")
        return self._id(self._l.rtw_image_texture(self._b, buf, nx, ny), "image_texture")

    # materials (material.h)
    def lambertian(self, tex): return self._id(self._l.rtw_lambertian(self._b, tex), "lambertian")
    def metal(self, albedo, fuzz): return self._id(self._l.rtw_metal(self._b, albedo[0], albedo[1], albedo[2], fuzz), "metal")
    def dielectric(self, ri): return self._id(self._l.rtw_dielectric(self._b, ri), "dielectric")
    def diffuse_light(self, tex): return self._id(self._l.rtw_diffuse_light(self._b, tex), "diffuse_light")
    def isotropic(self, tex): return self._id(self._l.rtw_isotropic(self._b, tex), "isotropic")

    # hittables (sphere.h, hittable.h, hittable_list.h)
    def sphere(self, cen, r, mat): return self._id(self._l.rtw_sphere(self._b, cen[0], cen[1], cen[2], r, mat), "sphere")

    def moving_sphere(self, cen, r, mat, center1, time0, time1):
        return self._id(self._l.rtw_moving_sphere(self._b, cen[0], cen[1], cen[2], r, mat, center1[0], center1[1], center1[2],
                                                  time0, time1), "moving_sphere")

    def xy_rect(self, x0, x1, y0, y1, k, mat): return self._id(self._l.rtw_xy_rect(self._b, x0, x1, y0, y1, k, mat), "xy_rect")
    def xz_rect(self, x0, x1, z0, z1, k, mat): return self._id(self._l.rtw_xz_rect(self._b, x0, x1, z0, z1, k, mat), "xz_rect")
    def yz_rect(self, y0, y1, z0, z1, k, mat): return self._id(self._l.rtw_yz_rect(self._b, y0, y1, z0, z1, k, mat), "yz_rect")
    def box(self, p0, p1, mat): return self._id(self._l.rtw_box(self._b, _d3(p0), _d3(p1), mat), "box")
    def flip_normals(self, h): return self._id(self._l.rtw_flip_normals(self._b, h), "flip_normals")
    def translate(self, h, off): return self._id(self._l.rtw_translate(self._b, h, off[0], off[1], off[2]), "translate")
    def rotate_y(self, h, angle): return self._id(self._l.rtw_rotate_y(self._b, h, angle), "rotate_y")

    def constant_medium(self, boundary, density, mat):
        return self._id(self._l.rtw_constant_medium(self._b, boundary, density, mat), "constant_medium")

    def hittable_list(self, objects: Sequence[int]):
        arr = (C.c_int32 * max(1, len(objects)))(*objects)
        return self._id(self._l.rtw_hittable_list(self._b, arr, len(objects)), "hittable_list")

    def bvh_node(self, objects: Sequence[int], time0=0.0, time1=1.0):
        arr = (C.c_int32 * len(objects))(*objects)
        return self._id(self._l.rtw_bvh_node(self._b, arr, len(objects), time0, time1), "bvh_node")

    def bounding_box(self, h, t0=0.0, t1=1.0):
        mn, mx = (C.c_double * 3)(), (C.c_double * 3)()
        r = self._l.rtw_bounding_box(self._b, h, t0, t1, mn, mx)
        if r < 0:
            raise _lib.RtwError("bounding_box: " + self._l.rtw_last_error().decode())
        return bool(r), list(mn), list(mx)

    # scene (Scene/scene.h) + camera (camera.h:13)
    def set_world(self, lst): check(self._l.rtw_set_world(self._b, lst), "set_world")
    def set_lights(self, lst): check(self._l.rtw_set_lights(self._b, -1 if lst is None else lst), "set_lights")
    def set_render_type(self, t): check(self._l.rtw_set_render_type(self._b, t), "set_render_type")
    def set_background_type(self, t): check(self._l.rtw_set_background_type(self._b, t), "set_background_type")

    def camera(self, lookfrom, lookat, vup, vfov, aspect, aperture, focus_dist, t0, t1):
        check(self._l.rtw_set_camera(self._b, _d3(lookfrom), _d3(lookat), _d3(vup), vfov, aspect, aperture, focus_dist, t0, t1), "camera")

    def named_scene(self, name: str, aspect: float, earth: Optional[np.ndarray] = None):
        """Scene/scene.h functions composed natively (rtw_build_named_scene)."""
        buf, nx, ny = None, 0, 0
        if earth is not None:
            e = np.ascontiguousarray(earth, dtype=np.uint8)
            ny, nx = e.shape[0], e.shape[1]
            buf = e.tobytes()
        check(self._l.rtw_build_named_scene(self._b, name.encode(), aspect, buf, nx, ny), f"named_scene({name})")
        return self

    def flatten(self):
        fs = self._l.rtw_builder_flatten(self._b)
        if not fs:
            raise _lib.RtwError("flatten: " + self._l.rtw_last_error().decode())
        return fs

    def serialize(self) -> bytes:
        fs = self.flatten()
        n = self._l.rtw_flat_serialize(fs, None, 0)
        buf = C.create_string_buffer(n)
        self._l.rtw_flat_serialize(fs, buf, n)
        return buf.raw


def load_earth(path: Optional[str] = None) -> np.ndarray:
    """RGB8 texels for image_texture (texture.h:73-99 takes caller-decoded bytes; the reference ships earth.jpg but no loader).
    `path`: a baseline JPEG, decoded by this library's own rtw_decode_jpeg (e.g. the reference's RayTracingWeekend/earth.jpg);
    None: the packaged 256x256 texels of that file (data/earth_256x256_rgb8.bin — byte-identical to decoding it, pinned by
    tests/test_imageio.py), so that the cfg2 / cfg3 scenes render where the reference tree is absent."""
    if path is not None:
        with open(path, "rb") as f:
            return decode_jpeg(f.read())
    return np.fromfile(EARTH_TEXELS, dtype=np.uint8).reshape(256, 256, 3)


class Scene:
    """Device-resident flattened scene (rtw_scene); immutable after creation."""

    def __init__(self, builder: Builder, device: int = 0):
        self._l = _lib.load()
        self.device = device
        fs = builder.flatten()
        self.n_prims = fs.contents.n_prims
        self._s = self._l.rtw_scene_create(fs, device)
        if not self._s:
            raise _lib.RtwError("rtw_scene_create: " + self._l.rtw_last_error().decode())

    @classmethod
    def named(cls, name: str, aspect: float, device: int = 0) -> "Scene":
        earth = load_earth() if name in ("two_perlin_spheres_earth", "book2_final_scene") else None
        b = Builder().named_scene(name, aspect, earth)
        try:
            return cls(b, device)
        finally:
            b.close()

    def close(self):
        if self._s:
            self._l.rtw_scene_destroy(self._s)
            self._s = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @staticmethod
    def _params(nx, ny, spp, spp_begin, max_depth, seed, flags) -> RenderParams:
        return RenderParams(nx, ny, spp_begin, spp, max_depth, flags, seed)

    def render(self, nx, ny, spp, spp_begin=0, max_depth=100, seed=0, flags=0, out: Optional[np.ndarray] = None):
        """Host-buffer render through rtw_render: returns (accum[ny,nx,4] float32 sums, stats dict)."""
        acc = out if out is not None else np.empty((ny, nx, 4), dtype=np.float32)
        p, st = self._params(nx, ny, spp, spp_begin, max_depth, seed, flags), RenderStats()
        check(self._l.rtw_render(self._s, C.byref(p), acc.ctypes.data_as(C.c_void_p), C.byref(st)), "rtw_render")
        return acc, _stats(st)

    def render_device(self, accum_ptr: int, nx, ny, spp, spp_begin=0, max_depth=100, seed=0, flags=0, stream: int = 0):
        """Asynchronous render into a device buffer (e.g. torch tensor .data_ptr()) on `stream`."""
        p = self._params(nx, ny, spp, spp_begin, max_depth, seed, flags)
        check(self._l.rtw_render_device(self._s, C.byref(p), C.c_void_p(accum_ptr), C.c_void_p(stream)), "rtw_render_device")

    def render_sharded(self, comm, nx, ny, spp, spp_begin=0, max_depth=100, seed=0, flags=0, root=0,
                       out: Optional[np.ndarray] = None):
        """rtw_render_sharded: this rank renders its share of the `spp` samples, the sums are reduced onto `root`
        (ncclReduce inside the library), and the root alone gets the image.  `comm` is a dist.Communicator (or None for a
        single rank).  Returns (accum on the root / None elsewhere, stats of the local shard)."""
        rank = comm.rank if comm is not None else 0
        acc = None
        if rank == root:
            acc = out if out is not None else np.empty((ny, nx, 4), dtype=np.float32)
        p, st = self._params(nx, ny, spp, spp_begin, max_depth, seed, flags), RenderStats()
        check(self._l.rtw_render_sharded(self._s, comm.handle if comm is not None else None, C.byref(p), root,
                                         acc.ctypes.data_as(C.c_void_p) if acc is not None else None, C.byref(st)), "rtw_render_sharded")
        return acc, _stats(st)

    def render_device_noalloc(self, accum_ptr: int, nx, ny, spp, spp_begin=0, max_depth=100, seed=0, flags=0, pool_slots: int = 0,
                              stream: int = 0) -> dict:
        """rtw_render_device_noalloc: the wavefront render on the library's static device scratch (allocates nothing;
        pool_slots <= RTW_NOALLOC_MAX_POOL_SLOTS, 0 = largest useful).  Synchronous; returns the stats."""
        p, st = self._params(nx, ny, spp, spp_begin, max_depth, seed, flags), RenderStats()
        check(self._l.rtw_render_device_noalloc(self._s, C.byref(p), C.c_void_p(accum_ptr), pool_slots, C.c_void_p(stream), C.byref(st)),
              "rtw_render_device_noalloc")
        return _stats(st)

    def finish(self) -> dict:
        st = RenderStats()
        check(self._l.rtw_render_device_finish(self._s, C.byref(st)), "rtw_render_device_finish")
        return _stats(st)

    def intersect_batch(self, rays: np.ndarray, t_min=0.001, t_max=3.4028234663852886e38):
        """rays: [n,7] or [n,8] float32 (o, d, time). Returns structured array of rtw_hit."""
        r = np.zeros((rays.shape[0], 8), dtype=np.float32)
        r[:, :7] = rays[:, :7]
        hits = np.zeros(rays.shape[0], dtype=HIT_DTYPE)
        check(self._l.rtw_intersect_batch(self._s, r.ctypes.data_as(C.c_void_p), hits.ctypes.data_as(C.c_void_p), r.shape[0],
                                          t_min, t_max), "rtw_intersect_batch")
        return hits


HIT_DTYPE = np.dtype([("prim", "<i4"), ("material", "<i4"), ("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("nx", "<f4"),
                      ("ny", "<f4"), ("nz", "<f4")])


def _stats(st: RenderStats) -> dict:
    return {"paths": int(st.paths), "rays": int(st.rays), "kernel_ms": float(st.kernel_ms), "kernel_launches": int(st.kernel_launches)}


def tonemap(accum: np.ndarray) -> np.ndarray:
    """RayTracingWeekend.cpp:241-244 via rtw_tonemap: [ny,nx,4] float32 sums -> [ny,nx,3] float64 canvas."""
    l = _lib.load()
    ny, nx = accum.shape[:2]
    a = np.ascontiguousarray(accum, dtype=np.float32)
    out = np.empty((ny, nx, 3), dtype=np.float64)
    check(l.rtw_tonemap(a.ctypes.data_as(C.c_void_p), nx, ny, out.ctypes.data_as(C.c_void_p)), "rtw_tonemap")
    return out


def write_ppm_binary(path: str, canvas: np.ndarray) -> None:
    """"P6" twin of write_ppm (same rows, same quantisation) via rtw_write_ppm_binary."""
    l = _lib.load()
    ny, nx = canvas.shape[:2]
    c = np.ascontiguousarray(canvas, dtype=np.float64)
    check(l.rtw_write_ppm_binary(path.encode(), c.ctypes.data_as(C.c_void_p), nx, ny), "rtw_write_ppm_binary")


def write_png(path: str, canvas: np.ndarray) -> None:
    """8-bit RGB PNG of the canvas (replaces the reference's `magick 1.ppm 1.png`, RayTracingWeekend.cpp:285)."""
    l = _lib.load()
    ny, nx = canvas.shape[:2]
    c = np.ascontiguousarray(canvas, dtype=np.float64)
    check(l.rtw_write_png(path.encode(), c.ctypes.data_as(C.c_void_p), nx, ny), "rtw_write_png")


def decode_jpeg(data: bytes) -> np.ndarray:
    """Baseline JPEG -> [ny, nx, 3] uint8, top row first: the byte array image_texture takes (texture.h:73-99)."""
    l = _lib.load()
    buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
    nx, ny = C.c_int(), C.c_int()
    check(l.rtw_decode_jpeg(buf, len(data), None, 0, C.byref(nx), C.byref(ny)), "rtw_decode_jpeg")
    out = np.empty((ny.value, nx.value, 3), dtype=np.uint8)
    check(l.rtw_decode_jpeg(buf, len(data), out.ctypes.data_as(C.c_void_p), out.size, C.byref(nx), C.byref(ny)), "rtw_decode_jpeg")
    return out


def write_ppm(path: str, canvas: np.ndarray) -> None:
    """RayTracingWeekend.cpp:252-275 via rtw_write_ppm."""
    l = _lib.load()
    ny, nx = canvas.shape[:2]
    c = np.ascontiguousarray(canvas, dtype=np.float64)
    check(l.rtw_write_ppm(path.encode(), c.ctypes.data_as(C.c_void_p), nx, ny), "rtw_write_ppm")
