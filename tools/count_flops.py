"""Algorithmic FP32 flops per path/ray of the reference's math from the instrumented CPU oracle.
Prints per-scene event rates; bench.py's FLOP_PER_PATH and DESIGN.md §5 quote the Cornell numbers."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as ol  # noqa: E402
import raytracingweekend_b200 as rtw  # noqa: E402

NAMES = ("CAMERA LENS RECT_TEST RECT_ACCEPT SPHERE_TEST SPHERE_ROOT SPHERE_ACCEPT MOVING_CENTER AABB TRANSLATE ROTATE FLIP MEDIUM "
         "BACKGROUND EMITTED ONB COSINE_GEN COSINE_VALUE SCATTERING_PDF LIGHT_PICK RECT_RANDOM RECT_PDF SPHERE_RANDOM SPHERE_PDF MIXTURE "
         "LAMBERT_UPDATE SPECULAR_UPDATE METAL DIELECTRIC ISOTROPIC CHECKER IMAGE_TEX PERLIN_NOISE NOISE_TEX RAY").split()

o = ol.oracle()
o.oracle_events_read.restype = C.c_double
o.oracle_events_read.argtypes = [C.c_void_p, C.c_int]
for scene, nx, ny, spp in [("cornell_box_aluminum", 64, 64, 64), ("cornell_box_scene", 64, 64, 64), ("dielectric_scene", 64, 32, 64),
                           ("random_balls_scene", 48, 32, 8), ("light_sample", 48, 48, 32), ("two_perlin_spheres_earth", 48, 48, 32),
                           ("book2_final_scene", 32, 32, 8)]:
    earth = rtw.load_earth() if scene in ol.EARTH_SCENES else None
    b = rtw.Builder().named_scene(scene, nx / ny, earth)
    o.oracle_events_enable(1)
    _, _, rays = ol.oracle_render(b, nx, ny, spp, mode="philox", seed=1)
    cnt = np.zeros(len(NAMES), dtype=np.uint64)
    fl = o.oracle_events_read(cnt.ctypes.data, len(NAMES))
    o.oracle_events_enable(0)
    paths = nx * ny * spp
    print(f"{scene}: flop/path {fl / paths:.1f}  flop/ray {fl / rays:.1f}  rays/path {rays / paths:.3f}")
    print("    per ray:", {n: round(float(c) / rays, 3) for n, c in zip(NAMES, cnt) if c})
