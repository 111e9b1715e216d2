"""Generate the golden fixtures under tests/golden/ from the UNMODIFIED reference (oracle/_ref/ref_harness).

Run here (container with /root/reference), after `make -C oracle ref`:
    python tests/golden/make_golden.py [--images]

Writes:
    earth_256x256_rgb8.bin     RGB8 texels of the reference's earth.jpg (Pillow decode; SHA-256 pinned in tests)
    funcs.npz                  per-function sequences + Perlin tables (ref_harness funcs / perlin)
    kat.json                   CppTest/unittest1.cpp assertions evaluated by the reference's vec3/ray/aabb
    flat_sha256.json           SHA-256 of the adapter-flattened reference scenes (aspect 1.5)
    hits_<scene>.npz           fixed ray batch + the reference's world.hit() records (float64)
    serial_<scene>.npz         serial (1-thread) reference render, tiny: bit-comparable with the oracle
    image_<scene>.npz          (--images) converged linear radiance: per-pixel sum and sum of squares, float64
"""
from __future__ import annotations

import hashlib
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_lib as ol  # noqa: E402
import raytracingweekend_b200 as rtw  # noqa: E402

SCENES = rtw.SCENE_NAMES
ASPECT = 1.5
SERIAL = {"cornell_box_scene": (32, 32, 4), "cornell_box_aluminum": (24, 24, 4), "dielectric_scene": (32, 16, 4),
          "random_balls_scene": (24, 16, 2), "light_sample": (24, 24, 4), "two_perlin_spheres_earth": (24, 24, 4),
          "book2_final_scene": (24, 24, 4), "cornell_smoke_scene": (24, 24, 4), "checker_spheres_scene": (24, 24, 4),
          "cornell_two_boxes": (24, 24, 4), "cornell_aluminum_box": (24, 24, 4)}
# converged images: (nx, ny, spp)
# north_star: "converged images (>= 4096 spp)"; serial renders (1 thread) so that the samples are independent draws of
# one engine sequence — the parallel reference shares its engines across threads without synchronisation
IMAGES = {"cornell_box_scene": (64, 64, 4096), "cornell_box_aluminum": (64, 64, 4096), "dielectric_scene": (64, 32, 4096),
          "random_balls_scene": (48, 32, 4096), "light_sample": (48, 48, 4096), "two_perlin_spheres_earth": (48, 48, 4096),
          "book2_final_scene": (48, 48, 4096), "cornell_smoke_scene": (48, 48, 4096), "checker_spheres_scene": (48, 48, 4096),
          "cornell_two_boxes": (64, 64, 4096), "cornell_aluminum_box": (64, 64, 4096)}
# the same at the aspect ratio and an integer box-filter factor of the BASELINE.json shapes whose full-size tests need one
# (tests/test_gpu_parity.py test_full_size_*): file image_<scene>_<nx>x<ny>.npz.  cfg1 (1200x800 -> 48x32) and cfg4
# (1024x1024 -> 64x64) use the per-scene files above.
IMAGES_SHAPED = {"two_perlin_spheres_earth": (40, 40, 4096),   # cfg2 800x800, factor 20
                 "book2_final_scene": (40, 40, 4096),          # cfg3 800x800, factor 20
                 "cornell_box_aluminum": (64, 36, 4096)}       # cfg5 3840x2160, factor 60


fixed_rays = ol.fixed_rays  # the SURVEY 8d batch construction, shared with tests/test_gpu_parity.py


def builder_for(name: str, aspect: float):
    earth = rtw.load_earth() if name in ol.EARTH_SCENES else None
    return rtw.Builder().named_scene(name, aspect, earth)


def write_image(s: str, shaped: bool = False):
    nx, ny, spp = (IMAGES_SHAPED if shaped else IMAGES)[s]
    t0 = time.time()
    rs, rq, info = ol.ref_render(s, nx, ny, spp, threads=1)
    stem = f"image_{s}_{nx}x{ny}.npz" if shaped else f"image_{s}.npz"
    np.savez_compressed(os.path.join(HERE, stem), sum=rs, sumsq=rq, nx=nx, ny=ny, spp=spp,
                        paths_per_s=info["paths_per_s"])
    print(s, "image", nx, ny, spp, f"{time.time() - t0:.1f}s", info, flush=True)


def main():
    only = [a.split("=", 1)[1].split(",") for a in sys.argv if a.startswith("--only=")]
    global SCENES
    if only:  # regenerate the per-scene files of these scenes only (flat_sha256.json is merged)
        SCENES = tuple(only[0])
    assert ol.have_ref(), "build oracle/_ref/ref_harness first (make -C oracle ref)"
    earth_jpg = "/root/reference/RayTracingWeekend/earth.jpg"
    if os.path.exists(earth_jpg):
        from PIL import Image
        a = np.asarray(Image.open(earth_jpg).convert("RGB"), dtype=np.uint8)
        a.tofile(os.path.join(HERE, "earth_256x256_rgb8.bin"))
        print("earth sha256", hashlib.sha256(a.tobytes()).hexdigest())
    if not only:
        np.savez_compressed(os.path.join(HERE, "funcs.npz"), funcs=ol.ref_funcs(), perlin=ol.ref_perlin())
        with open(os.path.join(HERE, "kat.json"), "w") as f:
            json.dump(json.loads(ol.ref_run(["kat"])), f, indent=1)
    if "--shaped-only" in sys.argv:
        for s in SCENES:
            if s in IMAGES_SHAPED:
                write_image(s, shaped=True)
        return
    if "--images-only" in sys.argv:  # regenerate image_<scene>.npz alone (python make_golden.py --images-only --only=a,b)
        for s in SCENES:
            write_image(s)
        return
    sha = {}
    if only and os.path.exists(os.path.join(HERE, "flat_sha256.json")):
        sha = json.load(open(os.path.join(HERE, "flat_sha256.json")))
    sha.update({s: hashlib.sha256(ol.ref_flat(s, ASPECT)).hexdigest() for s in SCENES})
    with open(os.path.join(HERE, "flat_sha256.json"), "w") as f:
        json.dump(sha, f, indent=1)
    for s in SCENES:
        b = builder_for(s, ASPECT)
        rays = fixed_rays(b, 4000, 1)
        hits = ol.ref_hits(s, ASPECT, rays)
        np.savez_compressed(os.path.join(HERE, f"hits_{s}.npz"), rays=rays, hits=hits, aspect=ASPECT)
        nx, ny, spp = SERIAL[s]
        rs, rq, info = ol.ref_render(s, nx, ny, spp, threads=1)
        np.savez_compressed(os.path.join(HERE, f"serial_{s}.npz"), sum=rs, sumsq=rq, nx=nx, ny=ny, spp=spp)
        print(s, "hits", hits[:, 0].mean().round(3), "serial ok")
    if "--images" in sys.argv:
        for s in SCENES:
            write_image(s)


if __name__ == "__main__":
    main()
