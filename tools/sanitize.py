"""Small renders of every kernel variant / pipeline for `compute-sanitizer --tool memcheck python tools/sanitize.py`."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytracingweekend_b200 as r
from raytracingweekend_b200 import _lib

for name in ("cornell_box_aluminum", "two_perlin_spheres_earth", "random_balls_scene", "book2_final_scene", "cornell_smoke_scene"):
    sc = r.Scene.named(name, 37 / 21)
    for flags, tag in ((0, "sorted"), (_lib.RTW_FLAG_WAVEFRONT, "wavefront"), (_lib.RTW_FLAG_NEE, "nee")):
        acc, st = sc.render(37, 21, 3, seed=1, flags=flags)
        print(name, tag, st["rays"], bool(np.isfinite(acc[..., 3]).all()), flush=True)
    rays = np.random.default_rng(1).normal(size=(1000, 7)).astype(np.float32)
    h = sc.intersect_batch(rays)
    print(name, "intersect", int((h["prim"] >= 0).sum()), flush=True)
    sc.close()
os.environ["RTW_DEVICE_BVH"] = "1"
sc = r.Scene.named("random_balls_scene", 1.5)
print("device bvh", sc.render(24, 16, 2)[1]["rays"])
