"""A/B helper: render a few scenes with the current environment (e.g. RTW_MEGA_SORTED=1) and dump sums + ray counts.

    python tools/ab_dump.py out.npz          # render
    python tools/ab_dump.py a.npz b.npz      # compare two dumps
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

CASES = [("cornell_box_aluminum", 96, 64, 16, 100), ("cornell_box_aluminum", 37, 21, 3, 100), ("random_balls_scene", 48, 32, 4, 100),
         ("two_perlin_spheres_earth", 40, 40, 8, 100), ("book2_final_scene", 32, 32, 4, 50), ("cornell_smoke_scene", 32, 32, 8, 100)]

if len(sys.argv) == 2:
    import raytracingweekend_b200 as r

    out = {}
    for name, nx, ny, spp, depth in CASES:
        sc = r.Scene.named(name, nx / ny)
        acc, st = sc.render(nx, ny, spp, max_depth=depth, seed=21)
        acc2, _ = sc.render(nx, ny, spp, max_depth=depth, seed=21)
        key = f"{name}_{nx}x{ny}x{spp}"
        out[key] = acc
        out[key + "_rays"] = np.array([st["rays"], int(np.array_equal(acc, acc2, equal_nan=True))])
    np.savez(sys.argv[1], **out)
else:
    a, b = np.load(sys.argv[1]), np.load(sys.argv[2])
    for k in a.files:
        if k.endswith("_rays"):
            continue
        x, y = a[k], b[k]
        fin = np.isfinite(x).all(axis=2) & np.isfinite(y).all(axis=2)
        close = np.isclose(x, y, rtol=1e-4, atol=1e-4).all(axis=2) | ~fin
        print(f"{k:44s} rays {a[k + '_rays'][0]} / {b[k + '_rays'][0]}  reproducible {a[k + '_rays'][1]} / {b[k + '_rays'][1]}  "
              f"nonfinite {(~np.isfinite(x).all(axis=2)).sum()} / {(~np.isfinite(y).all(axis=2)).sum()}  close {close.mean():.4f}  max|d| {np.abs(x[fin] - y[fin]).max():.3g}")
