"""Compare wavefront vs megakernel renders of one scene and print the largest per-pixel differences."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracingweekend_b200 as r

name = sys.argv[1] if len(sys.argv) > 1 else "book2_final_scene"
nx = ny = int(sys.argv[2]) if len(sys.argv) > 2 else 32
spp = int(sys.argv[3]) if len(sys.argv) > 3 else 4
sc = r.Scene.named(name, 1.0)
m, sm = sc.render(nx, ny, spp, seed=21)
w, sw = sc.render(nx, ny, spp, seed=21, flags=1)
print("rays", sm["rays"], sw["rays"])
d = np.abs(m - w)
rel = d / (np.abs(m) + 1e-6)
idx = np.argsort(d[..., :3].max(axis=2).ravel())[::-1][:8]
for k in idx:
    j, i = divmod(int(k), nx)
    print((i, j), "mega", m[j, i], "wave", w[j, i])
print("pixels differing > 1e-4 rel:", int((rel[..., :3].max(axis=2) > 1e-4).sum()), "of", nx * ny)
