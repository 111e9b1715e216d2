import os, sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, oracle_lib as ol, raytracingweekend_b200 as rtw
for scene in ("book2_final_scene", "two_perlin_spheres_earth"):
    b = rtw.Builder().named_scene(scene, 1.5, rtw.load_earth()); sc = rtw.Scene(b)
    rays = ol.fixed_rays(b, 1 << 21, 2)
    gpu = sc.intersect_batch(rays.astype(np.float32)); orc = ol.oracle_hits(b, rays)
    oh, gh = orc[:, 0] == 1, gpu["prim"] >= 0
    both = oh & gh & (gpu["prim"] == orc[:, 11].astype(np.int32))
    rel = np.abs(gpu["t"][both] - orc[both, 1]) / np.abs(orc[both, 1])
    nerr = np.abs(np.stack([gpu["nx"][both], gpu["ny"][both], gpu["nz"][both]], 1) - orc[both, 5:8]).max()
    print(scene, "flips", int((oh != gh).sum()), "idmis", int((oh & gh).sum() - both.sum()), "t_rel max", rel.max(), ">1e-5:", int((rel > 1e-5).sum()), "n_err", nerr, flush=True)
