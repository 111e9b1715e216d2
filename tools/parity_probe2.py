"""Where the tails of the 2^21-ray batch and of the path-by-path comparison come from (diagnostics, GPU)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import oracle_lib as ol
import raytracingweekend_b200 as rtw

EPS = 2.0 ** -23
for scene in ("cornell_two_boxes", "book2_final_scene", "checker_spheres_scene", "random_balls_scene", "cornell_box_aluminum"):
    earth = rtw.load_earth() if scene in ol.EARTH_SCENES else None
    b = rtw.Builder().named_scene(scene, 1.5, earth)
    sc = rtw.Scene(b)
    rays = ol.fixed_rays(b, 1 << 21, 2)
    gpu = sc.intersect_batch(rays.astype(np.float32))
    orc = ol.oracle_hits(b, rays)
    both = (orc[:, 0] == 1) & (gpu["prim"] >= 0) & (gpu["prim"] == orc[:, 11].astype(np.int32))
    half = len(rays) // 2
    m = int(np.sqrt(half)) ** 2
    for name, sl in (("pinhole", slice(0, m)), ("random", slice(m, None))):
        idx = np.nonzero(both[sl])[0] + (sl.start or 0)
        t, tg = orc[idx, 1], gpu["t"][idx].astype(np.float64)
        rel = np.abs(tg - t) / np.abs(t)
        o, d = rays[idx, :3], rays[idx, 3:6]
        p = orc[idx, 2:5]
        dn = np.linalg.norm(d, axis=1)
        scale = np.maximum(np.abs(o).max(axis=1), np.abs(p).max(axis=1))
        gran = np.abs(tg - t) * dn / (EPS * scale)  # error in units of the coordinate granularity
        bad = rel > 1e-5
        nerr = np.abs(np.stack([gpu["nx"][idx], gpu["ny"][idx], gpu["nz"][idx]], 1) - orc[idx, 5:8]).max(axis=1)
        print(f"{scene} {name}: n={len(idx)} t_rel max {rel.max():.2e} >1e-5: {bad.sum()} ({bad.mean():.2e}); |dt||d|/(eps*scale) max {gran.max():.2f} "
              f"(over bad: {gran[bad].max() if bad.any() else 0:.2f}); t over bad: min {t[bad].min() if bad.any() else 0:.3g} max {t[bad].max() if bad.any() else 0:.3g}; "
              f"n_err max {nerr.max():.2e} at scale {scale[nerr.argmax()]:.0f} t {t[nerr.argmax()]:.0f}", flush=True)
    sc.close()

# path by path: book2
scene, nx, ny, spp = "book2_final_scene", 32, 32, 4
b = rtw.Builder().named_scene(scene, 1.0, rtw.load_earth())
sc = rtw.Scene(b)
acc, st = sc.render(nx, ny, spp, seed=11)
osum, _, orays = ol.oracle_render(b, nx, ny, spp, mode="philox", seed=11)
g = acc[..., :3].astype(np.float64)
div = ~np.isclose(g, osum, rtol=2e-3, atol=1e-3 * spp).all(axis=2)
for (j, i) in np.argwhere(div):
    print("div pixel", i, j, "gpu", g[j, i].round(4), "oracle", osum[j, i].round(4), "rel", (np.abs(g[j, i] - osum[j, i]) / (np.abs(osum[j, i]) + 1e-9)).max().round(4))
for e in (-22, -20, -18, -16):
    fr, sens = [], np.zeros_like(div)
    for k in range(8):
        ol.set_jitter(2.0 ** e, k)
        js, _, _ = ol.oracle_render(b, nx, ny, spp, mode="philox", seed=11)
        dd = ~np.isclose(js, osum, rtol=2e-3, atol=1e-3 * spp).all(axis=2)
        sens |= dd
        fr.append(dd.mean())
    ol.set_jitter(0.0)
    print(f"jitter 2^{e}: mean frac {np.mean(fr):.4f}, union {sens.mean():.4f}, gpu-divergent outside union {(div & ~sens).sum()} of {div.sum()}")
