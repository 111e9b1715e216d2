"""Exploratory: error distributions GPU intersect_batch vs oracle on the golden ray batches."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as ol
import raytracingweekend_b200 as rtw

for scene in rtw.SCENE_NAMES:
    g = np.load(os.path.join(ROOT, "tests", "golden", f"hits_{scene}.npz"))
    earth = rtw.load_earth() if scene in ol.EARTH_SCENES else None
    b = rtw.Builder().named_scene(scene, float(g["aspect"]), earth)
    sc = rtw.Scene(b)
    rays = g["rays"]
    gpu = sc.intersect_batch(rays.astype(np.float32))
    orc = ol.oracle_hits(b, rays)
    ohit, ghit = orc[:, 0] == 1, gpu["prim"] >= 0
    both = ohit & ghit
    same_id = gpu["prim"][both] == orc[both, 11].astype(np.int32)
    idx = np.nonzero(both)[0][same_id]
    t_rel = np.abs(gpu["t"][idx] - orc[idx, 1]) / np.abs(orc[idx, 1])
    n_err = np.abs(np.stack([gpu["nx"][idx], gpu["ny"][idx], gpu["nz"][idx]], 1) - orc[idx, 5:8]).max(1)
    uv_err = np.abs(np.stack([gpu["u"][idx], gpu["v"][idx]], 1) - orc[idx, 8:10]).max(1)
    mat_ok = (gpu["material"][idx] == orc[idx, 10]).all()
    print(f"{scene:26s} n={len(rays)} hitmiss_diff={(ohit != ghit).sum()} id_diff={(~same_id).sum()} mat_ok={mat_ok} "
          f"t_rel max={t_rel.max():.2e} p999={np.quantile(t_rel, 0.999):.2e} n_err max={n_err.max():.2e} p999={np.quantile(n_err,0.999):.2e} "
          f"uv_err max={uv_err.max():.2e} p999={np.quantile(uv_err,0.999):.2e}", flush=True)
    bad = np.nonzero(ohit != ghit)[0][:3]
    for i in bad: print("   hit/miss diff ray", i, rays[i], "oracle t", orc[i,1], "prim", orc[i,11], "gpu", gpu[i])
    bad = np.nonzero(both)[0][~same_id][:3]
    for i in bad: print("   id diff ray", i, "oracle t", orc[i,1], "prim", orc[i,11], "gpu", gpu[i])
