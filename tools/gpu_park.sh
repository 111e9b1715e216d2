cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out; O=gpurun_out
timeout 300 python - <<'PY' > $O/park_check.txt 2>&1
import os, subprocess, sys, numpy as np
code = r"""
import sys, numpy as np, raytracingweekend_b200 as rtw
sc = rtw.Scene.named("book2_final_scene", 1.0)
out = {}
for nx, ny, spp in ((96, 96, 8), (37, 21, 6), (200, 200, 16)):
    a, st = sc.render(nx, ny, spp, seed=11)
    out[f"{nx}_{ny}_{spp}"] = a; out[f"{nx}_{ny}_{spp}_rays"] = np.array([st["rays"]])
np.savez(sys.argv[1], **out)
"""
for v in ("1", "0"):
    subprocess.run([sys.executable, "-c", code, f"/tmp/park{v}.npz"], check=True, env=dict(os.environ, RTW_PARK=v), timeout=200)
a, b = np.load("/tmp/park1.npz"), np.load("/tmp/park0.npz")
for k in a.files:
    print(k, "identical" if np.array_equal(a[k], b[k], equal_nan=True) else "DIFFERENT", a[k].shape)
PY
echo "check rc=$?"; cat $O/park_check.txt
for v in 0 1; do echo "== RTW_PARK=$v"; for s in "book2_final_scene 800 800 128" "book2_final_scene 800 800 32"; do RTW_PARK=$v timeout 200 python tools/prof_render.py $s 3 2>&1 | tail -1; done; done 2>&1 | tee $O/ab_park.txt
