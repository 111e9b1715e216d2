cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out; O=gpurun_out
L=$PWD/raytracingweekend_b200/lib
# correctness first, small and bounded
timeout 120 python - <<'PY' > $O/persist_check.txt 2>&1
import numpy as np, raytracingweekend_b200 as rtw, os, sys
sys.path.insert(0,'tests')
import oracle_lib as ol
for name,nx,ny,spp in (("cornell_box_aluminum",64,64,8),("cornell_box_aluminum",37,21,5),("cornell_box_aluminum",1,1,7),("random_balls_scene",48,32,4),("book2_final_scene",32,32,4),("two_perlin_spheres_earth",40,40,8)):
    earth = rtw.load_earth() if name in ol.EARTH_SCENES else None
    b = rtw.Builder().named_scene(name, nx/ny, earth); sc = rtw.Scene(b)
    a,st = sc.render(nx,ny,spp,seed=11); a2,st2 = sc.render(nx,ny,spp,seed=11)
    osum,_,orays = ol.oracle_render(b,nx,ny,spp,mode="philox",seed=11)
    close = np.isclose(a[...,:3].astype(np.float64), osum, rtol=2e-3, atol=1e-3*spp).all(axis=2)
    print(name,nx,ny,spp,"repro",np.array_equal(a,a2,equal_nan=True),"w ok",bool(np.all(a[...,3]==spp)),"rays",st["rays"],orays,"path-exact",close.mean(), flush=True)
PY
echo "check rc=$?"; cat $O/persist_check.txt
ab() { for s in "cornell_box_aluminum 1024 1024 1024" "cornell_box_aluminum 1024 1024 128" "cornell_box_aluminum 1024 1024 16" "cornell_smoke_scene 600 600 200" "random_balls_scene 1200 800 100" "two_perlin_spheres_earth 800 800 256" "book2_final_scene 800 800 128"; do timeout 120 python tools/prof_render.py $s 3 2>&1 | tail -1; done; }
{ echo "== v2 (tile per block)"; RTW_B200_LIB=$L/librtw_b200_v2.so ab; echo "== v3 persistent, double-buffered tiles, 32-bit sums"; ab; } > $O/ab_persist.txt 2>&1
cat $O/ab_persist.txt
