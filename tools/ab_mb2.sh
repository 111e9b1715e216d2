for t in _mb2 _b _c _d _e; do echo "== $t"; 
for s in "book2_final_scene 800 800 128" "random_balls_scene 1200 800 100"; do
RTW_DEBUG_TREES=1 RTW_B200_LIB=$PWD/raytracingweekend_b200/lib/librtw_b200$t.so timeout 60 python tools/prof_render.py $s 3 2>&1 | grep -E "render_sorted|Gpaths" | tail -2; done; done
