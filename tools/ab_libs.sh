# A/B of tagged library builds: bash tools/ab_libs.sh "<tag1> <tag2> ..." ["scene nx ny spp" ...]   ('' = the default build)
# Build a tagged library with: python -c "from raytracingweekend_b200.build import build; build(extra=['-DX=1'], tag='_x')"
tags="$1"; shift
if [ $# -eq 0 ]; then set -- "cornell_box_aluminum 1024 1024 1024" "cornell_box_scene 1024 1024 256" "two_perlin_spheres_earth 800 800 256" "random_balls_scene 1200 800 100" "book2_final_scene 800 800 128"; fi
for t in $tags; do
  [ "$t" = "base" ] && t=""
  echo "== librtw_b200$t.so"
  for s in "$@"; do
    RTW_B200_LIB=$PWD/raytracingweekend_b200/lib/librtw_b200$t.so timeout 60 python tools/prof_render.py $s 4 2>&1 | tail -1
  done
done
