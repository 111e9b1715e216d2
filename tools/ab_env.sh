# A/B by environment switch or tagged library: bash tools/ab_env.sh "<label>=<ENV=VAL ...>" ...   (label alone = defaults)
# each case renders the standard set of scenes with tools/prof_render.py (event-timed, 3 launches, last line kept)
SCENES=("cornell_box_aluminum 1024 1024 1024" "cornell_box_aluminum 1024 1024 128" "cornell_box_scene 1024 1024 256" "cornell_smoke_scene 600 600 200" "random_balls_scene 1200 800 100" "two_perlin_spheres_earth 800 800 256" "book2_final_scene 800 800 128")
for c in "$@"; do
  label="${c%%=*}"; envs=""; [ "$c" != "$label" ] && envs="${c#*=}"
  echo "== $label ($envs)"
  for s in "${SCENES[@]}"; do
    env $envs timeout 120 python tools/prof_render.py $s 3 2>&1 | tail -1
  done
done
