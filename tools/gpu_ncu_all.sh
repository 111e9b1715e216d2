cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out; O=gpurun_out
cap() { name=$1; shift; timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_sorted -s 1 -c 1 -f -o $O/r2_$name python tools/prof_render.py "$@" > $O/ncu_$name.log 2>&1; echo "$name rc=$?"; }
cap cfg1_random_balls random_balls_scene 1200 800 100 2
cap cfg2_perlin_earth two_perlin_spheres_earth 800 800 256 2
cap cfg3_book2_final book2_final_scene 800 800 32 2
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-per-config > $O/ncu_bench.log 2>&1; echo "launch list rc=$?"
