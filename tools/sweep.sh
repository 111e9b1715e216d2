# Full-size sweep of the BASELINE.json configs and the other scenes on one GPU (event-timed, no profiler)
for s in "random_balls_scene 1200 800 100" "two_perlin_spheres_earth 800 800 256" "book2_final_scene 800 800 128" "cornell_box_aluminum 1024 1024 1024" "cornell_smoke_scene 600 600 200" "checker_spheres_scene 800 400 200" "cornell_box_scene 1024 1024 256" "dielectric_scene 800 400 200" "light_sample 800 400 200"; do
  python tools/prof_render.py $s 3 2>&1 | tail -1
done
