# A/B of the render-tree variants.  Build the tagged libraries first:
#   python -c "from raytracingweekend_b200.build import build; build(extra=['-DRTW_BVH4=2'], tag=''); build(extra=['-DRTW_BVH4=0'], tag='_bvh2'); build(extra=['-DRTW_BVH4=2','-DRTW_BVH4_SORT=1'], tag='_sort')"
for lib in librtw_b200.so librtw_b200_bvh2.so librtw_b200_nosort.so librtw_b200_nooct.so; do
  echo "== $lib"
  for s in "random_balls_scene 1200 800 100" "book2_final_scene 800 800 128" "cornell_smoke_scene 600 600 200"; do
    RTW_B200_LIB=$PWD/raytracingweekend_b200/lib/$lib python tools/prof_render.py $s 3 2>&1 | tail -1
  done
done
ncu --set full --clock-control none --import-source on -k regex:render_sorted -s 1 -c 1 -o gpurun_out/sorted_book2_bvh4 python tools/prof_render.py book2_final_scene 800 800 32 2 > gpurun_out/ncu3.log 2>&1
