# bash tools/gpu_ab_tags.sh "<tag> ..." ["scene nx ny spp" ...]   ('base' = default library)
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
tags="$1"; shift
keep=1; [ -n "$RTW_PROF_HASH" ] && keep=2   # RTW_PROF_HASH=1: prof_render.py also prints a hash of the image
if [ $# -eq 0 ]; then set -- "cornell_box_aluminum 1024 1024 1024" "cornell_box_aluminum 1024 1024 128" "cornell_box_aluminum 1024 1024 16"; fi
for t in $tags; do
  lib=$PWD/raytracingweekend_b200/lib/librtw_b200_$t.so; [ "$t" = "base" ] && lib=$PWD/raytracingweekend_b200/lib/librtw_b200.so
  echo "== $t"
  for s in "$@"; do RTW_B200_LIB=$lib timeout 120 python tools/prof_render.py $s 3 2>&1 | tail -$keep; done
done
