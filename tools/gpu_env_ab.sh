# A/B by environment: bash tools/gpu_env_ab.sh "<scene nx ny spp>" "<ENV=VAL ...>" ...   ('-' = defaults); keeps time + image hash
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
scene="$1"; shift
for c in "$@"; do
  envs="$c"; [ "$c" = "-" ] && envs=""
  echo "== $scene [$envs]"
  env RTW_PROF_HASH=1 $envs timeout 120 python tools/prof_render.py $scene 3 2>&1 | tail -2
done
