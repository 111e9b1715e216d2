# A/B of the refill kernel (RTW_REFILL=0/1, same library): bit-identical images, time.  bash tools/gpu_refill.sh [spp ...]
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
out=gpurun_out/ab_refill.txt; : > $out
for spp in "${@:-32 128}"; do
  for r in 0 1; do
    echo "== RTW_REFILL=$r book2_final_scene 800x800x$spp" >> $out
    RTW_PROF_HASH=1 RTW_REFILL=$r timeout 120 python tools/prof_render.py book2_final_scene 800 800 $spp 3 2>&1 | tail -2 >> $out || echo "FAILED/timeout rc=$?" >> $out
  done
done
cat $out
