# bash tools/gpu_scale.sh N : the default bench line (headline + per_config) at N GPUs
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out; N=$1
NCCL_DEBUG=WARN timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n$N.jsonl 2> gpurun_out/bench_n$N.err; echo "bench n$N rc=$?"; tail -3 gpurun_out/bench_n$N.err | cut -c1-300
python - $N <<'PY'
import json,sys
l=json.loads(open(f'gpurun_out/bench_n{sys.argv[1]}.jsonl').read())
print('cfg4', round(l['value']/1e9,3), 'e2e', round(l['e2e']['value']/1e9,3), 'ms', round(l['ms_per_step'],2), 'kernel ms', round(l['roofline']['kernel_ms_per_launch'],2))
for k,v in sorted(l.get('per_config',{}).items()): print(k, round(v['value']/1e9,3), 'e2e', round(v['e2e']['value']/1e9,3), 'ms', round(v['ms_per_step'],2))
PY
