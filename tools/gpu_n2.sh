cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out; O=gpurun_out
nvidia-smi -L
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > $O/pytest_n2.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest_n2.log
NCCL_DEBUG=WARN timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 --no-per-config > $O/bench_n2.jsonl 2> $O/bench_n2.err; echo "bench n2 rc=$?"; tail -5 $O/bench_n2.err; head -c 700 $O/bench_n2.jsonl; echo
timeout 600 python bench.py --gpus 1 --steps 3 --warmup 3 --no-per-config > $O/bench_n1.jsonl 2> $O/bench_n1.err; echo "bench n1 rc=$?"; head -c 300 $O/bench_n1.jsonl; echo
