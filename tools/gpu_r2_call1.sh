# round 2, GPU call: smoke, gpu tests, bench (headline + per_config), launch list
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out; O=gpurun_out
nvidia-smi -L
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/smoke.log
timeout 2400 python -m pytest tests -m gpu -x -q -s > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -25 $O/pytest.log
timeout 1200 python bench.py --steps 5 --warmup 3 > $O/bench.jsonl 2> $O/bench.err; echo "bench rc=$?"; tail -c 800 $O/bench.err; head -c 1200 $O/bench.jsonl
