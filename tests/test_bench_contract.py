"""bench.py's reference arm (`--impl reference`) runs on host cores only, so its side of the driver contract is checked here:
stdout is ONE JSON line with the agreed keys, and the CPU baseline really is the unmodified reference (oracle/_ref) when it
was built.  The GPU arm shares the line-building code and is exercised on the GPU box."""
import json
import os
import subprocess
import sys

import oracle_lib as ol

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_json_line_with_the_contract_keys():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, cwd=ROOT, timeout=600, env=dict(os.environ, RANK="0"))
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, r.stdout
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["metric"] == "paths/s" and line["unit"] == "paths/s"
    assert line["higher_is_better"] is True and line["n_gpus"] == 1 and line["steps"] == 1 and line["warmup"] == 0
    assert line["value"] > 1e4 and line["vs_baseline"] is None and line["data"] == "synthetic"
    cfg = line["config"]
    assert "cornell_box_aluminum" in cfg["workload"] and (cfg["nx"], cfg["ny"], cfg["spp"]) == (1024, 1024, 1024)
    cpu = line["cpu_baseline"]
    assert cpu["kind"] == ("reference" if ol.have_ref() else "port") and cpu["cores"] >= 1 and cpu["value"] == line["value"]
    assert line["e2e"] == {"value": line["value"], "unit": "paths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_is_silent_on_other_ranks():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, cwd=ROOT, timeout=60, env=dict(os.environ, RANK="1", WORLD_SIZE="2"))
    assert r.returncode == 0 and r.stdout.strip() == ""
