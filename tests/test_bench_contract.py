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
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                        "--cpu-seconds", "2", "--config", "cfg4"], capture_output=True, text=True, cwd=ROOT, timeout=900, env=dict(os.environ, RANK="0"))
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
    # the headline line also carries the other four BASELINE.json configs at their stated shapes
    per = line["per_config"]
    assert sorted(per) == ["cfg1", "cfg2", "cfg3", "cfg5"]
    shapes = {"cfg1": (1200, 800, 100), "cfg2": (800, 800, 256), "cfg3": (800, 800, 1024), "cfg5": (3840, 2160, 4096)}
    for k, v in per.items():
        c = v["config"]
        assert (c["nx"], c["ny"], c["spp"]) == shapes[k] and c["max_depth"] == 100 and v["value"] > 1e3


def test_bench_configs_are_the_baseline_json_configs():
    import importlib.util

    spec = importlib.util.spec_from_file_location("bench", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    base = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    assert len(base["configs"]) == len(bench.CONFIGS) == 5
    for text, (name, c) in zip(base["configs"], sorted(bench.CONFIGS.items())):
        t = text.replace("\u00d7", "x").replace("×", "x")
        assert f"{c['nx']}x{c['ny']}" in t and f"{c['spp']} spp" in t, (name, text)
    fl = bench.algorithmic_flops("cfg4")
    assert abs(fl["flop_per_path"] - 1344.5) < 40 and abs(fl["rays_per_path"] - 5.11) < 0.1


def test_reference_arm_is_silent_on_other_ranks():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, cwd=ROOT, timeout=60, env=dict(os.environ, RANK="1", WORLD_SIZE="2"))
    assert r.returncode == 0 and r.stdout.strip() == ""
