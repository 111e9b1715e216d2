"""Ownership check of the sorted kernel's slot rings (sorted_kernel.cuh, -DRTW_SORTED_CHECK).

    python tools/ring_check.py build                      # here (no GPU): raytracingweekend_b200/lib/librtw_b200_ringcheck.so
    python tools/ring_check.py run [scene nx ny spp reps] # on the GPU box: renders with that library, reports anomalies

The check build keeps, per path slot, who holds it and where it was last queued / taken, and prints the first 40 anomalies:
a slot taken from a ring while another warp holds it, a slot queued by a warp that does not hold it, a block that ends with
queued entries or unclaimed samples.  This is the tool that found the one-lap race of nearly full rings (profiles/r2_ab_phased.txt).
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
LIB = os.path.join(ROOT, "raytracingweekend_b200", "lib", "librtw_b200_ringcheck.so")
DEFAULT = [("book2_final_scene", 800, 800, 256, 3), ("cornell_box_aluminum", 1024, 1024, 128, 2), ("random_balls_scene", 1200, 800, 50, 2),
           ("cornell_smoke_scene", 600, 600, 100, 2), ("two_perlin_spheres_earth", 800, 800, 64, 2)]

if sys.argv[1:2] == ["build"]:
    from raytracingweekend_b200.build import build

    print(build(extra=["-DRTW_SORTED_CHECK=1"], tag="_ringcheck"))
    sys.exit(0)
cases = [tuple(sys.argv[2:3]) + tuple(int(x) for x in sys.argv[3:7])] if len(sys.argv) > 6 else DEFAULT
bad = 0
for scene, nx, ny, spp, reps in cases:
    env = dict(os.environ, RTW_B200_LIB=LIB, RTW_PROF_HASH="2")
    try:
        out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "prof_render.py"), scene, str(nx), str(ny), str(spp), str(reps)], env=env,
                             capture_output=True, text=True, timeout=120).stdout
    except subprocess.TimeoutExpired:
        print(f"{scene}: TIMEOUT")
        bad += 1
        continue
    anomalies = [l for l in out.splitlines() if l.startswith("block ")]
    digests = sorted(set(l for l in out.splitlines() if l.startswith("sha256")))
    print(f"{scene} {nx}x{ny}x{spp} x{reps}: {len(anomalies)} anomalies, {len(digests)} distinct image(s) {digests[:2]}")
    for l in anomalies[:6]:
        print("   ", l)
    bad += bool(anomalies) or len(digests) != 1
sys.exit(1 if bad else 0)
