"""Short render loop for ncu captures: `python tools/prof_render.py [scene] [nx] [ny] [spp] [reps] [flags]`.

Same kernel and launch shape as bench.py's step, but with few samples so that ncu's ~40 replays stay cheap.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracingweekend_b200 as r

name = sys.argv[1] if len(sys.argv) > 1 else "cornell_box_aluminum"
nx = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
ny = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
spp = int(sys.argv[4]) if len(sys.argv) > 4 else 64
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 4
flags = int(sys.argv[6]) if len(sys.argv) > 6 else 0  # 1 = RTW_FLAG_WAVEFRONT
sc = r.Scene.named(name, nx / ny)
tag = " [wavefront]" if flags & 1 else ""
for _ in range(reps):
    acc, st = sc.render(nx, ny, spp, flags=flags)
    if os.environ.get("RTW_PROF_HASH") == "2":  # every repetition
        import hashlib

        print(f"sha256(accum)={hashlib.sha256(acc.tobytes()).hexdigest()[:16]} rays={st['rays']}", flush=True)
    print(f"{name}{tag} {nx}x{ny}x{spp}: launches={st['kernel_launches']} kernel_ms={st['kernel_ms']:.2f} Gpaths/s={st['paths'] / st['kernel_ms'] / 1e6:.3f} "
          f"Grays/s={st['rays'] / st['kernel_ms'] / 1e6:.3f}", flush=True)
if os.environ.get("RTW_PROF_HASH"):  # A/B of kernels that must give bit-identical images
    import hashlib

    print(f"sha256(accum)={hashlib.sha256(acc.tobytes()).hexdigest()[:16]} rays={st['rays']}", flush=True)
