"""GPU sanity: render each named scene, print mean 8-bit RGB and throughput; write PPMs to gpurun_out/."""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracingweekend_b200 as r

os.makedirs("gpurun_out", exist_ok=True)
for name, nx, ny, spp in [("cornell_box_scene", 400, 400, 64), ("cornell_box_aluminum", 400, 400, 64),
                          ("dielectric_scene", 400, 200, 64), ("random_balls_scene", 300, 200, 16),
                          ("light_sample", 200, 200, 16), ("two_perlin_spheres_earth", 200, 200, 16),
                          ("book2_final_scene", 200, 200, 16)]:
    t0 = time.time()
    sc = r.Scene.named(name, nx / ny)
    acc, st = sc.render(nx, ny, spp)
    canvas = r.tonemap(acc)
    img = (255.99 * canvas).astype(int)
    nan = int(np.isnan(acc).any(axis=2).sum())
    print(f"{name:26s} {nx}x{ny}x{spp} mean8={np.nanmean(img, axis=(0,1)).round(1)} black={np.mean(img.sum(2)==0):.3f} "
          f"nanpix={nan} rays/path={st['rays']/st['paths']:.3f} kernel_ms={st['kernel_ms']:.2f} "
          f"Mpaths/s={st['paths']/st['kernel_ms']/1e3:.1f} Mrays/s={st['rays']/st['kernel_ms']/1e3:.1f} wall={time.time()-t0:.2f}s", flush=True)
    r.write_ppm(f"gpurun_out/{name}.ppm", canvas)
sc = r.Scene.named("cornell_box_aluminum", 1.0)
for spp in (16, 256, 1024):
    acc, st = sc.render(1024, 1024, spp)
    print(f"cornell_aluminum 1024x1024x{spp}: kernel_ms={st['kernel_ms']:.1f} Gpaths/s={st['paths']/st['kernel_ms']/1e6:.3f} Grays/s={st['rays']/st['kernel_ms']/1e6:.3f}", flush=True)
