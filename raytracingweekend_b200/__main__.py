"""Run-time replacement for the reference's compile-time configuration (RayTracingWeekend.cpp:32-36 `const int` image
size / sample count, :199-202 scene `typedef`): pick scene, size, samples and output at the command line.

    python -m raytracingweekend_b200 --scene cornell_box_aluminum --nx 1024 --ny 1024 --spp 1024 --out cornell.png

Prints `Trace:` / `Write:` milliseconds like the reference (RayTracingWeekend.cpp:279-280).  Under torchrun
(`python -m torch.distributed.run --nproc-per-node N -m raytracingweekend_b200 ...`) the samples are sharded across the
ranks' GPUs and rank 0 writes the image.  Python is plumbing: every step is a call into librtw_b200.so.
"""
from __future__ import annotations

import argparse
import os
import time

import numpy as np

from . import Scene, SCENE_NAMES, _lib, tonemap, write_png, write_ppm, write_ppm_binary


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="python -m raytracingweekend_b200", description=__doc__.split("\n\n")[0])
    ap.add_argument("--scene", default="cornell_box_scene", choices=SCENE_NAMES)
    ap.add_argument("--nx", type=int, default=400)           # RayTracingWeekend.cpp:35
    ap.add_argument("--ny", type=int, default=400)           # :36
    ap.add_argument("--spp", type=int, default=64)           # :33 subPixelCount
    ap.add_argument("--max-depth", type=int, default=100)    # :42
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--out", default="1.ppm", help=".ppm (ASCII P3 like the reference), .pbm6/.p6 (binary PPM) or .png")
    ap.add_argument("--pipeline", default="megakernel", choices=["megakernel", "wavefront", "auto"])
    ap.add_argument("--debug-ray", action="store_true", help="RayTracingWeekend.cpp:221-225: every sample aims at the centre pixel")
    ap.add_argument("--normals", action="store_true", help="RenderType::Normal (RayTracingWeekend.cpp:135-136)")
    ap.add_argument("--earth", default=None, metavar="JPEG", help="baseline JPEG for the image_texture scenes, decoded by rtw_decode_jpeg "
                    "(e.g. the reference's RayTracingWeekend/earth.jpg); default: the packaged texels of that file")
    a = ap.parse_args(argv)

    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    flags = {"megakernel": 0, "wavefront": _lib.RTW_FLAG_WAVEFRONT, "auto": _lib.RTW_FLAG_AUTO}[a.pipeline]
    if a.debug_ray:
        flags |= _lib.RTW_FLAG_DEBUG_RAY
    if a.normals:
        from . import Builder, load_earth
        earth = load_earth(a.earth) if a.scene in ("two_perlin_spheres_earth", "book2_final_scene") else None
        b = Builder().named_scene(a.scene, a.nx / a.ny, earth)
        b.set_render_type(1)
        scene = Scene(b, device=local)
    else:
        scene = Scene.named(a.scene, a.nx / a.ny, device=local)

    t0 = time.perf_counter()
    if world > 1:
        import torch
        import torch.distributed as dist
        from .dist import render_sharded
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        acc_t, st = render_sharded(scene, a.nx, a.ny, a.spp, max_depth=a.max_depth, seed=a.seed, flags=flags)
        torch.cuda.synchronize()
        acc = acc_t.cpu().numpy() if rank == 0 else None
        rays = torch.tensor([float(st["rays"])], dtype=torch.float64, device=acc_t.device)
        dist.all_reduce(rays)
        st = dict(st, rays=int(rays.item()))
    else:
        acc, st = scene.render(a.nx, a.ny, a.spp, max_depth=a.max_depth, seed=a.seed, flags=flags)
    trace_ms = (time.perf_counter() - t0) * 1e3
    if rank == 0:
        t1 = time.perf_counter()
        canvas = tonemap(acc)  # :241-244
        ext = os.path.splitext(a.out)[1].lower()
        {".png": write_png, ".p6": write_ppm_binary, ".pbm6": write_ppm_binary}.get(ext, write_ppm)(a.out, canvas)
        write_ms = (time.perf_counter() - t1) * 1e3
        paths = a.nx * a.ny * a.spp
        print(f"Trace: {trace_ms:.0f}ms")  # :279
        print(f"Write: {write_ms:.0f}ms")  # :280
        print(f"{a.scene} {a.nx}x{a.ny}x{a.spp} on {world} GPU(s): {paths / trace_ms / 1e6:.3f} Gpaths/s, "
              f"{st['rays'] / trace_ms / 1e6:.3f} Grays/s (wall clock incl. copies), NaN pixels: {int(np.isnan(acc).any(axis=2).sum())} -> {a.out}")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
