#!/usr/bin/env python
"""bench.py — benchmark of the path-tracing hot path on the five BASELINE.json configs (metric: paths/s & Mrays/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config cfg1..cfg5] [--impl reference] [--no-per-config]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...     (N > 1, one rank per GPU)

One "step" = one pass of the hot path over the whole workload of the chosen config.  The headline (default) is cfg4 =
BASELINE.json configs[3]: the Book-3 Cornell box with glass sphere and aluminium box at 1024x1024, 1024 samples per pixel
= 1.07 G paths.  At N GPUs the samples of every pixel are sharded into disjoint Philox streams (strong scaling: total
work fixed) and the float4 accumulators are summed onto rank 0 inside the timed region (rtw_reduce_accumulators =
ncclReduce through the C ABI).

The JSON line carries: `value` (paths/s, device-timed, accumulators resident in HBM), `e2e` (same metric through the
public host-buffer C-ABI call incl. scene flatten + upload + D2H of the image), `roofline` (FP32-issue bound, algorithmic
flops counted in this run by the instrumented oracle), `cpu_baseline` (the unmodified reference on the host cores),
`clocks`, `gpu_launches`, and — on the default line — `per_config`: a bounded run (3 warm-up + 2 timed steps, full stated
shape) of the other four configs with the same fields.  `--impl reference` times the reference's own CPU implementation.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

MAX_DEPTH = 100
FP32_PEAK_SPEC_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12  # 74.4

# BASELINE.json `configs`, in order.  Scenes: Scene/scene.h:98-174 (cfg1), SURVEY App. C (cfg2, cfg3), Scene/scene.h:176-249
# with :228 enabled (cfg4, cfg5).  `flop_sample` = (nx, ny, spp) of the instrumented-oracle run that counts the algorithmic
# FP32 flops per path (same aspect ratio as the config, so the mix of rays is the config's).
CONFIGS = {
    "cfg1": dict(scene="random_balls_scene", nx=1200, ny=800, spp=100, book="Book-1 random scene, 485 spheres, linear hittable_list",
                 note="BASELINE.json configs[0]", flop_sample=(72, 48, 8)),
    "cfg2": dict(scene="two_perlin_spheres_earth", nx=800, ny=800, spp=256, book="Book-2 two Perlin spheres + earth image_texture sphere",
                 note="BASELINE.json configs[1]", flop_sample=(64, 64, 32)),
    "cfg3": dict(scene="book2_final_scene", nx=800, ny=800, spp=1024, book="Book-2 final scene: bvh_node over 400 boxes + 1000 spheres, "
                 "moving sphere, two constant_medium", note="BASELINE.json configs[2]", flop_sample=(48, 48, 8)),
    "cfg4": dict(scene="cornell_box_aluminum", nx=1024, ny=1024, spp=1024, book="Book-3 Cornell box, glass sphere + aluminium box",
                 note="BASELINE.json configs[3]", flop_sample=(64, 64, 64)),
    "cfg5": dict(scene="cornell_box_aluminum", nx=3840, ny=2160, spp=4096, book="Book-3 Cornell box, glass sphere + aluminium box",
                 note="BASELINE.json configs[4], spp-sharded scaling sweep", flop_sample=(128, 72, 32)),
}
HEADLINE = "cfg4"
EARTH_SCENES = ("two_perlin_spheres_earth", "book2_final_scene")
NCU_SUMMARY = os.path.join(ROOT, "profiles", "ncu_summary.json")  # written by tools/ncu_summary.py from the committed captures


def workload_config(cfg_name: str, n_gpus: int) -> dict:
    c = CONFIGS[cfg_name]
    return {"workload": f"{cfg_name}: {c['scene']} ({c['book']}; {c['note']})",
            "nx": c["nx"], "ny": c["ny"], "spp": c["spp"], "max_depth": MAX_DEPTH, "paths_per_step": c["nx"] * c["ny"] * c["spp"],
            "parallelism": f"spp-shard x{n_gpus} + ncclReduce (rtw_reduce_accumulators)" if n_gpus > 1 else "single GPU",
            "l2": "flushed between timed steps (256 MiB write); scene data is <= 0.5 MB and cache-resident by design",
            "rng": "Philox4x32-10 keyed (pixel, sample, bounce)"}


def make_builder(cfg_name: str):
    import raytracingweekend_b200 as rtw

    c = CONFIGS[cfg_name]
    earth = rtw.load_earth() if c["scene"] in EARTH_SCENES else None
    return rtw.Builder().named_scene(c["scene"], c["nx"] / c["ny"], earth)


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.gpu)], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, power, reasons = [], [], [], set()
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])), mx.append(float(f[2])), power.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.path)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------ CPU reference
def cpu_model() -> str:
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def cpu_reference(cfg_name: str, target_seconds: float = 15.0, min_spp: int = 1, courtesy: bool = False) -> dict:
    """The reference's own image loop + color() (oracle/_ref/ref_harness: unmodified sources, std::thread ppl shim)
    on all host cores, on a bounded sample of the SAME workload: same scene, same resolution, reduced spp (paths/s does
    not depend on spp).  `courtesy` adds the 1-thread serial_for figure and, when the -O3 -march=native build of the same
    sources exists (oracle/_ref/ref_harness_native, `make -C oracle ref_native`), its all-core figure.
    Falls back to the oracle port when the reference binary is absent."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol

    c = CONFIGS[cfg_name]
    scene, nx, ny = c["scene"], c["nx"], c["ny"]
    cores = os.cpu_count() or 1
    if ol.have_ref():
        # calibrate on a quarter-resolution frame of the same aspect, then size the full-resolution sample for ~target_seconds
        _, _, cal = ol.ref_render(scene, max(1, nx // 4), max(1, ny // 4), 1, max_depth=MAX_DEPTH, threads=cores, aspect=nx / ny)
        spp = int(max(min_spp, min(256, round(target_seconds * cal["paths_per_s"] / (nx * ny)))))
        _, _, info = ol.ref_render(scene, nx, ny, spp, max_depth=MAX_DEPTH, threads=cores)
        out = {"value": info["paths_per_s"], "unit": "paths/s", "cores": cores, "kind": "reference", "cpu_model": cpu_model(),
               "build": "g++ -std=c++17 -O2 (oracle/Makefile)",
               "sample": f"{scene} {nx}x{ny} at {info['spp']} spp ({info['paths']:.0f} paths, {info['trace_ms']} ms), "
                         f"unmodified reference via std::thread parallel_for shim, {cores} threads sharing the "
                         f"reference's unsynchronised static RNG engines (as written)"}
        if courtesy:
            # 1 thread = the reference's serial_for (RayTracingWeekend.cpp:186-192): no RNG contention
            sspp = max(1, spp // (2 * cores))
            t_budget = max(1, min(sspp, int(round(0.5 * target_seconds * info["paths_per_s"] / cores / (nx * ny))) or 1))
            _, _, s_info = ol.ref_render(scene, nx, ny, t_budget, max_depth=MAX_DEPTH, threads=1)
            out["serial_1_thread"] = {"value": s_info["paths_per_s"], "spp": s_info["spp"],
                                      "threads_speedup": info["paths_per_s"] / s_info["paths_per_s"]}
            if os.path.exists(ol.REF_HARNESS + "_native"):
                _, _, n_info = ol.ref_render(scene, nx, ny, spp, max_depth=MAX_DEPTH, threads=cores, binary=ol.REF_HARNESS + "_native")
                out["o3_march_native"] = {"value": n_info["paths_per_s"], "spp": n_info["spp"], "cores": cores}
        return out
    b = make_builder(cfg_name)
    t0 = time.time()
    ol.oracle_render(b, 256, 256 * ny // nx, 2, max_depth=MAX_DEPTH, mode="philox", threads=cores)
    rate = 256 * (256 * ny // nx) * 2 / (time.time() - t0)
    spp = int(max(min_spp, min(256, round(target_seconds * rate / (nx * ny)))))
    t0 = time.time()
    ol.oracle_render(b, nx, ny, spp, max_depth=MAX_DEPTH, mode="philox", threads=cores)
    dt = time.time() - t0
    return {"value": nx * ny * spp / dt, "unit": "paths/s", "cores": cores, "kind": "port", "cpu_model": cpu_model(),
            "sample": f"{scene} {nx}x{ny} at {spp} spp, FP64 oracle port (oracle/rtw_oracle.cpp), {cores} threads"}


def algorithmic_flops(cfg_name: str) -> dict:
    """F_path = sum_e N_e C_e / paths — event counts N_e from the instrumented CPU oracle on this config's scene
    and aspect ratio (the reference's math: linear list walks, both bvh_node children visited), costs C_e = Appendix D."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ctypes as C

    import numpy as np
    import oracle_lib as ol

    o = ol.oracle()
    o.oracle_events_read.restype = C.c_double
    o.oracle_events_read.argtypes = [C.c_void_p, C.c_int]
    nx, ny, spp = CONFIGS[cfg_name]["flop_sample"]
    b = make_builder(cfg_name)
    o.oracle_events_enable(1)
    _, _, rays = ol.oracle_render(b, nx, ny, spp, mode="philox", seed=1)
    n_ev = o.oracle_event_count()
    cnt = np.zeros(n_ev, dtype=np.uint64)
    fl = o.oracle_events_read(cnt.ctypes.data, n_ev)
    o.oracle_events_enable(0)
    b.close()
    paths = nx * ny * spp
    return {"flop_per_path": fl / paths, "flop_per_ray": fl / rays, "rays_per_path": rays / paths,
            "sample": f"instrumented oracle, {nx}x{ny}x{spp} spp", "events": cnt}


def tree_flops(cfg_name: str, oracle_events) -> dict | None:
    """cfg1 / cfg3: the render kernels walk trees built at upload (SAH over the long list / over the reference's leaves,
    front to back with a shrinking interval), so the reference-math count above over-states the work the kernel's
    ALGORITHM needs (cfg1: a 485-sphere linear list per ray).  Count the kernel's own traversal events with the
    instrumented library build (librtw_b200_count.so, -DRTW_COUNT_EVENTS) and price them with the same Appendix-D costs:
    traversal events from the device, everything else (shading, camera, accepts) from the oracle."""
    lib = os.path.join(ROOT, "raytracingweekend_b200", "lib", "librtw_b200_count.so")
    if not os.path.exists(lib):
        return None
    c = CONFIGS[cfg_name]
    nx, ny = c["nx"] // 4, c["ny"] // 4
    try:
        out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "count_tree_events.py"), c["scene"], str(nx), str(ny), "8"],
                             env=dict(os.environ, RTW_B200_LIB=lib), capture_output=True, text=True, timeout=300, check=True).stdout
        dev = json.loads(out.strip().splitlines()[-1])
    except Exception as e:  # the count build is diagnostics only
        return {"error": str(e)[:200]}
    # Appendix D: aabb::hit 21, sphere test 23 (+4 root, +12 moving centre), rect test 6, translate 3 / rotate 12 per transformed ray
    cost = {"aabb": 21.0, "sphere_test": 23.0, "sphere_root": 4.0, "moving_center": 12.0, "rect_test": 6.0, "box_test": 21.0, "xf": 15.0}
    trav = sum(cost[k] * dev["per_path"].get(k, 0.0) for k in cost)
    names = ("CAMERA LENS RECT_TEST RECT_ACCEPT SPHERE_TEST SPHERE_ROOT SPHERE_ACCEPT MOVING_CENTER AABB TRANSLATE ROTATE FLIP MEDIUM "
             "BACKGROUND EMITTED ONB COSINE_GEN COSINE_VALUE SCATTERING_PDF LIGHT_PICK RECT_RANDOM RECT_PDF SPHERE_RANDOM SPHERE_PDF MIXTURE "
             "LAMBERT_UPDATE SPECULAR_UPDATE METAL DIELECTRIC ISOTROPIC CHECKER IMAGE_TEX PERLIN_NOISE NOISE_TEX RAY").split()
    costs = [34, 25, 6, 12, 23, 4, 19, 12, 21, 3, 12, 3, 20, 21, 5, 36, 25, 15, 15, 3, 9, 43, 76, 60, 3, 12, 3, 48, 75, 21, 8, 7, 140, 48, 0]
    replaced = {"RECT_TEST", "SPHERE_TEST", "SPHERE_ROOT", "MOVING_CENTER", "AABB", "TRANSLATE", "ROTATE"}
    fs = CONFIGS[cfg_name]["flop_sample"]
    paths = fs[0] * fs[1] * fs[2]
    rest = sum(float(n) * cst for name, n, cst in zip(names, oracle_events, costs) if name not in replaced) / paths
    return {"flop_per_path": trav + rest, "traversal_flop_per_path": trav, "other_flop_per_path": rest, "device_events_per_path": dev["per_path"],
            "sample": f"librtw_b200_count.so, {nx}x{ny}x8 spp on the kernel's own trees"}


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = CONFIGS[args.config]
    # >= 16 spp and >= 20 s of CPU work per sample; the driver's --steps/--warmup (20/5) would make that 10
    # minutes, so at most 1 warm-up and 4 timed samples are actually run and the line says so
    n_warm, n_timed = min(args.warmup, 1), max(1, min(args.steps, 4))
    vals, last = [], None
    for i in range(n_warm + n_timed):
        r = cpu_reference(args.config, args.cpu_seconds, min_spp=16 if cfg["nx"] * cfg["ny"] <= 1200 * 1024 else 2, courtesy=(i == n_warm + n_timed - 1))
        if i >= n_warm:
            vals.append(r["value"])
        last = r
    v = sum(vals) / len(vals)
    paths = cfg["nx"] * cfg["ny"] * cfg["spp"]
    line = {"impl": "reference", "metric": "paths/s", "value": v, "unit": "paths/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": paths / v * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(args.config, args.gpus),
            "cpu_baseline": dict(last, value=v),
            "e2e": {"value": v, "unit": "paths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "samples_run": {"warmup": n_warm, "timed": n_timed, "values": vals},
            "note": f"each step is a bounded sample (same scene and resolution, reduced spp) of the {cfg['spp']}-spp workload; "
                    f"ms_per_step is extrapolated to the full step; {n_timed} timed samples stand for the {args.steps} requested steps"}
    if args.config == HEADLINE and not args.no_per_config:
        per = {}
        for name in CONFIGS:
            if name != args.config:
                r = cpu_reference(name, min(8.0, args.cpu_seconds))
                per[name] = {"value": r["value"], "unit": "paths/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"],
                             "config": workload_config(name, 1)}
        line["per_config"] = per
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------ our arm
class Ctx:
    """Per-process state shared by the per-config runs."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist

        from raytracingweekend_b200 import _lib

        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus and self.world == 1 and args.gpus > 1:
            raise SystemExit("launch with torchrun --nproc-per-node N for --gpus N")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.lib = _lib.load()  # fails loudly when the CUDA library is missing
        self.flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=self.dev)
        self.stream = torch.cuda.current_stream(self.dev)
        self.comm = None
        if self.world > 1:
            from raytracingweekend_b200.dist import Communicator

            self.comm = Communicator.from_torch(self.rank, self.world, self.local)  # ncclCommInitRank inside librtw_b200.so


def bench_config(ctx: Ctx, cfg_name: str, steps: int, warmup: int, with_cpu: bool, cpu_seconds: float, courtesy: bool) -> dict:
    import ctypes as C

    import raytracingweekend_b200 as rtw
    from raytracingweekend_b200 import _lib
    from raytracingweekend_b200.dist import shard_range

    torch, dist = ctx.torch, ctx.dist
    world, rank, local, dev, stream = ctx.world, ctx.rank, ctx.local, ctx.dev, ctx.stream
    c = CONFIGS[cfg_name]
    NX, NY, SPP = c["nx"], c["ny"], c["spp"]
    builder = make_builder(cfg_name)
    scene = rtw.Scene(builder, device=local)
    begin, count = shard_range(SPP, rank, world)
    accum = torch.empty((NY, NX, 4), dtype=torch.float32, device=dev)

    def step():
        scene.render_device(accum.data_ptr(), NX, NY, count, spp_begin=begin, max_depth=MAX_DEPTH, seed=0, stream=stream.cuda_stream)
        st = scene.finish()
        if world > 1:
            ctx.comm.reduce(accum.data_ptr(), NX, NY, root=0, stream=stream.cuda_stream)
        return st

    for _ in range(warmup):
        step()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    times, kernel_ms, rays, launches = [], [], [], 0
    for _ in range(steps):
        ctx.flush.fill_(1.0)  # L2 flush between timed iterations
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        st = step()
        e1.record(stream)
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
        kernel_ms.append(st["kernel_ms"])
        rays.append(st["rays"])
        launches += st["kernel_launches"]
    clocks = sampler.stop() if sampler else None
    t = torch.tensor([sum(times), sum(kernel_ms), float(sum(rays)), float(launches)], dtype=torch.float64, device=dev)
    if world > 1:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone()
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        total_ms, kern_ms, total_rays, total_launches = float(tmax[0]), float(tmax[1]), float(tsum[2]), int(tsum[3])
    else:
        total_ms, kern_ms, total_rays, total_launches = float(t[0]), float(t[1]), float(t[2]), int(t[3])
    paths_per_step = NX * NY * SPP
    value = paths_per_step * steps / (total_ms * 1e-3)
    rays_per_s = total_rays / (total_ms * 1e-3)

    # ---- end to end through the public API with HOST buffers, at N GPUs, every step from scratch:
    #      scene description -> flatten -> H2D upload -> render (this rank's sample shard) -> [ncclReduce to rank 0] ->
    #      D2H of the float4 image into pinned host memory on rank 0.  Wall clock between barriers, max over ranks.
    #      N = 1 goes through the C-ABI host-buffer call rtw_render itself (Scene.render); N > 1 through rtw_render_sharded.
    fs = builder.flatten().contents
    h2d = (fs.n_items * 32 + fs.n_prims * 36 + fs.n_xfs * 32 + fs.n_nodes * 32 + fs.n_leaves * 8 + fs.n_materials * 32 +
           fs.n_textures * 48 + fs.n_lights * 32 + fs.n_texel_bytes + C.sizeof(_lib.RenderParams))
    host_t = torch.empty((NY, NX, 4), dtype=torch.float32).pin_memory()
    host = host_t.numpy()
    e2e_steps = max(1, min(steps, 3))
    e2e_times = []
    for it in range(e2e_steps + 1):  # the first pass is an untimed warm-up (NCCL channels, allocator, page-ins)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        b2 = make_builder(cfg_name)                      # scene description -> flatten
        s2 = rtw.Scene(b2, device=local)                 # H2D upload of the flat arrays
        if world == 1:
            s2.render(NX, NY, SPP, max_depth=MAX_DEPTH, seed=0, out=host)  # kernel + D2H of the float4 image
        else:
            s2.render_sharded(ctx.comm, NX, NY, SPP, max_depth=MAX_DEPTH, seed=0, out=host if rank == 0 else None)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        s2.close()
        b2.close()
        if os.environ.get("RTW_BENCH_DEBUG"):
            sys.stderr.write(f"{cfg_name} e2e pass {it}: {dt * 1e3:.2f} ms\n")
        if it > 0:
            e2e_times.append(dt)
    te = torch.tensor([sum(e2e_times)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te[0]) / e2e_steps
    scene.close()
    builder.close()
    del accum, host_t
    if rank != 0:
        return {}
    e2e = {"value": paths_per_step / e2e_s, "unit": "paths/s", "h2d_bytes_per_step": int(h2d) * world,
           "d2h_bytes_per_step": NX * NY * 16, "ms_per_step": e2e_s * 1e3, "n_gpus": world,
           "what": ("Builder.named_scene + Scene() + rtw_render into a host buffer (flatten, H2D, kernel, D2H)" if world == 1 else
                    "per rank: Builder.named_scene + Scene() + rtw_render_sharded (its sample shard, ncclReduce to rank 0, D2H of the "
                    "image into pinned host memory on rank 0)")}

    # ---- roofline: FP32 issue ----
    peak = C.c_double(0)
    ctx.lib.rtw_measure_fp32_peak(local, C.byref(peak))
    fl = algorithmic_flops(cfg_name)
    fpp = fl["flop_per_path"]
    kern_paths_per_s = paths_per_step * steps / world / (kern_ms * 1e-3)  # one GPU's kernel
    achieved = fpp * kern_paths_per_s / 1e12
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    ncu = {}
    try:
        ncu = json.load(open(NCU_SUMMARY)).get(cfg_name, {})
    except Exception:
        pass
    sorted_kernel = os.environ.get("RTW_MEGA_SORTED", "1") != "0"
    roofline = {"bound": "fp32-issue", "achieved": achieved, "peak": peak.value, "unit": "TFLOP/s",
                "frac": achieved / peak.value if peak.value else None,
                # dram__bytes_read.sum + dram__bytes_write.sum of ONE launch, from the committed ncu --set full capture of this
                # config's kernel (profiles/ncu_summary.json names the file); null when no capture of this config exists.
                "traffic": ncu.get("dram_bytes_per_launch") if world == 1 and sorted_kernel else None,
                "traffic_source": ncu.get("source"),
                "peak_source": "FFMA micro-benchmark run in this process (rtw_measure_fp32_peak); MEASURED_PEAKS.json has no FP32 "
                               f"entry; spec-derived {FP32_PEAK_SPEC_TFLOPS:.1f} TFLOP/s at 1965 MHz",
                "frac_of_spec": achieved / FP32_PEAK_SPEC_TFLOPS,
                "algorithmic_flop_per_path": fpp, "algorithmic_flop_per_ray": fl["flop_per_ray"], "flop_count_sample": fl["sample"],
                # the roofline this kernel actually sits under is warp-instruction ISSUE (one per scheduler per cycle, every
                # opcode): ncu sm__inst_issued / issue slots of the committed capture; the algorithmic count above excludes
                # compares, selects, address/loop arithmetic and the Philox integer rounds
                "issue_slots_busy_ncu": ncu.get("issue_slots_busy") if sorted_kernel else None,
                "kernel": (("render_sorted, phased form (shade units / closest_hit units alternate per block: the code outgrows the SM "
                            "instruction cache; csrc/sorted_kernel.cuh, sorted_big.cu)") if cfg_name == "cfg3" and os.environ.get("RTW_PHASED", "1") != "0"
                           else "render_sorted (material-sorted megakernel, csrc/sorted_kernel.cuh)") if sorted_kernel else "render_megakernel",
                "kernel_ms_per_launch": kern_ms / max(1, total_launches // world),
                "hbm": {"algorithmic_bytes_per_launch": NX * NY * 16, "achieved_gbs": NX * NY * 16 / (kern_ms / steps * 1e-3) / 1e9,
                        "peak_gbs": peaks.get("hbm_gbs"), "note": "not HBM-bound: one 16-byte store per pixel per launch"}}
    if cfg_name in ("cfg1", "cfg3") and world == 1:
        tf = tree_flops(cfg_name, fl["events"])
        if tf and "flop_per_path" in tf:
            roofline["kernel_tree"] = dict(tf, achieved=tf["flop_per_path"] * kern_paths_per_s / 1e12,
                                           frac=tf["flop_per_path"] * kern_paths_per_s / 1e12 / peak.value if peak.value else None,
                                           note="algorithmic flops of the kernel's own acceleration structure (trees built at upload); the "
                                                "reference-math count above walks the list linearly / visits both bvh_node children")
        elif tf:
            roofline["kernel_tree"] = tf
    line = {"metric": "paths/s", "value": value, "unit": "paths/s", "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": total_ms / steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(cfg_name, world), "mrays_per_s": rays_per_s / 1e6,
            "rays_per_path": total_rays / (paths_per_step * steps), "e2e": e2e, "gpu_launches": total_launches,
            "clocks": clocks, "roofline": roofline}
    if with_cpu:
        cpu = cpu_reference(cfg_name, cpu_seconds, courtesy=courtesy)
        line["cpu_baseline"] = cpu
        line["speedup_vs_cpu_reference"] = {"device": value / cpu["value"], "e2e": e2e["value"] / cpu["value"]}
    return line


def run_ours(args) -> None:
    # stdout must carry ONE JSON line: NCCL prints its version banner to the process's stdout (C stdio) at communicator set-up,
    # so file descriptor 1 points at stderr for the whole run and the line is written to the saved descriptor at the end
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    ctx = Ctx(args)
    line = bench_config(ctx, args.config, args.steps, args.warmup, with_cpu=ctx.world == 1, cpu_seconds=15.0, courtesy=True)
    if args.config == HEADLINE and not args.no_per_config:
        per = {}
        for name in CONFIGS:
            if name == args.config:
                continue
            r = bench_config(ctx, name, steps=2, warmup=3, with_cpu=ctx.world == 1, cpu_seconds=6.0, courtesy=False)
            if ctx.rank == 0:
                per[name] = r
        if ctx.rank == 0:
            line["per_config"] = per
            line["gpu_launches_per_config"] = {k: v["gpu_launches"] for k, v in per.items()}
    if ctx.rank == 0:
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if ctx.world > 1:
        ctx.dist.barrier()
        if ctx.comm:
            ctx.comm.close()
        ctx.dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default=HEADLINE, choices=sorted(CONFIGS),
                    help="cfg4 = the headline workload (default, carries a per_config block for the others); cfg1..cfg5 = BASELINE.json configs in order")
    ap.add_argument("--no-per-config", action="store_true", help="headline line only")
    ap.add_argument("--cpu-seconds", type=float, default=22.0, help="reference arm: CPU seconds per sample (SURVEY 8d: >= 20 s, >= 16 spp)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
