#!/usr/bin/env python
"""bench.py — headline benchmark of the path-tracing hot path (BASELINE.json: paths/s & Mrays/s, Cornell box).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...     (N > 1, one rank per GPU)

One "step" = one pass of the hot path over the whole workload: the Book-3 Cornell box with glass sphere and
aluminium box (BASELINE.json configs[3]) at 1024x1024, 1024 samples per pixel = 1.07 G paths.  At N GPUs the 1024
samples of every pixel are sharded into disjoint Philox streams (strong scaling: total work fixed) and the float4
accumulators are reduced to rank 0 with one ncclReduce inside the timed region.

The JSON line carries: `value` (paths/s, device-timed, accumulators resident in HBM), `e2e` (same metric through
the public host-buffer C-ABI call incl. scene flatten + upload + D2H of the image), `roofline` (FP32-issue bound,
algorithmic flops from the oracle's event counts), `cpu_baseline` (the unmodified reference on the host cores),
`clocks`, `gpu_launches`.  `--impl reference` times the reference's own CPU implementation instead.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SCENE = "cornell_box_aluminum"
NX = NY = 1024
SPP = 1024
MAX_DEPTH = 100
# SURVEY §8d / DESIGN.md §5: algorithmic FP32 flops per path of the reference's math on this scene, counted by the
# instrumented oracle (tools/count_flops.py: oracle_events_*, 64x64x64 spp sample, Appendix D cost table).
NCU_DRAM_BYTES_PER_LAUNCH = 49664 + 1280  # read + write (the image is still in L2 when the kernel ends), see roofline.traffic below
FLOP_PER_PATH = 1344.5
FP32_PEAK_SPEC_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12  # 74.4


CONFIGS = {  # --config: the headline workload (default) and the scaling-sweep shape of BASELINE.json configs[4]
    # nx, ny, spp, note, algorithmic flop/path (tools/count_flops.py; the 16:9 frame sees past the box: 3.31 rays/path, not 5.11)
    "cfg4": (1024, 1024, 1024, "BASELINE.json configs[3]", 1344.5),
    "cfg5": (3840, 2160, 4096, "BASELINE.json configs[4], spp-sharded scaling sweep", 819.3),
}
CONFIG_NOTE = CONFIGS["cfg4"][3]


def workload_config(n_gpus: int) -> dict:
    return {"workload": f"{SCENE} (Book-3 Cornell box, glass sphere + aluminium box; {CONFIG_NOTE})",
            "nx": NX, "ny": NY, "spp": SPP, "max_depth": MAX_DEPTH, "paths_per_step": NX * NY * SPP,
            "parallelism": f"spp-shard x{n_gpus} + ncclReduce" if n_gpus > 1 else "single GPU",
            "l2": "flushed between timed steps (256 MiB write); scene data is ~1 KB and cache-resident by design",
            "rng": "Philox4x32-10 keyed (pixel, sample, bounce)"}


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
def cpu_reference(target_seconds: float = 15.0) -> dict:
    """The reference's own image loop + color() (oracle/_ref/ref_harness: unmodified sources, std::thread ppl shim)
    on all host cores, on a bounded sample of the SAME workload: same scene, same resolution (1024x1024), reduced spp
    (paths/s does not depend on spp).  Falls back to the oracle port when the reference binary is absent."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol

    cores = os.cpu_count() or 1
    if ol.have_ref():
        # calibrate with 1 spp, then size the sample for ~target_seconds
        _, _, info = ol.ref_render(SCENE, NX, NY, 1, max_depth=MAX_DEPTH, threads=cores)
        spp = int(max(1, min(64, round(target_seconds * info["paths_per_s"] / (NX * NY)))))
        if spp > 1:
            _, _, info = ol.ref_render(SCENE, NX, NY, spp, max_depth=MAX_DEPTH, threads=cores)
        return {"value": info["paths_per_s"], "unit": "paths/s", "cores": cores, "kind": "reference",
                "sample": f"{SCENE} {NX}x{NY} at {info['spp']} spp ({info['paths']:.0f} paths, {info['trace_ms']} ms), "
                          f"unmodified reference via std::thread parallel_for shim, {cores} threads sharing the "
                          f"reference's unsynchronised static RNG engines (as written)"}
    import raytracingweekend_b200 as rtw

    b = rtw.Builder().named_scene(SCENE, NX / NY)
    t0 = time.time()
    ol.oracle_render(b, 256, 256, 4, max_depth=MAX_DEPTH, mode="philox", threads=cores)
    rate = 256 * 256 * 4 / (time.time() - t0)
    spp = int(max(1, min(64, round(target_seconds * rate / (NX * NY)))))
    t0 = time.time()
    ol.oracle_render(b, NX, NY, spp, max_depth=MAX_DEPTH, mode="philox", threads=cores)
    dt = time.time() - t0
    return {"value": NX * NY * spp / dt, "unit": "paths/s", "cores": cores, "kind": "port",
            "sample": f"{SCENE} {NX}x{NY} at {spp} spp, FP64 oracle port (oracle/rtw_oracle.cpp), {cores} threads"}


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    vals, sample, kind, cores = [], "", "reference", os.cpu_count() or 1
    per_step = max(4.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    for i in range(args.warmup + args.steps):
        r = cpu_reference(per_step)
        if i >= args.warmup:
            vals.append(r["value"])
        sample, kind, cores = r["sample"], r["kind"], r["cores"]
    v = sum(vals) / len(vals)
    line = {"impl": "reference", "metric": "paths/s", "value": v, "unit": "paths/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": NX * NY * SPP / v * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(args.gpus),
            "cpu_baseline": {"value": v, "unit": "paths/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": "paths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": f"ms_per_step is extrapolated to the full {SPP}-spp step from the bounded sample"}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------ our arm
def run_ours(args) -> None:
    import numpy as np
    import torch
    import torch.distributed as dist

    import raytracingweekend_b200 as rtw
    from raytracingweekend_b200 import _lib
    from raytracingweekend_b200.dist import reduce_accumulators, shard_range

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torchrun --nproc-per-node N for --gpus N")
    torch.cuda.set_device(local)
    # stdout must carry ONE JSON line: NCCL prints its version banner to the process's stdout (C stdio) at communicator set-up,
    # so file descriptor 1 points at stderr for the whole run and the line is written to the saved descriptor at the end
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = _lib.load()  # fails loudly when the CUDA library is missing

    builder = rtw.Builder().named_scene(SCENE, NX / NY)
    scene = rtw.Scene(builder, device=local)
    begin, count = shard_range(SPP, rank, world)
    dev = torch.device("cuda", local)
    accum = torch.empty((NY, NX, 4), dtype=torch.float32, device=dev)
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream(dev)
    rays_total = 0

    def step():
        nonlocal rays_total
        scene.render_device(accum.data_ptr(), NX, NY, count, spp_begin=begin, max_depth=MAX_DEPTH, seed=0, stream=stream.cuda_stream)
        st = scene.finish()
        if world > 1:
            reduce_accumulators(accum, dst=0)
        return st

    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    times, kernel_ms, rays = [], [], []
    for _ in range(args.steps):
        flush.fill_(1.0)  # L2 flush between timed iterations
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
    clocks = sampler.stop() if sampler else None
    t = torch.tensor([sum(times), sum(kernel_ms), float(sum(rays))], dtype=torch.float64, device=dev)
    if world > 1:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone()
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        total_ms, kern_ms, total_rays = float(tmax[0]), float(tmax[1]), float(tsum[2])
    else:
        total_ms, kern_ms, total_rays = float(t[0]), float(t[1]), float(t[2])
    paths_per_step = NX * NY * SPP
    value = paths_per_step * args.steps / (total_ms * 1e-3)
    rays_per_s = total_rays / (total_ms * 1e-3)

    # ---- end to end through the public API with HOST buffers, at N GPUs, every step from scratch:
    #      scene description -> flatten -> H2D upload -> render (this rank's sample shard) -> [ncclReduce to rank 0] ->
    #      D2H of the float4 image into pinned host memory on rank 0.  Wall clock between barriers, max over ranks.
    #      N = 1 goes through the C-ABI host-buffer call rtw_render itself (Scene.render).
    import ctypes as C
    from raytracingweekend_b200.dist import render_sharded
    fs = builder.flatten().contents
    h2d = (fs.n_items * 32 + fs.n_prims * 36 + fs.n_xfs * 32 + fs.n_nodes * 32 + fs.n_leaves * 8 + fs.n_materials * 32 +
           fs.n_textures * 48 + fs.n_lights * 32 + fs.n_texel_bytes + C.sizeof(_lib.RenderParams))
    host_t = torch.empty((NY, NX, 4), dtype=torch.float32).pin_memory()
    host = host_t.numpy()
    e2e_steps = max(1, min(args.steps, 3))
    e2e_times = []
    for it in range(e2e_steps + 1):  # the first pass is an untimed warm-up (NCCL reduce channels, allocator, page-ins)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        b2 = rtw.Builder().named_scene(SCENE, NX / NY)   # scene description -> flatten
        s2 = rtw.Scene(b2, device=local)                 # H2D upload of the flat arrays
        if world == 1:
            s2.render(NX, NY, SPP, max_depth=MAX_DEPTH, seed=0, out=host)  # kernel + D2H of the float4 image
        else:
            acc2, _ = render_sharded(s2, NX, NY, SPP, max_depth=MAX_DEPTH, seed=0, accum=accum)  # shard + ncclReduce
            if rank == 0:
                host_t.copy_(acc2, non_blocking=False)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        s2.close()
        b2.close()
        if os.environ.get("RTW_BENCH_DEBUG"):
            sys.stderr.write(f"e2e pass {it}: {dt * 1e3:.2f} ms\n")
        if it > 0:
            e2e_times.append(dt)
    te = torch.tensor([sum(e2e_times)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te[0]) / e2e_steps

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    e2e = {"value": paths_per_step / e2e_s, "unit": "paths/s", "h2d_bytes_per_step": int(h2d) * world,
           "d2h_bytes_per_step": NX * NY * 16, "ms_per_step": e2e_s * 1e3, "n_gpus": world,
           "what": ("Builder.named_scene + Scene() + rtw_render into a host buffer (flatten, H2D, kernel, D2H)" if world == 1 else
                    "per rank: Builder.named_scene + Scene() + rtw_render_device of its sample shard; ncclReduce to rank 0; "
                    "D2H of the image into pinned host memory")}

    # ---- roofline: FP32 issue ----
    peak = C.c_double(0)
    lib.rtw_measure_fp32_peak(local, C.byref(peak))
    fpp = FLOP_PER_PATH  # counted offline by tools/count_flops.py (oracle event counts x Appendix-D costs)
    kern_paths_per_s = paths_per_step * args.steps / world / (kern_ms * 1e-3)  # one GPU's kernel
    achieved = fpp * kern_paths_per_s / 1e12
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    sorted_kernel = os.environ.get("RTW_MEGA_SORTED", "1") != "0"
    roofline = {"bound": "fp32-issue", "achieved": achieved, "peak": peak.value, "unit": "TFLOP/s",
                "frac": achieved / peak.value if peak.value else None,
                # dram__bytes_read.sum + dram__bytes_write.sum of ONE launch over the whole 1024x1024 image (256 spp; the traffic
                # is the scene read and the image store and does not grow with spp), from the ncu
                # --set full capture summarised in profiles/r1_sorted_v11_cornell_ncu_details.txt; the 16 MB image was still
                # in L2 when the kernel ended.  Only valid for the single-GPU shape of that capture.
                "traffic": NCU_DRAM_BYTES_PER_LAUNCH if world == 1 and sorted_kernel and (NX, NY) == (1024, 1024) else None,
                "traffic_unit": "bytes per launch (ncu, profiles/r1_sorted_v11_cornell_ncu_details.txt)",
                "peak_source": "FFMA micro-benchmark run in this process (rtw_measure_fp32_peak); spec-derived "
                               f"{FP32_PEAK_SPEC_TFLOPS:.1f} TFLOP/s at 1965 MHz",
                "frac_of_spec": achieved / FP32_PEAK_SPEC_TFLOPS,
                "algorithmic_flop_per_path": fpp,
                # the roofline this kernel actually sits under is warp-instruction ISSUE (one per scheduler per cycle, every
                # opcode): ncu sm__inst_issued / issue slots of the cfg4 capture (profiles/r1_sorted_v11_cornell_ncu_details.txt);
                # the algorithmic count above excludes compares, selects, address/loop arithmetic and the Philox integer rounds
                "issue_slots_busy_ncu": 0.801 if sorted_kernel and (NX, NY) == (1024, 1024) else None,
                "kernel": "render_sorted (material-sorted megakernel, csrc/sorted.cu)" if sorted_kernel else "render_megakernel",
                "kernel_ms_per_launch": kern_ms / args.steps,
                "hbm": {"algorithmic_bytes_per_launch": NX * NY * 16, "achieved_gbs": NX * NY * 16 / (kern_ms / args.steps * 1e-3) / 1e9,
                        "peak_gbs": peaks.get("hbm_gbs"), "note": "not HBM-bound: one 16-byte store per pixel per launch"}}
    cpu = cpu_reference(15.0) if world == 1 else None

    line = {"metric": "paths/s", "value": value, "unit": "paths/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(world), "mrays_per_s": rays_per_s / 1e6,
            "rays_per_path": total_rays / (paths_per_step * args.steps), "e2e": e2e, "gpu_launches": args.steps * world,
            "clocks": clocks, "roofline": roofline}
    if cpu:
        line["cpu_baseline"] = cpu
        line["speedup_vs_cpu_reference"] = {"device": value / cpu["value"], "e2e": e2e["value"] / cpu["value"]}
    sys.stdout.flush()
    os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="cfg4", choices=sorted(CONFIGS),
                    help="cfg4 = the headline workload (default); cfg5 = Cornell 3840x2160x4096, the scaling-sweep shape")
    args = ap.parse_args()
    global NX, NY, SPP, CONFIG_NOTE, FLOP_PER_PATH
    NX, NY, SPP, CONFIG_NOTE, FLOP_PER_PATH = CONFIGS[args.config]
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
