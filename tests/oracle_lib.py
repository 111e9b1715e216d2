"""Test helper: ctypes binding of oracle/librtw_oracle.so (CPU restatement) and runner for oracle/_ref/ref_harness
(the unmodified reference).  TEST INFRASTRUCTURE ONLY — nothing in the product imports this."""
from __future__ import annotations

import ctypes as C
import json
import os
import subprocess
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "librtw_oracle.so")
REF_HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
EARTH = os.path.join(ROOT, "tests", "golden", "earth_256x256_rgb8.bin")
EARTH_SCENES = ("two_perlin_spheres_earth", "book2_final_scene")

_o = None


def oracle():
    global _o
    if _o is None:
        if not os.path.exists(ORACLE_SO):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], stdout=subprocess.DEVNULL)
        _o = C.CDLL(ORACLE_SO)
        _o.oracle_hit_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_double, C.c_double, C.c_void_p]
        _o.oracle_hit_batch.restype = None
        _o.oracle_render.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_int,
                                     C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]
        _o.oracle_render.restype = None
        _o.oracle_funcs.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        _o.oracle_funcs.restype = C.c_int
        _o.oracle_hardware_threads.restype = C.c_int
    return _o


HIT_COLS = ("hit", "t", "px", "py", "pz", "nx", "ny", "nz", "u", "v", "material", "prim")


def oracle_hits(builder, rays: np.ndarray, tmin=float(np.float32(0.001)), tmax=1.7976931348623157e308) -> np.ndarray:
    """rays [n,7] float64 -> [n,12] float64 (HIT_COLS). `builder` is a raytracingweekend_b200.Builder."""
    fs = builder.flatten()
    r = np.ascontiguousarray(rays, dtype=np.float64)
    out = np.zeros((r.shape[0], 12), dtype=np.float64)
    oracle().oracle_hit_batch(C.cast(fs, C.c_void_p), r.ctypes.data, r.shape[0], tmin, tmax, out.ctypes.data)
    return out


def oracle_render(builder, nx, ny, spp, spp_begin=0, max_depth=100, mode="philox", seed=0, threads=None):
    """Returns (sum[ny,nx,3], sumsq[ny,nx,3], rays)."""
    fs = builder.flatten()
    s = np.zeros((ny, nx, 3), dtype=np.float64)
    q = np.zeros((ny, nx, 3), dtype=np.float64)
    rays = C.c_uint64(0)
    if threads is None:
        threads = oracle().oracle_hardware_threads()
    oracle().oracle_render(C.cast(fs, C.c_void_p), nx, ny, spp_begin, spp, max_depth, 0 if mode == "minstd" else 1, seed, threads,
                           s.ctypes.data, q.ctypes.data, C.byref(rays))
    return s, q, int(rays.value)


def have_ref() -> bool:
    return os.path.exists(REF_HARNESS)


def _earth_arg(scene: str) -> str:
    return EARTH if scene in EARTH_SCENES else "-"


def ref_run(args, threads=1, timeout=3600, binary=None) -> str:
    env = dict(os.environ, RTW_PPL_THREADS=str(threads))
    return subprocess.run([binary or REF_HARNESS] + [str(a) for a in args], check=True, capture_output=True, text=True, env=env, timeout=timeout).stdout


def ref_flat(scene: str, aspect: float) -> bytes:
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "flat.bin")
        ref_run(["flat", scene, aspect, _earth_arg(scene), p])
        return open(p, "rb").read()


def ref_hits(scene: str, aspect: float, rays: np.ndarray, tmin="0.0010000000474974513", tmax="inf") -> np.ndarray:
    """Reference world.hit() -> [n,11]: hit, t, p(3), n(3), u, v, material."""
    with tempfile.TemporaryDirectory() as d:
        pi, po = os.path.join(d, "rays.bin"), os.path.join(d, "hits.bin")
        np.ascontiguousarray(rays, dtype=np.float64).tofile(pi)
        ref_run(["hits", scene, aspect, _earth_arg(scene), pi, po, tmin, tmax])
        return np.fromfile(po, dtype=np.float64).reshape(-1, 11)


def ref_render(scene: str, nx, ny, spp, max_depth=100, threads=1, aspect=0.0, binary=None):
    """Reference image loop -> (sum[ny,nx,3], sumsq[ny,nx,3], info dict).  `binary`: another build of the same harness."""
    with tempfile.TemporaryDirectory() as d:
        po = os.path.join(d, "img.bin")
        out = ref_run(["render", scene, aspect, _earth_arg(scene), nx, ny, spp, max_depth, po], threads=threads, binary=binary)
        a = np.fromfile(po, dtype=np.float64).reshape(ny, nx, 6)
        return a[..., :3].copy(), a[..., 3:].copy(), json.loads(out.strip().splitlines()[-1])


def ref_perlin() -> np.ndarray:
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "perlin.bin")
        ref_run(["perlin", p])
        return np.fromfile(p, dtype=np.float64)


def ref_funcs() -> np.ndarray:
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "funcs.bin")
        ref_run(["funcs", p])
        return np.fromfile(p, dtype=np.float64)


def oracle_funcs(tables: np.ndarray) -> np.ndarray:
    out = np.zeros(4096, dtype=np.float64)
    t = np.ascontiguousarray(tables, dtype=np.float64)
    n = oracle().oracle_funcs(t.ctypes.data, out.ctypes.data, out.size)
    return out[:n]


def set_jitter(eps: float, seed: int = 0) -> None:
    """oracle_set_jitter: per-ray relative perturbation of philox-mode renders (sensitivity probe); 0 switches it off."""
    o = oracle()
    o.oracle_set_jitter.argtypes = [C.c_double, C.c_uint64]
    o.oracle_set_jitter.restype = None
    o.oracle_set_jitter(float(eps), int(seed))


def fixed_rays(builder, n: int, seed: int) -> np.ndarray:
    """SURVEY 8d ray batch: half pinhole-grid primary rays (pixel centres of an m x m grid, m*m = n/2), half rays with random
    origins around the view volume and uniform directions; times uniform in [0, 1).  Every value is exactly representable in
    FP32, so GPU and CPU see the same rays.  (tests/golden/make_golden.py draws the reference-held 4000-ray batches the same way.)"""
    rng = np.random.default_rng(seed)
    cam = builder.flatten().contents.camera
    o, llc = np.array(cam.origin), np.array(cam.lower_left_corner)
    h, v = np.array(cam.horizontal), np.array(cam.vertical)
    m = int(np.sqrt(n // 2))
    s, t = np.meshgrid((np.arange(m) + 0.5) / m, (np.arange(m) + 0.5) / m)
    d = llc + s.reshape(-1, 1) * h + t.reshape(-1, 1) * v - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    prim = np.concatenate([np.tile(o, (m * m, 1)), d, rng.random((m * m, 1))], axis=1)
    ctr = o + d.mean(0) * np.linalg.norm(llc + 0.5 * h + 0.5 * v - o) * 60
    k = n - m * m
    ro = ctr + (rng.random((k, 3)) - 0.5) * np.linalg.norm(ctr - o) * 1.5
    rd = rng.normal(size=(k, 3))
    rays = np.concatenate([prim, np.concatenate([ro, rd, rng.random((k, 1))], axis=1)])
    return rays.astype(np.float32).astype(np.float64)


def record_stats(name: str, stats: dict) -> None:
    """Observed parity numbers of a GPU test run (flip counts, error maxima, divergent-pixel fractions): appended to
    gpurun_out/parity_stats.jsonl so that the thresholds in the tests can be set from — and checked against — measurements."""
    try:
        d = os.path.join(ROOT, "gpurun_out")
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, "parity_stats.jsonl"), "a") as f:
            f.write(json.dumps(dict(stats, test=name)) + "\n")
    except OSError:
        pass
