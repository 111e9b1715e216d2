"""In-tree build of the native libraries (nvcc cross-compiles sm_100a without a GPU).

    raytracingweekend_b200/lib/librtw_b200_host.so   builder + flattener + scenes + PPM (no CUDA)
    raytracingweekend_b200/lib/librtw_b200.so        the above + CUDA kernels + C-ABI hot path

Both are git-ignored (*.so) and travel to the GPU box with the gpurun snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "lib")
OBJ = os.path.join(HERE, "build")

HOST_SRCS = ["builder.cpp", "scenes.cpp", "hostio.cpp", "imageio.cpp"]
CUDA_SRCS = ["kernels.cu", "sorted.cu", "sorted_big.cu", "wavefront.cu", "lbvh.cu", "api.cu"]
NVCC_ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
CXXFLAGS = ["-std=c++17", "-O2", "-fPIC", "-Wall"]
# FP32 division / sqrt as reciprocal-multiply and MUFU forms (2 ulp), denormals flushed; sin/cos/log/exp stay accurate
NVCC_MATH = ["-prec-div=false", "-prec-sqrt=false", "-ftz=true"]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _stale(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd: list[str]) -> None:
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        raise RuntimeError("build failed: " + cmd[0])


def _headers() -> list[str]:
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".hpp", ".cuh", ".h"))]
    hs.append(os.path.join(HERE, "..", "include", "rtw_b200.h"))
    return hs


def build(force: bool = False, verbose: bool = False, extra: list[str] | None = None, tag: str = "") -> str:
    """Compile everything that is out of date; returns the path of librtw_b200.so."""
    os.makedirs(LIB, exist_ok=True)
    os.makedirs(OBJ, exist_ok=True)
    hdrs = _headers()
    host_objs = []
    for s in HOST_SRCS:
        src, obj = os.path.join(CSRC, s), os.path.join(OBJ, s + ".o")
        if force or _stale(obj, [src] + hdrs):
            _run(["g++", *CXXFLAGS, "-c", src, "-o", obj])
        host_objs.append(obj)
    host_so = os.path.join(LIB, "librtw_b200_host.so")
    if force or _stale(host_so, host_objs):
        _run(["g++", "-shared", "-o", host_so, *host_objs])
    nvcc = _nvcc()
    cuda_objs, cmds = [], []
    for s in CUDA_SRCS:
        src, obj = os.path.join(CSRC, s), os.path.join(OBJ, s + tag + ".o")
        if force or _stale(obj, [src] + hdrs):
            cmd = [nvcc, *NVCC_ARCH, "-O3", "-lineinfo", "-std=c++17", *NVCC_MATH, *(extra or []), "-Xcompiler", "-fPIC", "-c", src, "-o", obj]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            cmds.append(cmd)
        cuda_objs.append(obj)
    if cmds:  # the translation units are independent: compile them side by side (the kernels' TUs take a minute each)
        from concurrent.futures import ThreadPoolExecutor

        with ThreadPoolExecutor(max_workers=min(len(cmds), os.cpu_count() or 1)) as pool:
            list(pool.map(_run, cmds))
    full_so = os.path.join(LIB, f"librtw_b200{tag}.so")
    if force or _stale(full_so, host_objs + cuda_objs):
        _run([nvcc, *NVCC_ARCH, "-shared", "-o", full_so, *host_objs, *cuda_objs, "-cudart", "static"])
    return full_so


def build_count(force: bool = False) -> str:
    """Diagnostics twin of the library with device-side traversal event counters (tools/count_tree_events.py; bench.py uses
    it for roofline.kernel_tree).  Never loaded by the product path."""
    return build(force=force, extra=["-DRTW_COUNT_EVENTS"], tag="_count")


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
